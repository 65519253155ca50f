// The small operators around the tree cost: SURVEY §8a rows T3-T7 and their reverse modes.
//   T3 update_seq                 src/trex/tree.py:110-130, src/trex/evals/benchmark.py:210-232
//   T4 update_tree                src/trex/tree.py:50-107   (Gumbel noise supplied by the caller)
//   T5 enforce_graph_constraints  src/trex/tree.py:133-160
//   T6 compute_cost               src/trex/tree.py:269-296
//   T7 discretize_tree_topology   src/trex/tree.py:31-47
#include "common.cuh"

namespace trex {

// ---- T3: rows n_leaves.. of S = softmax(logits * temperature) over the last axis ------------
__global__ void update_seq_fwd_kernel(const float* __restrict__ logits, long long rows, int Q,
                                      float temperature, float* __restrict__ out) {
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  const float* x = logits + r * Q;
  float m = -INFINITY;
  for (int q = 0; q < Q; ++q) m = fmaxf(m, x[q] * temperature);
  if (!(m > -INFINITY && m < INFINITY)) m = 0.f;
  float z = 0.f;
  for (int q = 0; q < Q; ++q) z += expf(x[q] * temperature - m);
  const float inv = 1.f / z;
  for (int q = 0; q < Q; ++q) out[r * Q + q] = expf(x[q] * temperature - m) * inv;
}

// dlogits = temperature * S * (dS - sum_q dS*S)
__global__ void update_seq_bwd_kernel(const float* __restrict__ S, const float* __restrict__ dS,
                                      long long rows, int Q, float temperature,
                                      float* __restrict__ dlogits) {
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  float dot = 0.f;
  for (int q = 0; q < Q; ++q) dot = fmaf(S[r * Q + q], dS[r * Q + q], dot);
  for (int q = 0; q < Q; ++q) dlogits[r * Q + q] = temperature * S[r * Q + q] * (dS[r * Q + q] - dot);
}

// ---- T4: masked logits + row softmax ---------------------------------------------------------
// row i < n_leaves: columns n_leaves.. get p[i][:]; ancestor row i (< N-1): only columns with
// (j - n_leaves) > (i - n_leaves); root row: only [N-1][N-1] = 1.0; everything else -inf.
__device__ __forceinline__ float tree_logit(const float* __restrict__ params, const float* __restrict__ noise,
                                            const float* __restrict__ gates, float temperature, int N,
                                            int n_leaves, int i, int j) {
  const int n_anc = N - n_leaves;
  if (i == N - 1) return j == N - 1 ? 1.0f : -INFINITY;
  if (j < n_leaves) return -INFINITY;
  const int a = j - n_leaves;
  if (i >= n_leaves && a <= i - n_leaves) return -INFINITY;
  float p = params[(size_t)i * n_anc + a];
  if (noise) p += noise[(size_t)i * n_anc + a];
  if (gates) p *= gates[(size_t)i * n_anc + a];
  return p / temperature;
}

__global__ void __launch_bounds__(128)
    update_tree_fwd_kernel(const float* __restrict__ params, const float* __restrict__ noise,
                           const float* __restrict__ gates, float temperature, int N, int n_leaves,
                           float* __restrict__ adj) {
  const int i = blockIdx.x;
  __shared__ float red[4];
  float m = -INFINITY;
  for (int j = threadIdx.x; j < N; j += 128) m = fmaxf(m, tree_logit(params, noise, gates, temperature, N, n_leaves, i, j));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
  __syncthreads();
  m = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
  if (!(m > -INFINITY && m < INFINITY)) m = 0.f;
  __syncthreads();
  float z = 0.f;
  for (int j = threadIdx.x; j < N; j += 128) z += expf(tree_logit(params, noise, gates, temperature, N, n_leaves, i, j) - m);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) z += __shfl_xor_sync(0xffffffffu, z, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = z;
  __syncthreads();
  z = red[0] + red[1] + red[2] + red[3];
  for (int j = threadIdx.x; j < N; j += 128)
    adj[(size_t)i * N + j] = expf(tree_logit(params, noise, gates, temperature, N, n_leaves, i, j) - m) / z;
}

// dparams[i][a] = gates * Adj_ij (dAdj_ij - sum_j dAdj_ij Adj_ij) / temperature on unmasked entries
__global__ void __launch_bounds__(128)
    update_tree_bwd_kernel(const float* __restrict__ adj, const float* __restrict__ dadj,
                           const float* __restrict__ gates, float temperature, int N, int n_leaves,
                           float* __restrict__ dparams) {
  const int i = blockIdx.x;  // 0 .. N-2
  const int n_anc = N - n_leaves;
  __shared__ float red[4];
  float dot = 0.f;
  for (int j = threadIdx.x; j < N; j += 128) dot = fmaf(adj[(size_t)i * N + j], dadj[(size_t)i * N + j], dot);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = dot;
  __syncthreads();
  dot = red[0] + red[1] + red[2] + red[3];
  for (int a = threadIdx.x; a < n_anc; a += 128) {
    const int j = n_leaves + a;
    float gval = 0.f;
    if (!(i >= n_leaves && a <= i - n_leaves)) {
      gval = adj[(size_t)i * N + j] * (dadj[(size_t)i * N + j] - dot) / temperature;
      if (gates) gval *= gates[(size_t)i * n_anc + a];
    }
    dparams[(size_t)i * n_anc + a] = gval;
  }
}

// ---- T5: scale * sum_anc (colsum(adj[:-1, anc]) - 2)^2 and its gradient ----------------------
__global__ void __launch_bounds__(128)
    graph_constraint_kernel(const float* __restrict__ adj, int N, int n_anc, float scale,
                            double* __restrict__ acc, float* __restrict__ dadj) {
  const int a = blockIdx.x;  // ancestor column index 0..n_anc-1 -> column N - n_anc + a
  const int col = N - n_anc + a;
  __shared__ float red[4];
  float s = 0.f;
  for (int i = threadIdx.x; i < N - 1; i += 128) s += adj[(size_t)i * N + col];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  s = red[0] + red[1] + red[2] + red[3];
  const float d = s - 2.f;
  if (threadIdx.x == 0) atomicAdd(acc, (double)scale * (double)d * (double)d);
  if (dadj)
    for (int i = threadIdx.x; i < N - 1; i += 128) dadj[(size_t)i * N + col] = 2.f * scale * d;
}

__global__ void finish_scalar_kernel(const double* __restrict__ acc, float* __restrict__ out) {
  *out = (float)*acc;
}

// ---- T7 / T6 -------------------------------------------------------------------------------
__global__ void row_argmax_kernel(const float* __restrict__ x, long long rows, int cols,
                                  int* __restrict__ idx) {
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  const float* p = x + r * cols;
  float b = p[0];
  int bi = 0;
  for (int c = 1; c < cols; ++c)
    if (p[c] > b) { b = p[c]; bi = c; }  // first maximum (jnp.argmax)
  idx[r] = bi;
}

__global__ void one_hot_rows_kernel(const int* __restrict__ idx, int N, float* __restrict__ out) {
  const int i = blockIdx.x;
  for (int j = threadIdx.x; j < N; j += blockDim.x) out[(size_t)i * N + j] = (idx[i] == j) ? 1.f : 0.f;
}

// sum over nodes 0..N-2 and sites of subst[state[parent[node]][l]][state[node][l]] (tree.py:296)
__global__ void __launch_bounds__(256)
    compute_cost_kernel(const int* __restrict__ state, const int* __restrict__ parent,
                        const float* __restrict__ subst, int N, long long L, int Q,
                        double* __restrict__ acc) {
  const long long l = (long long)blockIdx.x * 256 + threadIdx.x;
  const int node = blockIdx.y;
  double v = 0.0;
  if (l < L) {
    const int ps = state[(size_t)parent[node] * L + l], cs = state[(size_t)node * L + l];
    v = (double)subst[ps * Q + cs];
  }
  block_atomic_add(v, acc);
}

}  // namespace trex

extern "C" {

int trex_update_seq_fwd(const float* logits, int n_ancestors, int64_t n_sites, int n_states,
                        float temperature, float* seqs_ancestor_rows, void* stream) {
  using namespace trex;
  TREX_CHECK_ARG(logits && seqs_ancestor_rows && n_ancestors >= 0 && n_sites >= 1 && n_states >= 1,
                 "trex_update_seq_fwd: bad arguments");
  const long long rows = (long long)n_ancestors * n_sites;
  if (rows == 0) return TREX_OK;
  update_seq_fwd_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      logits, rows, n_states, temperature, seqs_ancestor_rows);
  note_launches(1);
  TREX_CHECK_LAUNCH("update_seq_fwd_kernel");
  return TREX_OK;
}

int trex_update_seq_bwd(const float* seqs_ancestor_rows, const float* d_seqs_ancestor_rows,
                        int n_ancestors, int64_t n_sites, int n_states, float temperature,
                        float* d_logits, void* stream) {
  using namespace trex;
  TREX_CHECK_ARG(seqs_ancestor_rows && d_seqs_ancestor_rows && d_logits && n_sites >= 1 && n_states >= 1,
                 "trex_update_seq_bwd: bad arguments");
  const long long rows = (long long)n_ancestors * n_sites;
  if (rows == 0) return TREX_OK;
  update_seq_bwd_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      seqs_ancestor_rows, d_seqs_ancestor_rows, rows, n_states, temperature, d_logits);
  note_launches(1);
  TREX_CHECK_LAUNCH("update_seq_bwd_kernel");
  return TREX_OK;
}

int trex_update_tree_fwd(const float* tree_params, const float* gumbel_noise, const float* gates,
                         float temperature, int n_nodes, int n_leaves, float* adjacency, void* stream) {
  using namespace trex;
  TREX_CHECK_ARG(tree_params && adjacency && n_nodes > n_leaves && n_leaves >= 1 && temperature != 0.f,
                 "trex_update_tree_fwd: bad arguments (n_ancestors == 0 is handled by the caller: eye)");
  update_tree_fwd_kernel<<<n_nodes, 128, 0, (cudaStream_t)stream>>>(tree_params, gumbel_noise, gates, temperature,
                                                                    n_nodes, n_leaves, adjacency);
  note_launches(1);
  TREX_CHECK_LAUNCH("update_tree_fwd_kernel");
  return TREX_OK;
}

int trex_update_tree_bwd(const float* adjacency, const float* d_adjacency, const float* gates,
                         float temperature, int n_nodes, int n_leaves, float* d_tree_params, void* stream) {
  using namespace trex;
  TREX_CHECK_ARG(adjacency && d_adjacency && d_tree_params && n_nodes > n_leaves && n_leaves >= 1,
                 "trex_update_tree_bwd: bad arguments");
  update_tree_bwd_kernel<<<n_nodes - 1, 128, 0, (cudaStream_t)stream>>>(adjacency, d_adjacency, gates, temperature,
                                                                        n_nodes, n_leaves, d_tree_params);
  note_launches(1);
  TREX_CHECK_LAUNCH("update_tree_bwd_kernel");
  return TREX_OK;
}

int trex_graph_constraint(const float* adjacency, int n_nodes, float scaling_factor, float* out_loss,
                          float* d_adjacency, double* scratch, void* stream) {
  using namespace trex;
  TREX_CHECK_ARG(adjacency && out_loss && scratch && n_nodes >= 1, "trex_graph_constraint: bad arguments");
  cudaStream_t s = (cudaStream_t)stream;
  int n_anc = (n_nodes - 1) / 2;              // tree.py:157
  if (n_anc == 0) n_anc = n_nodes;            // adjacency[:-1, -0:] is the full slice
  int rc = check_cuda(cudaMemsetAsync(scratch, 0, sizeof(double), s), "memset");
  if (rc) return rc;
  if (d_adjacency) {
    rc = check_cuda(cudaMemsetAsync(d_adjacency, 0, sizeof(float) * (size_t)n_nodes * n_nodes, s), "memset");
    if (rc) return rc;
  }
  graph_constraint_kernel<<<n_anc, 128, 0, s>>>(adjacency, n_nodes, n_anc, scaling_factor, scratch, d_adjacency);
  note_launches(1);
  TREX_CHECK_LAUNCH("graph_constraint_kernel");
  finish_scalar_kernel<<<1, 1, 0, s>>>(scratch, out_loss);
  note_launches(1);
  TREX_CHECK_LAUNCH("finish_scalar_kernel");
  return TREX_OK;
}

int trex_discretize_topology(const float* adjacency, int n_nodes, int32_t* parent_scratch,
                             float* one_hot, void* stream) {
  using namespace trex;
  TREX_CHECK_ARG(adjacency && parent_scratch && one_hot && n_nodes >= 1, "trex_discretize_topology: bad arguments");
  cudaStream_t s = (cudaStream_t)stream;
  row_argmax_kernel<<<(n_nodes + 127) / 128, 128, 0, s>>>(adjacency, n_nodes, n_nodes, parent_scratch);
  note_launches(1);
  TREX_CHECK_LAUNCH("row_argmax_kernel");
  one_hot_rows_kernel<<<n_nodes, 128, 0, s>>>(parent_scratch, n_nodes, one_hot);
  note_launches(1);
  TREX_CHECK_LAUNCH("one_hot_rows_kernel");
  return TREX_OK;
}

int trex_compute_cost(const float* seqs_one_hot, const float* adjacency, const float* subst,
                      int n_nodes, int64_t n_sites, int n_states, int32_t* state_scratch,
                      int32_t* parent_scratch, double* acc_scratch, float* out_cost, void* stream) {
  using namespace trex;
  TREX_CHECK_ARG(seqs_one_hot && adjacency && subst && state_scratch && parent_scratch && acc_scratch && out_cost,
                 "trex_compute_cost: null pointer");
  TREX_CHECK_ARG(n_nodes >= 1 && n_nodes <= 65535 && n_sites >= 1 && n_states >= 1, "trex_compute_cost: bad sizes");
  cudaStream_t s = (cudaStream_t)stream;
  const long long rows = (long long)n_nodes * n_sites;
  int rc = check_cuda(cudaMemsetAsync(acc_scratch, 0, sizeof(double), s), "memset");
  if (rc) return rc;
  row_argmax_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, s>>>(seqs_one_hot, rows, n_states, state_scratch);
  note_launches(1);
  TREX_CHECK_LAUNCH("row_argmax_kernel");
  row_argmax_kernel<<<(n_nodes + 127) / 128, 128, 0, s>>>(adjacency, n_nodes, n_nodes, parent_scratch);
  note_launches(1);
  TREX_CHECK_LAUNCH("row_argmax_kernel");
  if (n_nodes > 1) {
    dim3 grid((unsigned)((n_sites + 255) / 256), n_nodes - 1);  // [:-1] drops the last node
    compute_cost_kernel<<<grid, 256, 0, s>>>(state_scratch, parent_scratch, subst, n_nodes, n_sites, n_states, acc_scratch);
    note_launches(1);
    TREX_CHECK_LAUNCH("compute_cost_kernel");
  }
  finish_scalar_kernel<<<1, 1, 0, s>>>(acc_scratch, out_cost);
  note_launches(1);
  TREX_CHECK_LAUNCH("finish_scalar_kernel");
  return TREX_OK;
}


// ---- no-allocation variant of trex_compute_cost -------------------------------------------------------
// No counterpart in the reference: added for callers that cannot call
// cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).  The three scratch
// buffers of trex_compute_cost live in ONE file-scope device array of a fixed, modest size, resident on every
// GPU that loads the library until the process exits:
//   [0, 256)                      f64 accumulator (must start at zero: cleared by every call)
//   [256, 256 + 4 * kNaMaxNodes)  i32 parent[n_nodes]   (fully overwritten by every call)
//   [.., + 4 * kNaMaxCells)       i32 state[n_nodes][n_sites] (fully overwritten by every call)
// In host code the array's name is only a host-side stand-in for the symbol, not a device address: the address
// is looked up once per device with cudaGetSymbolAddress (that device current) and cached.
}  // extern "C" (closed for the file-scope definitions below)

#include <mutex>

namespace {
constexpr int kNaMaxNodes = TREX_NOALLOC_MAX_NODES;
constexpr long long kNaMaxCells = TREX_NOALLOC_MAX_CELLS;
constexpr size_t kNaParentOff = 256;
constexpr size_t kNaStateOff = kNaParentOff + sizeof(int32_t) * (size_t)kNaMaxNodes;
constexpr size_t kNaBytes = kNaStateOff + sizeof(int32_t) * (size_t)kNaMaxCells;
static_assert(kNaStateOff % 256 == 0, "state table must stay 256-byte aligned");
}  // namespace

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNaBytes];

namespace {
std::mutex g_na_mutex;
unsigned char* g_na_ptr[64] = {nullptr};  // device address of g_noalloc_scratch, per device ordinal

// device address of the scratch array on the CURRENT device (looked up once per device); nullptr + error set on failure
unsigned char* noalloc_scratch_ptr() {
  int dev = -1;
  if (trex::check_cuda(cudaGetDevice(&dev), "cudaGetDevice")) return nullptr;
  if (dev < 0 || dev >= 64) {
    trex::set_error("trex_compute_cost_noalloc: device ordinal %d out of range", dev);
    return nullptr;
  }
  std::lock_guard<std::mutex> lock(g_na_mutex);
  if (!g_na_ptr[dev]) {
    void* p = nullptr;
    if (trex::check_cuda(cudaGetSymbolAddress(&p, g_noalloc_scratch), "cudaGetSymbolAddress(g_noalloc_scratch)"))
      return nullptr;
    g_na_ptr[dev] = static_cast<unsigned char*>(p);
  }
  return g_na_ptr[dev];
}
}  // namespace

extern "C" {

size_t trex_compute_cost_noalloc_scratch_bytes(void) { return kNaBytes; }

int trex_compute_cost_noalloc(const float* seqs_one_hot, const float* adjacency, const float* subst, int n_nodes,
                              int64_t n_sites, int n_states, float* out_cost, void* stream) {
  using namespace trex;
  TREX_CHECK_ARG(seqs_one_hot && adjacency && subst && out_cost, "trex_compute_cost_noalloc: null pointer");
  TREX_CHECK_ARG(n_nodes >= 1 && n_sites >= 1 && n_states >= 1, "trex_compute_cost_noalloc: bad sizes");
  // size limits first: nothing below runs, and no CUDA call is made, for a problem the static scratch cannot hold
  TREX_CHECK_ARG(n_nodes <= kNaMaxNodes,
                 "trex_compute_cost_noalloc: n_nodes %d exceeds the static scratch limit TREX_NOALLOC_MAX_NODES = %d "
                 "(use trex_compute_cost with caller-provided scratch)", n_nodes, kNaMaxNodes);
  TREX_CHECK_ARG((long long)n_nodes * (long long)n_sites <= kNaMaxCells,
                 "trex_compute_cost_noalloc: n_nodes * n_sites = %lld exceeds the static scratch limit "
                 "TREX_NOALLOC_MAX_CELLS = %lld (use trex_compute_cost with caller-provided scratch)",
                 (long long)n_nodes * (long long)n_sites, kNaMaxCells);
  unsigned char* base = noalloc_scratch_ptr();
  if (!base) return TREX_ERR_CUDA;
  double* acc = reinterpret_cast<double*>(base);
  int32_t* parent = reinterpret_cast<int32_t*>(base + kNaParentOff);
  int32_t* state = reinterpret_cast<int32_t*>(base + kNaStateOff);
  cudaStream_t s = (cudaStream_t)stream;
  const long long rows = (long long)n_nodes * n_sites;
  // the array keeps the previous call's contents: the accumulator is the one part that must start at zero
  int rc = check_cuda(cudaMemsetAsync(acc, 0, sizeof(double), s), "memset");
  if (rc) return rc;
  row_argmax_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, s>>>(seqs_one_hot, rows, n_states, state);
  note_launches(1);
  TREX_CHECK_LAUNCH("row_argmax_kernel");
  row_argmax_kernel<<<(n_nodes + 127) / 128, 128, 0, s>>>(adjacency, n_nodes, n_nodes, parent);
  note_launches(1);
  TREX_CHECK_LAUNCH("row_argmax_kernel");
  if (n_nodes > 1) {
    dim3 grid((unsigned)((n_sites + 255) / 256), n_nodes - 1);  // [:-1] drops the last node
    compute_cost_kernel<<<grid, 256, 0, s>>>(state, parent, subst, n_nodes, n_sites, n_states, acc);
    note_launches(1);
    TREX_CHECK_LAUNCH("compute_cost_kernel");
  }
  finish_scalar_kernel<<<1, 1, 0, s>>>(acc, out_cost);
  note_launches(1);
  TREX_CHECK_LAUNCH("finish_scalar_kernel");
  return TREX_OK;
}

}  // extern "C"
