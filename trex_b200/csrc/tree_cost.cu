// Soft-adjacency tree cost: value and all gradients in one call.
//
// Replaces compute_surrogate_cost / compute_soft_cost (src/trex/tree.py:163-209, 212-266)
// and the backward XLA derives from them under jax.value_and_grad
// (src/trex/evals/benchmark.py:167-173).  With r/c the row/column sums of Adj and
//     M = diag((r+c)/2) - Adj,   W = S | S*c | S@C  (tree.py:248-253),  G = S_flat W_flat^T
// the reference's (term1+term2+term3)/2 (tree.py:262-266) is  cost = <M, G>, and
//     d/dAdj_ij = (G_ii + G_jj)/2 - G_ij
//     d/dS      = (M S) C + (M^T S) C^T          (C = I: (M + M^T) S)
//     d/dC_qp   = sum_{j,l} S_jlq (M^T S)_jlp     (vector c: the diagonal of that)
//
// This translation unit is the exact-f32 CUDA-core implementation (split over the contraction
// axis K = L*Q, f64 cross-CTA accumulation of G).  It is the parity baseline for the tcgen05
// path and the fallback for shapes the tensor-core kernel does not take.
#include "common.cuh"
#include "tree_cost_tc.cuh"

namespace trex {

constexpr int kTile = 64;    // output tile edge
constexpr int kKcMax = 64;   // contraction columns staged per step

struct TcWorkspace {
  double* G;        // [N][N]
  double* dC;       // [Q][Q]
  double* cost;     // [1]
  unsigned* ticket; // [1]
  float* d;         // [N] (r+c)/2
};

__host__ __device__ inline size_t tc_align(size_t x) { return (x + 255) & ~(size_t)255; }

__host__ __device__ inline TcWorkspace tc_carve(void* ws, int N, int Q) {
  char* p = (char*)ws;
  TcWorkspace w;
  w.G = (double*)p;          p += tc_align(sizeof(double) * (size_t)N * N);
  w.dC = (double*)p;         p += tc_align(sizeof(double) * (size_t)Q * Q);
  w.cost = (double*)p;       p += 256;
  w.ticket = (unsigned*)p;   p += 256;
  w.d = (float*)p;
  return w;
}

// ---- 1: d_i = (rowsum_i + colsum_i)/2; zero the accumulators -------------------------------
__global__ void __launch_bounds__(128)
    tc_prologue_kernel(const float* __restrict__ adj, int N, int Q, TcWorkspace w) {
  const int i = blockIdx.x;
  float acc = 0.f;
  for (int j = threadIdx.x; j < N; j += 128) {
    acc += adj[(size_t)i * N + j] + adj[(size_t)j * N + i];
    w.G[(size_t)i * N + j] = 0.0;
  }
  __shared__ float red[4];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) w.d[i] = 0.5f * (red[0] + red[1] + red[2] + red[3]);
  if (i == 0) {
    for (int j = threadIdx.x; j < Q * Q; j += 128) w.dC[j] = 0.0;
    if (threadIdx.x == 0) { *w.cost = 0.0; *w.ticket = 0u; }
  }
}

// weighted element W[row][col] of tree.py:248-253 straight from global S
__device__ __forceinline__ float weighted_elem(const float* __restrict__ srow, long long col, int Q,
                                               const float* __restrict__ sC, int kind) {
  if (kind == 0) return srow[col];
  const int p = (int)(col % Q);
  if (kind == 1) return srow[col] * sC[p];
  const long long base = col - p;
  float a = 0.f;
  for (int q = 0; q < Q; ++q) a = fmaf(srow[base + q], sC[q * Q + p], a);
  return a;
}

// ---- 2: partial Gram tiles, split over K ---------------------------------------------------
__global__ void __launch_bounds__(256)
    tc_gram_kernel(const float* __restrict__ S, const float* __restrict__ cost, int kind, int N,
                   long long K, int Q, int kc, long long cols_per_part, TcWorkspace w) {
  __shared__ float sS[kTile][kKcMax + 1];
  __shared__ float sW[kTile][kKcMax + 1];
  __shared__ float sC[32 * 32];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int i0 = blockIdx.y * kTile, j0 = blockIdx.z * kTile;
  const int nC = kind == 2 ? Q * Q : (kind == 1 ? Q : 0);
  for (int i = tid; i < nC; i += 256) sC[i] = cost[i];
  const long long kbeg = (long long)blockIdx.x * cols_per_part;
  const long long kend = min(K, kbeg + cols_per_part);
  float acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
  __syncthreads();
  for (long long k0 = kbeg; k0 < kend; k0 += kc) {
    const int kw = (int)min((long long)kc, kend - k0);
    for (int e = tid; e < kTile * kc; e += 256) {
      const int r = e / kc, c = e % kc;
      float sv = 0.f, wv = 0.f;
      if (c < kw) {
        if (i0 + r < N) sv = S[(size_t)(i0 + r) * K + k0 + c];
        if (j0 + r < N) wv = weighted_elem(S + (size_t)(j0 + r) * K, k0 + c, Q, sC, kind);
      }
      sS[r][c] = sv;
      sW[r][c] = wv;
    }
    __syncthreads();
    for (int k = 0; k < kw; ++k) {
      float a[4], b[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) { a[q] = sS[ty * 4 + q][k]; b[q] = sW[tx * 4 + q][k]; }
#pragma unroll
      for (int x = 0; x < 4; ++x)
#pragma unroll
        for (int y = 0; y < 4; ++y) acc[x][y] = fmaf(a[x], b[y], acc[x][y]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int x = 0; x < 4; ++x)
#pragma unroll
    for (int y = 0; y < 4; ++y) {
      const int i = i0 + ty * 4 + x, j = j0 + tx * 4 + y;
      if (i < N && j < N) atomicAdd(w.G + (size_t)i * N + j, (double)acc[x][y]);
    }
}

// ---- 3: U = M S, Ut = M^T S on a (64 rows x chunk) panel; dS and dC from them -----------------
__global__ void __launch_bounds__(256)
    tc_ds_kernel(const float* __restrict__ S, const float* __restrict__ adj,
                 const float* __restrict__ cost, int kind, int N, long long K, int Q, int kc,
                 float* __restrict__ dS, bool want_dc, TcWorkspace w) {
  extern __shared__ __align__(16) float ds_smem[];
  float (*sS)[kKcMax + 1] = reinterpret_cast<float (*)[kKcMax + 1]>(ds_smem);  // S[j-block][chunk]; later S[i-tile][chunk]
  float (*sM)[kTile + 1] = reinterpret_cast<float (*)[kTile + 1]>(ds_smem + kTile * (kKcMax + 1));  // M[i-tile][j-block]; later U
  float (*sMt)[kTile + 1] = reinterpret_cast<float (*)[kTile + 1]>(ds_smem + kTile * (kKcMax + 1) + kTile * (kTile + 1));  // M^T; later Ut
  float* sC = ds_smem + kTile * (kKcMax + 1) + 2 * kTile * (kTile + 1);
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int i0 = blockIdx.y * kTile;
  const long long k0 = (long long)blockIdx.x * kc;
  const int kw = (int)min((long long)kc, K - k0);
  const int nC = kind == 2 ? Q * Q : (kind == 1 ? Q : 0);
  for (int i = tid; i < nC; i += 256) sC[i] = cost[i];
  float u[4][4], ut[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) u[a][b] = ut[a][b] = 0.f;
  for (int j0 = 0; j0 < N; j0 += kTile) {
    __syncthreads();
    for (int e = tid; e < kTile * kc; e += 256) {
      const int r = e / kc, c = e % kc;
      sS[r][c] = (c < kw && j0 + r < N) ? S[(size_t)(j0 + r) * K + k0 + c] : 0.f;
    }
    for (int e = tid; e < kTile * kTile; e += 256) {
      const int r = e / kTile, c = e % kTile;  // r: i offset, c: j offset
      const int i = i0 + r, j = j0 + c;
      float m = 0.f, mt = 0.f;
      if (i < N && j < N) {
        const float dd = (i == j) ? w.d[i] : 0.f;
        m = dd - adj[(size_t)i * N + j];
        mt = dd - adj[(size_t)j * N + i];
      }
      sM[r][c] = m;
      sMt[r][c] = mt;
    }
    __syncthreads();
    for (int j = 0; j < kTile; ++j) {
      float sv[4], mv[4], mtv[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        sv[q] = sS[j][tx * 4 + q];
        mv[q] = sM[ty * 4 + q][j];
        mtv[q] = sMt[ty * 4 + q][j];
      }
#pragma unroll
      for (int x = 0; x < 4; ++x)
#pragma unroll
        for (int y = 0; y < 4; ++y) {
          u[x][y] = fmaf(mv[x], sv[y], u[x][y]);
          ut[x][y] = fmaf(mtv[x], sv[y], ut[x][y]);
        }
    }
  }
  __syncthreads();
  // park U / Ut ([row][col]) and reload S[i-tile][chunk] for dC
#pragma unroll
  for (int x = 0; x < 4; ++x)
#pragma unroll
    for (int y = 0; y < 4; ++y) {
      sM[ty * 4 + x][tx * 4 + y] = u[x][y];
      sMt[ty * 4 + x][tx * 4 + y] = ut[x][y];
    }
  for (int e = tid; e < kTile * kc; e += 256) {
    const int r = e / kc, c = e % kc;
    sS[r][c] = (c < kw && i0 + r < N) ? S[(size_t)(i0 + r) * K + k0 + c] : 0.f;
  }
  __syncthreads();
  if (dS) {
    for (int e = tid; e < kTile * kc; e += 256) {
      const int r = e / kc, c = e % kc;
      if (c >= kw || i0 + r >= N) continue;
      float v;
      if (kind == 0) {
        v = sM[r][c] + sMt[r][c];
      } else if (kind == 1) {
        v = (sM[r][c] + sMt[r][c]) * sC[(k0 + c) % Q];
      } else {
        const int q = (int)((k0 + c) % Q);
        const int base = c - q;  // chunk starts on a site boundary
        v = 0.f;
        for (int pp = 0; pp < Q; ++pp) {
          v = fmaf(sM[r][base + pp], sC[pp * Q + q], v);    // (U @ C)[q]
          v = fmaf(sMt[r][base + pp], sC[q * Q + pp], v);   // (Ut @ C^T)[q]
        }
      }
      dS[(size_t)(i0 + r) * K + k0 + c] = v;
    }
  }
  if (want_dc && kind != 0) {
    const int sites = kw / Q;
    const int npair = kind == 2 ? Q * Q : Q;
    for (int pr = tid; pr < npair; pr += 256) {
      const int q = kind == 2 ? pr / Q : pr, pp = kind == 2 ? pr % Q : pr;
      float a = 0.f;
      for (int r = 0; r < kTile; ++r)
        for (int l = 0; l < sites; ++l) a = fmaf(sS[r][l * Q + q], sMt[r][l * Q + pp], a);
      if (a != 0.f) atomicAdd(w.dC + pr, (double)a);
    }
  }
}

// ---- 4: cost = <M, G>, dAdj, dC; the last block writes the scalars ----------------------------
__global__ void __launch_bounds__(256)
    tc_finalize_kernel(const float* __restrict__ adj, int N, int Q, int kind,
                       float* __restrict__ out_cost, float* __restrict__ d_adj,
                       float* __restrict__ d_cost, TcWorkspace w) {
  const long long e = (long long)blockIdx.x * 256 + threadIdx.x;
  double part = 0.0;
  if (e < (long long)N * N) {
    const int i = (int)(e / N), j = (int)(e % N);
    const double g = w.G[e];
    const double m = (i == j ? (double)w.d[i] : 0.0) - (double)adj[e];
    part = m * g;
    if (d_adj) d_adj[e] = (float)(0.5 * (w.G[(size_t)i * N + i] + w.G[(size_t)j * N + j]) - g);
  }
  block_atomic_add(part, w.cost);
  __shared__ bool last;
  __threadfence();
  if (threadIdx.x == 0) last = (atomicAdd(w.ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (last) {
    __threadfence();
    if (threadIdx.x == 0) *out_cost = (float)(*(volatile double*)w.cost);
    if (d_cost) {
      const int n = kind == 2 ? Q * Q : (kind == 1 ? Q : 0);
      for (int i = threadIdx.x; i < n; i += 256) d_cost[i] = (float)w.dC[i];
    }
  }
}

}  // namespace trex

extern "C" {

size_t trex_tree_cost_workspace_bytes(int n_nodes, int64_t k, int n_states) {
  using namespace trex;
  if (n_nodes < 1 || n_states < 1) return 0;
  const size_t tcb = n_nodes <= tc::kMaxNodes ? tc::workspace_bytes(n_nodes, k) : 0;  // tcgen05 path (either path may run)
  const size_t simt = tc_align(sizeof(double) * (size_t)n_nodes * n_nodes) +
         tc_align(sizeof(double) * (size_t)n_states * n_states) + 512 +
         tc_align(sizeof(float) * (size_t)n_nodes);
  return simt > tcb ? simt : tcb;
}

int trex_tree_cost_fwd_bwd(const float* seqs, const float* adj, const float* cost, int cost_kind,
                           int n_nodes, int64_t n_sites, int n_states, float* out_cost,
                           float* d_seqs, float* d_adj, float* d_cost, void* workspace,
                           void* stream) {
  using namespace trex;
  TREX_CHECK_ARG(seqs && adj && out_cost && workspace, "trex_tree_cost_fwd_bwd: null pointer");
  TREX_CHECK_ARG(cost_kind >= 0 && cost_kind <= 2, "trex_tree_cost_fwd_bwd: cost_kind must be 0, 1 or 2");
  TREX_CHECK_ARG((cost_kind == 0) == (cost == nullptr), "trex_tree_cost_fwd_bwd: cost pointer / cost_kind mismatch");
  TREX_CHECK_ARG(n_nodes >= 1 && n_nodes <= 65535 * 8 && n_sites >= 1 && n_states >= 1 && n_states <= 32,
                 "trex_tree_cost_fwd_bwd: bad sizes (n_states must be <= 32)");
  const int N = n_nodes, Q = n_states;
  const long long K = (long long)n_sites * Q;
  cudaStream_t s = (cudaStream_t)stream;
  // identity cost, N <= 256: tcgen05 / TMEM / TMA path (3xTF32); everything else: exact-f32 CUDA cores
  if (tc::supported(seqs, N, K, cost_kind)) return tc::run(seqs, adj, N, K, out_cost, d_seqs, d_adj, workspace, s);
  TcWorkspace w = tc_carve(workspace, N, Q);
  tc_prologue_kernel<<<N, 128, 0, s>>>(adj, N, Q, w);
  note_launches(1);
  TREX_CHECK_LAUNCH("tc_prologue_kernel");
  const int spc = kKcMax / Q > 0 ? kKcMax / Q : 1;  // whole sites per staged chunk
  const int kc = spc * Q;
  const int tiles = (N + kTile - 1) / kTile;
  {
    const long long nchunks = (K + kc - 1) / kc;
    long long parts = (2LL * sm_count() + (long long)tiles * tiles - 1) / ((long long)tiles * tiles);
    if (parts > nchunks) parts = nchunks;
    if (parts < 1) parts = 1;
    const long long chunks_per_part = (nchunks + parts - 1) / parts;
    parts = (nchunks + chunks_per_part - 1) / chunks_per_part;
    dim3 grid((unsigned)parts, tiles, tiles);
    tc_gram_kernel<<<grid, 256, 0, s>>>(seqs, cost, cost_kind, N, K, Q, kc, chunks_per_part * kc, w);
    note_launches(1);
    TREX_CHECK_LAUNCH("tc_gram_kernel");
  }
  const bool want_dc = d_cost != nullptr && cost_kind != 0;
  if (d_seqs || want_dc) {
    const long long nchunks = (K + kc - 1) / kc;
    TREX_CHECK_ARG(nchunks <= 2147483647LL, "trex_tree_cost_fwd_bwd: K too large");
    dim3 grid((unsigned)nchunks, tiles);
    const size_t smem = sizeof(float) * (kTile * (kKcMax + 1) + 2 * kTile * (kTile + 1) + 32 * 32);
    int rc = check_cuda(cudaFuncSetAttribute(tc_ds_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                        "cudaFuncSetAttribute(tc_ds_kernel)");
    if (rc) return rc;
    tc_ds_kernel<<<grid, 256, smem, s>>>(seqs, adj, cost, cost_kind, N, K, Q, kc, d_seqs, want_dc, w);
    note_launches(1);
    TREX_CHECK_LAUNCH("tc_ds_kernel");
  }
  {
    const long long n2 = (long long)N * N;
    tc_finalize_kernel<<<(unsigned)((n2 + 255) / 256), 256, 0, s>>>(adj, N, Q, cost_kind, out_cost, d_adj,
                                                                    d_cost, w);
    note_launches(1);
    TREX_CHECK_LAUNCH("tc_finalize_kernel");
  }
  return TREX_OK;
}

}  // extern "C"
