// Hard-Sankoff companions: leaf encoding, ancestral-state traceback (+ edge histogram) and
// the exports from the native site-contiguous tables to the reference's array layouts.
#include <cstdlib>
#include <type_traits>

#include "common.cuh"
#include "fwd_params.cuh"

namespace trex {

// ---- leaf encoding: sequences[i].astype(int32) + scatter index rules (sankoff.py:49-52) ----
__global__ void encode_leaf_kernel(const float* __restrict__ seqs, int n_leaves, long long n_sites,
                                   long long in_stride, int A, uint8_t* __restrict__ out,
                                   long long stride) {
  const long long l = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int i = blockIdx.y;
  if (l >= stride) return;
  int code = 255;
  if (l < n_sites) {
    int s = (int)seqs[(size_t)i * in_stride + l];  // f32 -> int32 truncates toward zero
    if (s < 0) s += A;                             // negative indices wrap
    if (s >= 0 && s < A) code = s;                 // out-of-bounds updates are dropped
  }
  out[(size_t)i * stride + l] = (uint8_t)code;
}

// ---- traceback -----------------------------------------------------------------------------
// Top-down pointer chase over the reverse post-order of the root's subtree; equivalent to the
// explicit-stack DFS of backtrack_sankoff_jit (sankoff.py:212-265) on a tree.
struct TraceParams {
  const int32_t* plan;
  size_t plan_words;
  int n_all, n_leaves, n_states;
  long long n_sites, stride;
  const uint8_t* leaf;
  const float* dp;
  const uint8_t* arg;
  uint8_t* anc;
  unsigned long long* hist;
  int phase;  // node list of this launch (common.cuh: Walk): 0 whole tree, 1 segment blockIdx.z, 2 top part
  // trex_sankoff_traceback_cost: the forward's cost matrix and mode, to repeat its choice of table format
  const float* cost;
  int mode;
  int compact;  // the forward was asked for pointer words (TREX_MODE_COMPACT_ARG) and n_states allows them
};

template <int V, int THREADS>
__global__ void __launch_bounds__(THREADS) traceback_kernel(const TraceParams p) {
  extern __shared__ unsigned int s_hist[];  // [A*A] when p.hist
  const int A = p.n_states;
  const int tree = blockIdx.y;
  const int n_int = p.n_all - p.n_leaves;
  PlanView plan(p.plan + (size_t)tree * p.plan_words, n_int);
  if (p.phase == 1 && (int)blockIdx.z >= plan.hdr()[2]) return;  // empty segment slice (CTA-uniform)
  // pointer words (sankoff_uniform.cu) or the byte table: the test the forward kernels made, on the same matrix
  bool words = false;
  if (p.compact) {
    float a_uni;
    words = hard_uniform_applies(p.cost, A, p.mode, p.n_all, a_uni);
  }
  if (p.hist) {
    for (int i = threadIdx.x; i < A * A; i += THREADS) s_hist[i] = 0;
    __syncthreads();
  }
  const long long site0 = ((long long)blockIdx.x * THREADS + threadIdx.x) * V;
  if (site0 < p.n_sites) {
    const float* dp = p.dp + (size_t)tree * n_int * A * p.stride + site0;
    const uint8_t* arg = p.arg + ((size_t)tree * n_int * A * p.stride + site0) * 2;
    uint8_t* anc = p.anc + (size_t)tree * n_int * p.stride + site0;
    const uint8_t* leaf = p.leaf + site0;
    const Walk wk = make_walk(plan, p.phase, blockIdx.z);
    const int32_t* order = wk.ord;
    const int n_reach = wk.n_rev;
    const bool has_root = n_reach > 0 && __ldg(order + n_reach - 1) == p.n_all - 1;
    if (has_root) {
      // nodes the root does not reach keep the reference's zero initialisation (sankoff.py:215)
      int zero[V];
#pragma unroll
      for (int v = 0; v < V; ++v) zero[v] = 0;
      for (int i = plan.hdr()[1]; i < n_int; ++i)
        st_u8<V>(anc + (size_t)(__ldg(plan.order() + i) - p.n_leaves) * p.stride, zero);
      // root state: first argmin of the root row (sankoff.py:172)
      const int k = n_int - 1;
      float best[V];
      int st[V];
      ld_f<V>(dp + (size_t)k * A * p.stride, best);
#pragma unroll
      for (int v = 0; v < V; ++v) st[v] = 0;
      for (int s = 1; s < A; ++s) {
        float x[V];
        ld_f<V>(dp + ((size_t)k * A + s) * p.stride, x);
#pragma unroll
        for (int v = 0; v < V; ++v)
          if (x[v] < best[v]) { best[v] = x[v]; st[v] = s; }
      }
      st_u8<V>(anc + (size_t)k * p.stride, st);
    }
    for (int i = n_reach - 1; i >= 0; --i) {
      const int node = __ldg(order + i);
      const int k = node - p.n_leaves;
      int st[V];
      ld_u8<V>(anc + (size_t)k * p.stride, st);
#pragma unroll
      for (int v = 0; v < V; ++v) st[v] = min(st[v], A - 1);  // pad lanes may hold anything
      // both children's back-pointers of (state, site) are one 16-bit word
      const int ch0 = __ldg(plan.child(0) + k), ch1 = __ldg(plan.child(1) + k);
      unsigned pair[V];
#pragma unroll
      for (int v = 0; v < V; ++v) pair[v] = 0;
      if (ch0 >= p.n_leaves || ch1 >= p.n_leaves) {
        if (words) {  // rows 0-3 of the node's table: the 16-bit quarters of (word of slot 0, word of slot 1), coalesced
                      // and the same for every parent state; pointer(s) = bit 5+s ? s : bits 0-4.  (Requesting the next
                      // node's words a node ahead and handing the child's state over in registers was tried: slower,
                      // 0.36 vs 0.34 ms — the walk is bound by instructions, not by the latency of its reads.)
          unsigned q[4][V];  // one access per row: the V 16-bit quarters of this lane's sites
#pragma unroll
          for (int r = 0; r < 4; ++r) {
            const uint8_t* src = arg + ((size_t)k * A + r) * p.stride * 2;
            if constexpr (V == 1) {
              q[r][0] = __ldcs(reinterpret_cast<const unsigned short*>(src));
            } else if constexpr (V == 2) {
              const unsigned x = __ldcs(reinterpret_cast<const unsigned*>(src));
              q[r][0] = x & 0xffffu; q[r][1] = x >> 16;
            } else {
              const uint2 x = __ldcs(reinterpret_cast<const uint2*>(src));
              q[r][0] = x.x & 0xffffu; q[r][1] = x.x >> 16; q[r][2] = x.y & 0xffffu; q[r][3] = x.y >> 16;
            }
          }
#pragma unroll
          for (int v = 0; v < V; ++v) {
            const unsigned w0 = q[0][v] | (q[1][v] << 16), w1 = q[2][v] | (q[3][v] << 16);
            const unsigned p0 = ((w0 >> (5 + st[v])) & 1u) ? (unsigned)st[v] : (w0 & 31u);
            const unsigned p1 = ((w1 >> (5 + st[v])) & 1u) ? (unsigned)st[v] : (w1 & 31u);
            pair[v] = p0 | (p1 << 8);
          }
        } else {
#pragma unroll
        for (int v = 0; v < V; ++v)
          pair[v] = __ldcs(reinterpret_cast<const unsigned short*>(arg + (((size_t)k * A + st[v]) * p.stride + v) * 2));
        }
      }
#pragma unroll
      for (int slot = 0; slot < 2; ++slot) {
        const int c = slot ? ch1 : ch0;
        int cs[V];
        if (c >= p.n_leaves) {
#pragma unroll
          for (int v = 0; v < V; ++v) cs[v] = (int)((pair[v] >> (8 * slot)) & 0xffu);
          st_u8<V>(anc + (size_t)(c - p.n_leaves) * p.stride, cs);
        } else if (c >= 0 && p.hist) {
          ld_u8<V>(leaf + (size_t)c * p.stride, cs);
        }
        if (p.hist && c >= 0) {
#pragma unroll
          for (int v = 0; v < V; ++v)
            if (site0 + v < p.n_sites && cs[v] < A) atomicAdd(&s_hist[st[v] * A + cs[v]], 1u);
        }
      }
    }
  }
  if (p.hist) {
    __syncthreads();
    for (int i = threadIdx.x; i < A * A; i += THREADS)
      if (s_hist[i]) atomicAdd(p.hist + (size_t)tree * A * A + i, (unsigned long long)s_hist[i]);
  }
}

// f64 -> f32 cast of small result vectors (totals, dC) for bindings that run without 64-bit types
__global__ void cast_f64_f32_kernel(const double* __restrict__ src, float* __restrict__ dst, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = (float)src[i];
}

// ---- exports -------------------------------------------------------------------------------
// dp_ref[l][node][s] from dp[k][s][l]: 32x32 shared-memory transpose, both sides coalesced.
__global__ void __launch_bounds__(256)
    export_dp_kernel(int n_all, int n_leaves, int A, long long n_sites, long long stride,
                     const uint8_t* __restrict__ leaf, const float* __restrict__ dp,
                     float* __restrict__ out) {
  __shared__ float tile[32][33];
  const long long J = (long long)n_all * A;
  const long long l0 = (long long)blockIdx.x * 32, j0 = (long long)blockIdx.y * 32;
  for (int r = threadIdx.y; r < 32; r += 8) {
    const long long j = j0 + r, l = l0 + threadIdx.x;
    float v = 0.f;
    if (j < J && l < n_sites) {
      const int node = (int)(j / A), s = (int)(j % A);
      if (node < n_leaves) v = (leaf[(size_t)node * stride + l] == s) ? 0.f : kSentinel;
      else v = dp[((size_t)(node - n_leaves) * A + s) * stride + l];
    }
    tile[r][threadIdx.x] = v;
  }
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += 8) {
    const long long l = l0 + r, j = j0 + threadIdx.x;
    if (l < n_sites && j < J) out[(size_t)l * J + j] = tile[threadIdx.x][r];
  }
}

__global__ void export_recon_kernel(int n_all, int n_leaves, long long n_sites, long long stride,
                                    const float* __restrict__ leaf_f32, long long leaf_in_stride,
                                    const uint8_t* __restrict__ leaf, const uint8_t* __restrict__ anc,
                                    float* __restrict__ out) {
  const long long l = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int node = blockIdx.y;
  if (l >= n_sites) return;
  float v = 0.f;
  if (node < n_leaves) {
    v = leaf_f32 ? leaf_f32[(size_t)node * leaf_in_stride + l] : (float)leaf[(size_t)node * stride + l];
  } else if (anc) {
    v = (float)anc[(size_t)(node - n_leaves) * stride + l];
  }
  out[(size_t)node * n_sites + l] = v;
}

// bt_ref[l][node][s][0..3] = (child0 id, argmin0, child1 id, argmin1) as f32 (sankoff.py:80-83)
__global__ void export_bt_kernel(const int32_t* __restrict__ plan, int n_all, int n_leaves, int A,
                                 long long n_sites, long long stride,
                                 const uint8_t* __restrict__ arg, float* __restrict__ out) {
  const long long l = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int j = blockIdx.y;  // node * A + s
  if (l >= n_sites) return;
  const int node = j / A, s = j % A;
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (node >= n_leaves) {
    const int n_int = n_all - n_leaves, k = node - n_leaves;
    PlanView pv(plan, n_int);
    v.x = (float)pv.raw(0)[k];
    v.z = (float)pv.raw(1)[k];
    v.y = (float)arg[(((size_t)k * A + s) * stride + l) * 2 + 0];
    v.w = (float)arg[(((size_t)k * A + s) * stride + l) * 2 + 1];
  }
  reinterpret_cast<float4*>(out)[(size_t)l * n_all * A + j] = v;
}

// ---- traceback over a reference-layout back-pointer table -----------------------------------
// backtrack_sankoff_jit (src/trex/sankoff.py:191-267) for callers that hold the reference's own
// (n_all, A, 4) f32 table per site.  Top-down over node ids (children have smaller ids than
// their parent), equivalent to the explicit stack DFS on trees.  recon: [n_sites][n_all] i32.
__global__ void backtrack_ref_kernel(const float* __restrict__ bt, int n_all, int n_leaves, int A,
                                     long long n_sites, const int* __restrict__ root_node,
                                     const int* __restrict__ root_state, int* __restrict__ recon) {
  const long long l = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (l >= n_sites) return;
  int* rec = recon + (size_t)l * n_all;
  const float* b = bt + (size_t)l * n_all * A * 4;
  for (int i = 0; i < n_all; ++i) rec[i] = -1;
  const int root = root_node[0];
  if (root >= n_leaves && root < n_all) rec[root] = min(max(root_state[l], 0), A - 1);
  for (int node = min(root, n_all - 1); node >= n_leaves; --node) {
    const int st = rec[node];
    if (st < 0) continue;
    const float4 info = *reinterpret_cast<const float4*>(b + ((size_t)node * A + st) * 4);
    const int c0 = (int)info.x, s0 = (int)info.y, c1 = (int)info.z, s1 = (int)info.w;
    // stack order of the reference: child 2 is popped (and written) before child 1
    if (c1 >= n_leaves && c1 < node) rec[c1] = min(max(s1, 0), A - 1);
    if (c0 >= n_leaves && c0 < node) rec[c0] = min(max(s0, 0), A - 1);
  }
  for (int i = 0; i < n_all; ++i)
    if (rec[i] < 0) rec[i] = 0;
}

}  // namespace trex

extern "C" {

int trex_encode_leaf_states(const float* seqs, int n_leaves, int64_t n_sites, int64_t in_stride,
                            int n_states, uint8_t* leaf_states, int64_t site_stride, void* stream) {
  using namespace trex;
  TREX_CHECK_ARG(seqs && leaf_states, "trex_encode_leaf_states: null pointer");
  TREX_CHECK_ARG(n_leaves >= 1 && n_leaves <= 65535 && n_sites >= 1 && in_stride >= n_sites &&
                     site_stride >= n_sites && site_stride % 16 == 0 && n_states >= 1 && n_states <= 32,
                 "trex_encode_leaf_states: bad sizes");
  dim3 grid((unsigned)((site_stride + 255) / 256), n_leaves);
  encode_leaf_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(seqs, n_leaves, n_sites, in_stride,
                                                              n_states, leaf_states, site_stride);
  note_launches(1);
  TREX_CHECK_LAUNCH("encode_leaf_kernel");
  return TREX_OK;
}

int trex_sankoff_traceback_cost(const int32_t* plan, int n_all, int n_leaves, int n_states,
                                int64_t n_sites, int64_t site_stride, int batch,
                                const uint8_t* leaf_states, const float* cost, int mode, const float* dp,
                                const uint8_t* argmin, uint8_t* anc_states, unsigned long long* edge_hist,
                                void* stream) {
  using namespace trex;
  TREX_CHECK_ARG(plan && leaf_states && dp && argmin && anc_states, "trex_sankoff_traceback: null pointer");
  const int want_compact = mode & TREX_MODE_COMPACT_ARG;
  mode &= ~TREX_MODE_COMPACT_ARG;
  TREX_CHECK_ARG(mode == TREX_MODE_AUTO || mode == TREX_MODE_GENERAL, "trex_sankoff_traceback_cost: bad mode");
  TREX_CHECK_ARG(!want_compact || cost, "trex_sankoff_traceback_cost: TREX_MODE_COMPACT_ARG needs the cost matrix");
  TREX_CHECK_ARG(n_leaves >= 1 && n_all > n_leaves && n_states >= 1 && n_states <= 32,
                 "trex_sankoff_traceback: bad tree sizes");
  TREX_CHECK_ARG(n_sites >= 1 && site_stride >= n_sites && site_stride % 16 == 0,
                 "trex_sankoff_traceback: site_stride must be a multiple of 16 and >= n_sites");
  TREX_CHECK_ARG(batch >= 1 && batch <= 65535, "trex_sankoff_traceback: batch out of range");
  cudaStream_t s = (cudaStream_t)stream;
  if (edge_hist) {
    int rc = check_cuda(cudaMemsetAsync(edge_hist, 0, sizeof(unsigned long long) * batch * n_states * n_states, s),
                        "memset edge_hist");
    if (rc) return rc;
  }
  TraceParams p{plan, trex_sankoff_plan_words(n_all, n_leaves), n_all, n_leaves, n_states,
                (long long)n_sites, (long long)site_stride, leaf_states, dp, argmin, anc_states, edge_hist};
  p.phase = 0;
  p.cost = cost;
  p.mode = mode;
  p.compact = want_compact && compact_arg_ok(n_states);
  constexpr int THREADS = 128;
  const size_t smem = edge_hist ? sizeof(unsigned int) * n_states * n_states : 0;
  auto run = [&](auto vc) {
    constexpr int V = decltype(vc)::value;
    const unsigned gx = (unsigned)(((n_sites + V - 1) / V + THREADS - 1) / THREADS);
    if (use_segments(n_leaves)) {
      p.phase = 2;  // top-down: the nodes above the segments first, then the independent subtrees
      traceback_kernel<V, THREADS><<<dim3(gx, batch), THREADS, smem, s>>>(p);
      p.phase = 1;
      traceback_kernel<V, THREADS><<<dim3(gx, batch, kMaxSeg), THREADS, smem, s>>>(p);
      note_launches(2);
    } else {
      traceback_kernel<V, THREADS><<<dim3(gx, batch), THREADS, smem, s>>>(p);
      note_launches(1);
    }
  };
  // Sites per lane.  The walk is instruction-bound (order / child look-ups and addresses per node and lane), so wide
  // lanes win as long as enough warps remain to hide the chain of dependent reads: measured on the 1024-leaf tree
  // (tools/probe_trace.py, pointer words, with histogram) 100k sites 0.60 / 0.45 / 0.36 ms at 1 / 2 / 4 sites per lane,
  // 12.5k sites 0.128 / 0.102 / 0.115, 1k sites 0.066 / 0.081 / 0.098.  TREX_TRACE_V overrides (probe).
  const long long work = (long long)n_sites * batch;
  int v = work >= 50000 ? 4 : work >= 5000 ? 2 : 1;
  if (const char* e = getenv("TREX_TRACE_V")) v = atoi(e);
  if (v == 4) run(std::integral_constant<int, 4>{});
  else if (v == 2) run(std::integral_constant<int, 2>{});
  else run(std::integral_constant<int, 1>{});
  TREX_CHECK_LAUNCH("traceback_kernel");
  return TREX_OK;
}

int trex_sankoff_traceback(const int32_t* plan, int n_all, int n_leaves, int n_states,
                           int64_t n_sites, int64_t site_stride, int batch,
                           const uint8_t* leaf_states, const float* dp, const uint8_t* argmin,
                           uint8_t* anc_states, unsigned long long* edge_hist, void* stream) {
  return trex_sankoff_traceback_cost(plan, n_all, n_leaves, n_states, n_sites, site_stride, batch, leaf_states, nullptr,
                                     TREX_MODE_AUTO, dp, argmin, anc_states, edge_hist, stream);
}

int trex_sankoff_export_dp(int n_all, int n_leaves, int n_states, int64_t n_sites,
                           int64_t site_stride, const uint8_t* leaf_states, const float* dp,
                           float* dp_ref, void* stream) {
  using namespace trex;
  TREX_CHECK_ARG(leaf_states && dp && dp_ref, "trex_sankoff_export_dp: null pointer");
  const long long J = (long long)n_all * n_states;
  TREX_CHECK_ARG((J + 31) / 32 <= 65535, "trex_sankoff_export_dp: n_all * n_states too large");
  dim3 grid((unsigned)((n_sites + 31) / 32), (unsigned)((J + 31) / 32));
  export_dp_kernel<<<grid, dim3(32, 8), 0, (cudaStream_t)stream>>>(n_all, n_leaves, n_states, n_sites,
                                                                  site_stride, leaf_states, dp, dp_ref);
  note_launches(1);
  TREX_CHECK_LAUNCH("export_dp_kernel");
  return TREX_OK;
}

int trex_sankoff_export_recon(int n_all, int n_leaves, int64_t n_sites, int64_t site_stride,
                              const float* leaf_seqs_f32, int64_t leaf_in_stride,
                              const uint8_t* leaf_states, const uint8_t* anc_states,
                              float* recon_ref, void* stream) {
  using namespace trex;
  TREX_CHECK_ARG((leaf_seqs_f32 || leaf_states) && recon_ref, "trex_sankoff_export_recon: null pointer");
  TREX_CHECK_ARG(n_all <= 65535, "trex_sankoff_export_recon: n_all too large");
  dim3 grid((unsigned)((n_sites + 255) / 256), n_all);
  export_recon_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(n_all, n_leaves, n_sites, site_stride,
                                                               leaf_seqs_f32, leaf_in_stride, leaf_states,
                                                               anc_states, recon_ref);
  note_launches(1);
  TREX_CHECK_LAUNCH("export_recon_kernel");
  return TREX_OK;
}

int trex_sankoff_export_bt(const int32_t* plan, int n_all, int n_leaves, int n_states,
                           int64_t n_sites, int64_t site_stride, const uint8_t* argmin,
                           float* bt_ref, void* stream) {
  using namespace trex;
  TREX_CHECK_ARG(plan && argmin && bt_ref, "trex_sankoff_export_bt: null pointer");
  TREX_CHECK_ARG((long long)n_all * n_states <= 65535, "trex_sankoff_export_bt: n_all * n_states too large");
  dim3 grid((unsigned)((n_sites + 255) / 256), n_all * n_states);
  export_bt_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(plan, n_all, n_leaves, n_states, n_sites,
                                                            site_stride, argmin, bt_ref);
  note_launches(1);
  TREX_CHECK_LAUNCH("export_bt_kernel");
  return TREX_OK;
}

int trex_sankoff_backtrack_ref(const float* bt_ref, int n_all, int n_leaves, int n_states,
                               int64_t n_sites, const int32_t* root_node, const int32_t* root_state,
                               int32_t* recon, void* stream) {
  using namespace trex;
  TREX_CHECK_ARG(bt_ref && root_node && root_state && recon, "trex_sankoff_backtrack_ref: null pointer");
  TREX_CHECK_ARG(n_leaves >= 1 && n_all >= n_leaves && n_states >= 1 && n_sites >= 1,
                 "trex_sankoff_backtrack_ref: bad sizes");
  backtrack_ref_kernel<<<(unsigned)((n_sites + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
      bt_ref, n_all, n_leaves, n_states, n_sites, root_node, root_state, recon);
  note_launches(1);
  TREX_CHECK_LAUNCH("backtrack_ref_kernel");
  return TREX_OK;
}

int trex_cast_f64_to_f32(const double* src, float* dst, int64_t n, void* stream) {
  using namespace trex;
  TREX_CHECK_ARG(src && dst && n >= 0, "trex_cast_f64_to_f32: bad arguments");
  if (n == 0) return TREX_OK;
  cast_f64_f32_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(src, dst, (long long)n);
  note_launches(1);
  TREX_CHECK_LAUNCH("cast_f64_f32_kernel");
  return TREX_OK;
}

}  // extern "C"
