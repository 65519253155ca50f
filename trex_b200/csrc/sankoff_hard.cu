// Hard (min-plus) Sankoff forward over the whole tree — one thread owns V consecutive sites
// and walks the DFS post-order plan; sites are independent (src/trex/sankoff.py:97), so there
// is no inter-thread communication at all.
//
// Replaces run_dp / vectorized_dp and the table + score part of run_sankoff
// (src/trex/sankoff.py:24-97, 141-160, 187).
//
// HBM layout: dp[tree][k][state][site] f32, argmin[tree][k][state][site][slot] u8, leaf states
// [leaf][site] u8.  Leaf dp rows (1e5 with a 0 at the observed state, :49-52,152) are never
// materialised: min_t(C[s,t] + leafrow[t]) is a (A+1) x A lookup table in shared memory built
// with exactly the reference's f32 arithmetic (row A = "no state": all-1e5).
//
// Two inner loops, bit-identical wherever the second is enabled:
//  * exact-f32: v = C[s,t] + d[t]; strict '<' keeps the first minimum (jnp.argmin, :69).
//  * packed key (integer-valued 0 <= C, max(C)*n_all < 2^18): key = 32*(C[s,t]+d[t]) + t - 15.5
//    is exact in f32, so one fmin chain (FMNMX3 on sm_100a) yields min AND first argmin.
#include "common.cuh"
#include "fwd_params.cuh"

namespace trex {


template <int AP>
struct HardSmem {
  __align__(16) float C[AP * AP];            // [s][t], t >= A padded with 0
  __align__(16) float leaf_cost[(AP + 1) * AP];  // [o][s]
  __align__(4) uint8_t leaf_arg[(AP + 1) * AP];
};

// min and first-argmin over t of crow[t] + d[t][v]
template <int AP, int V, bool FAST>
__device__ __forceinline__ void minplus(const float (&crow)[AP], const float (&d)[AP][V],
                                        float (&best)[V], int (&idx)[V]) {
  if constexpr (FAST) {
    // d holds keys 32*d + (t - 15.5); crow holds C (multiplied by 32 in the FFMA)
#pragma unroll
    for (int v = 0; v < V; ++v) {
      float b = fmaf(crow[0], 32.f, d[0][v]);
#pragma unroll
      for (int t = 1; t < AP; ++t) b = fminf(b, fmaf(crow[t], 32.f, d[t][v]));
      const float val = fmaf(b, 0.03125f, 12582912.f) - 12582912.f;  // round(b/32): exact value
      const float w = fmaf(val, -32.f, b) + 4194319.5f;              // 2^22 + t, ulp 0.5
      best[v] = val;
      idx[v] = (__float_as_int(w) >> 1) & 31;
    }
  } else {
#pragma unroll
    for (int v = 0; v < V; ++v) {
      float b = crow[0] + d[0][v];
      int bi = 0;
#pragma unroll
      for (int t = 1; t < AP; ++t) {
        const float x = crow[t] + d[t][v];
        if (x < b) { b = x; bi = t; }
      }
      best[v] = b;
      idx[v] = bi;
    }
  }
}

// FA: n_states == AP (no padded states: 4, 8, ..., 20, ...) — the state count becomes a compile-time constant, so the row
// loads lose their predicates and INFINITY defaults (170 of the 440 instructions of a node's load phase) and the
// state loop its remainder handling.
template <int AP, int V, bool L0, bool L1, bool FAST, bool FA>
__device__ __forceinline__ void hard_node(const HardSmem<AP>& sm, int A_rt, long long stride,
                                          const float* __restrict__ dp, const uint8_t* __restrict__ leaf,
                                          int n_leaves, int c0, int c1, float* __restrict__ out,
                                          uint8_t* __restrict__ arg_out, bool is_root,
                                          float (&rmin)[V], const int (&o0)[V], const int (&o1)[V]) {
  const int A = FA ? AP : A_rt;
  float d0[L0 ? 1 : AP][V], d1[L1 ? 1 : AP][V];
  const unsigned row_b = (unsigned)stride * 4u;  // bytes between two state rows of dp
  if constexpr (!L0) {
    const float* src = dp + (size_t)(c0 - n_leaves) * A * stride;
#pragma unroll
    for (int t = 0; t < AP; ++t) {
      if (t < A) ld_f<V>(row_ptr(src, row_b, t), d0[t]);
      else {
#pragma unroll
        for (int v = 0; v < V; ++v) d0[t][v] = INFINITY;
      }
    }
  }
  if constexpr (!L1) {
    const float* src = dp + (size_t)(c1 - n_leaves) * A * stride;
#pragma unroll
    for (int t = 0; t < AP; ++t) {
      if (t < A) ld_f<V>(row_ptr(src, row_b, t), d1[t]);
      else {
#pragma unroll
        for (int v = 0; v < V; ++v) d1[t][v] = INFINITY;
      }
    }
  }
  if constexpr (FAST) {
#pragma unroll
    for (int t = 0; t < AP; ++t)
#pragma unroll
      for (int v = 0; v < V; ++v) {
        if constexpr (!L0) d0[t][v] = fmaf(d0[t][v], 32.f, (float)t - 15.5f);
        if constexpr (!L1) d1[t][v] = fmaf(d1[t][v], 32.f, (float)t - 15.5f);
      }
  }
#pragma unroll 2
  for (int s = 0; s < A; ++s) {
    float crow[AP];
    if constexpr (!L0 || !L1) {
      if constexpr (AP % 4 == 0) {
#pragma unroll
        for (int t = 0; t < AP; t += 4) {
          const float4 c4 = *reinterpret_cast<const float4*>(&sm.C[s * AP + t]);
          crow[t] = c4.x; crow[t + 1] = c4.y; crow[t + 2] = c4.z; crow[t + 3] = c4.w;
        }
      } else {
#pragma unroll
        for (int t = 0; t < AP; ++t) crow[t] = sm.C[s * AP + t];
      }
    }
    float b0[V], b1[V], o[V];
    int i0[V], i1[V];
    if constexpr (L0) {
#pragma unroll
      for (int v = 0; v < V; ++v) {
        b0[v] = sm.leaf_cost[o0[v] * AP + s];
        i0[v] = sm.leaf_arg[o0[v] * AP + s];
      }
    } else {
      minplus<AP, V, FAST>(crow, d0, b0, i0);
    }
    if constexpr (L1) {
#pragma unroll
      for (int v = 0; v < V; ++v) {
        b1[v] = sm.leaf_cost[o1[v] * AP + s];
        i1[v] = sm.leaf_arg[o1[v] * AP + s];
      }
    } else {
      minplus<AP, V, FAST>(crow, d1, b1, i1);
    }
#pragma unroll
    for (int v = 0; v < V; ++v) o[v] = b0[v] + b1[v];  // (0 + cost0) + cost1, sankoff.py:70
    st_f<V>(row_ptr(out, row_b, s), o);
    if (arg_out) st_arg_pair<V>(row_ptr(arg_out, row_b >> 1, s), i0, i1, true);
    if (is_root) {
#pragma unroll
      for (int v = 0; v < V; ++v) rmin[v] = fminf(rmin[v], o[v]);
    }
  }
}

// A node whose two children are leaves ("cherry", half of the internal nodes of a balanced tree): both rows come from the
// leaf lookup table, whole table rows at a time (LDS.128 for the scores, one word per 4 back-pointers) instead of one
// LDS + one LDS.U8 with its own address per (child, site, state) — 850 -> ~250 instructions per node and lane.
template <int AP, int V, bool FA>
__device__ __forceinline__ void hard_cherry(const HardSmem<AP>& sm, int A_rt, unsigned row_b, float* __restrict__ out,
                                            uint8_t* __restrict__ arg_out, const int (&o0)[V], const int (&o1)[V]) {
  static_assert(AP % 4 == 0, "hard_cherry: whole float4 table rows");
  const int A = FA ? AP : A_rt;
  float c0[V][AP], c1[V][AP];
  unsigned a0[V][AP / 4], a1[V][AP / 4];
#pragma unroll
  for (int v = 0; v < V; ++v) {
#pragma unroll
    for (int j = 0; j < AP / 4; ++j) {
      const float4 x = *reinterpret_cast<const float4*>(&sm.leaf_cost[o0[v] * AP + 4 * j]);
      const float4 y = *reinterpret_cast<const float4*>(&sm.leaf_cost[o1[v] * AP + 4 * j]);
      c0[v][4 * j] = x.x; c0[v][4 * j + 1] = x.y; c0[v][4 * j + 2] = x.z; c0[v][4 * j + 3] = x.w;
      c1[v][4 * j] = y.x; c1[v][4 * j + 1] = y.y; c1[v][4 * j + 2] = y.z; c1[v][4 * j + 3] = y.w;
      a0[v][j] = *reinterpret_cast<const unsigned*>(&sm.leaf_arg[o0[v] * AP + 4 * j]);
      a1[v][j] = *reinterpret_cast<const unsigned*>(&sm.leaf_arg[o1[v] * AP + 4 * j]);
    }
  }
#pragma unroll
  for (int s = 0; s < AP; ++s) {
    if (s < A) {
      float o[V];
#pragma unroll
      for (int v = 0; v < V; ++v) o[v] = c0[v][s] + c1[v][s];  // (0 + cost0) + cost1, sankoff.py:70
      st_f<V>(row_ptr(out, row_b, s), o);
      if (arg_out) {
        if constexpr (V == 2) {  // bytes [slot0, slot1] of site 0, then of site 1: three PRMTs
          const unsigned sel = (s % 4) | ((4 + s % 4) << 4);
          const unsigned lo = __byte_perm(a0[0][s / 4], a1[0][s / 4], sel);
          const unsigned hi = __byte_perm(a0[1][s / 4], a1[1][s / 4], sel);
          __stcs(reinterpret_cast<unsigned*>(row_ptr(arg_out, row_b >> 1, s)), __byte_perm(lo, hi, 0x5410u));
        } else {
          int i0[V], i1[V];
#pragma unroll
          for (int v = 0; v < V; ++v) {
            i0[v] = (a0[v][s / 4] >> (8 * (s % 4))) & 0xff;
            i1[v] = (a1[v][s / 4] >> (8 * (s % 4))) & 0xff;
          }
          st_arg_pair<V>(row_ptr(arg_out, row_b >> 1, s), i0, i1, true);
        }
      }
    }
  }
}

template <int AP, int V, int THREADS>
__global__ void __launch_bounds__(THREADS, 4) hard_fwd_kernel(const HardFwdParams p) {
  __shared__ HardSmem<AP> sm;
  const int A = p.n_states;
  const int tid = threadIdx.x;
  const int tree = blockIdx.y;
  const int n_int = p.n_all - p.n_leaves;
  PlanView plan(p.plan + (size_t)tree * p.plan_words, n_int);
  if (p.phase == 1 && (int)blockIdx.z >= plan.hdr()[2]) return;  // empty segment slice: skip the prologue (CTA-uniform)
  {
    float a_uni;
    if (hard_uniform_applies(p, a_uni)) return;  // hard_fwd_uniform_kernel (sankoff_uniform.cu) owns this call
  }

  // ---- prologue: cost matrix, leaf lookup table, fast-path eligibility --------------------
  bool c_ok = true;
  for (int i = tid; i < AP * AP; i += THREADS) {
    const int s = i / AP, t = i % AP;
    float c = 0.f;
    if (s < A && t < A) {
      c = __ldg(p.cost + s * A + t);
      c_ok = c_ok && fast_cost_ok(c, p.n_all);
    }
    sm.C[i] = c;
  }
  const int fast_ok = __syncthreads_and(c_ok) && p.mode == TREX_MODE_AUTO &&
                      !(plan.hdr()[3] & TREX_PLAN_HAS_SENTINEL_CHILD);
  for (int i = tid; i < (A + 1) * A; i += THREADS) {
    const int o = i / A, s = i % A;
    float b = sm.C[s * AP + 0] + (o == 0 ? 0.f : kSentinel);  // sankoff.py:67 on a leaf row
    int bi = 0;
    for (int t = 1; t < A; ++t) {
      const float x = sm.C[s * AP + t] + (o == t ? 0.f : kSentinel);
      if (x < b) { b = x; bi = t; }
    }
    sm.leaf_cost[o * AP + s] = b;
    sm.leaf_arg[o * AP + s] = (uint8_t)bi;
  }
  __syncthreads();

  const long long site0 = ((long long)blockIdx.x * THREADS + tid) * V;
  const bool active = site0 < p.n_sites;
  float rmin[V];
#pragma unroll
  for (int v = 0; v < V; ++v) rmin[v] = INFINITY;

  // The top part consumes rows the subtree segments wrote (another launch): a stateless leaf anywhere below disables
  // the packed-key path exactly as in the single walk.  One look at all leaves of this CTA's sites, 16 bytes per load,
  // the whole CTA sharing the verdict (coarser than the per-site flag of the single walk; both paths give the same
  // rows while the keys are exact, so only speed depends on it).
  bool cta_bad = false;
  if (p.phase == 2 && fast_ok) {
    constexpr int kChunks = THREADS * V / 16;  // 16-byte chunks per leaf row of this CTA's site range
    static_assert(THREADS % kChunks == 0, "leaf scan: threads per row");
    const long long off = (long long)blockIdx.x * THREADS * V + (tid % kChunks) * 16;
    unsigned any = 0;
    if (off < p.n_sites) {
      const unsigned lim = (unsigned)A * 0x01010101u;
      if ((reinterpret_cast<uintptr_t>(p.leaf) & 15) == 0) {
        const uint8_t* q = p.leaf + off;
#pragma unroll 4
        for (int c = tid / kChunks; c < p.n_leaves; c += THREADS / kChunks) {
          const uint4 w = __ldg(reinterpret_cast<const uint4*>(row_ptr(q, (unsigned)p.stride, c)));
          any |= __vcmpgeu4(w.x, lim) | __vcmpgeu4(w.y, lim) | __vcmpgeu4(w.z, lim) | __vcmpgeu4(w.w, lim);
        }
      } else {  // a leaf table that is not 16-byte aligned: bytes
        for (int c = tid / kChunks; c < p.n_leaves; c += THREADS / kChunks)
          for (int j = 0; j < 16; ++j) any |= p.leaf[(size_t)c * p.stride + off + j] >= A;
      }
    }
    cta_bad = __syncthreads_or(any != 0);
  }

  if (active) {
    float* dp = p.dp + (size_t)tree * n_int * A * p.stride + site0;
    uint8_t* arg = p.arg ? p.arg + ((size_t)tree * n_int * A * p.stride + site0) * 2 : nullptr;
    const uint8_t* leaf = p.leaf + site0;
    const Walk wk = make_walk(plan, p.phase, blockIdx.z);
    const int32_t* order = wk.ord;
    const int32_t* ch0 = plan.child(0);
    const int32_t* ch1 = plan.child(1);
    bool bad = cta_bad;  // an out-of-range leaf state puts ~1e5 values into internal rows
    for (int i = 0; i < wk.n_fwd; ++i) {
      const int node = __ldg(order + i);
      const int k = node - p.n_leaves;
      const int c0 = __ldg(ch0 + k), c1 = __ldg(ch1 + k);
      const bool l0 = c0 < p.n_leaves, l1 = c1 < p.n_leaves;
      int o0[V], o1[V];
#pragma unroll
      for (int v = 0; v < V; ++v) o0[v] = o1[v] = A;
      if (l0 && c0 >= 0) {
        ld_u8<V>(leaf + (size_t)c0 * p.stride, o0);
#pragma unroll
        for (int v = 0; v < V; ++v) { if (o0[v] >= A) { o0[v] = A; bad = true; } }
      }
      if (l1 && c1 >= 0) {
        ld_u8<V>(leaf + (size_t)c1 * p.stride, o1);
#pragma unroll
        for (int v = 0; v < V; ++v) { if (o1[v] >= A) { o1[v] = A; bad = true; } }
      }
      float* out = dp + (size_t)k * A * p.stride;
      uint8_t* aout = arg ? arg + (size_t)k * 2 * A * p.stride : nullptr;
      const bool is_root = node == p.n_all - 1;
#define TREX_NODE(L0, L1, F)                                                                                          \
  do {                                                                                                                \
    if (A == AP)                                                                                                      \
      hard_node<AP, V, L0, L1, F, true>(sm, A, p.stride, dp, leaf, p.n_leaves, c0, c1, out, aout, is_root, rmin, o0, o1);  \
    else                                                                                                              \
      hard_node<AP, V, L0, L1, F, false>(sm, A, p.stride, dp, leaf, p.n_leaves, c0, c1, out, aout, is_root, rmin, o0, o1); \
  } while (0)
      if (l0 && l1) {
        if constexpr (AP % 4 == 0) {
          if (is_root) TREX_NODE(true, true, false);
          else if (A == AP) hard_cherry<AP, V, true>(sm, A, (unsigned)p.stride * 4u, out, aout, o0, o1);
          else hard_cherry<AP, V, false>(sm, A, (unsigned)p.stride * 4u, out, aout, o0, o1);
        } else {
          TREX_NODE(true, true, false);
        }
      } else if (fast_ok && !bad) {
        if (l0) TREX_NODE(true, false, true);
        else if (l1) TREX_NODE(false, true, true);
        else TREX_NODE(false, false, true);
      } else {
        if (l0) TREX_NODE(true, false, false);
        else if (l1) TREX_NODE(false, true, false);
        else TREX_NODE(false, false, false);
      }
#undef TREX_NODE
    }
    if (p.root_min && p.phase != 1) st_f<V>(p.root_min + (size_t)tree * p.stride + site0, rmin);
  }
  if (p.total && p.phase != 1) {
    double acc = 0.0;
    if (active) {
#pragma unroll
      for (int v = 0; v < V; ++v)
        if (site0 + v < p.n_sites) acc += (double)rmin[v];
    }
    block_atomic_add(acc, p.total + tree);
  }
}

template <int AP, int V>
static int launch_hard(const HardFwdParams& p, int batch, cudaStream_t s) {
  constexpr int THREADS = 128;
  const long long nvec = (p.n_sites + V - 1) / V;
  const unsigned gx = (unsigned)((nvec + THREADS - 1) / THREADS);
  if (use_segments(p.n_leaves)) {
    HardFwdParams q = p;
    q.phase = 1;  // independent subtrees side by side ...
    hard_fwd_kernel<AP, V, THREADS><<<dim3(gx, batch, kMaxSeg), THREADS, 0, s>>>(q);
    q.phase = 2;  // ... then the nodes above them
    hard_fwd_kernel<AP, V, THREADS><<<dim3(gx, batch), THREADS, 0, s>>>(q);
    note_launches(2);
  } else {
    hard_fwd_kernel<AP, V, THREADS><<<dim3(gx, batch), THREADS, 0, s>>>(p);
    note_launches(1);
  }
  return check_cuda(cudaPeekAtLastError(), "hard_fwd_kernel");
}

}  // namespace trex

extern "C" int trex_sankoff_hard_fwd(const int32_t* plan, int n_all, int n_leaves, int n_states,
                                     int64_t n_sites, int64_t site_stride, int batch,
                                     const uint8_t* leaf_states, const float* cost, float* dp,
                                     uint8_t* argmin, float* root_min, double* total, int mode,
                                     void* stream) {
  using namespace trex;
  TREX_CHECK_ARG(plan && leaf_states && cost && dp, "trex_sankoff_hard_fwd: null pointer");
  TREX_CHECK_ARG(n_leaves >= 1 && n_all > n_leaves, "trex_sankoff_hard_fwd: need n_all > n_leaves >= 1");
  TREX_CHECK_ARG(n_states >= 1 && n_states <= 32, "trex_sankoff_hard_fwd: n_states %d not in 1..32", n_states);
  TREX_CHECK_ARG(n_sites >= 1 && site_stride >= n_sites && site_stride % 16 == 0 && site_stride < (1LL << 29),
                 "trex_sankoff_hard_fwd: site_stride must be a multiple of 16, >= n_sites and < 2^29");
  TREX_CHECK_ARG(batch >= 1 && batch <= 65535, "trex_sankoff_hard_fwd: batch out of range");
  const int want_compact = mode & TREX_MODE_COMPACT_ARG;
  mode &= ~TREX_MODE_COMPACT_ARG;
  TREX_CHECK_ARG(mode == TREX_MODE_AUTO || mode == TREX_MODE_GENERAL, "trex_sankoff_hard_fwd: bad mode");
  cudaStream_t s = (cudaStream_t)stream;
  if (total) {
    int rc = check_cuda(cudaMemsetAsync(total, 0, sizeof(double) * batch, s), "memset total");
    if (rc) return rc;
  }
  HardFwdParams p{plan, trex_sankoff_plan_words(n_all, n_leaves), n_all, n_leaves, n_states,
                  (long long)n_sites, (long long)site_stride, leaf_states, cost, dp, argmin,
                  root_min, total, mode};
  p.phase = 0;
  p.compact = want_compact && argmin && compact_arg_ok(n_states);
  const int A = n_states;
  {
    // device-side dispatch on the cost matrix: the O(A) kernel returns at once unless C is uniform
    // (and mode == AUTO); the general kernels below return at once when it is
    int rc = launch_hard_fwd_uniform(p, batch, s);
    if (rc) return rc;
  }
  if (A <= 2) return launch_hard<2, 4>(p, batch, s);
  if (A <= 4) return launch_hard<4, 4>(p, batch, s);
  if (A <= 8) return launch_hard<8, 2>(p, batch, s);
  if (A <= 20) {
    // few sites per GPU (site-sharded runs): one site per thread keeps more warps in flight
    if (n_sites * (long long)batch < 148LL * 512) return launch_hard<20, 1>(p, batch, s);
    return launch_hard<20, 2>(p, batch, s);
  }
  return launch_hard<32, 1>(p, batch, s);
}
