// Relaxed (softmin) Sankoff forward — SURVEY §8a row R1.  The reference has no relaxed code
// (SURVEY F3); the recurrence is src/trex/sankoff.py:62-79 and :187 with
//     min_t x   ->   softmin_T(x) = -T * log(sum_t exp(-x_t / T)).
//
// Same thread/site decomposition and HBM layout as the hard kernel.  The softmin of a child
// row d[.] against the cost matrix is evaluated in the exp domain (Felsenstein-pruning form):
//     softmin_t(C[s,t] + d[t]) = m - T*log( sum_t K[s,t] * e[t] ),
//     K = exp(-C/T) (shared memory),  m = min_t d[t],  e[t] = exp(-(d[t]-m)/T)
// i.e. A exp2 + A*A FFMA + A log2 per child instead of A*A exp2 (MUFU-bound).  Since
// sum_t K[s,t] e[t] >= exp(-max(C)/T), it cannot underflow while max|C|/T <= 80; otherwise a
// direct (A*A exp2) evaluation is used.  Leaf rows given as uint8 states are a lookup table.
#include "common.cuh"
#include "fwd_params.cuh"

namespace trex {


template <int AP>
struct SoftFwdSmem {
  __align__(16) float K[AP * AP];  // exp(-C/T), [s][t]
  __align__(16) float C[AP * AP];
  __align__(16) float leaf_soft[(AP + 1) * AP];  // [o][s]: softmin_t(C[s,t] + leafrow_o[t])
};

template <int AP, int V>
__device__ __forceinline__ void load_row(const float* __restrict__ src, int A, long long stride,
                                         float (&d)[AP][V]) {
#pragma unroll
  for (int t = 0; t < AP; ++t) {
    if (t < A) ld_f<V>(row_ptr(src, (unsigned)stride * 4u, t), d[t]);
    else {
#pragma unroll
      for (int v = 0; v < V; ++v) d[t][v] = INFINITY;
    }
  }
}

// in place: d[t] -> e[t] = exp(-(d[t]-m)/T); returns m
template <int AP, int V>
__device__ __forceinline__ void to_exp_domain(float (&d)[AP][V], float (&m)[V], float c1) {
#pragma unroll
  for (int v = 0; v < V; ++v) {
    float mm = d[0][v];
#pragma unroll
    for (int t = 1; t < AP; ++t) mm = fminf(mm, d[t][v]);
    m[v] = mm;
#pragma unroll
    for (int t = 0; t < AP; ++t) d[t][v] = ex2_ftz((mm - d[t][v]) * c1);
  }
}

// FA: n_states == AP, the state count is a compile-time constant (no load predicates, no loop remainder)
template <int AP, int V, bool R0, bool R1, bool EXPDOM, bool FA>
__device__ __forceinline__ void soft_node(const SoftFwdSmem<AP>& sm, int A_rt, long long stride,
                                          const float* __restrict__ src0, const float* __restrict__ src1,
                                          const int (&o0)[V], const int (&o1)[V], float c1, float tln2,
                                          float* __restrict__ out) {
  const int A = FA ? AP : A_rt;
  float d0[R0 ? AP : 1][V], d1[R1 ? AP : 1][V];
  float m0[V], m1[V];
  if constexpr (R0) load_row<AP, V>(src0, A, stride, d0);
  if constexpr (R1) load_row<AP, V>(src1, A, stride, d1);
  if constexpr (EXPDOM) {
    if constexpr (R0) to_exp_domain<AP, V>(d0, m0, c1);
    if constexpr (R1) to_exp_domain<AP, V>(d1, m1, c1);
  }
#pragma unroll 2
  for (int s = 0; s < A; ++s) {
    float row[AP];
    if constexpr (R0 || R1) {
      const float* base = EXPDOM ? sm.K : sm.C;
      if constexpr (AP % 4 == 0) {
#pragma unroll
        for (int t = 0; t < AP; t += 4) {
          const float4 c4 = *reinterpret_cast<const float4*>(&base[s * AP + t]);
          row[t] = c4.x; row[t + 1] = c4.y; row[t + 2] = c4.z; row[t + 3] = c4.w;
        }
      } else {
#pragma unroll
        for (int t = 0; t < AP; ++t) row[t] = base[s * AP + t];
      }
    }
    float o[V];
#pragma unroll
    for (int v = 0; v < V; ++v) {
      float a, b;
      if constexpr (!R0) {
        a = sm.leaf_soft[o0[v] * AP + s];
      } else if constexpr (EXPDOM) {
        float z = 0.f;
#pragma unroll
        for (int t = 0; t < AP; ++t) z = fmaf(row[t], d0[t][v], z);
        a = fmaf(-tln2, lg2_ftz(z), m0[v]);  // z >= exp(-80): the term of the smallest row entry (c_ok), not subnormal
      } else {
        float mm = INFINITY;
#pragma unroll
        for (int t = 0; t < AP; ++t) mm = fminf(mm, row[t] + d0[t][v]);
        float z = 0.f;
#pragma unroll
        for (int t = 0; t < AP; ++t) z += ex2_ftz((mm - (row[t] + d0[t][v])) * c1);
        a = fmaf(-tln2, lg2_ftz(z), mm);  // z >= 1: the minimum's own term
      }
      if constexpr (!R1) {
        b = sm.leaf_soft[o1[v] * AP + s];
      } else if constexpr (EXPDOM) {
        float z = 0.f;
#pragma unroll
        for (int t = 0; t < AP; ++t) z = fmaf(row[t], d1[t][v], z);
        b = fmaf(-tln2, lg2_ftz(z), m1[v]);
      } else {
        float mm = INFINITY;
#pragma unroll
        for (int t = 0; t < AP; ++t) mm = fminf(mm, row[t] + d1[t][v]);
        float z = 0.f;
#pragma unroll
        for (int t = 0; t < AP; ++t) z += ex2_ftz((mm - (row[t] + d1[t][v])) * c1);
        b = fmaf(-tln2, lg2_ftz(z), mm);
      }
      o[v] = a + b;
    }
    st_f<V>(row_ptr(out, (unsigned)stride * 4u, s), o);
  }
}

template <int AP, int V, int THREADS>
__global__ void __launch_bounds__(THREADS, 4) soft_fwd_kernel(const SoftFwdParams p) {
  __shared__ SoftFwdSmem<AP> sm;
  const int A = p.n_states;
  const int tid = threadIdx.x;
  const int tree = blockIdx.y;
  const int n_int = p.n_all - p.n_leaves;
  PlanView plan(p.plan + (size_t)tree * p.plan_words, n_int);
  const float c1 = 1.4426950408889634f / p.T;   // log2(e)/T
  const float tln2 = p.T * 0.6931471805599453f;  // T*ln(2)
  if (p.phase == 1 && (int)blockIdx.z >= plan.hdr()[2]) return;  // empty segment slice: skip the prologue (CTA-uniform)
  {
    float a_uni;
    if (soft_uniform_applies(p, a_uni)) return;  // soft_fwd_uniform_kernel (sankoff_uniform.cu) owns this call
  }

  bool c_ok = true;
  for (int i = tid; i < AP * AP; i += THREADS) {
    const int s = i / AP, t = i % AP;
    float c = 0.f, k = 0.f;
    if (s < A && t < A) {
      c = __ldg(p.cost + s * A + t);
      c_ok = c_ok && (fabsf(c) <= 80.f * p.T);  // false for NaN/inf as well
      k = exp2f(-c * c1);
    } else if (s < A) {
      c = INFINITY;  // padded states never contribute (direct path); K stays 0
    }
    sm.C[i] = c;
    sm.K[i] = k;
  }
  const int expdom = __syncthreads_and(c_ok);
  for (int i = tid; i < (A + 1) * A; i += THREADS) {
    const int o = i / A, s = i % A;
    double mm = 1e300;
    for (int t = 0; t < A; ++t) mm = fmin(mm, (double)sm.C[s * AP + t] + (o == t ? 0.0 : (double)kSentinel));
    double z = 0.0;
    for (int t = 0; t < A; ++t)
      z += exp(-(((double)sm.C[s * AP + t] + (o == t ? 0.0 : (double)kSentinel)) - mm) / (double)p.T);
    sm.leaf_soft[o * AP + s] = (float)(mm - (double)p.T * log(z));
  }
  __syncthreads();

  const long long site0 = ((long long)blockIdx.x * THREADS + tid) * V;
  const bool active = site0 < p.n_sites;
  const Walk wk = make_walk(plan, p.phase, blockIdx.z);
  const bool has_root = p.phase != 1;  // the whole-tree walk and the top part end at the root
  double acc = 0.0;
  if (active) {
    float* dp = p.dp + (size_t)tree * n_int * A * p.stride + site0;
    const int32_t* order = wk.ord;
    const int32_t* ch0 = plan.child(0);
    const int32_t* ch1 = plan.child(1);
    for (int i = 0; i < wk.n_fwd; ++i) {
      const int node = __ldg(order + i);
      const int k = node - p.n_leaves;
      const int c0 = __ldg(ch0 + k), c1n = __ldg(ch1 + k);
      const float* src0 = nullptr;
      const float* src1 = nullptr;
      int o0[V], o1[V];
#pragma unroll
      for (int v = 0; v < V; ++v) o0[v] = o1[v] = A;
      if (c0 >= p.n_leaves) src0 = dp + (size_t)(c0 - p.n_leaves) * A * p.stride;
      else if (c0 >= 0) {
        if (p.leaf_cost) src0 = p.leaf_cost + (size_t)c0 * A * p.stride + site0;
        else {
          ld_u8<V>(p.leaf + (size_t)c0 * p.stride + site0, o0);
#pragma unroll
          for (int v = 0; v < V; ++v) o0[v] = min(o0[v], A);
        }
      }
      if (c1n >= p.n_leaves) src1 = dp + (size_t)(c1n - p.n_leaves) * A * p.stride;
      else if (c1n >= 0) {
        if (p.leaf_cost) src1 = p.leaf_cost + (size_t)c1n * A * p.stride + site0;
        else {
          ld_u8<V>(p.leaf + (size_t)c1n * p.stride + site0, o1);
#pragma unroll
          for (int v = 0; v < V; ++v) o1[v] = min(o1[v], A);
        }
      }
      float* out = dp + (size_t)k * A * p.stride;
#define TREX_SNODE(R0, R1, E)                                                                       \
  do {                                                                                              \
    if (A == AP) soft_node<AP, V, R0, R1, E, true>(sm, A, p.stride, src0, src1, o0, o1, c1, tln2, out);  \
    else soft_node<AP, V, R0, R1, E, false>(sm, A, p.stride, src0, src1, o0, o1, c1, tln2, out);         \
  } while (0)
      if (expdom) {
        if (src0 && src1) TREX_SNODE(true, true, true);
        else if (src0) TREX_SNODE(true, false, true);
        else if (src1) TREX_SNODE(false, true, true);
        else TREX_SNODE(false, false, true);
      } else {
        if (src0 && src1) TREX_SNODE(true, true, false);
        else if (src0) TREX_SNODE(true, false, false);
        else if (src1) TREX_SNODE(false, true, false);
        else TREX_SNODE(false, false, false);
      }
#undef TREX_SNODE
    }
    // total += softmin_s dp[root][s]   (sankoff.py:187 relaxed)
    if (has_root) {
    float x[AP][V];
    load_row<AP, V>(dp + (size_t)(n_int - 1) * A * p.stride, A, p.stride, x);
#pragma unroll
    for (int v = 0; v < V; ++v) {
      float mm = x[0][v];
#pragma unroll
      for (int s = 1; s < AP; ++s) mm = fminf(mm, x[s][v]);
      float z = 0.f;
#pragma unroll
      for (int s = 0; s < AP; ++s) z += ex2_ftz((mm - x[s][v]) * c1);
      if (site0 + v < p.n_sites) acc += (double)fmaf(-tln2, lg2_ftz(z), mm);
    }
    }
  }
  if (p.total && has_root) block_atomic_add(acc, p.total + tree);
}

template <int AP, int V>
static int launch_soft_fwd(const SoftFwdParams& p, int batch, cudaStream_t s) {
  constexpr int THREADS = 128;
  const long long nvec = (p.n_sites + V - 1) / V;
  const unsigned gx = (unsigned)((nvec + THREADS - 1) / THREADS);
  if (use_segments(p.n_leaves)) {
    SoftFwdParams q = p;
    q.phase = 1;  // independent subtrees side by side ...
    soft_fwd_kernel<AP, V, THREADS><<<dim3(gx, batch, kMaxSeg), THREADS, 0, s>>>(q);
    q.phase = 2;  // ... then the nodes above them
    soft_fwd_kernel<AP, V, THREADS><<<dim3(gx, batch), THREADS, 0, s>>>(q);
    note_launches(2);
  } else {
    soft_fwd_kernel<AP, V, THREADS><<<dim3(gx, batch), THREADS, 0, s>>>(p);
    note_launches(1);
  }
  return check_cuda(cudaPeekAtLastError(), "soft_fwd_kernel");
}

}  // namespace trex

extern "C" int trex_sankoff_soft_fwd(const int32_t* plan, int n_all, int n_leaves, int n_states,
                                     int64_t n_sites, int64_t site_stride, int batch,
                                     const uint8_t* leaf_states, const float* leaf_cost,
                                     const float* cost, float temperature, float* dp,
                                     double* total, void* stream) {
  using namespace trex;
  TREX_CHECK_ARG(plan && cost && dp, "trex_sankoff_soft_fwd: null pointer");
  TREX_CHECK_ARG((leaf_states != nullptr) != (leaf_cost != nullptr),
                 "trex_sankoff_soft_fwd: exactly one of leaf_states / leaf_cost must be given");
  TREX_CHECK_ARG(n_leaves >= 1 && n_all > n_leaves, "trex_sankoff_soft_fwd: need n_all > n_leaves >= 1");
  TREX_CHECK_ARG(n_states >= 1 && n_states <= 32, "trex_sankoff_soft_fwd: n_states %d not in 1..32", n_states);
  TREX_CHECK_ARG(n_sites >= 1 && site_stride >= n_sites && site_stride % 16 == 0 && site_stride < (1LL << 29),
                 "trex_sankoff_soft_fwd: site_stride must be a multiple of 16, >= n_sites and < 2^29");
  TREX_CHECK_ARG(batch >= 1 && batch <= 65535, "trex_sankoff_soft_fwd: batch out of range");
  TREX_CHECK_ARG(temperature > 0.f && temperature < 1e30f, "trex_sankoff_soft_fwd: temperature must be > 0");
  cudaStream_t s = (cudaStream_t)stream;
  if (total) {
    int rc = check_cuda(cudaMemsetAsync(total, 0, sizeof(double) * batch, s), "memset total");
    if (rc) return rc;
  }
  SoftFwdParams p{plan, trex_sankoff_plan_words(n_all, n_leaves), n_all, n_leaves, n_states,
                  (long long)n_sites, (long long)site_stride, leaf_states, leaf_cost, cost,
                  temperature, dp, total};
  const int A = n_states;
  {
    int rc = launch_soft_fwd_uniform(p, batch, s);  // returns at once on the device unless C is uniform
    if (rc) return rc;
  }
  if (A <= 2) return launch_soft_fwd<2, 4>(p, batch, s);
  if (A <= 4) return launch_soft_fwd<4, 4>(p, batch, s);
  if (A <= 8) return launch_soft_fwd<8, 2>(p, batch, s);
  if (A <= 20) {
    if (n_sites * (long long)batch < 148LL * 512) return launch_soft_fwd<20, 1>(p, batch, s);
    return launch_soft_fwd<20, 2>(p, batch, s);
  }
  return launch_soft_fwd<32, 1>(p, batch, s);
}
