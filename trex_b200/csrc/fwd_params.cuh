// Launch parameters of the Sankoff forward kernels and the "uniform cost" predicate shared by the
// general kernels (sankoff_hard.cu, sankoff_soft_fwd.cu) and the O(A) kernels (sankoff_uniform.cu).
#pragma once
#include "common.cuh"

namespace trex {

struct HardFwdParams {
  const int32_t* plan;
  size_t plan_words;
  int n_all, n_leaves, n_states;
  long long n_sites, stride;
  const uint8_t* leaf;
  const float* cost;
  float* dp;
  uint8_t* arg;
  float* root_min;
  double* total;
  int mode;
  int phase;  // node list of this launch (common.cuh: Walk): 0 whole tree, 1 segment blockIdx.z, 2 top part
  int compact;  // TREX_MODE_COMPACT_ARG asked for and compact_arg_ok(n_states): the uniform kernel writes pointer words
};

// Pointer words (sankoff_uniform.cu) need 5 + n_states bits of a 32-bit word and, living at the start of the byte
// table's buffer, 8 bytes per (node, site) of its 2 * n_states: 4 <= n_states <= 20 (the 20-state tile is the widest
// one below 27).
__host__ __device__ inline bool compact_arg_ok(int n_states) { return n_states >= 4 && n_states <= 20; }

struct SoftFwdParams {
  const int32_t* plan;
  size_t plan_words;
  int n_all, n_leaves, n_states;
  long long n_sites, stride;
  const uint8_t* leaf;
  const float* leaf_cost;
  const float* cost;
  float T;
  float* dp;
  double* total;
  int phase;  // as HardFwdParams::phase
};

// Uniform substitution cost: C[s][s] == 0 and C[s][t] == a >= 0 for every s != t (Fitch-style
// parsimony, e.g. the unit cost 1 - I).  Then min_t / softmin_t over C[s,t] + d[t] needs only d[s]
// and one reduction over t:  O(A) per parent state row instead of O(A^2).
// Call with all threads of the CTA; `cost` is the device A x A matrix.  Returns the CTA-wide verdict
// and the off-diagonal value.
__device__ __forceinline__ bool cost_is_uniform(const float* __restrict__ cost, int A, float& a_out) {
  const float a = A > 1 ? __ldg(cost + 1) : 0.f;
  bool ok = a >= 0.f && a < 65536.f;
  for (int i = threadIdx.x; i < A * A; i += blockDim.x) {
    const float c = __ldg(cost + i);
    ok = ok && (c == ((i / A == i % A) ? 0.f : a));
  }
  a_out = a;
  return __syncthreads_and(ok);
}

// packed-key / exact-integer condition of the hard kernels: C integral, >= 0, and small enough that
// every partial score stays far below 2^24
__device__ __forceinline__ bool fast_cost_ok(float c, int n_all) {
  return (c >= 0.f) && (c == truncf(c)) && (c * (float)n_all < 262000.f);
}

// true (CTA-wide) when the O(A) kernels of sankoff_uniform.cu own this call
__device__ __forceinline__ bool hard_uniform_applies(const float* __restrict__ cost, int n_states, int mode, int n_all,
                                                     float& a) {
  const bool uni = cost_is_uniform(cost, n_states, a);
  return uni && mode == TREX_MODE_AUTO && fast_cost_ok(a, n_all);
}
__device__ __forceinline__ bool hard_uniform_applies(const HardFwdParams& p, float& a) {
  return hard_uniform_applies(p.cost, p.n_states, p.mode, p.n_all, a);
}
__device__ __forceinline__ bool soft_uniform_applies(const SoftFwdParams& p, float& a) {
  const bool uni = cost_is_uniform(p.cost, p.n_states, a);
  return uni && a <= 80.f * p.T;  // exp-domain evaluation valid (no underflow of exp(-a/T))
}

// sankoff_uniform.cu
int launch_hard_fwd_uniform(const HardFwdParams& p, int batch, cudaStream_t s);
int launch_soft_fwd_uniform(const SoftFwdParams& p, int batch, cudaStream_t s);

}  // namespace trex
