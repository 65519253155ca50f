// Launch parameters shared by the two relaxed-Sankoff reverse kernels.
#pragma once
#include "common.cuh"
#include "fwd_params.cuh"

namespace trex {

struct SoftBwdParams {
  const int32_t* plan;
  size_t plan_words;
  int n_all, n_leaves, n_states;
  long long n_sites, stride;
  const uint8_t* leaf;
  const float* leaf_cost;
  const float* cost;
  float T;
  const float* dp;
  const float* gscale;
  float* gbar;
  float* dleaf;
  float* dleaf_obs;  // compact leaf gradient [tree][leaf][site] (leaves given as states) or nullptr
  double* dcost;
  int only_direct;  // 1: return immediately when the exp-domain (warp-autonomous) kernel applies
  int phase;        // node list of this launch (common.cuh: Walk): 0 whole tree, 1 segment blockIdx.z, 2 top part
};

// exp-domain evaluation is valid while exp(-C/T) cannot underflow a whole row sum
__device__ __forceinline__ bool expdom_ok(float c, float T) { return fabsf(c) <= 80.f * T; }

// true (CTA-wide) when the O(A) kernel of sankoff_soft_bwd_uniform.cu owns this call: uniform cost
// (fwd_params.cuh), exp-domain evaluation valid, alphabet within the warp kernels' range
__device__ __forceinline__ bool soft_bwd_uniform_applies(const SoftBwdParams& p, float& a) {
  const bool uni = cost_is_uniform(p.cost, p.n_states, a);
  return uni && p.n_states <= 20 && a <= 80.f * p.T;
}

// sankoff_soft_bwd_uniform.cu: O(A) per edge for a uniform cost (returns at once otherwise)
int launch_soft_bwd_uniform(const SoftBwdParams& p, int batch, cudaStream_t s);

// sankoff_soft_bwd_warp.cu: warp-autonomous exp-domain kernel (returns at once when !expdom)
int launch_soft_bwd_warp(const SoftBwdParams& p, int batch, cudaStream_t s);

}  // namespace trex
