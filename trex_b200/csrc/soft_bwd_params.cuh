// Launch parameters shared by the two relaxed-Sankoff reverse kernels.
#pragma once
#include "common.cuh"

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
  double* dcost;
  int only_direct;  // 1: return immediately when the exp-domain (warp-autonomous) kernel applies
};

// exp-domain evaluation is valid while exp(-C/T) cannot underflow a whole row sum
__device__ __forceinline__ bool expdom_ok(float c, float T) { return fabsf(c) <= 80.f * T; }

// sankoff_soft_bwd_warp.cu: warp-autonomous exp-domain kernel (returns at once when !expdom)
int launch_soft_bwd_warp(const SoftBwdParams& p, int batch, cudaStream_t s);

}  // namespace trex
