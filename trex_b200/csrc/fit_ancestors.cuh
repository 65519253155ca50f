// Shared by fit_ancestors.cu (generic kernel, setup, C entry point) and fit_ancestors_fast.cu
// (vectorised Q = 4 / 20 kernel with the sparse-adjacency product).
#pragma once
#include "common.cuh"

namespace trex {
namespace fit {

constexpr int kThreads = 256;
constexpr int kMaxAnc = 128;   // up to here A2[anc, anc] is staged in shared memory (generic kernel: hard limit)
constexpr int kMaxAncFast = 4096;  // fast kernel beyond kMaxAnc: ELL from shared memory, or the dense A2 read from L2
constexpr int kTileCols = 64;  // columns (sites x states) of one product tile of the generic kernel
constexpr int kEll = 8;        // a row of A2[anc, anc] with at most this many non-zeros is "sparse" (a tree has <= 4)
constexpr int kFastThreads = 512;
constexpr int kSlotDoubles = 16;  // workspace doubles per CTA and parity buffer: one 128-byte line per slot, so the polling
                                 // loads of a barrier spread over the L2 slices (the kernels use 2: cost, |g|^2)

enum { OPT_ADAM = 0, OPT_SGD = 1, OPT_RMSPROP = 2, OPT_ADAMW = 3 };

struct FitParams {
  const float* a2;    // [N][N]
  const float* bias;  // B [n_anc][K]
  const double* c0;
  const int32_t* ell_idx;  // [n_anc][kEll] columns (ancestor index) of the non-zeros of A2[anc, anc], -1 padded
  const float* ell_val;    // [n_anc][kEll]
  const int32_t* ell_ok;   // 1 when every row fits in kEll non-zeros
  float* theta;       // [n_anc][K] logits, in/out
  float* mom1;        // [n_anc][K] first moment (adam, adamw) / trace (sgd); unused by rmsprop
  float* mom2;        // [n_anc][K] second moment (adam, adamw, rmsprop); unused by sgd
  float* grad;        // [n_anc][K] scratch (only touched when the state is not shared-memory resident)
  float* losses;      // [n_iter] or null
  int32_t* argmax;    // [n_anc][L] or null
  double* slots;      // [2][grid][kSlotDoubles]
  unsigned long long* bar;
  int N, n_leaves, n_anc, Q;
  long long L, K;
  int opt;
  float lr, clip, weight_decay;
  int n_iter;
  long long count0;
  int sites_per_cta;
  int tile_sites;
  int resident;  // bit mask of the state arrays kept in shared memory: theta, grad, mom1, mom2, bias (generic kernel: 0 or 31)
  int debug;  // timing probes (TREX_FIT_DEBUG): 1 no barrier wait, 2 no product, 4 no update, 8 no softmax/VJP
};

__device__ __forceinline__ unsigned long long ld_acquire(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// fast kernel (fit_ancestors_fast.cu): Q in {4, 20}, 16-byte aligned arrays
size_t fit_fast_smem(int n_anc, int Q, int sites_per_cta, int tile_sites, int n_resident);
int launch_fit_fast(const FitParams& p, int grid, size_t smem, cudaStream_t s);

}  // namespace fit
}  // namespace trex
