// Whole ancestor-fitting loop on the device — SURVEY §8(f) rank 1.
//   run_trex_optimization_configurable   src/trex/evals/benchmark.py:75-200
//     loss_fn      = compute_surrogate_cost / compute_soft_cost(identity) of _update_seq_stacked  (:152-165, :210-232)
//     training_step = value_and_grad -> clip_by_global_norm(1.0) -> adam | sgd | rmsprop | adamw   (:41-72, :167-200)
// The reference runs n_iterations (default 10 000) of this under lax.fori_loop / lax.scan; every iteration
// is a handful of tiny XLA kernels, so at the benchmark shapes (c2: 127 nodes x 1000 sites x 4 states) it is
// bounded by launch latency.  Here ONE persistent cooperative kernel runs all iterations:
//  * leaves never change, so with A2 = diag(r + c) - Adj - Adj^T (symmetric; dS = A2 S) the leaf terms
//    are folded once:  B = A2[anc, leaf] S_leaf,  c0 = 1/2 <S_leaf, A2[leaf, leaf] S_leaf>;  per iteration
//        Y = A2[anc, anc] S_anc,   dS_anc = B + Y,   cost = c0 + <S_anc, B> + 1/2 <S_anc, Y>
//    — one (n_anc x n_anc) x (n_anc x K) product instead of the N x N x K Gram + dS pair;
//  * sites are independent in everything except the global gradient norm and the loss, so each CTA owns a
//    contiguous block of sites, keeps its logits / optimiser moments / gradient in shared memory for the whole
//    run (when they fit) and meets the others once per iteration: a monotonic-counter grid barrier that also
//    carries the per-CTA (cost, |g|^2) partials, summed by every CTA in the same fixed order (bit-identical
//    clip factor everywhere, run-to-run deterministic);
//  * softmax, the product, the softmax VJP, the clip and the optimiser update never leave the SM.
#include <algorithm>
#include <cstdlib>

#include "fit_ancestors.cuh"

namespace trex {

using namespace fit;

namespace {

// fixed-order block sum of two doubles (every thread gets the result)
__device__ __forceinline__ void block_sum2(double& a, double& b, double* red /* [2][kThreads/32] */) {
  a = warp_sum(a);
  b = warp_sum(b);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) { red[w] = a; red[kThreads / 32 + w] = b; }
  __syncthreads();
  a = b = 0.0;
#pragma unroll
  for (int i = 0; i < kThreads / 32; ++i) { a += red[i]; b += red[kThreads / 32 + i]; }
}

// ---- one-time folding of the adjacency and the leaves --------------------------------------
// A2 = diag(r + c) - Adj - Adj^T   (d cost / d S = A2 S for cost = 1/2 sum_ij Adj_ij |S_i - S_j|^2, tree.py:199-209)
__global__ void fit_a2_kernel(const float* __restrict__ adj, int N, float* __restrict__ a2) {
  const int i = blockIdx.x;
  __shared__ float red[2][32];
  float r = 0.f, c = 0.f;
  for (int j = threadIdx.x; j < N; j += blockDim.x) {
    r += adj[(size_t)i * N + j];
    c += adj[(size_t)j * N + i];
  }
  for (int o = 16; o > 0; o >>= 1) {
    r += __shfl_xor_sync(0xffffffffu, r, o);
    c += __shfl_xor_sync(0xffffffffu, c, o);
  }
  if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = r; red[1][threadIdx.x >> 5] = c; }
  __syncthreads();
  r = c = 0.f;
  for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { r += red[0][w]; c += red[1][w]; }
  for (int j = threadIdx.x; j < N; j += blockDim.x)
    a2[(size_t)i * N + j] = (i == j ? r + c : 0.f) - adj[(size_t)i * N + j] - adj[(size_t)j * N + i];
}

// B[a][col] = sum_{b < n_leaves} A2[n_leaves + a][b] S_leaf[b][col];   c0 += 1/2 S_leaf[b][col] (A2[leaf, leaf] S_leaf)[b][col]
__global__ void fit_bias_kernel(const float* __restrict__ a2, const float* __restrict__ s_leaf, int N, int n_leaves,
                                long long K, float* __restrict__ bias, double* __restrict__ c0) {
  const long long col = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int row = blockIdx.y;  // 0..N-1: leaf rows feed c0, ancestor rows B
  double part = 0.0;
  if (col < K) {
    const float* arow = a2 + (size_t)row * N;
    float acc = 0.f;
    for (int b = 0; b < n_leaves; ++b) acc = fmaf(arow[b], s_leaf[(size_t)b * K + col], acc);
    if (row >= n_leaves) bias[(size_t)(row - n_leaves) * K + col] = acc;
    else part = 0.5 * (double)acc * (double)s_leaf[(size_t)row * K + col];
  }
  if (row < n_leaves) block_atomic_add(part, c0);
}

// ELL form of A2[anc, anc]: a hard tree has at most 4 non-zeros per row (diagonal, parent, two children)
__global__ void fit_ell_kernel(const float* __restrict__ a2, int N, int n_leaves, int32_t* __restrict__ idx,
                               float* __restrict__ val, int32_t* __restrict__ ok) {
  const int a = blockIdx.x * blockDim.x + threadIdx.x;
  const int n_anc = N - n_leaves;
  if (a >= n_anc) return;
  const float* row = a2 + (size_t)(n_leaves + a) * N + n_leaves;
  int k = 0;
  bool fits = true;
  for (int b = 0; b < n_anc; ++b) {
    const float v = row[b];
    if (v != 0.f) {
      if (k < kEll) { idx[a * kEll + k] = b; val[a * kEll + k] = v; ++k; }
      else fits = false;
    }
  }
  for (; k < kEll; ++k) { idx[a * kEll + k] = -1; val[a * kEll + k] = 0.f; }
  if (!fits) *ok = 0;
}

// ---- the loop --------------------------------------------------------------------------------
template <int MR>
__device__ __forceinline__ void product_tile(const float* __restrict__ sA, int lda, const float* __restrict__ sS,
                                             float* __restrict__ sD, int ldt, int n_anc, int tc4,
                                             const float* __restrict__ bias_t, int ldb, int tcols) {
  // D[a][c] = B[a][c] + sum_b A[b][a] S[b][c]   (A symmetric, stored [b][a] so a micro-tile row is contiguous)
  const int RT = (n_anc + MR - 1) / MR;
  for (int id = threadIdx.x; id < RT * tc4; id += kThreads) {
    const int rt = id / tc4, ct = id - rt * tc4;
    float acc[MR][4];
#pragma unroll
    for (int i = 0; i < MR; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    const float* ap = sA + rt * MR;
    const float* sp = sS + ct * 4;
#pragma unroll 4
    for (int b = 0; b < n_anc; ++b) {
      float av[MR];
      if constexpr (MR == 4) {
        const float4 t = *reinterpret_cast<const float4*>(ap + b * lda);
        av[0] = t.x; av[1] = t.y; av[2] = t.z; av[3] = t.w;
      } else {
        const float2 t = *reinterpret_cast<const float2*>(ap + b * lda);
        av[0] = t.x; av[1] = t.y;
      }
      const float4 sv = *reinterpret_cast<const float4*>(sp + b * ldt);
#pragma unroll
      for (int i = 0; i < MR; ++i) {
        acc[i][0] = fmaf(av[i], sv.x, acc[i][0]);
        acc[i][1] = fmaf(av[i], sv.y, acc[i][1]);
        acc[i][2] = fmaf(av[i], sv.z, acc[i][2]);
        acc[i][3] = fmaf(av[i], sv.w, acc[i][3]);
      }
    }
#pragma unroll
    for (int i = 0; i < MR; ++i) {
      const int a = rt * MR + i;
      if (a >= n_anc) break;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int c = ct * 4 + j;
        if (c < tcols) sD[a * ldt + c] = acc[i][j] + bias_t[(size_t)a * ldb + c];
      }
    }
  }
}

__global__ void __launch_bounds__(kThreads) fit_ancestors_kernel(const FitParams p) {
  extern __shared__ __align__(16) float smem[];
  __shared__ double red[2 * (kThreads / 32)];
  const int n_anc = p.n_anc, Q = p.Q;
  const int lda = (n_anc + 3) & ~3;
  const int ldt = p.tile_sites * Q + 4 - ((p.tile_sites * Q) & 3);  // tile pitch: multiple of 4, >= tile columns + 1
  float* sA = smem;                       // [n_anc][lda]   A2[anc, anc] (symmetric)
  float* sS = sA + (size_t)n_anc * lda;   // [n_anc][ldt]   softmax of the tile
  float* sD = sS + (size_t)n_anc * ldt;   // [n_anc][ldt]   dS = B + A S
  float* sRes = sD + (size_t)n_anc * ldt; // resident state: theta, mom1, mom2, grad, bias  [n_anc][ldc] each

  const long long site_lo = (long long)blockIdx.x * p.sites_per_cta;
  const long long site_hi = min(p.L, site_lo + p.sites_per_cta);
  const int nsites = (int)(site_hi - site_lo);
  const int ncols = nsites * Q;
  const long long col_lo = site_lo * Q;
  const int G = gridDim.x;

  // state accessors: shared-memory slabs [n_anc][ldc] or the global arrays [n_anc][K] at this CTA's columns
  const int ldc = (ncols + 3) & ~3;
  const int ldc_max = (p.sites_per_cta * Q + 3) & ~3;
  float *th, *m1, *m2, *gr;
  const float* bi;
  long long ld;
  if (p.resident) {
    th = sRes; m1 = th + (size_t)n_anc * ldc_max; m2 = m1 + (size_t)n_anc * ldc_max;
    gr = m2 + (size_t)n_anc * ldc_max;
    float* sb = gr + (size_t)n_anc * ldc_max;
    bi = sb;
    ld = ldc;
    for (int i = threadIdx.x; i < n_anc * ncols; i += kThreads) {
      const int a = i / ncols, c = i - a * ncols;
      const size_t gi = (size_t)a * p.K + col_lo + c;
      th[a * ldc + c] = p.theta[gi];
      m1[a * ldc + c] = p.mom1[gi];
      m2[a * ldc + c] = p.mom2[gi];
      sb[a * ldc + c] = p.bias[gi];
    }
  } else {
    th = p.theta + col_lo; m1 = p.mom1 + col_lo; m2 = p.mom2 + col_lo; gr = p.grad + col_lo;
    bi = p.bias + col_lo;
    ld = p.K;
  }
  for (int i = threadIdx.x; i < n_anc * lda; i += kThreads) {
    const int b = i / lda, a = i - b * lda;
    sA[i] = a < n_anc ? p.a2[(size_t)(p.n_leaves + b) * p.N + p.n_leaves + a] : 0.f;
  }
  for (int i = threadIdx.x; i < 2 * n_anc * ldt; i += kThreads) sS[i] = 0.f;  // pad columns stay zero
  __syncthreads();
  const double c0 = *p.c0;

  for (int it = 0; it < p.n_iter; ++it) {
    double cost = 0.0, norm2 = 0.0;
    for (int s0 = 0; s0 < nsites; s0 += p.tile_sites) {
      const int ts = min(p.tile_sites, nsites - s0);
      const int tcols = ts * Q;
      const int tc4 = (tcols + 3) >> 2;
      const int cbase = s0 * Q;
      // softmax of the logits (temperature 1: _update_seq_stacked, evals/benchmark.py:231)
      for (int i = threadIdx.x; i < ((p.debug & 8) ? 0 : n_anc * ts); i += kThreads) {
        const int a = i / ts, s = i - a * ts;
        const float* x = th + (size_t)a * ld + cbase + s * Q;
        float* out = sS + a * ldt + s * Q;
        float mx = -INFINITY;
        for (int q = 0; q < Q; ++q) mx = fmaxf(mx, x[q]);
        if (!(mx > -INFINITY && mx < INFINITY)) mx = 0.f;
        float z = 0.f;
        for (int q = 0; q < Q; ++q) { const float e = expf(x[q] - mx); out[q] = e; z += e; }
        const float inv = 1.f / z;
        for (int q = 0; q < Q; ++q) out[q] *= inv;
      }
      if (tcols & 3)  // a narrower last tile: clear what the previous one left in the pad columns
        for (int i = threadIdx.x; i < n_anc * (4 - (tcols & 3)); i += kThreads) {
          const int w = 4 - (tcols & 3);
          sS[(i / w) * ldt + tcols + i % w] = 0.f;
        }
      __syncthreads();
      const int RT4 = (n_anc + 3) >> 2;
      if (p.debug & 2) {}
      else if (RT4 * tc4 >= kThreads * 3 / 4) product_tile<4>(sA, lda, sS, sD, ldt, n_anc, tc4, bi + cbase, (int)ld, tcols);
      else product_tile<2>(sA, lda, sS, sD, ldt, n_anc, tc4, bi + cbase, (int)ld, tcols);
      __syncthreads();
      // softmax VJP, the loss and |g|^2
      float fc = 0.f, fn = 0.f;
      for (int i = threadIdx.x; i < ((p.debug & 8) ? 0 : n_anc * ts); i += kThreads) {
        const int a = i / ts, s = i - a * ts;
        const float* S = sS + a * ldt + s * Q;
        const float* D = sD + a * ldt + s * Q;
        const float* Bv = bi + (size_t)a * ld + cbase + s * Q;
        float* gout = gr + (size_t)a * ld + cbase + s * Q;
        float dot = 0.f;
        for (int q = 0; q < Q; ++q) {
          dot = fmaf(S[q], D[q], dot);
          fc = fmaf(S[q], 0.5f * (D[q] + Bv[q]), fc);  // <S, B> + 1/2 <S, Y>
        }
        for (int q = 0; q < Q; ++q) {
          const float g = S[q] * (D[q] - dot);
          gout[q] = g;
          fn = fmaf(g, g, fn);
        }
      }
      cost += (double)fc;
      norm2 += (double)fn;
      __syncthreads();
    }
    // ---- meet the other CTAs: (cost, |g|^2) partials + grid barrier ----------------------------
    block_sum2(cost, norm2, red);
    double* slot = p.slots + (size_t)(it & 1) * 2 * G;
    if (threadIdx.x == 0) {
      slot[2 * blockIdx.x] = cost;
      slot[2 * blockIdx.x + 1] = norm2;
      __threadfence();
      atomicAdd(p.bar, 1ULL);
      const unsigned long long want = (unsigned long long)(it + 1) * G;
      if (!(p.debug & 1)) while (ld_acquire(p.bar) < want) {}
    }
    __syncthreads();
    cost = norm2 = 0.0;
    for (int i = threadIdx.x; i < G; i += kThreads) {
      cost += __ldcg(slot + 2 * i);
      norm2 += __ldcg(slot + 2 * i + 1);
    }
    block_sum2(cost, norm2, red);
    if (p.losses && blockIdx.x == 0 && threadIdx.x == 0) p.losses[it] = (float)(c0 + cost);

    // ---- optax.clip_by_global_norm + the optimiser (evals/benchmark.py:41-72) --------------------
    const float gnorm = sqrtf((float)norm2);
    const bool clipped = p.clip > 0.f && !(gnorm < p.clip);
    const float t = (float)(p.count0 + it + 1);
    const float b1 = 0.9f, b2 = 0.999f, eps = 1e-8f;
    const float bc1 = 1.f - powf(b1, t), bc2 = 1.f - powf(b2, t);
    for (int i = threadIdx.x; i < ((p.debug & 4) ? 0 : n_anc * ncols); i += kThreads) {
      const int a = i / ncols, c = i - a * ncols;
      const size_t o = (size_t)a * ld + c;
      float g = gr[o];
      if (clipped) g = (g / gnorm) * p.clip;
      float x = th[o], u;
      if (p.opt == OPT_SGD) {  // optax.sgd(lr, momentum=0.9): trace
        const float tr = g + 0.9f * m1[o];
        m1[o] = tr;
        u = tr;
      } else if (p.opt == OPT_RMSPROP) {  // optax.rmsprop(lr): decay 0.9, eps 1e-8 inside the sqrt
        const float nu = 0.9f * m2[o] + (1.f - 0.9f) * g * g;
        m2[o] = nu;
        u = g * rsqrtf(nu + eps);
      } else {  // optax.adam / adamw(weight_decay=0.01)
        const float mu = (1.f - b1) * g + b1 * m1[o];
        const float nu = (1.f - b2) * (g * g) + b2 * m2[o];
        m1[o] = mu;
        m2[o] = nu;
        u = (mu / bc1) / (sqrtf(nu / bc2) + eps);
        if (p.opt == OPT_ADAMW) u += p.weight_decay * x;
      }
      th[o] = x - p.lr * u;
    }
    __syncthreads();
  }

  if (p.resident) {
    for (int i = threadIdx.x; i < n_anc * ncols; i += kThreads) {
      const int a = i / ncols, c = i - a * ncols;
      const size_t gi = (size_t)a * p.K + col_lo + c;
      p.theta[gi] = th[a * ldc + c];
      p.mom1[gi] = m1[a * ldc + c];
      p.mom2[gi] = m2[a * ldc + c];
    }
  }
  if (p.argmax) {  // jnp.argmax(ancestors, axis=-1): first maximum (evals/benchmark.py:200)
    for (int i = threadIdx.x; i < n_anc * nsites; i += kThreads) {
      const int a = i / nsites, s = i - a * nsites;
      const float* x = th + (size_t)a * ld + s * Q;
      int best = 0;
      float bv = x[0];
      for (int q = 1; q < Q; ++q)
        if (x[q] > bv) { bv = x[q]; best = q; }
      p.argmax[(size_t)a * p.L + site_lo + s] = best;
    }
  }
}

struct FitLayout {
  size_t a2, bias, grad, ell_idx, ell_val, slots, c0, bar, ell_ok, tail, total;
  int grid, sites_per_cta, tile_sites, resident, fast;
  size_t smem;
};

FitLayout fit_layout(int N, int n_leaves, long long L, int Q) {
  FitLayout f{};
  const int n_anc = N - n_leaves;
  const long long K = L * Q;
  int g = (int)std::min<long long>(L, sm_count());
  f.sites_per_cta = (int)((L + g - 1) / g);
  f.grid = (int)((L + f.sites_per_cta - 1) / f.sites_per_cta);
  const size_t budget = 220 * 1024;
  f.fast = (Q == 4 || Q == 20) && n_anc <= kMaxAncFast;
  if (const char* e = getenv("TREX_FIT_PATH")) {  // tests: "generic" forces the scalar kernel
    if (e[0] == 'g') f.fast = 0;
  }
  if (f.fast) {
    const size_t per_site = sizeof(float) * (size_t)n_anc * Q;
    f.tile_sites = (int)std::max<size_t>(1, std::min<size_t>(f.sites_per_cta, (64 * 1024) / per_site));
    // as many of the five state arrays as fit stay in shared memory (theta, grad, mom1, mom2, bias: bits 0..4);
    // a smaller softmax tile is traded for one more resident array
    f.resident = 0;
    const int tile_min = std::min(f.sites_per_cta, (kFastThreads + n_anc - 1) / n_anc);  // a tile keeps every thread busy
    int n_res = 0;
    while (n_res < 5 && fit_fast_smem(n_anc, Q, f.sites_per_cta, tile_min, n_res + 1) <= budget) ++n_res;
    f.resident = (1 << n_res) - 1;
    f.tile_sites = std::max(f.tile_sites, tile_min);
    while (f.tile_sites > tile_min && fit_fast_smem(n_anc, Q, f.sites_per_cta, f.tile_sites, n_res) > budget) --f.tile_sites;
    f.smem = fit_fast_smem(n_anc, Q, f.sites_per_cta, f.tile_sites, n_res);
  } else {
    f.tile_sites = std::max(1, std::min(f.sites_per_cta, kTileCols / Q));
    const int lda = (n_anc + 3) & ~3;
    const int ldt = f.tile_sites * Q + 4 - ((f.tile_sites * Q) & 3);
    const size_t base = sizeof(float) * ((size_t)n_anc * lda + 2 * (size_t)n_anc * ldt);
    const size_t res = sizeof(float) * 5 * (size_t)n_anc * (((size_t)f.sites_per_cta * Q + 3) & ~(size_t)3);
    f.resident = base + res <= budget ? 31 : 0;
    f.smem = base + (f.resident ? res : 0);
  }
  auto up = [](size_t x) { return (x + 255) & ~(size_t)255; };
  f.a2 = 0;
  f.bias = up(sizeof(float) * (size_t)N * N);
  f.grad = f.bias + up(sizeof(float) * (size_t)n_anc * K);
  f.ell_idx = f.grad + up(sizeof(float) * (size_t)n_anc * K);
  f.ell_val = f.ell_idx + up(sizeof(int32_t) * (size_t)n_anc * kEll);
  f.slots = f.ell_val + up(sizeof(float) * (size_t)n_anc * kEll);  // zeroed per call from here on
  f.c0 = f.slots + up(sizeof(double) * 2 * kSlotDoubles * (size_t)f.grid);
  f.bar = f.c0 + 64;
  f.ell_ok = f.bar + 64;
  f.tail = f.ell_ok + 64 - f.slots;
  f.total = f.ell_ok + 64;
  return f;
}

}  // namespace

}  // namespace trex

using namespace trex;

extern "C" {

size_t trex_fit_ancestors_workspace_bytes(int n_nodes, int n_leaves, int64_t n_sites, int n_states) {
  if (n_nodes <= n_leaves || n_leaves < 1 || n_sites < 1 || n_states < 1) return 0;
  return fit_layout(n_nodes, n_leaves, n_sites, n_states).total;
}

int trex_fit_ancestors(const float* adjacency, const float* leaf_seqs, int n_nodes, int n_leaves, int64_t n_sites,
                       int n_states, float* ancestor_logits, float* opt_mom1, float* opt_mom2, int optimizer,
                       float learning_rate, float clip_norm, int n_iterations, int64_t step_count,
                       float* out_losses, int32_t* out_argmax, void* workspace, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  TREX_CHECK_ARG(adjacency && leaf_seqs && ancestor_logits && opt_mom1 && opt_mom2 && workspace,
                 "trex_fit_ancestors: null pointer");
  TREX_CHECK_ARG(n_leaves >= 1 && n_nodes > n_leaves, "trex_fit_ancestors: need n_nodes > n_leaves >= 1");
  TREX_CHECK_ARG(n_sites >= 1 && n_states >= 1 && n_states <= kTileCols,
                 "trex_fit_ancestors: n_sites >= 1 and 1 <= n_states <= %d required", kTileCols);
  TREX_CHECK_ARG(optimizer >= OPT_ADAM && optimizer <= OPT_ADAMW, "trex_fit_ancestors: unknown optimizer %d", optimizer);
  TREX_CHECK_ARG(n_iterations >= 0 && step_count >= 0, "trex_fit_ancestors: negative iteration count");
  const int n_anc = n_nodes - n_leaves;
  const FitLayout f = fit_layout(n_nodes, n_leaves, n_sites, n_states);
  if (n_anc > (f.fast ? kMaxAncFast : kMaxAnc) || f.smem > 225 * 1024) {
    set_error("trex_fit_ancestors: %d ancestors x %d states does not fit (limit %d ancestors for 4 / 20 states, %d "
              "otherwise)", n_anc, n_states, kMaxAncFast, kMaxAnc);
    return TREX_ERR_UNSUPPORTED;
  }
  char* ws = (char*)workspace;
  FitParams p{};
  p.a2 = (const float*)(ws + f.a2);
  p.bias = (const float*)(ws + f.bias);
  p.c0 = (const double*)(ws + f.c0);
  p.theta = ancestor_logits;
  p.mom1 = opt_mom1;
  p.mom2 = opt_mom2;
  p.grad = (float*)(ws + f.grad);
  p.losses = out_losses;
  p.argmax = out_argmax;
  p.slots = (double*)(ws + f.slots);
  p.bar = (unsigned long long*)(ws + f.bar);
  p.N = n_nodes; p.n_leaves = n_leaves; p.n_anc = n_anc; p.Q = n_states;
  p.L = n_sites; p.K = (long long)n_sites * n_states;
  p.opt = optimizer; p.lr = learning_rate; p.clip = clip_norm; p.weight_decay = 0.01f;
  p.n_iter = n_iterations; p.count0 = step_count;
  p.sites_per_cta = f.sites_per_cta; p.tile_sites = f.tile_sites; p.resident = f.resident;
  { const char* e = getenv("TREX_FIT_DEBUG"); p.debug = e ? atoi(e) : 0; }

  p.ell_idx = (const int32_t*)(ws + f.ell_idx);
  p.ell_val = (const float*)(ws + f.ell_val);
  p.ell_ok = (const int32_t*)(ws + f.ell_ok);
  const bool aligned = (((uintptr_t)ancestor_logits | (uintptr_t)opt_mom1 | (uintptr_t)opt_mom2 | (uintptr_t)workspace) & 15) == 0;
  const bool fast = f.fast && aligned;
  if (f.fast && !aligned) {
    set_error("trex_fit_ancestors: ancestor_logits / opt_mom1 / opt_mom2 / workspace must be 16-byte aligned");
    return TREX_ERR_INVALID_ARGUMENT;
  }

  int rc = check_cuda(cudaMemsetAsync(ws + f.slots, 0, f.tail, s), "cudaMemsetAsync(fit slots/c0/barrier)");
  if (rc) return rc;
  rc = check_cuda(cudaMemsetAsync(ws + f.ell_ok, 1, 1, s), "cudaMemsetAsync(fit ell_ok)");  // little-endian 1
  if (rc) return rc;
  fit_a2_kernel<<<n_nodes, 128, 0, s>>>(adjacency, n_nodes, (float*)(ws + f.a2));
  note_launches(1);
  TREX_CHECK_LAUNCH("fit_a2_kernel");
  fit_bias_kernel<<<dim3((unsigned)((p.K + 255) / 256), n_nodes), 256, 0, s>>>(
      p.a2, leaf_seqs, n_nodes, n_leaves, p.K, (float*)(ws + f.bias), (double*)(ws + f.c0));
  note_launches(1);
  TREX_CHECK_LAUNCH("fit_bias_kernel");

  if (fast) {
    fit_ell_kernel<<<(n_anc + 127) / 128, 128, 0, s>>>(p.a2, n_nodes, n_leaves, (int32_t*)(ws + f.ell_idx),
                                                      (float*)(ws + f.ell_val), (int32_t*)(ws + f.ell_ok));
    note_launches(1);
    TREX_CHECK_LAUNCH("fit_ell_kernel");
    return launch_fit_fast(p, f.grid, f.smem, s);
  }
  rc = check_cuda(cudaFuncSetAttribute(fit_ancestors_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)f.smem),
                  "cudaFuncSetAttribute(fit_ancestors_kernel)");
  if (rc) return rc;
  void* args[] = {(void*)&p};
  rc = check_cuda(cudaLaunchCooperativeKernel((const void*)fit_ancestors_kernel, dim3(f.grid), dim3(kThreads), args,
                                              f.smem, s),
                  "cudaLaunchCooperativeKernel(fit_ancestors_kernel)");
  if (rc) return rc;
  note_launches(1);
  return TREX_OK;
}

}  // extern "C"
