// Ancestor-fitting loop, fast path: alphabets of 4 (DNA) and 20 (protein) states — every benchmark config.
// Same loop as fit_ancestors.cu (evals/benchmark.py:75-200), organised around ONE work item = one
// (ancestor, site) row of Q contiguous floats handled by one thread with 128-bit accesses:
//   A1  x = logits row -> softmax -> S tile in shared memory
//   A2  Y = sum_b A2[a][b] S[b] for this row (ELL-sparse when every row of A2[anc, anc] has <= kEll
//       non-zeros — a hard tree has 4 — dense from shared memory otherwise), dS = B + Y, softmax VJP,
//       g row stored, partial loss and |g|^2
//   --  partials published in a per-CTA slot whose tag the other CTAs poll directly: the barrier and
//       the reduction cost one L2 round trip; every CTA sums the slots in the same order
//   B   clip + optimiser update of the row
// Lanes of a warp are consecutive sites of one ancestor: S rows are read conflict-free, A2 entries are
// broadcast, and state that does not fit in shared memory streams from L2 fully coalesced.
#include <cstdio>

#include "fit_ancestors.cuh"

namespace trex {
namespace fit {

namespace {

constexpr int NT = kFastThreads;
constexpr int NW = NT / 32;
constexpr int kPoll = 5;  // 160 slots per polling round (B200: 148 CTAs)

template <int Q>
__device__ __forceinline__ void ld_row(const float* __restrict__ p, float (&v)[Q]) {
#pragma unroll
  for (int j = 0; j < Q / 4; ++j) {
    const float4 t = *reinterpret_cast<const float4*>(p + 4 * j);
    v[4 * j] = t.x; v[4 * j + 1] = t.y; v[4 * j + 2] = t.z; v[4 * j + 3] = t.w;
  }
}
template <int Q>
__device__ __forceinline__ void st_row(float* __restrict__ p, const float (&v)[Q]) {
#pragma unroll
  for (int j = 0; j < Q / 4; ++j)
    *reinterpret_cast<float4*>(p + 4 * j) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
}
template <int Q>
__device__ __forceinline__ void axpy_row(float w, const float* __restrict__ row, float (&acc)[Q]) {
  float r[Q];
  ld_row<Q>(row, r);
#pragma unroll
  for (int q = 0; q < Q; ++q) acc[q] = fmaf(w, r[q], acc[q]);
}

// One row of the optimiser update (evals/benchmark.py:41-72); g is the raw gradient, gscale the clip factor.
// mu: first moment (adam, adamw) or momentum trace (sgd); nu: second moment (adam, adamw, rmsprop).
template <int Q>
__device__ __forceinline__ void opt_row(int opt, float lr, float wd, float ibc1, float ibc2, float gscale,
                                        const float (&g0)[Q], float (&x)[Q], float (&mu)[Q], float (&nu)[Q]) {
  const float b1 = 0.9f, b2 = 0.999f, eps = 1e-8f;
#pragma unroll
  for (int q = 0; q < Q; ++q) {
    const float g = g0[q] * gscale;
    if (opt == OPT_SGD) {  // optax.sgd(lr, momentum=0.9): trace
      mu[q] = fmaf(0.9f, mu[q], g);
      x[q] = fmaf(-lr, mu[q], x[q]);
    } else if (opt == OPT_RMSPROP) {  // optax.rmsprop(lr): decay 0.9, eps 1e-8 inside the sqrt
      nu[q] = fmaf(0.9f, nu[q], (1.f - 0.9f) * g * g);
      x[q] = fmaf(-lr, g * rsqrtf(nu[q] + eps), x[q]);
    } else {  // optax.adam / adamw(weight_decay=0.01)
      mu[q] = fmaf(b1, mu[q], (1.f - b1) * g);
      nu[q] = fmaf(b2, nu[q], (1.f - b2) * (g * g));
      float u = __fdividef(mu[q] * ibc1, sqrtf(nu[q] * ibc2) + eps);
      if (opt == OPT_ADAMW) u = fmaf(wd, x[q], u);
      x[q] = fmaf(-lr, u, x[q]);
    }
  }
}

template <int Q>
__global__ void __launch_bounds__(NT, 1) fit_fast_kernel(const FitParams p) {
  extern __shared__ __align__(16) float smem[];
  __shared__ double red[2 * NW + 2];
  const int n_anc = p.n_anc;
  const bool big = n_anc > kMaxAnc;  // no room for the dense A2[anc, anc]: read it from global memory (L2)
  const int lda = big ? p.N : (n_anc + 3) & ~3;
  const int ldt = p.tile_sites * Q;
  float* sA_smem = smem;                                          // [n_anc][lda]   dense A2[anc, anc] (!big)
  const float* sA = big ? p.a2 + (size_t)p.n_leaves * p.N + p.n_leaves : sA_smem;
  int* sIdx = reinterpret_cast<int*>(sA_smem + (big ? 0 : (size_t)n_anc * lda));   // [n_anc][kEll]
  float* sVal = reinterpret_cast<float*>(sIdx + n_anc * kEll);    // [n_anc][kEll]
  float* sS = sVal + n_anc * kEll;                                // [n_anc][ldt]   softmax of the tile
  float* sRes = sS + (size_t)n_anc * ldt;                         // resident state, [n_anc][ldc] each

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long long site_lo = (long long)blockIdx.x * p.sites_per_cta;
  const int nsites = (int)(min(p.L, site_lo + p.sites_per_cta) - site_lo);
  const long long col_lo = site_lo * Q;
  const int G = gridDim.x;
  const bool sparse = *p.ell_ok == 0 ? false : true;

  // Per-array placement: bit k of p.resident puts array k (theta, grad, mom1, mom2, bias — the order of
  // their accesses per iteration) in a shared-memory slab [n_anc][ldc]; the others stream from L2.
  const int ldc = p.sites_per_cta * Q;
  float* slab = sRes;
  auto place = [&](int bit, float* garr, long long& ld_out) -> float* {
    if ((p.resident >> bit) & 1) {
      float* q = slab;
      slab += (size_t)n_anc * ldc;
      ld_out = ldc;
      return q;
    }
    ld_out = p.K;
    return garr + col_lo;
  };
  long long ld_th, ld_gr, ld_m1, ld_m2, ld_bi;
  float* th = place(0, p.theta, ld_th);
  float* gr = place(1, p.grad, ld_gr);
  float* m1 = place(2, p.mom1, ld_m1);
  float* m2 = place(3, p.mom2, ld_m2);
  float* sb = place(4, const_cast<float*>(p.bias), ld_bi);
  const float* bi = sb;
  for (int i = tid; i < n_anc * nsites; i += NT) {
    const int a = i / nsites, s = i - a * nsites;
    const size_t gi = (size_t)a * p.K + col_lo + s * Q;
    const int o = a * ldc + s * Q;
    float v[Q];
    if (p.resident & 1) { ld_row<Q>(p.theta + gi, v); st_row<Q>(th + o, v); }
    if (p.resident & 4) { ld_row<Q>(p.mom1 + gi, v); st_row<Q>(m1 + o, v); }
    if (p.resident & 8) { ld_row<Q>(p.mom2 + gi, v); st_row<Q>(m2 + o, v); }
    if (p.resident & 16) { ld_row<Q>(p.bias + gi, v); st_row<Q>(sb + o, v); }
  }
  if (sparse) {
    for (int i = tid; i < n_anc * kEll; i += NT) { sIdx[i] = p.ell_idx[i]; sVal[i] = p.ell_val[i]; }
  } else if (!big) {
    for (int i = tid; i < n_anc * lda; i += NT) {
      const int b = i / lda, a = i - b * lda;
      sA_smem[i] = a < n_anc ? p.a2[(size_t)(p.n_leaves + b) * p.N + p.n_leaves + a] : 0.f;
    }
  }
  __syncthreads();
  const double c0 = *p.c0;
  const float b1 = 0.9f, b2 = 0.999f;
  double pw1 = pow((double)b1, (double)p.count0), pw2 = pow((double)b2, (double)p.count0);  // b^t, kept running

#ifdef TREX_FIT_PROFILE  // per-phase cycle counts of tools/fit_probe.py (build with -DTREX_FIT_PROFILE, run with TREX_FIT_DEBUG=16)
  long long prof[6] = {0, 0, 0, 0, 0, 0}, tk = clock64();
  auto tick = [&](int k) { if (p.debug & 16) { const long long t = clock64(); prof[k] += t - tk; tk = t; } };
#else
  auto tick = [](int) {};
#endif
  for (int it = 0; it < p.n_iter; ++it) {
    float fc = 0.f, fn = 0.f;
    for (int s0 = 0; s0 < nsites; s0 += p.tile_sites) {
      const int ts = min(p.tile_sites, nsites - s0);
      const int nitems = (p.debug & 8) ? 0 : n_anc * ts;
      // A1: softmax of the logits (temperature 1: _update_seq_stacked, evals/benchmark.py:231)
      for (int i = tid; i < nitems; i += NT) {
        const int a = i / ts, s = i - a * ts;
        float x[Q];
        ld_row<Q>(th + (size_t)a * ld_th + (s0 + s) * Q, x);
        float mx = x[0];
#pragma unroll
        for (int q = 1; q < Q; ++q) mx = fmaxf(mx, x[q]);
        if (!(mx > -INFINITY && mx < INFINITY)) mx = 0.f;
        float z = 0.f;
#pragma unroll
        for (int q = 0; q < Q; ++q) { x[q] = expf(x[q] - mx); z += x[q]; }
        const float inv = 1.f / z;
#pragma unroll
        for (int q = 0; q < Q; ++q) x[q] *= inv;
        st_row<Q>(sS + a * ldt + s * Q, x);
      }
      __syncthreads();
      // A2: dS row = B + (A2 S) row, softmax VJP, loss and |g|^2 partials
      for (int i = tid; i < nitems; i += NT) {
        const int a = i / ts, s = i - a * ts;
        float y[Q];
#pragma unroll
        for (int q = 0; q < Q; ++q) y[q] = 0.f;
        const float* scol = sS + s * Q;
        if (p.debug & 2) {
        } else if (sparse) {
#pragma unroll
          for (int j = 0; j < kEll; ++j) {
            const int b = sIdx[a * kEll + j];
            if (b < 0) break;
            axpy_row<Q>(sVal[a * kEll + j], scol + b * ldt, y);
          }
        } else {
#pragma unroll 2
          for (int b = 0; b < n_anc; ++b) axpy_row<Q>(sA[b * lda + a], scol + b * ldt, y);
        }
        float S[Q], bv[Q];
        ld_row<Q>(scol + a * ldt, S);
        ld_row<Q>(bi + (size_t)a * ld_bi + (s0 + s) * Q, bv);
        float dot = 0.f;
#pragma unroll
        for (int q = 0; q < Q; ++q) {
          fc = fmaf(S[q], fmaf(0.5f, y[q], bv[q]), fc);  // <S, B> + 1/2 <S, Y>
          y[q] += bv[q];
          dot = fmaf(S[q], y[q], dot);
        }
#pragma unroll
        for (int q = 0; q < Q; ++q) {
          y[q] = S[q] * (y[q] - dot);
          fn = fmaf(y[q], y[q], fn);
        }
        st_row<Q>(gr + (size_t)a * ld_gr + (s0 + s) * Q, y);
      }
      __syncthreads();
    }

    tick(0);
    // ---- publish (cost, |g|^2), meet the other CTAs, sum all slots in a fixed order -------------
    // Nothing but these two numbers ever crosses CTAs, so no fence is needed: each double carries a 2-bit
    // iteration tag in its lowest mantissa bits (8-byte accesses are single-copy atomic; the payload moves
    // by < 2^-50 relative, identically for every reader).  Warp 0 publishes the CTA's sums, then polls the
    // slots of all CTAs — lane l owns slots l, l + 32, ... — until every tag is this iteration's.
    double cost, norm2;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      fc += __shfl_xor_sync(0xffffffffu, fc, o);
      fn += __shfl_xor_sync(0xffffffffu, fn, o);
    }
    if (lane == 0) { red[warp] = (double)fc; red[NW + warp] = (double)fn; }
    __syncthreads();
    tick(1);
    pw1 *= (double)b1;
    pw2 *= (double)b2;
    const float ibc1 = 1.f / (1.f - (float)pw1), ibc2 = 1.f / (1.f - (float)pw2);
    const int nupd = (p.debug & 4) ? 0 : n_anc * nsites;
    // Row state of the update in registers.  When a CTA has at most one row per thread (c2), the warps that
    // do not poll run the update while the barrier is in flight, assuming the gradient is not clipped
    // (|g| < clip: the common case away from the first iterations); a clipped step is simply redone.
    float ug[Q], ux[Q], umu[Q], unu[Q];
    size_t og, ox, o1, o2;
    auto load_rows = [&](int i) {
      const int a = i / nsites, s = i - a * nsites;
      og = (size_t)a * ld_gr + s * Q; ox = (size_t)a * ld_th + s * Q;
      o1 = (size_t)a * ld_m1 + s * Q; o2 = (size_t)a * ld_m2 + s * Q;
      ld_row<Q>(gr + og, ug);
      ld_row<Q>(th + ox, ux);
      if (p.opt != OPT_RMSPROP) ld_row<Q>(m1 + o1, umu);
      if (p.opt != OPT_SGD) ld_row<Q>(m2 + o2, unu);
    };
    auto store_rows = [&](int) {
      if (p.opt != OPT_RMSPROP) st_row<Q>(m1 + o1, umu);
      if (p.opt != OPT_SGD) st_row<Q>(m2 + o2, unu);
      st_row<Q>(th + ox, ux);
    };
    // (warp 0 polls, so it owns no row: row r belongs to thread r + 32)
    const bool spec = Q <= 4 && n_anc * nsites <= NT - 32;
    bool spec_done = false;
    if (spec && warp != 0 && tid - 32 < nupd) {
      load_rows(tid - 32);
      opt_row<Q>(p.opt, p.lr, p.weight_decay, ibc1, ibc2, 1.f, ug, ux, umu, unu);
      spec_done = true;
    }
    if (warp == 0) {
      const unsigned long long tag = (unsigned long long)((it + 1) & 3);
      unsigned long long* slots = reinterpret_cast<unsigned long long*>(p.slots) + (size_t)(it & 1) * kSlotDoubles * G;
      cost = lane < NW ? red[lane] : 0.0;
      norm2 = lane < NW ? red[NW + lane] : 0.0;
      cost = warp_sum(cost);
      norm2 = warp_sum(norm2);
      if (lane == 0) {
        const unsigned long long c = ((unsigned long long)__double_as_longlong(cost) & ~3ull) | tag;
        const unsigned long long n = ((unsigned long long)__double_as_longlong(norm2) & ~3ull) | tag;
        asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(slots + (size_t)kSlotDoubles * blockIdx.x), "l"(c), "l"(n) : "memory");
      }
      __syncwarp();
      tick(2);
      // kPoll slots per lane are in flight together: one L2 round trip per polling round
      cost = norm2 = 0.0;
      for (int i0 = 0; i0 < G; i0 += 32 * kPoll) {
        unsigned long long c[kPoll], n[kPoll];
        bool again;
        do {
          again = false;
#pragma unroll
          for (int k = 0; k < kPoll; ++k) {
            const int i = min(i0 + k * 32 + lane, G - 1);
            asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(c[k]), "=l"(n[k]) : "l"(slots + (size_t)kSlotDoubles * i) : "memory");
          }
#pragma unroll
          for (int k = 0; k < kPoll; ++k) again = again || (c[k] & 3ull) != tag || (n[k] & 3ull) != tag;
        } while (again && !(p.debug & 1));
#pragma unroll
        for (int k = 0; k < kPoll; ++k)
          if (i0 + k * 32 + lane < G) {
            cost += __longlong_as_double((long long)(c[k] & ~3ull));
            norm2 += __longlong_as_double((long long)(n[k] & ~3ull));
          }
      }
      cost = warp_sum(cost);
      norm2 = warp_sum(norm2);
      if (lane == 0) { red[2 * NW] = cost; red[2 * NW + 1] = norm2; }
      tick(3);
    }
    __syncthreads();
    tick(4);
    cost = red[2 * NW];
    norm2 = red[2 * NW + 1];
    if (p.losses && blockIdx.x == 0 && tid == 0) p.losses[it] = (float)(c0 + cost);

    // ---- B: optax.clip_by_global_norm + the optimiser (evals/benchmark.py:41-72) -------------------
    const float gnorm = sqrtf((float)norm2);
    const float gscale = (p.clip > 0.f && !(gnorm < p.clip)) ? p.clip / gnorm : 1.f;
    if (spec) {
      // the rows were updated ahead of the barrier on the assumption that nothing is clipped (see above)
      if (spec_done) {
        if (gscale != 1.f) {
          load_rows(tid - 32);
          opt_row<Q>(p.opt, p.lr, p.weight_decay, ibc1, ibc2, gscale, ug, ux, umu, unu);
        }
        store_rows(tid - 32);
      }
    } else {
      for (int i = tid; i < nupd; i += NT) {
        load_rows(i);
        opt_row<Q>(p.opt, p.lr, p.weight_decay, ibc1, ibc2, gscale, ug, ux, umu, unu);
        store_rows(i);
      }
    }
    __syncthreads();
    tick(5);
  }
#ifdef TREX_FIT_PROFILE
  if ((p.debug & 16) && tid == 0 && (blockIdx.x == 0 || blockIdx.x == G - 1 || (p.debug & 32)))
    { unsigned smid; asm("mov.u32 %0, %smid;" : "=r"(smid)); printf("fit_fast sm %u cta %d cycles/iter: phaseA %lld reduce %lld publish %lld poll %lld sync %lld update %lld\n", smid, blockIdx.x,
           prof[0] / p.n_iter, prof[1] / p.n_iter, prof[2] / p.n_iter, prof[3] / p.n_iter, prof[4] / p.n_iter, prof[5] / p.n_iter); }
#endif

  for (int i = tid; i < n_anc * nsites; i += NT) {
    const int a = i / nsites, s = i - a * nsites;
    const size_t gi = (size_t)a * p.K + col_lo + s * Q;
    const int o = a * ldc + s * Q;
    float v[Q];
    if (p.resident & 1) { ld_row<Q>(th + o, v); st_row<Q>(p.theta + gi, v); }
    if (p.resident & 4) { ld_row<Q>(m1 + o, v); st_row<Q>(p.mom1 + gi, v); }
    if (p.resident & 8) { ld_row<Q>(m2 + o, v); st_row<Q>(p.mom2 + gi, v); }
  }
  if (p.argmax) {  // jnp.argmax(ancestors, axis=-1): first maximum (evals/benchmark.py:200)
    for (int i = tid; i < n_anc * nsites; i += NT) {
      const int a = i / nsites, s = i - a * nsites;
      float x[Q];
      ld_row<Q>(th + (size_t)a * ld_th + s * Q, x);
      int best = 0;
      float bv = x[0];
#pragma unroll
      for (int q = 1; q < Q; ++q)
        if (x[q] > bv) { bv = x[q]; best = q; }
      p.argmax[(size_t)a * p.L + site_lo + s] = best;
    }
  }
}

}  // namespace

size_t fit_fast_smem(int n_anc, int Q, int sites_per_cta, int tile_sites, int n_resident) {
  const int lda = (n_anc + 3) & ~3;
  size_t fl = (n_anc > kMaxAnc ? 0 : (size_t)n_anc * lda) + 2 * (size_t)n_anc * kEll + (size_t)n_anc * tile_sites * Q;
  fl += (size_t)n_resident * n_anc * sites_per_cta * Q;
  return fl * sizeof(float);
}

int launch_fit_fast(const FitParams& p, int grid, size_t smem, cudaStream_t s) {
  const void* fn = p.Q == 4 ? (const void*)fit_fast_kernel<4> : (const void*)fit_fast_kernel<20>;
  int rc = check_cuda(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                      "cudaFuncSetAttribute(fit_fast_kernel)");
  if (rc) return rc;
  void* args[] = {(void*)&p};
  rc = check_cuda(cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(NT), args, smem, s),
                  "cudaLaunchCooperativeKernel(fit_fast_kernel)");
  if (rc) return rc;
  note_launches(1);
  return TREX_OK;
}

}  // namespace fit
}  // namespace trex
