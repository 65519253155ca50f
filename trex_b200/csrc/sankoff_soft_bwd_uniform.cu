// Relaxed Sankoff reverse mode for a UNIFORM substitution cost (C[s][s] = 0, C[s][t] = a: Fitch-style
// parsimony, the unit cost 1 - I of the reference's tests and benchmarks) — SURVEY §8a row R2.
// Same math and outputs as sankoff_soft_bwd_warp.cu, but with K = exp(-C/T) = k + (1-k) I the two
// mat-vecs of an edge collapse to O(A):
//     e[t] = exp(-(d[t]-m)/T),  E = sum_t e[t],  Z[s] = (1-k) e[s] + k E,  r[s] = g[s] / Z[s],
//     R = sum_s r[s],           gbar[child][t] = e[t] ((1-k) r[t] + k R),
//     dC[s,t] += K[s,t] * sum_sites r[s] e[t]            (the only O(A^2) part, wanted or not by the caller).
//
// Every warp is autonomous (CTA = one warp, lanes = sites, V sites per lane) and walks the reverse
// post-order without any CTA-wide synchronisation.  Three tricks keep it near the HBM roofline:
//  * cotangent forwarding: in reverse post-order the next node is this node's last internal child,
//    so that child's gbar row is written to HBM (it is an output) AND left in the warp's shared
//    slab as the next node's cotangent — no read-back behind the warp's own store;
//  * dp rows never depend on this kernel, so the next child's row is prefetched into the registers
//    the current one just vacated, behind the outer-product sweep;
//  * children given as observed states contribute in "g space" (their softmin weight is one-hot, so
//    K r == g): both leaf children of a node share one sweep with e' = onehot(o0) + onehot(o1).
// The dC outer product is the pair-tile sweep of the warp kernel: lane = (pair tile, site group),
// TR x TC grid of RS x CS register tiles, 128-bit conflict-free shared loads, f32 partial sums
// flushed every kFlush nodes into a warp-private f64 table, f64 atomics once at the end.
#include <cstdlib>

#include "fwd_params.cuh"
#include "soft_bwd_params.cuh"

namespace trex {

namespace {

constexpr int kFlush = 16;
constexpr unsigned kFull = 0xffffffffu;

template <int AP, int V>
struct UniCfg {
  static constexpr int TR = 4, TC = 4;
  static constexpr int RS = AP / TR, CS = AP / TC;
  static constexpr int NT = TR * TC;
  static constexpr int SG = 32 / NT;
  static constexpr int NS = 32 * V;
  static constexpr int PITCH = NS + 4;
  static constexpr int NCH = NS / (4 * SG);
  static constexpr int kWarpFloats = 3 * AP * PITCH + 2 * AP * AP;  // G, R, E rows; f64 pair table
  static_assert(AP % TR == 0 && AP % TC == 0 && 32 % NT == 0 && NS % (4 * SG) == 0, "bad sweep tiling");
};

__device__ __forceinline__ float ex2_fast(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_fast(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// global -> shared copy of V floats without a register round trip; bytes past `src_bytes` are zero-filled
template <int V>
__device__ __forceinline__ void cp_async_v(float* smem_dst, const float* gsrc, int src_bytes) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], %2, %3;" ::"r"(d), "l"(gsrc), "n"(4 * V), "r"(src_bytes)
               : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// acc[a][b] += sum over this lane's site group of r[a][site] * e[b][site]
template <int AP, int V, int UNROLL>
__device__ __forceinline__ void pair_sweep(const float* __restrict__ rbase, const float* __restrict__ ebase,
                                           float (&acc)[UniCfg<AP, V>::RS][UniCfg<AP, V>::CS]) {
  using Cfg = UniCfg<AP, V>;
  constexpr int RS = Cfg::RS, CS = Cfg::CS, PITCH = Cfg::PITCH, SG = Cfg::SG;
#pragma unroll(UNROLL)
  for (int it = 0; it < Cfg::NCH; ++it) {
    float4 rr[RS], ee[CS];
#pragma unroll
    for (int a = 0; a < RS; ++a) rr[a] = *reinterpret_cast<const float4*>(rbase + a * PITCH + it * (SG * 4));
#pragma unroll
    for (int b = 0; b < CS; ++b) ee[b] = *reinterpret_cast<const float4*>(ebase + b * PITCH + it * (SG * 4));
#pragma unroll
    for (int a = 0; a < RS; ++a)
#pragma unroll
      for (int b = 0; b < CS; ++b) {
        float x = acc[a][b];
        x = fmaf(rr[a].x, ee[b].x, x);
        x = fmaf(rr[a].y, ee[b].y, x);
        x = fmaf(rr[a].z, ee[b].z, x);
        x = fmaf(rr[a].w, ee[b].w, x);
        acc[a][b] = x;
      }
  }
}

}  // namespace

// FULL: n_states == AP (no padded state rows, no per-row guards)
// WIDE: the long-alignment build of the dC kernel.  Registers are allocated per SM sub-partition (16 384 each), so
// with 9-12 one-warp CTAs per SM some sub-partition hosts three warps and ptxas caps the kernel at 168 registers —
// where this kernel spills and the dC sweep cannot be unrolled (every one of its 8 iterations then exposes a full
// LDS latency: 17 % of the stall samples, profiles/r02_g_soft_bwd_uniform_hot_lines.txt).  At 8 CTAs per SM the cap
// is 255: no spills, a 4x unrolled sweep whose loads run ahead of the FMAs, and room to keep the cotangent / r rows
// of the row math in registers.  Measured at 1024 taxa x 100k sites, A = 20 (profiles/r02_h_soft_bwd_variants.txt):
// 8.30 -> 6.74 ms; without dC the narrow build is the faster one (5.11 vs 5.56 ms: it needs no registers for the
// sweep), so WIDE is the dC kernel only.
template <int AP, int V, bool WANT_DC, bool FULL, bool WIDE = false>
__global__ void __launch_bounds__(32, AP > 8 ? (WIDE ? 8 : 11) : 16) soft_bwd_uniform_kernel(const SoftBwdParams p) {
  using Cfg = UniCfg<AP, V>;
  constexpr int kSweepUnroll = WIDE ? 4 : 1;
  constexpr int RS = Cfg::RS, CS = Cfg::CS, NT = Cfg::NT, PITCH = Cfg::PITCH;
  if (p.phase == 1 && (int)blockIdx.z >= (p.plan + (size_t)blockIdx.y * p.plan_words)[2]) return;  // empty segment slice (CTA-uniform): before any other work
  float a_off;
  if (!soft_bwd_uniform_applies(p, a_off)) return;  // the general kernels own every other cost matrix
  extern __shared__ __align__(16) float smem[];
  float* sG = smem;               // cotangent of the current node, [state][site]
  float* sR = sG + AP * PITCH;    // r = g / Z
  float* sE = sR + AP * PITCH;    // e (or the one-hot counts of the g-space sweep)
  double* sAcc = reinterpret_cast<double*>(sE + AP * PITCH);

  const int A = FULL ? AP : p.n_states;
  const int lane = threadIdx.x;
  const int tree = blockIdx.y;
  const int n_int = p.n_all - p.n_leaves;
  PlanView plan(p.plan + (size_t)tree * p.plan_words, n_int);
  const float c1 = 1.4426950408889634f / p.T;
  const float kk = exp2f(-a_off * c1);  // exp(-a/T): every off-diagonal entry of K
  const float km1 = 1.f - kk;
  const float row_k = 1.f + (float)(A - 1) * kk;  // row sum of K

  // Lanes past the end mirror the last valid site vector: their loads stay in bounds and finite,
  // their stores are predicated off and their cotangent / e are masked to zero.
  const long long site_raw = ((long long)blockIdx.x * 32 + lane) * V;
  const bool active = site_raw < p.n_sites;
  const long long site0 = active ? site_raw : ((p.n_sites - 1) / V) * V;
  bool ok[V];
#pragma unroll
  for (int v = 0; v < V; ++v) ok[v] = site_raw + v < p.n_sites;
  int ok_bytes = 0;
#pragma unroll
  for (int v = 0; v < V; ++v) ok_bytes += ok[v] ? 4 : 0;
  const float gs = p.gscale ? __ldg(p.gscale) : 1.f;
  const size_t stride = (size_t)p.stride;
  const float* dp = p.dp + (size_t)tree * n_int * A * stride + site0;
  float* gbar = p.gbar + (size_t)tree * n_int * A * stride + site0;
  float* dleaf = p.dleaf ? p.dleaf + (size_t)tree * p.n_leaves * A * stride + site0 : nullptr;
  float* dleaf_obs = p.dleaf_obs ? p.dleaf_obs + (size_t)tree * p.n_leaves * stride + site0 : nullptr;
  const float* leaf_cost = p.leaf_cost ? p.leaf_cost + site0 : nullptr;
  const Walk wk = make_walk(plan, p.phase, blockIdx.z);
  const int32_t* order = wk.ord;  // this launch walks order[n_reach-1] down to order[0]
  const int n_reach = wk.n_rev;
  if (n_reach == 0) return;
  const bool has_root = __ldg(order + n_reach - 1) == p.n_all - 1;  // the root is the last reachable node
  const int32_t* ch0 = plan.child(0);
  const int32_t* ch1 = plan.child(1);
  float* myG = sG + lane * V;  // this lane's column of the staged rows
  float* myR = sR + lane * V;
  float* myE = sE + lane * V;

  // pair-tile ownership in the sweeps
  const int tile = lane % NT, sg = lane / NT;
  const int rb = tile / Cfg::TC, cb = tile % Cfg::TC;
  const float* gbase = sG + rb * RS * PITCH + sg * 4;  // (the G and R slabs swap roles, see below)
  const float* rbase = sR + rb * RS * PITCH + sg * 4;
  const float* ebase = sE + cb * CS * PITCH + sg * 4;
  float acc[RS][CS], accG[RS][CS];  // r-space (times K at flush) and g-space pair sums
#pragma unroll
  for (int x = 0; x < RS; ++x)
#pragma unroll
    for (int y = 0; y < CS; ++y) acc[x][y] = accG[x][y] = 0.f;
  for (int i = lane; i < AP * AP; i += 32) sAcc[i] = 0.0;
  for (int i = lane; i < 3 * AP * PITCH; i += 32) sG[i] = 0.f;  // pad rows / columns stay zero
  __syncwarp();

  // rows of a child that lives in memory: a dp node, or a leaf of a float leaf-cost table
  auto row_src = [&](int c) -> const float* {
    if (c >= p.n_leaves) return dp + (size_t)(c - p.n_leaves) * A * stride;
    if (c >= 0 && leaf_cost) return leaf_cost + (size_t)c * A * stride;
    return nullptr;
  };
  // what the walk needs to know about a node; decoded one node ahead so that the plan lookups and
  // the leaf-state loads of node i-1 are in flight while node i is processed
  struct Desc {
    int k, cA, cB;      // children are taken "forwarded child last"
    bool fwdB;          // cB is the next node of the walk: its gbar row stays in sG as the next cotangent
    const float *srcA, *srcB;
    unsigned rawA, rawB;  // observed-state bytes of state children, still packed (unpacked where they are used)
  };
  auto decode = [&](int j, Desc& q) {
    const int node = __ldg(order + j);
    const int nq = j > 0 ? __ldg(order + j - 1) : -1;
    q.k = node - p.n_leaves;
    const int q0 = __ldg(ch0 + q.k), q1 = __ldg(ch1 + q.k);
    // in reverse post-order the next node is this node's last internal child
    const bool sw = q0 == nq && q0 >= p.n_leaves;
    q.cA = sw ? q1 : q0;
    q.cB = sw ? q0 : q1;
    q.fwdB = q.cB == nq && q.cB >= p.n_leaves;
    q.srcA = row_src(q.cA);
    q.srcB = row_src(q.cB);
    q.rawA = q.rawB = 0xffffffffu;  // "no state" in every byte (missing child / leaf without a state table)
    if (!q.srcA && q.cA >= 0 && p.leaf) q.rawA = ld_u8_raw<V>(p.leaf + (size_t)q.cA * stride + site0);
    if (!q.srcB && q.cB >= 0 && p.leaf) q.rawB = ld_u8_raw<V>(p.leaf + (size_t)q.cB * stride + site0);
  };

  // ---- nodes the root does not reach get zero cotangent; root: gscale * softmax(-dp/T) --------
  const bool have_g0 = has_root;  // sG holds the cotangent of the node about to be processed
  if (has_root) {  // (the walk that owns the root)
    float zero[V];
#pragma unroll
    for (int v = 0; v < V; ++v) zero[v] = 0.f;
    if (active) {
      for (int i = plan.hdr()[1]; i < n_int; ++i) {
        const int ku = __ldg(plan.order() + i) - p.n_leaves;
        float* row = gbar + (size_t)ku * A * stride;
        for (int s = 0; s < A; ++s) st_f<V>(row + s * stride, zero);
        for (int slot = 0; slot < 2; ++slot) {  // ... and so do the leaf-gradient rows of their leaf children
          const int c = __ldg(plan.child(slot) + ku);
          if (c < 0 || c >= p.n_leaves) continue;
          if (dleaf)
            for (int s = 0; s < A; ++s) st_f<V>(dleaf + ((size_t)c * A + s) * stride, zero);
          if (dleaf_obs) st_f<V>(dleaf_obs + (size_t)c * stride, zero);
        }
      }
    }
    const float* rrow = dp + (size_t)(n_int - 1) * A * stride;
    float* grow = gbar + (size_t)(n_int - 1) * A * stride;
    float mm[V], z[V];
#pragma unroll
    for (int v = 0; v < V; ++v) { mm[v] = INFINITY; z[v] = 0.f; }
    for (int s = 0; s < A; ++s) {
      float x[V];
      ld_f<V>(rrow + s * stride, x);
#pragma unroll
      for (int v = 0; v < V; ++v) mm[v] = fminf(mm[v], x[v]);
    }
    for (int s = 0; s < A; ++s) {
      float x[V];
      ld_f<V>(rrow + s * stride, x);
#pragma unroll
      for (int v = 0; v < V; ++v) z[v] += exp2f((mm[v] - x[v]) * c1);
    }
    for (int s = 0; s < A; ++s) {
      float x[V];
      ld_f<V>(rrow + s * stride, x);
#pragma unroll
      for (int v = 0; v < V; ++v) x[v] = exp2f((mm[v] - x[v]) * c1) * (gs / z[v]);
      if (active) st_f<V>(grow + s * stride, x);
#pragma unroll
      for (int v = 0; v < V; ++v) x[v] = ok[v] ? x[v] : 0.f;
      st_f<V>(myG + s * PITCH, x);
    }
  }

  bool have_g = have_g0;
  float d[AP][V];  // dp row of the next child with a cost row (prefetched), e in place once used
  const float* pf_src = nullptr;

  // One child with a cost row: reads d (prefetched, or loaded here), the cotangent from sG; writes the
  // child's gbar / dleaf row, stages r and e for the sweep, and prefetches `next_src` behind it.
  // Lanes past the end need no masking of e: their cotangent is zero, so r and every product are.
  auto row_child = [&](const float* src, float* dst, bool forward, const float* next_src) {
    if (pf_src != src) {
#pragma unroll
      for (int t = 0; t < AP; ++t)
        if (FULL || t < A) ld_f_stream<V>(src + (size_t)t * stride, d[t]);
    }
    // WIDE: the cotangent rows go to registers up front (all LDS in flight behind the exp math below) and r stays
    // there; the narrow build has no registers for that and round-trips r through the R slab, where every LDS
    // waits behind the previous STS (the G and R slabs swap roles, so the compiler cannot reorder them).
    float gr[WIDE ? AP : 1][V];
    if constexpr (WIDE) {
#pragma unroll
      for (int s = 0; s < AP; ++s) ld_f<V>(myG + s * PITCH, gr[s]);
    }
    float kE[V], R[V];
#pragma unroll
    for (int v = 0; v < V; ++v) {
      float mm = d[0][v];
#pragma unroll
      for (int t = 1; t < AP; ++t)
        if (FULL || t < A) mm = fminf(mm, d[t][v]);
      const float mc = mm * c1;
      float E0 = 0.f, E1 = 0.f;
#pragma unroll
      for (int t = 0; t < AP; ++t)
        if (FULL || t < A) {
          d[t][v] = ex2_fast(fmaf(d[t][v], -c1, mc));
          if (t & 1) E1 += d[t][v]; else E0 += d[t][v];
        }
      kE[v] = kk * (E0 + E1);
      R[v] = 0.f;
    }
    if constexpr (WIDE) {  // (WIDE implies FULL and WANT_DC)
#pragma unroll
      for (int s = 0; s < AP; ++s) {
#pragma unroll
        for (int v = 0; v < V; ++v) {
          gr[s][v] *= rcp_fast(fmaf(km1, d[s][v], kE[v]));  // r[s] = g[s] / Z[s],  Z >= k > 0
          R[v] += gr[s][v];
        }
        st_f<V>(myR + s * PITCH, gr[s]);
        st_f<V>(myE + s * PITCH, d[s]);
      }
#pragma unroll
      for (int v = 0; v < V; ++v) R[v] *= kk;
#pragma unroll
      for (int t = 0; t < AP; ++t) {
        float r[V];
#pragma unroll
        for (int v = 0; v < V; ++v) r[v] = d[t][v] * fmaf(km1, gr[t][v], R[v]);  // e[t] ((1-k) r[t] + k R)
        if (dst && active) st_f<V>(dst + (size_t)t * stride, r);
        if (forward) st_f<V>(myG + t * PITCH, r);
      }
    } else {
#pragma unroll
    for (int s = 0; s < AP; ++s)
      if (FULL || s < A) {
        float gv[V];
        ld_f<V>(myG + s * PITCH, gv);
#pragma unroll
        for (int v = 0; v < V; ++v) {
          gv[v] *= rcp_fast(fmaf(km1, d[s][v], kE[v]));  // r[s] = g[s] / Z[s],  Z >= k > 0
          R[v] += gv[v];
        }
        st_f<V>(myR + s * PITCH, gv);
        if (WANT_DC) st_f<V>(myE + s * PITCH, d[s]);
      }
#pragma unroll
    for (int v = 0; v < V; ++v) R[v] *= kk;
#pragma unroll
    for (int t = 0; t < AP; ++t)
      if (FULL || t < A) {
        float r[V];
        ld_f<V>(myR + t * PITCH, r);
#pragma unroll
        for (int v = 0; v < V; ++v) r[v] = d[t][v] * fmaf(km1, r[v], R[v]);  // e[t] ((1-k) r[t] + k R)
        if (dst && active) st_f<V>(dst + (size_t)t * stride, r);
        if (forward) st_f<V>(myG + t * PITCH, r);
      }
    }
    pf_src = next_src;
    if (next_src) {
#pragma unroll
      for (int t = 0; t < AP; ++t)
        if (FULL || t < A) ld_f_stream<V>(next_src + (size_t)t * stride, d[t]);
    }
    if (WANT_DC) {
      __syncwarp();
      pair_sweep<AP, V, kSweepUnroll>(rbase, ebase, acc);
      __syncwarp();
    }
  };

  Desc cur;
  if (n_reach > 0) decode(n_reach - 1, cur);
  bool g_pref = false;  // the R slab receives (cp.async) the cotangent of the next node
  int since_flush = 0;
  for (int i = n_reach - 1;; --i) {
    if constexpr (WANT_DC) {
      // f32 pair sums -> warp-private f64 table (K applied to the r-space part)
      if (i < 0 || since_flush == kFlush) {
        since_flush = 0;
#pragma unroll
        for (int x = 0; x < RS; ++x)
#pragma unroll
          for (int y = 0; y < CS; ++y) {
            float vr = acc[x][y], vg = accG[x][y];
            acc[x][y] = accG[x][y] = 0.f;
#pragma unroll
            for (int o = NT; o < 32; o <<= 1) {
              vr += __shfl_xor_sync(kFull, vr, o);
              vg += __shfl_xor_sync(kFull, vg, o);
            }
            if (sg == 0) {
              const int row = rb * RS + x, col = cb * CS + y;
              sAcc[row * AP + col] += (double)vr * (double)(row == col ? 1.f : kk) + (double)vg;
            }
          }
      }
    }
    if (i < 0) break;
    ++since_flush;
    Desc nx;
    const bool more = i > 0;
    if (more) decode(i - 1, nx);
    const float* after = more ? (nx.srcA ? nx.srcA : nx.srcB) : nullptr;  // next row the walk will ask for
    const int k = cur.k, cA = cur.cA, cB = cur.cB;
    const bool fwdB = cur.fwdB;
    const float* srcA = cur.srcA;
    const float* srcB = cur.srcB;

    if (!have_g) {
      if (g_pref) {
        // the cotangent was copied into the R slab behind the previous node: the slabs swap roles
        cp_async_wait_all();
        float* tf = myG; myG = myR; myR = tf;
        const float* tb = gbase; gbase = rbase; rbase = tb;
      } else {  // cotangent written earlier by this lane (zero-filled past the end)
        const float* grow = gbar + (size_t)k * A * stride;
#pragma unroll 4
        for (int s = 0; s < A; ++s) {
          float gv[V];
          ld_f<V>(grow + s * stride, gv);
#pragma unroll
          for (int v = 0; v < V; ++v) gv[v] = ok[v] ? gv[v] : 0.f;
          st_f<V>(myG + s * PITCH, gv);
        }
      }
    }
    g_pref = false;
    if (more && !fwdB && !srcA && !srcB) {
      // No child to forward (both children are states): the next node's cotangent, stored by this lane
      // when its parent was processed, is copied global -> R slab while this node is worked on.
      const float* nrow = gbar + (size_t)nx.k * A * stride;
#pragma unroll 4
      for (int s = 0; s < A; ++s) cp_async_v<V>(myR + s * PITCH, nrow + s * stride, ok_bytes);
      g_pref = true;
    }

    if (!srcA || !srcB) {
      // ---- children given as states (or missing): one g-space sweep for both -----------------
      int oA[V], oB[V];  // A: no state, -1: the child has a cost row (handled below, not here)
      bool nA = false, nB = false;
      unpack_u8<V>(cur.rawA, oA);
      unpack_u8<V>(cur.rawB, oB);
#pragma unroll
      for (int v = 0; v < V; ++v) {
        oA[v] = srcA ? -1 : min(oA[v], A);
        oB[v] = srcB ? -1 : min(oB[v], A);
        nA = nA || (oA[v] == A && ok[v]);
        nB = nB || (oB[v] == A && ok[v]);
      }
      const bool nsA = __any_sync(kFull, nA), nsB = __any_sync(kFull, nB);
      float tot[V];
#pragma unroll
      for (int v = 0; v < V; ++v) tot[v] = 0.f;
#pragma unroll
      for (int s = 0; s < AP; ++s)
        if (FULL || s < A) {
          float gv[V], cnt[V];
          ld_f<V>(myG + s * PITCH, gv);
#pragma unroll
          for (int v = 0; v < V; ++v) {
            tot[v] += gv[v];
            cnt[v] = (oA[v] == s ? 1.f : 0.f) + (oB[v] == s ? 1.f : 0.f);
          }
          if (WANT_DC) st_f<V>(myE + s * PITCH, cnt);
        }
      // d total / d leaf_cost of leaves with a state: all of the parent's mass on that state
      if (dleaf && active) {
#pragma unroll
        for (int slot = 0; slot < 2; ++slot) {
          const int c = slot ? cB : cA;
          if ((slot ? srcB : srcA) || c < 0) continue;
          float* dst = dleaf + (size_t)c * A * stride;
#pragma unroll 4
          for (int t = 0; t < A; ++t) {
            float out[V];
#pragma unroll
            for (int v = 0; v < V; ++v) out[v] = (slot ? oB[v] : oA[v]) == t ? tot[v] : 0.f;
            st_f_stream<V>(dst + (size_t)t * stride, out);
          }
        }
      }
      // compact form: one value per (leaf, site) = the mass on the leaf's row (its observed-state entry)
      if (dleaf_obs && active) {
#pragma unroll
        for (int slot = 0; slot < 2; ++slot) {
          const int c = slot ? cB : cA;
          if ((slot ? srcB : srcA) || c < 0) continue;
          st_f_stream<V>(dleaf_obs + (size_t)c * stride, tot);
        }
      }
      if (WANT_DC) {
        __syncwarp();
        pair_sweep<AP, V, kSweepUnroll>(gbase, ebase, accG);
        __syncwarp();
      }
      // rare: lanes whose child has no state (missing child, state out of range): all-1e5 child row,
      // e uniform, softmin weights K / rowsum(K)  (r-space)
#pragma unroll 1
      for (int slot = 0; slot < 2; ++slot) {
        if (!(slot ? nsB : nsA)) continue;
        if (g_pref) {  // the R slab is needed here: give the prefetched cotangent up (re-read at the next node)
          cp_async_wait_all();
          g_pref = false;
        }
        const int c = slot ? cB : cA;
        bool ns[V];
        bool any = false;
        float R[V];
#pragma unroll
        for (int v = 0; v < V; ++v) {
          ns[v] = (slot ? oB[v] : oA[v]) == A && ok[v];
          any = any || ns[v];
          R[v] = 0.f;
        }
#pragma unroll 1
        for (int s = 0; s < A; ++s) {
          float gv[V], one[V];
          ld_f<V>(myG + s * PITCH, gv);
#pragma unroll
          for (int v = 0; v < V; ++v) {
            gv[v] = ns[v] ? __fdividef(gv[v], row_k) : 0.f;
            one[v] = ns[v] ? 1.f : 0.f;
            R[v] += gv[v];
          }
          st_f<V>(myR + s * PITCH, gv);
          st_f<V>(myE + s * PITCH, one);
        }
        if (c >= 0 && dleaf && any && active) {
          float* dst = dleaf + (size_t)c * A * stride;
#pragma unroll 1
          for (int t = 0; t < A; ++t) {
            float out[V], r[V];
            ld_f<V>(dst + (size_t)t * stride, out);  // lanes with a state keep what was stored above
            ld_f<V>(myR + t * PITCH, r);
#pragma unroll
            for (int v = 0; v < V; ++v)
              if (ns[v]) out[v] = fmaf(km1, r[v], kk * R[v]);
            st_f<V>(dst + (size_t)t * stride, out);
          }
        }
        if (WANT_DC) {
          __syncwarp();
          pair_sweep<AP, V, kSweepUnroll>(rbase, ebase, acc);
          __syncwarp();
        }
      }
    }

    // ---- children with a cost row (internal nodes, soft leaves) --------------------------------
    auto dst_of = [&](int c) -> float* {
      if (c >= p.n_leaves) return gbar + (size_t)(c - p.n_leaves) * A * stride;
      return dleaf ? dleaf + (size_t)c * A * stride : nullptr;
    };
#pragma unroll 1
    for (int slot = 0; slot < 2; ++slot) {  // one code site for both children
      const float* src = slot ? srcB : srcA;
      if (!src) continue;
      row_child(src, dst_of(slot ? cB : cA), slot && fwdB, (slot == 0 && srcB) ? srcB : after);
    }
    have_g = fwdB;
    cur = nx;
  }
  if constexpr (WANT_DC) {
    __syncwarp();
    const bool not_tree = (plan.hdr()[3] & (TREX_PLAN_NOT_A_TREE | TREX_PLAN_SHARED_LEAF)) != 0;
    for (int i = lane; i < A * A; i += 32) {
      const double v = not_tree ? (double)NAN : sAcc[(i / A) * AP + i % A];  // reverse mode is defined on trees only
      if (v != 0.0) atomicAdd(p.dcost + (size_t)tree * A * A + i, v);
    }
  }
}

namespace {
template <int AP, int V>
int launch_uniform(const SoftBwdParams& p0, int batch, cudaStream_t s) {
  using Cfg = UniCfg<AP, V>;
  const size_t smem = sizeof(float) * (size_t)Cfg::kWarpFloats;
  const unsigned gx = (unsigned)((p0.n_sites + Cfg::NS - 1) / Cfg::NS);
  const bool segs = use_segments(p0.n_leaves);
  auto go = [&](auto kern) -> cudaError_t {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    SoftBwdParams p = p0;
    if (segs) {
      p.phase = 2;  // the nodes above the segments first (root downwards) ...
      kern<<<dim3(gx, batch), 32, smem, s>>>(p);
      p.phase = 1;  // ... then the independent subtrees side by side
      kern<<<dim3(gx, batch, kMaxSeg), 32, smem, s>>>(p);
      note_launches(2);
    } else {
      kern<<<dim3(gx, batch), 32, smem, s>>>(p);
      note_launches(1);
    }
    return e;
  };
  const SoftBwdParams& p = p0;
  const bool full = p.n_states == AP;
  // the 255-register build of the dC kernel: faster at every alignment length measured once its row math keeps the
  // cotangent rows in registers (12.5k sites 1.16 -> 1.01 ms, 25k 2.27 -> 1.86, 100k 8.27 -> 6.74;
  // profiles/r02_j_probe_step.txt)
  bool wide = AP > 8 && full;
  if (const char* e = getenv("TREX_BWD_WIDE")) {  // tests: force the choice (the builds are parity-tested against each other's oracle)
    if (e[0] == '0' || e[0] == '1') wide = AP > 8 && full && e[0] == '1';
  }
  cudaError_t e;
  if (p.dcost && wide) e = go(soft_bwd_uniform_kernel<AP, V, true, true, true>);
  else if (p.dcost) e = full ? go(soft_bwd_uniform_kernel<AP, V, true, true>) : go(soft_bwd_uniform_kernel<AP, V, true, false>);
  else e = full ? go(soft_bwd_uniform_kernel<AP, V, false, true>) : go(soft_bwd_uniform_kernel<AP, V, false, false>);
  if (e != cudaSuccess) return check_cuda(e, "cudaFuncSetAttribute(soft_bwd_uniform)");
  return check_cuda(cudaPeekAtLastError(), "soft_bwd_uniform_kernel");
}
}  // namespace

// A <= 20 only (as the warp kernel); returns at once on the device unless the cost is uniform
int launch_soft_bwd_uniform(const SoftBwdParams& p, int batch, cudaStream_t s) {
  const int A = p.n_states;
  if (A <= 4) return launch_uniform<4, 4>(p, batch, s);
  if (A <= 8) return launch_uniform<8, 2>(p, batch, s);
  return launch_uniform<20, 2>(p, batch, s);
}

}  // namespace trex
