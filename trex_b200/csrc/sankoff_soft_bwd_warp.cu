// Relaxed Sankoff reverse mode, exp-domain fast path — SURVEY §8a row R2 (no reference code
// exists).  Same math as sankoff_soft_bwd.cu:
//     e[t] = exp(-(d[t]-m)/T),  Z[s] = sum_t K[s,t] e[t],  r[s] = g[s]/Z[s],
//     gbar[child][t] = e[t] * sum_s K[s,t] r[s],   dC[s,t] += K[s,t] * r[s] * e[t].
//
// Every WARP is autonomous: it owns 32*V consecutive sites, walks the whole reverse post-order
// and never synchronises with another warp, so work balances at warp granularity over the SMs
// (c5: 1563 warps over 148 SMs, all resident) and there is no __syncthreads in the tree walk.
//
// Per edge the warp runs two phases over its own shared-memory slab:
//  1. lanes = sites: exp-domain mat-vecs with K / K^T rows broadcast from shared memory; r and e
//     are staged [state][site] (the cotangent row g and r never live in registers).
//  2. the dC outer product, re-partitioned: lane = (pair tile, site group); the AP x AP pairs
//     form a TR x TC grid of RS x CS register tiles, each lane sweeps its site group with
//     128-bit conflict-free loads (RS+CS LDS.128 per 4*RS*CS FFMA).
// Children given as observed states (hard leaves) contribute in "g space": their softmin weight
// is one-hot, so K*r == g and both leaf children of a node share ONE sweep with
// e' = onehot(o0) + onehot(o1).  Pair sums are flushed every kFlush nodes into a warp-private
// f64 table in shared memory and leave the SM once, as f64 atomics, at the end.
#include "soft_bwd_params.cuh"

namespace trex {

namespace {

constexpr int kFlush = 16;  // nodes between f32 -> f64 flushes of the register pair sums
constexpr unsigned kFull = 0xffffffffu;

template <int AP> struct SweepGrid { static constexpr int TR = 4, TC = 4; };

template <int AP, int V>
struct WarpCfg {
  static constexpr int TR = SweepGrid<AP>::TR, TC = SweepGrid<AP>::TC;
  static constexpr int RS = AP / TR, CS = AP / TC;  // register tile of pair sums
  static constexpr int NT = TR * TC;                // tiles == lanes per site group
  static constexpr int SG = 32 / NT;                // site groups
  static constexpr int NS = 32 * V;                 // sites per warp
  static constexpr int PITCH = NS + 4;              // floats per staged row (bank-conflict-free)
  static constexpr int NCH = NS / (4 * SG);         // float4 chunks per lane and sweep
  static constexpr int kSharedFloats = 2 * AP * AP + ((AP + 3) / 4) * 4;  // K, K^T, rowsum(K)
  static constexpr int kWarpFloats = 2 * AP * PITCH + 2 * AP * AP;        // R, E rows; f64 pair table
  static_assert(AP % TR == 0 && AP % TC == 0 && 32 % NT == 0 && NS % (4 * SG) == 0, "bad sweep tiling");
};

template <int V>
__device__ __forceinline__ void sts_v(float* p, const float (&v)[V]) {
  st_f<V>(p, v);
}

// single-MUFU forms: results below 2^-126 flush to zero, which is what the callers want
__device__ __forceinline__ float ex2_fast(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// global -> shared copy of V floats without a register round trip; bytes past `src_bytes` are
// zero-filled (masks the cotangent of lanes past the end)
template <int V>
__device__ __forceinline__ void cp_async_v(float* smem_dst, const float* gsrc, int src_bytes) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], %2, %3;" ::"r"(d), "l"(gsrc), "n"(4 * V), "r"(src_bytes)
               : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
__device__ __forceinline__ float rcp_fast(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

}  // namespace

template <int AP, int V, int W, bool WANT_DC>
__global__ void __launch_bounds__(32 * W, (AP > 8 ? 11 : 16) / W) soft_bwd_warp_kernel(const SoftBwdParams p) {
  using Cfg = WarpCfg<AP, V>;
  constexpr int RS = Cfg::RS, CS = Cfg::CS, NT = Cfg::NT, SG = Cfg::SG, PITCH = Cfg::PITCH;
  extern __shared__ __align__(16) float smem[];
  float* sK = smem;            // [s][t]
  float* sKt = sK + AP * AP;   // [t][s]
  float* sRowK = sKt + AP * AP;

  const int A = p.n_states;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tree = blockIdx.y;
  const int n_int = p.n_all - p.n_leaves;
  PlanView plan(p.plan + (size_t)tree * p.plan_words, n_int);
  const float c1 = 1.4426950408889634f / p.T;

  if (p.phase == 1 && (int)blockIdx.z >= plan.hdr()[2]) return;  // empty segment slice: skip the prologue (CTA-uniform)
  {
    float a_uni;
    if (soft_bwd_uniform_applies(p, a_uni)) return;  // soft_bwd_uniform_kernel owns this call
  }
  // ---- prologue (whole CTA): K = exp(-C/T), its transpose and row sums ----------------------
  bool c_ok = true;
  for (int i = tid; i < AP * AP; i += 32 * W) {
    const int s = i / AP, t = i % AP;
    float k = 0.f;
    if (s < A && t < A) {
      const float c = __ldg(p.cost + s * A + t);
      c_ok = c_ok && expdom_ok(c, p.T);
      k = exp2f(-c * c1);
    }
    sK[i] = k;
    sKt[t * AP + s] = k;
  }
  if (!__syncthreads_and(c_ok)) return;  // direct-evaluation kernel (sankoff_soft_bwd.cu) handles it
  for (int s = tid; s < AP; s += 32 * W) {
    float z = 0.f;
    for (int t = 0; t < AP; ++t) z += sK[s * AP + t];
    sRowK[s] = z;
  }
  __syncthreads();

  const long long gw = (long long)blockIdx.x * W + warp;
  if (gw * Cfg::NS >= p.n_sites) return;  // no __syncthreads below this line
  float* sR = smem + Cfg::kSharedFloats + warp * Cfg::kWarpFloats;
  float* sE = sR + AP * PITCH;
  double* sAcc = reinterpret_cast<double*>(sE + AP * PITCH);


  // Lanes past the end mirror the last valid site vector: their loads stay in bounds and finite,
  // their stores are predicated off and their cotangent is masked to zero (okf).
  const long long site_raw = (gw * 32 + lane) * V;
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
  const Walk wk = make_walk(plan, p.phase, blockIdx.z);
  const int32_t* order = wk.ord;  // this launch walks order[n_reach-1] down to order[0]
  const int n_reach = wk.n_rev;
  if (n_reach == 0) return;
  const bool has_root = __ldg(order + n_reach - 1) == p.n_all - 1;  // the root is the last reachable node
  const int32_t* ch0 = plan.child(0);
  const int32_t* ch1 = plan.child(1);
  float* myR = sR + lane * V;  // this lane's column of the staged rows
  float* myE = sE + lane * V;

  // pair-tile ownership in the sweep
  const int tile = lane % NT, sg = lane / NT;
  const int rb = tile / Cfg::TC, cb = tile % Cfg::TC;
  const float* rbase = sR + rb * RS * PITCH + sg * 4;
  const float* ebase = sE + cb * CS * PITCH + sg * 4;
  float acc[RS][CS], accG[RS][CS];  // r-space (times K at flush) and g-space pair sums
#pragma unroll
  for (int a = 0; a < RS; ++a)
#pragma unroll
    for (int b = 0; b < CS; ++b) acc[a][b] = accG[a][b] = 0.f;
  if constexpr (WANT_DC) {
    for (int i = lane; i < AP * AP; i += 32) sAcc[i] = 0.0;
    for (int i = lane; i < 2 * AP * PITCH; i += 32) sR[i] = 0.f;  // pad rows / columns stay zero
    __syncwarp();
  }

  // ---- nodes the root does not reach get zero cotangent; root: gscale * softmax(-dp/T) --------
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
    }
  }

  int since_flush = 0;
  for (int i = n_reach - 1;; --i) {
    if constexpr (WANT_DC) {
      // f32 pair sums -> warp-private f64 table (K applied to the r-space part); single code site
      if (i < 0 || since_flush == kFlush) {
        since_flush = 0;
#pragma unroll
        for (int a = 0; a < RS; ++a)
#pragma unroll
          for (int b = 0; b < CS; ++b) {
            float vr = acc[a][b], vg = accG[a][b];
            acc[a][b] = accG[a][b] = 0.f;
#pragma unroll
            for (int o = NT; o < 32; o <<= 1) {
              vr += __shfl_xor_sync(kFull, vr, o);
              vg += __shfl_xor_sync(kFull, vg, o);
            }
            if (sg == 0) {
              const int idx = (rb * RS + a) * AP + cb * CS + b;
              sAcc[idx] += (double)vr * (double)sK[idx] + (double)vg;
            }
          }
      }
    }
    if (i < 0) break;
    ++since_flush;
    const int node = __ldg(order + i);
    const int k = node - p.n_leaves;
    const float* grow = gbar + (size_t)k * A * stride;  // cotangent of this node (own earlier store)
    const int cc0 = __ldg(ch0 + k), cc1 = __ldg(ch1 + k);
    const bool gen0 = cc0 >= p.n_leaves || (cc0 >= 0 && p.leaf_cost != nullptr);
    const bool gen1 = cc1 >= p.n_leaves || (cc1 >= 0 && p.leaf_cost != nullptr);
    int o0[V], o1[V];  // observed states of children handled in g space (A: none, -1: not handled)
    bool ns0 = false, ns1 = false;

    // phases: 0 = children given as states (one g-space sweep for both); 1,2 = stateless leaves /
    // missing children of slot 0,1 (rare); 3,4 = children with a cost row.  One sweep site.
#pragma unroll 1
    for (int ph = 0; ph < 5; ++ph) {
      bool gspace = false;
      if (ph == 0) {
        if (gen0 && gen1) { ph = 2; continue; }
        gspace = true;
#pragma unroll
        for (int v = 0; v < V; ++v) { o0[v] = gen0 ? -1 : A; o1[v] = gen1 ? -1 : A; }
        if (!gen0 && cc0 >= 0) ld_u8<V>(p.leaf + (size_t)cc0 * stride + site0, o0);
        if (!gen1 && cc1 >= 0) ld_u8<V>(p.leaf + (size_t)cc1 * stride + site0, o1);
        bool n0 = false, n1 = false;
#pragma unroll
        for (int v = 0; v < V; ++v) {
          o0[v] = min(o0[v], A);
          o1[v] = min(o1[v], A);
          n0 = n0 || (o0[v] == A && ok[v]);
          n1 = n1 || (o1[v] == A && ok[v]);
        }
        ns0 = __any_sync(kFull, n0);
        ns1 = __any_sync(kFull, n1);
        float tot[V];
#pragma unroll
        for (int v = 0; v < V; ++v) tot[v] = 0.f;
#pragma unroll 4
        for (int s = 0; s < A; ++s) {
          float gv[V], cnt[V];
          ld_f<V>(grow + s * stride, gv);
#pragma unroll
          for (int v = 0; v < V; ++v) {
            gv[v] = ok[v] ? gv[v] : 0.f;
            tot[v] += gv[v];
            cnt[v] = (o0[v] == s ? 1.f : 0.f) + (o1[v] == s ? 1.f : 0.f);
          }
          if constexpr (WANT_DC) {
            sts_v<V>(myR + s * PITCH, gv);
            sts_v<V>(myE + s * PITCH, cnt);
          }
        }
        // d total / d leaf_cost of leaves with a state: all of the parent's mass on that state
        if (dleaf && active) {
#pragma unroll
          for (int slot = 0; slot < 2; ++slot) {
            const int c = slot ? cc1 : cc0;
            if ((slot ? gen1 : gen0) || c < 0) continue;
            float* dst = dleaf + (size_t)c * A * stride;
            for (int t = 0; t < A; ++t) {
              float out[V];
#pragma unroll
              for (int v = 0; v < V; ++v) out[v] = (slot ? o1[v] : o0[v]) == t ? tot[v] : 0.f;
              st_f_stream<V>(dst + t * stride, out);
            }
          }
        }
        // compact form: one value per (leaf, site) = the mass on the leaf's row (its observed-state entry)
        if (dleaf_obs && active) {
#pragma unroll
          for (int slot = 0; slot < 2; ++slot) {
            const int c = slot ? cc1 : cc0;
            if ((slot ? gen1 : gen0) || c < 0) continue;
            st_f_stream<V>(dleaf_obs + (size_t)c * stride, tot);
          }
        }
      } else if (ph < 3) {
        // all-1e5 child row: uniform e, softmin weights K / rowsum(K)  (r-space)
        const int slot = ph - 1;
        if (!(slot ? ns1 : ns0)) continue;
        const int c = slot ? cc1 : cc0;
        bool ns[V];
        bool any = false;
#pragma unroll
        for (int v = 0; v < V; ++v) { ns[v] = (slot ? o1[v] : o0[v]) == A && ok[v]; any = any || ns[v]; }
        for (int s = 0; s < A; ++s) {
          float gv[V], one[V];
          ld_f<V>(grow + s * stride, gv);
#pragma unroll
          for (int v = 0; v < V; ++v) {
            gv[v] = ns[v] ? __fdividef(gv[v], sRowK[s]) : 0.f;
            one[v] = ns[v] ? 1.f : 0.f;
          }
          sts_v<V>(myR + s * PITCH, gv);
          sts_v<V>(myE + s * PITCH, one);
        }
        if (c >= 0 && dleaf && any && active) {
          float* dst = dleaf + (size_t)c * A * stride;
          for (int t = 0; t < A; ++t) {
            float out[V];
            ld_f<V>(dst + t * stride, out);  // lanes with a state keep what phase 0 stored
#pragma unroll
            for (int v = 0; v < V; ++v) {
              if (!ns[v]) continue;
              float a = 0.f;
              for (int s = 0; s < A; ++s) a = fmaf(sKt[t * AP + s], myR[s * PITCH + v], a);
              out[v] = a;
            }
            st_f<V>(dst + t * stride, out);
          }
        }
      } else {
        // ---- child with a cost row (internal node or soft leaf) ---------------------------------
        const int slot = ph - 3;
        if (!(slot ? gen1 : gen0)) continue;
        const int c = slot ? cc1 : cc0;
        const float* src;
        float* dst;
        if (c >= p.n_leaves) {
          src = dp + (size_t)(c - p.n_leaves) * A * stride;
          dst = gbar + (size_t)(c - p.n_leaves) * A * stride;
        } else {
          src = p.leaf_cost + (size_t)c * A * stride + site0;
          dst = dleaf ? dleaf + (size_t)c * A * stride : nullptr;
        }
        float e[AP][V];
#pragma unroll
        for (int t = 0; t < AP; ++t) {
          if (t < A) ld_f_stream<V>(src + t * stride, e[t]);
          else {
#pragma unroll
            for (int v = 0; v < V; ++v) e[t][v] = INFINITY;
          }
        }
        // cotangent rows go straight to this lane's column of R (zero-filled past the end)
        for (int s = 0; s < A; ++s) cp_async_v<V>(myR + s * PITCH, grow + s * stride, ok_bytes);
#pragma unroll
        for (int v = 0; v < V; ++v) {
          float mm = e[0][v];
#pragma unroll
          for (int t = 1; t < AP; ++t) mm = fminf(mm, e[t][v]);
#pragma unroll
          for (int t = 0; t < AP; ++t) e[t][v] = ok[v] ? ex2_fast((mm - e[t][v]) * c1) : 0.f;
        }
#pragma unroll
        for (int t = 0; t < AP; ++t) sts_v<V>(myE + t * PITCH, e[t]);
        cp_async_wait_all();
        // r[s] = g[s] / Z[s], four rows of K at a time (8 independent FFMA chains at V = 2)
#pragma unroll 1
        for (int sb = 0; sb < AP; sb += 4) {
          float z[4][V];
#pragma unroll
          for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int v = 0; v < V; ++v) z[r][v] = 0.f;
          const float* krow = sK + sb * AP;
#pragma unroll
          for (int t = 0; t < AP; t += 4) {
            float4 k4[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) k4[r] = *reinterpret_cast<const float4*>(krow + r * AP + t);
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
              for (int v = 0; v < V; ++v) {
                z[r][v] = fmaf(k4[r].x, e[t][v], z[r][v]);
                z[r][v] = fmaf(k4[r].y, e[t + 1][v], z[r][v]);
                z[r][v] = fmaf(k4[r].z, e[t + 2][v], z[r][v]);
                z[r][v] = fmaf(k4[r].w, e[t + 3][v], z[r][v]);
              }
          }
#pragma unroll
          for (int r = 0; r < 4; ++r) {
            float gv[V];
            ld_f<V>(myR + (sb + r) * PITCH, gv);
#pragma unroll
            for (int v = 0; v < V; ++v) gv[v] *= rcp_fast(fmaxf(z[r][v], 1e-37f));
            sts_v<V>(myR + (sb + r) * PITCH, gv);
          }
        }
        // gbar[child][t] = e[t] * sum_s K[s,t] r[s]; r comes back from this lane's own column
        {
          float r[AP][V];
#pragma unroll
          for (int s = 0; s < AP; ++s) ld_f<V>(myR + s * PITCH, r[s]);
#pragma unroll 1
          for (int tb = 0; tb < AP; tb += 4) {
            float a[4][V];
#pragma unroll
            for (int q = 0; q < 4; ++q)
#pragma unroll
              for (int v = 0; v < V; ++v) a[q][v] = 0.f;
            const float* ktrow = sKt + tb * AP;
#pragma unroll
            for (int s = 0; s < AP; s += 4) {
              float4 k4[4];
#pragma unroll
              for (int q = 0; q < 4; ++q) k4[q] = *reinterpret_cast<const float4*>(ktrow + q * AP + s);
#pragma unroll
              for (int q = 0; q < 4; ++q)
#pragma unroll
                for (int v = 0; v < V; ++v) {
                  a[q][v] = fmaf(k4[q].x, r[s][v], a[q][v]);
                  a[q][v] = fmaf(k4[q].y, r[s + 1][v], a[q][v]);
                  a[q][v] = fmaf(k4[q].z, r[s + 2][v], a[q][v]);
                  a[q][v] = fmaf(k4[q].w, r[s + 3][v], a[q][v]);
                }
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              float et[V];
              ld_f<V>(myE + (tb + q) * PITCH, et);
#pragma unroll
              for (int v = 0; v < V; ++v) a[q][v] *= et[v];
              if (dst && active && tb + q < A) st_f<V>(dst + (size_t)(tb + q) * stride, a[q]);
            }
          }
        }
      }
      if constexpr (WANT_DC) {
        // dC outer product over this warp's sites: lane = (pair tile, site group)
        __syncwarp();
        float tmp[RS][CS];
#pragma unroll
        for (int a = 0; a < RS; ++a)
#pragma unroll
          for (int b = 0; b < CS; ++b) tmp[a][b] = 0.f;
#pragma unroll 2
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
              float x = tmp[a][b];
              x = fmaf(rr[a].x, ee[b].x, x);
              x = fmaf(rr[a].y, ee[b].y, x);
              x = fmaf(rr[a].z, ee[b].z, x);
              x = fmaf(rr[a].w, ee[b].w, x);
              tmp[a][b] = x;
            }
        }
        if (gspace) {
#pragma unroll
          for (int a = 0; a < RS; ++a)
#pragma unroll
            for (int b = 0; b < CS; ++b) accG[a][b] += tmp[a][b];
        } else {
#pragma unroll
          for (int a = 0; a < RS; ++a)
#pragma unroll
            for (int b = 0; b < CS; ++b) acc[a][b] += tmp[a][b];
        }
        __syncwarp();
      }
    }
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

template <int AP, int V, int W>
static int launch_warp(const SoftBwdParams& p, int batch, cudaStream_t s) {
  using Cfg = WarpCfg<AP, V>;
  const size_t smem = sizeof(float) * (Cfg::kSharedFloats + (size_t)W * Cfg::kWarpFloats);
  const long long nwarps = (p.n_sites + Cfg::NS - 1) / Cfg::NS;
  const unsigned gx = (unsigned)((nwarps + W - 1) / W);
  const bool segs = use_segments(p.n_leaves);
  auto go = [&](auto kern) -> cudaError_t {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    SoftBwdParams q = p;
    if (segs) {
      q.phase = 2;  // the nodes above the segments first (root downwards) ...
      kern<<<dim3(gx, batch), 32 * W, smem, s>>>(q);
      q.phase = 1;  // ... then the independent subtrees side by side
      kern<<<dim3(gx, batch, kMaxSeg), 32 * W, smem, s>>>(q);
      note_launches(2);
    } else {
      kern<<<dim3(gx, batch), 32 * W, smem, s>>>(q);
      note_launches(1);
    }
    return e;
  };
  const cudaError_t e = p.dcost ? go(soft_bwd_warp_kernel<AP, V, W, true>) : go(soft_bwd_warp_kernel<AP, V, W, false>);
  if (e != cudaSuccess) return check_cuda(e, "cudaFuncSetAttribute(soft_bwd_warp)");
  return check_cuda(cudaPeekAtLastError(), "soft_bwd_warp_kernel");
}

// A <= 20 only (larger alphabets stay on the CTA-synchronous kernel)
int launch_soft_bwd_warp(const SoftBwdParams& p, int batch, cudaStream_t s) {
  const int A = p.n_states;
  if (A <= 4) return launch_warp<4, 4, 2>(p, batch, s);
  if (A <= 8) return launch_warp<8, 2, 2>(p, batch, s);
  return launch_warp<20, 2, 1>(p, batch, s);
}

}  // namespace trex
