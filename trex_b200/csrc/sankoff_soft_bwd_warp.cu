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

  const long long site0 = (gw * 32 + lane) * V;
  const bool active = site0 < p.n_sites;
  bool ok[V];
#pragma unroll
  for (int v = 0; v < V; ++v) ok[v] = site0 + v < p.n_sites;
  const float gs = p.gscale ? __ldg(p.gscale) : 1.f;
  const float* dp = p.dp + (size_t)tree * n_int * A * p.stride + site0;
  float* gbar = p.gbar + (size_t)tree * n_int * A * p.stride + site0;
  float* dleaf = p.dleaf ? p.dleaf + (size_t)tree * p.n_leaves * A * p.stride + site0 : nullptr;
  const int32_t* order = plan.order();
  const int32_t* ch0 = plan.child(0);
  const int32_t* ch1 = plan.child(1);
  const int n_reach = plan.hdr()[1];
  float* myR = sR + lane * V;  // this lane's column of the staged rows
  float* myE = sE + lane * V;

  // pair-tile ownership in the sweep
  const int tile = lane % NT, sg = lane / NT;
  const int rb = tile / Cfg::TC, cb = tile % Cfg::TC;
  float acc[RS][CS], accG[RS][CS];  // r-space (times K at flush) and g-space pair sums
#pragma unroll
  for (int a = 0; a < RS; ++a)
#pragma unroll
    for (int b = 0; b < CS; ++b) acc[a][b] = accG[a][b] = 0.f;
  if constexpr (WANT_DC) {
    for (int i = lane; i < AP * AP; i += 32) sAcc[i] = 0.0;
  }

  auto sweep = [&](float (&dst)[RS][CS]) {
    const float* rbase = sR + rb * RS * PITCH + sg * 4;
    const float* ebase = sE + cb * CS * PITCH + sg * 4;
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
          float x = dst[a][b];
          x = fmaf(rr[a].x, ee[b].x, x);
          x = fmaf(rr[a].y, ee[b].y, x);
          x = fmaf(rr[a].z, ee[b].z, x);
          x = fmaf(rr[a].w, ee[b].w, x);
          dst[a][b] = x;
        }
    }
  };

  auto flush = [&](float (&src)[RS][CS], bool rspace) {
#pragma unroll
    for (int a = 0; a < RS; ++a)
#pragma unroll
      for (int b = 0; b < CS; ++b) {
        float v = src[a][b];
        src[a][b] = 0.f;
#pragma unroll
        for (int o = NT; o < 32; o <<= 1) v += __shfl_xor_sync(kFull, v, o);
        if (sg == 0) {
          const int idx = (rb * RS + a) * AP + cb * CS + b;
          sAcc[idx] += (double)v * (rspace ? (double)sK[idx] : 1.0);
        }
      }
  };

  // ---- nodes the root does not reach get zero cotangent; root: gscale * softmax(-dp/T) --------
  if (active) {
    float zero[V];
#pragma unroll
    for (int v = 0; v < V; ++v) zero[v] = 0.f;
    for (int i = n_reach; i < n_int; ++i) {
      float* row = gbar + (size_t)(__ldg(order + i) - p.n_leaves) * A * p.stride;
      for (int s = 0; s < A; ++s) st_f<V>(row + (size_t)s * p.stride, zero);
    }
    float x[AP][V];
    const float* rrow = dp + (size_t)(n_int - 1) * A * p.stride;
#pragma unroll
    for (int s = 0; s < AP; ++s) {
      if (s < A) ld_f<V>(rrow + (size_t)s * p.stride, x[s]);
      else {
#pragma unroll
        for (int v = 0; v < V; ++v) x[s][v] = INFINITY;
      }
    }
#pragma unroll
    for (int v = 0; v < V; ++v) {
      float mm = x[0][v];
#pragma unroll
      for (int s = 1; s < AP; ++s) mm = fminf(mm, x[s][v]);
      float z = 0.f;
#pragma unroll
      for (int s = 0; s < AP; ++s) { x[s][v] = exp2f((mm - x[s][v]) * c1); z += x[s][v]; }
      const float inv = gs / z;
#pragma unroll
      for (int s = 0; s < AP; ++s) x[s][v] *= inv;
    }
    float* grow = gbar + (size_t)(n_int - 1) * A * p.stride;
#pragma unroll
    for (int s = 0; s < AP; ++s)
      if (s < A) st_f<V>(grow + (size_t)s * p.stride, x[s]);
  }

  int since_flush = 0;
  for (int i = n_reach - 1; i >= 0; --i) {
    const int node = __ldg(order + i);
    const int k = node - p.n_leaves;
    const float* grow = gbar + (size_t)k * A * p.stride;  // cotangent of this node (own earlier store)
    int cc[2] = {__ldg(ch0 + k), __ldg(ch1 + k)};
    bool gen[2];
#pragma unroll
    for (int slot = 0; slot < 2; ++slot)
      gen[slot] = cc[slot] >= p.n_leaves || (cc[slot] >= 0 && p.leaf_cost != nullptr);

    // ---- children given as observed states (or missing): one g-space sweep for both ----------
    if (!gen[0] || !gen[1]) {
      int o[2][V];  // observed state; A = leaf without state / missing child; -1 = slot not handled here
      bool nostate = false;
#pragma unroll
      for (int slot = 0; slot < 2; ++slot) {
#pragma unroll
        for (int v = 0; v < V; ++v) o[slot][v] = gen[slot] ? -1 : A;
        if (!gen[slot] && cc[slot] >= 0 && active) {
          ld_u8<V>(p.leaf + (size_t)cc[slot] * p.stride + site0, o[slot]);
#pragma unroll
          for (int v = 0; v < V; ++v) o[slot][v] = min(o[slot][v], A);
        }
#pragma unroll
        for (int v = 0; v < V; ++v) nostate = nostate || (o[slot][v] == A && ok[v]);
      }
      float tot[V];
#pragma unroll
      for (int v = 0; v < V; ++v) tot[v] = 0.f;
#pragma unroll 4
      for (int s = 0; s < A; ++s) {
        float gv[V], cnt[V];
#pragma unroll
        for (int v = 0; v < V; ++v) gv[v] = 0.f;
        if (active) ld_f<V>(grow + (size_t)s * p.stride, gv);
#pragma unroll
        for (int v = 0; v < V; ++v) {
          gv[v] = ok[v] ? gv[v] : 0.f;
          tot[v] += gv[v];
          cnt[v] = (o[0][v] == s ? 1.f : 0.f) + (o[1][v] == s ? 1.f : 0.f);
        }
        if constexpr (WANT_DC) {
          sts_v<V>(myR + s * PITCH, gv);
          sts_v<V>(myE + s * PITCH, cnt);
        }
      }
      // d total / d leaf_cost of leaves with a state: all of the parent's mass on the observed state
#pragma unroll
      for (int slot = 0; slot < 2; ++slot) {
        if (gen[slot] || cc[slot] < 0 || !dleaf || !active) continue;
        float* dst = dleaf + (size_t)cc[slot] * A * p.stride;
        for (int t = 0; t < A; ++t) {
          float out[V];
#pragma unroll
          for (int v = 0; v < V; ++v) out[v] = o[slot][v] == t ? tot[v] : 0.f;
          st_f_stream<V>(dst + (size_t)t * p.stride, out);
        }
      }
      if constexpr (WANT_DC) {
        __syncwarp();
        sweep(accG);
        __syncwarp();
      }
      // leaves without a state / missing children: the all-1e5 row has uniform e, w = K / rowsum(K)
      if (__any_sync(kFull, nostate)) {
#pragma unroll
        for (int slot = 0; slot < 2; ++slot) {
          bool ns[V];
          bool any = false;
#pragma unroll
          for (int v = 0; v < V; ++v) { ns[v] = o[slot][v] == A && ok[v]; any = any || ns[v]; }
          if (!__any_sync(kFull, any)) continue;
          float* dst = (cc[slot] >= 0 && dleaf) ? dleaf + (size_t)cc[slot] * A * p.stride : nullptr;
          for (int s = 0; s < A; ++s) {
            float gv[V], one[V];
#pragma unroll
            for (int v = 0; v < V; ++v) gv[v] = 0.f;
            if (active) ld_f<V>(grow + (size_t)s * p.stride, gv);
#pragma unroll
            for (int v = 0; v < V; ++v) {
              gv[v] = ns[v] ? __fdividef(gv[v], sRowK[s]) : 0.f;
              one[v] = ns[v] ? 1.f : 0.f;
            }
            sts_v<V>(myR + s * PITCH, gv);
            sts_v<V>(myE + s * PITCH, one);
          }
          if (dst && any) {
            for (int t = 0; t < A; ++t) {
              float out[V];
              ld_f<V>(dst + (size_t)t * p.stride, out);  // hot lanes keep what was stored above
#pragma unroll
              for (int v = 0; v < V; ++v) {
                if (!ns[v]) continue;
                float a = 0.f;
                for (int s = 0; s < A; ++s) a = fmaf(sKt[t * AP + s], myR[s * PITCH + v], a);
                out[v] = a;
              }
              st_f<V>(dst + (size_t)t * p.stride, out);
            }
          }
          if constexpr (WANT_DC) {
            __syncwarp();
            sweep(acc);
            __syncwarp();
          }
        }
      }
    }

    // ---- children with a cost row (internal nodes, soft leaves) --------------------------------
#pragma unroll 1
    for (int slot = 0; slot < 2; ++slot) {
      if (!(slot ? gen[1] : gen[0])) continue;
      const int c = slot ? cc[1] : cc[0];
      const float* src;
      float* dst;
      if (c >= p.n_leaves) {
        src = dp + (size_t)(c - p.n_leaves) * A * p.stride;
        dst = gbar + (size_t)(c - p.n_leaves) * A * p.stride;
      } else {
        src = p.leaf_cost + (size_t)c * A * p.stride + site0;
        dst = dleaf ? dleaf + (size_t)c * A * p.stride : nullptr;
      }
      float e[AP][V];
#pragma unroll
      for (int t = 0; t < AP; ++t) {
#pragma unroll
        for (int v = 0; v < V; ++v) e[t][v] = t < A ? 0.f : INFINITY;
        if (t < A && active) ld_f_stream<V>(src + (size_t)t * p.stride, e[t]);
      }
#pragma unroll
      for (int v = 0; v < V; ++v) {
        float mm = e[0][v];
#pragma unroll
        for (int t = 1; t < AP; ++t) mm = fminf(mm, e[t][v]);
        mm = ok[v] ? mm : 0.f;
#pragma unroll
        for (int t = 0; t < AP; ++t) e[t][v] = ok[v] ? exp2f((mm - e[t][v]) * c1) : 0.f;
      }
#pragma unroll
      for (int t = 0; t < AP; ++t) sts_v<V>(myE + t * PITCH, e[t]);
      // r[s] = g[s] / Z[s]  -> staged; never kept in registers
#pragma unroll 4
      for (int s = 0; s < A; ++s) {
        float gv[V], z[V];
#pragma unroll
        for (int v = 0; v < V; ++v) { gv[v] = 0.f; z[v] = 0.f; }
        if (active) ld_f<V>(grow + (size_t)s * p.stride, gv);
        const float* krow = sK + s * AP;
        if constexpr (AP % 4 == 0) {
#pragma unroll
          for (int t = 0; t < AP; t += 4) {
            const float4 k4 = *reinterpret_cast<const float4*>(krow + t);
#pragma unroll
            for (int v = 0; v < V; ++v) {
              z[v] = fmaf(k4.x, e[t][v], z[v]);
              z[v] = fmaf(k4.y, e[t + 1][v], z[v]);
              z[v] = fmaf(k4.z, e[t + 2][v], z[v]);
              z[v] = fmaf(k4.w, e[t + 3][v], z[v]);
            }
          }
        } else {
#pragma unroll
          for (int t = 0; t < AP; ++t)
#pragma unroll
            for (int v = 0; v < V; ++v) z[v] = fmaf(krow[t], e[t][v], z[v]);
        }
#pragma unroll
        for (int v = 0; v < V; ++v) gv[v] = ok[v] ? __fdividef(gv[v], z[v]) : 0.f;
        sts_v<V>(myR + s * PITCH, gv);
      }
      // gbar[child][t] = e[t] * sum_s K[s,t] r[s]; r comes back from this lane's own column
      {
        float r[AP][V];
#pragma unroll
        for (int s = 0; s < AP; ++s) {
          if (s < A) ld_f<V>(myR + s * PITCH, r[s]);
          else {
#pragma unroll
            for (int v = 0; v < V; ++v) r[s][v] = 0.f;
          }
        }
#pragma unroll 4
        for (int t = 0; t < A; ++t) {
          float a[V], et[V];
#pragma unroll
          for (int v = 0; v < V; ++v) a[v] = 0.f;
          ld_f<V>(myE + t * PITCH, et);
          const float* ktrow = sKt + t * AP;
          if constexpr (AP % 4 == 0) {
#pragma unroll
            for (int s = 0; s < AP; s += 4) {
              const float4 k4 = *reinterpret_cast<const float4*>(ktrow + s);
#pragma unroll
              for (int v = 0; v < V; ++v) {
                a[v] = fmaf(k4.x, r[s][v], a[v]);
                a[v] = fmaf(k4.y, r[s + 1][v], a[v]);
                a[v] = fmaf(k4.z, r[s + 2][v], a[v]);
                a[v] = fmaf(k4.w, r[s + 3][v], a[v]);
              }
            }
          } else {
#pragma unroll
            for (int s = 0; s < AP; ++s)
#pragma unroll
              for (int v = 0; v < V; ++v) a[v] = fmaf(ktrow[s], r[s][v], a[v]);
          }
#pragma unroll
          for (int v = 0; v < V; ++v) a[v] *= et[v];
          if (dst && active) st_f<V>(dst + (size_t)t * p.stride, a);
        }
      }
      if constexpr (WANT_DC) {
        __syncwarp();
        sweep(acc);
        __syncwarp();
      }
    }
    if constexpr (WANT_DC) {
      if (++since_flush == kFlush) {
        since_flush = 0;
        flush(acc, true);
        flush(accG, false);
      }
    }
  }
  if constexpr (WANT_DC) {
    flush(acc, true);
    flush(accG, false);
    __syncwarp();
    for (int i = lane; i < A * A; i += 32) {
      const double v = sAcc[(i / A) * AP + i % A];
      if (v != 0.0) atomicAdd(p.dcost + (size_t)tree * A * A + i, v);
    }
  }
}

template <int AP, int V, int W>
static int launch_warp(const SoftBwdParams& p, int batch, cudaStream_t s) {
  using Cfg = WarpCfg<AP, V>;
  const size_t smem = sizeof(float) * (Cfg::kSharedFloats + (size_t)W * Cfg::kWarpFloats);
  const long long nwarps = (p.n_sites + Cfg::NS - 1) / Cfg::NS;
  dim3 grid((unsigned)((nwarps + W - 1) / W), batch);
  cudaError_t e;
  if (p.dcost) {
    auto kern = soft_bwd_warp_kernel<AP, V, W, true>;
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return check_cuda(e, "cudaFuncSetAttribute(soft_bwd_warp)");
    kern<<<grid, 32 * W, smem, s>>>(p);
  } else {
    auto kern = soft_bwd_warp_kernel<AP, V, W, false>;
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return check_cuda(e, "cudaFuncSetAttribute(soft_bwd_warp)");
    kern<<<grid, 32 * W, smem, s>>>(p);
  }
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
