// Relaxed Sankoff reverse mode — SURVEY §8a row R2 (no reference code exists).
//
// Top-down over the reverse post-order of the root's subtree.  For a node with cotangent
// g[s] = d total / d dp[node][s] and a child row d[.]:
//     w[s,t] = exp(-(C[s,t] + d[t] - softmin_s)/T) = K[s,t] e[t] / Z[s],  Z[s] = sum_t K[s,t] e[t]
//     gbar[child][t] = e[t] * sum_s (g[s]/Z[s]) K[s,t]
//     dC[s,t]       += K[s,t] * (g[s]/Z[s]) * e[t]          (summed over sites and edges)
// gbar rows of internal nodes are the soft ancestral marginals; for leaves they are
// d total / d leaf_cost.  The dC outer products are accumulated by re-partitioning the CTA:
// r = g/Z and e are staged in shared memory [state][site]; every warp owns an RS x CS block of
// (s,t) pairs in registers and sweeps the CTA's sites (lane = site, conflict-free), so the
// pair sums live in registers for the whole tree walk and are flushed with f64 atomics.
#include "common.cuh"
#include "soft_bwd_params.cuh"

namespace trex {


template <int AP> struct DcTile;  // RS x CS pair block per warp, NBLK blocks cover AP x AP
template <> struct DcTile<2> { static constexpr int RS = 2, CS = 2; };
template <> struct DcTile<4> { static constexpr int RS = 4, CS = 4; };
template <> struct DcTile<8> { static constexpr int RS = 4, CS = 8; };
template <> struct DcTile<20> { static constexpr int RS = 5, CS = 10; };
template <> struct DcTile<32> { static constexpr int RS = 8, CS = 8; };

__device__ __forceinline__ float warp_sum_f(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

template <int AP, int V, int THREADS, bool WANT_DC>
__global__ void __launch_bounds__(THREADS) soft_bwd_kernel(const SoftBwdParams p) {
  constexpr int RS = DcTile<AP>::RS, CS = DcTile<AP>::CS;
  constexpr int NBLK = (AP / RS) * (AP / CS);
  constexpr int NW = THREADS / 32;
  static_assert(NW % NBLK == 0, "warps must tile the pair blocks");
  constexpr int SS = NW / NBLK;       // site split between warps that own the same block
  constexpr int NS = THREADS * V;     // sites per CTA
  constexpr int FLUSH = 64;           // nodes between f64 flushes of the register pair sums

  extern __shared__ __align__(16) float smem[];
  float* sK = smem;                    // [s][t]
  float* sKt = sK + AP * AP;           // [t][s]
  float* sC = sKt + AP * AP;           // [s][t] (direct path)
  float* sRowK = sC + AP * AP;         // [s] sum_t K[s,t]
  float* sR = sRowK + AP;              // [AP][NS]  (WANT_DC)
  float* sE = sR + (WANT_DC ? AP * NS : 0);
  float* sP = sE + (WANT_DC ? AP * NS : 0);  // [AP*AP] direct-path pair sums (WANT_DC)

  const int A = p.n_states;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tree = blockIdx.y;
  const int n_int = p.n_all - p.n_leaves;
  PlanView plan(p.plan + (size_t)tree * p.plan_words, n_int);
  const float c1 = 1.4426950408889634f / p.T;

  bool c_ok = true;
  for (int i = tid; i < AP * AP; i += THREADS) {
    const int s = i / AP, t = i % AP;
    float c = 0.f, k = 0.f;
    if (s < A && t < A) {
      c = __ldg(p.cost + s * A + t);
      c_ok = c_ok && (fabsf(c) <= 80.f * p.T);
      k = exp2f(-c * c1);
    } else if (s < A) {
      c = INFINITY;
    }
    sC[i] = c;
    sK[i] = k;
    sKt[t * AP + s] = k;
    if (WANT_DC) sP[i] = 0.f;
  }
  const int expdom = __syncthreads_and(c_ok);
  if (p.only_direct && expdom) return;  // soft_bwd_warp_kernel owns this case
  for (int s = tid; s < AP; s += THREADS) {
    float z = 0.f;
    for (int t = 0; t < AP; ++t) z += sK[s * AP + t];
    sRowK[s] = z;
  }
  __syncthreads();

  const long long site0 = ((long long)blockIdx.x * THREADS + tid) * V;
  const bool active = site0 < p.n_sites;
  const float gs = p.gscale ? __ldg(p.gscale) : 1.f;
  const float* dp = p.dp + (size_t)tree * n_int * A * p.stride + site0;
  float* gbar = p.gbar + (size_t)tree * n_int * A * p.stride + site0;
  float* dleaf = p.dleaf ? p.dleaf + (size_t)tree * p.n_leaves * A * p.stride + site0 : nullptr;
  float* dleaf_obs = p.dleaf_obs ? p.dleaf_obs + (size_t)tree * p.n_leaves * p.stride + site0 : nullptr;
  const int32_t* order = plan.order();
  const int n_reach = plan.hdr()[1];

  // pair-block ownership of this warp
  const int blk = warp % NBLK, ss = warp / NBLK;
  const int s0 = (blk / (AP / CS)) * RS, t0 = (blk % (AP / CS)) * CS;
  float acc[RS][CS];
#pragma unroll
  for (int a = 0; a < RS; ++a)
#pragma unroll
    for (int b = 0; b < CS; ++b) acc[a][b] = 0.f;

  auto flush = [&]() {
#pragma unroll
    for (int a = 0; a < RS; ++a)
#pragma unroll
      for (int b = 0; b < CS; ++b) {
        const float v = warp_sum_f(acc[a][b]);
        acc[a][b] = 0.f;
        if (lane == 0 && s0 + a < A && t0 + b < A && v != 0.f)
          atomicAdd(p.dcost + ((size_t)tree * A + s0 + a) * A + t0 + b,
                    (double)v * (double)sK[(s0 + a) * AP + t0 + b]);
      }
  };

  if (active) {
    float zero[V];
#pragma unroll
    for (int v = 0; v < V; ++v) zero[v] = 0.f;
    for (int i = n_reach; i < n_int; ++i) {
      const int ku = __ldg(order + i) - p.n_leaves;
      float* row = gbar + (size_t)ku * A * p.stride;
      for (int s = 0; s < A; ++s) st_f<V>(row + (size_t)s * p.stride, zero);
      for (int slot = 0; slot < 2; ++slot) {  // ... and so do the leaf-gradient rows of their leaf children
        const int c = __ldg(plan.child(slot) + ku);
        if (c < 0 || c >= p.n_leaves) continue;
        if (dleaf)
          for (int s = 0; s < A; ++s) st_f<V>(dleaf + ((size_t)c * A + s) * p.stride, zero);
        if (dleaf_obs) st_f<V>(dleaf_obs + (size_t)c * p.stride, zero);
      }
    }
    // root cotangent: gscale * softmax_s(-dp[root]/T)
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
    float g[AP][V];
    if (active) {
      const float* grow = gbar + (size_t)k * A * p.stride;
#pragma unroll
      for (int s = 0; s < AP; ++s) {
        if (s < A) ld_f<V>(grow + (size_t)s * p.stride, g[s]);
        else {
#pragma unroll
          for (int v = 0; v < V; ++v) g[s][v] = 0.f;
        }
      }
    }
#pragma unroll 1
    for (int slot = 0; slot < 2; ++slot) {
      const int c = __ldg(plan.child(slot) + k);
      float e[AP][V], r[AP][V];
      if (active) {
        const float* src = nullptr;
        float* dst = nullptr;
        int o[V];
#pragma unroll
        for (int v = 0; v < V; ++v) o[v] = A;
        if (c >= p.n_leaves) {
          src = dp + (size_t)(c - p.n_leaves) * A * p.stride;
          dst = gbar + (size_t)(c - p.n_leaves) * A * p.stride;
        } else if (c >= 0) {
          if (dleaf) dst = dleaf + (size_t)c * A * p.stride;
          if (p.leaf_cost) src = p.leaf_cost + (size_t)c * A * p.stride + site0;
          else {
            ld_u8<V>(p.leaf + (size_t)c * p.stride + site0, o);
#pragma unroll
            for (int v = 0; v < V; ++v) o[v] = min(o[v], A);
          }
        }
        float gc[AP][V];
        if (expdom) {
          if (src) {
#pragma unroll
            for (int t = 0; t < AP; ++t) {
              if (t < A) ld_f_stream<V>(src + (size_t)t * p.stride, e[t]);
              else {
#pragma unroll
                for (int v = 0; v < V; ++v) e[t][v] = INFINITY;
              }
            }
#pragma unroll
            for (int v = 0; v < V; ++v) {
              float mm = e[0][v];
#pragma unroll
              for (int t = 1; t < AP; ++t) mm = fminf(mm, e[t][v]);
#pragma unroll
              for (int t = 0; t < AP; ++t) e[t][v] = exp2f((mm - e[t][v]) * c1);
            }
            // Z[s] and r[s] = g[s]/Z[s]
#pragma unroll
            for (int s = 0; s < AP; ++s) {
#pragma unroll
              for (int v = 0; v < V; ++v) {
                float z = 0.f;
#pragma unroll
                for (int t = 0; t < AP; ++t) z = fmaf(sK[s * AP + t], e[t][v], z);
                r[s][v] = s < A ? __fdividef(g[s][v], z) : 0.f;
              }
            }
#pragma unroll
            for (int t = 0; t < AP; ++t) {
#pragma unroll
              for (int v = 0; v < V; ++v) {
                float a = 0.f;
#pragma unroll
                for (int s = 0; s < AP; ++s) a = fmaf(sKt[t * AP + s], r[s][v], a);
                gc[t][v] = a * e[t][v];
              }
            }
          } else {
            // leaf given as a state (or the all-1e5 row): e is one-hot (or all ones)
#pragma unroll
            for (int v = 0; v < V; ++v) {
              const bool hot = o[v] < A;
#pragma unroll
              for (int t = 0; t < AP; ++t) e[t][v] = hot ? (t == o[v] ? 1.f : 0.f) : (t < A ? 1.f : 0.f);
              float tot = 0.f;
#pragma unroll
              for (int s = 0; s < AP; ++s) {
                const float z = hot ? sKt[o[v] * AP + s] : sRowK[s];
                r[s][v] = s < A ? __fdividef(g[s][v], z) : 0.f;
                tot += s < A ? g[s][v] : 0.f;
              }
              if (hot) {
#pragma unroll
                for (int t = 0; t < AP; ++t) gc[t][v] = t == o[v] ? tot : 0.f;
              } else if (dst) {
#pragma unroll
                for (int t = 0; t < AP; ++t) {
                  float a = 0.f;
#pragma unroll
                  for (int s = 0; s < AP; ++s) a = fmaf(sKt[t * AP + s], r[s][v], a);
                  gc[t][v] = a;
                }
              }
            }
          }
        } else {
          // direct evaluation (max|C|/T > 80): w[s,t] = exp(-(x - mm)/T)/Z, x = C[s,t] + d[t]
          if (src) {
#pragma unroll
            for (int t = 0; t < AP; ++t) {
              if (t < A) ld_f_stream<V>(src + (size_t)t * p.stride, e[t]);
              else {
#pragma unroll
                for (int v = 0; v < V; ++v) e[t][v] = INFINITY;
              }
            }
          } else {
#pragma unroll
            for (int v = 0; v < V; ++v)
#pragma unroll
              for (int t = 0; t < AP; ++t) e[t][v] = t < A ? ((t == o[v]) ? 0.f : kSentinel) : INFINITY;
          }
#pragma unroll
          for (int t = 0; t < AP; ++t)
#pragma unroll
            for (int v = 0; v < V; ++v) gc[t][v] = 0.f;
#pragma unroll
          for (int s = 0; s < AP; ++s) {
            if (s >= A) break;
#pragma unroll
            for (int v = 0; v < V; ++v) {
              float mm = INFINITY;
#pragma unroll
              for (int t = 0; t < AP; ++t) mm = fminf(mm, sC[s * AP + t] + e[t][v]);
              float w[AP];
              float z = 0.f;
#pragma unroll
              for (int t = 0; t < AP; ++t) { w[t] = exp2f((mm - (sC[s * AP + t] + e[t][v])) * c1); z += w[t]; }
              const float gz = __fdividef(g[s][v], z);
#pragma unroll
              for (int t = 0; t < AP; ++t) {
                const float contrib = gz * w[t];
                gc[t][v] += contrib;
                if (WANT_DC && t < A && site0 + v < p.n_sites && contrib != 0.f)
                  atomicAdd(&sP[s * AP + t], contrib);
              }
            }
          }
        }
        if (dst) {
#pragma unroll
          for (int t = 0; t < AP; ++t)
            if (t < A) st_f<V>(dst + (size_t)t * p.stride, gc[t]);
        }
        if (dleaf_obs && c >= 0 && c < p.n_leaves && !p.leaf_cost) {
          // compact leaf gradient: the whole cotangent mass of the parent lands on the leaf's row
          // (on its observed-state entry when it has one) because the softmin weights of a row sum to 1
          float mass[V];
#pragma unroll
          for (int v = 0; v < V; ++v) {
            mass[v] = 0.f;
#pragma unroll
            for (int s = 0; s < AP; ++s) mass[v] += s < A ? g[s][v] : 0.f;
          }
          st_f<V>(dleaf_obs + (size_t)c * p.stride, mass);
        }
      }
      if constexpr (WANT_DC) {
        if (expdom) {
          // stage r and e [state][site]; pad lanes / inactive threads contribute zeros
#pragma unroll
          for (int s = 0; s < AP; ++s) {
            float rv[V], ev[V];
#pragma unroll
            for (int v = 0; v < V; ++v) {
              const bool ok = active && (site0 + v < p.n_sites);
              rv[v] = ok ? r[s][v] : 0.f;
              ev[v] = ok ? e[s][v] : 0.f;
            }
            st_f<V>(sR + s * NS + tid * V, rv);
            st_f<V>(sE + s * NS + tid * V, ev);
          }
          __syncthreads();
          for (int j = ss * 32 + lane; j < NS; j += 32 * SS) {
            float rr[RS], ee[CS];
#pragma unroll
            for (int a = 0; a < RS; ++a) rr[a] = sR[(s0 + a) * NS + j];
#pragma unroll
            for (int b = 0; b < CS; ++b) ee[b] = sE[(t0 + b) * NS + j];
#pragma unroll
            for (int a = 0; a < RS; ++a)
#pragma unroll
              for (int b = 0; b < CS; ++b) acc[a][b] = fmaf(rr[a], ee[b], acc[a][b]);
          }
          __syncthreads();
        }
      }
    }
    if constexpr (WANT_DC) {
      if (++since_flush == FLUSH) {
        since_flush = 0;
        if (expdom) flush();
      }
    }
  }
  if constexpr (WANT_DC) {
    if (expdom) {
      flush();
    } else {
      __syncthreads();
      for (int i = tid; i < A * A; i += THREADS) {
        const int s = i / A, t = i % A;
        const float v = sP[s * AP + t];
        if (v != 0.f) atomicAdd(p.dcost + (size_t)tree * A * A + i, (double)v);
      }
    }
    // reverse mode is defined on trees only (include/trex_b200.h, TREX_PLAN_NOT_A_TREE / _SHARED_LEAF): poison dC
    if (blockIdx.x == 0 && (plan.hdr()[3] & (TREX_PLAN_NOT_A_TREE | TREX_PLAN_SHARED_LEAF)))
      for (int i = tid; i < A * A; i += THREADS) atomicAdd(p.dcost + (size_t)tree * A * A + i, (double)NAN);
  }
}

template <int AP, int V, int THREADS>
static int launch_soft_bwd(const SoftBwdParams& p, int batch, cudaStream_t s) {
  const bool want_dc = p.dcost != nullptr;
  const size_t base = (size_t)(3 * AP * AP + AP) * sizeof(float);
  const size_t smem = base + (want_dc ? (size_t)(2 * AP * THREADS * V + AP * AP) * sizeof(float) : 0);
  const long long nvec = (p.n_sites + V - 1) / V;
  dim3 grid((unsigned)((nvec + THREADS - 1) / THREADS), batch);
  cudaError_t e;
  if (want_dc) {
    auto kern = soft_bwd_kernel<AP, V, THREADS, true>;
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return check_cuda(e, "cudaFuncSetAttribute(soft_bwd)");
    kern<<<grid, THREADS, smem, s>>>(p);
    note_launches(1);
  } else {
    auto kern = soft_bwd_kernel<AP, V, THREADS, false>;
    kern<<<grid, THREADS, smem, s>>>(p);
    note_launches(1);
  }
  return check_cuda(cudaPeekAtLastError(), "soft_bwd_kernel");
}

}  // namespace trex

namespace trex {
static int soft_bwd_entry(const int32_t* plan, int n_all, int n_leaves, int n_states, int64_t n_sites,
                          int64_t site_stride, int batch, const uint8_t* leaf_states, const float* leaf_cost,
                          const float* cost, float temperature, const float* dp, const float* gscale,
                          float* gbar, float* dleaf, float* dleaf_obs, double* dcost, void* stream) {
  TREX_CHECK_ARG(plan && cost && dp && gbar, "trex_sankoff_soft_bwd: null pointer");
  TREX_CHECK_ARG((leaf_states != nullptr) != (leaf_cost != nullptr),
                 "trex_sankoff_soft_bwd: exactly one of leaf_states / leaf_cost must be given");
  TREX_CHECK_ARG(!(dleaf_obs && leaf_cost), "trex_sankoff_soft_bwd_compact: the compact leaf gradient needs leaf_states");
  TREX_CHECK_ARG(n_leaves >= 1 && n_all > n_leaves, "trex_sankoff_soft_bwd: need n_all > n_leaves >= 1");
  TREX_CHECK_ARG(n_states >= 1 && n_states <= 32, "trex_sankoff_soft_bwd: n_states %d not in 1..32", n_states);
  TREX_CHECK_ARG(n_sites >= 1 && site_stride >= n_sites && site_stride % 16 == 0,
                 "trex_sankoff_soft_bwd: site_stride must be a multiple of 16 and >= n_sites");
  TREX_CHECK_ARG(batch >= 1 && batch <= 65535, "trex_sankoff_soft_bwd: batch out of range");
  TREX_CHECK_ARG(temperature > 0.f && temperature < 1e30f, "trex_sankoff_soft_bwd: temperature must be > 0");
  cudaStream_t s = (cudaStream_t)stream;
  if (dcost) {
    int rc = check_cuda(cudaMemsetAsync(dcost, 0, sizeof(double) * batch * n_states * n_states, s), "memset dcost");
    if (rc) return rc;
  }
  SoftBwdParams p{plan, trex_sankoff_plan_words(n_all, n_leaves), n_all, n_leaves, n_states,
                  (long long)n_sites, (long long)site_stride, leaf_states, leaf_cost, cost,
                  temperature, dp, gscale, gbar, dleaf, dleaf_obs, dcost, 0, 0};
  const int A = n_states;
  if (A <= 20) {
    // Device-side dispatch (the cost matrix is device data): the O(A) kernel returns at once unless
    // the cost is uniform, the warp-autonomous exp-domain kernel returns at once when it is or when
    // max|C|/T > 80, the direct-evaluation kernel returns at once unless max|C|/T > 80.
    int rc = launch_soft_bwd_uniform(p, batch, s);
    if (rc) return rc;
    rc = launch_soft_bwd_warp(p, batch, s);
    if (rc) return rc;
    p.only_direct = 1;
  }
  if (A <= 2) return launch_soft_bwd<2, 2, 256>(p, batch, s);
  if (A <= 4) return launch_soft_bwd<4, 2, 256>(p, batch, s);
  if (A <= 8) return launch_soft_bwd<8, 1, 256>(p, batch, s);
  if (A <= 20) return launch_soft_bwd<20, 1, 256>(p, batch, s);
  return launch_soft_bwd<32, 1, 512>(p, batch, s);
}
}  // namespace trex

extern "C" int trex_sankoff_soft_bwd(const int32_t* plan, int n_all, int n_leaves, int n_states,
                                     int64_t n_sites, int64_t site_stride, int batch,
                                     const uint8_t* leaf_states, const float* leaf_cost,
                                     const float* cost, float temperature, const float* dp,
                                     const float* gscale, float* gbar, float* dleaf,
                                     double* dcost, void* stream) {
  return trex::soft_bwd_entry(plan, n_all, n_leaves, n_states, n_sites, site_stride, batch, leaf_states, leaf_cost,
                              cost, temperature, dp, gscale, gbar, dleaf, nullptr, dcost, stream);
}

extern "C" int trex_sankoff_soft_bwd_compact(const int32_t* plan, int n_all, int n_leaves, int n_states,
                                             int64_t n_sites, int64_t site_stride, int batch,
                                             const uint8_t* leaf_states, const float* cost, float temperature,
                                             const float* dp, const float* gscale, float* gbar,
                                             float* dleaf_obs, double* dcost, void* stream) {
  return trex::soft_bwd_entry(plan, n_all, n_leaves, n_states, n_sites, site_stride, batch, leaf_states, nullptr,
                              cost, temperature, dp, gscale, gbar, nullptr, dleaf_obs, dcost, stream);
}
