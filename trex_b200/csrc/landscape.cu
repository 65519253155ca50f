// Landscape-aware ("parental guidance") loss term — SURVEY §8(f) rank 2.
//   compute_parental_logits                 src/trex/evals/benchmark.py:586-663
//   _compute_loss_landscape_aware_stacked   src/trex/evals/benchmark.py:235-306  (the fitness_cost part, :283-303)
// For node n with parent p = argmax_j Adj[n][j] and site i with neighbours nb_0..nb_{k-1} = interactions[i]:
//     joint[j]  = prod_m P[p][nb_m][digit_m(j)]          j in [0, Q^k), first neighbour most significant (:633-640)
//     logits[s] = sum_j F_i[s * Q^k + j] joint[j]         F_i = fitness_tables[i]                        (:642-652)
//     ce        = - mask[i] sum_s S[n][i][s] log_softmax(logits)[s]                                     (:293-298)
//     fitness_cost = sum ce / (#non-root nodes * #valid sites)                                         (:300-301)
// The reference walks the sites with lax.map in chunks of 64 (safe_map) to bound XLA's memory; here one CTA
// owns one site — its fitness table is staged once in shared memory and broadcast to the nodes (threads) —
// and the value and d fitness_cost / d S come out of the same pass (the child path -w log p, and the parent
// path through the softmax and the table, scattered to the neighbour sites of the parent row with atomics).
#include "common.cuh"

namespace trex {

namespace {

constexpr int kMaxK = 3;
constexpr int kMaxQ = 32;
constexpr int kLsThreads = 128;

// parent[i] = argmax_j adj[i][j] (first maximum, jnp.argmax); scal[0] = ce accumulator (zeroed),
// scal[1] = 1 / (#non-root * #valid)
__global__ void landscape_setup_kernel(const float* __restrict__ adj, const int32_t* __restrict__ parent_in, int N,
                                       const float* __restrict__ mask, long long L, int32_t* __restrict__ parent,
                                       double* __restrict__ scal) {
  __shared__ int s_nonroot;
  __shared__ double s_valid;
  if (threadIdx.x == 0) { s_nonroot = 0; s_valid = 0.0; }
  __syncthreads();
  int nonroot = 0;
  for (int i = threadIdx.x; i < N; i += blockDim.x) {
    int best = 0;
    if (parent_in) {
      best = parent_in[i];
    } else {
      float bv = adj[(size_t)i * N];
      for (int j = 1; j < N; ++j) {
        const float v = adj[(size_t)i * N + j];
        if (v > bv) { bv = v; best = j; }
      }
    }
    parent[i] = best;
    nonroot += best != i;
  }
  double valid = 0.0;
  for (long long l = threadIdx.x; l < L; l += blockDim.x) valid += mask ? (double)mask[l] : 1.0;
  atomicAdd(&s_nonroot, nonroot);
  atomicAdd(&s_valid, valid);
  __syncthreads();
  if (threadIdx.x == 0) {
    scal[0] = 0.0;
    scal[1] = 1.0 / ((double)s_nonroot * s_valid);
  }
}

// QT: compile-time alphabet (4, 20: every per-state array lives in registers) or 0 = run-time Q <= kMaxQ
template <int K, int QT>
__global__ void __launch_bounds__(kLsThreads)
    landscape_kernel(const float* __restrict__ S, const int32_t* __restrict__ parent, const int32_t* __restrict__ inter,
                     const float* __restrict__ tables, const float* __restrict__ mask, int N, long long L, int q_rt,
                     int table_in_smem, float* __restrict__ out_logits, float* __restrict__ dS,
                     double* __restrict__ scal) {
  extern __shared__ __align__(16) float sF[];
  constexpr int QA = QT ? QT : kMaxQ;
  const int Q = QT ? QT : q_rt;
  const long long site = blockIdx.x;
  int QK = 1;
#pragma unroll
  for (int m = 0; m < K; ++m) QK *= Q;
  const size_t tsize = (size_t)QK * Q;
  const float* F = tables + (size_t)site * tsize;
  if (table_in_smem) {
    for (size_t i = threadIdx.x; i < tsize; i += kLsThreads) sF[i] = F[i];
    __syncthreads();
    F = sF;
  }
  int nb[K > 0 ? K : 1];
#pragma unroll
  for (int m = 0; m < K; ++m) nb[m] = inter[site * K + m];
  const float w_site = mask ? mask[site] : 1.f;
  const float wn = (float)scal[1];
  double ce_sum = 0.0;

  for (int node = threadIdx.x; node < N; node += kLsThreads) {
    const int p = parent[node];
    float P[K > 0 ? K : 1][QA], dP[K > 0 ? K : 1][QA];
#pragma unroll
    for (int m = 0; m < K; ++m)
      for (int t = 0; t < Q; ++t) {
        P[m][t] = S[((size_t)p * L + nb[m]) * Q + t];
        dP[m][t] = 0.f;
      }
    // logits[s] = sum_j F[s][j] joint[j]
    float logit[QA];
    _Pragma("unroll") for (int s = 0; s < Q; ++s) logit[s] = 0.f;
    if constexpr (K == 1 && QT > 0) {  // the benchmark's K: a Q x Q mat-vec, everything in registers
#pragma unroll
      for (int j = 0; j < QT; ++j)
        _Pragma("unroll") for (int s = 0; s < Q; ++s) logit[s] = fmaf(F[s * QT + j], P[0][j], logit[s]);
    } else {
      int d[K > 0 ? K : 1];
#pragma unroll
      for (int m = 0; m < K; ++m) d[m] = 0;
      for (int j = 0; j < QK; ++j) {
        float joint = 1.f;
#pragma unroll
        for (int m = 0; m < K; ++m) joint *= P[m][d[m]];
        _Pragma("unroll") for (int s = 0; s < Q; ++s) logit[s] = fmaf(F[(size_t)s * QK + j], joint, logit[s]);
#pragma unroll
        for (int m = K - 1; m >= 0; --m) {  // next digit vector, last neighbour fastest
          if (++d[m] < Q) break;
          d[m] = 0;
        }
      }
    }
    if (out_logits)
      _Pragma("unroll") for (int s = 0; s < Q; ++s) out_logits[((size_t)node * L + site) * Q + s] = logit[s];
    // masked cross-entropy of the child's distribution against log_softmax(logits)
    float mx = logit[0];
    _Pragma("unroll") for (int s = 1; s < Q; ++s) mx = fmaxf(mx, logit[s]);
    float z = 0.f;
    _Pragma("unroll") for (int s = 0; s < Q; ++s) z += expf(logit[s] - mx);
    const float lz = mx + logf(z);
    const float* child = S + ((size_t)node * L + site) * Q;
    float csum = 0.f, ce = 0.f;
    _Pragma("unroll") for (int s = 0; s < Q; ++s) {
      const float c = child[s];
      csum += c;
      ce -= c * (logit[s] - lz);
    }
    ce_sum += (double)(ce * w_site);
    if (dS && w_site != 0.f) {
      const float w = w_site * wn;
      float* dchild = dS + ((size_t)node * L + site) * Q;
      // child path: -w log p;  parent path: d logits = w (softmax * sum(child) - child)
      float dl[QA];
      _Pragma("unroll") for (int s = 0; s < Q; ++s) {
        const float lp = logit[s] - lz;
        atomicAdd(dchild + s, -w * lp);
        dl[s] = w * (expf(lp) * csum - child[s]);
      }
      if constexpr (K == 1 && QT > 0) {
#pragma unroll
        for (int j = 0; j < QT; ++j) {
          float dj = 0.f;
          _Pragma("unroll") for (int s = 0; s < Q; ++s) dj = fmaf(dl[s], F[s * QT + j], dj);
          if (dj != 0.f) atomicAdd(dS + ((size_t)p * L + nb[0]) * Q + j, dj);
        }
      } else if (K > 0) {
        int d[K > 0 ? K : 1];
#pragma unroll
        for (int m = 0; m < K; ++m) d[m] = 0;
        for (int j = 0; j < QK; ++j) {
          float dj = 0.f;
          _Pragma("unroll") for (int s = 0; s < Q; ++s) dj = fmaf(dl[s], F[(size_t)s * QK + j], dj);
#pragma unroll
          for (int m = 0; m < K; ++m) {
            float others = dj;
#pragma unroll
            for (int m2 = 0; m2 < K; ++m2)
              if (m2 != m) others *= P[m2][d[m2]];
            dP[m][d[m]] += others;
          }
#pragma unroll
          for (int m = K - 1; m >= 0; --m) {
            if (++d[m] < Q) break;
            d[m] = 0;
          }
        }
#pragma unroll
        for (int m = 0; m < K; ++m)
          for (int t = 0; t < Q; ++t)
            if (dP[m][t] != 0.f) atomicAdd(dS + ((size_t)p * L + nb[m]) * Q + t, dP[m][t]);
      }
    }
  }
  block_atomic_add(ce_sum, scal);
}

__global__ void landscape_final_kernel(const double* __restrict__ scal, float* __restrict__ out) {
  *out = (float)(scal[0] * scal[1]);
}

}  // namespace

}  // namespace trex

using namespace trex;

extern "C" {

int trex_landscape_loss_fwd_bwd(const float* seqs, const float* adjacency, const int32_t* parent_index,
                                const int32_t* interactions, const float* fitness_tables, const float* seq_mask,
                                int n_nodes, int64_t n_sites, int n_states, int k, float* out_fitness_cost,
                                float* out_logits, float* d_seqs, int32_t* parent_scratch, double* acc_scratch,
                                void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  TREX_CHECK_ARG(seqs && fitness_tables && parent_scratch && acc_scratch, "trex_landscape_loss_fwd_bwd: null pointer");
  TREX_CHECK_ARG(adjacency || parent_index, "trex_landscape_loss_fwd_bwd: adjacency or parent_index must be given");
  TREX_CHECK_ARG(n_nodes >= 1 && n_sites >= 1, "trex_landscape_loss_fwd_bwd: empty problem");
  TREX_CHECK_ARG(n_states >= 1 && n_states <= kMaxQ, "trex_landscape_loss_fwd_bwd: n_states %d not in 1..%d", n_states, kMaxQ);
  TREX_CHECK_ARG(k >= 0 && k <= kMaxK, "trex_landscape_loss_fwd_bwd: k = %d not in 0..%d", k, kMaxK);
  TREX_CHECK_ARG(k == 0 || interactions, "trex_landscape_loss_fwd_bwd: interactions missing");
  TREX_CHECK_ARG(n_sites <= 0x7fffffffLL, "trex_landscape_loss_fwd_bwd: too many sites for one launch");
  size_t tsize = n_states;
  for (int m = 0; m < k; ++m) tsize *= n_states;
  const size_t smem = tsize * sizeof(float);
  const int in_smem = smem <= 160 * 1024;

  landscape_setup_kernel<<<1, 256, 0, s>>>(adjacency, parent_index, n_nodes, seq_mask, n_sites, parent_scratch,
                                          acc_scratch);
  note_launches(1);
  TREX_CHECK_LAUNCH("landscape_setup_kernel");
  if (d_seqs) {
    int rc = check_cuda(cudaMemsetAsync(d_seqs, 0, sizeof(float) * (size_t)n_nodes * n_sites * n_states, s),
                        "cudaMemsetAsync(d_seqs)");
    if (rc) return rc;
  }
  auto go = [&](auto kern) -> int {
    if (in_smem && smem > 48 * 1024) {
      int rc = check_cuda(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                          "cudaFuncSetAttribute(landscape_kernel)");
      if (rc) return rc;
    }
    kern<<<(unsigned)n_sites, kLsThreads, in_smem ? smem : 0, s>>>(seqs, parent_scratch, interactions, fitness_tables,
                                                                  seq_mask, n_nodes, n_sites, n_states, in_smem,
                                                                  out_logits, d_seqs, acc_scratch);
    note_launches(1);
    return check_cuda(cudaPeekAtLastError(), "landscape_kernel");
  };
  int rc;
#define TREX_LS_CASE(KK)                                                   \
  case KK:                                                                 \
    rc = n_states == 4    ? go(landscape_kernel<KK, 4>)                    \
         : n_states == 20 ? go(landscape_kernel<KK, 20>)                   \
                          : go(landscape_kernel<KK, 0>);                   \
    break;
  switch (k) {
    TREX_LS_CASE(0)
    TREX_LS_CASE(1)
    TREX_LS_CASE(2)
    default:
      rc = n_states == 4 ? go(landscape_kernel<3, 4>) : n_states == 20 ? go(landscape_kernel<3, 20>) : go(landscape_kernel<3, 0>);
      break;
  }
#undef TREX_LS_CASE
  if (rc) return rc;
  if (out_fitness_cost) {
    landscape_final_kernel<<<1, 1, 0, s>>>(acc_scratch, out_fitness_cost);
    note_launches(1);
    TREX_CHECK_LAUNCH("landscape_final_kernel");
  }
  return TREX_OK;
}

}  // extern "C"
