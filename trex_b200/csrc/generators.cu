// On-device input generators — SURVEY 8(f) rank 3: the step immediately before the hot path.
//
//   generate_groundtruth / mutate        src/trex/ground_truth.py:20-197
//   create_nk_model_landscape            src/trex/nk_model.py:17-43
//   get_fitness / batched_get_fitness    src/trex/nk_model.py:46-113
//   generate_tree_data (MH evolution)    src/trex/nk_model.py:116-278
//   create_balanced_binary_tree          src/trex/evals/benchmark.py:781-791
//
// Random numbers.  The reference draws from jax.random (threefry), whose streams cannot be reproduced outside JAX
// (SURVEY 8c), so every generator takes its draws in one of two ways: EXPLICIT arrays supplied by the caller (the
// parity tests: the same arrays go to the NumPy restatement of the reference code, results must be identical), or
// NULL pointers, in which case the kernels draw from a counter-based Philox4x32-10 stream keyed by (seed, stream id,
// counter) — statistically equivalent, reproducible, no state, any number of threads.
//
// B200 design (not a translation of the reference's fori_loop over nodes):
//  * generate_groundtruth: given each edge's mutation set, sites are independent — one thread owns one site and walks
//    the whole tree top-down (parent index > child index, root last), fully coalesced u8 rows.  The mutation set of an
//    edge is a bit row; "choose n_mutations distinct sites" is a keyed Feistel permutation of [0, L) (distinct by
//    construction, no rejection loop, parallel).
//  * NK evolution: every Metropolis-Hastings step accepts or rejects a whole proposal on the mean fitness over all
//    sites, so one CTA owns one node (threads stride over the sites, deterministic f64 block reductions); nodes are
//    ordered by depth and a CTA spins on its parent's ready flag — its parent always has a lower block index, so the
//    wait cannot deadlock; one launch for the whole tree instead of one per level.
#include <cstdint>

#include "common.cuh"

namespace trex {

// ---- Philox4x32-10 (Salmon et al. 2011): (key0, key1) x (c0..c3) -> 4 x u32 ---------------------------------------
struct Philox {
  uint32_t k0, k1;
  __host__ __device__ Philox(uint64_t seed, uint32_t stream) : k0((uint32_t)seed ^ (stream * 0x9E3779B9u)), k1((uint32_t)(seed >> 32) + stream) {}
  __device__ __forceinline__ uint4 operator()(uint64_t counter, uint32_t sub = 0) const {
    uint32_t c0 = (uint32_t)counter, c1 = (uint32_t)(counter >> 32), c2 = sub, c3 = 0x1BD11BDAu;
    uint32_t a = k0, b = k1;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
      const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
      const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
      c0 = hi1 ^ c1 ^ a; c1 = lo1; c2 = hi0 ^ c3 ^ b; c3 = lo0;
      a += 0x9E3779B9u; b += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
  }
};
__device__ __forceinline__ float u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }  // [0, 1)

enum : uint32_t { kStreamOffsets = 1, kStreamPerm = 2, kStreamTable = 3, kStreamInter = 4, kStreamNk = 5, kStreamNkSite = 6 };

// ---- keyed permutation of [0, L): 4-round Feistel on 2^(2h) >= L with cycle walking -----------------------------------
__device__ __forceinline__ uint32_t feistel_perm(uint32_t x, uint32_t L, uint32_t half_bits, const uint32_t (&rk)[4]) {
  const uint32_t mask = (1u << half_bits) - 1u;
  do {
    uint32_t l = x >> half_bits, r = x & mask;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint32_t f = (r + rk[i]) * 0x9E3779B1u;
      f ^= f >> 15; f *= 0x85EBCA77u; f ^= f >> 13;
      const uint32_t nl = r, nr = l ^ (f & mask);
      l = nl; r = nr;
    }
    x = (l << half_bits) | r;
  } while (x >= L);
  return x;
}

// ---- generate_groundtruth (ground_truth.py:113-197) ---------------------------------------------------------------------
// mut_sites [n_edges][n_mut] i32: the `random.choice(..., replace=False)` draw of every mutate call (:40-45); edge = child
// node index (children 2 p_i, 2 p_i + 1 of parent n_leaves + p_i, :163-166).  Filled here when the caller gave none.
__global__ void gt_draw_sites_kernel(int n_edges, int n_mut, uint32_t L, uint64_t seed, int32_t* __restrict__ mut_sites) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)n_edges * n_mut) return;
  const int e = (int)(i / n_mut), j = (int)(i % n_mut);
  uint32_t half_bits = 1;
  while ((1ull << (2 * half_bits)) < L) ++half_bits;
  const uint4 k = Philox(seed, kStreamPerm)((uint64_t)e);
  const uint32_t rk[4] = {k.x, k.y, k.z, k.w};
  mut_sites[i] = (int32_t)feistel_perm((uint32_t)j, L, half_bits, rk);
}

// mutation_mask.at[mutations].set(True) (:46) as bit rows [n_edges][words]
__global__ void gt_scatter_mask_kernel(int n_edges, int n_mut, long long L, const int32_t* __restrict__ mut_sites,
                                       uint32_t* __restrict__ mask, int words) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)n_edges * n_mut) return;
  const int e = (int)(i / n_mut);
  long long s = mut_sites[i];
  if (s < 0) s += L;            // negative indices wrap (jnp scatter)
  if (s < 0 || s >= L) return;  // out-of-bounds updates are dropped
  atomicOr(&mask[(size_t)e * words + (s >> 5)], 1u << (s & 31));
}

// one thread = one site: root state 0 (:148), then every parent from the root down hands its state to its two
// children, mutated where the edge's mask says so: (state + offset) % n_states, offset in 1..n_states-1 (:47-49)
__global__ void __launch_bounds__(256)
    gt_walk_kernel(int n_leaves, int n_states, long long L, long long stride, uint64_t seed,
                   const uint32_t* __restrict__ mask, int words, const uint8_t* __restrict__ offsets,
                   uint8_t* __restrict__ states) {
  const long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= L) return;
  const int n_all = 2 * n_leaves - 1;
  const Philox rng(seed, kStreamOffsets);
  states[(size_t)(n_all - 1) * stride + s] = 0;
  for (int parent = n_all - 1; parent >= n_leaves; --parent) {
    const int ps = states[(size_t)parent * stride + s];
    const int p_i = parent - n_leaves;
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const int child = 2 * p_i + c;
      int st = ps;
      if ((mask[(size_t)child * words + (s >> 5)] >> (s & 31)) & 1u) {
        int off;
        if (offsets) off = offsets[(size_t)child * L + s];
        else off = n_states > 1 ? 1 + (int)(rng((uint64_t)child * (uint64_t)L + (uint64_t)s).x % (uint32_t)(n_states - 1)) : 0;
        st = (ps + off) % n_states;
      }
      states[(size_t)child * stride + s] = (uint8_t)st;
    }
  }
}

// adjacency of generate_groundtruth (:180-191: children 2i, 2i+1 -> parent n_leaves + i) or of
// create_balanced_binary_tree (evals/benchmark.py:781-791: leaf i -> n + i/2, ancestor k -> n + (k + n)/2)
__global__ void tree_adjacency_kernel(int n_leaves, int kind, float* __restrict__ adj) {
  const int n_all = 2 * n_leaves - 1;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)n_all * n_all) return;
  const int child = (int)(i / n_all), parent = (int)(i % n_all);
  int want = -1;
  if (child < n_all - 1) {
    if (kind == 0) want = n_leaves + child / 2;
    else want = child < n_leaves ? n_leaves + child / 2 : n_leaves + (child - n_leaves + n_leaves) / 2;
  }
  adj[i] = parent == want ? 1.f : 0.f;
}

// ---- NK landscape (nk_model.py:17-43): interactions ~ randint(0, n), fitness tables ~ U[0, 1) ---------------------------
__global__ void nk_landscape_kernel(long long n, int k, long long table, uint64_t seed, int32_t* __restrict__ interactions,
                                    float* __restrict__ tables) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n * k) interactions[i] = (int32_t)(Philox(seed, kStreamInter)((uint64_t)i).x % (uint32_t)n);
  if (i < n * table) tables[i] = u01(Philox(seed, kStreamTable)((uint64_t)i).x);
}

// table index of site i for sequence `seq` (nk_model.py:96-104): sum_j s_j q^j, s_0 = seq[i], s_1.. = partners
__device__ __forceinline__ long long nk_index(const uint8_t* __restrict__ seq, long long i, int k, int q,
                                              const int32_t* __restrict__ inter) {
  long long idx = seq[i], pw = q;
  for (int j = 0; j < k; ++j) {
    idx += (long long)seq[inter[i * k + j]] * pw;
    pw *= q;
  }
  return idx;
}

// deterministic block-wide f64 sum (fixed tree order: the result does not depend on scheduling)
template <int THREADS>
__device__ __forceinline__ double block_sum(double v, double* red) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  double t = 0.0;
  for (int i = 0; i < THREADS / 32; ++i) t += red[i];
  return t;
}

// batched_get_fitness (nk_model.py:46-113): mean over (masked) sites of the table entries; one CTA per sequence
template <int THREADS>
__global__ void __launch_bounds__(THREADS)
    nk_fitness_kernel(const uint8_t* __restrict__ seqs, long long stride, long long n, int k, int q, long long table,
                      const int32_t* __restrict__ inter, const float* __restrict__ tables,
                      const float* __restrict__ seq_mask, float* __restrict__ out) {
  __shared__ double red[THREADS / 32];
  const uint8_t* seq = seqs + (size_t)blockIdx.x * stride;
  double f = 0.0, m = 0.0;
  for (long long i = threadIdx.x; i < n; i += THREADS) {
    const float w = seq_mask ? seq_mask[i] : 1.f;
    f += (double)(tables[i * table + nk_index(seq, i, k, q, inter)] * w);
    m += (double)w;
  }
  f = block_sum<THREADS>(f, red);
  m = block_sum<THREADS>(m, red);
  if (threadIdx.x == 0) out[blockIdx.x] = (float)(f / m);
}

// ---- generate_tree_data (nk_model.py:116-278) ---------------------------------------------------------------------------
// parents = argmax(adjacency, axis=1) (:148), root = the node that is its own argmax (:149); depth by walking up.
__global__ void nk_parents_kernel(const float* __restrict__ adj, int N, int32_t* __restrict__ parent) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= N) return;
  const float* row = adj + (size_t)v * N;
  float b = row[0];
  int bi = 0;
  for (int j = 1; j < N; ++j)
    if (row[j] > b) { b = row[j]; bi = j; }
  // The reference finds the root as "the node that is its own argmax" (:149), which needs a root self-loop; its own
  // callers pass adjacencies without one (evals/benchmark.py:917, tests/test_nk_model_new.py:29), for which the
  // reference degenerates (jnp.where(size=1) falls back to node 0 and only nodes 0 and N-1 ever get a sequence).
  // That defect is documented, not reproduced: a node whose row has no positive entry is a root as well.
  parent[v] = b > 0.f ? bi : v;
}
// order[] = nodes sorted by (depth, index): the BFS order of :151-185 up to the order inside a level, which does not
// matter (a node only needs its parent).  One CTA; N is small (thousands).
__global__ void nk_order_kernel(const int32_t* __restrict__ parent, int N, int32_t* __restrict__ depth,
                                int32_t* __restrict__ order, int32_t* __restrict__ ready) {
  for (int v = threadIdx.x; v < N; v += blockDim.x) {
    int d = 0, u = v;
    while (parent[u] != u && d <= N) { u = parent[u]; ++d; }
    depth[v] = d <= N ? d : N;  // (a cycle that never reaches a root: parked at the end)
    ready[v] = 0;
  }
  __syncthreads();
  if (threadIdx.x == 0) {  // counting sort by depth, stable in the index
    int pos = 0;
    for (int d = 0; d <= N && pos < N; ++d)
      for (int v = 0; v < N; ++v)
        if (depth[v] == d) order[pos++] = v;
  }
}

struct NkDraws {  // explicit draws (parity tests) or all-NULL (Philox); indexed [node][step] (+ [site])
  const float* noise;        // [N]           N(0,1) of the per-edge rate noise (:199)
  const float* u_coupled;    // [N][B]        < coupled_mutation_prob -> coupled mutation (:230)
  const int32_t* site;       // [N][B]        the site of a coupled mutation (:221)
  const float* u_mask;       // [N][B][L]     < rate -> the site mutates (random mutation, :214)
  const uint8_t* new_vals;   // [N][B][L]     proposed states (both kinds, :216 / :225)
  const float* u_accept;     // [N][B]        < min(1, exp(dF)) -> accept (:243-244)
};

template <int THREADS>
__global__ void __launch_bounds__(THREADS)
    nk_evolve_kernel(const int32_t* __restrict__ order, const int32_t* __restrict__ parent, int N, long long L,
                     long long stride, int k, int q, long long table, const int32_t* __restrict__ inter,
                     const float* __restrict__ tables, const uint8_t* __restrict__ root_seq, float mutation_rate,
                     float coupled_prob, float noise_std, int branch_length, uint64_t seed, NkDraws dr,
                     uint8_t* __restrict__ states, uint8_t* __restrict__ scratch, volatile int32_t* __restrict__ ready) {
  __shared__ double red[THREADS / 32];
  __shared__ int s_flag;
  const int node = order[blockIdx.x];
  const int par = parent[node];
  uint8_t* cur = states + (size_t)node * stride;
  uint8_t* prop = scratch + (size_t)blockIdx.x * stride;  // the proposal of this CTA
  const int tid = threadIdx.x;
  if (par == node) {  // the root keeps the caller's sequence (:189, :262-266)
    for (long long i = tid; i < L; i += THREADS) cur[i] = root_seq[i];
  } else {
    if (tid == 0) {
      while (ready[par] == 0) __nanosleep(64);  // parent = lower block index: already running or done
      __threadfence();
    }
    __syncthreads();
    const uint8_t* pseq = states + (size_t)par * stride;
    for (long long i = tid; i < L; i += THREADS) cur[i] = pseq[i];
    __syncthreads();
    const Philox rng(seed, kStreamNk), rng_site(seed, kStreamNkSite);
    // per-edge mutation rate: base * exp(N(0,1) * std), clamped to 1 (:198-204)
    float z;
    if (dr.noise) z = dr.noise[node];
    else {
      const uint4 r = rng((uint64_t)node, 0xffffffffu);
      z = sqrtf(-2.f * logf(fmaxf(u01(r.x), 1e-12f))) * cospif(2.f * u01(r.y));
    }
    const float rate = fminf(mutation_rate * expf(z * noise_std), 1.0f);
    for (int step = 0; step < branch_length; ++step) {
      const size_t ns = (size_t)node * branch_length + step;
      const uint4 r = rng((uint64_t)ns, 1);
      const float uc = dr.u_coupled ? dr.u_coupled[ns] : u01(r.x);
      const long long site = dr.site ? dr.site[ns] : (long long)(r.y % (uint32_t)L);
      const float ua = dr.u_accept ? dr.u_accept[ns] : u01(r.z);
      const bool coupled = uc < coupled_prob;
      // proposal (:212-236)
      for (long long i = tid; i < L; i += THREADS) {
        bool hit;
        if (coupled) {
          hit = i == site;
          for (int j = 0; j < k && !hit; ++j) hit = inter[site * k + j] == i;
        } else {
          const float um = dr.u_mask ? dr.u_mask[ns * L + i] : u01(rng_site((uint64_t)(ns * L + i), 0).x);
          hit = um < rate;
        }
        uint8_t v = cur[i];
        if (hit) v = dr.new_vals ? dr.new_vals[ns * L + i] : (uint8_t)(rng_site((uint64_t)(ns * L + i), 1).x % (uint32_t)q);
        prop[i] = v;
      }
      __syncthreads();
      // Metropolis-Hastings (:238-246): mean fitness of both, accept with min(1, exp(dF))
      double f0 = 0.0, f1 = 0.0;
      for (long long i = tid; i < L; i += THREADS) {
        f0 += (double)tables[i * table + nk_index(cur, i, k, q, inter)];
        f1 += (double)tables[i * table + nk_index(prop, i, k, q, inter)];
      }
      f0 = block_sum<THREADS>(f0, red);
      f1 = block_sum<THREADS>(f1, red);
      if (tid == 0) {
        const float d = (float)(f1 / (double)L) - (float)(f0 / (double)L);
        s_flag = ua < fminf(1.0f, expf(d)) ? 1 : 0;
      }
      __syncthreads();
      if (s_flag)
        for (long long i = tid; i < L; i += THREADS) cur[i] = prop[i];
      __syncthreads();
    }
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) ready[node] = 1;
}

}  // namespace trex

extern "C" {

int trex_tree_adjacency(int n_leaves, int kind, float* adjacency, void* stream) {
  using namespace trex;
  TREX_CHECK_ARG(adjacency && n_leaves >= 1 && (kind == 0 || kind == 1), "trex_tree_adjacency: bad arguments");
  const long long n_all = 2LL * n_leaves - 1;
  tree_adjacency_kernel<<<(unsigned)((n_all * n_all + 255) / 256), 256, 0, (cudaStream_t)stream>>>(n_leaves, kind, adjacency);
  note_launches(1);
  TREX_CHECK_LAUNCH("tree_adjacency_kernel");
  return TREX_OK;
}

size_t trex_generate_groundtruth_workspace_bytes(int n_leaves, int n_mutations, int64_t n_sites) {
  if (n_leaves < 1 || n_mutations < 0 || n_sites < 1) return 0;
  const size_t n_edges = 2 * (size_t)n_leaves - 2, words = (size_t)((n_sites + 31) / 32);
  return 256 + sizeof(uint32_t) * n_edges * words + sizeof(int32_t) * n_edges * (size_t)n_mutations;
}

int trex_generate_groundtruth(int n_leaves, int n_states, int n_mutations, int64_t n_sites, int64_t site_stride,
                              uint64_t seed, const int32_t* mut_sites, const uint8_t* offsets, uint8_t* all_states,
                              float* adjacency, void* workspace, void* stream) {
  using namespace trex;
  TREX_CHECK_ARG(all_states && workspace, "trex_generate_groundtruth: null pointer");
  // ground_truth.py:142-144 raises ValueError for anything but a power of two
  TREX_CHECK_ARG(n_leaves > 0 && (n_leaves & (n_leaves - 1)) == 0, "trex_generate_groundtruth: n_leaves must be a power of 2.");
  TREX_CHECK_ARG(n_states >= 1 && n_states <= 255 && n_sites >= 1 && site_stride >= n_sites && n_sites < (1LL << 31),
                 "trex_generate_groundtruth: bad sizes");
  TREX_CHECK_ARG(n_mutations >= 0 && n_mutations <= n_sites, "trex_generate_groundtruth: cannot choose %d distinct sites of %lld",
                 n_mutations, (long long)n_sites);
  cudaStream_t s = (cudaStream_t)stream;
  const int n_edges = 2 * n_leaves - 2, words = (int)((n_sites + 31) / 32);
  uint32_t* mask = reinterpret_cast<uint32_t*>(workspace);
  int32_t* sites_ws = reinterpret_cast<int32_t*>(mask + (size_t)n_edges * words);
  int rc = check_cuda(cudaMemsetAsync(mask, 0, sizeof(uint32_t) * (size_t)n_edges * words, s), "memset mask");
  if (rc) return rc;
  const long long draws = (long long)n_edges * n_mutations;
  if (draws > 0) {
    if (!mut_sites) {
      gt_draw_sites_kernel<<<(unsigned)((draws + 255) / 256), 256, 0, s>>>(n_edges, n_mutations, (uint32_t)n_sites, seed, sites_ws);
      note_launches(1);
      TREX_CHECK_LAUNCH("gt_draw_sites_kernel");
      mut_sites = sites_ws;
    }
    gt_scatter_mask_kernel<<<(unsigned)((draws + 255) / 256), 256, 0, s>>>(n_edges, n_mutations, n_sites, mut_sites, mask, words);
    note_launches(1);
    TREX_CHECK_LAUNCH("gt_scatter_mask_kernel");
  }
  gt_walk_kernel<<<(unsigned)((n_sites + 255) / 256), 256, 0, s>>>(n_leaves, n_states, n_sites, site_stride, seed, mask, words,
                                                                    offsets, all_states);
  note_launches(1);
  TREX_CHECK_LAUNCH("gt_walk_kernel");
  if (adjacency) return trex_tree_adjacency(n_leaves, 0, adjacency, stream);
  return TREX_OK;
}

int trex_nk_landscape(int64_t n, int k, int n_states, uint64_t seed, int32_t* interactions, float* fitness_tables,
                      void* stream) {
  using namespace trex;
  TREX_CHECK_ARG((interactions || k == 0) && fitness_tables && n >= 1 && k >= 0 && k <= 6 && n_states >= 1 && n_states <= 255,
                 "trex_nk_landscape: bad arguments");
  long long table = 1;
  for (int j = 0; j <= k; ++j) table *= n_states;
  const long long total = n * (table > k ? table : k);
  nk_landscape_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(n, k, table, seed, interactions,
                                                                                       fitness_tables);
  note_launches(1);
  TREX_CHECK_LAUNCH("nk_landscape_kernel");
  return TREX_OK;
}

int trex_nk_fitness(const uint8_t* seqs, int64_t seq_stride, int batch, int64_t n, int k, int n_states,
                    const int32_t* interactions, const float* fitness_tables, const float* seq_mask, float* out_fitness,
                    void* stream) {
  using namespace trex;
  TREX_CHECK_ARG(seqs && (interactions || k == 0) && fitness_tables && out_fitness, "trex_nk_fitness: null pointer");
  TREX_CHECK_ARG(batch >= 1 && n >= 1 && seq_stride >= n && k >= 0 && k <= 6 && n_states >= 1, "trex_nk_fitness: bad sizes");
  long long table = 1;
  for (int j = 0; j <= k; ++j) table *= n_states;
  nk_fitness_kernel<256><<<batch, 256, 0, (cudaStream_t)stream>>>(seqs, seq_stride, n, k, n_states, table, interactions,
                                                                 fitness_tables, seq_mask, out_fitness);
  note_launches(1);
  TREX_CHECK_LAUNCH("nk_fitness_kernel");
  return TREX_OK;
}

size_t trex_nk_generate_tree_workspace_bytes(int n_nodes, int64_t site_stride) {
  if (n_nodes < 1 || site_stride < 1) return 0;
  return 256 + 4 * sizeof(int32_t) * (size_t)n_nodes + (size_t)n_nodes * (size_t)site_stride;
}

int trex_nk_generate_tree(const float* adjacency, int n_nodes, const uint8_t* root_sequence, int64_t n_sites,
                          int64_t site_stride, int k, int n_states, const int32_t* interactions,
                          const float* fitness_tables, float mutation_rate, float coupled_mutation_prob,
                          float mutation_rate_noise_std, int branch_length, uint64_t seed, const float* draw_noise,
                          const float* draw_u_coupled, const int32_t* draw_site, const float* draw_u_mask,
                          const uint8_t* draw_new_vals, const float* draw_u_accept, uint8_t* all_states, void* workspace,
                          void* stream) {
  using namespace trex;
  TREX_CHECK_ARG(adjacency && root_sequence && (interactions || k == 0) && fitness_tables && all_states && workspace,
                 "trex_nk_generate_tree: null pointer");
  TREX_CHECK_ARG(n_nodes >= 1 && n_sites >= 1 && site_stride >= n_sites && k >= 0 && k <= 6 && n_states >= 1 &&
                     n_states <= 255 && branch_length >= 0,
                 "trex_nk_generate_tree: bad sizes");
  cudaStream_t s = (cudaStream_t)stream;
  long long table = 1;
  for (int j = 0; j <= k; ++j) table *= n_states;
  int32_t* parent = reinterpret_cast<int32_t*>(workspace);
  int32_t* depth = parent + n_nodes;
  int32_t* order = depth + n_nodes;
  int32_t* ready = order + n_nodes;
  uint8_t* scratch = reinterpret_cast<uint8_t*>(ready + n_nodes);
  nk_parents_kernel<<<(n_nodes + 127) / 128, 128, 0, s>>>(adjacency, n_nodes, parent);
  note_launches(1);
  TREX_CHECK_LAUNCH("nk_parents_kernel");
  nk_order_kernel<<<1, 256, 0, s>>>(parent, n_nodes, depth, order, ready);
  note_launches(1);
  TREX_CHECK_LAUNCH("nk_order_kernel");
  const NkDraws dr{draw_noise, draw_u_coupled, draw_site, draw_u_mask, draw_new_vals, draw_u_accept};
  nk_evolve_kernel<256><<<n_nodes, 256, 0, s>>>(order, parent, n_nodes, n_sites, site_stride, k, n_states, table, interactions,
                                                fitness_tables, root_sequence, mutation_rate, coupled_mutation_prob,
                                                mutation_rate_noise_std, branch_length, seed, dr, all_states, scratch, ready);
  note_launches(1);
  TREX_CHECK_LAUNCH("nk_evolve_kernel");
  return TREX_OK;
}

}  // extern "C"
