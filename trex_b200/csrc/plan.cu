// Tree schedule built on the device from the adjacency matrix.
//
// Replaces `jnp.where(adj[:, node] == 1, size=2, fill_value=-1)` evaluated inside every
// fori_loop iteration of the reference (src/trex/sankoff.py:60, :87-92).  The adjacency is a
// traced value under jit, so nothing here may touch the host.
#include <cstdarg>
#include <cstdio>

#include "common.cuh"
#include <atomic>
#include <cstdlib>

namespace trex {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_cuda(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return TREX_OK;
  set_error("%s: %s", what, cudaGetErrorString(e));
  return TREX_ERR_CUDA;
}

static std::atomic<unsigned long long> g_kernel_launches{0};
void note_launches(int n) { g_kernel_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }
unsigned long long kernel_launches() { return g_kernel_launches.load(std::memory_order_relaxed); }

bool use_segments(int n_leaves) {
  const char* e = getenv("TREX_SEGMENTS");
  if (e && (e[0] == '0' || e[0] == '1')) return e[0] == '1';
  return n_leaves >= 64;
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

// ---- kernel 1: first two rows with adj[row, col] == 1 for every internal column -----------
// CTA = 32 columns x 8 row chunks; a warp reads 32 consecutive columns of one row (coalesced).
constexpr int kColsPerCta = 32;
constexpr int kRowChunks = 32;  // (8: 44 us per 2047-node adjacency, latency-bound on 32 CTAs of 256 threads)

__global__ void __launch_bounds__(kColsPerCta* kRowChunks)
    plan_children_kernel(const float* __restrict__ adj, int n_all, int n_leaves,
                         int zero_root_selfloop, int32_t* __restrict__ plan, size_t plan_words) {
  const int n_int = n_all - n_leaves;
  const int b = blockIdx.y;
  const float* a = adj + (size_t)b * n_all * n_all;
  int32_t* p = plan + (size_t)b * plan_words;
  const int c = threadIdx.x, r = threadIdx.y;
  const int k = blockIdx.x * kColsPerCta + c;  // internal node index
  const int col = n_leaves + k;
  __shared__ int found[kRowChunks][kColsPerCta][2];
  int f0 = -1, f1 = -1;
  if (k < n_int) {
    const int rows_per = (n_all + kRowChunks - 1) / kRowChunks;
    const int r0 = r * rows_per;
    const int r1 = min(n_all, r0 + rows_per);
    // eight independent loads in flight per thread (the scan is latency-bound); stop after the second hit
    for (int i0 = r0; i0 < r1 && f1 < 0; i0 += 8) {
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = i0 + u < r1 ? __ldg(a + (size_t)(i0 + u) * n_all + col) : 0.f;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = i0 + u;
        if (zero_root_selfloop && i == n_all - 1 && col == n_all - 1) v[u] = 0.f;  // sankoff.py:141
        if (v[u] == 1.0f && f1 < 0) {
          if (f0 < 0) f0 = i; else f1 = i;
        }
      }
    }
  }
  found[r][c][0] = f0;
  found[r][c][1] = f1;
  __syncthreads();
  if (r == 0 && k < n_int) {
    int c0 = -1, c1 = -1;
    for (int q = 0; q < kRowChunks && c1 < 0; ++q) {
      for (int e = 0; e < 2 && c1 < 0; ++e) {
        const int v = found[q][c][e];
        if (v >= 0) {
          if (c0 < 0) c0 = v; else c1 = v;
        }
      }
    }
    // raw ids as the reference stores them in its back-pointer table (sankoff.py:80,82)
    p[kPlanHdr + 2 * n_int + k] = c0;
    p[kPlanHdr + 3 * n_int + k] = c1;
    // canonical: a child row that is not yet computed when `col` is processed in index order
    // (id >= col, or the -1 fill that wraps to the root row) still holds the 1e5 fill.
    p[kPlanHdr + 0 * n_int + k] = (c0 >= 0 && c0 < col) ? c0 : -1;
    p[kPlanHdr + 1 * n_int + k] = (c1 >= 0 && c1 < col) ? c1 : -1;
  }
}

// ---- kernel 2: DFS post-order, subtree segments (one CTA per tree) ---------------------------
// Works on shared-memory copies of the per-node arrays when the tree fits (n_int <= kSmemNodes),
// directly on the plan's global arrays otherwise.  Subtree sizes are propagated bottom-up and post-order
// offsets top-down in parallel rounds (one round per tree level, all threads of the CTA; a balanced
// 1024-leaf tree takes 10 + 10 rounds instead of two 1023-step dependent chains); the few inherently
// ordered steps (numbering the segments, compacting the top list) are light single-thread loops.
constexpr int kSmemNodes = 8192;  // 6 int32 arrays of this many nodes = 192 KB of dynamic shared memory
constexpr int kPlanThreads = 256;
constexpr int kLeafSeenWords = 1024;  // shared-leaf detection covers the first 32 768 leaves

__global__ void __launch_bounds__(kPlanThreads)
    plan_order_kernel(int n_all, int n_leaves, int32_t* __restrict__ plan, size_t plan_words, int use_smem,
                      int seg_target) {
  extern __shared__ int32_t sm[];
  const int n_int = n_all - n_leaves;
  const int tid = threadIdx.x;
  int32_t* p = plan + (size_t)blockIdx.x * plan_words;
  int32_t* g_ch0 = p + kPlanHdr;
  int32_t* g_ch1 = p + kPlanHdr + n_int;
  int32_t* g_order = p + kPlanHdr + 4 * n_int;
  int32_t* g_size = p + kPlanHdr + 5 * n_int;
  int32_t* g_off = p + kPlanHdr + 6 * n_int;
  int32_t* g_top = p + kPlanHdr + 7 * n_int;
  int32_t* g_seg = p + kPlanHdr + 8 * n_int;  // seg_start[kMaxSeg], seg_len[kMaxSeg]
  const int32_t* ch0 = g_ch0;
  const int32_t* ch1 = g_ch1;
  int32_t *order = g_order, *size = g_size, *off = g_off;
  int32_t* par = p + kPlanHdr + 8 * n_int + 2 * kMaxSeg;  // parent (internal index) or -1; later: in-segment map
  if (use_smem) {
    int32_t* s0 = sm;
    int32_t* s1 = sm + n_int;
    for (int k = tid; k < n_int; k += kPlanThreads) { s0[k] = g_ch0[k]; s1[k] = g_ch1[k]; }
    ch0 = s0;
    ch1 = s1;
    order = sm + 2 * n_int;
    size = sm + 3 * n_int;
    off = sm + 4 * n_int;
    par = sm + 5 * n_int;
  }
  __shared__ int s_hdr[8];
  __shared__ int s_status;
  __shared__ unsigned s_leaf_seen[kLeafSeenWords];  // one bit per leaf: already the child of some node
  for (int w = tid; w < kLeafSeenWords; w += kPlanThreads) s_leaf_seen[w] = 0u;
  if (tid == 0) s_status = 0;
  for (int k = tid; k < n_int; k += kPlanThreads) { size[k] = 0; off[k] = -1; par[k] = -1; }
  __syncthreads();
  // parents; shared children / sentinel children
  for (int k = tid; k < n_int; k += kPlanThreads) {
    const int a = ch0[k], b = ch1[k];
    int st = 0;
    if (a < 0 || b < 0) st |= TREX_PLAN_HAS_SENTINEL_CHILD;
    if (a >= n_leaves && atomicExch(&par[a - n_leaves], k) != -1) st |= TREX_PLAN_NOT_A_TREE;
    if (b >= n_leaves && atomicExch(&par[b - n_leaves], k) != -1) st |= TREX_PLAN_NOT_A_TREE;
    // a leaf with two parents (or listed twice): harmless for the forward pass, but the reverse kernels STORE a
    // leaf's gradient row instead of accumulating it
    for (int c : {a, b})
      if (c >= 0 && c < n_leaves && c < 32 * kLeafSeenWords) {
        const unsigned bit = 1u << (c & 31);
        if (atomicOr(&s_leaf_seen[c >> 5], bit) & bit) st |= TREX_PLAN_SHARED_LEAF;
      }
    if (st) atomicOr(&s_status, st);
  }
  __syncthreads();
  // bottom-up rounds: a node's size is known once its internal children's are (children have smaller
  // indices, so this terminates; also valid on a DAG)
  for (;;) {
    int changed = 0;
    for (int k = tid; k < n_int; k += kPlanThreads) {
      if (size[k]) continue;
      const int a = ch0[k], b = ch1[k];
      const int sa = a >= n_leaves ? size[a - n_leaves] : -1, sb = b >= n_leaves ? size[b - n_leaves] : -1;
      if (sa == 0 || sb == 0) continue;
      size[k] = 1 + (sa > 0 ? sa : 0) + (sb > 0 ? sb : 0);
      changed = 1;
    }
    if (!__syncthreads_or(changed)) break;
  }
  const int status = s_status;
  if (!(status & TREX_PLAN_NOT_A_TREE)) {
    const int n_reach = size[n_int - 1];
    const int cap = max(1, (n_reach + seg_target - 1) / seg_target);
    // offsets of the forest roots: the root first (its subtree is order[0 .. n_reach)); roots the root
    // does not reach take disjoint ranges behind it (any order is valid for them)
    __shared__ int s_next;
    if (tid == 0) { s_next = n_reach; off[n_int - 1] = 0; }
    __syncthreads();
    for (int k = tid; k < n_int - 1; k += kPlanThreads)
      if (par[k] < 0) off[k] = atomicAdd(&s_next, size[k]);
    // top-down rounds: a node with a known offset places itself and hands offsets to its children; the
    // larger internal subtree goes first so that few rows wait for their sibling (re-read from L2).
    // order[] doubles as the "expanded" marker: it is filled exactly once per node.
    for (int k = tid; k < n_int; k += kPlanThreads) order[k] = -1;
    __syncthreads();
    for (;;) {
      int changed = 0;
      for (int k = tid; k < n_int; k += kPlanThreads) {
        int o = off[k];
        if (o < 0 || order[o + size[k] - 1] == n_leaves + k) continue;
        order[o + size[k] - 1] = n_leaves + k;
        int a = ch0[k] >= n_leaves ? ch0[k] - n_leaves : -1;
        int b = ch1[k] >= n_leaves ? ch1[k] - n_leaves : -1;
        if (a >= 0 && b >= 0 && size[b] > size[a]) {
          const int t = a; a = b; b = t;
        }
        if (a >= 0) {
          off[a] = o;
          o += size[a];
        }
        if (b >= 0) off[b] = o;
        changed = 1;
      }
      if (!__syncthreads_or(changed)) break;
    }
    // Ordered compactions over post-order positions (thread t owns a contiguous chunk of positions;
    // chunk counts are combined with a block-wide exclusive scan).
    __shared__ int s_scan[kPlanThreads];
    const int chunk = (n_int + kPlanThreads - 1) / kPlanThreads;
    const int i_lo = min(n_int, tid * chunk), i_hi = min(n_int, i_lo + chunk);
    auto block_exclusive = [&](int count, int& total) -> int {
      s_scan[tid] = count;
      __syncthreads();
      for (int d = 1; d < kPlanThreads; d <<= 1) {
        const int v = tid >= d ? s_scan[tid - d] : 0;
        __syncthreads();
        s_scan[tid] += v;
        __syncthreads();
      }
      total = s_scan[kPlanThreads - 1];
      const int excl = s_scan[tid] - count;
      __syncthreads();
      return excl;
    };
    // segments: maximal subtrees of at most `cap` internal nodes inside the root's subtree, numbered by
    // position; at most kMaxSeg (later ones stay in the top part)
    auto is_seg_root = [&](int i) {
      if (n_reach <= 1 || i >= n_reach) return false;
      const int k = order[i] - n_leaves;
      return size[k] <= cap && (par[k] < 0 || size[par[k]] > cap);
    };
    int cnt = 0;
    for (int i = i_lo; i < i_hi; ++i) cnt += is_seg_root(i) ? 1 : 0;
    int total_seg;
    int idx = block_exclusive(cnt, total_seg);
    for (int i = i_lo; i < i_hi; ++i) {
      if (!is_seg_root(i)) continue;
      if (idx < kMaxSeg) {
        const int k = order[i] - n_leaves;
        g_seg[idx] = off[k];
        g_seg[kMaxSeg + idx] = size[k];
      }
      ++idx;
    }
    const int n_seg = min(total_seg, kMaxSeg);
    __syncthreads();
    // par[] is no longer needed: reuse it as the "position is inside a segment" map
    for (int k = tid; k < n_int; k += kPlanThreads) par[k] = 0;
    __syncthreads();
    for (int sgi = tid / 32; sgi < n_seg; sgi += kPlanThreads / 32) {  // one warp per segment
      const int s0 = g_seg[sgi], l0 = g_seg[kMaxSeg + sgi];
      for (int i = tid % 32; i < l0; i += 32) par[s0 + i] = 1;
    }
    __syncthreads();
    // top part: every node outside the segments in post-order, reachable ones first
    cnt = 0;
    int cnt_reach = 0;
    for (int i = i_lo; i < i_hi; ++i) {
      if (par[i]) continue;
      ++cnt;
      if (i < n_reach) ++cnt_reach;
    }
    int n_top, n_top_reach;
    idx = block_exclusive(cnt, n_top);
    (void)block_exclusive(cnt_reach, n_top_reach);
    for (int i = i_lo; i < i_hi; ++i)
      if (!par[i]) g_top[idx++] = order[i];
    if (tid == 0) {
      s_hdr[1] = n_reach;
      s_hdr[2] = n_seg;
      s_hdr[4] = n_top_reach;
      s_hdr[5] = n_top;
    }
  } else if (tid == 0) {
    // not a tree (a child with two parents): plain index order (always valid for the forward walk),
    // reachable nodes first; no segments
    int n_reach = 0;
    for (int k = 0; k < n_int; ++k) off[k] = 0;
    off[n_int - 1] = 1;
    for (int k = n_int - 1; k >= 0; --k) {
      if (!off[k]) continue;
      ++n_reach;
      if (ch0[k] >= n_leaves) off[ch0[k] - n_leaves] = 1;
      if (ch1[k] >= n_leaves) off[ch1[k] - n_leaves] = 1;
    }
    int w0 = 0, w1 = n_reach;
    for (int k = 0; k < n_int; ++k) {
      if (off[k]) order[w0++] = n_leaves + k; else order[w1++] = n_leaves + k;
    }
    for (int i = 0; i < n_int; ++i) g_top[i] = order[i];
    s_hdr[1] = n_reach;
    s_hdr[2] = 0;
    s_hdr[4] = n_reach;
    s_hdr[5] = n_int;
  }
  if (tid == 0) {
    s_hdr[0] = n_int;
    s_hdr[3] = status;
    s_hdr[6] = s_hdr[7] = 0;
  }
  __syncthreads();
  if (tid < 8) p[tid] = s_hdr[tid];
  if (use_smem) {
    for (int k = tid; k < n_int; k += kPlanThreads) {
      g_order[k] = order[k];
      g_size[k] = size[k];
      g_off[k] = off[k];
    }
  }
}

}  // namespace trex

extern "C" {

const char* trex_last_error(void) { return trex::g_err; }
int trex_version(void) { return 100; }
unsigned long long trex_kernel_launches(void) { return trex::kernel_launches(); }

size_t trex_sankoff_plan_words(int n_all, int n_leaves) {
  const int n_int = n_all - n_leaves;
  if (n_int < 0) return 0;
  return (size_t)trex::kPlanHdr + 9 * (size_t)n_int + 2 * (size_t)trex::kMaxSeg;  // (+ n_int words of builder scratch)
}

// Segment count the walk kernels get the most out of: measured on B200 in round 2 (tools/probe_step.py, 1024 taxa,
// A = 20, ms per step = hard fwd + traceback + relaxed fwd + bwd; profiles/r02_j_probe_step.txt):
//   12.5k sites: 8 segments 2.07, 16: 1.82, 32: 1.93      25k sites: 3.31  3.28  3.33      100k: flat between 8 and 32
// (the forward kernels alone would take 8 on short alignments, the reverse kernel 16; one plan serves them all).
static int segment_target(long long n_sites, int batch) {
  (void)n_sites;
  (void)batch;
  int target = 16;
  if (const char* e = getenv("TREX_SEG_TARGET")) {  // experiments / tests
    const int v = atoi(e);
    if (v >= 1 && v <= trex::kMaxSeg) target = v;
  }
  return target;
}

int trex_sankoff_plan_for_sites(const float* adj, int n_all, int n_leaves, int batch, int zero_root_selfloop,
                                int64_t n_sites, int32_t* plan, void* stream) {
  using namespace trex;
  TREX_CHECK_ARG(adj && plan, "trex_sankoff_plan: null pointer");
  TREX_CHECK_ARG(n_leaves >= 1 && n_all > n_leaves, "trex_sankoff_plan: need n_all > n_leaves >= 1 (got %d, %d)", n_all, n_leaves);
  TREX_CHECK_ARG(batch >= 1 && batch <= 65535, "trex_sankoff_plan: batch %d out of range", batch);
  const int n_int = n_all - n_leaves;
  const size_t words = trex_sankoff_plan_words(n_all, n_leaves);
  cudaStream_t s = (cudaStream_t)stream;
  dim3 grid((n_int + kColsPerCta - 1) / kColsPerCta, batch);
  dim3 block(kColsPerCta, kRowChunks);
  plan_children_kernel<<<grid, block, 0, s>>>(adj, n_all, n_leaves, zero_root_selfloop, plan, words);
  note_launches(1);
  TREX_CHECK_LAUNCH("plan_children_kernel");
  const int use_smem = n_int <= kSmemNodes;
  const size_t smem = use_smem ? 6 * sizeof(int32_t) * (size_t)n_int : 0;
  if (smem > 48 * 1024) {
    int rc = check_cuda(cudaFuncSetAttribute(plan_order_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                        "cudaFuncSetAttribute(plan_order_kernel)");
    if (rc) return rc;
  }
  plan_order_kernel<<<batch, kPlanThreads, smem, s>>>(n_all, n_leaves, plan, words, use_smem,
                                                     segment_target((long long)n_sites, batch));
  note_launches(1);
  TREX_CHECK_LAUNCH("plan_order_kernel");
  return TREX_OK;
}

int trex_sankoff_plan(const float* adj, int n_all, int n_leaves, int batch,
                      int zero_root_selfloop, int32_t* plan, void* stream) {
  return trex_sankoff_plan_for_sites(adj, n_all, n_leaves, batch, zero_root_selfloop, 0, plan, stream);
}

}  // extern "C"
