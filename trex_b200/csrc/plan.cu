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

bool use_segments(long long n_site_warps) {
  const char* e = getenv("TREX_SEGMENTS");
  if (e && (e[0] == '0' || e[0] == '1')) return e[0] == '1';
  return n_site_warps < 6LL * sm_count();
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
constexpr int kRowChunks = 8;

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
    for (int i = r0; i < r1 && f1 < 0; ++i) {
      float v = __ldg(a + (size_t)i * n_all + col);
      if (zero_root_selfloop && i == n_all - 1 && col == n_all - 1) v = 0.f;  // sankoff.py:141
      if (v == 1.0f) {
        if (f0 < 0) f0 = i; else f1 = i;
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
// The passes are sequential (thread 0) but run on shared-memory copies of the child arrays when the
// tree fits (n_int <= kSmemNodes), which takes the dependent-load latency from ~L2 to ~30 cycles per
// node; larger trees use the global arrays directly.  Two linear passes, no stack.
constexpr int kSmemNodes = 8192;  // 5 int32 arrays of this many nodes = 160 KB of dynamic shared memory

__global__ void __launch_bounds__(256)
    plan_order_kernel(int n_all, int n_leaves, int32_t* __restrict__ plan, size_t plan_words, int use_smem) {
  extern __shared__ int32_t sm[];
  const int n_int = n_all - n_leaves;
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
  if (use_smem) {
    int32_t* s0 = sm;
    int32_t* s1 = sm + n_int;
    for (int k = threadIdx.x; k < n_int; k += blockDim.x) { s0[k] = g_ch0[k]; s1[k] = g_ch1[k]; }
    ch0 = s0;
    ch1 = s1;
    order = sm + 2 * n_int;
    size = sm + 3 * n_int;
    off = sm + 4 * n_int;
    __syncthreads();
  }
  __shared__ int s_hdr[8];
  if (threadIdx.x == 0) {
    int status = 0;
    for (int k = 0; k < n_int; ++k) off[k] = -1;
    // pass 1 (ascending): subtree sizes in internal nodes; detect shared children / sentinels
    for (int k = 0; k < n_int; ++k) {
      const int a = ch0[k], b = ch1[k];
      int s = 1;
      if (a < 0 || b < 0) status |= TREX_PLAN_HAS_SENTINEL_CHILD;
      if (a >= n_leaves) {
        s += size[a - n_leaves];
        if (off[a - n_leaves] == -2) status |= TREX_PLAN_NOT_A_TREE;
        off[a - n_leaves] = -2;
      }
      if (b >= n_leaves) {
        s += size[b - n_leaves];
        if (off[b - n_leaves] == -2) status |= TREX_PLAN_NOT_A_TREE;
        off[b - n_leaves] = -2;
      }
      size[k] = s;
    }
    int n_reach = 0, n_seg = 0, n_top_reach = 0, n_top = 0;
    if (!(status & TREX_PLAN_NOT_A_TREE)) {
      for (int k = 0; k < n_int; ++k) off[k] = -1;
      // pass 2 (descending): post-order offsets; the larger internal subtree goes first so the
      // number of rows waiting for their sibling stays small (they are re-read from L2).
      // Segments: a node whose subtree has at most `cap` internal nodes while its parent's has more
      // roots a segment (its whole subtree is the contiguous run order[off .. off + size)).  `off`
      // of a node inside a segment is flagged by adding kInSeg to its parent's decision below.
      n_reach = size[n_int - 1];
      const int cap = max(1, (n_reach + kSegTarget - 1) / kSegTarget);
      int next = 0;
      for (int k = n_int - 1; k >= 0; --k) {
        int o = off[k];
        bool in_seg = false;
        if (o <= -4) {  // parent already belongs to a segment: encoded offset
          in_seg = true;
          o = -o - 4;
        } else if (o < 0) {  // not reached from a parent: a root of the forest
          o = next;
          next += size[k];
        }
        const bool reach = o < n_reach;  // the root's subtree occupies order[0 .. n_reach)
        if (!in_seg && reach && size[k] <= cap && n_seg < kMaxSeg && n_reach > 1) {
          g_seg[n_seg] = o;
          g_seg[kMaxSeg + n_seg] = size[k];
          ++n_seg;
          in_seg = true;
        }
        order[o + size[k] - 1] = n_leaves + k;
        if (in_seg) size[k] = -size[k];  // sign = "inside a segment" (restored below)
        int a = ch0[k] >= n_leaves ? ch0[k] - n_leaves : -1;
        int b = ch1[k] >= n_leaves ? ch1[k] - n_leaves : -1;
        if (a >= 0 && b >= 0 && size[b] > size[a]) {
          const int t = a; a = b; b = t;
        }
        if (a >= 0) {
          off[a] = in_seg ? -(o + 4) : o;
          o += size[a];
        }
        if (b >= 0) off[b] = in_seg ? -(o + 4) : o;
      }
      // top part: every node outside the segments, in post-order (= ascending position in `order`):
      // reachable ones first, then the nodes the root does not reach (forward walks still fill them)
      for (int i = 0; i < n_int; ++i) {
        const int k = order[i] - n_leaves;
        if (size[k] > 0) {
          g_top[n_top++] = order[i];
          if (i < n_reach) ++n_top_reach;
        }
      }
      for (int k = 0; k < n_int; ++k) {
        size[k] = size[k] < 0 ? -size[k] : size[k];
        off[k] = off[k] <= -4 ? -off[k] - 4 : off[k];
      }
    } else {
      // fallback: plain index order (always valid for the forward); reachable nodes first; no segments
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
      n_top = n_int;
      n_top_reach = n_reach;
    }
    s_hdr[0] = n_int;
    s_hdr[1] = n_reach;
    s_hdr[2] = n_seg;
    s_hdr[3] = status;
    s_hdr[4] = n_top_reach;
    s_hdr[5] = n_top;
    s_hdr[6] = s_hdr[7] = 0;
  }
  __syncthreads();
  if (threadIdx.x < 8) p[threadIdx.x] = s_hdr[threadIdx.x];
  if (use_smem) {
    for (int k = threadIdx.x; k < n_int; k += blockDim.x) {
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
  return (size_t)trex::kPlanHdr + 8 * (size_t)n_int + 2 * (size_t)trex::kMaxSeg;
}

int trex_sankoff_plan(const float* adj, int n_all, int n_leaves, int batch,
                      int zero_root_selfloop, int32_t* plan, void* stream) {
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
  const size_t smem = use_smem ? 5 * sizeof(int32_t) * (size_t)n_int : 0;
  if (smem > 48 * 1024) {
    int rc = check_cuda(cudaFuncSetAttribute(plan_order_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                        "cudaFuncSetAttribute(plan_order_kernel)");
    if (rc) return rc;
  }
  plan_order_kernel<<<batch, 256, smem, s>>>(n_all, n_leaves, plan, words, use_smem);
  note_launches(1);
  TREX_CHECK_LAUNCH("plan_order_kernel");
  return TREX_OK;
}

}  // extern "C"
