// Tree schedule built on the device from the adjacency matrix.
//
// Replaces `jnp.where(adj[:, node] == 1, size=2, fill_value=-1)` evaluated inside every
// fori_loop iteration of the reference (src/trex/sankoff.py:60, :87-92).  The adjacency is a
// traced value under jit, so nothing here may touch the host.
#include <cstdarg>
#include <cstdio>

#include "common.cuh"
#include <atomic>

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

// ---- kernel 2: DFS post-order (one thread per tree; two linear passes, no stack) -----------
__global__ void plan_order_kernel(int n_all, int n_leaves, int32_t* __restrict__ plan,
                                  size_t plan_words) {
  const int n_int = n_all - n_leaves;
  int32_t* p = plan + (size_t)blockIdx.x * plan_words;
  if (threadIdx.x != 0) return;
  const int32_t* ch0 = p + kPlanHdr;
  const int32_t* ch1 = p + kPlanHdr + n_int;
  int32_t* order = p + kPlanHdr + 4 * n_int;
  int32_t* size = p + kPlanHdr + 5 * n_int;
  int32_t* off = p + kPlanHdr + 6 * n_int;
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
  int n_reach = 0;
  if (!(status & TREX_PLAN_NOT_A_TREE)) {
    for (int k = 0; k < n_int; ++k) off[k] = -1;
    // pass 2 (descending): post-order offsets; the larger internal subtree goes first so the
    // number of rows waiting for their sibling stays small (they are re-read from L2).
    int next = 0;
    for (int k = n_int - 1; k >= 0; --k) {
      int o = off[k];
      if (o < 0) {
        o = next;
        next += size[k];
      }
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
    }
    n_reach = size[n_int - 1];
  } else {
    // fallback: plain index order (always valid for the forward); reachable nodes first
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
  }
  p[0] = n_int;
  p[1] = n_reach;
  p[2] = 0;
  p[3] = status;
  p[4] = p[5] = p[6] = p[7] = 0;
}

}  // namespace trex

extern "C" {

const char* trex_last_error(void) { return trex::g_err; }
int trex_version(void) { return 100; }
unsigned long long trex_kernel_launches(void) { return trex::kernel_launches(); }

size_t trex_sankoff_plan_words(int n_all, int n_leaves) {
  const int n_int = n_all - n_leaves;
  if (n_int < 0) return 0;
  return (size_t)trex::kPlanHdr + 7 * (size_t)n_int;
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
  plan_order_kernel<<<batch, 32, 0, s>>>(n_all, n_leaves, plan, words);
  note_launches(1);
  TREX_CHECK_LAUNCH("plan_order_kernel");
  return TREX_OK;
}

}  // extern "C"
