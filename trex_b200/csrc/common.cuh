// Shared helpers for the libtrex_b200 kernels (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <type_traits>

#include "../../include/trex_b200.h"

namespace trex {

constexpr float kSentinel = 1e5f;  // src/trex/sankoff.py:152
constexpr int kPlanHdr = 8;        // int32 header words per tree plan
constexpr int kMaxSeg = 128;       // independent subtrees ("segments") a plan can expose
constexpr int kSegTarget = 32;     // the builder aims at this many segments of ~n_reach / kSegTarget nodes

// plan word offsets (per tree); n_int = n_all - n_leaves
struct PlanView {
  const int32_t* base;
  int n_int;
  __host__ __device__ PlanView(const int32_t* b, int n) : base(b), n_int(n) {}
  __device__ const int32_t* hdr() const { return base; }
  __device__ const int32_t* child(int k) const { return base + kPlanHdr + k * n_int; }      // k=0,1
  __device__ const int32_t* raw(int k) const { return base + kPlanHdr + (2 + k) * n_int; }  // k=0,1
  __device__ const int32_t* order() const { return base + kPlanHdr + 4 * n_int; }
  // Subtree decomposition for short alignments (few sites -> few warps): maximal subtrees of at most
  // ~n_reach / kSegTarget internal nodes are contiguous runs order[seg_start .. +seg_len) of the
  // post-order and independent of each other; top_order lists the remaining nodes in post-order
  // (reachable ones first, hdr[4] of them; hdr[5] in total with the nodes the root does not reach).
  __device__ const int32_t* top_order() const { return base + kPlanHdr + 7 * n_int; }
  __device__ const int32_t* seg_start() const { return base + kPlanHdr + 8 * n_int; }
  __device__ const int32_t* seg_len() const { return base + kPlanHdr + 8 * n_int + kMaxSeg; }
};

// The node list one kernel launch walks.  phase 0: the whole tree (order, all n_int nodes forward /
// the n_reach reachable ones in reverse); phase 1: segment `z` of the plan; phase 2: the top part.
struct Walk {
  const int32_t* ord;
  int n_fwd;  // nodes a forward (post-order) walk visits
  int n_rev;  // nodes a reverse walk visits: ord[n_rev-1] down to ord[0]
};
__device__ __forceinline__ Walk make_walk(const PlanView& plan, int phase, int z) {
  Walk w;
  const int32_t* h = plan.hdr();
  if (phase == 0) {
    w.ord = plan.order();
    w.n_fwd = plan.n_int;
    w.n_rev = h[1];
  } else if (phase == 1) {
    const bool on = z < h[2];
    w.ord = plan.order() + (on ? plan.seg_start()[z] : 0);
    w.n_fwd = w.n_rev = on ? plan.seg_len()[z] : 0;
  } else {
    w.ord = plan.top_order();
    w.n_fwd = h[5];
    w.n_rev = h[4];
  }
  return w;
}

void set_error(const char* fmt, ...);
int check_cuda(cudaError_t e, const char* what);
int sm_count();
void note_launches(int n);  // process-wide count of kernel launches (trex_kernel_launches)
// Walk the plan's independent subtrees in parallel (two launches: segments, then the top part) instead
// of the whole tree in one launch.  On short alignments this supplies the parallelism the sites cannot
// (12.5k sites: 5x faster); on long ones the many small work items balance the SMs better than one
// resident wave of site blocks (100k sites: 8-40% faster per kernel).  Small trees have nothing to split.
// TREX_SEGMENTS=0/1 forces the choice (tests).
bool use_segments(int n_leaves);

#define TREX_CHECK_ARG(cond, ...)        \
  do {                                   \
    if (!(cond)) {                       \
      ::trex::set_error(__VA_ARGS__);    \
      return TREX_ERR_INVALID_ARGUMENT;  \
    }                                    \
  } while (0)

#define TREX_CHECK_LAUNCH(what)                                                 \
  do {                                                                          \
    int _rc = ::trex::check_cuda(cudaPeekAtLastError(), what);                  \
    if (_rc) return _rc;                                                        \
  } while (0)

// ---- vector load/store of V consecutive sites --------------------------------------------
template <int V> struct VecF;
template <> struct VecF<1> { using type = float; };
template <> struct VecF<2> { using type = float2; };
template <> struct VecF<4> { using type = float4; };

template <int V>
__device__ __forceinline__ void ld_f(const float* __restrict__ p, float (&out)[V]) {
  if constexpr (V == 1) {
    out[0] = *p;
  } else if constexpr (V == 2) {
    float2 t = *reinterpret_cast<const float2*>(p);
    out[0] = t.x; out[1] = t.y;
  } else {
    float4 t = *reinterpret_cast<const float4*>(p);
    out[0] = t.x; out[1] = t.y; out[2] = t.z; out[3] = t.w;
  }
}

// row r of a [row][site] table whose rows are `row_bytes` apart: base + r * row_bytes as ONE 32x32->64 multiply-add
// (IMAD.WIDE.U32) instead of the 64-bit multiply a `long long` stride costs (5-6 instructions per address).
// Entry points check site_stride < 2^29, so a row pitch in bytes always fits 32 bits.
template <typename T>
__device__ __forceinline__ T* row_ptr(T* base, unsigned row_bytes, unsigned r) {
  using B = typename std::conditional<std::is_const<T>::value, const char, char>::type;
  return reinterpret_cast<T*>(reinterpret_cast<B*>(base) + (size_t)row_bytes * r);
}

// 2^x, one MUFU.EX2 (flush-to-zero: a term below 2^-126 of the largest one is nothing at float32)
__device__ __forceinline__ float ex2_ftz(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// log2(x), one MUFU.LG2, for x that is never subnormal (callers say why)
__device__ __forceinline__ float lg2_ftz(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// streaming read of data that is not re-read by this kernel (evict-first, bypass L1)
template <int V>
__device__ __forceinline__ void ld_f_stream(const float* __restrict__ p, float (&out)[V]) {
  if constexpr (V == 1) {
    out[0] = __ldcs(p);
  } else if constexpr (V == 2) {
    float2 t = __ldcs(reinterpret_cast<const float2*>(p));
    out[0] = t.x; out[1] = t.y;
  } else {
    float4 t = __ldcs(reinterpret_cast<const float4*>(p));
    out[0] = t.x; out[1] = t.y; out[2] = t.z; out[3] = t.w;
  }
}

template <int V>
__device__ __forceinline__ void st_f(float* __restrict__ p, const float (&v)[V]) {
  if constexpr (V == 1) {
    *p = v[0];
  } else if constexpr (V == 2) {
    *reinterpret_cast<float2*>(p) = make_float2(v[0], v[1]);
  } else {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  }
}

// write-once outputs that this kernel never reads back (streaming store)
template <int V>
__device__ __forceinline__ void st_f_stream(float* __restrict__ p, const float (&v)[V]) {
  if constexpr (V == 1) {
    __stcs(p, v[0]);
  } else if constexpr (V == 2) {
    __stcs(reinterpret_cast<float2*>(p), make_float2(v[0], v[1]));
  } else {
    __stcs(reinterpret_cast<float4*>(p), make_float4(v[0], v[1], v[2], v[3]));
  }
}

template <int V>
__device__ __forceinline__ void ld_u8(const uint8_t* __restrict__ p, int (&out)[V]) {
  if constexpr (V == 1) {
    out[0] = *p;
  } else if constexpr (V == 2) {
    uchar2 t = *reinterpret_cast<const uchar2*>(p);
    out[0] = t.x; out[1] = t.y;
  } else {
    uchar4 t = *reinterpret_cast<const uchar4*>(p);
    out[0] = t.x; out[1] = t.y; out[2] = t.z; out[3] = t.w;
  }
}

// The V state bytes of a lane as ONE raw word, unpacked later: a prefetch that is consumed a node ahead must not
// convert the bytes at the load (the conversion would wait for the load right there).
template <int V>
__device__ __forceinline__ unsigned ld_u8_raw(const uint8_t* __restrict__ p) {
  if constexpr (V == 1) return *p;
  else if constexpr (V == 2) return *reinterpret_cast<const unsigned short*>(p);
  else return *reinterpret_cast<const unsigned*>(p);
}
template <int V>
__device__ __forceinline__ void unpack_u8(unsigned raw, int (&out)[V]) {
#pragma unroll
  for (int v = 0; v < V; ++v) out[v] = (int)((raw >> (8 * v)) & 0xffu);
}

template <int V>
__device__ __forceinline__ void st_u8(uint8_t* __restrict__ p, const int (&v)[V]) {
  if constexpr (V == 1) {
    *p = (uint8_t)v[0];
  } else if constexpr (V == 2) {
    *reinterpret_cast<uchar2*>(p) = make_uchar2((uint8_t)v[0], (uint8_t)v[1]);
  } else {
    *reinterpret_cast<uchar4*>(p) =
        make_uchar4((uint8_t)v[0], (uint8_t)v[1], (uint8_t)v[2], (uint8_t)v[3]);
  }
}

template <int V>
__device__ __forceinline__ void st_u8_stream(uint8_t* __restrict__ p, const int (&v)[V]) {
  if constexpr (V == 1) {
    __stcs(p, (uint8_t)v[0]);
  } else if constexpr (V == 2) {
    __stcs(reinterpret_cast<uchar2*>(p), make_uchar2((uint8_t)v[0], (uint8_t)v[1]));
  } else {
    __stcs(reinterpret_cast<uchar4*>(p),
           make_uchar4((uint8_t)v[0], (uint8_t)v[1], (uint8_t)v[2], (uint8_t)v[3]));
  }
}

// ---- back-pointer (argmin) table: [tree][k][state][site][slot] u8 --------------------------------
// The two children's back-pointers of a (state, site) are adjacent bytes: the forward kernels store the 2V
// bytes of a lane with ONE 16/32/64-bit store per state row (128 B per warp and row at V = 2 instead of two
// 64 B rows), and the traceback fetches both children's pointers with one 16-bit load.
// p points at (state row, site0, slot 0); a / b are the indices of the two children, a_is_slot0 says which
// of them is the first child (slot 0 = first child, sankoff.py:80-83).
template <int V>
__device__ __forceinline__ void st_arg_pair(uint8_t* __restrict__ p, const int (&a)[V], const int (&b)[V],
                                            bool a_is_slot0) {
  if constexpr (V == 1) {
    const unsigned w = __byte_perm((unsigned)a[0], (unsigned)b[0], a_is_slot0 ? 0x0040u : 0x0004u);
    __stcs(reinterpret_cast<unsigned short*>(p), (unsigned short)w);
  } else if constexpr (V == 2) {
    const unsigned pa = (unsigned)a[0] | ((unsigned)a[1] << 8), pb = (unsigned)b[0] | ((unsigned)b[1] << 8);
    __stcs(reinterpret_cast<unsigned*>(p), __byte_perm(pa, pb, a_is_slot0 ? 0x5140u : 0x1504u));
  } else {
    const unsigned pa = (unsigned)a[0] | ((unsigned)a[1] << 8) | ((unsigned)a[2] << 16) | ((unsigned)a[3] << 24);
    const unsigned pb = (unsigned)b[0] | ((unsigned)b[1] << 8) | ((unsigned)b[2] << 16) | ((unsigned)b[3] << 24);
    const unsigned lo = __byte_perm(pa, pb, a_is_slot0 ? 0x5140u : 0x1504u);
    const unsigned hi = __byte_perm(pa, pb, a_is_slot0 ? 0x7362u : 0x3726u);
    __stcs(reinterpret_cast<uint2*>(p), make_uint2(lo, hi));
  }
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// block-wide sum of doubles -> one atomicAdd per block (blockDim.x multiple of 32, <= 1024)
__device__ __forceinline__ void block_atomic_add(double v, double* dst) {
  __shared__ double red[32];
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) red[w] = v;
  __syncthreads();
  if (w == 0) {
    const int nw = (blockDim.x + 31) >> 5;
    v = lane < nw ? red[lane] : 0.0;
    v = warp_sum(v);
    if (lane == 0) atomicAdd(dst, v);
  }
}

}  // namespace trex
