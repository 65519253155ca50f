// Sankoff forward kernels for a UNIFORM substitution cost: C[s][s] = 0, C[s][t] = a for s != t
// (Fitch-style parsimony; the unit cost 1 - I of the reference's tests and benchmarks,
// tests/test_sankoff.py:50, tests/test_convergence.py:47).  Same recurrence, outputs and layouts as
// sankoff_hard.cu / sankoff_soft_fwd.cu (src/trex/sankoff.py:24-97, 141-160, 187), but
//     min_t (C[s,t] + d[t])      = min(d[s], a + min_t d[t])                         (hard)
//     sum_t K[s,t] e[t]          = e[s] + k (sum_t e[t] - e[s]),  k = exp(-a/T)       (relaxed)
// so a parent row costs O(A) instead of O(A^2) and the kernels are bound by HBM, not by issue rate.
// The hard form reproduces jnp.argmin's first-minimum rule exactly:
//     with (m1, i1) the first minimum of d,   candidate "self" = (d[s], s),  "other" = (a + m1, i1);
//     self wins when d[s] < a + m1, or on a tie when s <= i1      (for s == i1 self always wins).
// Leaf rows (1e5 with a 0 at the observed state, sankoff.py:49-52,152) are materialised in
// registers and take the same code path, so no lookup table is needed.
//
// Dispatch is on the device (the cost matrix is device data under jit): these kernels return at
// once unless the matrix is uniform, the general kernels return at once when it is.
#include "fwd_params.cuh"
#include <utility>

namespace trex {

namespace {

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ float lg2_approx(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// child row as the recurrence sees it: internal -> dp row, leaf / missing -> 1e5 with 0 at the state
template <int AP, int V, bool FULL = false>
__device__ __forceinline__ void load_child_row(const float* __restrict__ dp, const uint8_t* __restrict__ leaf,
                                               int c, int n_leaves, int A, size_t stride, float (&d)[AP][V]) {
  if (c >= n_leaves) {
    const float* src = dp + (size_t)(c - n_leaves) * A * stride;
#pragma unroll
    for (int t = 0; t < AP; ++t) {
      if (FULL || t < A) ld_f<V>(src + (size_t)t * stride, d[t]);
      else {
#pragma unroll
        for (int v = 0; v < V; ++v) d[t][v] = INFINITY;
      }
    }
  } else {
    int o[V];
#pragma unroll
    for (int v = 0; v < V; ++v) o[v] = A;  // missing child (-1 fill): all-1e5 row (sankoff.py:60,152)
    if (c >= 0) ld_u8<V>(leaf + (size_t)c * stride, o);
#pragma unroll
    for (int t = 0; t < AP; ++t)
#pragma unroll
      for (int v = 0; v < V; ++v) d[t][v] = (FULL || t < A) ? (o[v] == t ? 0.f : kSentinel) : INFINITY;
  }
}

template <int... I, class F>
__device__ __forceinline__ void static_for_impl(std::integer_sequence<int, I...>, F&& f) {
  (f(std::integral_constant<int, I>{}), ...);
}
// f(integral_constant<int, 0>) ... f(integral_constant<int, N-1>): a loop whose index is a constant expression
template <int N, class F>
__device__ __forceinline__ void static_for(F&& f) {
  static_for_impl(std::make_integer_sequence<int, N>{}, f);
}

// argmin of {self = (d, s), other = (am, j)} under jnp.argmin's first-minimum rule: self wins when
// strictly smaller, or on a tie when s is not after j.  Three chained SETPs + one SELP.
template <int S>
__device__ __forceinline__ int pick_index(float d, float am, int j) {
  int idx;
  asm("{\n\t.reg .pred p;\n\t"
      "setp.le.s32 p, %3, %4;\n\t"
      "setp.eq.and.f32 p, %1, %2, p;\n\t"
      "setp.lt.or.f32 p, %1, %2, p;\n\t"
      "selp.s32 %0, %3, %4, p;\n\t}"
      : "=r"(idx) : "f"(d), "f"(am), "n"(S), "r"(j));
  return idx;
}

// The same decision with ONE compare when the scores are small integers (they are integers whenever this
// kernel runs: hard_uniform_applies demands an integral cost, and leaf rows are 0 / 1e5): with
// key(d, s) = d + s/64, exact for am < 2^17, "self wins" <=> key(d, S) <= key(am, j).  d itself may be large
// (a lineage of stateless leaves): rounding is monotone, so a d >= 2^18 still compares above any valid key.
// The add runs on the FMA pipe; the ALU pipe — the one this kernel saturates — sees FSETP + SEL instead of
// three SETPs + SEL.
template <int S>
__device__ __forceinline__ int pick_index_key(float d, float keyj, int j) {
  int idx;
  asm("{\n\t.reg .pred p;\n\t.reg .f32 k;\n\t"
      "add.f32 k, %1, %2;\n\t"
      "setp.le.f32 p, k, %3;\n\t"
      "selp.s32 %0, %4, %5, p;\n\t}"
      : "=r"(idx) : "f"(d), "f"((float)S * 0.015625f), "f"(keyj), "n"(S), "r"(j));
  return idx;
}

// Compact back-pointers ("pointer words", TREX_MODE_COMPACT_ARG): under a uniform cost the pointer of parent state s is
// either s itself ("stay") or the child's first minimum j, whatever s is — so one 32-bit word per (node, site, child)
// holds all of them: bits 0-4 = j, bit 5+s = "state s stays" (n_states <= 27).  2 * n_states bytes of table per node
// and site become 8 (rows 0-3 of the node's byte table, the other rows are not written), and the traceback reads those
// 4 rows coalesced instead of gathering 2 bytes from one of n_states rows (32-byte sectors: 6.8x its useful bytes).  Same decisions as pick_index / pick_index_key: the bit is set exactly when they return S.
template <int S>
__device__ __forceinline__ void stay_bit(float d, float am, int j, unsigned& mask) {
  asm("{\n\t.reg .pred p;\n\t"
      "setp.le.s32 p, %3, %4;\n\t"
      "setp.eq.and.f32 p, %1, %2, p;\n\t"
      "setp.lt.or.f32 p, %1, %2, p;\n\t"
      "@p or.b32 %0, %0, %5;\n\t}"
      : "+r"(mask) : "f"(d), "f"(am), "n"(S), "r"(j), "n"(1u << (5 + S)));
}
template <int S>
__device__ __forceinline__ void stay_bit_key(float d, float keyj, unsigned& mask) {
  asm("{\n\t.reg .pred p;\n\t.reg .f32 k;\n\t"
      "add.f32 k, %1, %2;\n\t"
      "setp.le.f32 p, k, %3;\n\t"
      "@p or.b32 %0, %0, %4;\n\t}"
      : "+r"(mask) : "f"(d), "f"((float)S * 0.015625f), "f"(keyj), "n"(1u << (5 + S)));
}

// A rows of a dp node, state-major
template <int AP, int V, bool FULL>
__device__ __forceinline__ void load_rows(const float* __restrict__ src, int A, size_t stride, float (&d)[AP][V]) {
#pragma unroll
  for (int t = 0; t < AP; ++t) {
    if (FULL || t < A) ld_f<V>(src + (size_t)t * stride, d[t]);
    else {
#pragma unroll
      for (int v = 0; v < V; ++v) d[t][v] = INFINITY;
    }
  }
}

// leaf / missing child as the recurrence sees it: 1e5 with a 0 at the observed state (sankoff.py:49-52,152)
template <int AP, int V, bool FULL>
__device__ __forceinline__ void leaf_rows(const int (&o)[V], int A, float (&d)[AP][V]) {
#pragma unroll
  for (int t = 0; t < AP; ++t)
#pragma unroll
    for (int v = 0; v < V; ++v) d[t][v] = (FULL || t < A) ? (o[v] == t ? 0.f : kSentinel) : INFINITY;
}

// a + (minimum of the row) and the index of its FIRST minimum (jnp.argmin's rule)
template <int AP, int V, bool WANT_IDX>
__device__ __forceinline__ void row_first_min(const float (&d)[AP][V], float a, float (&am)[V], int (&j)[V]) {
#pragma unroll
  for (int v = 0; v < V; ++v) {
    float m = d[0][v];
#pragma unroll
    for (int t = 1; t + 1 < AP; t += 2) m = fminf(m, fminf(d[t][v], d[t + 1][v]));  // FMNMX3
    if ((AP & 1) == 0) m = fminf(m, d[AP - 1][v]);
    int jj = 0;
    if (WANT_IDX) {
#pragma unroll
      for (int t = AP - 1; t >= 1; --t) jj = d[t][v] == m ? t : jj;
      jj = d[0][v] == m ? 0 : jj;
    }
    am[v] = a + m;  // C[s,t] + dp[child,t] for t != s at the minimum (sankoff.py:67)
    j[v] = jj;
  }
}

// The same for a child given by its observed state o, whose row is 0 at o and 1e5 elsewhere: minimum 0 at o
// (an all-1e5 row when o >= A: minimum 1e5, first at 0) — no 20-element min / first-index chain.  Half of the
// children of a binary tree are leaves, and the kernel is bound by the ALU pipe.
template <int V>
__device__ __forceinline__ void leaf_first_min(const int (&o)[V], int A, float a, float (&am)[V], int (&j)[V]) {
#pragma unroll
  for (int v = 0; v < V; ++v) {
    const bool none = o[v] >= A;
    am[v] = a + (none ? kSentinel : 0.f);
    j[v] = none ? 0 : o[v];
  }
}

}  // namespace

// FULL: n_states == AP (no padded state rows); WANT_ARG: the traceback tables are written; COMPACT: as pointer words.
template <int AP, int V, int THREADS, bool FULL, bool WANT_ARG, bool COMPACT = false>
__global__ void __launch_bounds__(THREADS, 4) hard_fwd_uniform_kernel(const HardFwdParams p) {
  static_assert(!COMPACT || (WANT_ARG && AP <= 27), "pointer words: 5 + n_states bits");
  float a;
  if (p.phase == 1 && (int)blockIdx.z >= (p.plan + (size_t)blockIdx.y * p.plan_words)[2]) return;  // empty segment slice (CTA-uniform): before any other work
  if (!hard_uniform_applies(p, a)) return;
  const int A = FULL ? AP : p.n_states;
  const int tree = blockIdx.y;
  const int n_int = p.n_all - p.n_leaves;
  PlanView plan(p.plan + (size_t)tree * p.plan_words, n_int);
  const Walk wk = make_walk(plan, p.phase, blockIdx.z);
  const int n_walk = wk.n_fwd;
  if (n_walk == 0) return;
  // the root is the last reachable node of the walk that contains it
  const bool has_root = wk.n_rev > 0 && __ldg(wk.ord + wk.n_rev - 1) == p.n_all - 1;
  const size_t stride = (size_t)p.stride;
  const long long site0 = ((long long)blockIdx.x * THREADS + threadIdx.x) * V;
  const bool active = site0 < p.n_sites;
  float rmin[V];
#pragma unroll
  for (int v = 0; v < V; ++v) rmin[v] = INFINITY;
  const unsigned walkers = __ballot_sync(0xffffffffu, active);  // the lanes of this warp that walk the tree

  if (active) {
    float* dp = p.dp + (size_t)tree * n_int * A * stride + site0;
    uint8_t* arg = WANT_ARG ? p.arg + ((size_t)tree * n_int * A * stride + site0) * 2 : nullptr;
    const uint8_t* leaf = p.leaf + site0;
    const int32_t* order = wk.ord;
    const int32_t* ch0 = plan.child(0);
    const int32_t* ch1 = plan.child(1);
    // In DFS post-order the node computed last is a child of the next node: its row stays in
    // registers (dx, updated in place), so only the other child's row (dy) is read back, and no
    // load ever waits on this thread's own preceding store.  dy is software-pipelined in place:
    // as soon as state s of the current node is done, dy[s] is reloaded with row s of the NEXT
    // node's other child, so the L2/HBM latency hides behind the remaining states.
    float dx[AP][V], dy[AP][V];
    int ox[V], oy[V];  // prefetched leaf states of a node's leaf children (A = "no state")
    int prev = -1;
    int node, k, cx, cy;
    bool x_is_0, fwd;
    auto decode = [&](int i) {
      node = __ldg(order + i);
      k = node - p.n_leaves;
      const int c0 = __ldg(ch0 + k), c1 = __ldg(ch1 + k);
      x_is_0 = prev >= p.n_leaves && c0 == prev;            // dx already holds child 0
      fwd = x_is_0 || (prev >= p.n_leaves && c1 == prev);
      cx = x_is_0 ? c0 : c1;                                // dx <-> cx, dy <-> cy
      cy = x_is_0 ? c1 : c0;
#pragma unroll
      for (int v = 0; v < V; ++v) ox[v] = oy[v] = A;        // missing child (-1 fill): all-1e5 row
      if (!fwd && cx >= 0 && cx < p.n_leaves) ld_u8<V>(leaf + (size_t)cx * stride, ox);
      if (cy >= 0 && cy < p.n_leaves) ld_u8<V>(leaf + (size_t)cy * stride, oy);
    };
    {
      decode(0);
      if (cy >= p.n_leaves) load_rows<AP, V, FULL>(dp + (size_t)(cy - p.n_leaves) * A * stride, A, stride, dy);
    }
    for (int i = 0; i < n_walk; ++i) {
      const bool x_leaf = !fwd && cx < p.n_leaves, y_leaf = cy < p.n_leaves;  // warp-uniform (plan data)
      if (!fwd) {
        if (cx >= p.n_leaves) load_rows<AP, V, FULL>(dp + (size_t)(cx - p.n_leaves) * A * stride, A, stride, dx);
        else leaf_rows<AP, V, FULL>(ox, A, dx);
      }
      if (y_leaf) leaf_rows<AP, V, FULL>(oy, A, dy);
      float amx[V], amy[V];
      int jx[V], jy[V];
      if (x_leaf) leaf_first_min<V>(ox, A, a, amx, jx);
      else row_first_min<AP, V, WANT_ARG>(dx, a, amx, jx);
      if (y_leaf) leaf_first_min<V>(oy, A, a, amy, jy);
      else row_first_min<AP, V, WANT_ARG>(dy, a, amy, jy);
      float* out = dp + (size_t)k * A * stride;
      // back-pointers of child x / child y, interleaved per site (slot 0 = first child, sankoff.py:80-83);
      // x_first is taken NOW: decode(i + 1) below overwrites x_is_0 with the next node's
      uint8_t* aout = WANT_ARG ? arg + (size_t)k * 2 * A * stride : nullptr;
      const bool x_first = x_is_0;
      // the next node (its leaf states are fetched now, its dy rows inside the state loop)
      prev = node;
      const bool more = i + 1 < n_walk;
      if (more) decode(i + 1);
      const bool reload = more && cy >= p.n_leaves;
      const float* nsrc = dp + (size_t)(reload ? cy - p.n_leaves : 0) * A * stride;
      // one-compare back-pointers when every lane's a + min is a small integer (warp-uniform choice)
      bool small = WANT_ARG;
      float kjx[V], kjy[V];
#pragma unroll
      for (int v = 0; v < V; ++v) {
        small = small && amx[v] < 131072.f && amy[v] < 131072.f;
        kjx[v] = fmaf((float)jx[v], 0.015625f, amx[v]);
        kjy[v] = fmaf((float)jy[v], 0.015625f, amy[v]);
      }
      small = WANT_ARG && __all_sync(walkers, small);
      unsigned mx[V], my[V];  // COMPACT: the "stays" bits of the two children
#pragma unroll
      for (int v = 0; v < V; ++v) mx[v] = my[v] = 0u;
      if constexpr (COMPACT) {
        // Before the state loop (which overwrites dx and reloads dy).  A leaf child needs no compares: its row is 0 at
        // the observed state o and 1e5 elsewhere, so for a > 0 only state o stays (0 < a; 1e5 > a + 0), and a
        // stateless leaf (all 1e5 < a + 1e5) makes every state stay.  Half of a binary tree's children are leaves.
        auto masks = [&](const float (&d)[AP][V], const float (&am)[V], const float (&kj)[V], const int (&j)[V],
                         bool leaf, unsigned (&m)[V]) {
          if (leaf && a > 0.f) {
#pragma unroll
            for (int v = 0; v < V; ++v) m[v] = am[v] >= kSentinel ? (((1u << A) - 1u) << 5) : (1u << (5 + j[v]));
          } else if (small) {
            static_for<AP>([&](auto sc) {
              constexpr int s = decltype(sc)::value;
              if (FULL || s < A) {
#pragma unroll
                for (int v = 0; v < V; ++v) stay_bit_key<s>(d[s][v], kj[v], m[v]);
              }
            });
          } else {
            static_for<AP>([&](auto sc) {
              constexpr int s = decltype(sc)::value;
              if (FULL || s < A) {
#pragma unroll
                for (int v = 0; v < V; ++v) stay_bit<s>(d[s][v], am[v], j[v], m[v]);
              }
            });
          }
        };
        masks(dx, amx, kjx, jx, x_leaf, mx);
        masks(dy, amy, kjy, jy, y_leaf, my);
      }
      auto state_rows = [&](auto keyed) {
      static_for<AP>([&](auto sc) {
        constexpr int s = decltype(sc)::value;
        if (FULL || s < A) {
          int ix[V], iy[V];
#pragma unroll
          for (int v = 0; v < V; ++v) {
            if constexpr (COMPACT) {
              // the words were taken before this loop
            } else if (WANT_ARG) {
              if constexpr (decltype(keyed)::value) {
                ix[v] = pick_index_key<s>(dx[s][v], kjx[v], jx[v]);
                iy[v] = pick_index_key<s>(dy[s][v], kjy[v], jy[v]);
              } else {
                ix[v] = pick_index<s>(dx[s][v], amx[v], jx[v]);
                iy[v] = pick_index<s>(dy[s][v], amy[v], jy[v]);
              }
            }
            // (0 + cost0) + cost1, sankoff.py:70 (f32 addition commutes)
            dx[s][v] = fminf(dx[s][v], amx[v]) + fminf(dy[s][v], amy[v]);
          }
          st_f<V>(out + (size_t)s * stride, dx[s]);
          if (WANT_ARG && !COMPACT) st_arg_pair<V>(aout + (size_t)s * 2 * stride, ix, iy, x_first);
          if (reload) ld_f<V>(nsrc + (size_t)s * stride, dy[s]);
        }
      });
      };
      if (small && !COMPACT) state_rows(std::true_type{});
      else state_rows(std::false_type{});
      if constexpr (COMPACT) {
        // rows 0-3 of this node's byte table take the four 16-bit quarters of (word of slot 0, word of slot 1): any
        // pointer arithmetic that is right for the byte table (tree batches, site blocks) is right for the words.
        // slot 0 = first child (sankoff.py:80-83); jx / jy are this node's (decode(i + 1) does not touch them)
        unsigned w0[V], w1[V];
#pragma unroll
        for (int v = 0; v < V; ++v) {
          const unsigned wx = mx[v] | (unsigned)jx[v], wy = my[v] | (unsigned)jy[v];
          w0[v] = x_first ? wx : wy;
          w1[v] = x_first ? wy : wx;
        }
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          uint8_t* dst = aout + (size_t)r * 2 * stride;
          const unsigned sel = (r & 1) ? 0x7632u : 0x5410u;  // high / low halves of two words
          if constexpr (V == 1) {
            __stcs(reinterpret_cast<unsigned short*>(dst), (unsigned short)(((r < 2 ? w0[0] : w1[0]) >> (16 * (r & 1))) & 0xffffu));
          } else if constexpr (V == 2) {
            __stcs(reinterpret_cast<unsigned*>(dst), r < 2 ? __byte_perm(w0[0], w0[1], sel) : __byte_perm(w1[0], w1[1], sel));
          } else {
            __stcs(reinterpret_cast<uint2*>(dst),
                   r < 2 ? make_uint2(__byte_perm(w0[0], w0[1], sel), __byte_perm(w0[2], w0[3], sel))
                         : make_uint2(__byte_perm(w1[0], w1[1], sel), __byte_perm(w1[2], w1[3], sel)));
          }
        }
      }
    }
    // per-site root minimum (sankoff.py:187): the root's row is still in dx unless nodes the root does
    // not reach followed it in the walk
    if (has_root) {
      if (prev != p.n_all - 1) load_rows<AP, V, FULL>(dp + (size_t)(n_int - 1) * A * stride, A, stride, dx);
#pragma unroll
      for (int s = 0; s < AP; ++s)
        if (FULL || s < A) {
#pragma unroll
          for (int v = 0; v < V; ++v) rmin[v] = fminf(rmin[v], dx[s][v]);
        }
      if (p.root_min) st_f<V>(p.root_min + (size_t)tree * stride + site0, rmin);
    }
  }
  if (p.total && has_root) {
    double acc = 0.0;
    if (active) {
#pragma unroll
      for (int v = 0; v < V; ++v)
        if (site0 + v < p.n_sites) acc += (double)rmin[v];
    }
    block_atomic_add(acc, p.total + tree);
  }
}

// ---- relaxed (softmin) forward, uniform cost — SURVEY §8a row R1 -------------------------------
// softmin_t(C[s,t] + d[t]) = m - T log( e[s] + k (E - e[s]) ),  e[t] = exp(-(d[t]-m)/T), E = sum_t e[t]
// Same walk as the hard kernel: the row computed last stays in registers (dx), the other child's
// row (dy) is reloaded in place, state by state, for the next node.  Per node and site this is
// 2A ex2 + A lg2 on the SFU (the two children's logs are merged: log z0 + log z1 = log z0 z1).
namespace {

// d -> e = exp(-(d - m)/T) in place; returns m and k * sum(e)
template <int AP, int V>
__device__ __forceinline__ void to_exp_domain(float (&d)[AP][V], float c1, float kk, float (&m)[V], float (&kE)[V]) {
#pragma unroll
  for (int v = 0; v < V; ++v) {
    float mm = d[0][v];
#pragma unroll
    for (int t = 1; t + 1 < AP; t += 2) mm = fminf(mm, fminf(d[t][v], d[t + 1][v]));
    if ((AP & 1) == 0) mm = fminf(mm, d[AP - 1][v]);
    const float mc = mm * c1;
    float E0 = 0.f, E1 = 0.f;
#pragma unroll
    for (int t = 0; t < AP; ++t) {
      d[t][v] = ex2_approx(fmaf(d[t][v], -c1, mc));  // padded rows: ex2(-inf) = 0
      if (t & 1) E1 += d[t][v]; else E0 += d[t][v];
    }
    m[v] = mm;
    kE[v] = kk * (E0 + E1);
  }
}

// The same for a child given by its observed state o (one-hot 1e5 row) without touching the SFU: the row is
// 0 at o and 1e5 elsewhere, so m = 0, e[o] = 1 and e[t != o] = e_off = exp(-1e5/T) (0 unless T is huge) —
// exactly what to_exp_domain computes from the materialised row.  o >= A (no state, an all-1e5 row): m = 1e5,
// e = 1 everywhere.  Half of the children of a binary tree are leaves: half of the kernel's ex2.
template <int AP, int V, bool FULL>
__device__ __forceinline__ void leaf_exp_domain(const int (&o)[V], int A, float e_off, float kk, float (&d)[AP][V],
                                                float (&m)[V], float (&kE)[V]) {
#pragma unroll
  for (int v = 0; v < V; ++v) {
    const bool none = o[v] >= A;
#pragma unroll
    for (int t = 0; t < AP; ++t) d[t][v] = (FULL || t < A) ? ((none || o[v] == t) ? 1.f : e_off) : 0.f;
    m[v] = none ? kSentinel : 0.f;
    kE[v] = kk * (none ? (float)A : fmaf((float)(A - 1), e_off, 1.f));
  }
}

}  // namespace

template <int AP, int V, int THREADS, bool FULL>
__global__ void __launch_bounds__(THREADS, 4) soft_fwd_uniform_kernel(const SoftFwdParams p) {
  float a;
  if (p.phase == 1 && (int)blockIdx.z >= (p.plan + (size_t)blockIdx.y * p.plan_words)[2]) return;  // empty segment slice (CTA-uniform): before any other work
  if (!soft_uniform_applies(p, a)) return;
  const int A = FULL ? AP : p.n_states;
  const int tree = blockIdx.y;
  const int n_int = p.n_all - p.n_leaves;
  PlanView plan(p.plan + (size_t)tree * p.plan_words, n_int);
  const Walk wk = make_walk(plan, p.phase, blockIdx.z);
  const int n_walk = wk.n_fwd;
  if (n_walk == 0) return;
  const bool has_root = wk.n_rev > 0 && __ldg(wk.ord + wk.n_rev - 1) == p.n_all - 1;
  const float c1 = 1.4426950408889634f / p.T;    // log2(e)/T
  const float tln2 = p.T * 0.6931471805599453f;  // T*ln(2)
  const float kk = exp2f(-a * c1);               // exp(-a/T)
  const float km1 = 1.f - kk;
  const float e_off = ex2_approx(-kSentinel * c1);  // exp(-1e5/T): every unobserved state of a leaf row
  const size_t stride = (size_t)p.stride;
  const long long site0 = ((long long)blockIdx.x * THREADS + threadIdx.x) * V;
  const bool active = site0 < p.n_sites;
  double acc = 0.0;
  if (active) {
    float* dp = p.dp + (size_t)tree * n_int * A * stride + site0;
    const uint8_t* leaf = p.leaf ? p.leaf + site0 : nullptr;
    const float* leaf_cost = p.leaf_cost ? p.leaf_cost + site0 : nullptr;
    const int32_t* order = wk.ord;
    const int32_t* ch0 = plan.child(0);
    const int32_t* ch1 = plan.child(1);
    // rows of a child that lives in memory (dp node, or a leaf of a float leaf-cost table); null for
    // a leaf given by its state (one-hot 1e5 row) or a missing child (all-1e5 row), sankoff.py:49-60
    auto row_src = [&](int c) -> const float* {
      if (c >= p.n_leaves) return dp + (size_t)(c - p.n_leaves) * A * stride;
      if (c >= 0 && leaf_cost) return leaf_cost + (size_t)c * A * stride;
      return nullptr;
    };
    float dx[AP][V], dy[AP][V];
    int ox[V], oy[V];
    int prev = -1;
    int node, k;
    const float *sx, *sy;
    bool fwd;
    auto decode = [&](int i) {
      node = __ldg(order + i);
      k = node - p.n_leaves;
      const int c0 = __ldg(ch0 + k), c1_ = __ldg(ch1 + k);
      const bool x_is_0 = prev >= p.n_leaves && c0 == prev;
      fwd = x_is_0 || (prev >= p.n_leaves && c1_ == prev);
      const int cx = x_is_0 ? c0 : c1_, cy = x_is_0 ? c1_ : c0;   // the sum over children commutes
      sx = row_src(cx);
      sy = row_src(cy);
#pragma unroll
      for (int v = 0; v < V; ++v) ox[v] = oy[v] = A;
      if (!fwd && !sx && cx >= 0 && leaf) ld_u8<V>(leaf + (size_t)cx * stride, ox);
      if (!sy && cy >= 0 && leaf) ld_u8<V>(leaf + (size_t)cy * stride, oy);
    };
    {
      decode(0);
      if (sy) load_rows<AP, V, FULL>(sy, A, stride, dy);
    }
    for (int i = 0; i < n_walk; ++i) {
      if (!fwd && !sx && !sy) {
        // Both children given by their states (half of the nodes of a binary tree): e is 1 at the observed state
        // and e_off elsewhere, so z0 z1 takes four values per site — four lg2 instead of A.  A stateless leaf
        // (o >= A: e = 1 everywhere, m = 1e5) is the same formula with e_off -> 1.  Bit-identical to the general
        // path: the same fmaf / product / lg2 on the same operands.
        float l11[V], l10[V], l01[V], l00[V];
        int lx[V], ly[V];
#pragma unroll
        for (int v = 0; v < V; ++v) {
          lx[v] = ox[v];
          ly[v] = oy[v];
          const bool nx = ox[v] >= A, ny = oy[v] >= A;
          const float kEx = kk * (nx ? (float)A : fmaf((float)(A - 1), e_off, 1.f));
          const float kEy = kk * (ny ? (float)A : fmaf((float)(A - 1), e_off, 1.f));
          const float zx1 = fmaf(km1, 1.f, kEx), zx0 = fmaf(km1, nx ? 1.f : e_off, kEx);
          const float zy1 = fmaf(km1, 1.f, kEy), zy0 = fmaf(km1, ny ? 1.f : e_off, kEy);
          const float mm = (nx ? kSentinel : 0.f) + (ny ? kSentinel : 0.f);
          l11[v] = fmaf(-tln2, lg2_approx(zx1 * zy1), mm);
          l10[v] = fmaf(-tln2, lg2_approx(zx1 * zy0), mm);
          l01[v] = fmaf(-tln2, lg2_approx(zx0 * zy1), mm);
          l00[v] = fmaf(-tln2, lg2_approx(zx0 * zy0), mm);
        }
        float* out = dp + (size_t)k * A * stride;
        prev = node;
        const bool more = i + 1 < n_walk;
        if (more) decode(i + 1);
        const bool reload = more && sy != nullptr;
        const float* nsrc = reload ? sy : dp;
#pragma unroll
        for (int s = 0; s < AP; ++s) {
          if (FULL || s < A) {
#pragma unroll
            for (int v = 0; v < V; ++v)
              dx[s][v] = s == lx[v] ? (s == ly[v] ? l11[v] : l10[v]) : (s == ly[v] ? l01[v] : l00[v]);
            st_f<V>(out + (size_t)s * stride, dx[s]);
            if (reload) ld_f<V>(nsrc + (size_t)s * stride, dy[s]);
          } else {
#pragma unroll
            for (int v = 0; v < V; ++v) dx[s][v] = dy[s][v] = INFINITY;
          }
        }
        continue;
      }
      float mx[V], my[V], kEx[V], kEy[V];
      if (!fwd && sx) load_rows<AP, V, FULL>(sx, A, stride, dx);
      if (!fwd && !sx) leaf_exp_domain<AP, V, FULL>(ox, A, e_off, kk, dx, mx, kEx);
      else to_exp_domain<AP, V>(dx, c1, kk, mx, kEx);
      if (!sy) leaf_exp_domain<AP, V, FULL>(oy, A, e_off, kk, dy, my, kEy);
      else to_exp_domain<AP, V>(dy, c1, kk, my, kEy);
      float mxy[V];
#pragma unroll
      for (int v = 0; v < V; ++v) mxy[v] = mx[v] + my[v];
      float* out = dp + (size_t)k * A * stride;
      prev = node;
      const bool more = i + 1 < n_walk;
      if (more) decode(i + 1);
      const bool reload = more && sy != nullptr;
      const float* nsrc = reload ? sy : dp;
#pragma unroll
      for (int s = 0; s < AP; ++s) {
        if (FULL || s < A) {
#pragma unroll
          for (int v = 0; v < V; ++v) {
            const float z0 = fmaf(km1, dx[s][v], kEx[v]);
            const float z1 = fmaf(km1, dy[s][v], kEy[v]);
            dx[s][v] = fmaf(-tln2, lg2_approx(z0 * z1), mxy[v]);
          }
          st_f<V>(out + (size_t)s * stride, dx[s]);
          if (reload) ld_f<V>(nsrc + (size_t)s * stride, dy[s]);
        } else {
#pragma unroll
          for (int v = 0; v < V; ++v) dx[s][v] = dy[s][v] = INFINITY;  // padded rows: back from e = 0 to d = inf
        }
      }
    }
    // total += softmin_s dp[root][s]   (sankoff.py:187 relaxed); the root row is in dx when it was
    // the last node of the plan
    if (has_root) {
      if (prev != p.n_all - 1) load_rows<AP, V, FULL>(dp + (size_t)(n_int - 1) * A * stride, A, stride, dx);
#pragma unroll
      for (int v = 0; v < V; ++v) {
        float mm = INFINITY;
#pragma unroll
        for (int s = 0; s < AP; ++s)
          if (FULL || s < A) mm = fminf(mm, dx[s][v]);
        float z = 0.f;
#pragma unroll
        for (int s = 0; s < AP; ++s)
          if (FULL || s < A) z += exp2f((mm - dx[s][v]) * c1);
        if (site0 + v < p.n_sites) acc += (double)fmaf(-tln2, __log2f(z), mm);
      }
    }
  }
  if (p.total && has_root) block_atomic_add(acc, p.total + tree);
}

namespace {
template <int AP, int V>
void launch_hard_uniform(const HardFwdParams& p0, int batch, cudaStream_t s) {
  constexpr int T = 128;
  const unsigned gx = (unsigned)((p0.n_sites + (long long)T * V - 1) / ((long long)T * V));
  const bool full = p0.n_states == AP;
  auto run = [&](const HardFwdParams& p, dim3 grid) {
    if (p.arg && p.compact) {
      if constexpr (AP <= 27) {
        if (full) { hard_fwd_uniform_kernel<AP, V, T, true, true, true><<<grid, T, 0, s>>>(p); note_launches(1); }
        else { hard_fwd_uniform_kernel<AP, V, T, false, true, true><<<grid, T, 0, s>>>(p); note_launches(1); }
      }
    } else if (p.arg) {
      if (full) { hard_fwd_uniform_kernel<AP, V, T, true, true><<<grid, T, 0, s>>>(p); note_launches(1); }
      else { hard_fwd_uniform_kernel<AP, V, T, false, true><<<grid, T, 0, s>>>(p); note_launches(1); }
    } else {
      if (full) { hard_fwd_uniform_kernel<AP, V, T, true, false><<<grid, T, 0, s>>>(p); note_launches(1); }
      else { hard_fwd_uniform_kernel<AP, V, T, false, false><<<grid, T, 0, s>>>(p); note_launches(1); }
    }
  };
  if (use_segments(p0.n_leaves)) {
    HardFwdParams p = p0;
    p.phase = 1;  // independent subtrees side by side ...
    run(p, dim3(gx, batch, kMaxSeg));
    p.phase = 2;  // ... then the nodes above them
    run(p, dim3(gx, batch));
  } else {
    run(p0, dim3(gx, batch));
  }
}
}  // namespace

int launch_hard_fwd_uniform(const HardFwdParams& p, int batch, cudaStream_t s) {
  const int A = p.n_states;
  if (A <= 4) launch_hard_uniform<4, 4>(p, batch, s);
  else if (A <= 8) launch_hard_uniform<8, 2>(p, batch, s);
  else if (A <= 20) launch_hard_uniform<20, 2>(p, batch, s);
  else launch_hard_uniform<32, 1>(p, batch, s);
  return check_cuda(cudaPeekAtLastError(), "hard_fwd_uniform_kernel");
}

namespace {
template <int AP, int V>
void launch_soft_uniform(const SoftFwdParams& p0, int batch, cudaStream_t s) {
  constexpr int T = 128;
  const unsigned gx = (unsigned)((p0.n_sites + (long long)T * V - 1) / ((long long)T * V));
  auto run = [&](const SoftFwdParams& p, dim3 grid) {
    if (p.n_states == AP) { soft_fwd_uniform_kernel<AP, V, T, true><<<grid, T, 0, s>>>(p); note_launches(1); }
    else { soft_fwd_uniform_kernel<AP, V, T, false><<<grid, T, 0, s>>>(p); note_launches(1); }
  };
  if (use_segments(p0.n_leaves)) {
    SoftFwdParams p = p0;
    p.phase = 1;
    run(p, dim3(gx, batch, kMaxSeg));
    p.phase = 2;
    run(p, dim3(gx, batch));
  } else {
    run(p0, dim3(gx, batch));
  }
}
}  // namespace

int launch_soft_fwd_uniform(const SoftFwdParams& p, int batch, cudaStream_t s) {
  const int A = p.n_states;
  if (A <= 4) launch_soft_uniform<4, 4>(p, batch, s);
  else if (A <= 8) launch_soft_uniform<8, 2>(p, batch, s);
  else if (A <= 20) launch_soft_uniform<20, 2>(p, batch, s);
  else launch_soft_uniform<32, 1>(p, batch, s);
  return check_cuda(cudaPeekAtLastError(), "soft_fwd_uniform_kernel");
}

}  // namespace trex
