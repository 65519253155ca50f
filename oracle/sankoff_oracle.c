/*
 * TEST INFRASTRUCTURE ONLY — plain-C restatement of /root/reference/src/trex/sankoff.py, used
 * (a) as a second checker beside the NumPy restatement and (b) as the timed CPU baseline of
 * bench.py (`cpu_baseline`, `--impl reference`): the reference itself is JAX and JAX is not
 * installable here, so this port stands in for it (kind = "port").  Never linked into, called
 * from or shipped with trex_b200.
 *
 * It keeps the reference's data structures and control flow: per site a (n_all, A) f32 DP table
 * filled with 1e5 (sankoff.py:152), a (n_all, A, 4) f32 back-pointer table (:151), index-order
 * loop over internal nodes (:87-92), children = first two rows with adj[:, node] == 1 (:60,
 * hoisted out of the site loop exactly as vmap hoists it), first-min argmin (:69), and the
 * explicit-stack traceback (:212-265).  Sites are the vmap axis (:97) -> the OpenMP loop.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define SENTINEL 1e5f

int oracle_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* jnp.where(adj[:, node] == 1, size=2, fill_value=-1)[0] for every node (sankoff.py:60) */
static void find_children(const float* adj, int n_all, int zero_root_selfloop, int* kids) {
  for (int node = 0; node < n_all; ++node) {
    int c = 0;
    kids[2 * node] = kids[2 * node + 1] = -1;
    for (int i = 0; i < n_all && c < 2; ++i) {
      float v = adj[(size_t)i * n_all + node];
      if (zero_root_selfloop && i == n_all - 1 && node == n_all - 1) v = 0.f; /* :141 */
      if (v == 1.0f) kids[2 * node + c++] = i;
    }
  }
}

static inline int wrap(int idx, int n) { return idx < 0 ? idx + n : idx; }

/* run_sankoff (sankoff.py:114-188).  seqs [n_rows][L] f32 (row pitch L); outputs in the
 * reference's layouts: recon [n_all][L] f32, dp [L][n_all][A] f32 (may be NULL), root_min [L]
 * (may be NULL), total f32.  Returns 0. */
int oracle_run_sankoff(const float* adj, const float* cost, const float* seqs, int n_all, int A,
                       int n_leaves, long L, int return_path, float* recon, float* dp_out,
                       float* root_min, float* total) {
  int* kids = (int*)malloc(sizeof(int) * 2 * (size_t)n_all);
  find_children(adj, n_all, 1, kids);
  const int dp_leaves = (n_all + 1) / 2; /* run_dp derives it from the shape, :45-46 */
  double tot = 0.0;
#pragma omp parallel reduction(+ : tot)
  {
    float* dp = (float*)malloc(sizeof(float) * (size_t)n_all * A);
    float* bt = (float*)malloc(sizeof(float) * (size_t)n_all * A * 4);
    int* stack = (int*)malloc(sizeof(int) * 2 * ((size_t)n_all + 2));
    int* rec = (int*)malloc(sizeof(int) * (size_t)n_all);
#pragma omp for schedule(static)
    for (long l = 0; l < L; ++l) {
      for (int i = 0; i < n_all * A; ++i) dp[i] = SENTINEL;
      memset(bt, 0, sizeof(float) * (size_t)n_all * A * 4);
      for (int i = 0; i < dp_leaves; ++i) { /* :49-52 */
        int s = (int)seqs[(size_t)i * L + l];
        if (s < 0) s += A;
        if (s >= 0 && s < A) dp[i * A + s] = 0.f;
      }
      for (int node = dp_leaves; node < n_all; ++node) { /* :87-92 */
        float* out = dp + (size_t)node * A;
        float acc[64];
        for (int s = 0; s < A; ++s) acc[s] = 0.f;
        for (int k = 0; k < 2; ++k) { /* scan over the two children, :73-77 */
          const int child = wrap(kids[2 * node + k], n_all);
          const float* d = dp + (size_t)child * A;
          for (int s = 0; s < A; ++s) {
            const float* c = cost + (size_t)s * A;
            float best = c[0] + d[0];
            int bi = 0;
            for (int t = 1; t < A; ++t) {
              const float x = c[t] + d[t];
              if (x < best) { best = x; bi = t; }
            }
            acc[s] += best;
            bt[((size_t)node * A + s) * 4 + 2 * k] = (float)kids[2 * node + k];
            bt[((size_t)node * A + s) * 4 + 2 * k + 1] = (float)bi;
          }
        }
        for (int s = 0; s < A; ++s) out[s] = acc[s]; /* :79 */
      }
      /* score: dp[:, -1].min(axis=1) summed over sites, :187 */
      const float* root = dp + (size_t)(n_all - 1) * A;
      float rm = root[0];
      int rs = 0;
      for (int s = 1; s < A; ++s)
        if (root[s] < rm) { rm = root[s]; rs = s; }
      tot += (double)rm;
      if (root_min) root_min[l] = rm;
      if (dp_out) memcpy(dp_out + (size_t)l * n_all * A, dp, sizeof(float) * (size_t)n_all * A);
      for (int i = 0; i < n_all; ++i)
        recon[(size_t)i * L + l] = i < n_leaves ? seqs[(size_t)i * L + l] : 0.f; /* :161-162 */
      if (return_path) { /* :164-185, 212-265 */
        for (int i = 0; i < n_all; ++i) rec[i] = 0;
        int ptr = 1;
        stack[0] = n_all - 1;
        stack[1] = rs;
        while (ptr > 0) {
          --ptr;
          const int node = stack[2 * ptr], state = stack[2 * ptr + 1];
          if (node >= n_leaves) {
            rec[node] = state;
            const float* info = bt + ((size_t)node * A + state) * 4;
            stack[2 * ptr] = (int)info[0];
            stack[2 * ptr + 1] = (int)info[1];
            stack[2 * ptr + 2] = (int)info[2];
            stack[2 * ptr + 3] = (int)info[3];
            ptr += 2;
          }
        }
        for (int i = n_leaves; i < n_all; ++i) recon[(size_t)i * L + l] = (float)rec[i];
      }
    }
    free(dp); free(bt); free(stack); free(rec);
  }
  *total = (float)tot;
  free(kids);
  return 0;
}

/* Relaxed forward + reverse (not in the reference — restatement of :55-92 with
 * min -> -T*logsumexp(-x/T)).  Hard leaves (seqs) or, when leaf_cost != NULL, a float leaf table
 * leaf_cost [n_leaves][L][A] (REAL).  Outputs: the total, dC [A][A] f64 and (optionally) the dp table
 * dp_out [L][n_all][A] and the cotangents / marginals gbar [L][n_all][A] (rows < n_leaves = d total /
 * d leaf_cost).  Two instances of the same body: float32 like the reference's default dtype (the timed
 * CPU baseline) and float64 (the checker of the 1e-5 parity tests at the benchmark shapes). */
#define DEFINE_SOFT_SANKOFF(NAME, REAL, EXPF, LOGF)                                                      \
  int NAME(const float* adj, const float* cost, const float* seqs, const REAL* leaf_cost, int n_all,    \
           int A, int n_leaves, long L, REAL T, double* total, double* dC, REAL* dp_out,                \
           REAL* gbar_out) {                                                                             \
    int* kids = (int*)malloc(sizeof(int) * 2 * (size_t)n_all);                                           \
    find_children(adj, n_all, 1, kids);                                                                  \
    double tot = 0.0;                                                                                    \
    for (int i = 0; i < A * A; ++i) dC[i] = 0.0;                                                         \
    _Pragma("omp parallel reduction(+ : tot)") {                                                         \
      REAL* dp = (REAL*)malloc(sizeof(REAL) * (size_t)n_all * A);                                        \
      REAL* g = (REAL*)malloc(sizeof(REAL) * (size_t)n_all * A);                                         \
      double* dcl = (double*)calloc((size_t)A * A, sizeof(double));                                      \
      REAL w[64];                                                                                        \
      _Pragma("omp for schedule(static)") for (long l = 0; l < L; ++l) {                                 \
        for (int i = 0; i < n_all * A; ++i) dp[i] = (REAL)SENTINEL;                                      \
        for (int i = 0; i < n_leaves; ++i) {                                                             \
          if (leaf_cost) {                                                                               \
            for (int s = 0; s < A; ++s) dp[i * A + s] = leaf_cost[((size_t)i * L + l) * A + s];          \
          } else {                                                                                       \
            int s = (int)seqs[(size_t)i * L + l];                                                        \
            if (s < 0) s += A;                                                                           \
            if (s >= 0 && s < A) dp[i * A + s] = (REAL)0;                                                \
          }                                                                                              \
        }                                                                                                \
        for (int node = n_leaves; node < n_all; ++node) {                                                \
          for (int s = 0; s < A; ++s) {                                                                  \
            REAL acc = (REAL)0;                                                                          \
            for (int k = 0; k < 2; ++k) {                                                                \
              const int child = wrap(kids[2 * node + k], n_all);                                         \
              const int valid = child < node;                                                            \
              const float* c = cost + (size_t)s * A;                                                     \
              REAL m = (REAL)INFINITY;                                                                   \
              for (int t = 0; t < A; ++t) {                                                              \
                w[t] = (REAL)c[t] + (valid ? dp[(size_t)child * A + t] : (REAL)SENTINEL);                \
                if (w[t] < m) m = w[t];                                                                  \
              }                                                                                          \
              REAL z = (REAL)0;                                                                          \
              for (int t = 0; t < A; ++t) z += EXPF(-(w[t] - m) / T);                                    \
              acc += m - T * LOGF(z);                                                                    \
            }                                                                                            \
            dp[(size_t)node * A + s] = acc;                                                              \
          }                                                                                              \
        }                                                                                                \
        const REAL* root = dp + (size_t)(n_all - 1) * A;                                                 \
        REAL m = root[0];                                                                                \
        for (int s = 1; s < A; ++s)                                                                      \
          if (root[s] < m) m = root[s];                                                                  \
        REAL z = (REAL)0;                                                                                \
        for (int s = 0; s < A; ++s) z += EXPF(-(root[s] - m) / T);                                       \
        tot += (double)(m - T * LOGF(z));                                                                \
        memset(g, 0, sizeof(REAL) * (size_t)n_all * A);                                                  \
        for (int s = 0; s < A; ++s) g[(size_t)(n_all - 1) * A + s] = EXPF(-(root[s] - m) / T) / z;       \
        for (int node = n_all - 1; node >= n_leaves; --node) {                                           \
          for (int k = 0; k < 2; ++k) {                                                                  \
            const int child = wrap(kids[2 * node + k], n_all);                                           \
            const int valid = child < node;                                                              \
            for (int s = 0; s < A; ++s) {                                                                \
              const REAL gs = g[(size_t)node * A + s];                                                   \
              if (gs == (REAL)0) continue;                                                               \
              const float* c = cost + (size_t)s * A;                                                     \
              REAL mm = (REAL)INFINITY;                                                                  \
              for (int t = 0; t < A; ++t) {                                                              \
                w[t] = (REAL)c[t] + (valid ? dp[(size_t)child * A + t] : (REAL)SENTINEL);                \
                if (w[t] < mm) mm = w[t];                                                                \
              }                                                                                          \
              REAL zz = (REAL)0;                                                                         \
              for (int t = 0; t < A; ++t) {                                                              \
                w[t] = EXPF(-(w[t] - mm) / T);                                                           \
                zz += w[t];                                                                              \
              }                                                                                          \
              for (int t = 0; t < A; ++t) {                                                              \
                const REAL contrib = gs * w[t] / zz;                                                     \
                dcl[(size_t)s * A + t] += (double)contrib;                                               \
                if (valid) g[(size_t)child * A + t] += contrib;                                          \
              }                                                                                          \
            }                                                                                            \
          }                                                                                              \
        }                                                                                                \
        if (dp_out) memcpy(dp_out + (size_t)l * n_all * A, dp, sizeof(REAL) * (size_t)n_all * A);        \
        if (gbar_out) memcpy(gbar_out + (size_t)l * n_all * A, g, sizeof(REAL) * (size_t)n_all * A);     \
      }                                                                                                  \
      _Pragma("omp critical") for (int i = 0; i < A * A; ++i) dC[i] += dcl[i];                           \
      free(dp);                                                                                          \
      free(g);                                                                                           \
      free(dcl);                                                                                         \
    }                                                                                                    \
    *total = tot;                                                                                        \
    free(kids);                                                                                          \
    return 0;                                                                                            \
  }

DEFINE_SOFT_SANKOFF(oracle_soft_sankoff_f32, float, expf, logf)
DEFINE_SOFT_SANKOFF(oracle_soft_sankoff_f64, double, exp, log)

/* the timed CPU baseline: float32, hard leaves */
int oracle_soft_sankoff(const float* adj, const float* cost, const float* seqs, int n_all, int A,
                        int n_leaves, long L, float T, double* total, double* dC, float* gbar_out) {
  return oracle_soft_sankoff_f32(adj, cost, seqs, NULL, n_all, A, n_leaves, L, T, total, dC, NULL, gbar_out);
}

/* thread count of the OpenMP legs (torchrun exports OMP_NUM_THREADS=1: bench.py sets it back) */
void oracle_set_num_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}
