/*
 * trex_b200.h — C ABI of libtrex_b200.so: hand-written sm_100a CUDA kernels for the
 * differentiable tree-scoring hot path of maraxen/trex.
 *
 * The reference has no native/FFI interface (it is pure Python/JAX): the boundary
 * a maintainer binds is the Python signature of each jit-ed function.  Every entry point below
 * names the reference function (file:line under /root/reference) whose body it replaces; the
 * binding stubs (jax.ffi / ctypes) are shown in INTEGRATION.md.
 *
 * Conventions
 *  - Plain pointers and sizes only.  All data pointers are DEVICE pointers unless the name ends
 *    in `_host`.  Inputs are read-only; outputs and workspaces are caller-owned; the library
 *    never allocates, frees or retains device memory and never synchronises the device.
 *  - `stream` is a cudaStream_t passed as void*; all work is enqueued on it (graph-capturable).
 *  - Return value: 0 = ok, non-zero = error; `trex_last_error()` returns a thread-local message.
 *  - Native table layout is site-contiguous  [tree][node][state][site]  with row pitch
 *    `site_stride` (elements, multiple of 16, >= n_sites).  Pad lanes are written with
 *    unspecified values and are excluded from every reduction.  The kernels move up to four sites per
 *    lane with 8- and 16-byte accesses: every table pointer (dp, argmin, leaf/ancestral states, gradients)
 *    must be 16-byte aligned — true of cudaMalloc / framework allocations and of any offset by whole trees or by
 *    site blocks that start at a multiple of 64 sites.
 *  - Leaf states are uint8 codes: 0..A-1 observed state, >= A "no state" (the reference's
 *    dropped out-of-bounds scatter, sankoff.py:50: the leaf row stays all-1e5).
 *  - Internal node k (0-based) is node id n_leaves + k.  n_int = n_all - n_leaves.
 */
#ifndef TREX_B200_H_
#define TREX_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define TREX_API __attribute__((visibility("default")))
#else
#define TREX_API
#endif

#define TREX_OK 0
#define TREX_ERR_INVALID_ARGUMENT 1
#define TREX_ERR_CUDA 2
#define TREX_ERR_UNSUPPORTED 3

/* inner-loop selection of the hard Sankoff forward (testing knob; AUTO is the product path) */
#define TREX_MODE_AUTO 0    /* packed-key path when the cost matrix allows it, else exact-f32 */
#define TREX_MODE_GENERAL 1 /* always the exact-f32 compare/select path */
/* OR-ed into `mode` of trex_sankoff_hard_fwd: when the cost matrix is uniform (C[s][s] = 0, C[s][t] = a: the O(A)
 * kernels run) and 4 <= n_states <= 20, the back-pointers of a node are written as two POINTER WORDS per site —
 *   word of slot c (c = 0: first child): bits 0-4 = that child's first-minimum state j, bit 5+s = "parent state s
 *   keeps its own state", i.e. pointer(s) = bit ? s : j —
 * into state rows 0-3 of the node's [n_states][site][2] byte table (row r, site i: bits 16r..16r+15 of
 * word0 | word1 << 32, little-endian; rows 4.. are not written).  8 bytes per (node, site) instead of 2 * n_states,
 * coalesced on both sides; buffer size, alignment and every offset (tree batches, site blocks) stay those of the byte
 * table.  For any other cost matrix or state count the byte table is written as without the flag.  Only
 * trex_sankoff_traceback_cost — which is given the same cost matrix and mode and repeats the device-side decision —
 * can read the result; trex_sankoff_traceback and trex_sankoff_export_bt need a table written WITHOUT this flag.
 * No counterpart in the reference (its backtracking table is (n_all, A, 4) f32 per site, sankoff.py:80-83). */
#define TREX_MODE_COMPACT_ARG 0x100

/* plan status bits (word 3 of each tree's plan header) */
#define TREX_PLAN_HAS_SENTINEL_CHILD 1 /* some node has < 2 children, or a child id >= its own */
#define TREX_PLAN_NOT_A_TREE 2         /* a node is the child of two parents; index order used */
#define TREX_PLAN_SHARED_LEAF 4        /* a leaf is the child of two parents (or listed twice) */
/* The forward passes accept such adjacencies and match the reference on them.  The relaxed REVERSE pass is defined
 * for trees only (it stores each child's cotangent row once instead of accumulating): when bit 2 or 4 is set,
 * trex_sankoff_soft_bwd / _compact write NaN into every entry of `dcost` (when given), so the condition is visible
 * in the results without a host synchronisation; gbar / dleaf are then unspecified. */

TREX_API const char* trex_last_error(void);
TREX_API int trex_version(void);
/* Diagnostics: number of CUDA kernels this library has launched in this process (every entry point
 * counts its own launches).  bench.py reports the difference over its timed region as gpu_launches. */
TREX_API unsigned long long trex_kernel_launches(void);

/* ---- tree schedule ------------------------------------------------------------------------
 * Replaces the per-node `jnp.where(adj[:, node] == 1, size=2, fill_value=-1)` of
 * src/trex/sankoff.py:60 and the fori_loop node order of :87-92 by a device-built plan:
 * children of every internal node (first two rows equal to 1, ascending, -1 when missing)
 * and a DFS post-order in which a child's row is consumed right after it is produced.
 * `zero_root_selfloop` applies run_sankoff's `adj[-1,-1] = 0` (:141).  adj: [batch][n_all][n_all] f32.
 * The plan also holds a decomposition into independent subtrees ("segments": contiguous runs of the
 * post-order) plus the list of the nodes above them; on short alignments (too few sites to fill the
 * GPU) the walk kernels process the segments side by side in one launch and the top part in a second.
 * plan: [batch][trex_sankoff_plan_words()] int32 (opaque to the caller). */
TREX_API size_t trex_sankoff_plan_words(int n_all, int n_leaves);
TREX_API int trex_sankoff_plan(const float* adj, int n_all, int n_leaves, int batch,
                      int zero_root_selfloop, int32_t* plan, void* stream);
/* Same plan, with the number of subtree segments tuned to the alignment the walks will run on (few sites per
 * GPU -> fewer, larger segments and a shorter serial top part; n_sites <= 0 = unknown = trex_sankoff_plan).
 * Any plan is valid for any n_sites: the hint only moves time.  Replaces the same reference lines. */
TREX_API int trex_sankoff_plan_for_sites(const float* adj, int n_all, int n_leaves, int batch,
                                int zero_root_selfloop, int64_t n_sites, int32_t* plan, void* stream);

/* ---- leaf encoding ------------------------------------------------------------------------
 * `sequences[i].astype(int32)` + scatter index rules of src/trex/sankoff.py:49-52 (negative
 * wraps, out-of-range dropped) applied once: f32 [n_leaves][in_stride] -> u8 [n_leaves][site_stride]. */
TREX_API int trex_encode_leaf_states(const float* seqs, int n_leaves, int64_t n_sites, int64_t in_stride,
                            int n_states, uint8_t* leaf_states, int64_t site_stride, void* stream);

/* ---- hard Sankoff forward -----------------------------------------------------------------
 * Replaces run_dp / vectorized_dp / the table part of run_sankoff
 * (src/trex/sankoff.py:24-97, 141-160, 187).
 *   leaf_states [n_leaves][site_stride] u8 (shared by all trees of the batch)
 *   cost        [A][A] f32
 *   dp          [batch][n_int][A][site_stride] f32 out (leaf rows are implicit)
 *   argmin      [batch][n_int][A][site_stride][2] u8 out, or NULL (score/dp only): the back-pointers of the two
 *               children (slot 0 = first child) of a (state, site) are adjacent bytes
 *   root_min    [batch][site_stride] f32 out = min_s dp[root][s] per site, or NULL
 *   total       [batch] f64 out = sum over sites of root_min, or NULL */
TREX_API int trex_sankoff_hard_fwd(const int32_t* plan, int n_all, int n_leaves, int n_states,
                          int64_t n_sites, int64_t site_stride, int batch,
                          const uint8_t* leaf_states, const float* cost, float* dp,
                          uint8_t* argmin, float* root_min, double* total, int mode,
                          void* stream);

/* ---- hard backward: ancestral-state traceback (+ optional edge histogram) ------------------
 * Replaces the root argmin and backtrack_sankoff_jit (src/trex/sankoff.py:164-185, 191-267).
 *   anc_states [batch][n_int][site_stride] u8 out (nodes unreachable from the root: 0)
 *   edge_hist  [batch][A][A] u64 out or NULL: # of (parent state s, child state t) over all
 *              edges and sites of the reconstruction = sub-gradient of the score w.r.t. cost. */
TREX_API int trex_sankoff_traceback(const int32_t* plan, int n_all, int n_leaves, int n_states,
                           int64_t n_sites, int64_t site_stride, int batch,
                           const uint8_t* leaf_states, const float* dp, const uint8_t* argmin,
                           uint8_t* anc_states, unsigned long long* edge_hist, void* stream);
/* The same for an `argmin` buffer written by trex_sankoff_hard_fwd(..., cost, ..., mode): `cost` [A][A] f32 and `mode`
 * must be the ones of that call (mode may carry TREX_MODE_COMPACT_ARG; without the flag this is
 * trex_sankoff_traceback).  The kernel repeats the forward's device-side test of the cost matrix to know whether it
 * reads pointer words or the byte table. */
TREX_API int trex_sankoff_traceback_cost(const int32_t* plan, int n_all, int n_leaves, int n_states,
                                int64_t n_sites, int64_t site_stride, int batch,
                                const uint8_t* leaf_states, const float* cost, int mode, const float* dp,
                                const uint8_t* argmin, uint8_t* anc_states, unsigned long long* edge_hist,
                                void* stream);

/* ---- exports to the reference's layouts (what run_sankoff returns, sankoff.py:161-188) ------
 *   dp_ref    [L][n_all][A] f32 incl. the explicit leaf rows (1e5 / 0)
 *   recon_ref [n_all][L] f32: leaf rows = leaf_seqs_f32 (may be NULL -> decoded states),
 *             ancestor rows = anc_states (or 0 when anc_states is NULL: return_path=False) */
TREX_API int trex_sankoff_export_dp(int n_all, int n_leaves, int n_states, int64_t n_sites,
                           int64_t site_stride, const uint8_t* leaf_states, const float* dp,
                           float* dp_ref, void* stream);
TREX_API int trex_sankoff_export_recon(int n_all, int n_leaves, int64_t n_sites, int64_t site_stride,
                              const float* leaf_seqs_f32, int64_t leaf_in_stride,
                              const uint8_t* leaf_states, const uint8_t* anc_states,
                              float* recon_ref, void* stream);
/* backtracking table of run_dp (sankoff.py:80-83): bt_ref [L][n_all][A][4] f32 */
TREX_API int trex_sankoff_export_bt(const int32_t* plan, int n_all, int n_leaves, int n_states,
                           int64_t n_sites, int64_t site_stride, const uint8_t* argmin,
                           float* bt_ref, void* stream);

/* backtrack_sankoff_jit (sankoff.py:191-267) on the reference's own table layout:
 *   bt_ref [n_sites][n_all][A][4] f32, root_node [1] i32, root_state [n_sites] i32
 *   recon  [n_sites][n_all] i32 out (rows of leaves / unreached nodes are 0, as in the reference) */
TREX_API int trex_sankoff_backtrack_ref(const float* bt_ref, int n_all, int n_leaves, int n_states,
                               int64_t n_sites, const int32_t* root_node, const int32_t* root_state,
                               int32_t* recon, void* stream);

/* The library accumulates totals and dC in f64.  The reference runs with 64-bit types disabled
 * (src/trex/types.py:12), so a jax.ffi binding cannot declare F64 results: this casts a small f64 vector to the
 * f32 the reference returns (`total_cost` of sankoff.py:187 is an f32 sum).  dst[i] = (float)src[i]. */
TREX_API int trex_cast_f64_to_f32(const double* src, float* dst, int64_t n, void* stream);

/* ---- relaxed (softmin) Sankoff forward / reverse (not in the reference) --
 * min_t x -> -T*logsumexp_t(-x/T) in the recurrence of src/trex/sankoff.py:62-79 and :187.
 * Leaves: either `leaf_states` (u8, hard leaves = 1e5*(1-onehot)) or `leaf_cost`
 * [n_leaves][A][site_stride] f32 (soft leaves); exactly one is non-NULL.
 *   dp     [batch][n_int][A][site_stride] f32 out
 *   total  [batch] f64 out = sum over sites of softmin_s dp[root][s] */
TREX_API int trex_sankoff_soft_fwd(const int32_t* plan, int n_all, int n_leaves, int n_states,
                          int64_t n_sites, int64_t site_stride, int batch,
                          const uint8_t* leaf_states, const float* leaf_cost, const float* cost,
                          float temperature, float* dp, double* total, void* stream);
/*   gscale  device f32 scalar (upstream cotangent) or NULL = 1
 *   gbar    [batch][n_int][A][site_stride] f32 out: d total / d dp[node] = soft ancestral marginals
 *   dleaf   [batch][n_leaves][A][site_stride] f32 out or NULL: d total / d leaf_cost (rows of leaves under nodes
 *           the root does not reach are written as zeros; rows of leaves that are no node's child are not written)
 *   dcost   [batch][A][A] f64 out or NULL */
TREX_API int trex_sankoff_soft_bwd(const int32_t* plan, int n_all, int n_leaves, int n_states,
                          int64_t n_sites, int64_t site_stride, int batch,
                          const uint8_t* leaf_states, const float* leaf_cost, const float* cost,
                          float temperature, const float* dp, const float* gscale, float* gbar,
                          float* dleaf, double* dcost, void* stream);

/* Same reverse pass for leaves given as STATE CODES, with the leaf gradient in compact form.  For such a
 * leaf d total / d leaf_cost[leaf][t] is non-zero only at the observed state t = o (every other entry of the
 * hard leaf row 1e5*(1-onehot) has softmin weight exp(-1e5/T) = 0), so the dense [leaf][A][site] table of
 * trex_sankoff_soft_bwd would be (A-1)/A zeros — 7.8 of the 16.3 GB that pass writes at 1024 taxa x 100k
 * sites x 20 states.  Here:
 *   dleaf_obs [batch][n_leaves][site_stride] f32 out or NULL: sum_t d total / d leaf_cost[leaf][site][t]
 *             = the entry at the observed state (for a leaf without a state: the sum over its row);
 *             rows of leaves the root does not reach are not written. */
TREX_API int trex_sankoff_soft_bwd_compact(const int32_t* plan, int n_all, int n_leaves, int n_states,
                                  int64_t n_sites, int64_t site_stride, int batch,
                                  const uint8_t* leaf_states, const float* cost, float temperature,
                                  const float* dp, const float* gscale, float* gbar, float* dleaf_obs,
                                  double* dcost, void* stream);

/* ---- soft-adjacency tree cost, value + gradients in one pass --------------------------------
 * Replaces compute_surrogate_cost / compute_soft_cost (src/trex/tree.py:163-209, 212-266)
 * and what jax.value_and_grad derives from them (src/trex/evals/benchmark.py:167-173).
 *   seqs [N][K] f32, K = L*Q row-major (the reference's (N,L,Q) array, flat)
 *   adj  [N][N] f32
 *   cost NULL | [Q] (cost_kind 1) | [Q][Q] (cost_kind 2)
 *   out_cost f32 scalar; d_seqs [N][K] or NULL; d_adj [N][N] or NULL; d_cost [Q]/[Q][Q] or NULL
 *   workspace: trex_tree_cost_workspace_bytes(N, K, Q) bytes */
TREX_API size_t trex_tree_cost_workspace_bytes(int n_nodes, int64_t k, int n_states);
TREX_API int trex_tree_cost_fwd_bwd(const float* seqs, const float* adj, const float* cost, int cost_kind,
                           int n_nodes, int64_t n_sites, int n_states, float* out_cost,
                           float* d_seqs, float* d_adj, float* d_cost, void* workspace,
                           void* stream);

/* ---- the small operators around the tree cost and their reverse modes ------
 * T3 update_seq (src/trex/tree.py:110-130; stacked twin src/trex/evals/benchmark.py:210-232):
 *    seqs_ancestor_rows [n_anc][L][Q] = softmax(logits * temperature) over the last axis. */
TREX_API int trex_update_seq_fwd(const float* logits, int n_ancestors, int64_t n_sites, int n_states,
                        float temperature, float* seqs_ancestor_rows, void* stream);
TREX_API int trex_update_seq_bwd(const float* seqs_ancestor_rows, const float* d_seqs_ancestor_rows,
                        int n_ancestors, int64_t n_sites, int n_states, float temperature,
                        float* d_logits, void* stream);
/* T4 update_tree (src/trex/tree.py:50-107): tree_params/gumbel_noise/gates [(N-1)][n_anc]
 *    (noise and gates may be NULL; jax.random streams are not reproducible outside JAX, so the
 *    Gumbel draw is an input), adjacency [N][N] out, rows sum to 1. */
TREX_API int trex_update_tree_fwd(const float* tree_params, const float* gumbel_noise, const float* gates,
                         float temperature, int n_nodes, int n_leaves, float* adjacency, void* stream);
TREX_API int trex_update_tree_bwd(const float* adjacency, const float* d_adjacency, const float* gates,
                         float temperature, int n_nodes, int n_leaves, float* d_tree_params, void* stream);
/* T5 enforce_graph_constraints (src/trex/tree.py:133-160); d_adjacency [N][N] out or NULL;
 *    scratch: one f64. */
TREX_API int trex_graph_constraint(const float* adjacency, int n_nodes, float scaling_factor, float* out_loss,
                          float* d_adjacency, double* scratch, void* stream);
/* T7 discretize_tree_topology (src/trex/tree.py:31-47); parent_scratch [N] i32. */
TREX_API int trex_discretize_topology(const float* adjacency, int n_nodes, int32_t* parent_scratch,
                             float* one_hot, void* stream);
/* T6 compute_cost (src/trex/tree.py:269-296); state_scratch [N][L] i32, parent_scratch [N] i32,
 *    acc_scratch one f64. */
TREX_API int trex_compute_cost(const float* seqs_one_hot, const float* adjacency, const float* subst,
                      int n_nodes, int64_t n_sites, int n_states, int32_t* state_scratch,
                      int32_t* parent_scratch, double* acc_scratch, float* out_cost, void* stream);

/* ---- allocation-free variant of trex_compute_cost --------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).  Same result as
 * trex_compute_cost, but the three scratch buffers (state [N][L] i32, parent [N] i32, one f64 accumulator) live
 * in a static device array inside the library instead of being passed in: 256 + 4 * TREX_NOALLOC_MAX_NODES +
 * 4 * TREX_NOALLOC_MAX_CELLS bytes (= 4 MiB + 16.25 KiB), resident on every GPU that loads the library until
 * the process exits.  The call makes no cudaMalloc / cudaFree and calls no allocating library routine; the caller
 * still passes inputs and the output.
 *   Limits: n_nodes <= TREX_NOALLOC_MAX_NODES and n_nodes * n_sites <= TREX_NOALLOC_MAX_CELLS; anything larger is
 *   rejected with TREX_ERR_INVALID_ARGUMENT (trex_last_error names the limit) before any CUDA call is made.
 *   Sharing: every call on a device uses the SAME array, which keeps the previous call's contents (the
 *   accumulator is cleared by each call, the two tables are fully overwritten).  Calls that overlap on one GPU
 *   (different streams, or concurrently replayed graphs) RACE on it: serialise them, or use trex_compute_cost with
 *   per-call scratch. */
#define TREX_NOALLOC_MAX_NODES 4096
#define TREX_NOALLOC_MAX_CELLS (1LL << 20)
TREX_API size_t trex_compute_cost_noalloc_scratch_bytes(void);
TREX_API int trex_compute_cost_noalloc(const float* seqs_one_hot, const float* adjacency, const float* subst,
                              int n_nodes, int64_t n_sites, int n_states, float* out_cost, void* stream);

/* ---- whole ancestor-fitting loop on the device ------------------------------
 * Replaces the body of run_trex_optimization_configurable (src/trex/evals/benchmark.py:75-200):
 * n_iterations of { loss, grads = value_and_grad(loss_fn)(ancestors); optimizer.update; apply_updates }
 * with loss_fn = compute_surrogate_cost(_update_seq_stacked(ancestors, masked_sequences, n_leaves, 1.0), adj)
 * (:152-165; compute_soft_cost with the identity cost matrix is the same function) and
 * optimizer = optax.chain(clip_by_global_norm(clip_norm), adam | sgd(momentum 0.9) | rmsprop | adamw(0.01))
 * (create_optimizer, :41-72), as ONE persistent cooperative kernel (one grid barrier per iteration).
 *   adjacency [N][N] f32 (any soft or hard adjacency); leaf_seqs [n_leaves][L][Q] f32 (one-hot or soft rows)
 *   ancestor_logits [n_anc][L][Q] f32 in/out; opt_mom1 / opt_mom2 [n_anc][L][Q] f32 in/out (zeros at step 0):
 *     first / second moment (adam, adamw), momentum trace in opt_mom1 (sgd), mean square in opt_mom2 (rmsprop)
 *   optimizer: TREX_OPT_*; clip_norm <= 0 disables clipping; step_count = updates already applied (bias correction)
 *   out_losses [n_iterations] f32 or NULL (loss BEFORE each update, as lax.scan returns them, :176-188)
 *   out_argmax [n_anc][L] i32 or NULL (jnp.argmax(ancestors, -1) after the last update, :200)
 *   workspace: trex_fit_ancestors_workspace_bytes(...) bytes.  Limits (else TREX_ERR_UNSUPPORTED): with 4 or 20
 *   states up to 4096 ancestors as long as one site of all ancestors fits in shared memory (n_anc * n_states
 *   * 4 B + 64 B * n_anc <= 220 KB: 1023 ancestors at 20 states do); other alphabets: n_anc <= 128. */
#define TREX_OPT_ADAM 0
#define TREX_OPT_SGD 1
#define TREX_OPT_RMSPROP 2
#define TREX_OPT_ADAMW 3
TREX_API size_t trex_fit_ancestors_workspace_bytes(int n_nodes, int n_leaves, int64_t n_sites, int n_states);
TREX_API int trex_fit_ancestors(const float* adjacency, const float* leaf_seqs, int n_nodes, int n_leaves,
                       int64_t n_sites, int n_states, float* ancestor_logits, float* opt_mom1, float* opt_mom2,
                       int optimizer, float learning_rate, float clip_norm, int n_iterations, int64_t step_count,
                       float* out_losses, int32_t* out_argmax, void* workspace, void* stream);

/* ---- landscape-aware ("parental guidance") loss term --------------------------
 * compute_parental_logits (src/trex/evals/benchmark.py:586-663) and the fitness_cost part of
 * _compute_loss_landscape_aware_stacked (:283-303):  parent p = argmax_j Adj[n][j];
 * logits[n][i][s] = sum_j F_i[s * Q^k + j] * prod_m S[p][interactions[i][m]][digit_m(j)];
 * fitness_cost = -sum_{n,i} mask[i] sum_s S[n][i][s] log_softmax(logits[n][i])[s] / (#non-root nodes * sum(mask)).
 *   seqs [N][L][Q] f32 (the updated sequences); adjacency [N][N] f32 or NULL when parent_index [N] i32 is given
 *   (compute_parental_logits itself takes the parents' rows: pass them as seqs with parent_index = 0..N-1)
 *   interactions [L][k] i32; fitness_tables [L][Q^(k+1)] f32; seq_mask [L] f32 (0/1) or NULL; 0 <= k <= 3, Q <= 32
 *   out_fitness_cost f32 scalar or NULL; out_logits [N][L][Q] or NULL; d_seqs [N][L][Q] or NULL
 *   (d fitness_cost / d seqs, both the child and the parent path); parent_scratch [N] i32; acc_scratch 2 f64. */
TREX_API int trex_landscape_loss_fwd_bwd(const float* seqs, const float* adjacency, const int32_t* parent_index,
                                const int32_t* interactions, const float* fitness_tables, const float* seq_mask,
                                int n_nodes, int64_t n_sites, int n_states, int k, float* out_fitness_cost,
                                float* out_logits, float* d_seqs, int32_t* parent_scratch, double* acc_scratch,
                                void* stream);

/* ---- on-device input generators ---------------------------------------------------------
 * Random draws: jax.random (threefry) streams cannot be reproduced outside JAX, so every generator takes its draws
 * either as EXPLICIT device arrays (the parity tests feed the same arrays to the NumPy restatement of the reference
 * code and demand identical results) or, when the draw pointers are NULL, from a counter-based Philox4x32-10 stream
 * keyed by `seed` (reproducible; not JAX's stream).  State tables are u8 [node][site_stride].
 *
 * create_balanced_binary_tree (src/trex/evals/benchmark.py:781-791; kind 1) and the adjacency of
 * generate_groundtruth (src/trex/ground_truth.py:180-191; kind 0): adjacency [n_all][n_all] f32, no root self-loop. */
TREX_API int trex_tree_adjacency(int n_leaves, int kind, float* adjacency, void* stream);
/* generate_groundtruth (src/trex/ground_truth.py:113-197) with mutate (:20-53): root sequence all zeros, every node's
 * two children (2 p_i, 2 p_i + 1 of parent n_leaves + p_i) are the parent with `n_mutations` distinct sites moved to
 * (state + offset) % n_states, offset in 1..n_states-1.  n_leaves must be a power of two (the reference raises
 * ValueError, :142-144; here TREX_ERR_INVALID_ARGUMENT).
 *   mut_sites  [n_all-1][n_mutations] i32 or NULL: the sites each edge (= child node) mutates (negative wrap, OOB dropped)
 *   offsets    [n_all-1][n_sites] u8 or NULL:      the offset draw of every (edge, site)
 *   all_states [n_all][site_stride] u8 out (rows < n_leaves are the leaves = the reference's masked_sequences rows)
 *   adjacency  [n_all][n_all] f32 out or NULL */
TREX_API size_t trex_generate_groundtruth_workspace_bytes(int n_leaves, int n_mutations, int64_t n_sites);
TREX_API int trex_generate_groundtruth(int n_leaves, int n_states, int n_mutations, int64_t n_sites, int64_t site_stride,
                              uint64_t seed, const int32_t* mut_sites, const uint8_t* offsets, uint8_t* all_states,
                              float* adjacency, void* workspace, void* stream);
/* create_nk_model_landscape (src/trex/nk_model.py:17-43): interactions [n][k] i32 ~ randint(0, n) (may be NULL when k == 0),
 * fitness_tables [n][n_states^(k+1)] f32 ~ U[0,1). */
TREX_API int trex_nk_landscape(int64_t n, int k, int n_states, uint64_t seed, int32_t* interactions, float* fitness_tables,
                      void* stream);
/* get_fitness / batched_get_fitness (src/trex/nk_model.py:46-113): out[b] = sum_i mask_i f_i(seq_i, partners) / sum mask;
 * seqs [batch][seq_stride] u8, seq_mask [n] f32 or NULL. */
TREX_API int trex_nk_fitness(const uint8_t* seqs, int64_t seq_stride, int batch, int64_t n, int k, int n_states,
                    const int32_t* interactions, const float* fitness_tables, const float* seq_mask, float* out_fitness,
                    void* stream);
/* generate_tree_data (src/trex/nk_model.py:116-278): Metropolis-Hastings evolution down the tree, `branch_length` steps
 * per edge, each a coupled mutation (one site + its k partners redrawn, probability coupled_mutation_prob) or a random
 * mutation (every site redrawn with the edge's rate = min(1, mutation_rate * exp(N(0,1) * noise_std))), accepted with
 * probability min(1, exp(fitness(proposal) - fitness(current))).  Parents = argmax of each adjacency row.
 * ROOT: the reference takes "the node that is its own argmax", which needs a root self-loop; without one (the
 * adjacency of its own callers) it falls back to node 0 and produces a degenerate tree.  That defect is not
 * reproduced: here a node whose row has no positive entry is a root too.
 *   draw_noise [N]; draw_u_coupled / draw_site / draw_u_accept [N][branch_length]; draw_u_mask (f32) / draw_new_vals (u8)
 *   [N][branch_length][n_sites] — each may be NULL (Philox).  all_states [N][site_stride] u8 out. */
TREX_API size_t trex_nk_generate_tree_workspace_bytes(int n_nodes, int64_t site_stride);
TREX_API int trex_nk_generate_tree(const float* adjacency, int n_nodes, const uint8_t* root_sequence, int64_t n_sites,
                          int64_t site_stride, int k, int n_states, const int32_t* interactions,
                          const float* fitness_tables, float mutation_rate, float coupled_mutation_prob,
                          float mutation_rate_noise_std, int branch_length, uint64_t seed, const float* draw_noise,
                          const float* draw_u_coupled, const int32_t* draw_site, const float* draw_u_mask,
                          const uint8_t* draw_new_vals, const float* draw_u_accept, uint8_t* all_states, void* workspace,
                          void* stream);

/* ---- file formats ----------------------------------------------------------------------
 * The reference has no on-disk or wire format at all (its inputs exist only as in-memory arrays), so these replace
 * nothing and have nothing to be parity-checked against; what they guarantee is the INPUT CONTRACT of the hot path
 * (src/trex/sankoff.py:45-52, 60, 87-92; src/trex/utils/types.py:33).  Pure host code: every pointer is a HOST
 * pointer, no CUDA call is made.
 *
 * FASTA / PHYLIP (sequential or interleaved, relaxed names) -> u8 states [taxon][site_stride]: index of the character
 * in `alphabet` (case-insensitive, e.g. "ACGT" or "ARNDCQEGHILKMFPSTWYV"), 255 = no state (gap, ambiguity code,
 * anything else; pad columns too).  Call once with states_host = names_host = NULL to get n_taxa / n_sites.
 *   names_host: the taxon names, '\n'-separated and NUL-terminated (first word of each header), or NULL. */
#define TREX_FORMAT_FASTA 0
#define TREX_FORMAT_PHYLIP 1
TREX_API int trex_parse_alignment_host(const char* text, size_t text_len, int format, const char* alphabet, int* n_taxa,
                              int64_t* n_sites, uint8_t* states_host, int64_t site_stride, char* names_host,
                              size_t names_capacity);
/* Newick -> rooted BINARY tree in the node order run_dp requires: leaves 0..n-1 (row order of `taxon_names`, a
 * '\n'-separated list as returned above, or order of appearance when NULL), internal nodes n..2n-2 in post-order —
 * every child index is below its parent's, the root is the last node.  Branch lengths, comments and internal labels
 * are ignored; unary nodes are transparent; a node with more than two children is an error (the reference's DP reads
 * exactly two children, sankoff.py:60).
 *   parent_host [2n-1] i32 (root: -1) or NULL; adjacency_host [2n-1][2n-1] f32 with adjacency[i][parent[i]] = 1 and
 *   no root self-loop, or NULL; names_host: leaf names in leaf-index order, or NULL.  n_leaves is always written. */
TREX_API int trex_parse_newick_host(const char* text, size_t text_len, const char* taxon_names, int* n_leaves,
                           int32_t* parent_host, float* adjacency_host, char* names_host, size_t names_capacity);

#ifdef __cplusplus
}
#endif
#endif /* TREX_B200_H_ */
