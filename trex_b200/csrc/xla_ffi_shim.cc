// XLA-FFI shim over the plain C ABI of libtrex_b200.so (include/trex_b200.h).
//
// OPTIONAL translation unit: it needs jaxlib's headers (`python -c "import jax; print(jax.ffi.
// include_dir())"`), which do not exist in the build container, so it is NOT part of
// `make all`; `make ffi JAX_FFI_INCLUDE=<dir>` builds libtrex_b200_ffi.so where they do.  Every
// handler only unpacks XLA buffers/attributes and forwards to the C entry point on XLA's stream;
// scratch comes from XLA's ScratchAllocator, nothing is allocated, freed or synchronised here,
// so the handlers are CUDA-graph (command-buffer) compatible.  Python side: trex_b200/jax_ffi.py.
//
// UNVERIFIED: neither jaxlib nor its headers exist on any machine of this project, so this file has never been
// compiled or run.  Conventions chosen so that it has a chance to work first time where they do exist:
//  * vmap: the Python side uses vmap_method="expand_dims", so under vmap every operand arrives with ONE leading
//    batch dimension (size B for mapped operands, size 1 for unmapped ones) and every result carries a leading B.
//    Handlers accept both ranks.  vmap over candidate trees with a shared alignment and cost matrix (the use the
//    reference designs for, src/trex/evals/benchmark.py:92-93) is ONE batched launch of the library; anything
//    else loops over the batch inside the handler.
//  * no F64 results: the reference runs with 64-bit types disabled (src/trex/types.py:12), so totals and dC are
//    accumulated in f64 scratch by the library and returned as f32 (trex_cast_f64_to_f32).
#include <cuda_runtime.h>

#include <cstdint>

#include "../../include/trex_b200.h"
#include "xla/ffi/api/ffi.h"

namespace ffi = xla::ffi;

namespace {

ffi::Error status(int rc) {
  if (rc == TREX_OK) return ffi::Error::Success();
  if (rc == TREX_ERR_INVALID_ARGUMENT) return ffi::Error::InvalidArgument(trex_last_error());
  return ffi::Error::Internal(trex_last_error());
}

#define TREX_TRY(expr)                      \
  do {                                      \
    int _rc = (expr);                       \
    if (_rc != TREX_OK) return status(_rc); \
  } while (0)

inline int64_t pad16(int64_t x) { return (x + 15) / 16 * 16; }

// leading batch size of a buffer whose unbatched rank is `rank` (expand_dims convention: at most one extra dim)
template <typename B>
inline int64_t lead(const B& b, size_t rank) { return b.dimensions().size() > rank ? b.dimensions()[0] : 1; }
template <typename B>
inline int64_t dim_from_end(const B& b, size_t i) { return b.dimensions()[b.dimensions().size() - 1 - i]; }
template <typename B>
inline int64_t elems(const B& b, size_t rank) {
  int64_t n = 1;
  for (size_t i = 0; i < rank; ++i) n *= dim_from_end(b, i);
  return n;
}

// run_sankoff (src/trex/sankoff.py:114-188): adj ([B,] n_all, n_all) f32, cost ([B|1,] A, A) f32,
// seqs ([B|1,] rows >= n_leaves, L) f32 -> recon ([B,] n_all, L) f32, dp ([B,] L, n_all, A) f32, total ([B]) f32.
// Native tables live in XLA scratch for the duration of the call.
ffi::Error RunSankoffImpl(cudaStream_t stream, ffi::ScratchAllocator scratch,
                          ffi::Buffer<ffi::F32> adj, ffi::Buffer<ffi::F32> cost,
                          ffi::Buffer<ffi::F32> seqs, int32_t n_leaves, bool return_path,
                          ffi::ResultBuffer<ffi::F32> recon, ffi::ResultBuffer<ffi::F32> dp_ref,
                          ffi::ResultBuffer<ffi::F32> total) {
  const int n_all = (int)dim_from_end(adj, 0);
  const int A = (int)dim_from_end(cost, 0);
  const int64_t L = dim_from_end(seqs, 0);
  const int64_t rows = dim_from_end(seqs, 1);
  const int64_t stride = pad16(L);
  const int n_int = n_all - n_leaves;
  const int64_t Ba = lead(adj, 2), Bc = lead(cost, 2), Bs = lead(seqs, 2);
  const int64_t B = Ba > Bc ? (Ba > Bs ? Ba : Bs) : (Bc > Bs ? Bc : Bs);
  if ((Ba != 1 && Ba != B) || (Bc != 1 && Bc != B) || (Bs != 1 && Bs != B))
    return ffi::Error::InvalidArgument("trex run_sankoff: inconsistent batch dimensions");
  // candidate trees on one alignment and one cost matrix: a single batched launch; otherwise one launch per element
  const bool shared = Bc == 1 && Bs == 1;
  const int64_t nb = shared ? B : 1;  // trees per launch
  auto take = [&](size_t bytes) -> void* {
    auto p = scratch.Allocate(bytes);
    return p.has_value() ? *p : nullptr;
  };
  const size_t plan_words = trex_sankoff_plan_words(n_all, n_leaves);
  auto* plan = (int32_t*)take(sizeof(int32_t) * plan_words * nb);
  auto* leaf = (uint8_t*)take((size_t)n_leaves * stride);
  auto* dp = (float*)take(sizeof(float) * (size_t)nb * n_int * A * stride);
  auto* tot64 = (double*)take(sizeof(double) * (size_t)B);
  auto* arg = return_path ? (uint8_t*)take((size_t)nb * n_int * 2 * A * stride) : nullptr;
  auto* anc = return_path ? (uint8_t*)take((size_t)nb * n_int * stride) : nullptr;
  if (!plan || !leaf || !dp || !tot64 || (return_path && (!arg || !anc)))
    return ffi::Error::Internal("trex run_sankoff: scratch allocation failed");
  const size_t adj_sz = (size_t)n_all * n_all, cost_sz = (size_t)A * A, seq_sz = (size_t)rows * L;
  const size_t recon_sz = (size_t)n_all * L, dpref_sz = (size_t)L * n_all * A;
  for (int64_t b0 = 0; b0 < B; b0 += nb) {
    const float* adj_b = adj.typed_data() + (Ba == 1 ? 0 : b0 * adj_sz);
    const float* cost_b = cost.typed_data() + (Bc == 1 ? 0 : b0 * cost_sz);
    const float* seq_b = seqs.typed_data() + (Bs == 1 ? 0 : b0 * seq_sz);
    if (Ba == 1 && nb > 1) return ffi::Error::InvalidArgument("trex run_sankoff: batch without a batched adjacency");
    TREX_TRY(trex_sankoff_plan_for_sites(adj_b, n_all, n_leaves, (int)nb, /*zero_root_selfloop=*/1, L, plan, stream));
    TREX_TRY(trex_encode_leaf_states(seq_b, n_leaves, L, L, A, leaf, stride, stream));
    // run_sankoff returns the reconstruction, not the backtracking table: let the library write pointer words when the
    // (traced) cost matrix turns out uniform; the traceback is given the same matrix and mode and reads either form
    const int mode = TREX_MODE_AUTO | TREX_MODE_COMPACT_ARG;
    TREX_TRY(trex_sankoff_hard_fwd(plan, n_all, n_leaves, A, L, stride, (int)nb, leaf, cost_b, dp, arg, nullptr,
                                   tot64 + b0, mode, stream));
    if (return_path)
      TREX_TRY(trex_sankoff_traceback_cost(plan, n_all, n_leaves, A, L, stride, (int)nb, leaf, cost_b, mode, dp, arg, anc,
                                           nullptr, stream));
    for (int64_t t = 0; t < nb; ++t) {  // exports are per tree
      TREX_TRY(trex_sankoff_export_recon(n_all, n_leaves, L, stride, seq_b, L, leaf,
                                         anc ? anc + (size_t)t * n_int * stride : nullptr,
                                         recon->typed_data() + (size_t)(b0 + t) * recon_sz, stream));
      TREX_TRY(trex_sankoff_export_dp(n_all, n_leaves, A, L, stride, leaf, dp + (size_t)t * n_int * A * stride,
                                      dp_ref->typed_data() + (size_t)(b0 + t) * dpref_sz, stream));
    }
  }
  TREX_TRY(trex_cast_f64_to_f32(tot64, total->typed_data(), B, stream));
  return ffi::Error::Success();
}

// Relaxed Sankoff forward: adj ([B,] n_all, n_all), cost ([B|1,] A, A), leaf_cost native
// ([B|1,] n_leaves, A, stride) f32 -> plan ([B,] words) i32, dp native ([B,] n_int, A, stride), total ([B]) f32.
ffi::Error SoftFwdImpl(cudaStream_t stream, ffi::ScratchAllocator scratch, ffi::Buffer<ffi::F32> adj,
                       ffi::Buffer<ffi::F32> cost, ffi::Buffer<ffi::F32> leaf_cost, int32_t n_leaves,
                       int64_t n_sites, float temperature, ffi::ResultBuffer<ffi::S32> plan,
                       ffi::ResultBuffer<ffi::F32> dp, ffi::ResultBuffer<ffi::F32> total) {
  const int n_all = (int)dim_from_end(adj, 0);
  const int A = (int)dim_from_end(cost, 0);
  const int64_t stride = dim_from_end(leaf_cost, 0);
  const int n_int = n_all - n_leaves;
  const int64_t Ba = lead(adj, 2), Bc = lead(cost, 2), Bl = lead(leaf_cost, 3);
  const int64_t B = Ba > Bc ? (Ba > Bl ? Ba : Bl) : (Bc > Bl ? Bc : Bl);
  const size_t plan_words = trex_sankoff_plan_words(n_all, n_leaves);
  auto t64 = scratch.Allocate(sizeof(double) * (size_t)B);
  if (!t64.has_value()) return ffi::Error::Internal("trex soft_fwd: scratch allocation failed");
  auto* tot64 = (double*)*t64;
  for (int64_t b = 0; b < B; ++b) {
    int32_t* plan_b = plan->typed_data() + (size_t)b * plan_words;
    TREX_TRY(trex_sankoff_plan_for_sites(adj.typed_data() + (Ba == 1 ? 0 : (size_t)b * n_all * n_all), n_all, n_leaves,
                                         1, 1, n_sites, plan_b, stream));
    TREX_TRY(trex_sankoff_soft_fwd(plan_b, n_all, n_leaves, A, n_sites, stride, 1, nullptr,
                                   leaf_cost.typed_data() + (Bl == 1 ? 0 : (size_t)b * n_leaves * A * stride),
                                   cost.typed_data() + (Bc == 1 ? 0 : (size_t)b * A * A), temperature,
                                   dp->typed_data() + (size_t)b * n_int * A * stride, tot64 + b, stream));
  }
  TREX_TRY(trex_cast_f64_to_f32(tot64, total->typed_data(), B, stream));
  return ffi::Error::Success();
}

// Relaxed Sankoff reverse: residuals (plan, dp) + cotangent g ([B]) f32 -> gbar native (soft ASR
// marginals), d leaf_cost native, d cost ([B,] A, A) f32.  Every operand carries the same leading B (or none).
ffi::Error SoftBwdImpl(cudaStream_t stream, ffi::ScratchAllocator scratch, ffi::Buffer<ffi::S32> plan,
                       ffi::Buffer<ffi::F32> cost, ffi::Buffer<ffi::F32> leaf_cost, ffi::Buffer<ffi::F32> dp,
                       ffi::Buffer<ffi::F32> g, int32_t n_all, int32_t n_leaves, int64_t n_sites, float temperature,
                       ffi::ResultBuffer<ffi::F32> gbar, ffi::ResultBuffer<ffi::F32> dleaf,
                       ffi::ResultBuffer<ffi::F32> dcost) {
  const int A = (int)dim_from_end(cost, 0);
  const int64_t stride = dim_from_end(leaf_cost, 0);
  const int n_int = n_all - n_leaves;
  const int64_t Bp = lead(plan, 1), Bc = lead(cost, 2), Bl = lead(leaf_cost, 3), Bd = lead(dp, 3), Bg = lead(g, 1);
  int64_t B = Bp;
  for (int64_t x : {Bc, Bl, Bd, Bg}) B = x > B ? x : B;
  const size_t plan_words = trex_sankoff_plan_words(n_all, n_leaves);
  auto d64 = scratch.Allocate(sizeof(double) * (size_t)B * A * A);
  if (!d64.has_value()) return ffi::Error::Internal("trex soft_bwd: scratch allocation failed");
  auto* dc64 = (double*)*d64;
  for (int64_t b = 0; b < B; ++b) {
    TREX_TRY(trex_sankoff_soft_bwd(plan.typed_data() + (Bp == 1 ? 0 : (size_t)b * plan_words), n_all, n_leaves, A, n_sites,
                                   stride, 1, nullptr,
                                   leaf_cost.typed_data() + (Bl == 1 ? 0 : (size_t)b * n_leaves * A * stride),
                                   cost.typed_data() + (Bc == 1 ? 0 : (size_t)b * A * A), temperature,
                                   dp.typed_data() + (Bd == 1 ? 0 : (size_t)b * n_int * A * stride),
                                   g.typed_data() + (Bg == 1 ? 0 : b), gbar->typed_data() + (size_t)b * n_int * A * stride,
                                   dleaf->typed_data() + (size_t)b * n_leaves * A * stride, dc64 + (size_t)b * A * A, stream));
  }
  TREX_TRY(trex_cast_f64_to_f32(dc64, dcost->typed_data(), B * A * A, stream));
  return ffi::Error::Success();
}

// compute_soft_cost / compute_surrogate_cost value + gradients (src/trex/tree.py:163-266), operands with an
// optional leading batch dimension (expand_dims convention).
// cost_kind: 0 none (cost buffer ignored, pass a 1-element dummy), 1 vector (Q,), 2 matrix (Q,Q).
ffi::Error TreeCostImpl(cudaStream_t stream, ffi::ScratchAllocator scratch, ffi::Buffer<ffi::F32> seqs,
                        ffi::Buffer<ffi::F32> adj, ffi::Buffer<ffi::F32> cost, int32_t cost_kind,
                        ffi::ResultBuffer<ffi::F32> out_cost, ffi::ResultBuffer<ffi::F32> d_seqs,
                        ffi::ResultBuffer<ffi::F32> d_adj, ffi::ResultBuffer<ffi::F32> d_cost) {
  const int N = (int)dim_from_end(seqs, 2);
  const int64_t L = dim_from_end(seqs, 1);
  const int Q = (int)dim_from_end(seqs, 0);
  const size_t crank = cost_kind == 2 ? 2 : 1;
  const int64_t Bs = lead(seqs, 3), Ba = lead(adj, 2), Bc = lead(cost, crank);
  const int64_t B = Bs > Ba ? (Bs > Bc ? Bs : Bc) : (Ba > Bc ? Ba : Bc);
  const size_t csz = (size_t)elems(cost, crank);
  auto ws = scratch.Allocate(trex_tree_cost_workspace_bytes(N, L * Q, Q));
  if (!ws.has_value()) return ffi::Error::Internal("trex tree_cost: scratch allocation failed");
  for (int64_t b = 0; b < B; ++b) {
    TREX_TRY(trex_tree_cost_fwd_bwd(seqs.typed_data() + (Bs == 1 ? 0 : (size_t)b * N * L * Q),
                                    adj.typed_data() + (Ba == 1 ? 0 : (size_t)b * N * N),
                                    cost_kind ? cost.typed_data() + (Bc == 1 ? 0 : b * csz) : nullptr, cost_kind, N, L, Q,
                                    out_cost->typed_data() + b, d_seqs->typed_data() + (size_t)b * N * L * Q,
                                    d_adj->typed_data() + (size_t)b * N * N,
                                    cost_kind ? d_cost->typed_data() + b * csz : nullptr, *ws, stream));
  }
  return ffi::Error::Success();
}

// run_trex_optimization_configurable (src/trex/evals/benchmark.py:75-200): the whole fori_loop / scan.
// `logits` is aliased to result 0 by the Python wrapper (input_output_aliases), the moments start at zero.
ffi::Error FitAncestorsImpl(cudaStream_t stream, ffi::ScratchAllocator scratch, ffi::Buffer<ffi::F32> adj,
                            ffi::Buffer<ffi::F32> leaf_seqs, ffi::Buffer<ffi::F32> logits, int32_t optimizer,
                            float learning_rate, float clip_norm, int32_t n_iterations,
                            ffi::ResultBuffer<ffi::F32> out_logits, ffi::ResultBuffer<ffi::F32> mom1,
                            ffi::ResultBuffer<ffi::F32> mom2, ffi::ResultBuffer<ffi::F32> losses,
                            ffi::ResultBuffer<ffi::S32> argmax) {
  const int N = (int)adj.dimensions()[0];
  const int n_leaves = (int)leaf_seqs.dimensions()[0];
  const int64_t L = leaf_seqs.dimensions()[1];
  const int Q = (int)leaf_seqs.dimensions()[2];
  const size_t bytes = sizeof(float) * (size_t)(N - n_leaves) * L * Q;
  if (out_logits->typed_data() != logits.typed_data())
    cudaMemcpyAsync(out_logits->typed_data(), logits.typed_data(), bytes, cudaMemcpyDeviceToDevice, stream);
  cudaMemsetAsync(mom1->typed_data(), 0, bytes, stream);
  cudaMemsetAsync(mom2->typed_data(), 0, bytes, stream);
  auto ws = scratch.Allocate(trex_fit_ancestors_workspace_bytes(N, n_leaves, L, Q));
  if (!ws.has_value()) return ffi::Error::Internal("trex fit_ancestors: scratch allocation failed");
  TREX_TRY(trex_fit_ancestors(adj.typed_data(), leaf_seqs.typed_data(), N, n_leaves, L, Q, out_logits->typed_data(),
                              mom1->typed_data(), mom2->typed_data(), optimizer, learning_rate, clip_norm,
                              n_iterations, 0, losses->typed_data(), argmax->typed_data(), *ws, stream));
  return ffi::Error::Success();
}

}  // namespace

XLA_FFI_DEFINE_HANDLER_SYMBOL(trex_run_sankoff_ffi, RunSankoffImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Ctx<ffi::ScratchAllocator>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Attr<int32_t>("n_leaves")
                                  .Attr<bool>("return_path")
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>());

XLA_FFI_DEFINE_HANDLER_SYMBOL(trex_soft_fwd_ffi, SoftFwdImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Ctx<ffi::ScratchAllocator>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Attr<int32_t>("n_leaves")
                                  .Attr<int64_t>("n_sites")
                                  .Attr<float>("temperature")
                                  .Ret<ffi::Buffer<ffi::S32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>());

XLA_FFI_DEFINE_HANDLER_SYMBOL(trex_soft_bwd_ffi, SoftBwdImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Ctx<ffi::ScratchAllocator>()
                                  .Arg<ffi::Buffer<ffi::S32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Attr<int32_t>("n_all")
                                  .Attr<int32_t>("n_leaves")
                                  .Attr<int64_t>("n_sites")
                                  .Attr<float>("temperature")
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>());

XLA_FFI_DEFINE_HANDLER_SYMBOL(trex_tree_cost_ffi, TreeCostImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Ctx<ffi::ScratchAllocator>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Attr<int32_t>("cost_kind")
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>());

XLA_FFI_DEFINE_HANDLER_SYMBOL(trex_fit_ancestors_ffi, FitAncestorsImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Ctx<ffi::ScratchAllocator>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Attr<int32_t>("optimizer")
                                  .Attr<float>("learning_rate")
                                  .Attr<float>("clip_norm")
                                  .Attr<int32_t>("n_iterations")
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::S32>>());
