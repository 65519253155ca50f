// XLA-FFI shim over the plain C ABI of libtrex_b200.so (include/trex_b200.h).
//
// OPTIONAL translation unit: it needs jaxlib's headers (`python -c "import jax; print(jax.ffi.
// include_dir())"`), which do not exist in the build container, so it is NOT part of
// `make all`; `make ffi JAX_FFI_INCLUDE=<dir>` builds libtrex_b200_ffi.so where they do.  Every
// handler only unpacks XLA buffers/attributes and forwards to the C entry point on XLA's stream;
// scratch comes from XLA's ScratchAllocator, nothing is allocated, freed or synchronised here,
// so the handlers are CUDA-graph (command-buffer) compatible.  Python side: trex_b200/jax_ffi.py.
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

// run_sankoff (src/trex/sankoff.py:114-188): adj (n_all,n_all) f32, cost (A,A) f32,
// seqs (rows>=n_leaves, L) f32 -> recon (n_all,L) f32, dp (L,n_all,A) f32, total () f32.
// Native tables live in XLA scratch for the duration of the call.
ffi::Error RunSankoffImpl(cudaStream_t stream, ffi::ScratchAllocator scratch,
                          ffi::Buffer<ffi::F32> adj, ffi::Buffer<ffi::F32> cost,
                          ffi::Buffer<ffi::F32> seqs, int32_t n_leaves, bool return_path,
                          ffi::ResultBuffer<ffi::F32> recon, ffi::ResultBuffer<ffi::F32> dp_ref,
                          ffi::ResultBuffer<ffi::F64> total) {
  const int n_all = (int)adj.dimensions()[0];
  const int A = (int)cost.dimensions()[0];
  const int64_t L = seqs.dimensions()[1];
  const int64_t stride = pad16(L);
  const int n_int = n_all - n_leaves;
  auto take = [&](size_t bytes) -> void* {
    auto p = scratch.Allocate(bytes);
    return p.has_value() ? *p : nullptr;
  };
  auto* plan = (int32_t*)take(sizeof(int32_t) * trex_sankoff_plan_words(n_all, n_leaves));
  auto* leaf = (uint8_t*)take((size_t)n_leaves * stride);
  auto* dp = (float*)take(sizeof(float) * (size_t)n_int * A * stride);
  auto* arg = return_path ? (uint8_t*)take((size_t)n_int * 2 * A * stride) : nullptr;
  auto* anc = return_path ? (uint8_t*)take((size_t)n_int * stride) : nullptr;
  if (!plan || !leaf || !dp || (return_path && (!arg || !anc)))
    return ffi::Error::Internal("trex run_sankoff: scratch allocation failed");
  TREX_TRY(trex_sankoff_plan(adj.typed_data(), n_all, n_leaves, 1, /*zero_root_selfloop=*/1, plan, stream));
  TREX_TRY(trex_encode_leaf_states(seqs.typed_data(), n_leaves, L, L, A, leaf, stride, stream));
  TREX_TRY(trex_sankoff_hard_fwd(plan, n_all, n_leaves, A, L, stride, 1, leaf, cost.typed_data(), dp, arg,
                                 nullptr, total->typed_data(), TREX_MODE_AUTO, stream));
  if (return_path)
    TREX_TRY(trex_sankoff_traceback(plan, n_all, n_leaves, A, L, stride, 1, leaf, dp, arg, anc, nullptr, stream));
  TREX_TRY(trex_sankoff_export_recon(n_all, n_leaves, L, stride, seqs.typed_data(), L, leaf, anc,
                                     recon->typed_data(), stream));
  TREX_TRY(trex_sankoff_export_dp(n_all, n_leaves, A, L, stride, leaf, dp, dp_ref->typed_data(), stream));
  return ffi::Error::Success();
}

// Relaxed Sankoff forward: leaf_cost native [n_leaves][A][stride] f32 -> dp native, total f64.
ffi::Error SoftFwdImpl(cudaStream_t stream, ffi::ScratchAllocator scratch, ffi::Buffer<ffi::F32> adj,
                       ffi::Buffer<ffi::F32> cost, ffi::Buffer<ffi::F32> leaf_cost, int32_t n_leaves,
                       int64_t n_sites, float temperature, ffi::ResultBuffer<ffi::S32> plan,
                       ffi::ResultBuffer<ffi::F32> dp, ffi::ResultBuffer<ffi::F64> total) {
  (void)scratch;
  const int n_all = (int)adj.dimensions()[0];
  const int A = (int)cost.dimensions()[0];
  const int64_t stride = leaf_cost.dimensions()[2];
  TREX_TRY(trex_sankoff_plan(adj.typed_data(), n_all, n_leaves, 1, 1, plan->typed_data(), stream));
  TREX_TRY(trex_sankoff_soft_fwd(plan->typed_data(), n_all, n_leaves, A, n_sites, stride, 1, nullptr,
                                 leaf_cost.typed_data(), cost.typed_data(), temperature, dp->typed_data(),
                                 total->typed_data(), stream));
  return ffi::Error::Success();
}

// Relaxed Sankoff reverse: residuals (plan, dp) + cotangent g () f32 ->
// gbar native (soft ASR marginals), d leaf_cost native, d cost (A,A) f64.
ffi::Error SoftBwdImpl(cudaStream_t stream, ffi::Buffer<ffi::S32> plan, ffi::Buffer<ffi::F32> cost,
                       ffi::Buffer<ffi::F32> leaf_cost, ffi::Buffer<ffi::F32> dp, ffi::Buffer<ffi::F32> g,
                       int32_t n_all, int32_t n_leaves, int64_t n_sites, float temperature,
                       ffi::ResultBuffer<ffi::F32> gbar, ffi::ResultBuffer<ffi::F32> dleaf,
                       ffi::ResultBuffer<ffi::F64> dcost) {
  const int A = (int)cost.dimensions()[0];
  const int64_t stride = leaf_cost.dimensions()[2];
  TREX_TRY(trex_sankoff_soft_bwd(plan.typed_data(), n_all, n_leaves, A, n_sites, stride, 1, nullptr,
                                 leaf_cost.typed_data(), cost.typed_data(), temperature, dp.typed_data(),
                                 g.typed_data(), gbar->typed_data(), dleaf->typed_data(), dcost->typed_data(),
                                 stream));
  return ffi::Error::Success();
}

// compute_soft_cost / compute_surrogate_cost value + gradients (src/trex/tree.py:163-266).
// cost_kind: 0 none (cost buffer ignored, pass a 1-element dummy), 1 vector (Q,), 2 matrix (Q,Q).
ffi::Error TreeCostImpl(cudaStream_t stream, ffi::ScratchAllocator scratch, ffi::Buffer<ffi::F32> seqs,
                        ffi::Buffer<ffi::F32> adj, ffi::Buffer<ffi::F32> cost, int32_t cost_kind,
                        ffi::ResultBuffer<ffi::F32> out_cost, ffi::ResultBuffer<ffi::F32> d_seqs,
                        ffi::ResultBuffer<ffi::F32> d_adj, ffi::ResultBuffer<ffi::F32> d_cost) {
  const int N = (int)seqs.dimensions()[0];
  const int64_t L = seqs.dimensions()[1];
  const int Q = (int)seqs.dimensions()[2];
  auto ws = scratch.Allocate(trex_tree_cost_workspace_bytes(N, L * Q, Q));
  if (!ws.has_value()) return ffi::Error::Internal("trex tree_cost: scratch allocation failed");
  TREX_TRY(trex_tree_cost_fwd_bwd(seqs.typed_data(), adj.typed_data(), cost_kind ? cost.typed_data() : nullptr,
                                  cost_kind, N, L, Q, out_cost->typed_data(), d_seqs->typed_data(),
                                  d_adj->typed_data(), cost_kind ? d_cost->typed_data() : nullptr, *ws, stream));
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
                                  .Ret<ffi::Buffer<ffi::F64>>());

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
                                  .Ret<ffi::Buffer<ffi::F64>>());

XLA_FFI_DEFINE_HANDLER_SYMBOL(trex_soft_bwd_ffi, SoftBwdImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
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
                                  .Ret<ffi::Buffer<ffi::F64>>());

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
