"""Host-side mirror of the reference's ``trex.tree`` (``/root/reference/src/trex/tree.py``) for
the soft-adjacency tree cost, on top of ``trex_tree_cost_fwd_bwd`` (include/trex_b200.h).

``compute_surrogate_cost`` / ``compute_soft_cost`` keep the reference signatures and are
differentiable (torch autograd plays the role of ``jax.custom_vjp``): the single library call
returns the value and every gradient, which the backward pass merely scales by the upstream
cotangent.  Arrays are torch CUDA tensors; there is no CPU fallback.
"""

from __future__ import annotations

import torch

from . import _lib
from ._lib import call, ptr, stream_ptr


def tree_cost_value_and_grad(sequences, adjacency, cost_matrix=None, *, want_dseq=True,
                             want_dadj=True, want_dcost=True, workspace=None, out=None):
  """One pass of value + gradients of ``compute_soft_cost`` (tree.py:212-266).

  Returns ``(cost () f32, d_sequences (N,L,Q)|None, d_adjacency (N,N)|None, d_cost|None)`` —
  what ``jax.value_and_grad`` yields in ``src/trex/evals/benchmark.py:167-173``.
  """
  lib = _lib.load()
  if not sequences.is_cuda:
    raise RuntimeError("trex_b200 needs CUDA tensors; there is no CPU fallback")
  S = sequences.to(torch.float32).contiguous()
  adj = adjacency.to(device=S.device, dtype=torch.float32).contiguous()
  N, L, Q = S.shape
  kind = 0
  C = None
  if cost_matrix is not None:
    C = cost_matrix.to(device=S.device, dtype=torch.float32).contiguous()
    kind = 1 if C.dim() == 1 else 2
  if workspace is None:
    nbytes = lib.trex_tree_cost_workspace_bytes(N, L * Q, Q)
    workspace = torch.empty(nbytes, dtype=torch.uint8, device=S.device)
  if out is None:
    out = {}
  cost = out.get("cost")
  if cost is None:
    cost = torch.empty((), dtype=torch.float32, device=S.device)
  dS = (out.get("dS") if out.get("dS") is not None else torch.empty_like(S)) if want_dseq else None
  dA = (out.get("dA") if out.get("dA") is not None else torch.empty_like(adj)) if want_dadj else None
  dC = None
  if want_dcost and C is not None:
    dC = out.get("dC") if out.get("dC") is not None else torch.empty_like(C)
  call("trex_tree_cost_fwd_bwd", ptr(S), ptr(adj), ptr(C), kind, N, L, Q, ptr(cost), ptr(dS),
       ptr(dA), ptr(dC), ptr(workspace), stream_ptr())
  return cost, dS, dA, dC


class _TreeCost(torch.autograd.Function):

  @staticmethod
  def forward(ctx, sequences, adjacency, cost_matrix):
    need = ctx.needs_input_grad
    cost, dS, dA, dC = tree_cost_value_and_grad(
      sequences, adjacency, cost_matrix, want_dseq=need[0], want_dadj=need[1],
      want_dcost=cost_matrix is not None and need[2])
    ctx.grads = (dS, dA, dC)
    ctx.in_dtypes = (sequences.dtype, adjacency.dtype, None if cost_matrix is None else cost_matrix.dtype)
    return cost

  @staticmethod
  def backward(ctx, g):
    dS, dA, dC = ctx.grads
    dt = ctx.in_dtypes
    return (None if dS is None else (dS * g).to(dt[0]), None if dA is None else (dA * g).to(dt[1]),
            None if dC is None else (dC * g).to(dt[2]))


def compute_surrogate_cost(sequences, adjacency):
  """``trex.tree.compute_surrogate_cost`` (tree.py:163-209)."""
  return _TreeCost.apply(sequences, adjacency, None)


def compute_soft_cost(sequences, adjacency, cost_matrix=None):
  """``trex.tree.compute_soft_cost`` (tree.py:212-266); ``cost_matrix`` (Q,Q), (Q,) or None."""
  return _TreeCost.apply(sequences, adjacency, cost_matrix)
