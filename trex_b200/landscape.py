"""Host-side mirror of the landscape-aware loss of the reference's benchmark harness
(``/root/reference/src/trex/evals/benchmark.py``: ``compute_parental_logits`` :586-663,
``_compute_loss_landscape_aware_stacked`` :235-306) on top of ``trex_landscape_loss_fwd_bwd``
(include/trex_b200.h) — SURVEY §8(f) rank 2.  Same names, argument order and return values; arrays are
torch CUDA tensors, ``landscape`` is the reference's dict with ``interactions`` (L, k) and
``fitness_tables`` (L, Q**(k+1)).  ``batch_size`` (the reference's ``safe_map`` chunk, an XLA memory
workaround) is accepted and ignored: one CTA per site needs no chunking.  No CPU fallback.
"""

from __future__ import annotations

import torch

from . import _lib
from . import tree as _tree
from ._lib import call, ptr, stream_ptr


def _landscape_arrays(landscape, real_k, device):
  tables = torch.as_tensor(landscape["fitness_tables"]).to(device=device, dtype=torch.float32).contiguous()
  inter = torch.as_tensor(landscape["interactions"]).to(device=device, dtype=torch.int32)
  k = 0 if real_k == 0 else int(inter.shape[1])  # k_eff = interactions.shape[1] (:621)
  inter = inter[:, :k].contiguous() if k else None
  return inter, tables, k


def _call(S, adj, parent_index, inter, tables, mask, k, want_cost, want_logits, want_grad):
  _lib.load()
  if not S.is_cuda:
    raise RuntimeError("trex_b200 needs CUDA tensors; there is no CPU fallback")
  N, L, Q = S.shape
  assert tables.shape == (L, Q ** (k + 1)), (tables.shape, (L, Q ** (k + 1)))
  dev = S.device
  cost = torch.empty((), dtype=torch.float32, device=dev) if want_cost else None
  logits = torch.empty_like(S) if want_logits else None
  dS = torch.empty_like(S) if want_grad else None
  parent = torch.empty(N, dtype=torch.int32, device=dev)
  acc = torch.empty(2, dtype=torch.float64, device=dev)
  call("trex_landscape_loss_fwd_bwd", ptr(S), ptr(adj), ptr(parent_index), ptr(inter), ptr(tables), ptr(mask), N, L, Q,
       k, ptr(cost), ptr(logits), ptr(dS), ptr(parent), ptr(acc), stream_ptr())
  return cost, logits, dS


def compute_parental_logits(parent_sequences, landscape, real_k, batch_size=64):
  """evals/benchmark.py:586-663 → (n_parents, L, Q) expected fitness of each state of each site."""
  del batch_size
  P = parent_sequences.to(torch.float32).contiguous()
  inter, tables, k = _landscape_arrays(landscape, real_k, P.device)
  ident = torch.arange(P.shape[0], dtype=torch.int32, device=P.device)
  return _call(P, None, ident, inter, tables, None, k, False, True, False)[1]


class _FitnessCost(torch.autograd.Function):

  @staticmethod
  def forward(ctx, sequences, adjacency, inter, tables, mask, k):
    S = sequences.to(torch.float32).contiguous()
    adj = adjacency.to(device=S.device, dtype=torch.float32).contiguous()
    cost, _, dS = _call(S, adj, None, inter, tables, mask, k, True, False, ctx.needs_input_grad[0])
    ctx.save_for_backward(dS if dS is not None else torch.empty(0, device=S.device))
    return cost

  @staticmethod
  def backward(ctx, g):
    (dS,) = ctx.saved_tensors
    return (g * dS if ctx.needs_input_grad[0] else None), None, None, None, None, None


def fitness_cost(sequences, adj_matrix, landscape, real_k, seq_mask=None):
  """The parental-guidance term of evals/benchmark.py:283-303 (before the lambda weight); differentiable
  w.r.t. ``sequences`` (child path and parent path)."""
  inter, tables, k = _landscape_arrays(landscape, real_k, sequences.device)
  mask = None if seq_mask is None else torch.as_tensor(seq_mask).to(device=sequences.device, dtype=torch.float32).contiguous()
  return _FitnessCost.apply(sequences, adj_matrix, inter, tables, mask, k)


def _compute_loss_landscape_aware_stacked(ancestors, masked_sequences, n_leaves, landscape, adj_matrix, n_all,
                                          lambda_val, real_k, temperature=1.0, seq_mask=None, batch_size=64):
  """evals/benchmark.py:235-306: ``compute_surrogate_cost + lambda_val * fitness_cost`` of the stacked
  ancestor logits; differentiable w.r.t. ``ancestors`` (torch autograd in the role of ``jax.grad``)."""
  del batch_size
  assert masked_sequences.shape[0] == n_all and n_leaves == (n_all + 1) // 2, "binary tree expected (tree.py:127-128)"
  S = _tree.update_seq(ancestors, masked_sequences, temperature)
  loss = _tree.compute_surrogate_cost(S, adj_matrix)
  if lambda_val > 0.0 and real_k > 0:  # :282, Python-level like the reference
    loss = loss + lambda_val * fitness_cost(S, adj_matrix, landscape, real_k, seq_mask)
  return loss
