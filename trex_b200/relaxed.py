"""Relaxed (softmin) Sankoff with reverse-mode gradients — SURVEY §8a rows R1/R2.

The reference has no relaxed Sankoff; this is the differentiable form named by
the north star: the recurrence of ``src/trex/sankoff.py:62-79,187`` with ``min`` replaced by
``-T*logsumexp(-x/T)``.  ``soft_sankoff_cost`` is the ``jax.custom_vjp`` analogue for torch:
the forward kernel keeps its dp table as the residual, the backward kernel produces the
cotangents of the leaf cost table and of the substitution-cost matrix in one tree walk.
"""

from __future__ import annotations

import torch

from .sankoff import SankoffEngine, pad_sites


def to_native_leaf_cost(leaf_cost_nla: torch.Tensor) -> torch.Tensor:
  """(n_leaves, L, A) reference-style table -> native [n_leaves][A][stride] f32 (zero pad)."""
  n, L, A = leaf_cost_nla.shape
  out = torch.zeros((n, A, pad_sites(L)), dtype=torch.float32, device=leaf_cost_nla.device)
  out[:, :, :L].copy_(leaf_cost_nla.permute(0, 2, 1))
  return out


class _SoftSankoff(torch.autograd.Function):

  @staticmethod
  def forward(ctx, cost, leaf_cost_native, eng, temperature, leaf_is_states):
    ctx.eng, ctx.T, ctx.leaf_is_states = eng, float(temperature), leaf_is_states
    lc = None if leaf_is_states else leaf_cost_native.contiguous()
    total = eng.soft_forward(cost, temperature, leaf_cost=lc)
    ctx.save_for_backward(cost, lc if lc is not None else torch.empty(0, device=cost.device))
    return total.to(torch.float32).sum() if eng.batch == 1 else total.to(torch.float32)

  @staticmethod
  def backward(ctx, g):
    cost, lc = ctx.saved_tensors
    eng = ctx.eng
    lc = None if ctx.leaf_is_states else lc
    if eng.batch != 1:
      raise NotImplementedError("per-tree cotangents: call SankoffEngine.soft_backward per tree")
    gs = g.detach().to(torch.float32).reshape(1).contiguous()
    eng.soft_backward(cost, ctx.T, leaf_cost=lc, gscale=gs, want_dleaf=lc is not None,
                      want_dcost=True)
    dcost = eng.dcost[0].to(torch.float32)
    dleaf = eng.dleaf[0].clone() if lc is not None else None
    return dcost, dleaf, None, None, None


def soft_sankoff_cost(adjacency_matrix, cost_matrix, leaves, n_all, n_states, n_leaves,
                      temperature, engine: SankoffEngine | None = None):
  """Relaxed parsimony score, differentiable w.r.t. ``cost_matrix`` and (soft) ``leaves``.

  ``leaves`` is either an integer/float (n_leaves, L) array of states (hard leaves, same
  encoding rules as ``run_sankoff``) or a float32 native leaf-cost table
  ``[n_leaves][A][pad_sites(L)]`` (see ``to_native_leaf_cost``).
  """
  leaves = torch.as_tensor(leaves)
  is_states = leaves.dim() == 2
  L = leaves.shape[1] if is_states else None
  if engine is None:
    if not is_states:
      raise ValueError("pass engine=SankoffEngine(...) with soft leaves (L is not recoverable from the padded table)")
    engine = SankoffEngine(n_all, n_leaves, n_states, L)
  engine.set_adjacency(torch.as_tensor(adjacency_matrix), zero_root_selfloop=True)
  if is_states:
    engine.set_sequences_f32(leaves)
    dummy = torch.empty(0, device=engine.device)
    return _SoftSankoff.apply(torch.as_tensor(cost_matrix).to(engine.device, torch.float32), dummy, engine,
                              temperature, True)
  return _SoftSankoff.apply(torch.as_tensor(cost_matrix).to(engine.device, torch.float32),
                            leaves.to(engine.device, torch.float32), engine, temperature, False)
