"""Multi-GPU plumbing for the site-/tree-sharded hot path.

The reference has no distributed code at all.  Sankoff sites are independent
(``src/trex/sankoff.py:97``) and candidate trees are independent, so ranks own contiguous
shards and the only exchange is ONE packed all-reduce per step of the scalar costs and the
small gradient blocks (dC is A*A, the edge histogram A*A).  ``torch.distributed`` (NCCL on
GPUs, gloo in the CPU tests) is used for exactly that call.
"""

from __future__ import annotations

import torch
import torch.distributed as dist


def shard_range(n_items: int, world: int, rank: int, align: int = 1) -> tuple[int, int]:
  """Contiguous [start, stop) of ``n_items`` for ``rank``; shard starts are multiples of
  ``align`` (16 keeps every rank's native row pitch aligned).  Ragged tails go to the last
  non-empty rank; ranks past the end get an empty range."""
  per = -(-n_items // world)
  per = -(-per // align) * align
  start = min(n_items, rank * per)
  stop = min(n_items, start + per)
  return start, stop


class PackedAllReduce:
  """Packs several small tensors into one float64 buffer, all-reduces it once (sum), unpacks.

  f64 keeps integer-valued sums (hard totals, edge counts < 2^53) exact."""

  def __init__(self, shapes, device):
    self.shapes = [tuple(s) for s in shapes]
    self.sizes = [int(torch.Size(s).numel()) if len(s) else 1 for s in self.shapes]
    self.buf = torch.zeros(sum(self.sizes), dtype=torch.float64, device=device)

  def pack(self, tensors):
    off = 0
    for t, n in zip(tensors, self.sizes):
      self.buf[off:off + n].copy_(t.reshape(-1))
      off += n

  def reduce(self, group=None):
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
      dist.all_reduce(self.buf, op=dist.ReduceOp.SUM, group=group)

  def unpack(self):
    out, off = [], 0
    for s, n in zip(self.shapes, self.sizes):
      out.append(self.buf[off:off + n].view(s))
      off += n
    return out

  def __call__(self, tensors, group=None):
    self.pack(tensors)
    self.reduce(group)
    return self.unpack()


def _world(group=None):
  if dist.is_available() and dist.is_initialized():
    return dist.get_world_size(group), dist.get_rank(group)
  return 1, 0


# ---------------------------------------------------------------------------------------------
# (e2) batched scoring of candidate trees, tree batch sharded over the ranks (BASELINE config c4;
# the vmap-over-trees use of src/trex/evals/benchmark.py:92-93, 216-219 on top of sankoff.py:97)
# ---------------------------------------------------------------------------------------------


def _score_trees_local(adjacencies, cost, leaf_states_u8, n_states):
  """Hard parsimony score of every tree of ``adjacencies`` (B, n_all, n_all) on one alignment: one batched
  plan + one batched forward launch of libtrex_b200 (dp tables of the local trees only)."""
  from .sankoff import SankoffEngine

  B, n_all, _ = adjacencies.shape
  n_leaves, L = leaf_states_u8.shape
  eng = SankoffEngine(n_all, n_leaves, n_states, L, batch=B, device=adjacencies.device)
  eng.set_adjacency(adjacencies).set_leaf_states_u8(leaf_states_u8)
  return eng.hard_forward(cost, want_argmin=False).clone()  # (B,) f64, exact integer sums


def score_trees_sharded(adjacencies, cost, leaf_states_u8, n_states, group=None, local_fn=None):
  """Scores of ALL ``B`` candidate trees on every rank.  ``adjacencies`` (B, n_all, n_all) is the full batch
  (replicated or a view of it — only this rank's slice ``shard_range(B, world, rank)`` is read); the
  alignment and the cost matrix are replicated (640 KB at c4).  No data-path collective: each rank scores
  its trees, one all-gather of the f64 scores (B values) assembles the result.
  ``local_fn(adj_slice, cost, leaf_states, n_states) -> (b,) f64`` defaults to the CUDA path (tests of the
  host logic inject a CPU checker)."""
  world, rank = _world(group)
  B = adjacencies.shape[0]
  lo, hi = shard_range(B, world, rank)
  fn = local_fn or _score_trees_local
  per = -(-B // world)
  mine = torch.zeros(per, dtype=torch.float64, device=adjacencies.device)
  if hi > lo:
    mine[: hi - lo] = fn(adjacencies[lo:hi], cost, leaf_states_u8, n_states).to(torch.float64)
  if world == 1:
    return mine[:B]
  parts = [torch.empty_like(mine) for _ in range(world)]
  dist.all_gather(parts, mine, group=group)
  return torch.cat(parts)[:B]  # rank r holds trees [r * per, ...): the concatenation is in tree order


# ---------------------------------------------------------------------------------------------
# (e3) soft-adjacency tree cost, sites (= the contraction axis K = L*Q) sharded over the ranks
# ---------------------------------------------------------------------------------------------


def _tree_cost_local(S_local, adjacency, cost_matrix):
  from .tree import tree_cost_value_and_grad

  return tree_cost_value_and_grad(S_local, adjacency, cost_matrix)


def tree_cost_value_and_grad_sharded(S_local, adjacency, cost_matrix=None, group=None, local_fn=None):
  """Value and gradients of ``compute_soft_cost`` (src/trex/tree.py:212-266) with the sites sharded:
  every rank holds ``S[:, its L/g block, :]``.  ``cost = <M, G>`` with ``G = S_flat W_flat^T`` summed over
  the sites (tree.py:257-264) and ``M`` a function of the adjacency alone, so the cost, ``dAdj = 1/2 (G_ii +
  G_jj) - G_ij`` and ``dC = sum_l S_l^T M S_l`` are all linear in the per-shard Gram matrices: each rank
  evaluates the library call on its block, ONE packed all-reduce of ``[cost, dAdj (N*N), dC]`` (f64)
  finishes them; ``dS`` needs only M and the local block and stays sharded with its sites.
  Returns ``(cost f32 scalar, dS_local, dAdj (N,N) f32, dC | None)``."""
  fn = local_fn or _tree_cost_local
  cost, dS, dA, dC = fn(S_local, adjacency, cost_matrix)
  shapes = [(1,), tuple(dA.shape)] + ([tuple(dC.shape)] if dC is not None else [])
  red = PackedAllReduce(shapes, device=dA.device)
  outs = red([cost.reshape(1), dA] + ([dC] if dC is not None else []), group=group)
  cost_all = outs[0][0].to(torch.float32)
  dA_all = outs[1].to(torch.float32)
  dC_all = outs[2].to(torch.float32) if dC is not None else None
  return cost_all, dS, dA_all, dC_all
