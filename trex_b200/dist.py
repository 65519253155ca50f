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

  def __call__(self, tensors, group=None):
    off = 0
    for t, n in zip(tensors, self.sizes):
      self.buf[off:off + n].copy_(t.reshape(-1))
      off += n
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
      dist.all_reduce(self.buf, op=dist.ReduceOp.SUM, group=group)
    out, off = [], 0
    for s, n in zip(self.shapes, self.sizes):
      out.append(self.buf[off:off + n].view(s))
      off += n
    return out
