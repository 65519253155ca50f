"""TEST INFRASTRUCTURE ONLY — an INDEPENDENT second oracle for the relaxed Sankoff VJP.

The reference has no relaxed Sankoff, so ``oracle/sankoff_oracle.py`` and the CUDA kernels
implement the same hand-derived reverse recurrence.  This file derives nothing: it transcribes the
reference's forward recurrence literally — ``/root/reference/src/trex/sankoff.py:60`` (children),
``:62-79`` (``cost_array = cost_matrix + dp[child]``, reduce over axis 1, sum over the two children),
``:141`` (root self-loop), ``:187`` (reduce the root row, sum over sites) — with ``min`` replaced by
``-T * logsumexp(-x / T)``, in float64 torch, and lets ``torch.autograd`` produce every gradient:
d total / d leaf_cost, d total / d cost_matrix and d total / d dp[node] (the soft ancestral marginals).
PyTorch here is the differentiation engine of the checker, nothing of it is shipped or timed.
"""

from __future__ import annotations

import numpy as np
import torch

SENTINEL = 1e5  # sankoff.py:152


def soft_sankoff_autograd(adj, cost, leaf_cost, n_all, n_leaves, T, g=1.0):
  """adj (n_all, n_all); cost (A, A); leaf_cost (n_leaves, L, A).
  Returns dict(total, dp (L, n_all, A), d_leaf (n_leaves, L, A), d_cost (A, A), gbar (L, n_all, A))
  as float64 NumPy arrays; ``gbar[:, node]`` is d (g * total) / d dp[node] from autograd (``retain_grad``)."""
  adj = np.array(adj, dtype=np.float32, copy=True)
  adj[-1, -1] = 0  # sankoff.py:141
  C = torch.tensor(np.asarray(cost, dtype=np.float64), requires_grad=True)
  leaf = torch.tensor(np.asarray(leaf_cost, dtype=np.float64), requires_grad=True)
  L, A = leaf.shape[1], leaf.shape[2]
  rows = [leaf[i] for i in range(n_leaves)]  # dp rows, (L, A) each
  for r in rows:
    r.retain_grad()
  filler = torch.full((L, A), SENTINEL, dtype=torch.float64)  # a row that was never written (1e5 fill)
  for node in range(n_leaves, n_all):  # fori_loop, sankoff.py:87-92
    idx = np.nonzero(adj[:, node] == 1)[0][:2]  # jnp.where(..., size=2, fill_value=-1), :60
    kids = [int(idx[k]) if k < idx.size else -1 for k in range(2)]
    total = torch.zeros((L, A), dtype=torch.float64)
    for c in kids:  # scan over the two children, :73-77
      if c < 0:
        c += n_all  # -1 wraps to the last row
      d = rows[c] if c < node else filler  # rows >= node still hold the fill
      cost_array = C[None, :, :] + d[:, None, :]  # [site, s, t] = C[s,t] + dp[child,t], :67
      total = total + (-T) * torch.logsumexp(-cost_array / T, dim=2)  # min over axis 1 -> softmin, :68
    total.retain_grad()
    rows.append(total)
  out = ((-T) * torch.logsumexp(-rows[n_all - 1] / T, dim=1)).sum()  # :187
  (g * out).backward()
  zeros = np.zeros((L, A))
  gbar = np.stack([zeros if r.grad is None else r.grad.numpy() for r in rows], axis=1)
  return {
    "total": float(out.detach()),
    "dp": np.stack([r.detach().numpy() for r in rows], axis=1),
    "d_leaf": np.zeros_like(leaf_cost, dtype=np.float64) if leaf.grad is None else leaf.grad.numpy(),
    "d_cost": C.grad.numpy(),
    "gbar": gbar,
  }
