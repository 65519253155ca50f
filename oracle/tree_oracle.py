"""TEST INFRASTRUCTURE ONLY — NumPy restatement of ``/root/reference/src/trex/tree.py`` and of
``naive_compute_soft_cost`` (``/root/reference/benchmark_tree_cost.py:10-32``).

``dtype`` selects the arithmetic: float32 mirrors the reference, float64 is the high-precision
yardstick the 1e-5 tolerance is measured against.  Gradients are the closed forms
derived from the cited forward and are themselves checked by central differences
in ``tests/test_oracle.py``.
"""

from __future__ import annotations

import numpy as np


def softmax(x, axis=-1):
  """``jax.nn.softmax``: subtracts the max along ``axis``; ``-inf`` entries give 0."""
  x = np.asarray(x)
  m = np.max(x, axis=axis, keepdims=True)
  m = np.where(np.isfinite(m), m, 0)
  e = np.exp(x - m)
  return e / e.sum(axis=axis, keepdims=True)


def discretize_tree_topology(adjacency, n_nodes):
  """tree.py:31-47 — ``one_hot(argmax(adj, axis=1), n_nodes)``."""
  idx = np.argmax(np.asarray(adjacency), axis=1)
  return np.eye(n_nodes, dtype=np.asarray(adjacency).dtype)[idx]


def update_seq(ancestor_logits, sequences, temperature=1.0):
  """tree.py:110-130 — rows n_leaves: = softmax(logits * temperature) (note: multiply)."""
  sequences = np.array(sequences, copy=True)
  n_total = sequences.shape[0]
  n_leaf = (n_total + 1) // 2  # tree.py:127-128
  logits = np.stack([np.asarray(a) for a in ancestor_logits]) if isinstance(
    ancestor_logits, (list, tuple)) else np.asarray(ancestor_logits)
  sequences[n_leaf:] = softmax(logits * temperature, axis=-1)  # tree.py:129-130
  return sequences


def update_tree(tree_params, gumbel_noise, temperature=1.0, gates=None):
  """tree.py:50-107 with the Gumbel noise passed explicitly (``jax.random`` streams cannot
  be reproduced without JAX)."""
  tree_params = np.asarray(tree_params)
  n_all_minus_1, n_anc = tree_params.shape  # :64
  n_total = n_all_minus_1 + 1
  n_leaves = n_total - n_anc
  if n_anc == 0:
    return np.eye(n_total, dtype=tree_params.dtype)  # :68-69
  p = tree_params + np.asarray(gumbel_noise, dtype=tree_params.dtype)  # :72
  if gates is not None:
    p = p * gates  # :75-76
  p = p / temperature  # :78
  logits = np.full((n_total, n_total), -np.inf, dtype=tree_params.dtype)  # :81
  logits[:n_leaves, n_leaves:] = p[:n_leaves]  # :84-85
  i = np.arange(n_anc - 1)[:, None]
  j = np.arange(n_anc)[None, :]
  logits[n_leaves:-1, n_leaves:] = np.where(j > i, p[n_leaves:], -np.inf)  # :90-102
  logits[-1, -1] = 1.0  # :105
  return softmax(logits, axis=1)  # :107


def enforce_graph_constraints(adjacency, scaling_factor):
  """tree.py:133-160."""
  adjacency = np.asarray(adjacency)
  n_total = adjacency.shape[0]
  n_anc = (n_total - 1) // 2  # :157
  cols = adjacency[:-1, -n_anc:] if n_anc > 0 else adjacency[:-1, :]  # `-0:` is the full slice
  child_sums = cols.sum(axis=0)
  return scaling_factor * ((child_sums - 2) ** 2).sum()


def weighted(sequences, cost_matrix):
  """tree.py:248-253 — W = S | S*c | S@C."""
  if cost_matrix is None:
    return sequences
  cost_matrix = np.asarray(cost_matrix, dtype=sequences.dtype)
  if cost_matrix.ndim == 1:
    return sequences * cost_matrix
  return sequences @ cost_matrix


def compute_soft_cost(sequences, adjacency, cost_matrix=None, dtype=np.float32):
  """tree.py:212-266 (``compute_surrogate_cost``, :163-209, is the ``cost_matrix=None`` case)."""
  S = np.asarray(sequences, dtype=dtype)
  A = np.asarray(adjacency, dtype=dtype)
  W = weighted(S, cost_matrix)
  e = (S * W).sum(axis=(-1, -2))  # :255
  flat_s = S.reshape(S.shape[0], -1)
  flat_w = W.reshape(S.shape[0], -1)
  G = flat_s @ flat_w.T  # :260
  term1 = (A * e[:, None]).sum()
  term2 = (A * e[None, :]).sum()
  term3 = -2 * (A * G).sum()
  return (term1 + term2 + term3) / 2, (term1, term2, term3)


def compute_surrogate_cost(sequences, adjacency, dtype=np.float32):
  return compute_soft_cost(sequences, adjacency, None, dtype)[0]


def naive_compute_soft_cost(sequences, adjacency, cost_matrix=None, dtype=np.float64):
  """benchmark_tree_cost.py:10-32 — direct (N,N,L,Q) form, the independent second oracle."""
  S = np.asarray(sequences, dtype=dtype)
  A = np.asarray(adjacency, dtype=dtype)
  Q = S.shape[-1]
  C = np.eye(Q, dtype=dtype) if cost_matrix is None else np.asarray(cost_matrix, dtype=dtype)
  if C.ndim == 1:
    C = np.diag(C)  # compute_soft_cost's (Q,) form is a diagonal cost (tree.py:250-251)
  total = dtype(0)
  for i in range(S.shape[0]):  # row-blocked so the (N,N,L,Q) tensor is never materialised
    diff = S[i][None] - S  # (N, L, Q)
    wd = diff @ C
    total += ((diff * wd).sum(axis=(-1, -2)) * A[i]).sum()
  return total / 2


def compute_cost(sequences, adjacency, substitution_matrix):
  """tree.py:269-296 — exact discrete cost."""
  seq_int = np.argmax(np.asarray(sequences), axis=2)  # :287
  disc = discretize_tree_topology(adjacency, np.asarray(adjacency).shape[0])  # :288
  parent = np.argmax(disc, axis=1)  # :291
  parent_seqs = seq_int[parent]  # :294
  sub = np.asarray(substitution_matrix)
  return sub[parent_seqs, seq_int][:-1, :].sum()  # :296


def compute_loss(ancestor_logits, tree_params, gumbel_noise, sequences, temperature, adjacency,
                 graph_constraint_scale=10.0, fix_seqs=False, fix_tree=False, dtype=np.float32):
  """tree.py:299-361 (non-verbose).  ``update_tree`` is called with its default temperature
  1.0 (:338) while ``update_seq`` receives ``temperature`` (:336)."""
  S = np.asarray(sequences, dtype=dtype)
  S2 = S if fix_seqs else update_seq(np.asarray(ancestor_logits, dtype=dtype), S, temperature)
  A2 = np.asarray(adjacency, dtype=dtype) if fix_tree else update_tree(
    np.asarray(tree_params, dtype=dtype), gumbel_noise)
  sur = compute_surrogate_cost(S2, A2, dtype)
  con = enforce_graph_constraints(A2, graph_constraint_scale)
  return sur + temperature * con  # :342


# ---------------------------------------------------------------------------------------
# Gradients of compute_soft_cost (closed forms).  M = diag((r+c)/2) - Adj with
# r/c the row/column sums of Adj, so that cost = sum_ij M_ij G_ij, G = S_flat W_flat^T.
# ---------------------------------------------------------------------------------------


def soft_cost_value_and_grad(sequences, adjacency, cost_matrix=None, dtype=np.float64):
  """Returns cost, dS (N,L,Q), dAdj (N,N), dC (Q,Q | Q, | None)."""
  S = np.asarray(sequences, dtype=dtype)
  A = np.asarray(adjacency, dtype=dtype)
  N = S.shape[0]
  W = weighted(S, cost_matrix)
  flat_s = S.reshape(N, -1)
  G = flat_s @ W.reshape(N, -1).T
  r = A.sum(axis=1)
  c = A.sum(axis=0)
  M = np.diag((r + c) / 2) - A
  cost = (M * G).sum()
  d = np.diag(G)
  dA = (d[:, None] + d[None, :]) / 2 - G
  MS = np.tensordot(M, S, axes=(1, 0))  # sum_j M_ij S_j
  MtS = np.tensordot(M.T, S, axes=(1, 0))
  if cost_matrix is None:
    dS = MS + MtS
    dC = None
  else:
    Cm = np.asarray(cost_matrix, dtype=dtype)
    if Cm.ndim == 1:
      dS = (MS + MtS) * Cm
      dC = np.einsum("ilq,ij,jlq->q", S, M, S)
    else:
      # G_ij = sum_l S_il^T C^T S_jl  (W = S @ C, tree.py:253)
      dS = MS @ Cm + MtS @ Cm.T
      dC = np.einsum("jlq,ij,ilp->qp", S, M, S)
  return cost, dS, dA, dC
