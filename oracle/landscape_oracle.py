"""TEST INFRASTRUCTURE ONLY — NumPy restatement of the landscape-aware loss term of
``/root/reference/src/trex/evals/benchmark.py``: ``compute_parental_logits`` (:586-663) and
``_compute_loss_landscape_aware_stacked`` (:235-306).  Pinned by golden vectors of the UNMODIFIED
reference functions run under the NumPy stand-in for jax (tests/golden/make_golden_landscape.py →
tests/golden/landscape_golden.npz); the gradient is the closed form of the restated forward, checked
by central differences in tests/test_landscape_oracle.py."""

from __future__ import annotations

import numpy as np

from . import tree_oracle


def compute_parental_logits(parent_sequences, interactions, fitness_tables, real_k, dtype=np.float64):
  """evals/benchmark.py:586-663 → (n_parents, L, Q)."""
  P = np.asarray(parent_sequences, dtype=dtype)
  n_parents, L, Q = P.shape
  F = np.asarray(fitness_tables, dtype=dtype)
  if real_k == 0:  # :613-618
    return np.broadcast_to(F[None, :, :], (n_parents, L, Q)).copy()
  inter = np.asarray(interactions)
  k_eff = inter.shape[1]  # :621
  out = np.zeros((L, n_parents, Q), dtype=dtype)
  for i in range(L):
    nb = inter[i, :k_eff]  # :628
    probs = P[:, nb, :]  # :632
    joint = probs[:, 0, :]  # :636
    for j in range(1, k_eff):  # :637-641
      joint = np.einsum("pc,ps->pcs", joint, probs[:, j, :]).reshape(n_parents, -1)
    table = F[i].reshape(Q, -1)  # :646
    out[i] = np.einsum("si,pi->ps", table, joint)  # :652
  return np.transpose(out, (1, 0, 2))  # :663


def log_softmax(x):
  m = x.max(axis=-1, keepdims=True)
  return x - m - np.log(np.exp(x - m).sum(axis=-1, keepdims=True))


def fitness_cost(sequences, adjacency, interactions, fitness_tables, real_k, seq_mask=None, dtype=np.float64):
  """The second term of :283-303 (before the lambda weight)."""
  S = np.asarray(sequences, dtype=dtype)
  n_all, L, _ = S.shape
  mask = np.ones(L, dtype=dtype) if seq_mask is None else np.asarray(seq_mask).astype(dtype)
  parent = np.argmax(np.asarray(adjacency), axis=1)  # :284
  logits = compute_parental_logits(S[parent], interactions, fitness_tables, real_k, dtype=dtype)  # :285-290
  logp = log_softmax(logits)  # :292
  ce = -(S * logp).sum(axis=-1) * mask[None, :]  # :295-297
  is_root = np.arange(n_all) == parent  # :300
  return ce.sum() / ((~is_root).sum() * mask.sum())  # :301


def loss_landscape_aware_stacked(ancestors, masked_sequences, n_leaves, interactions, fitness_tables, adj_matrix,
                                 lambda_val, real_k, temperature=1.0, seq_mask=None, dtype=np.float64):
  """_compute_loss_landscape_aware_stacked (:235-306)."""
  S = np.array(masked_sequences, dtype=dtype)
  S[n_leaves:] = tree_oracle.softmax(np.asarray(ancestors, dtype=dtype) * temperature, axis=-1)  # :270, :231
  cost = tree_oracle.compute_surrogate_cost(S, adj_matrix, dtype=dtype)  # :275
  if lambda_val > 0.0 and real_k > 0:  # :282
    return cost + lambda_val * fitness_cost(S, adj_matrix, interactions, fitness_tables, real_k, seq_mask, dtype)
  return cost


def fitness_cost_grad(sequences, adjacency, interactions, fitness_tables, real_k, seq_mask=None):
  """d fitness_cost / d sequences (float64), child path and parent path."""
  S = np.asarray(sequences, dtype=np.float64)
  n_all, L, Q = S.shape
  mask = np.ones(L) if seq_mask is None else np.asarray(seq_mask).astype(np.float64)
  parent = np.argmax(np.asarray(adjacency), axis=1)
  F = np.asarray(fitness_tables, dtype=np.float64)
  logits = compute_parental_logits(S[parent], interactions, F, real_k)
  logp = log_softmax(logits)
  w = mask / (((np.arange(n_all) != parent).sum()) * mask.sum())
  dS = -logp * w[None, :, None]
  if real_k == 0:
    return dS
  dlogits = (np.exp(logp) * S.sum(axis=-1, keepdims=True) - S) * w[None, :, None]
  inter = np.asarray(interactions)
  k = inter.shape[1]
  for i in range(L):
    nb = inter[i]
    table = F[i].reshape((Q,) + (Q,) * k)
    probs = [S[parent][:, nb[m], :] for m in range(k)]  # (n_all, Q) each
    # djoint[n, t0..tk-1] = sum_s dlogits[n, i, s] table[s, t0..]
    dj = np.tensordot(dlogits[:, i, :], table, axes=(1, 0))
    letters = "abcdefgh"[:k]
    for m in range(k):  # g[n, t_m] = sum over the other neighbours' states of dj * their probabilities
      others = [m2 for m2 in range(k) if m2 != m]
      expr = "n" + letters + "".join(",n" + letters[m2] for m2 in others) + "->n" + letters[m]
      g = np.einsum(expr, dj, *[probs[m2] for m2 in others])
      np.add.at(dS, (parent, nb[m]), g)
  return dS
