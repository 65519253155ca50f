"""Synthetic trees and alignments (host-side NumPy; explicit arrays, fixed seeds — SURVEY §8d).

Tree conventions follow the reference's input contract (``Adjacency[i, j] = 1`` iff i is a
child of j, ``src/trex/utils/types.py:33``; children have smaller indices than their parent,
the root is the last node):

* ``balanced_tree``       — ``create_balanced_binary_tree`` (``src/trex/evals/benchmark.py:781-791``)
* ``groundtruth_tree``    — ``generate_groundtruth``'s adjacency (``src/trex/ground_truth.py:180-191``)
* ``random_join_tree``    — random pairwise joins in index order
* ``caterpillar_tree``    — the n-1 level worst case
"""

from __future__ import annotations

import numpy as np


def parents_to_adjacency(parent: np.ndarray) -> np.ndarray:
  n_all = parent.shape[0]
  adj = np.zeros((n_all, n_all), dtype=np.float32)
  idx = np.nonzero(parent >= 0)[0]
  adj[idx, parent[idx]] = 1.0
  return adj


def balanced_parents(n_leaves: int) -> np.ndarray:
  n_anc = n_leaves - 1
  parent = np.full(n_leaves + n_anc, -1, dtype=np.int64)
  parent[:n_leaves] = n_leaves + np.arange(n_leaves) // 2
  parent[n_leaves : n_leaves + n_anc - 1] = n_leaves + (np.arange(n_anc - 1) + n_leaves) // 2
  return parent


def balanced_tree(n_leaves: int) -> np.ndarray:
  return parents_to_adjacency(balanced_parents(n_leaves))


def groundtruth_parents(n_leaves: int) -> np.ndarray:
  n_anc = n_leaves - 1
  parent = np.full(n_leaves + n_anc, -1, dtype=np.int64)
  for i in range(n_anc):
    parent[2 * i] = n_leaves + i
    parent[2 * i + 1] = n_leaves + i
  return parent


def groundtruth_tree(n_leaves: int) -> np.ndarray:
  return parents_to_adjacency(groundtruth_parents(n_leaves))


def random_join_parents(n_leaves: int, rng: np.random.Generator) -> np.ndarray:
  n_all = 2 * n_leaves - 1
  parent = np.full(n_all, -1, dtype=np.int64)
  pool = list(range(n_leaves))
  for new in range(n_leaves, n_all):
    i, j = rng.choice(len(pool), size=2, replace=False)
    a, b = pool[i], pool[j]
    parent[a] = new
    parent[b] = new
    pool = [p for k, p in enumerate(pool) if k not in (i, j)]
    pool.append(new)
  return parent


def random_join_tree(n_leaves: int, rng: np.random.Generator) -> np.ndarray:
  return parents_to_adjacency(random_join_parents(n_leaves, rng))


def caterpillar_parents(n_leaves: int) -> np.ndarray:
  n_all = 2 * n_leaves - 1
  parent = np.full(n_all, -1, dtype=np.int64)
  parent[0] = n_leaves
  parent[1] = n_leaves
  for k in range(2, n_leaves):
    parent[k] = n_leaves + k - 1
    parent[n_leaves + k - 2] = n_leaves + k - 1
  return parent


def caterpillar_tree(n_leaves: int) -> np.ndarray:
  return parents_to_adjacency(caterpillar_parents(n_leaves))


def evolve(parent: np.ndarray, n_sites: int, n_states: int, rate: float,
           rng: np.random.Generator) -> np.ndarray:
  """Edge-mutation generator: uniform root sequence; along each edge every site mutates with
  probability ``rate`` to a different uniform state.  Returns all states (n_all, L) uint8."""
  n_all = parent.shape[0]
  states = np.zeros((n_all, n_sites), dtype=np.uint8)
  roots = np.nonzero(parent < 0)[0]
  for r in roots:
    states[r] = rng.integers(0, n_states, size=n_sites, dtype=np.uint8)
  for v in range(n_all - 1, -1, -1):  # parents have larger indices than children
    p = parent[v]
    if p < 0:
      continue
    mut = rng.random(n_sites) < rate
    off = rng.integers(1, n_states, size=n_sites)
    states[v] = np.where(mut, (states[p].astype(np.int64) + off) % n_states, states[p]).astype(np.uint8)
  return states


def unit_cost(n_states: int) -> np.ndarray:
  return (np.ones((n_states, n_states)) - np.eye(n_states)).astype(np.float32)


def random_int_cost(n_states: int, hi: int, rng: np.random.Generator) -> np.ndarray:
  """Random symmetric integer matrix in {1..hi} with zero diagonal."""
  m = rng.integers(1, hi + 1, size=(n_states, n_states))
  m = np.triu(m, 1)
  m = m + m.T
  return m.astype(np.float32)


def nk_alignment(parent: np.ndarray, n_sites: int, n_states: int, rng: np.random.Generator,
                 k: int = 1, mutation_rate: float = 0.1, coupled_prob: float = 0.5) -> np.ndarray:
  """NK-landscape flavoured generator (NumPy restatement in spirit of
  ``src/trex/nk_model.py:116-278``: per-site fitness tables over the site and its K right
  neighbours, Metropolis-Hastings acceptance of proposed point mutations down the tree).
  Only used to PRODUCE inputs; it is not on the hot path (OUT OF SCOPE as a kernel).
  """
  n_all = parent.shape[0]
  table = rng.random((n_sites, n_states ** (k + 1)), dtype=np.float32)
  nbr = (np.arange(n_sites)[:, None] + np.arange(k + 1)[None, :]) % n_sites
  pw = n_states ** np.arange(k + 1)

  def fitness_sites(seq):
    idx = (seq[nbr].astype(np.int64) * pw).sum(axis=1)
    return table[np.arange(n_sites), idx]

  states = np.zeros((n_all, n_sites), dtype=np.uint8)
  roots = np.nonzero(parent < 0)[0]
  for r in roots:
    states[r] = rng.integers(0, n_states, size=n_sites, dtype=np.uint8)
  for v in range(n_all - 1, -1, -1):
    p = parent[v]
    if p < 0:
      continue
    cur = states[p].copy()
    prop = cur.copy()
    mut = rng.random(n_sites) < mutation_rate
    off = rng.integers(1, n_states, size=n_sites)
    prop[mut] = ((cur[mut].astype(np.int64) + off[mut]) % n_states).astype(np.uint8)
    # site-wise Metropolis-Hastings on the local fitness contribution
    df = fitness_sites(prop) - fitness_sites(cur)
    accept = (df >= 0) | (rng.random(n_sites) < np.exp(np.minimum(df, 0) / max(coupled_prob, 1e-3)))
    states[v] = np.where(mut & accept, prop, cur)
  return states
