"""TEST INFRASTRUCTURE ONLY — NumPy restatement of the reference's input generators
(``/root/reference/src/trex/ground_truth.py``, ``/root/reference/src/trex/nk_model.py``), line by line, with every
``jax.random`` call replaced by an EXPLICIT draw array: threefry streams are not reproducible outside JAX (SURVEY
8c), so parity is defined on the draws — the same arrays go to the CUDA generators and must give identical results.

Draw conventions (one array per ``jax.random`` call site of the reference):
* ``mutate``: ``mut_sites`` = the ``random.choice(key, L, (n_mutations,), replace=False)`` result (ground_truth.py:40-45),
  ``offsets`` = the ``random.randint(key, shape, 1, n_states)`` result (:47);
* ``generate_tree_data``, per (node, step): ``noise`` (:199), ``u_coupled`` — the uniform behind
  ``bernoulli(sk, coupled_mutation_prob)`` (:230; jax's bernoulli is ``uniform < p``), ``site`` (:221), ``u_mask`` /
  ``new_vals`` (:214-216, and :225 for the coupled proposal), ``u_accept`` (:243-244).
"""

from __future__ import annotations

import numpy as np


# ---- ground_truth.py ---------------------------------------------------------------------------------------------


def mutate(sequence, n_states, mut_sites, offsets):
  """``mutate`` (ground_truth.py:20-53) for one sequence (L,)."""
  sequence = np.asarray(sequence).astype(np.int64)
  L = sequence.shape[0]
  mask = np.zeros(L, dtype=bool)  # :38
  for s in np.asarray(mut_sites).astype(np.int64):  # mutation_mask.at[mutations].set(True), :46
    if s < 0:
      s += L
    if 0 <= s < L:
      mask[s] = True
  proposed = (sequence + np.asarray(offsets).astype(np.int64)) % n_states  # :48
  return np.where(mask, proposed, sequence).astype(np.int8)  # :49-50


def generate_groundtruth(n_leaves, n_states, seq_length, mut_sites, offsets):
  """``generate_groundtruth`` (ground_truth.py:113-197).  ``mut_sites`` (n_all-1, n_mutations) and ``offsets``
  (n_all-1, L) are indexed by the CHILD node the ``mutate`` call produces.  Returns
  ``(masked_sequences (n_all,L) f32, all_sequences (n_all,L) f32, adjacency (n_all,n_all) f32)``."""
  if not (n_leaves > 0 and (n_leaves & (n_leaves - 1) == 0)):  # :142-144
    raise ValueError("n_leaves must be a power of 2.")
  n_ancestors = n_leaves - 1
  n_all = n_leaves + n_ancestors
  all_seqs = np.zeros((n_all, seq_length), dtype=np.int8)  # :150-151 (root sequence = zeros, :146)
  for i in range(n_ancestors):  # fori_loop, :155-173
    parent_idx = n_all - 1 - i
    p_i = parent_idx - n_leaves
    for c in (2 * p_i, 2 * p_i + 1):  # :160-170
      all_seqs[c] = mutate(all_seqs[parent_idx], n_states, mut_sites[c], offsets[c])
  masked = np.zeros_like(all_seqs)
  masked[:n_leaves] = all_seqs[:n_leaves]  # :175-177
  tree = np.zeros((n_all, n_all), dtype=np.float32)
  for i in range(n_ancestors):  # :181-191
    tree[2 * i, n_leaves + i] = 1
    tree[2 * i + 1, n_leaves + i] = 1
  return masked.astype(np.float32), all_seqs.astype(np.float32), tree


def create_balanced_binary_tree(n_leaves):
  """``create_balanced_binary_tree`` (evals/benchmark.py:781-791)."""
  n_anc = n_leaves - 1
  n_total = n_leaves + n_anc
  adj = np.zeros((n_total, n_total), dtype=np.float32)
  adj[np.arange(n_leaves), n_leaves + np.arange(n_leaves) // 2] = 1
  adj[n_leaves + np.arange(n_anc - 1), n_leaves + (np.arange(n_anc - 1) + n_leaves) // 2] = 1
  return adj


# ---- nk_model.py ---------------------------------------------------------------------------------------------------


def get_fitness(sequence, interactions, fitness_tables, n_states, seq_mask=None):
  """``get_fitness`` (nk_model.py:46-110), float32 like the reference's default dtype."""
  sequence = np.asarray(sequence).astype(np.int64).reshape(-1)
  n = sequence.shape[0]
  if seq_mask is None:
    seq_mask = np.ones(n, dtype=bool)  # :91-92
  gather = np.concatenate([np.arange(n)[:, None], np.asarray(interactions)], axis=1)  # :94-95
  states = sequence[gather]  # :97
  k = np.asarray(interactions).shape[1]
  powers = n_states ** np.arange(k + 1)  # :100
  idx = states @ powers  # :102
  vals = np.asarray(fitness_tables, dtype=np.float32)[np.arange(n), idx]  # :104
  m = np.asarray(seq_mask, dtype=np.float32)
  return np.float32(np.sum(vals * m, dtype=np.float32) / np.sum(m, dtype=np.float32))  # :107


def _bfs_literal(adjacency, root):
  """The BFS of nk_model.py:151-185 EXACTLY as written: ``jnp.where(..., size=n_nodes)`` pads the children list with
  zeros (so node 0 is enqueued again and again until it is visited), ``.at[idx].set`` drops out-of-range writes (so a
  full queue loses real children) — the resulting ``sorted_nodes`` may miss nodes and keep -1 entries."""
  n = adjacency.shape[0]
  queue = np.full(n, -1)
  queue[0] = root
  visited = np.zeros(n, dtype=bool)
  sorted_nodes = np.full(n, -1)
  while np.any(queue != -1):  # :152-154
    cur = int(queue[0])  # :158
    queue[0] = -1
    queue = np.roll(queue, -1)  # :159-160
    pos = int(visited.sum())
    if pos < n:
      sorted_nodes[pos] = cur  # :162
    visited[cur] = True  # :163
    kids = np.nonzero(adjacency[:, cur] == 1)[0]
    children = np.zeros(n, dtype=np.int64)  # size=n_nodes, fill value 0, :165
    children[: kids.size] = kids
    for child in children:  # fori_loop(0, n_nodes, add_child_to_queue), :167-176
      if not visited[child]:
        idx = int(np.sum(queue != -1))
        if idx < n:
          queue[idx] = child
  return sorted_nodes


def _bfs_intended(parent, roots):
  """Nodes by (depth, index): every node after its parent — what the BFS is for."""
  n = parent.shape[0]
  depth = np.full(n, n)
  for v in range(n):
    d, u = 0, v
    while parent[u] != u and d <= n:
      u = parent[u]
      d += 1
    depth[v] = d if d <= n else n
  return np.array(sorted(range(n), key=lambda v: (depth[v], v)))


def generate_tree_data(interactions, fitness_tables, adjacency, root_sequence, mutation_rate, draw,
                       coupled_mutation_prob=0.5, n_states=20, mutation_rate_noise_std=0.0, branch_length=1,
                       literal_bfs=False):
  """``generate_tree_data`` (nk_model.py:116-278).  ``draw(kind, node, step)`` supplies the random draws
  (kind in noise / u_coupled / site / u_mask / new_vals / u_accept); it is called in the reference's call order.

  ``literal_bfs=True``: the reference exactly as written — root = first node that is its own argmax parent, else the
  ``size=1`` fill value 0 (:148-149); the BFS of ``_bfs_literal``; ``sorted_nodes`` entries of -1 index the last node
  as jnp does.  This mode is pinned against the unmodified reference (tests/golden/generators_golden.npz).
  ``literal_bfs=False``: the intended semantics the CUDA generator implements — a node whose adjacency row has no
  positive entry is a root as well, and every node is visited once, after its parent.  The per-node evolution
  (rate noise, coupled / random proposal, Metropolis-Hastings acceptance) is the same code in both modes."""
  adjacency = np.asarray(adjacency)
  n_nodes = adjacency.shape[0]
  root_sequence = np.asarray(root_sequence).reshape(-1)
  L = root_sequence.shape[0]
  interactions = np.asarray(interactions)
  parent = adjacency.argmax(axis=1)  # :148
  if literal_bfs:
    hit = np.nonzero(parent == np.arange(n_nodes))[0]
    root = int(hit[0]) if hit.size else 0  # :149
    sorted_nodes = _bfs_literal(adjacency, root)
    roots = {root}
  else:
    parent = np.where(adjacency.max(axis=1) > 0, parent, np.arange(n_nodes))
    roots = set(np.nonzero(parent == np.arange(n_nodes))[0].tolist())
    sorted_nodes = _bfs_intended(parent, roots)
  seqs = np.zeros((n_nodes, L), dtype=np.int64)  # :187
  for r in roots:
    seqs[r] = root_sequence  # :188
  for i in range(n_nodes):  # fori_loop over sorted_nodes, :190-268
    node = int(sorted_nodes[i])
    if node in roots:  # node_index != root_node, :262 (-1 is not the root index: it evolves, as in the reference)
      continue
    node_w = node % n_nodes  # negative indices wrap
    seq = seqs[int(parent[node_w])].copy()
    noise = np.float32(draw("noise", node_w, 0)) * np.float32(mutation_rate_noise_std)  # :199
    rate = np.minimum(np.float32(mutation_rate) * np.exp(noise), np.float32(1.0))  # :203-204
    for step in range(branch_length):  # :248
      if draw("u_coupled", node_w, step) < coupled_mutation_prob:  # :229-235
        site = int(draw("site", node_w, step))  # :221
        sites = np.append([site], interactions[site])  # :222-223
        mask = np.zeros(L, dtype=bool)
        mask[sites] = True
        mutated = np.where(mask, np.asarray(draw("new_vals", node_w, step)), seq)  # :225-226
      else:
        mask = np.asarray(draw("u_mask", node_w, step)) < rate  # :214
        mutated = np.where(mask, np.asarray(draw("new_vals", node_w, step)), seq)  # :216-217
      f0 = get_fitness(seq, interactions, fitness_tables, n_states)  # :239-240
      f1 = get_fitness(mutated, interactions, fitness_tables, n_states)
      accept_p = np.minimum(np.float32(1.0), np.exp(np.float32(f1 - f0)))  # :241
      if draw("u_accept", node_w, step) < accept_p:  # :243-246
        seq = mutated
    seqs[node_w] = seq  # :268
  return seqs.astype(np.int32)


def array_draws(draws):
  """``draw`` callable over per-node arrays: noise [N], u_coupled / site / u_accept [N][B], u_mask / new_vals [N][B][L]."""
  def draw(kind, node, step):
    a = draws[kind]
    return a[node] if kind == "noise" else a[node, step]
  return draw
