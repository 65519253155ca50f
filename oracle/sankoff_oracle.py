"""TEST INFRASTRUCTURE ONLY — NumPy restatement of ``/root/reference/src/trex/sankoff.py``.

Every function cites the reference lines it follows.  Arithmetic is float32 exactly as the
reference (``get_default_dtype()`` is float32, ``types.py:12``), the table sentinel is
``1e5`` (``sankoff.py:152``), ``argmin`` returns the FIRST minimum (jnp semantics ==
``np.argmin``), children of a node are the first two rows with ``adj[:, node] == 1`` in
ascending order padded with ``-1`` (``sankoff.py:60``) and ``-1`` wraps to the last row.
"""

from __future__ import annotations

import numpy as np

SENTINEL = np.float32(1e5)  # sankoff.py:152


def children_of(adj: np.ndarray, node: int) -> np.ndarray:
  """``jnp.where(adj[:, node] == 1, size=2, fill_value=-1)[0]`` — sankoff.py:60."""
  idx = np.nonzero(adj[:, node] == 1)[0][:2]
  out = np.full(2, -1, dtype=np.int64)
  out[: idx.size] = idx
  return out


def init_leaf_rows(dp: np.ndarray, seq: np.ndarray, n_leaves: int) -> np.ndarray:
  """``dp.at[i, int(seq[i])].set(0)`` for i < n_leaves — sankoff.py:49-52.

  JAX scatter semantics: negative indices wrap, out-of-bounds updates are dropped.
  ``dp`` is (n_all, A) for one site.
  """
  n_states = dp.shape[1]
  for i in range(n_leaves):
    s = int(np.asarray(seq[i]).astype(np.int32))
    if s < 0:
      s += n_states
    if 0 <= s < n_states:
      dp[i, s] = 0
  return dp


def run_dp(adj, dp, bt, seq, cost):
  """One site of the Sankoff postorder DP — sankoff.py:24-94 (line-by-line)."""
  adj = np.asarray(adj)
  dp = np.array(dp, dtype=np.float32, copy=True)
  bt = np.array(bt, dtype=np.float32, copy=True)
  cost = np.asarray(cost, dtype=np.float32)
  n_all = adj.shape[0]
  n_leaves = (n_all + 1) // 2  # sankoff.py:45-46
  dp = init_leaf_rows(dp, seq, n_leaves)
  for node in range(n_leaves, n_all):  # fori_loop, sankoff.py:87-92
    ch = children_of(adj, node)
    total = np.zeros_like(dp[0])  # scan carry, sankoff.py:75
    chars = []
    for k in range(2):  # scan over child_idx, sankoff.py:73-77
      c = int(ch[k])
      cost_array = cost + dp[c, :]  # (A,A): [s,t] = C[s,t] + dp[child,t], sankoff.py:67
      total = (total + cost_array.min(axis=1)).astype(np.float32)  # :68,:70
      chars.append(cost_array.argmin(axis=1))  # :69
    dp[node, :] = total  # :79
    bt[node, :, 0] = ch[0]  # :80
    bt[node, :, 1] = chars[0]  # :81
    bt[node, :, 2] = ch[1]  # :82
    bt[node, :, 3] = chars[1]  # :83
  return dp, bt


def backtrack(root_node, root_state, bt, n_all, n_leaves):
  """Explicit-stack DFS traceback for one site — sankoff.py:191-267 (line-by-line)."""
  stack = np.zeros((n_all, 2), dtype=np.int64)
  stack[0] = (root_node, root_state)  # :213
  ptr = 1
  recon = np.zeros(n_all, dtype=np.int32)  # :215
  while ptr > 0:  # :220-223
    ptr -= 1
    node, state = (int(x) for x in stack[ptr])  # :232-233
    if node >= n_leaves:  # :257-262
      recon[node] = state  # :237
      info = bt[node, state]  # :240
      stack[ptr] = (int(info[0]), int(info[1]))  # :245
      stack[ptr + 1] = (int(info[2]), int(info[3]))  # :246
      ptr += 2  # :249
  return recon


def run_sankoff(adj, cost, seqs, n_all, n_states, n_leaves, *, return_path=False):
  """``run_sankoff`` vectorised over sites — sankoff.py:114-188.

  Returns ``(recon (n_all,L) f32, dp (L,n_all,A) f32, total f32)`` like the reference.  The
  per-node body is ``run_dp`` with the site axis carried along (``vectorized_dp``, :97).
  """
  out = run_sankoff_tables(adj, cost, seqs, n_all, n_states, n_leaves)
  dp, arg = out["dp"], out["argmin"]
  L = dp.shape[0]
  recon = np.zeros((n_all, L), dtype=np.float32)  # :161
  recon[:n_leaves] = np.asarray(seqs, dtype=np.float32)[:n_leaves]  # :162
  if return_path:
    states = traceback_topdown(out, n_all, n_leaves)
    recon[n_leaves:] = states[n_leaves:].astype(np.float32)  # :182-185
  total = dp[:, -1].min(axis=1).sum(dtype=np.float32)  # :187
  return recon, dp, np.float32(total)


def run_sankoff_tables(adj, cost, seqs, n_all, n_states, n_leaves):
  """Forward tables of ``run_sankoff``: dp (L,n_all,A) f32, argmin (L,n_all,2,A) i32,
  children (n_all,2) — sankoff.py:141-160 with ``run_dp`` vmapped over sites."""
  adj = np.array(adj, dtype=np.float32, copy=True)
  adj[-1, -1] = 0  # :141
  cost = np.asarray(cost, dtype=np.float32)  # :146
  seqs = np.asarray(seqs, dtype=np.float32)  # :145
  L = seqs.shape[1]
  dp = np.full((L, n_all, n_states), SENTINEL, dtype=np.float32)  # :152
  # leaf init: run_dp derives its own n_leaves from the adjacency shape (:45-46)
  dp_leaves = (adj.shape[0] + 1) // 2
  st = seqs[:dp_leaves].astype(np.int32)  # (n, L)
  st = np.where(st < 0, st + n_states, st)
  ok = (st >= 0) & (st < n_states)
  li, si = np.nonzero(ok)
  dp[si, li, st[li, si]] = 0  # :49-52
  arg = np.zeros((L, n_all, 2, n_states), dtype=np.int32)
  kids = np.full((n_all, 2), -1, dtype=np.int64)
  for node in range(dp_leaves, n_all):
    ch = children_of(adj, node)
    kids[node] = ch
    total = np.zeros((L, n_states), dtype=np.float32)
    for k in range(2):
      c = int(ch[k])
      ca = cost[None, :, :] + dp[:, c, None, :]  # (L, s, t)
      total = (total + ca.min(axis=2)).astype(np.float32)
      arg[:, node, k, :] = ca.argmin(axis=2)
    dp[:, node, :] = total
  return {"dp": dp, "argmin": arg, "children": kids}


def traceback_topdown(tables, n_all, n_leaves):
  """Top-down pointer chase, equivalent to the stack DFS of sankoff.py:212-265 on trees.

  Returns int32 states (n_all, L); rows of nodes not reachable from the root stay 0, leaf
  rows stay 0 (the reference overwrites only rows n_leaves: with them, :182).
  """
  dp, arg, kids = tables["dp"], tables["argmin"], tables["children"]
  L = dp.shape[0]
  states = np.zeros((n_all, L), dtype=np.int32)
  root = n_all - 1
  states[root] = dp[:, root, :].argmin(axis=1)  # :172
  visited = np.zeros(n_all, dtype=bool)
  visited[root] = True
  site = np.arange(L)
  for node in range(n_all - 1, n_leaves - 1, -1):
    if not visited[node]:
      continue
    for k in range(2):
      c = int(kids[node, k])
      if c >= n_leaves and c < node:
        states[c] = arg[site, node, k, states[node]]
        visited[c] = True
  return states


def per_site_root_min(dp: np.ndarray) -> np.ndarray:
  """``dp[:, -1].min(axis=1)`` — the exact per-site score vector of sankoff.py:187."""
  return dp[:, -1].min(axis=1)


# ---------------------------------------------------------------------------------------
# Independent cross-checks (not restatements): brute force and Fitch.
# ---------------------------------------------------------------------------------------


def brute_force_score(adj, cost, leaf_states, n_all, n_states, n_leaves):
  """Min over all A^(n_internal) ancestor assignments of the sum of edge costs, per site."""
  import itertools

  adj = np.asarray(adj)
  cost = np.asarray(cost, dtype=np.float64)
  parent = np.full(n_all, -1)
  for node in range(n_leaves, n_all):
    for c in children_of(adj, node):
      if c >= 0:
        parent[c] = node
  leaf_states = np.asarray(leaf_states).astype(np.int64)
  L = leaf_states.shape[1]
  best = np.full(L, np.inf)
  n_int = n_all - n_leaves
  for assign in itertools.product(range(n_states), repeat=n_int):
    tot = np.zeros(L)
    for v in range(n_all - 1):
      p = parent[v]
      if p < 0:
        continue
      ps = assign[p - n_leaves]
      cs = leaf_states[v] if v < n_leaves else assign[v - n_leaves]
      tot = tot + cost[ps, cs]
    best = np.minimum(best, tot)
  return best


def fitch_score(adj, leaf_states, n_all, n_states, n_leaves):
  """Fitch small-parsimony count (valid for unit cost C = 1 - I), per site."""
  adj = np.asarray(adj)
  leaf_states = np.asarray(leaf_states).astype(np.int64)
  L = leaf_states.shape[1]
  sets = np.zeros((n_all, L, n_states), dtype=bool)
  sets[np.arange(n_leaves)[:, None], np.arange(L)[None, :], leaf_states] = True
  score = np.zeros(L, dtype=np.int64)
  for node in range(n_leaves, n_all):
    c0, c1 = children_of(adj, node)
    inter = sets[c0] & sets[c1]
    empty = ~inter.any(axis=1)
    sets[node] = np.where(empty[:, None], sets[c0] | sets[c1], inter)
    score += empty
  return score


# ---------------------------------------------------------------------------------------
# Relaxed (softmin) Sankoff — SURVEY §8a rows R1/R2.  NOT in the reference: parity unpinned.
# Restatement = sankoff.py:55-92 with  min_t x  ->  -T * logsumexp_t(-x / T).
# ---------------------------------------------------------------------------------------


def _softmin(x, T, axis):
  m = x.min(axis=axis, keepdims=True)
  z = np.exp(-(x - m) / T).sum(axis=axis, keepdims=True)
  return np.squeeze(m - T * np.log(z), axis=axis)


def leaf_cost_from_states(leaf_states, n_states, dtype=np.float64):
  """Hard leaves as a cost table: ``1e5 * (1 - onehot)`` shaped (n_leaves, L, A) (R1)."""
  leaf_states = np.asarray(leaf_states).astype(np.int64)
  n, L = leaf_states.shape
  lc = np.full((n, L, n_states), float(SENTINEL), dtype=dtype)
  ii, ll = np.nonzero((leaf_states >= 0) & (leaf_states < n_states))
  lc[ii, ll, leaf_states[ii, ll]] = 0
  return lc


def soft_sankoff_fwd(adj, cost, leaf_cost, n_all, n_leaves, T, dtype=np.float64):
  """Relaxed forward.  ``leaf_cost`` (n_leaves, L, A).  Returns dp (L, n_all, A), total."""
  adj = np.array(adj, dtype=np.float32, copy=True)
  adj[-1, -1] = 0
  cost = np.asarray(cost, dtype=dtype)
  leaf_cost = np.asarray(leaf_cost, dtype=dtype)
  _, L, A = leaf_cost.shape
  dp = np.full((L, n_all, A), float(SENTINEL), dtype=dtype)
  dp[:, :n_leaves, :] = np.transpose(leaf_cost, (1, 0, 2))
  T = dtype(T)
  for node in range(n_leaves, n_all):
    ch = children_of(adj, node)
    tot = np.zeros((L, A), dtype=dtype)
    for k in range(2):
      c = int(ch[k])
      row = dp[:, c, :] if (0 <= c < node) else np.full((L, A), float(SENTINEL), dtype=dtype)
      ca = cost[None, :, :] + row[:, None, :]
      tot = tot + _softmin(ca, T, axis=2)
    dp[:, node, :] = tot
  total = _softmin(dp[:, n_all - 1, :], T, axis=1).sum()
  return dp, total


def soft_sankoff_bwd(adj, cost, dp, n_all, n_leaves, T, g=1.0, dtype=np.float64):
  """Relaxed VJP (R2).  Returns gbar (L, n_all, A) = d total / d dp rows (for leaves this is
  d total / d leaf_cost; for internal nodes the posterior-like marginals = soft ancestral
  reconstruction) and dC (A, A)."""
  adj = np.array(adj, dtype=np.float32, copy=True)
  adj[-1, -1] = 0
  cost = np.asarray(cost, dtype=dtype)
  dp = np.asarray(dp, dtype=dtype)
  L, _, A = dp.shape
  T = dtype(T)
  gbar = np.zeros((L, n_all, A), dtype=dtype)
  root = dp[:, n_all - 1, :]
  m = root.min(axis=1, keepdims=True)
  w = np.exp(-(root - m) / T)
  gbar[:, n_all - 1, :] = g * w / w.sum(axis=1, keepdims=True)
  dC = np.zeros((A, A), dtype=dtype)
  reach = np.zeros(n_all, dtype=bool)
  reach[n_all - 1] = True
  for node in range(n_all - 1, n_leaves - 1, -1):
    if not reach[node]:
      continue
    ch = children_of(adj, node)
    gn = gbar[:, node, :]  # (L, s)
    for k in range(2):
      c = int(ch[k])
      valid = 0 <= c < node
      row = dp[:, c, :] if valid else np.full((L, A), float(SENTINEL), dtype=dtype)
      ca = cost[None, :, :] + row[:, None, :]  # (L, s, t)
      sm = _softmin(ca, T, axis=2)  # (L, s)
      wgt = np.exp(-(ca - sm[:, :, None]) / T)  # rows sum to 1 over t
      contrib = gn[:, :, None] * wgt  # (L, s, t)
      dC += contrib.sum(axis=0)
      if valid:
        gbar[:, c, :] += contrib.sum(axis=1)
        reach[c] = True
  return gbar, dC


def pointer_words(tables, n_states):
  """The compact back-pointer form the CUDA library writes under a UNIFORM cost (include/trex_b200.h,
  TREX_MODE_COMPACT_ARG), derived from the forward tables: per (site, internal node, child slot) a 32-bit word with
  bits 0-4 = first minimum j of the child's dp row and bit 5+s = "the pointer of parent state s is s itself".
  Returns uint32 (L, n_all, 2); rows of leaves stay 0.  ``decode_pointer_words`` inverts it; the round trip equals
  ``tables['argmin']`` exactly when every pointer is s or j — the property the format rests on."""
  dp, arg, kids = tables["dp"], tables["argmin"], tables["children"]
  L, n_all, _ = dp.shape
  assert n_states <= 27
  words = np.zeros((L, n_all, 2), dtype=np.uint32)
  s = np.arange(n_states)
  for node in range(n_all):
    for k in range(2):
      c = int(kids[node, k])
      if kids[node, 0] == -1 and kids[node, 1] == -1:
        continue  # a leaf row of the table
      # a missing child (-1) reads the LAST row as the reference's indexing does; the root is filled last, so that row
      # was still all 1e5 when this node was computed: first minimum 0
      j = dp[:, c, :].argmin(axis=1).astype(np.uint32) if c >= 0 else np.zeros(L, dtype=np.uint32)
      stay = (arg[:, node, k, :] == s[None, :]).astype(np.uint32)
      words[:, node, k] = j | (stay << (5 + s)[None, :].astype(np.uint32)).sum(axis=1, dtype=np.uint32)
  return words


def decode_pointer_words(words, n_states):
  """(L, n_all, 2) uint32 -> (L, n_all, 2, A) int32 pointers: bit 5+s ? s : bits 0-4."""
  s = np.arange(n_states, dtype=np.uint32)
  w = words[..., None]
  return np.where((w >> (5 + s)) & 1, s, w & 31).astype(np.int32)
