"""CPU check of the property the compact back-pointer form rests on (include/trex_b200.h, TREX_MODE_COMPACT_ARG;
sankoff_uniform.cu): under a uniform cost C = a (1 - I) the reference's argmin over t of C[s,t] + dp[child,t]
(sankoff.py:67-69, first minimum) is, for EVERY parent state s, either s itself or the first minimum j of the child's
row — so one word (j + one bit per s) per child holds the whole (A,) pointer row — and a leaf child's word has the
closed form the kernel uses.  Checked on the NumPy restatement of the reference (itself pinned to the golden vectors),
with stateless leaf sites, missing children and ties; the GPU tests compare the library's words with the same tables."""

import numpy as np
import pytest

from oracle import sankoff_oracle as so
from trex_b200 import synth


def _tables(kind, n, A, L, a, seed, gaps=True, drop_edge=False):
  rng = np.random.default_rng(seed)
  par = {"balanced": synth.balanced_parents, "cat": synth.caterpillar_parents}.get(kind, lambda k: synth.random_join_parents(k, rng))(n)
  adj = synth.parents_to_adjacency(par)
  states = synth.evolve(par, L, A, 0.3, rng)[:n].astype(np.float32)
  if gaps:
    states[rng.random(states.shape) < 0.1] = 255          # no state: an all-1e5 row (the reference's dropped scatter)
  if drop_edge:
    adj[0, :] = 0                                           # leaf 0 hangs nowhere: its parent has a missing child (-1)
  cost = (a * synth.unit_cost(A)).astype(np.float32)
  return so.run_sankoff_tables(adj, cost, states, 2 * n - 1, A, n), states, adj


@pytest.mark.parametrize("kind,n,A,L,a", [
  ("balanced", 16, 4, 200, 1.0), ("random", 13, 20, 150, 1.0), ("cat", 12, 5, 120, 3.0), ("random", 30, 20, 80, 2.0),
  ("random", 9, 4, 300, 0.0), ("balanced", 8, 27, 60, 1.0)])
@pytest.mark.parametrize("drop_edge", [False, True])
def test_every_pointer_is_the_state_itself_or_the_childs_first_minimum(kind, n, A, L, a, drop_edge):
  tabs, states, _ = _tables(kind, n, A, L, a, seed=n * 31 + A, drop_edge=drop_edge)
  words = so.pointer_words(tabs, A)
  internal = tabs["children"][:, 0] != -1
  internal |= tabs["children"][:, 1] != -1
  got = so.decode_pointer_words(words, A)
  assert np.array_equal(got[:, internal], tabs["argmin"][:, internal])     # the round trip loses nothing
  assert internal.sum() >= n - 2


@pytest.mark.parametrize("A,a", [(4, 1.0), (20, 1.0), (5, 7.0)])
def test_leaf_children_have_the_closed_form_word(A, a):
  """A leaf with observed state o: only state o keeps itself, every other state points to o (word = o | 1 << (5+o));
  a stateless leaf (or a missing child): every state keeps itself, j = 0 — for a > 0 (sankoff_uniform.cu: masks)."""
  n, L = 20, 250
  tabs, states, adj = _tables("random", n, A, L, a, seed=A, drop_edge=True)
  words = so.pointer_words(tabs, A)
  kids = tabs["children"]
  checked = 0
  for node in range(n, 2 * n - 1):
    for k in range(2):
      c = int(kids[node, k])
      if c >= n:
        continue
      if c < 0:
        want = np.full(L, ((1 << A) - 1) << 5, dtype=np.uint32)
      else:
        o = states[c].astype(np.int64)
        has = (o >= 0) & (o < A)
        want = np.where(has, o | (1 << (5 + np.where(has, o, 0))), ((1 << A) - 1) << 5).astype(np.uint32)
      assert np.array_equal(words[:, node, k], want), (node, k, c)
      checked += 1
  assert checked >= n


def test_a_general_cost_matrix_does_not_have_the_property():
  """Why the flag only applies to a uniform matrix: with C random the pointer of s can be a third state."""
  rng = np.random.default_rng(0)
  n, A, L = 16, 8, 400
  par = synth.random_join_parents(n, rng)
  adj = synth.parents_to_adjacency(par)
  states = synth.evolve(par, L, A, 0.4, rng)[:n].astype(np.float32)
  tabs = so.run_sankoff_tables(adj, synth.random_int_cost(A, 5, rng), states, 2 * n - 1, A, n)
  got = so.decode_pointer_words(so.pointer_words(tabs, A), A)
  assert not np.array_equal(got[:, n:], tabs["argmin"][:, n:])
