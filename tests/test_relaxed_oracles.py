"""CPU tests of the relaxed-Sankoff checkers against each other (SURVEY rows R1/R2 — parity unpinned: the
reference has no relaxed Sankoff, so the oracle is pinned by an independent derivation instead).

* ``oracle/autograd_oracle.py`` — literal transcription of the reference forward (sankoff.py:60-79, :187)
  with min -> softmin, differentiated by torch.autograd in float64: no hand-derived reverse recurrence.
* ``oracle/sankoff_oracle.py`` — NumPy forward + hand-derived VJP (what the kernels implement).
* ``oracle/sankoff_oracle.c`` — the same in C: float64 (checker at the benchmark shapes) and float32 (the
  timed CPU arm of bench.py).
"""

import numpy as np
import pytest

from oracle import autograd_oracle as ao
from oracle import c_oracle
from oracle import sankoff_oracle as so
from trex_b200 import synth


def _tree(kind, n, rng):
  par = synth.balanced_parents(n) if kind == "balanced" else (
    synth.caterpillar_parents(n) if kind == "cat" else synth.random_join_parents(n, rng))
  return synth.parents_to_adjacency(par), par


def _cost(kind, A, rng):
  if kind == "unit":
    return synth.unit_cost(A)
  if kind == "int":
    return synth.random_int_cost(A, 5, rng)
  return (rng.random((A, A)) * 2).astype(np.float32)  # asymmetric, non-zero diagonal


CASES = [
  ("balanced", 16, 4, 24, "unit", 1.0),
  ("random", 11, 20, 9, "int", 1.0),
  ("random", 11, 20, 9, "int", 0.1),
  ("cat", 9, 5, 12, "float", 0.5),
  ("random", 33, 20, 6, "unit", 0.3),
  ("balanced", 64, 20, 4, "int", 1.0),   # the size at which the kernels switch to segment launches
  ("random", 70, 7, 5, "float", 2.0),
]


@pytest.mark.parametrize("kind,n,A,L,ck,T", CASES)
@pytest.mark.parametrize("leaves", ["states", "soft"])
def test_numpy_vjp_equals_autograd(kind, n, A, L, ck, T, leaves):
  rng = np.random.default_rng(n * 31 + A + L)
  adj, par = _tree(kind, n, rng)
  n_all = 2 * n - 1
  cost = _cost(ck, A, rng)
  if leaves == "states":
    lc = so.leaf_cost_from_states(synth.evolve(par, L, A, 0.25, rng)[:n], A)
  else:
    lc = rng.random((n, L, A)) * 3
  ref = ao.soft_sankoff_autograd(adj, cost, lc, n_all, n, T, g=1.7)
  dp, tot = so.soft_sankoff_fwd(adj, cost, lc, n_all, n, T)
  gbar, dC = so.soft_sankoff_bwd(adj, cost, dp, n_all, n, T, g=1.7)
  assert abs(tot - ref["total"]) <= 1e-12 * abs(ref["total"])
  assert np.abs(dp[:, n:] - ref["dp"][:, n:]).max() <= 1e-9 * np.abs(ref["dp"][:, n:]).max()
  assert np.abs(gbar - ref["gbar"]).max() <= 1e-11            # every dp row's cotangent (marginals are O(1))
  assert np.abs(gbar[:, :n] - np.transpose(ref["d_leaf"], (1, 0, 2))).max() <= 1e-11
  assert np.abs(dC - ref["d_cost"]).max() <= 1e-11 * max(1.0, np.abs(ref["d_cost"]).max())


def test_autograd_oracle_on_a_malformed_adjacency():
  """A node with one child (the second index is the -1 fill, sankoff.py:60: wraps to the unwritten root row)."""
  rng = np.random.default_rng(3)
  n, A, L, T = 6, 4, 7, 0.7
  adj, par = _tree("random", n, rng)
  n_all = 2 * n - 1
  victim = int(np.nonzero(adj[:, n])[0][0])
  adj[victim, n] = 0
  lc = rng.random((n, L, A)) * 2
  cost = _cost("float", A, rng)
  ref = ao.soft_sankoff_autograd(adj, cost, lc, n_all, n, T)
  dp, tot = so.soft_sankoff_fwd(adj, cost, lc, n_all, n, T)
  gbar, dC = so.soft_sankoff_bwd(adj, cost, dp, n_all, n, T)
  assert abs(tot - ref["total"]) <= 1e-12 * abs(ref["total"])
  assert np.abs(gbar - ref["gbar"]).max() <= 1e-9  # (softmin over 1e5-sized entries: cancellation in float64)
  assert np.abs(dC - ref["d_cost"]).max() <= 1e-9 * max(1.0, np.abs(ref["d_cost"]).max())


@pytest.mark.parametrize("ck,T", [("unit", 1.0), ("int", 0.1), ("float", 0.5)])
def test_c_float64_restatement_equals_autograd(ck, T):
  rng = np.random.default_rng(11)
  n, A, L = 64, 20, 6
  adj, par = _tree("balanced", n, rng)
  n_all = 2 * n - 1
  cost = _cost(ck, A, rng)
  states = synth.evolve(par, L, A, 0.2, rng)[:n]
  ref = ao.soft_sankoff_autograd(adj, cost, so.leaf_cost_from_states(states, A), n_all, n, T)
  tot, dC, dp, gbar = c_oracle.soft_sankoff_f64(adj, cost, n_all, A, n, T, seqs=states.astype(np.float32))
  assert abs(tot - ref["total"]) <= 1e-12 * abs(ref["total"])
  assert np.abs(dp[:, n:] - ref["dp"][:, n:]).max() <= 1e-9 * np.abs(ref["dp"][:, n:]).max()
  assert np.abs(gbar - ref["gbar"]).max() <= 1e-11
  assert np.abs(dC - ref["d_cost"]).max() <= 1e-11 * np.abs(ref["d_cost"]).max()
  # soft leaves through the same entry point
  lc = rng.random((n, L, A)) * 3
  ref = ao.soft_sankoff_autograd(adj, cost, lc, n_all, n, T)
  tot, dC, dp, gbar = c_oracle.soft_sankoff_f64(adj, cost, n_all, A, n, T, leaf_cost=lc)
  assert abs(tot - ref["total"]) <= 1e-12 * abs(ref["total"])
  assert np.abs(gbar - ref["gbar"]).max() <= 1e-11
  assert np.abs(dC - ref["d_cost"]).max() <= 1e-11 * np.abs(ref["d_cost"]).max()


@pytest.mark.parametrize("kind,n,A,L,ck,T", [("balanced", 32, 20, 40, "unit", 1.0), ("random", 21, 4, 64, "int", 0.5),
                                            ("cat", 12, 20, 33, "int", 1.0)])
def test_c_float32_timed_arm_matches_the_float64_oracle(kind, n, A, L, ck, T):
  """``c_oracle.soft_sankoff`` is half of the timed CPU arm (bench.py cpu_step): float32 like the reference's
  default dtype, so it agrees with the float64 restatement to float32 accuracy."""
  rng = np.random.default_rng(n + A)
  adj, par = _tree(kind, n, rng)
  n_all = 2 * n - 1
  cost = _cost(ck, A, rng)
  states = synth.evolve(par, L, A, 0.2, rng)[:n]
  lc = so.leaf_cost_from_states(states, A)
  dp64, tot64 = so.soft_sankoff_fwd(adj, cost, lc, n_all, n, T)
  g64, dC64 = so.soft_sankoff_bwd(adj, cost, dp64, n_all, n, T)
  tot, dC, gbar = c_oracle.soft_sankoff(adj, cost, states.astype(np.float32), n_all, A, n, T, want_gbar=True)
  assert abs(tot - tot64) <= 2e-5 * abs(tot64)
  assert np.abs(gbar - g64).max() <= 2e-4
  assert np.abs(dC - dC64).max() <= 2e-4 * np.abs(dC64).max()


def test_c_oracle_thread_count_can_be_set():
  import os

  want = max(1, (os.cpu_count() or 1))
  assert c_oracle.set_num_threads(want) == want
  assert c_oracle.set_num_threads(1) == 1
  c_oracle.set_num_threads(want)
