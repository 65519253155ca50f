"""CPU tests of the oracle itself: the hand-derived known answer for the reference's own test
inputs, the assertions of the reference's tests, and independent cross-checks."""

import numpy as np
import pytest

from oracle import sankoff_oracle as so
from oracle import tree_oracle as to
from trex_b200 import synth


def _ref_test_tree():
  """Inputs of /root/reference/tests/test_sankoff.py:39-67."""
  adj = np.zeros((5, 5), dtype=np.int32)
  adj[0, 3] = adj[1, 3] = adj[2, 4] = adj[3, 4] = 1
  cost = np.ones((2, 2), np.float32) - np.eye(2, dtype=np.float32)
  seqs = np.array([[0, 1], [1, 0], [0, 0], [0, 0], [0, 0]], dtype=np.float32)
  return adj, cost, seqs


def test_hand_derived_known_answer():
  adj, cost, seqs = _ref_test_tree()
  recon, dp, total = so.run_sankoff(adj, cost, seqs[:3], 5, 2, 3, return_path=True)
  assert recon.shape == (5, 2) and dp.shape == (2, 5, 2)  # test_sankoff.py:68-69
  assert total == np.float32(2.0)
  for site in range(2):
    assert dp[site, 3].tolist() == [1.0, 1.0]
    assert dp[site, 4].tolist() == [1.0, 2.0]
  assert dp[0, 0].tolist() == [0.0, 1e5] and dp[0, 1].tolist() == [1e5, 0.0]
  assert recon.tolist() == [[0, 1], [1, 0], [0, 0], [0, 0], [0, 0]]
  assert np.all(recon[:3] == seqs[:3])  # test_sankoff.py:72


def test_run_dp_reference_unit_test():
  """/root/reference/tests/test_sankoff.py:9-36 (malformed adjacency, fill_value=-1 path)."""
  adj = np.array([[0, 1, 0], [0, 1, 0], [0, 0, 0]], dtype=np.float32)
  seqs = np.array([[0], [1], [0]], dtype=np.float32)
  cost = np.array([[0, 1], [1, 0]], dtype=np.float32)
  dp = np.full((3, 2), 1e5, dtype=np.float32)
  bt = np.zeros((3, 2, 4), dtype=np.float32)
  dp_out, bt_out = so.run_dp(adj, dp, bt, seqs[:, 0], cost)
  assert dp_out.shape == (3, 2) and bt_out.shape == (3, 2, 4)
  assert dp_out[0, 0] == 0 and dp_out[1, 1] == 0
  # node 2 has no children: both slots read row -1 (itself, still 1e5) -> 2 * (0 + 1e5)
  assert dp_out[2].tolist() == [2e5, 2e5]
  assert bt_out[2, :, 0].tolist() == [-1, -1]


def test_stack_backtrack_equals_topdown(rng):
  for n, A, L in [(4, 4, 7), (8, 3, 5), (16, 4, 9)]:
    par = synth.random_join_parents(n, rng)
    adj = synth.parents_to_adjacency(par)
    states = synth.evolve(par, L, A, 0.3, rng)
    cost = synth.random_int_cost(A, 4, rng)
    n_all = 2 * n - 1
    tabs = so.run_sankoff_tables(adj, cost, states[:n], n_all, A, n)
    td = so.traceback_topdown(tabs, n_all, n)
    adjz = adj.copy()
    for l in range(L):
      dp1, bt1 = so.run_dp(adjz, np.full((n_all, A), 1e5, np.float32),
                           np.zeros((n_all, A, 4), np.float32), states[:n, l].astype(np.float32), cost)
      assert np.array_equal(dp1, tabs["dp"][l])
      rs = int(np.argmin(dp1[n_all - 1]))
      rec = so.backtrack(n_all - 1, rs, bt1, n_all, n)
      assert np.array_equal(rec[n:], td[n:, l])


def test_score_equals_bruteforce_and_fitch(rng):
  for n, A in [(3, 2), (4, 3), (4, 4)]:
    par = synth.random_join_parents(n, rng)
    adj = synth.parents_to_adjacency(par)
    n_all = 2 * n - 1
    states = rng.integers(0, A, size=(n_all, 6)).astype(np.uint8)
    for cost in (synth.unit_cost(A), synth.random_int_cost(A, 5, rng)):
      _, dp, total = so.run_sankoff(adj, cost, states[:n], n_all, A, n)
      bf = so.brute_force_score(adj, cost, states[:n], n_all, A, n)
      assert np.array_equal(so.per_site_root_min(dp).astype(np.float64), bf)
      assert float(total) == bf.sum()
    _, dp, _ = so.run_sankoff(adj, synth.unit_cost(A), states[:n], n_all, A, n)
    assert np.array_equal(so.per_site_root_min(dp).astype(np.int64),
                          so.fitch_score(adj, states[:n], n_all, A, n))


def test_sankoff_score_equals_compute_cost(rng):
  """The one numeric consistency check of the reference: tests/test_convergence.py:69-73."""
  n, A, L = 4, 4, 20
  par = synth.groundtruth_parents(n)
  adj = synth.parents_to_adjacency(par)
  states = synth.evolve(par, L, A, 0.3, rng)
  cost = synth.unit_cost(A)
  recon, _, total = so.run_sankoff(adj, cost, states[:n], 2 * n - 1, A, n, return_path=True)
  onehot = np.eye(A, dtype=np.float32)[recon.astype(np.int64)]
  assert abs(float(total) - float(to.compute_cost(onehot, adj, cost))) <= 1e-3


def test_tree_reference_unit_tests():
  """Assertions of /root/reference/tests/test_tree.py."""
  soft = np.array([[0.1, 0.9, 0.0], [0.8, 0.1, 0.1], [0.3, 0.3, 0.4]])
  oh = to.discretize_tree_topology(soft, 3)
  assert oh.shape == (3, 3) and np.all(oh.sum(axis=1) == 1)
  out = to.update_tree(np.ones((2, 1)), np.zeros((2, 1)))
  assert out.shape == (3, 3) and np.allclose(out.sum(axis=1), 1.0)
  assert to.update_tree(np.ones((1, 0)), np.zeros((1, 0))).shape == (2, 2)
  assert to.update_seq(np.ones((2, 5, 4)), np.zeros((4, 5, 4)), 1.0).shape == (4, 5, 4)
  assert to.enforce_graph_constraints(np.eye(5), 10.0) >= 0
  seqs = np.eye(4)[np.array([[0, 1, 2], [3, 2, 1]])]
  assert to.compute_surrogate_cost(seqs, np.eye(2)) >= 0
  seqs = np.eye(2)[np.array([[0, 1], [1, 0], [0, 0]])]
  assert to.compute_cost(seqs, np.eye(3), np.ones((2, 2)) - np.eye(2)) >= 0
  loss = to.compute_loss(np.ones((2, 5, 4)), np.ones((3, 2)), np.zeros((3, 2)),
                         np.zeros((4, 5, 4)), 1.0, np.eye(4))
  assert np.isfinite(loss) and loss >= 0


def test_soft_cost_matches_naive_form(rng):
  N, L, Q = 9, 11, 5
  S = to.softmax(rng.normal(size=(N, L, Q)))
  adj = (rng.random((N, N)) < 0.3).astype(np.float64)
  Csym = synth.random_int_cost(Q, 3, rng).astype(np.float64)
  for C in (None, Csym):
    got, _ = to.compute_soft_cost(S, adj, C, dtype=np.float64)
    want = to.naive_compute_soft_cost(S, adj, C)
    assert abs(got - want) <= 1e-10 * abs(want)


def test_soft_cost_gradients_finite_differences(rng):
  N, L, Q = 5, 3, 4
  S = to.softmax(rng.normal(size=(N, L, Q)))
  adj = rng.random((N, N))
  for C in (None, rng.random((Q, Q)), rng.random(Q)):
    cost, dS, dA, dC = to.soft_cost_value_and_grad(S, adj, C)
    f = lambda s, a, c: to.compute_soft_cost(s, a, c, dtype=np.float64)[0]
    assert abs(cost - f(S, adj, C)) <= 1e-10 * abs(cost)
    h = 1e-6
    for _ in range(6):
      i = tuple(rng.integers(0, d) for d in S.shape)
      Sp, Sm = S.copy(), S.copy()
      Sp[i] += h
      Sm[i] -= h
      assert abs((f(Sp, adj, C) - f(Sm, adj, C)) / (2 * h) - dS[i]) <= 1e-6 * (1 + abs(dS[i]))
      j = tuple(rng.integers(0, d) for d in adj.shape)
      Ap, Am = adj.copy(), adj.copy()
      Ap[j] += h
      Am[j] -= h
      assert abs((f(S, Ap, C) - f(S, Am, C)) / (2 * h) - dA[j]) <= 1e-6 * (1 + abs(dA[j]))
      if C is not None:
        k = tuple(rng.integers(0, d) for d in C.shape)
        Cp, Cm = C.copy(), C.copy()
        Cp[k] += h
        Cm[k] -= h
        assert abs((f(S, adj, Cp) - f(S, adj, Cm)) / (2 * h) - dC[k]) <= 1e-6 * (1 + abs(dC[k]))


def test_relaxed_sankoff_limits_and_gradients(rng):
  n, A, L = 6, 4, 5
  par = synth.random_join_parents(n, rng)
  adj = synth.parents_to_adjacency(par)
  n_all = 2 * n - 1
  states = synth.evolve(par, L, A, 0.3, rng)
  cost = synth.random_int_cost(A, 3, rng).astype(np.float64)
  _, dp_h, total_h = so.run_sankoff(adj, cost, states[:n], n_all, A, n)
  lc = so.leaf_cost_from_states(states[:n], A)
  # T -> 0 limit equals the hard score
  _, tot = so.soft_sankoff_fwd(adj, cost, lc, n_all, n, 1e-3)
  assert abs(tot - float(total_h)) < 1e-2
  # gradients vs central differences (float64), soft leaves so that d leaf_cost is non-trivial
  lc = rng.random((n, L, A)) * 3
  T = 0.7
  dp, tot = so.soft_sankoff_fwd(adj, cost, lc, n_all, n, T)
  gbar, dC = so.soft_sankoff_bwd(adj, cost, dp, n_all, n, T)
  assert np.allclose(gbar[:, n:, :].sum(axis=2), 1.0)  # marginals
  h = 1e-6
  for _ in range(8):
    k = tuple(rng.integers(0, d) for d in cost.shape)
    cp, cm = cost.copy(), cost.copy()
    cp[k] += h
    cm[k] -= h
    fd = (so.soft_sankoff_fwd(adj, cp, lc, n_all, n, T)[1] - so.soft_sankoff_fwd(adj, cm, lc, n_all, n, T)[1]) / (2 * h)
    assert abs(fd - dC[k]) <= 1e-6 * (1 + abs(dC[k]))
    i = tuple(rng.integers(0, d) for d in lc.shape)
    lp, lm = lc.copy(), lc.copy()
    lp[i] += h
    lm[i] -= h
    fd = (so.soft_sankoff_fwd(adj, cost, lp, n_all, n, T)[1] - so.soft_sankoff_fwd(adj, cost, lm, n_all, n, T)[1]) / (2 * h)
    assert abs(fd - gbar[i[1], i[0], i[2]]) <= 1e-6 * (1 + abs(fd))
