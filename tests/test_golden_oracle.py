"""CPU: the oracle restatements against the golden vectors produced by the UNMODIFIED reference
source (tests/golden/make_golden.py; reference run under the NumPy stand-in for jax)."""

import os

import numpy as np
import pytest

from oracle import c_oracle
from oracle import sankoff_oracle as so
from oracle import tree_oracle as to

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_golden.npz"))
SK = sorted({k.split("/")[1] for k in G.files if k.startswith("sk/")})
TR = sorted({k.split("/")[1] for k in G.files if k.startswith("tree/")})


@pytest.mark.parametrize("name", SK)
def test_sankoff_restatements_match_reference(name):
  g = lambda k: G[f"sk/{name}/{k}"]
  n_all, A, n = (int(x) for x in g("sizes"))
  rec, dp, tot = so.run_sankoff(g("adj"), g("cost"), g("seqs"), n_all, A, n, return_path=True)
  assert np.array_equal(dp, g("dp")) and np.array_equal(rec, g("recon")) and tot == g("total")
  rec0, _, _ = so.run_sankoff(g("adj"), g("cost"), g("seqs"), n_all, A, n, return_path=False)
  assert np.array_equal(rec0, g("recon_nopath"))
  tabs = so.run_sankoff_tables(g("adj"), g("cost"), g("seqs"), n_all, A, n)
  bt = g("bt")  # (L, n_all, A, 4): child0, argmin0, child1, argmin1
  assert np.array_equal(bt[:, n:, :, 1], tabs["argmin"][:, n:, 0, :])
  assert np.array_equal(bt[:, n:, :, 3], tabs["argmin"][:, n:, 1, :])
  assert np.array_equal(bt[0, n:, 0, 0], tabs["children"][n:, 0]) and np.array_equal(bt[0, n:, 0, 2], tabs["children"][n:, 1])
  # the plain-C port (the timed CPU baseline) as well
  crec, cdp, ctot, _ = c_oracle.run_sankoff(g("adj"), g("cost"), g("seqs"), n_all, A, n, return_path=True)
  assert np.array_equal(cdp, g("dp")) and np.array_equal(crec, g("recon"))
  # the f32 sum over sites (sankoff.py:187) has no defined order: exact for integer costs, 1e-6 otherwise
  assert abs(float(ctot) - float(g("total"))) <= 1e-6 * abs(float(g("total")))
  if np.all(g("cost") == np.round(g("cost"))):
    assert ctot == g("total")


def test_run_dp_malformed_adjacency_matches_reference():
  adj = G["run_dp/malformed/adj"]
  dp, bt = so.run_dp(adj, np.full((3, 2), 1e5, np.float32), np.zeros((3, 2, 4), np.float32),
                     np.array([0, 1, 0], np.float32), np.array([[0, 1], [1, 0]], np.float32))
  assert np.array_equal(dp, G["run_dp/malformed/dp"]) and np.array_equal(bt, G["run_dp/malformed/bt"])


@pytest.mark.parametrize("name", TR)
def test_tree_restatements_match_reference(name):
  g = lambda k: G[f"tree/{name}/{k}"]
  S = to.update_seq(g("logits"), g("S0"), g("temp"))
  assert np.allclose(S, g("S"), rtol=1e-6, atol=1e-7)
  adj_soft = to.update_tree(g("tree_params"), g("gumbel"), 0.9)
  assert np.allclose(adj_soft, g("adj_soft"), rtol=1e-6, atol=1e-7)
  N = S.shape[0]
  for an in ("soft", "bern"):
    A_ = g(f"adj_{an}")
    close = lambda a, b: abs(float(a) - float(b)) <= 2e-5 * abs(float(b)) + 1e-4  # f32 sums, order differs
    assert close(to.compute_surrogate_cost(g("S"), A_), g(f"surrogate_{an}"))
    assert close(to.compute_soft_cost(g("S"), A_, None)[0], g(f"soft_none_{an}"))
    assert close(to.compute_soft_cost(g("S"), A_, g("Csym"))[0], g(f"soft_sym_{an}"))
    assert close(to.compute_soft_cost(g("S"), A_, g("Casym"))[0], g(f"soft_asym_{an}"))
    assert close(to.compute_soft_cost(g("S"), A_, g("cvec"))[0], g(f"soft_vec_{an}"))
    assert close(to.naive_compute_soft_cost(g("S"), A_, None), g(f"naive_none_{an}"))
    assert close(to.naive_compute_soft_cost(g("S"), A_, g("Csym")), g(f"naive_sym_{an}"))
    assert close(to.enforce_graph_constraints(A_, 10.0), g(f"constraint_{an}"))
    assert np.array_equal(to.discretize_tree_topology(A_, N), g(f"discrete_{an}"))
    # closed-form <M,G> (what the kernels evaluate) == the reference's three-term form
    assert close(to.soft_cost_value_and_grad(g("S"), A_, g("Casym"))[0], g(f"soft_asym_{an}"))
  assert float(to.compute_cost(g("onehot"), g("adj_soft"), g("Csym"))) == float(g("compute_cost"))
  loss = to.compute_loss(g("logits"), g("tree_params"), g("loss_gumbel"), g("S0"), g("temp"), g("adj_bern"))
  assert abs(float(loss) - float(g("loss"))) <= 2e-5 * abs(float(g("loss")))
  lf = to.compute_loss(g("logits"), g("tree_params"), None, g("S0"), g("temp"), g("adj_bern"), fix_tree=True)
  assert abs(float(lf) - float(g("loss_fix_tree"))) <= 2e-5 * abs(float(g("loss_fix_tree")))
