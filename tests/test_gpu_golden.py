"""GPU: libtrex_b200 (through the C ABI / the reference-signature wrappers) against the golden
vectors produced by the UNMODIFIED reference source (tests/golden/make_golden.py)."""

import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_golden.npz"))
SK = sorted({k.split("/")[1] for k in G.files if k.startswith("sk/")})
TR = sorted({k.split("/")[1] for k in G.files if k.startswith("tree/")})


def _t(a):
  import torch

  return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize("name", SK)
def test_run_sankoff_matches_reference_golden(name):
  from trex_b200.sankoff import run_sankoff, vectorized_dp

  g = lambda k: G[f"sk/{name}/{k}"]
  n_all, A, n = (int(x) for x in g("sizes"))
  rec, dp, tot = run_sankoff(_t(g("adj")), _t(g("cost")), _t(g("seqs")), n_all, A, n, return_path=True)
  assert np.array_equal(dp.cpu().numpy(), g("dp"))          # bit-exact, reference layout (L, n_all, A)
  assert np.array_equal(rec.cpu().numpy(), g("recon"))      # leaves passed through, ancestors traced back
  assert abs(float(tot) - float(g("total"))) <= 1e-6 * abs(float(g("total")))
  if np.all(g("cost") == np.round(g("cost"))):
    assert float(tot) == float(g("total"))
  rec0, _, _ = run_sankoff(_t(g("adj")), _t(g("cost")), _t(g("seqs")), n_all, A, n, return_path=False)
  assert np.array_equal(rec0.cpu().numpy(), g("recon_nopath"))
  L = g("seqs").shape[1]
  adj = g("adj").copy()
  adj[-1, -1] = 0
  dp2, bt = vectorized_dp(_t(adj), _t(np.full((L, n_all, A), 1e5, np.float32)),
                          _t(np.zeros((L, n_all, A, 4), np.float32)), _t(g("seqs")), _t(g("cost")))
  assert np.array_equal(dp2.cpu().numpy(), g("dp"))
  assert np.array_equal(bt.cpu().numpy(), g("bt"))          # back-pointer table incl. child ids


def test_run_dp_malformed_matches_reference_golden():
  from trex_b200.sankoff import run_dp

  dp, bt = run_dp(_t(G["run_dp/malformed/adj"]), _t(np.full((3, 2), 1e5, np.float32)),
                  _t(np.zeros((3, 2, 4), np.float32)), _t(np.array([0, 1, 0], np.float32)),
                  _t(np.array([[0, 1], [1, 0]], np.float32)))
  assert np.array_equal(dp.cpu().numpy(), G["run_dp/malformed/dp"])
  assert np.array_equal(bt.cpu().numpy(), G["run_dp/malformed/bt"])


@pytest.mark.parametrize("name", TR)
def test_tree_costs_match_reference_golden(name):
  from trex_b200 import tree as T

  g = lambda k: G[f"tree/{name}/{k}"]
  S = _t(g("S"))
  for an in ("soft", "bern"):
    A_ = _t(g(f"adj_{an}"))
    # reference f32 values; 1e-5 relative (north star) + an absolute floor for the cancelling form
    close = lambda a, b: abs(float(a) - float(b)) <= 1e-5 * abs(float(b)) + 2e-4
    assert close(T.compute_surrogate_cost(S, A_), g(f"surrogate_{an}"))
    assert close(T.compute_soft_cost(S, A_), g(f"soft_none_{an}"))
    assert close(T.compute_soft_cost(S, A_, _t(g("Csym"))), g(f"soft_sym_{an}"))
    assert close(T.compute_soft_cost(S, A_, _t(g("Casym"))), g(f"soft_asym_{an}"))
    assert close(T.compute_soft_cost(S, A_, _t(g("cvec"))), g(f"soft_vec_{an}"))
