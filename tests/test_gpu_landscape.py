"""GPU: trex_landscape_loss_fwd_bwd through the reference-shaped host API (trex_b200/landscape.py) against
the golden vectors of the unmodified reference and the float64 oracle (value, logits, gradients)."""

import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-5
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "landscape_golden.npz"))
CASES = sorted({k.split("/")[1] for k in G.files})


def _t(a):
  import torch

  return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize("name", CASES)
def test_matches_reference_golden(name):
  import torch
  from trex_b200 import landscape as LS

  g = lambda k: G[f"ls/{name}/{k}"]
  n, L, Q, k = [int(x) for x in g("sizes")]
  land = {"interactions": _t(g("inter")), "fitness_tables": _t(g("tables"))}
  N = g("adj").shape[0]
  S = _t(g("masked")).clone()
  S[n:] = torch.softmax(_t(g("anc")), -1)
  logits = LS.compute_parental_logits(S, land, k)
  assert np.abs(logits.cpu().numpy() - g("logits_of_S")).max() <= 2e-6
  mask = _t(g("mask")) if int(g("has_mask")) else None
  loss = LS._compute_loss_landscape_aware_stacked(_t(g("anc")), _t(g("masked")), n, land, _t(g("adj")), N,
                                                  float(g("lambda")), k, 1.0, mask)
  assert abs(float(loss) - float(g("loss"))) <= RTOL * abs(float(g("loss")))
  loss0 = LS._compute_loss_landscape_aware_stacked(_t(g("anc")), _t(g("masked")), n, land, _t(g("adj")), N, 0.0, k)
  assert abs(float(loss0) - float(g("loss_lambda0"))) <= RTOL * abs(float(g("loss_lambda0")))


@pytest.mark.parametrize("n,L,Q,k,masked", [(64, 1000, 4, 1, False), (16, 257, 20, 1, True), (8, 40, 20, 2, True),
                                              (8, 33, 4, 3, False), (4, 3, 3, 2, True),
                                              (4, 4, 20, 3, False)])  # the last: a 640 KB table per site, read from L2
def test_value_and_gradients_match_float64_oracle(n, L, Q, k, masked):
  import torch
  from oracle import landscape_oracle as lo
  from oracle import tree_oracle
  from trex_b200 import landscape as LS
  from trex_b200 import synth

  rng = np.random.default_rng(100 * k + Q + n)
  parent = synth.balanced_parents(n)
  N = parent.shape[0]
  adj = synth.parents_to_adjacency(parent).astype(np.float32)
  S = tree_oracle.softmax(rng.normal(size=(N, L, Q))).astype(np.float32)
  inter = rng.integers(0, L, size=(L, k)).astype(np.int32)
  tables = rng.random((L, Q ** (k + 1))).astype(np.float32)
  mask = (rng.random(L) < 0.7).astype(np.float32) if masked else None
  land = {"interactions": _t(inter), "fitness_tables": _t(tables)}
  St = _t(S).requires_grad_(True)
  cost = LS.fitness_cost(St, _t(adj), land, k, None if mask is None else _t(mask))
  cost.backward()
  ref = lo.fitness_cost(S, adj, inter, tables, k, mask)
  assert abs(float(cost) - ref) <= RTOL * abs(ref)
  gref = lo.fitness_cost_grad(S, adj, inter, tables, k, mask)
  got = St.grad.cpu().numpy().astype(np.float64)
  assert np.abs(got - gref).max() <= RTOL * np.abs(gref).max()


def test_gradient_flows_to_the_ancestor_logits():
  import torch
  from oracle import landscape_oracle as lo
  from trex_b200 import landscape as LS

  name = "k1_q4"
  g = lambda k: G[f"ls/{name}/{k}"]
  n, L, Q, k = [int(x) for x in g("sizes")]
  land = {"interactions": _t(g("inter")), "fitness_tables": _t(g("tables"))}
  N = g("adj").shape[0]
  anc = _t(g("anc")).requires_grad_(True)
  lam = float(g("lambda"))
  loss = LS._compute_loss_landscape_aware_stacked(anc, _t(g("masked")), n, land, _t(g("adj")), N, lam, k)
  loss.backward()
  a = g("anc").astype(np.float64)
  f = lambda x: lo.loss_landscape_aware_stacked(x, g("masked"), n, g("inter"), g("tables"), g("adj"), lam, k)
  h = 1e-5
  for idx in [(0, 0, 0), (3, 5, 2), (6, 11, 3)]:
    ap, am = a.copy(), a.copy()
    ap[idx] += h
    am[idx] -= h
    fd = (f(ap) - f(am)) / (2 * h)
    assert abs(float(anc.grad[idx]) - fd) <= 1e-4 * max(1.0, abs(fd))
