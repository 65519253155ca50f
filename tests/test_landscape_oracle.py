"""CPU: the restated landscape-aware loss (oracle/landscape_oracle.py) against golden vectors of the
UNMODIFIED reference functions (evals/benchmark.py:586-663, 235-306; tests/golden/make_golden_landscape.py)
and its closed-form gradient against central differences."""

import os

import numpy as np
import pytest

from oracle import landscape_oracle as lo

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "landscape_golden.npz"))
CASES = sorted({k.split("/")[1] for k in G.files})


def case(name):
  g = lambda k: G[f"ls/{name}/{k}"]
  n, L, Q, k = [int(x) for x in g("sizes")]
  mask = g("mask") if int(g("has_mask")) else None
  return g, n, L, Q, k, mask


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_golden(name):
  g, n, L, Q, k, mask = case(name)
  S = np.array(g("masked"), dtype=np.float64)
  a = g("anc").astype(np.float64)
  e = np.exp(a - a.max(-1, keepdims=True))
  S[n:] = e / e.sum(-1, keepdims=True)
  logits = lo.compute_parental_logits(S, g("inter"), g("tables"), k)
  assert np.abs(logits - g("logits_of_S")).max() <= 1e-6
  for lam, key in ((float(g("lambda")), "loss"), (0.0, "loss_lambda0")):
    m = mask if lam > 0 else None
    loss = lo.loss_landscape_aware_stacked(g("anc"), g("masked"), n, g("inter"), g("tables"), g("adj"), lam, k, 1.0, m)
    assert abs(loss - float(g(key))) <= 2e-6 * abs(float(g(key)))


@pytest.mark.parametrize("k", [1, 2, 3])
def test_gradient_by_central_differences(k):
  rng = np.random.default_rng(k)
  N, L, Q = 7, 5, 3
  S = rng.random((N, L, Q))
  S /= S.sum(-1, keepdims=True)
  adj = np.zeros((N, N))
  for i, p in enumerate([4, 4, 5, 5, 6, 6, 6]):
    adj[i, p] = 1
  inter = rng.integers(0, L, size=(L, k))
  F = rng.random((L, Q ** (k + 1)))
  mask = np.array([1, 0, 1, 1, 1.0])
  grad = lo.fitness_cost_grad(S, adj, inter, F, k, mask)
  h = 1e-6
  for idx in [(0, 0, 0), (4, 2, 1), (6, 3, 2), (5, 1, 0), (6, 4, 1), (2, 1, 1)]:
    sp, sm = S.copy(), S.copy()
    sp[idx] += h
    sm[idx] -= h
    fd = (lo.fitness_cost(sp, adj, inter, F, k, mask) - lo.fitness_cost(sm, adj, inter, F, k, mask)) / (2 * h)
    assert abs(fd - grad[idx]) <= 1e-8 + 1e-6 * abs(fd)
