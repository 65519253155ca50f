"""GPU: the small tree operators against the golden vectors of the unmodified
reference source, the reference's own unit-test assertions, and float64 finite differences."""

import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_golden.npz"))
TR = sorted({k.split("/")[1] for k in G.files if k.startswith("tree/")})


def _t(a):
  import torch

  return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize("name", TR)
def test_tree_operators_match_reference_golden(name):
  import torch
  from trex_b200 import tree as T

  g = lambda k: G[f"tree/{name}/{k}"]
  temp = float(g("temp"))
  S = T.update_seq({"ancestors": [_t(x) for x in g("logits")]}, _t(g("S0")), temp)
  assert np.abs(S.cpu().numpy() - g("S")).max() <= 1e-6
  S2 = T.update_seq(_t(g("logits")), _t(g("S0")), temp)  # stacked form (evals/benchmark.py:210-232)
  assert torch.equal(S, S2)
  adj = T.update_tree(None, {"tree_params": _t(g("tree_params"))}, 0.9, noise=_t(g("gumbel")))
  assert np.abs(adj.cpu().numpy() - g("adj_soft")).max() <= 1e-6
  assert np.allclose(adj.sum(dim=1).cpu().numpy(), 1.0, atol=1e-6)  # tests/test_tree.py:31
  N = adj.shape[0]
  for an in ("soft", "bern"):
    A_ = _t(g(f"adj_{an}"))
    c = float(T.enforce_graph_constraints(A_, 10.0))
    assert abs(c - float(g(f"constraint_{an}"))) <= 1e-5 * abs(float(g(f"constraint_{an}"))) + 1e-5
    assert np.array_equal(T.discretize_tree_topology(A_, N).cpu().numpy(), g(f"discrete_{an}"))
  cc = T.compute_cost(_t(g("onehot")), _t(g("adj_soft")), _t(g("Csym")))
  assert float(cc) == float(g("compute_cost"))
  params = {"ancestors": [_t(x) for x in g("logits")], "tree_params": _t(g("tree_params"))}
  loss = T.compute_loss(None, params, _t(g("S0")), {}, temp, _t(g("adj_bern")), noise=_t(g("loss_gumbel")))
  assert abs(float(loss) - float(g("loss"))) <= 1e-5 * abs(float(g("loss"))) + 2e-4
  lf = T.compute_loss(None, params, _t(g("S0")), {}, temp, _t(g("adj_bern")), fix_tree=True)
  assert abs(float(lf) - float(g("loss_fix_tree"))) <= 1e-5 * abs(float(g("loss_fix_tree"))) + 2e-4


def test_reference_tree_unit_tests_on_gpu():
  """The assertions of /root/reference/tests/test_tree.py, on the library."""
  import torch
  from trex_b200 import tree as T

  soft = torch.tensor([[0.1, 0.9, 0.0], [0.8, 0.1, 0.1], [0.3, 0.3, 0.4]], device="cuda")
  oh = T.discretize_tree_topology(soft, 3)
  assert oh.shape == (3, 3) and bool(((oh == 0) | (oh == 1)).all()) and bool((oh.sum(dim=1) == 1).all())
  out = T.update_tree(0, {"tree_params": torch.ones((2, 1), device="cuda")}, temperature=1.0)
  assert out.shape == (3, 3) and torch.allclose(out.sum(dim=1), torch.ones(3, device="cuda"))
  assert T.update_tree(0, {"tree_params": torch.ones((1, 0), device="cuda")}).shape == (2, 2)
  s = T.update_seq({"ancestors": torch.ones((2, 5, 4), device="cuda")}, torch.zeros((4, 5, 4), device="cuda"), 1.0)
  assert s.shape == (4, 5, 4)
  assert float(T.enforce_graph_constraints(torch.eye(5, device="cuda"), 10.0)) >= 0
  seqs = torch.nn.functional.one_hot(torch.tensor([[0, 1, 2], [3, 2, 1]]), 4).float().cuda()
  assert float(T.compute_surrogate_cost(seqs, torch.eye(2, device="cuda"))) >= 0
  seqs = torch.nn.functional.one_hot(torch.tensor([[0, 1], [1, 0], [0, 0]]), 2).float().cuda()
  assert float(T.compute_cost(seqs, torch.eye(3, device="cuda"), (torch.ones((2, 2)) - torch.eye(2)).cuda())) >= 0
  params = {"tree_params": torch.ones((3, 2), device="cuda"), "ancestors": torch.ones((2, 5, 4), device="cuda")}
  loss = T.compute_loss(0, params, torch.zeros((4, 5, 4), device="cuda"), {"n_all": 4}, 1.0, torch.eye(4, device="cuda"))
  assert not bool(torch.isnan(loss)) and float(loss) >= 0


def test_compute_loss_gradients_vs_float64_oracle():
  """jax.value_and_grad(compute_loss) of tests/test_convergence.py:240-253: gradients w.r.t. the
  ancestor logits and the tree parameters, against central differences of the oracle."""
  from oracle import tree_oracle as to
  from trex_b200 import tree as T

  g = lambda k: G[f"tree/t_small/{k}"]
  temp = float(g("temp"))
  logits = _t(g("logits")).requires_grad_(True)
  tp = _t(g("tree_params")).requires_grad_(True)
  noise = g("loss_gumbel")
  loss = T.compute_loss(None, {"ancestors": logits, "tree_params": tp}, _t(g("S0")), {}, temp, _t(g("adj_bern")),
                        noise=_t(noise))
  loss.backward()
  f = lambda lg, p: float(to.compute_loss(lg, p, noise.astype(np.float64), g("S0"), temp, g("adj_bern"), dtype=np.float64))
  rng = np.random.default_rng(0)
  h = 1e-5
  lg0, tp0 = g("logits").astype(np.float64), g("tree_params").astype(np.float64)
  for _ in range(10):
    i = tuple(rng.integers(0, d) for d in lg0.shape)
    a, b = lg0.copy(), lg0.copy()
    a[i] += h
    b[i] -= h
    fd = (f(a, tp0) - f(b, tp0)) / (2 * h)
    assert abs(float(logits.grad[i]) - fd) <= 1e-4 * (1 + abs(fd))
    j = tuple(rng.integers(0, d) for d in tp0.shape)
    a, b = tp0.copy(), tp0.copy()
    a[j] += h
    b[j] -= h
    fd = (f(lg0, a) - f(lg0, b)) / (2 * h)
    assert abs(float(tp.grad[j]) - fd) <= 1e-4 * (1 + abs(fd))
