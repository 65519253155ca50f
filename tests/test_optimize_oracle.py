"""CPU: the restated ancestor-fitting loop (oracle/optimize_oracle.py; evals/benchmark.py:41-200).
The optimiser arithmetic is parity-unpinned (optax absent); what can be pinned here is pinned: the
loss is the golden-checked tree-cost restatement, its gradient is checked by central differences,
and the loop must do what the reference's convergence tests expect of it (loss goes down)."""

import numpy as np
import pytest

from oracle import optimize_oracle as oo
from oracle import tree_oracle
from trex_b200 import synth


def _problem(n, L, Q, seed, soft_adj=False):
  rng = np.random.default_rng(seed)
  parent = synth.balanced_parents(n)
  N = parent.shape[0]
  states = synth.evolve(parent, L, Q, 0.2, rng)
  adj = synth.parents_to_adjacency(parent).astype(np.float32)
  if soft_adj:
    adj = tree_oracle.softmax(rng.normal(size=(N, N)), axis=1).astype(np.float32)
  init = rng.normal(size=(N - n, L, Q)).astype(np.float32)
  return states[:n], N, adj, init


def test_loss_gradient_by_central_differences():
  leaf, N, adj, init = _problem(4, 3, 3, 0, soft_adj=True)
  n = 4
  masked = np.concatenate([np.eye(3)[leaf.astype(int)], np.zeros((N - n, 3, 3))], axis=0)
  _, g = oo.loss_and_grad(init, masked, n, adj)
  h = 1e-6
  for idx in [(0, 0, 0), (1, 2, 1), (2, 1, 2)]:
    xp, xm = init.astype(np.float64).copy(), init.astype(np.float64).copy()
    xp[idx] += h
    xm[idx] -= h
    fd = (oo.loss_and_grad(xp, masked, n, adj)[0] - oo.loss_and_grad(xm, masked, n, adj)[0]) / (2 * h)
    assert abs(fd - g[idx]) <= 1e-6 * max(1.0, abs(fd))


def test_loss_is_the_reference_surrogate_cost():
  leaf, N, adj, init = _problem(8, 5, 4, 1)
  masked = np.concatenate([np.eye(4)[leaf.astype(int)], np.zeros((N - 8, 5, 4))], axis=0)
  S = tree_oracle.update_seq(init.astype(np.float64), masked, 1.0)
  ref = tree_oracle.compute_surrogate_cost(S, adj, dtype=np.float64)
  assert abs(oo.loss_and_grad(init, masked, 8, adj)[0] - ref) <= 1e-9 * abs(ref)


@pytest.mark.parametrize("name", oo.OPTIMIZERS)
def test_every_optimizer_reduces_the_loss(name):
  leaf, N, adj, init = _problem(8, 6, 4, 2)
  _, losses, x = oo.run_trex_optimization_configurable(leaf, N, 8, 4, adj, init, name, 0.05, 40)
  assert losses[-1] < losses[0]
  assert np.isfinite(x).all()


def test_first_adam_step_is_lr_times_sign_of_clipped_gradient():
  # bias-corrected first step: u = g / (|g| + eps)
  leaf, N, adj, init = _problem(4, 3, 3, 3)
  masked = np.concatenate([np.eye(3)[leaf.astype(int)], np.zeros((N - 4, 3, 3))], axis=0)
  _, g = oo.loss_and_grad(init, masked, 4, adj)
  _, _, x = oo.run_trex_optimization_configurable(leaf, N, 4, 3, adj, init, "adam", 1e-2, 1)
  big = np.abs(g) > 1e-4
  assert np.allclose((init - x)[big], 1e-2 * np.sign(g)[big], rtol=1e-3)
