"""TEST INFRASTRUCTURE ONLY — NumPy restatement of the ancestor-fitting loop of
``/root/reference/src/trex/evals/benchmark.py`` (``create_optimizer`` :41-72,
``run_trex_optimization_configurable`` :75-200, ``_update_seq_stacked`` :210-232).

**Parity unpinned for the optimiser arithmetic**: the update rules live in optax (pinned
``optax==0.2.6`` in ``uv.lock:1728-1729``), which is absent from ``/root/reference`` and from
this image, as is JAX.  The transformations are restated from optax's published definitions:

* ``clip_by_global_norm(max_norm)``: ``g_norm = sqrt(sum g^2)``; ``g`` if ``g_norm < max_norm`` else
  ``(g / g_norm) * max_norm``.
* ``adam(lr)`` = ``scale_by_adam(b1=0.9, b2=0.999, eps=1e-8, eps_root=0)`` then ``* -lr``:
  ``mu = (1-b1) g + b1 mu``, ``nu = (1-b2) g^2 + b2 nu``, ``mu_hat = mu / (1 - b1^t)``,
  ``nu_hat = nu / (1 - b2^t)``, ``u = mu_hat / (sqrt(nu_hat + eps_root) + eps)``.
* ``adamw(lr, weight_decay=0.01)``: ``u += weight_decay * params`` before ``* -lr``.
* ``sgd(lr, momentum=0.9)``: ``trace = g + 0.9 trace``; ``u = trace``.
* ``rmsprop(lr)``: ``nu = 0.9 nu + 0.1 g^2``; ``u = g / sqrt(nu + 1e-8)`` (``eps_in_sqrt=True``).

The loss and its gradient are the restatements of ``tree_oracle`` (pinned by the golden vectors
of the unmodified reference source, tests/golden).  ``jax.random.normal`` cannot be reproduced
without JAX, so the initial logits are an explicit input.
"""

from __future__ import annotations

import numpy as np

from . import tree_oracle

OPTIMIZERS = ("adam", "sgd", "rmsprop", "adamw")


def loss_and_grad(ancestors, masked_sequences, n_leaves, adjacency, dtype=np.float64):
  """``jax.value_and_grad(loss_fn)(ancestors)`` of evals/benchmark.py:152-173 with the closed-form
  gradient of ``tree_oracle.soft_cost_value_and_grad`` chained through the softmax."""
  anc = np.asarray(ancestors, dtype=dtype)
  S = np.array(masked_sequences, dtype=dtype)
  P = tree_oracle.softmax(anc * 1.0, axis=-1)  # _update_seq_stacked, temperature 1.0 (:231)
  S[n_leaves:] = P
  cost, dS, _, _ = tree_oracle.soft_cost_value_and_grad(S, adjacency, None, dtype=dtype)
  dP = dS[n_leaves:]
  g = P * (dP - (P * dP).sum(axis=-1, keepdims=True))
  return cost, g


def run_trex_optimization_configurable(leaf_sequences, n_all, n_leaves, n_states, adj_matrix,
                                       init_ancestors, optimizer_name="adam", learning_rate=1e-3,
                                       n_iterations=100, use_gradient_clipping=True, dtype=np.float64):
  """Returns ``(argmax (n_anc, L) int32, losses (n_iter,), ancestors (n_anc, L, Q))``."""
  assert optimizer_name in OPTIMIZERS
  leaf = np.asarray(leaf_sequences).astype(np.int64)
  L = leaf.shape[1]
  n_anc = n_all - n_leaves
  one_hot = np.eye(n_states, dtype=dtype)[leaf]  # jax.nn.one_hot (:128)
  masked = np.concatenate([one_hot, np.zeros((n_anc, L, n_states), dtype=dtype)], axis=0)  # :132-138
  x = np.array(init_ancestors, dtype=dtype)
  mu = np.zeros_like(x)
  nu = np.zeros_like(x)
  losses = np.zeros(n_iterations, dtype=dtype)
  b1, b2, eps = dtype(0.9), dtype(0.999), dtype(1e-8)
  lr = dtype(learning_rate)
  for it in range(n_iterations):
    loss, g = loss_and_grad(x, masked, n_leaves, adj_matrix, dtype=dtype)
    losses[it] = loss
    if use_gradient_clipping:
      g_norm = np.sqrt((g * g).sum())
      if not g_norm < 1.0:
        g = (g / g_norm) * dtype(1.0)
    t = it + 1
    if optimizer_name == "sgd":
      mu = g + dtype(0.9) * mu
      u = mu
    elif optimizer_name == "rmsprop":
      nu = dtype(0.9) * nu + (1 - dtype(0.9)) * g * g
      u = g / np.sqrt(nu + eps)
    else:
      mu = (1 - b1) * g + b1 * mu
      nu = (1 - b2) * (g * g) + b2 * nu
      u = (mu / (1 - b1 ** t)) / (np.sqrt(nu / (1 - b2 ** t)) + eps)
      if optimizer_name == "adamw":
        u = u + dtype(0.01) * x
    x = x - lr * u
  return np.argmax(x, axis=-1).astype(np.int32), losses, x
