"""Host-side mirror of the ancestor-fitting driver of the reference's benchmark harness
(``/root/reference/src/trex/evals/benchmark.py``: ``create_optimizer`` :41-72,
``run_trex_optimization_configurable`` :75-200) on top of ``trex_fit_ancestors``
(include/trex_b200.h) — SURVEY §8(f) rank 1.

The whole ``fori_loop`` / ``scan`` of the reference (softmax, tree cost value + gradient, global-norm
clip, optimiser update; ``n_iterations`` times) is ONE library call = one persistent cooperative
kernel.  Arrays are torch CUDA tensors; there is no CPU fallback.
"""

from __future__ import annotations

from dataclasses import dataclass

import torch

from . import _lib
from ._lib import call, ptr, stream_ptr

_OPT_CODES = {"adam": 0, "sgd": 1, "rmsprop": 2, "adamw": 3}


@dataclass(frozen=True)
class OptimizerSpec:
  """What ``create_optimizer`` returns here: the name/learning-rate/clip triple the kernel needs
  (the reference returns an ``optax.GradientTransformation``)."""
  name: str
  learning_rate: float
  clip_norm: float  # <= 0: no clipping


def create_optimizer(name, learning_rate, *, use_gradient_clipping=True):
  """evals/benchmark.py:41-72 — same names, same ``ValueError`` for an unknown optimizer."""
  if name not in _OPT_CODES:
    raise ValueError(f"Unknown optimizer: {name}. Choose from {list(_OPT_CODES.keys())}")
  return OptimizerSpec(name, float(learning_rate), 1.0 if use_gradient_clipping else 0.0)


class AncestorFit:
  """Resident state of one fit: logits and optimiser moments stay in HBM between calls, so the
  loop can be continued (``run(n)`` twice == ``run(2n)``)."""

  def __init__(self, leaf_sequences, n_all, n_leaves, n_states, adj_matrix, init_ancestors,
               optimizer: OptimizerSpec):
    if not torch.cuda.is_available():
      raise RuntimeError("trex_b200 needs a CUDA device; there is no CPU fallback")
    lib = _lib.load()
    dev = torch.device("cuda", torch.cuda.current_device())
    leaf = torch.as_tensor(leaf_sequences).to(dev)
    self.n_all, self.n_leaves, self.Q = int(n_all), int(n_leaves), int(n_states)
    self.n_anc = self.n_all - self.n_leaves
    self.L = int(leaf.shape[1])
    # jax.nn.one_hot (:128): out-of-range states give an all-zero row
    idx = leaf[: self.n_leaves].long()
    ok = (idx >= 0) & (idx < self.Q)
    self.leaf_seqs = (torch.nn.functional.one_hot(idx.clamp(0, self.Q - 1), self.Q).float()
                      * ok.unsqueeze(-1)).contiguous()
    self.adj = torch.as_tensor(adj_matrix).to(device=dev, dtype=torch.float32).contiguous()
    assert self.adj.shape == (self.n_all, self.n_all), self.adj.shape
    self.logits = torch.as_tensor(init_ancestors).to(device=dev, dtype=torch.float32).clone().contiguous()
    assert self.logits.shape == (self.n_anc, self.L, self.Q), self.logits.shape
    self.mom1 = torch.zeros_like(self.logits)
    self.mom2 = torch.zeros_like(self.logits)
    self.opt = optimizer
    self.steps = 0
    nbytes = lib.trex_fit_ancestors_workspace_bytes(self.n_all, self.n_leaves, self.L, self.Q)
    self.workspace = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    self.argmax = torch.empty((self.n_anc, self.L), dtype=torch.int32, device=dev)

  def run(self, n_iterations, losses=None):
    """``n_iterations`` more updates; ``losses``: optional (n_iterations,) f32 CUDA tensor to fill."""
    call("trex_fit_ancestors", ptr(self.adj), ptr(self.leaf_seqs), self.n_all, self.n_leaves, self.L,
         self.Q, ptr(self.logits), ptr(self.mom1), ptr(self.mom2), _OPT_CODES[self.opt.name],
         self.opt.learning_rate, self.opt.clip_norm, int(n_iterations), int(self.steps), ptr(losses),
         ptr(self.argmax), ptr(self.workspace), stream_ptr())
    self.steps += int(n_iterations)
    return self.argmax


def run_trex_optimization_configurable(leaf_sequences, n_all, n_leaves, n_states, adj_matrix, key,
                                       use_soft_cost=False, optimizer_name="adam", learning_rate=1e-3,
                                       n_iterations=10000, return_losses=False, dtype=None,
                                       mixed_precision=False, *, init_ancestors=None):
  """evals/benchmark.py:75-200.  ``key``: an int seed or ``torch.Generator`` for the normal
  initial logits (JAX's threefry stream is not reproducible outside JAX — pass
  ``init_ancestors`` (n_anc, L, Q) to start from given logits instead).  ``use_soft_cost`` picks
  between two spellings of the same function (identity cost matrix, :149-165).  Compute is float32
  (``dtype`` / ``mixed_precision`` other than float32 are not in any benchmark config).

  Returns ``argmax (n_anc, L) int32`` or ``(argmax, losses (n_iterations,) f32)``."""
  if dtype not in (None, torch.float32) or mixed_precision:
    raise NotImplementedError("trex_b200 fits ancestors in float32 only")
  del use_soft_cost
  leaf = torch.as_tensor(leaf_sequences)
  n_anc, L = int(n_all) - int(n_leaves), int(leaf.shape[1])
  if init_ancestors is None:
    gen = key if isinstance(key, torch.Generator) else torch.Generator(device="cuda").manual_seed(int(key))
    init_ancestors = torch.randn((n_anc, L, int(n_states)), generator=gen, device="cuda",
                                 dtype=torch.float32)
  fit = AncestorFit(leaf, n_all, n_leaves, n_states, adj_matrix, init_ancestors,
                    create_optimizer(optimizer_name, learning_rate))
  losses = torch.empty((int(n_iterations),), dtype=torch.float32, device="cuda") if return_losses else None
  out = fit.run(n_iterations, losses)
  return (out, losses) if return_losses else out


def run_trex_optimization_batched(leaf_sequences, n_all, n_leaves, n_states, adj_matrix, key,
                                  use_soft_cost=False, n_iterations=10000, *, init_ancestors=None):
  """evals/benchmark.py:459-542 — the vmap-friendly twin of the loop above with its fixed optimiser
  ``optax.chain(clip_by_global_norm(1.0), adam(1e-3))`` (:496).  Returns ``argmax (n_anc, L) int32``.
  (The reference vmaps it over candidate trees; here each tree is one persistent kernel launch.)"""
  return run_trex_optimization_configurable(leaf_sequences, n_all, n_leaves, n_states, adj_matrix, key,
                                            use_soft_cost=use_soft_cost, optimizer_name="adam", learning_rate=1e-3,
                                            n_iterations=n_iterations, init_ancestors=init_ancestors)
