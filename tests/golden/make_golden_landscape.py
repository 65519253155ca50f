#!/usr/bin/env python
"""Generates tests/golden/landscape_golden.npz by running the UNMODIFIED reference functions
``compute_parental_logits`` and ``_compute_loss_landscape_aware_stacked``
(/root/reference/src/trex/evals/benchmark.py:586-663, 235-306) under the NumPy stand-in for jax
(jax_numpy_shim.py; real JAX and optax are not installable here — optax is only touched by
``create_optimizer``, which is not called).

Run in THIS container only (needs /root/reference):  python tests/golden/make_golden_landscape.py"""

import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import jax_numpy_shim as shim  # noqa: E402

shim.install()
sys.modules.setdefault("optax", types.ModuleType("optax"))
sys.path.insert(0, "/root/reference/src")

import jax.numpy as jnp  # noqa: E402  (the stand-in)
from trex.evals import benchmark as rb  # noqa: E402  (reference, unmodified)

from trex_b200 import synth  # noqa: E402

out = {}
# (name, n_leaves, L, Q, k, masked?)
CASES = [("k1_q4", 8, 12, 4, 1, False), ("k2_q3", 4, 9, 3, 2, True), ("k1_q20", 8, 7, 20, 1, True),
         ("k3_q3", 4, 6, 3, 3, False), ("k0_q4", 4, 5, 4, 0, False)]
for name, n, L, Q, k, masked in CASES:
  rng = np.random.default_rng(sum(map(ord, name)))
  parent = synth.balanced_parents(n)
  N = parent.shape[0]
  adj = synth.parents_to_adjacency(parent).astype(np.float32)
  leaf = synth.evolve(parent, L, Q, 0.2, rng)[:n]
  anc = rng.normal(size=(N - n, L, Q)).astype(np.float32)
  masked_seqs = np.concatenate([np.eye(Q, dtype=np.float32)[leaf.astype(int)], np.zeros((N - n, L, Q), np.float32)], 0)
  inter = rng.integers(0, L, size=(L, max(k, 1))).astype(np.int32)[:, :k] if k else np.zeros((L, 0), np.int32)
  tables = rng.random((L, Q ** (k + 1))).astype(np.float32)
  seq_mask = (rng.random(L) < 0.7) if masked else None
  land = {"interactions": jnp.asarray(inter), "fitness_tables": jnp.asarray(tables)}
  S = np.array(masked_seqs)
  e = np.exp(anc - anc.max(-1, keepdims=True))
  S[n:] = e / e.sum(-1, keepdims=True)
  logits = np.asarray(rb.compute_parental_logits(jnp.asarray(S), land, k))
  lam = 0.5
  loss = rb._compute_loss_landscape_aware_stacked(jnp.asarray(anc), jnp.asarray(masked_seqs), n, land, jnp.asarray(adj), N,
                                                  lam, k, 1.0, None if seq_mask is None else jnp.asarray(seq_mask))
  loss0 = rb._compute_loss_landscape_aware_stacked(jnp.asarray(anc), jnp.asarray(masked_seqs), n, land, jnp.asarray(adj),
                                                   N, 0.0, k, 1.0, None)
  for key, val in [("adj", adj), ("anc", anc), ("masked", masked_seqs), ("inter", inter), ("tables", tables),
                   ("mask", np.ones(L, np.float32) if seq_mask is None else seq_mask.astype(np.float32)),
                   ("has_mask", np.int32(seq_mask is not None)), ("sizes", np.array([n, L, Q, k], np.int64)),
                   ("logits_of_S", logits.astype(np.float32)), ("loss", np.float32(loss)), ("loss_lambda0", np.float32(loss0)),
                   ("lambda", np.float32(lam))]:
    out[f"ls/{name}/{key}"] = val
np.savez_compressed(os.path.join(HERE, "landscape_golden.npz"), **out)
print("wrote", len(out), "arrays")
