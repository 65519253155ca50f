#!/usr/bin/env python
"""Generates tests/golden/generators_golden.npz by running the UNMODIFIED reference generators
(/root/reference/src/trex/ground_truth.py, nk_model.py) under the NumPy stand-in for jax (jax_numpy_shim.py), whose
jax.random records every draw in call order.  The recorded draws are stored with the reference's outputs: the oracle
(oracle/generators_oracle.py) and the CUDA generators consume the same draws and must reproduce the outputs.

Run in THIS container only (needs /root/reference):  python tests/golden/make_golden_generators.py"""

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import jax_numpy_shim as shim  # noqa: E402

shim.install()
sys.path.insert(0, "/root/reference/src")

import jax  # noqa: E402  (the stand-in)
import jax.numpy as jnp  # noqa: E402
from trex import ground_truth as gt  # noqa: E402  (reference, unmodified)
from trex import nk_model as nk  # noqa: E402

from oracle import generators_oracle as go  # noqa: E402

out = {}

# ---- generate_groundtruth: the log is [choice, randint] per mutate call, children 2 p_i, 2 p_i + 1 of the parents
# n_all-1, n_all-2, ... (ground_truth.py:155-170; n_mutations = 0 logs no choice, :39) -------------------------------
for name, (n_leaves, A, m, L, seed) in {"gt_a": (8, 4, 3, 40, 1), "gt_b": (16, 20, 5, 33, 2), "gt_c": (4, 2, 0, 9, 3),
                                        "gt_d": (2, 5, 7, 7, 4)}.items():
  shim.DRAWS.clear()
  t = gt.generate_groundtruth(n_leaves, A, m, L, seed=seed)
  n_all = 2 * n_leaves - 1
  log = list(shim.DRAWS)
  sites = np.zeros((n_all - 1, m), dtype=np.int32)
  offs = np.zeros((n_all - 1, L), dtype=np.uint8)
  pos = 0
  for i in range(n_leaves - 1):
    p_i = (n_all - 1 - i) - n_leaves
    for c in (2 * p_i, 2 * p_i + 1):
      if m > 0:
        assert log[pos][0] == "choice"
        sites[c] = log[pos][1]
        pos += 1
      assert log[pos][0] == "randint"
      offs[c] = log[pos][1]
      pos += 1
  assert pos == len(log)
  out[name + "_cfg"] = np.array([n_leaves, A, m, L])
  out[name + "_sites"], out[name + "_offsets"] = sites, offs
  out[name + "_all"] = np.asarray(t.all_sequences, dtype=np.float32)
  out[name + "_masked"] = np.asarray(t.masked_sequences, dtype=np.float32)
  out[name + "_adj"] = np.asarray(t.adjacency, dtype=np.float32)

# ---- NK landscape + get_fitness + generate_tree_data -------------------------------------------------------------------
rng = np.random.default_rng(7)
for name, (n_leaves, L, k, q, B, rate, std, selfloop) in {
    "nk_a": (4, 12, 2, 4, 2, 0.3, 0.2, True), "nk_b": (8, 30, 1, 20, 1, 0.1, 0.0, True),
    "nk_c": (4, 10, 0, 2, 3, 0.5, 0.5, False), "nk_d": (16, 20, 2, 4, 1, 0.05, 0.0, False)}.items():
  shim.DRAWS.clear()
  land = nk.create_nk_model_landscape(L, k, jax.random.PRNGKey(1), n_states=q)
  inter = np.asarray(land["interactions"], dtype=np.int32)
  tables = np.asarray(land["fitness_tables"], dtype=np.float32)
  adj = go.create_balanced_binary_tree(n_leaves)
  if selfloop:
    adj[-1, -1] = 1
  root = rng.integers(0, q, size=(L, 1))
  seqs_for_fitness = rng.integers(0, q, size=(5, L))
  fit = np.array([float(nk.get_fitness(jnp.asarray(s), land)) for s in seqs_for_fitness], dtype=np.float32)
  mask = (rng.random(L) < 0.7)
  fit_masked = np.array([float(nk.get_fitness(jnp.asarray(s), land, jnp.asarray(mask))) for s in seqs_for_fitness], dtype=np.float32)
  shim.DRAWS.clear()
  t = nk.generate_tree_data(land, jnp.asarray(adj), jnp.asarray(root), rate, jax.random.PRNGKey(2),
                            coupled_mutation_prob=0.5, mutation_rate_noise_std=std, branch_length=B)
  log = list(shim.DRAWS)
  # store the log itself (call order): kinds as small ints, values flattened with offsets
  kinds = {"normal": 0, "bernoulli_uniform": 1, "randint": 2}
  out[name + "_log_kind"] = np.array([kinds[a] for a, _ in log], dtype=np.int8)
  out[name + "_log_size"] = np.array([np.size(b) for _, b in log], dtype=np.int64)
  out[name + "_log_vals"] = np.concatenate([np.asarray(b, dtype=np.float64).reshape(-1) for _, b in log]) if log else np.zeros(0)
  out[name + "_cfg"] = np.array([n_leaves, L, k, q, B, int(selfloop)])
  out[name + "_rate_std"] = np.array([rate, std], dtype=np.float64)
  out[name + "_inter"], out[name + "_tables"], out[name + "_adj"], out[name + "_root"] = inter, tables, adj, root[:, 0].astype(np.uint8)
  out[name + "_fit_seqs"], out[name + "_fit"], out[name + "_fit_mask"], out[name + "_fit_masked"] = \
      seqs_for_fitness.astype(np.uint8), fit, mask.astype(np.float32), fit_masked
  out[name + "_all"] = np.asarray(t.all_sequences, dtype=np.float32)

path = os.path.join(HERE, "generators_golden.npz")
np.savez_compressed(path, **out)
print("wrote", path, len(out), "arrays", os.path.getsize(path), "bytes")
