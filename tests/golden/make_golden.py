#!/usr/bin/env python
"""Generates tests/golden/reference_golden.npz by running the UNMODIFIED reference modules
(/root/reference/src/trex/sankoff.py, tree.py and /root/reference/benchmark_tree_cost.py) under
the NumPy stand-in for jax (jax_numpy_shim.py; real JAX is not installable here).

Run in THIS container only (needs /root/reference):  python tests/golden/make_golden.py
The committed .npz is what the CPU and GPU tests read; /root/reference is never needed at test
time.  Inputs are explicit arrays from fixed NumPy seeds."""

import importlib.util
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

import jax.numpy as jnp  # noqa: E402  (the stand-in)
from trex import sankoff as rs  # noqa: E402  (reference, unmodified)
from trex import tree as rt  # noqa: E402

from trex_b200 import synth  # noqa: E402

spec = importlib.util.spec_from_file_location("benchmark_tree_cost", "/root/reference/benchmark_tree_cost.py")
btc = importlib.util.module_from_spec(spec)
spec.loader.exec_module(btc)

out = {}


def sankoff_case(name, adj, cost, seqs, n_all, A, n):
  rec, dp, tot = rs.run_sankoff(jnp.asarray(adj), jnp.asarray(cost), jnp.asarray(seqs), n_all, A, n, return_path=True)
  rec0, _, _ = rs.run_sankoff(jnp.asarray(adj), jnp.asarray(cost), jnp.asarray(seqs), n_all, A, n, return_path=False)
  out[f"sk/{name}/adj"] = np.asarray(adj, dtype=np.float32)
  out[f"sk/{name}/cost"] = np.asarray(cost, dtype=np.float32)
  out[f"sk/{name}/seqs"] = np.asarray(seqs, dtype=np.float32)
  out[f"sk/{name}/sizes"] = np.array([n_all, A, n], dtype=np.int64)
  out[f"sk/{name}/recon"] = np.asarray(rec, dtype=np.float32)
  out[f"sk/{name}/recon_nopath"] = np.asarray(rec0, dtype=np.float32)
  out[f"sk/{name}/dp"] = np.asarray(dp, dtype=np.float32)
  out[f"sk/{name}/total"] = np.float32(tot)
  # the vmapped back-pointer table of vectorized_dp (sankoff.py:97,151-160)
  L = np.asarray(seqs).shape[1]
  a = np.array(adj, dtype=np.float32)
  a[-1, -1] = 0
  dp0 = np.full((L, n_all, A), 1e5, dtype=np.float32)
  bt0 = np.zeros((L, n_all, A, 4), dtype=np.float32)
  _, bt = rs.vectorized_dp(jnp.asarray(a), jnp.asarray(dp0), jnp.asarray(bt0), jnp.asarray(seqs, dtype=np.float32),
                           jnp.asarray(cost, dtype=np.float32))
  out[f"sk/{name}/bt"] = np.asarray(bt, dtype=np.float32)


# 1. the reference's own test inputs (tests/test_sankoff.py:39-67)
adj = np.zeros((5, 5), dtype=np.int32)
adj[0, 3] = adj[1, 3] = adj[2, 4] = adj[3, 4] = 1
sankoff_case("ref_test", adj, np.ones((2, 2), np.float32) - np.eye(2, dtype=np.float32),
             np.array([[0, 1], [1, 0], [0, 0]], dtype=np.float32), 5, 2, 3)

# 2. BASELINE config c1: 16-leaf balanced tree, 100 sites, DNA, unit cost
rng = np.random.default_rng(1)
par = synth.balanced_parents(16)
st = synth.evolve(par, 100, 4, 0.1, rng)
sankoff_case("c1", synth.parents_to_adjacency(par), synth.unit_cost(4), st[:16].astype(np.float32), 31, 4, 16)

# 3. assorted topologies / alphabets / cost matrices, incl. ties, non-integer and asymmetric costs
specs = [("rand_a20_int", "random", 9, 20, 12, "int"), ("cat_a5_int", "cat", 7, 5, 9, "int"),
         ("gt_a4_unit", "gt", 8, 4, 20, "unit"), ("rand_a3_float", "random", 6, 3, 15, "float"),
         ("rand_a24_unit", "random", 5, 24, 6, "unit")]
for name, kind, n, A, L, ck in specs:
  rng = np.random.default_rng(abs(hash(name)) % (2 ** 31) if False else sum(map(ord, name)))
  par = {"cat": synth.caterpillar_parents, "gt": synth.groundtruth_parents}.get(kind, lambda k: synth.random_join_parents(k, rng))(n)
  st = synth.evolve(par, L, A, 0.3, rng)
  cost = synth.unit_cost(A) if ck == "unit" else (synth.random_int_cost(A, 5, rng) if ck == "int"
                                                  else (rng.random((A, A)) * 3).astype(np.float32))
  sankoff_case(name, synth.parents_to_adjacency(par), cost, st[:n].astype(np.float32), 2 * n - 1, A, n)

# 4. out-of-range / negative leaf states (dropped scatter, wrap) — sankoff.py:50
rng = np.random.default_rng(44)
par = synth.random_join_parents(6, rng)
seqs = synth.evolve(par, 14, 4, 0.3, rng)[:6].astype(np.float32)
seqs[1, 3] = 7.0
seqs[4, 9] = -1.0
seqs[0, 0] = 4.0
sankoff_case("oob_states", synth.parents_to_adjacency(par), synth.random_int_cost(4, 3, rng), seqs, 11, 4, 6)

# 5. run_dp on the malformed adjacency of tests/test_sankoff.py:9-36 (fill_value=-1 path)
adj = np.array([[0, 1, 0], [0, 1, 0], [0, 0, 0]], dtype=np.float32)
dp, bt = rs.run_dp(jnp.asarray(adj), jnp.full((3, 2), 1e5, dtype=jnp.float32), jnp.zeros((3, 2, 4), dtype=jnp.float32),
                   jnp.asarray(np.array([0, 1, 0], dtype=np.float32)), jnp.asarray(np.array([[0, 1], [1, 0]], dtype=np.float32)))
out["run_dp/malformed/adj"] = adj
out["run_dp/malformed/dp"] = np.asarray(dp, dtype=np.float32)
out["run_dp/malformed/bt"] = np.asarray(bt, dtype=np.float32)

# 6. tree.py
rng = np.random.default_rng(7)
for name, N, L, Q in [("t_small", 7, 5, 4), ("t_mid", 15, 12, 20), ("t_c2ish", 31, 40, 4)]:
  n = (N + 1) // 2
  logits = rng.normal(size=(N - n, L, Q)).astype(np.float32)
  S0 = np.zeros((N, L, Q), dtype=np.float32)
  S0[:n] = np.eye(Q, dtype=np.float32)[rng.integers(0, Q, size=(n, L))]
  temp = np.float32(0.7)
  S = np.asarray(rt.update_seq({"ancestors": [jnp.asarray(x) for x in logits]}, jnp.asarray(S0), temp), dtype=np.float32)
  tp = rng.normal(size=(N - 1, N - n)).astype(np.float32)
  adj_soft = np.asarray(rt.update_tree(None, {"tree_params": jnp.asarray(tp)}, 0.9), dtype=np.float32)
  noise = shim.LAST_GUMBEL["noise"].copy()
  adj_b = (rng.random((N, N)) < 0.2).astype(np.float32)
  Csym = synth.random_int_cost(Q, 3, rng)
  Casym = rng.random((Q, Q)).astype(np.float32)
  cvec = (rng.random(Q) + 0.5).astype(np.float32)
  out[f"tree/{name}/logits"], out[f"tree/{name}/S0"], out[f"tree/{name}/temp"] = logits, S0, temp
  out[f"tree/{name}/S"], out[f"tree/{name}/tree_params"] = S, tp
  out[f"tree/{name}/gumbel"], out[f"tree/{name}/adj_soft"], out[f"tree/{name}/adj_bern"] = noise, adj_soft, adj_b
  out[f"tree/{name}/Csym"], out[f"tree/{name}/Casym"], out[f"tree/{name}/cvec"] = Csym, Casym, cvec
  for an, A_ in [("soft", adj_soft), ("bern", adj_b)]:
    out[f"tree/{name}/surrogate_{an}"] = np.float32(rt.compute_surrogate_cost(jnp.asarray(S), jnp.asarray(A_)))
    out[f"tree/{name}/soft_none_{an}"] = np.float32(rt.compute_soft_cost(jnp.asarray(S), jnp.asarray(A_)))
    out[f"tree/{name}/soft_sym_{an}"] = np.float32(rt.compute_soft_cost(jnp.asarray(S), jnp.asarray(A_), jnp.asarray(Csym)))
    out[f"tree/{name}/soft_asym_{an}"] = np.float32(rt.compute_soft_cost(jnp.asarray(S), jnp.asarray(A_), jnp.asarray(Casym)))
    out[f"tree/{name}/soft_vec_{an}"] = np.float32(rt.compute_soft_cost(jnp.asarray(S), jnp.asarray(A_), jnp.asarray(cvec)))
    out[f"tree/{name}/naive_none_{an}"] = np.float32(btc.naive_compute_soft_cost(jnp.asarray(S), jnp.asarray(A_)))
    out[f"tree/{name}/naive_sym_{an}"] = np.float32(btc.naive_compute_soft_cost(jnp.asarray(S), jnp.asarray(A_), jnp.asarray(Csym)))
    out[f"tree/{name}/constraint_{an}"] = np.float32(rt.enforce_graph_constraints(jnp.asarray(A_), 10.0))
    out[f"tree/{name}/discrete_{an}"] = np.asarray(rt.discretize_tree_topology(jnp.asarray(A_), N), dtype=np.float32)
  onehot = np.eye(Q, dtype=np.float32)[rng.integers(0, Q, size=(N, L))]
  out[f"tree/{name}/onehot"] = onehot
  out[f"tree/{name}/compute_cost"] = np.float32(rt.compute_cost(jnp.asarray(onehot), jnp.asarray(adj_soft), jnp.asarray(Csym)))
  key = np.array([0, 1], dtype=np.uint32)
  loss = rt.compute_loss(key, {"ancestors": [jnp.asarray(x) for x in logits], "tree_params": jnp.asarray(tp)},
                         jnp.asarray(S0), {}, temp, jnp.asarray(adj_b))
  out[f"tree/{name}/loss"] = np.float32(loss)
  out[f"tree/{name}/loss_gumbel"] = shim.LAST_GUMBEL["noise"].copy()
  out[f"tree/{name}/loss_fix_tree"] = np.float32(rt.compute_loss(
    key, {"ancestors": [jnp.asarray(x) for x in logits], "tree_params": jnp.asarray(tp)}, jnp.asarray(S0), {}, temp,
    jnp.asarray(adj_b), fix_tree=True))

np.savez_compressed(os.path.join(HERE, "reference_golden.npz"), **out)
print("wrote", len(out), "arrays,", os.path.getsize(os.path.join(HERE, "reference_golden.npz")), "bytes")
