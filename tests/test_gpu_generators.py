"""GPU parity tests of the on-device input generators, through the C ABI.

Explicit-draw mode: the draws the UNMODIFIED reference consumed (tests/golden/generators_golden.npz) or fresh NumPy
draws go to the CUDA generator and to the oracle; results must be identical (integer states: bit-exact; fitness:
float32 round-off).  Philox mode (no draws given): the structural properties the reference's own tests check
(tests/test_ground_truth.py, tests/test_nk_model_new.py: shapes, mutation counts, determinism, divergence grows with
branch length)."""

import os

import numpy as np
import pytest

from oracle import generators_oracle as go

pytestmark = pytest.mark.gpu

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "generators_golden.npz"))


@pytest.mark.parametrize("name", ["gt_a", "gt_b", "gt_c", "gt_d"])
def test_generate_groundtruth_with_the_reference_draws_equals_the_reference(name):
  from trex_b200 import generators as G

  n_leaves, A, m, L = (int(x) for x in GOLD[name + "_cfg"])
  t = G.generate_groundtruth(n_leaves, A, m, L, seed=0, mut_sites=GOLD[name + "_sites"], offsets=GOLD[name + "_offsets"])
  assert np.array_equal(t.all_sequences.cpu().numpy(), GOLD[name + "_all"])
  assert np.array_equal(t.masked_sequences.cpu().numpy(), GOLD[name + "_masked"])
  assert np.array_equal(t.adjacency.cpu().numpy(), GOLD[name + "_adj"])
  assert t.all_sequences.dtype.is_floating_point and t.states_u8.dtype.itemsize == 1


@pytest.mark.parametrize("n_leaves,A,m,L", [(64, 20, 17, 1000), (1024, 20, 50, 4096), (8, 2, 1, 5), (16, 4, 0, 64)])
def test_generate_groundtruth_philox_mode_properties(n_leaves, A, m, L):
  from trex_b200 import generators as G
  from trex_b200 import synth

  t = G.generate_groundtruth(n_leaves, A, m, L, seed=7)
  s = t.all_sequences.cpu().numpy()
  n_all = 2 * n_leaves - 1
  assert s.shape == (n_all, L) and s.min() >= 0 and s.max() < A
  assert np.all(s[-1] == 0)                                              # root sequence = zeros (ground_truth.py:146)
  for c in range(n_all - 1):                                             # exactly n_mutations distinct sites per edge
    assert int((s[c] != s[n_leaves + c // 2]).sum()) == m
  assert np.array_equal(t.adjacency.cpu().numpy(), synth.parents_to_adjacency(synth.groundtruth_parents(n_leaves)))
  again = G.generate_groundtruth(n_leaves, A, m, L, seed=7).all_sequences.cpu().numpy()
  other = G.generate_groundtruth(n_leaves, A, m, L, seed=8).all_sequences.cpu().numpy()
  assert np.array_equal(s, again) and (m == 0 or not np.array_equal(s, other))
  if m >= 17:  # the mutated sites of an edge look uniform: no site bucket of an eighth of the alignment is starved
    hits = np.stack([s[c] != s[n_leaves + c // 2] for c in range(n_all - 1)]).sum(axis=0)
    eighths = hits[: L // 8 * 8].reshape(8, -1).sum(axis=1)
    assert eighths.min() > 0.6 * eighths.mean()


def test_generate_groundtruth_rejects_non_powers_of_two_and_balanced_tree_matches():
  import torch
  from trex_b200 import generators as G
  from trex_b200._lib import TrexError, call, ptr, stream_ptr

  with pytest.raises(ValueError, match="power of 2"):
    G.generate_groundtruth(6, 4, 1, 10)
  buf = torch.zeros(4096, dtype=torch.uint8, device="cuda")
  with pytest.raises(TrexError, match="power of 2"):                      # the C entry point itself refuses as well
    call("trex_generate_groundtruth", 6, 4, 1, 10, 16, 0, None, None, ptr(buf), None, ptr(buf), stream_ptr())
  for n in (2, 16, 1024):
    assert np.array_equal(G.create_balanced_binary_tree(n).cpu().numpy(), go.create_balanced_binary_tree(n))


@pytest.mark.parametrize("name", ["nk_a", "nk_b", "nk_c", "nk_d"])
def test_nk_fitness_equals_the_reference(name):
  import torch
  from trex_b200 import generators as G

  n_leaves, L, k, q, B, _ = (int(x) for x in GOLD[name + "_cfg"])
  land = {"interactions": torch.from_numpy(GOLD[name + "_inter"]).cuda(),
          "fitness_tables": torch.from_numpy(GOLD[name + "_tables"]).cuda(), "n_states": q, "k": k}
  seqs = GOLD[name + "_fit_seqs"]
  got = G.batched_get_fitness(seqs, land).cpu().numpy()
  assert np.abs(got - GOLD[name + "_fit"]).max() <= 2e-7                  # f32 results of an f64 sum vs jnp's f32 sum
  gotm = G.batched_get_fitness(seqs, land, GOLD[name + "_fit_mask"]).cpu().numpy()
  assert np.abs(gotm - GOLD[name + "_fit_masked"]).max() <= 2e-7
  assert abs(float(G.get_fitness(seqs[0], land)) - float(GOLD[name + "_fit"][0])) <= 2e-7


def _nk_case(n_leaves, L, k, q, B, seed, tree="balanced"):
  from trex_b200 import synth

  rng = np.random.default_rng(seed)
  if tree == "balanced":
    adj = go.create_balanced_binary_tree(n_leaves)
  else:
    adj = synth.parents_to_adjacency(synth.random_join_parents(n_leaves, rng))
  N = adj.shape[0]
  inter = rng.integers(0, L, size=(L, k)).astype(np.int32)
  tables = rng.random((L, q ** (k + 1))).astype(np.float32)
  root = rng.integers(0, q, size=L).astype(np.uint8)
  draws = {"noise": rng.normal(size=N).astype(np.float32), "u_coupled": rng.random((N, B)).astype(np.float32),
           "site": rng.integers(0, L, size=(N, B)).astype(np.int32), "u_mask": rng.random((N, B, L)).astype(np.float32),
           "new_vals": rng.integers(0, q, size=(N, B, L)).astype(np.uint8), "u_accept": rng.random((N, B)).astype(np.float32)}
  return adj, inter, tables, root, draws


@pytest.mark.parametrize("n_leaves,L,k,q,B,rate,std,tree,selfloop", [
  (8, 30, 2, 4, 2, 0.3, 0.2, "balanced", False), (16, 64, 1, 20, 1, 0.1, 0.0, "balanced", True),
  (13, 50, 0, 2, 3, 0.5, 0.5, "random", False), (32, 200, 2, 4, 2, 0.05, 0.3, "random", True),
  (64, 1000, 1, 20, 1, 0.1, 0.0, "balanced", False)])
def test_nk_generate_tree_with_explicit_draws_equals_the_oracle(n_leaves, L, k, q, B, rate, std, tree, selfloop):
  """Intended semantics (every node after its parent; a row without a positive entry is a root): the evolve body of
  the oracle is the one pinned against the reference in tests/test_generators_oracle.py."""
  import torch
  from trex_b200 import generators as G

  adj, inter, tables, root, draws = _nk_case(n_leaves, L, k, q, B, seed=n_leaves + L, tree=tree)
  if selfloop:
    adj[-1, -1] = 1
  want = go.generate_tree_data(inter, tables, adj, root, rate, go.array_draws(draws), coupled_mutation_prob=0.5,
                               n_states=q, mutation_rate_noise_std=std, branch_length=B, literal_bfs=False)
  land = {"interactions": torch.from_numpy(inter).cuda(), "fitness_tables": torch.from_numpy(tables).cuda(),
          "n_states": q, "k": k}
  t = G.generate_tree_data(land, adj, root, rate, seed=0, coupled_mutation_prob=0.5, mutation_rate_noise_std=std,
                           branch_length=B, draws=draws)
  got = t.all_sequences.cpu().numpy().astype(np.int32)
  assert np.array_equal(got, want)
  assert np.array_equal(got[-1], root.astype(np.int32))                   # the root keeps the caller's sequence
  assert float(t.masked_sequences.abs().sum()) == 0.0                     # nk_model.py:271


def test_nk_generate_tree_philox_mode_properties():
  """What tests/test_nk_model_new.py checks of the reference: determinism per key, more divergence from the root with
  longer branches — plus: every node gets a sequence (the reference's BFS loses nodes, see the oracle tests)."""
  from trex_b200 import generators as G

  n_leaves, L, k, q = 16, 2000, 2, 4
  land = G.create_nk_model_landscape(L, k, seed=42, n_states=q)
  inter = land["interactions"].cpu().numpy()
  tables = land["fitness_tables"].cpu().numpy()
  assert inter.min() >= 0 and inter.max() < L and tables.min() >= 0 and tables.max() < 1
  assert abs(tables.mean() - 0.5) < 0.01 and abs(inter.mean() - (L - 1) / 2) < 0.05 * L
  adj = G.create_balanced_binary_tree(n_leaves)
  root = np.random.default_rng(0).integers(0, q, size=L)

  def mean_dist(B, seed):
    s = G.generate_tree_data(land, adj, root, 0.05, seed=seed, branch_length=B).all_sequences.cpu().numpy()
    assert np.array_equal(s[-1], root)
    return float((s[:-1] != root[None, :]).mean()), s

  d1, s1 = mean_dist(1, 101)
  d10, _ = mean_dist(10, 101)
  assert d10 > d1 > 0                                                      # test_generate_tree_variable_branch_length
  _, s1b = mean_dist(1, 101)
  assert np.array_equal(s1, s1b)                                           # same key, same tree
  parent = go.create_balanced_binary_tree(n_leaves).argmax(axis=1)
  changed = np.array([(s1[v] != s1[parent[v]]).mean() for v in range(2 * n_leaves - 2)])
  assert changed.max() <= 0.2                                              # one step: rate 0.05 or a (k+1)-site coupled move
  assert (changed > 0).sum() >= n_leaves                                   # most edges accept something
