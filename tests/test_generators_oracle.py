"""CPU tests: the NumPy restatement of the reference's input generators (oracle/generators_oracle.py) against golden
vectors of the UNMODIFIED reference (tests/golden/generators_golden.npz, made by make_golden_generators.py under the
NumPy stand-in for jax with a recording jax.random).  The draws the reference consumed are part of the fixture."""

import os

import numpy as np
import pytest

from oracle import generators_oracle as go

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "generators_golden.npz"))
GT_CASES = ["gt_a", "gt_b", "gt_c", "gt_d"]
NK_CASES = ["nk_a", "nk_b", "nk_c", "nk_d"]


def log_draws(name):
  """``draw`` callable that pops the reference's recorded draws in call order and checks the kind of each."""
  kinds, sizes, vals = GOLD[name + "_log_kind"], GOLD[name + "_log_size"], GOLD[name + "_log_vals"]
  offs = np.concatenate([[0], np.cumsum(sizes)])
  want = {"noise": 0, "u_coupled": 1, "u_mask": 1, "u_accept": 1, "site": 2, "new_vals": 2}
  state = {"i": 0}

  def draw(kind, node, step):
    i = state["i"]
    assert i < len(kinds), "the restatement draws more often than the reference did"
    assert kinds[i] == want[kind], (i, kind, int(kinds[i]))
    v = vals[offs[i]:offs[i + 1]]
    state["i"] = i + 1
    if kind in ("site", "new_vals"):
      v = v.astype(np.int64)
    else:
      v = v.astype(np.float32)
    return v[0] if kind in ("noise", "u_coupled", "u_accept", "site") else v

  draw.state, draw.n = state, len(kinds)
  return draw


@pytest.mark.parametrize("name", GT_CASES)
def test_generate_groundtruth_restatement_equals_the_reference(name):
  n_leaves, A, m, L = (int(x) for x in GOLD[name + "_cfg"])
  masked, all_seqs, adj = go.generate_groundtruth(n_leaves, A, L, GOLD[name + "_sites"], GOLD[name + "_offsets"])
  assert np.array_equal(all_seqs, GOLD[name + "_all"])
  assert np.array_equal(masked, GOLD[name + "_masked"])
  assert np.array_equal(adj, GOLD[name + "_adj"])
  # mutate: exactly n_mutations sites differ between a child and its parent (ground_truth.py:39-50, offsets >= 1)
  n_all = 2 * n_leaves - 1
  for c in range(n_all - 1):
    assert int((all_seqs[c] != all_seqs[n_leaves + c // 2]).sum()) == (m if A > 1 else 0)


def test_generate_groundtruth_rejects_non_powers_of_two():
  with pytest.raises(ValueError, match="power of 2"):
    go.generate_groundtruth(6, 4, 5, np.zeros((10, 1), np.int32), np.ones((10, 5), np.uint8))


@pytest.mark.parametrize("name", NK_CASES)
def test_get_fitness_restatement_equals_the_reference(name):
  n_leaves, L, k, q, B, _ = (int(x) for x in GOLD[name + "_cfg"])
  inter, tables = GOLD[name + "_inter"], GOLD[name + "_tables"]
  got = np.array([go.get_fitness(s, inter, tables, q) for s in GOLD[name + "_fit_seqs"]])
  assert np.abs(got - GOLD[name + "_fit"]).max() <= 2e-7          # f32 sums in a different order
  gotm = np.array([go.get_fitness(s, inter, tables, q, GOLD[name + "_fit_mask"] > 0) for s in GOLD[name + "_fit_seqs"]])
  assert np.abs(gotm - GOLD[name + "_fit_masked"]).max() <= 2e-7


@pytest.mark.parametrize("name", NK_CASES)
def test_generate_tree_data_literal_restatement_equals_the_reference(name):
  n_leaves, L, k, q, B, selfloop = (int(x) for x in GOLD[name + "_cfg"])
  rate, std = (float(x) for x in GOLD[name + "_rate_std"])
  draw = log_draws(name)
  got = go.generate_tree_data(GOLD[name + "_inter"], GOLD[name + "_tables"], GOLD[name + "_adj"], GOLD[name + "_root"],
                              rate, draw, coupled_mutation_prob=0.5, n_states=q, mutation_rate_noise_std=std,
                              branch_length=B, literal_bfs=True)
  assert draw.state["i"] == draw.n                                 # every recorded draw consumed, in the same order
  assert np.array_equal(got, GOLD[name + "_all"].astype(np.int32))


def test_the_reference_bfs_loses_nodes_and_the_intended_order_does_not():
  """Documents the reference defect the CUDA generator does NOT reproduce (nk_model.py:149, :165): on the adjacency of
  the reference's own callers (balanced tree, no root self-loop) the literal BFS starts at node 0 and visits almost
  nothing; with a self-loop the zero-padded children list floods the queue with node 0 and real children are lost."""
  adj = go.create_balanced_binary_tree(8)
  literal = go._bfs_literal(adj, 0)                               # no self-loop -> root falls back to node 0
  assert sorted(set(literal.tolist())) == [-1, 0]
  adj[-1, -1] = 1
  literal = go._bfs_literal(adj, adj.shape[0] - 1)
  assert len(set(literal.tolist()) - {-1}) < adj.shape[0]          # some node is never visited
  parent = np.where(adj.max(axis=1) > 0, adj.argmax(axis=1), np.arange(adj.shape[0]))
  order = go._bfs_intended(parent, {adj.shape[0] - 1})
  assert sorted(order.tolist()) == list(range(adj.shape[0]))
  pos = {v: i for i, v in enumerate(order.tolist())}
  assert all(pos[int(parent[v])] < pos[v] for v in range(adj.shape[0] - 1))
