"""GPU parity tests of the hard Sankoff path: every check goes through the C ABI
(``libtrex_b200.so`` via ``trex_b200._lib``) and is compared with the oracle on the same inputs.
Bit-exact for dp tables, per-site scores, back-pointers and traceback states."""

import numpy as np
import pytest

from oracle import sankoff_oracle as so
from oracle import tree_oracle as to
from trex_b200 import synth

pytestmark = pytest.mark.gpu


def _torch():
  import torch

  return torch


def _case(kind, n, A, L, seed, cost_kind="unit", rate=0.2):
  rng = np.random.default_rng(seed)
  par = {"balanced": synth.balanced_parents, "gt": synth.groundtruth_parents,
         "cat": synth.caterpillar_parents}.get(kind, None)
  par = par(n) if par else synth.random_join_parents(n, rng)
  adj = synth.parents_to_adjacency(par)
  states = synth.evolve(par, L, A, rate, rng)
  if cost_kind == "unit":
    cost = synth.unit_cost(A)
  elif cost_kind == "unit3":  # uniform off-diagonal cost != 1 (O(A) kernels, sankoff_uniform.cu)
    cost = 3.0 * synth.unit_cost(A)
  elif cost_kind == "int":
    cost = synth.random_int_cost(A, 5, rng)
  else:  # non-integer, asymmetric: exercises the exact-f32 path
    cost = (rng.random((A, A)) * 3).astype(np.float32)
  return adj, cost, states, par


def _native_tables(eng, tree=0):
  """dp (L,n_all,A), argmin (L,n_int,2,A) in the oracle's layout."""
  dp = eng.export_dp(tree).cpu().numpy()
  arg = eng.argmin_table()[tree, :, :, : eng.L, :].permute(2, 0, 3, 1).cpu().numpy()  # (L, n_int, slot, A)
  return dp, arg


CASES = [
  ("balanced", 16, 4, 100, "unit"),   # BASELINE config c1
  ("random", 9, 2, 1, "unit"),
  ("random", 13, 4, 17, "int"),
  ("cat", 12, 5, 33, "int"),
  ("gt", 8, 20, 50, "unit"),
  ("random", 24, 20, 130, "int"),
  ("random", 10, 20, 64, "float"),
  ("random", 7, 3, 250, "float"),
  ("balanced", 32, 7, 40, "int"),
  ("random", 6, 24, 19, "int"),       # A in 21..32 -> padded AP=32 kernels
  ("random", 6, 24, 19, "unit"),
  ("cat", 21, 20, 70, "unit3"),
  ("random", 40, 5, 200, "unit3"),
  ("random", 33, 8, 129, "unit"),
]


@pytest.mark.parametrize("kind,n,A,L,ck", CASES)
@pytest.mark.parametrize("mode", [0, 1])
def test_hard_forward_and_traceback_bit_exact(kind, n, A, L, ck, mode):
  torch = _torch()
  from trex_b200.sankoff import SankoffEngine

  adj, cost, states, _ = _case(kind, n, A, L, seed=n * 131 + A * 7 + L, cost_kind=ck)
  n_all = 2 * n - 1
  tabs = so.run_sankoff_tables(adj, cost, states[:n], n_all, A, n)
  eng = SankoffEngine(n_all, n, A, L)
  eng.set_adjacency(torch.from_numpy(adj)).set_leaf_states_u8(torch.from_numpy(states[:n]))
  eng.hard_forward(torch.from_numpy(cost), mode=mode)
  dp, arg = _native_tables(eng)
  assert np.array_equal(dp, tabs["dp"])  # bit-exact f32, incl. the implicit leaf rows
  assert np.array_equal(arg, tabs["argmin"][:, n:].astype(np.uint8))
  assert np.array_equal(eng.root_min[0, :L].cpu().numpy(), so.per_site_root_min(tabs["dp"]))
  assert float(eng.total[0]) == float(so.per_site_root_min(tabs["dp"]).astype(np.float64).sum())
  eng.traceback(want_hist=True)
  want = so.traceback_topdown(tabs, n_all, n)
  got = eng.anc[0, :, :L].cpu().numpy()
  assert np.array_equal(got, want[n:].astype(np.uint8))
  # edge histogram == transitions of the reconstruction
  full = np.concatenate([states[:n], got], axis=0).astype(np.int64)
  hist = np.zeros((A, A), dtype=np.int64)
  for node in range(n, n_all):
    for c in tabs["children"][node]:
      if c >= 0:
        np.add.at(hist, (full[node], full[c]), 1)
  assert np.array_equal(eng.edge_hist[0].cpu().numpy(), hist)
  n_reach, status = eng.plan_status()
  assert n_reach == n - 1 and status == 0


def test_reference_test_inputs_known_answer():
  """/root/reference/tests/test_sankoff.py:39-72 + the hand-derived values of SURVEY §8c."""
  torch = _torch()
  from trex_b200.sankoff import run_sankoff

  adj = np.zeros((5, 5), dtype=np.int32)
  adj[0, 3] = adj[1, 3] = adj[2, 4] = adj[3, 4] = 1
  cost = np.ones((2, 2), np.float32) - np.eye(2, dtype=np.float32)
  seqs = np.array([[0, 1], [1, 0], [0, 0], [0, 0], [0, 0]], dtype=np.float32)
  rec, dp, total = run_sankoff(torch.from_numpy(adj), torch.from_numpy(cost), torch.from_numpy(seqs[:3]),
                               5, 2, 3, return_path=True)
  assert tuple(rec.shape) == (5, 2) and tuple(dp.shape) == (2, 5, 2)
  assert float(total) == 2.0 and total.dtype == torch.float32
  assert rec.cpu().numpy().tolist() == [[0, 1], [1, 0], [0, 0], [0, 0], [0, 0]]
  dpn = dp.cpu().numpy()
  for site in range(2):
    assert dpn[site, 3].tolist() == [1.0, 1.0] and dpn[site, 4].tolist() == [1.0, 2.0]
  assert dpn[0, 0].tolist() == [0.0, 1e5] and dpn[0, 1].tolist() == [1e5, 0.0]
  rec0, _, _ = run_sankoff(torch.from_numpy(adj), torch.from_numpy(cost), torch.from_numpy(seqs[:3]),
                           5, 2, 3, return_path=False)
  assert rec0.cpu().numpy().tolist() == [[0, 1], [1, 0], [0, 0], [0, 0], [0, 0]]


def test_reference_run_dp_unit_test_malformed_adjacency():
  """/root/reference/tests/test_sankoff.py:9-36: node 2 has no children (fill_value=-1 path)."""
  torch = _torch()
  from trex_b200.sankoff import run_dp

  adj = np.array([[0, 1, 0], [0, 1, 0], [0, 0, 0]], dtype=np.float32)
  seqs = np.array([[0], [1], [0]], dtype=np.float32)
  cost = np.array([[0, 1], [1, 0]], dtype=np.float32)
  dp0 = np.full((3, 2), 1e5, dtype=np.float32)
  bt0 = np.zeros((3, 2, 4), dtype=np.float32)
  dp, bt = run_dp(torch.from_numpy(adj), torch.from_numpy(dp0), torch.from_numpy(bt0),
                  torch.from_numpy(seqs), torch.from_numpy(cost))
  wdp, wbt = so.run_dp(adj, dp0, bt0, seqs[:, 0], cost)
  assert tuple(dp.shape) == (3, 2) and tuple(bt.shape) == (3, 2, 4)
  assert np.array_equal(dp.cpu().numpy(), wdp)
  assert np.array_equal(bt.cpu().numpy(), wbt)
  assert dp[0, 0] == 0 and dp[1, 1] == 0


def test_out_of_range_and_negative_leaf_states():
  torch = _torch()
  from trex_b200.sankoff import run_sankoff

  rng = np.random.default_rng(5)
  n, A, L = 8, 4, 60
  adj, cost, states, _ = _case("random", n, A, L, seed=77, cost_kind="int")
  seqs = states[:n].astype(np.float32)
  seqs[rng.integers(0, n, 12), rng.integers(0, L, 12)] = 9.0   # dropped scatter -> all-1e5 row
  seqs[rng.integers(0, n, 6), rng.integers(0, L, 6)] = -1.0    # wraps to A-1
  want = so.run_sankoff(adj, cost, seqs, 2 * n - 1, A, n, return_path=True)
  got = run_sankoff(torch.from_numpy(adj), torch.from_numpy(cost), torch.from_numpy(seqs), 2 * n - 1, A, n,
                    return_path=True)
  assert np.array_equal(got[1].cpu().numpy(), want[1])
  assert np.array_equal(got[0].cpu().numpy(), want[0])
  assert float(got[2]) == float(want[2])


def test_batched_trees_share_one_alignment():
  """BASELINE config c4 in miniature: vmap over candidate trees."""
  torch = _torch()
  from trex_b200.sankoff import SankoffEngine

  rng = np.random.default_rng(4)
  n, A, L, B = 12, 4, 70, 9
  n_all = 2 * n - 1
  states = rng.integers(0, A, size=(n, L)).astype(np.uint8)
  adjs = np.stack([synth.random_join_tree(n, rng) for _ in range(B)])
  cost = synth.unit_cost(A)
  eng = SankoffEngine(n_all, n, A, L, batch=B)
  eng.set_adjacency(torch.from_numpy(adjs)).set_leaf_states_u8(torch.from_numpy(states))
  eng.hard_forward(torch.from_numpy(cost))
  eng.traceback()
  for b in range(B):
    tabs = so.run_sankoff_tables(adjs[b], cost, states, n_all, A, n)
    dp, arg = _native_tables(eng, b)
    assert np.array_equal(dp, tabs["dp"])
    assert float(eng.total[b]) == float(so.per_site_root_min(tabs["dp"]).sum())
    want = so.traceback_topdown(tabs, n_all, n)
    assert np.array_equal(eng.anc[b, :, :L].cpu().numpy(), want[n:].astype(np.uint8))


def test_backtrack_sankoff_jit_on_reference_layout_table():
  torch = _torch()
  from trex_b200.sankoff import backtrack_sankoff_jit

  adj, cost, states, _ = _case("random", 9, 4, 5, seed=3, cost_kind="int")
  n, n_all, A = 9, 17, 4
  for l in range(5):
    dp1, bt1 = so.run_dp(adj, np.full((n_all, A), 1e5, np.float32), np.zeros((n_all, A, 4), np.float32),
                         states[:n, l].astype(np.float32), cost)
    rs = int(np.argmin(dp1[-1]))
    want = so.backtrack(n_all - 1, rs, bt1, n_all, n)
    got = backtrack_sankoff_jit(n_all - 1, rs, torch.from_numpy(bt1), n_all, n)
    assert got.dtype == torch.int32 and np.array_equal(got.cpu().numpy(), want)


def test_score_equals_compute_cost_of_reconstruction():
  """tests/test_convergence.py:69-73 — the reference's one numeric consistency check."""
  torch = _torch()
  from trex_b200.sankoff import run_sankoff

  adj, cost, states, _ = _case("gt", 4, 4, 20, seed=42)
  rec, _, total = run_sankoff(torch.from_numpy(adj), torch.from_numpy(cost),
                              torch.from_numpy(states[:4].astype(np.float32)), 7, 4, 4, return_path=True)
  onehot = np.eye(4, dtype=np.float32)[rec.cpu().numpy().astype(np.int64)]
  assert abs(float(total) - float(to.compute_cost(onehot, adj, cost))) <= 1e-3


@pytest.mark.parametrize("n,A,L", [(256, 20, 10_000), (1024, 20, 20_000), (128, 4, 5_000)])
def test_full_size_properties(n, A, L):
  """Sizes of BASELINE configs c3/c4/c5 (c5 on a 20k-site slice per call to bound test time):
  packed-key path == exact-f32 path bit for bit, score == Fitch count, and the cost of the
  traced-back reconstruction re-computed edge by edge equals the score."""
  torch = _torch()
  from trex_b200.sankoff import SankoffEngine

  rng = np.random.default_rng(n + A)
  par = synth.balanced_parents(n)
  adj = synth.parents_to_adjacency(par)
  states = synth.evolve(par, L, A, 0.1, rng)
  n_all = 2 * n - 1
  cost = synth.unit_cost(A)
  eng = SankoffEngine(n_all, n, A, L)
  eng.set_adjacency(torch.from_numpy(adj)).set_leaf_states_u8(torch.from_numpy(states[:n]))
  eng.hard_forward(torch.from_numpy(cost), mode=0)
  dp_fast = eng.dp.clone()
  arg_fast = eng.argmin_table().clone()
  rm = eng.root_min[0, :L].cpu().numpy()
  eng.hard_forward(torch.from_numpy(cost), mode=1)
  assert torch.equal(dp_fast[..., :L], eng.dp[..., :L])
  assert not eng.wrote_pointer_words()  # mode=1: the byte table, whatever the matrix
  assert torch.equal(arg_fast[..., :L, :], eng.argmin[..., :L, :])
  fitch = so.fitch_score(adj, states[:n], n_all, A, n)
  assert np.array_equal(rm.astype(np.int64), fitch)
  assert int(eng.total[0]) == int(fitch.sum())
  eng.traceback(want_hist=True)
  full = np.concatenate([states[:n], eng.anc[0, :, :L].cpu().numpy()], axis=0)
  edge = (full[par[: n_all - 1]] != full[: n_all - 1]).sum(axis=0)
  assert np.array_equal(edge, fitch)  # the reconstruction realises the optimal score
  hist = eng.edge_hist[0].cpu().numpy()
  assert hist.sum() == (n_all - 1) * L and hist.sum() - np.trace(hist) == fitch.sum()


@pytest.mark.parametrize("kind,n", [("balanced", 1024), ("random", 300), ("cat", 64), ("balanced", 4), ("random", 2)])
def test_plan_segments_partition_the_post_order(kind, n):
  """The plan's subtree decomposition (common.cuh: PlanView / Walk): segments are disjoint contiguous
  runs of the post-order, each closed under taking children; the top list holds every other node in
  post-order; together they cover each internal node exactly once."""
  torch = _torch()
  from trex_b200.sankoff import SankoffEngine

  adj, _, _, _ = _case(kind, n, 4, 8, seed=n)
  n_all, n_int = 2 * n - 1, n - 1
  eng = SankoffEngine(n_all, n, 4, 8)
  eng.set_adjacency(torch.from_numpy(adj))
  plan = eng.plan[0].cpu().numpy()
  hdr, body = plan[:8], plan[8:]
  ch = [body[0:n_int], body[n_int:2 * n_int]]
  order = body[4 * n_int:5 * n_int]
  top = body[7 * n_int:8 * n_int]
  seg_start, seg_len = body[8 * n_int:8 * n_int + 128], body[8 * n_int + 128:8 * n_int + 256]
  n_reach, n_seg, n_top_reach, n_top = int(hdr[1]), int(hdr[2]), int(hdr[4]), int(hdr[5])
  assert n_reach == n_int and sorted(order.tolist()) == list(range(n, n_all))
  pos = {int(v): i for i, v in enumerate(order)}
  for k in range(n_int):  # post-order: children before parents
    for c in (ch[0][k], ch[1][k]):
      if c >= n:
        assert pos[int(c)] < pos[n + k]
  covered = []
  for s in range(n_seg):
    run = order[seg_start[s]:seg_start[s] + seg_len[s]].tolist()
    members = set(run)
    for v in run:  # closed under children, and the run's last node is the only one with an outside parent
      for c in (ch[0][v - n], ch[1][v - n]):
        if c >= n:
          assert int(c) in members
    covered += run
  assert len(covered) == len(set(covered))
  assert n_top == n_top_reach == n_int - len(covered)
  assert sorted(covered + top[:n_top].tolist()) == list(range(n, n_all))
  tpos = [pos[int(v)] for v in top[:n_top]]
  assert tpos == sorted(tpos)
  if n >= 64 and kind == "balanced":
    assert n_seg >= 16  # enough independent subtrees to fill the machine on short alignments


@pytest.mark.parametrize("ck", ["unit", "int"])
def test_internal_node_the_root_does_not_reach(ck):
  """A forest: internal node 4 (child: leaf 3) hangs off nothing, the root 6 = (5, 2), 5 = (0, 1).  The
  forward pass still fills node 4's row (the reference computes every node in index order) and the root
  minimum must come from the root's row although node 4 follows the root in the plan's node list."""
  torch = _torch()
  from trex_b200.sankoff import SankoffEngine

  n, A, L = 4, 5, 37
  n_all = 2 * n - 1
  adj = np.zeros((n_all, n_all), dtype=np.float32)
  for c, par in [(0, 5), (1, 5), (5, 6), (2, 6), (3, 4)]:
    adj[c, par] = 1
  rng = np.random.default_rng(11)
  states = rng.integers(0, A, size=(n, L)).astype(np.uint8)
  cost = synth.unit_cost(A) if ck == "unit" else synth.random_int_cost(A, 4, rng)
  tabs = so.run_sankoff_tables(adj, cost, states, n_all, A, n)
  eng = SankoffEngine(n_all, n, A, L)
  eng.set_adjacency(torch.from_numpy(adj)).set_leaf_states_u8(torch.from_numpy(states))
  eng.hard_forward(torch.from_numpy(cost))
  dp, arg = _native_tables(eng)
  assert np.array_equal(dp, tabs["dp"])
  assert np.array_equal(eng.root_min[0, :L].cpu().numpy(), so.per_site_root_min(tabs["dp"]))
  assert float(eng.total[0]) == float(so.per_site_root_min(tabs["dp"]).astype(np.float64).sum())
  eng.traceback()
  want = so.traceback_topdown(tabs, n_all, n)
  assert np.array_equal(eng.anc[0, :, :L].cpu().numpy(), want[n:].astype(np.uint8))


def test_site_blocks_add_up_to_the_whole_alignment():
  """Streaming an alignment in site blocks (SankoffEngine.site_block: views of the same tables) gives the
  same tables, and per-block totals / dC / edge histograms that add up to the single-call results."""
  import torch
  from trex_b200 import synth
  from trex_b200.sankoff import SankoffEngine

  rng = np.random.default_rng(11)
  n, A, L = 32, 20, 1000
  parent = synth.balanced_parents(n)
  states = synth.evolve(parent, L, A, 0.15, rng)[:n]
  adj = torch.from_numpy(synth.parents_to_adjacency(parent).astype(np.float32)).cuda()
  cost = torch.from_numpy(synth.unit_cost(A)).cuda()
  st = torch.from_numpy(states).cuda()

  def run(e):
    e.hard_forward(cost, want_argmin=True)
    e.traceback(want_hist=True)
    e.soft_forward(cost, 0.7)
    e.soft_backward(cost, 0.7, want_dleaf=True, want_dcost=True)

  whole = SankoffEngine(2 * n - 1, n, A, L)
  whole.set_adjacency(adj)
  whole.set_leaf_states_u8(st)
  run(whole)
  ref = {k: getattr(whole, k).clone() for k in ("dp", "anc", "soft_dp", "gbar", "dleaf", "total", "soft_total", "dcost", "edge_hist")}

  eng = SankoffEngine(2 * n - 1, n, A, L)
  eng.set_adjacency(adj)
  blocks = eng.site_blocks(3)
  assert sum(b.L for b in blocks) == L and len(blocks) == 3
  tot = {k: torch.zeros_like(ref[k]) for k in ("total", "soft_total", "dcost", "edge_hist")}
  lo = 0
  for b in blocks:
    b.set_leaf_states_u8(st[:, lo:lo + b.L])
    lo += b.L
    run(b)
    for k in tot:
      tot[k] += getattr(b, k)
  for k in ("dp", "anc", "soft_dp", "gbar", "dleaf"):
    assert torch.equal(getattr(eng, k)[..., :L], ref[k][..., :L]), k
  assert torch.equal(tot["total"], ref["total"]) and torch.equal(tot["edge_hist"], ref["edge_hist"])
  assert abs(float(tot["soft_total"]) - float(ref["soft_total"])) <= 1e-9 * abs(float(ref["soft_total"]))
  assert torch.allclose(tot["dcost"], ref["dcost"], rtol=1e-9, atol=0)


@pytest.mark.parametrize("A,ck", [(20, "unit"), (4, "unit3"), (5, "unit")])
def test_stateless_leaves_take_the_exact_back_pointer_path(A, ck):
  """Leaves without a state (code >= A: the reference's dropped scatter, sankoff.py:50 -> an all-1e5 row) push
  a + min above 2^17, where the one-compare back-pointer keys of the uniform kernel are no longer exact: warps
  holding such a site must fall back to the three-compare rule.  Sites with 0, 1, 2 and many stateless leaves
  are mixed, so neighbouring warps take different paths; dp and back-pointers stay bit-exact."""
  torch = _torch()
  from trex_b200.sankoff import SankoffEngine

  n, L = 32, 300
  adj, cost, states, _ = _case("random", n, A, L, seed=77 + A, cost_kind=ck)
  rng = np.random.default_rng(5)
  st = states[:n].copy()
  for l in range(L):
    k = [0, 1, 2, 3, 8, 20][l % 6]
    st[rng.choice(n, size=k, replace=False), l] = 255 if l % 2 else A
  n_all = 2 * n - 1
  seqs = st.astype(np.float32)
  tabs = so.run_sankoff_tables(adj, cost, seqs, n_all, A, n)
  eng = SankoffEngine(n_all, n, A, L)
  eng.set_adjacency(torch.from_numpy(adj)).set_leaf_states_u8(torch.from_numpy(st))
  eng.hard_forward(torch.from_numpy(cost))
  dp, arg = _native_tables(eng)
  assert np.array_equal(dp, tabs["dp"])
  assert np.array_equal(arg, tabs["argmin"][:, n:].astype(np.uint8))
  eng.traceback()
  want = so.traceback_topdown(tabs, n_all, n)
  assert np.array_equal(eng.anc[0, :, :L].cpu().numpy(), want[n:].astype(np.uint8))


@pytest.mark.parametrize("seg", [0, 1])
@pytest.mark.parametrize("kind,n,A,L,ck,words", [
  ("balanced", 64, 20, 700, "unit", True), ("random", 37, 4, 333, "unit3", True), ("cat", 20, 5, 129, "unit", True),
  ("random", 50, 20, 400, "int", False),    # not uniform: the byte table, whatever the flag says
  ("random", 21, 3, 100, "unit", False),    # 3 states: 8 bytes of words would not fit the 6 bytes of table
  ("random", 21, 24, 100, "unit", False)])  # 24 states run in the 32-wide tile: no words
def test_pointer_words_and_byte_table_are_the_same_back_pointers(kind, n, A, L, ck, words, seg, monkeypatch):
  """TREX_MODE_COMPACT_ARG (include/trex_b200.h): the decoded words equal the byte table for EVERY parent state
  (= the oracle's argmin, sankoff.py:80-83), and the traceback is the same from either form; batch of two trees."""
  torch = _torch()
  from trex_b200.sankoff import SankoffEngine

  monkeypatch.setenv("TREX_SEGMENTS", str(seg))
  adj, cost, states, _ = _case(kind, n, A, L, seed=n * 7 + A, cost_kind=ck)
  adj2, _, _, _ = _case("random", n, A, L, seed=n * 7 + A + 1, cost_kind=ck)
  states[1, ::17] = 255                                        # stateless leaf sites
  n_all = 2 * n - 1
  both = torch.from_numpy(np.stack([adj, adj2]))
  out = {}
  for compact in (False, True):
    eng = SankoffEngine(n_all, n, A, L, batch=2)
    eng.set_adjacency(both).set_leaf_states_u8(torch.from_numpy(states[:n]))
    eng.argmin.fill_(0xEE)                                      # rows the words do not cover must not matter
    eng.hard_forward(torch.from_numpy(cost), want_argmin=True, compact_arg=compact)
    assert eng.wrote_pointer_words() == (compact and words)
    table = eng.argmin_table()[..., :L, :].cpu().numpy()
    eng.traceback(want_hist=True)
    out[compact] = (table, eng.anc[..., :L].cpu().numpy(), eng.edge_hist.cpu().numpy(), eng.dp[..., :L].cpu().numpy())
  for a, b in zip(out[False], out[True]):
    assert np.array_equal(a, b)
  for t, ad in enumerate((adj, adj2)):
    tabs = so.run_sankoff_tables(ad, cost, states[:n], n_all, A, n)
    want = np.transpose(tabs["argmin"][:, n:].astype(np.uint8), (1, 3, 0, 2))  # (L, n_int, slot, A) -> (n_int, A, L, slot)
    assert np.array_equal(out[True][0][t], want)
    assert np.array_equal(out[True][1][t], so.traceback_topdown(tabs, n_all, n)[n:].astype(np.uint8))
