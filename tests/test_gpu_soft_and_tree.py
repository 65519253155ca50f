"""GPU parity tests: relaxed (softmin) Sankoff forward/backward and the soft-adjacency tree cost,
through the C ABI, against the float64 oracle.  Tolerance: 1e-5 relative (BASELINE north star),
written out in each assertion."""

import numpy as np
import pytest

from oracle import sankoff_oracle as so
from oracle import tree_oracle as to
from trex_b200 import synth

pytestmark = pytest.mark.gpu

RTOL = 1e-5


def _relerr(got, want):
  got, want = np.asarray(got, dtype=np.float64), np.asarray(want, dtype=np.float64)
  return np.abs(got - want).max() / max(np.abs(want).max(), 1e-30)


def _soft_case(kind, n, A, L, seed, cost_kind):
  rng = np.random.default_rng(seed)
  par = synth.balanced_parents(n) if kind == "balanced" else (
    synth.caterpillar_parents(n) if kind == "cat" else synth.random_join_parents(n, rng))
  adj = synth.parents_to_adjacency(par)
  states = synth.evolve(par, L, A, 0.2, rng)
  cost = synth.unit_cost(A) if cost_kind == "unit" else 2.5 * synth.unit_cost(A) if cost_kind == "unit2.5" else (
    synth.random_int_cost(A, 5, rng) if cost_kind == "int" else (rng.random((A, A)) * 2).astype(np.float32))
  return adj, cost, states


SOFT_CASES = [
  ("balanced", 16, 4, 100, "unit", 1.0),
  ("random", 11, 20, 37, "int", 1.0),
  ("random", 11, 20, 37, "int", 0.1),
  ("cat", 9, 5, 20, "float", 0.5),
  ("random", 14, 2, 300, "unit", 0.3),
  ("random", 8, 4, 50, "int", 0.01),   # max(C)/T > 80 -> direct (non exp-domain) evaluation
  ("random", 6, 24, 21, "unit", 0.7),  # padded AP=32 kernels
  # uniform off-diagonal cost: the O(A) forward kernels of sankoff_uniform.cu
  ("cat", 19, 20, 150, "unit", 1.0),
  ("random", 37, 20, 70, "unit2.5", 0.6),
  ("random", 23, 7, 130, "unit2.5", 2.0),
  ("random", 9, 5, 40, "unit", 0.01),   # a/T > 80 -> general kernel, direct evaluation
]


@pytest.mark.parametrize("kind,n,A,L,ck,T", SOFT_CASES)
def test_soft_forward_backward_hard_leaves(kind, n, A, L, ck, T):
  import torch
  from trex_b200.sankoff import SankoffEngine

  adj, cost, states = _soft_case(kind, n, A, L, seed=n * 17 + A + L, cost_kind=ck)
  n_all = 2 * n - 1
  lc = so.leaf_cost_from_states(states[:n], A)
  dp64, tot64 = so.soft_sankoff_fwd(adj, cost, lc, n_all, n, T)
  g64, dC64 = so.soft_sankoff_bwd(adj, cost, dp64, n_all, n, T)
  eng = SankoffEngine(n_all, n, A, L)
  eng.set_adjacency(torch.from_numpy(adj)).set_leaf_states_u8(torch.from_numpy(states[:n]))
  eng.soft_forward(torch.from_numpy(cost), T)
  dp = eng.export_dp(soft=True).cpu().numpy()
  assert _relerr(dp[:, n:], dp64[:, n:]) <= RTOL
  assert abs(float(eng.soft_total[0]) - tot64) <= RTOL * abs(tot64)
  eng.soft_backward(torch.from_numpy(cost), T)
  gbar = eng.gbar[0, :, :, :L].permute(2, 0, 1).cpu().numpy()  # (L, n_int, A)
  assert np.abs(gbar - g64[:, n:]).max() <= RTOL  # marginals are O(1) probabilities
  dleaf = eng.dleaf[0, :, :, :L].permute(2, 0, 1).cpu().numpy()
  assert np.abs(dleaf - g64[:, :n]).max() <= RTOL
  assert _relerr(eng.dcost[0].cpu().numpy(), dC64) <= RTOL


@pytest.mark.parametrize("ck,n,A,L,T", [("int", 10, 4, 45, 0.8), ("unit", 13, 20, 77, 0.8), ("unit2.5", 8, 3, 66, 1.5)])
def test_soft_leaves_and_autograd_scaling(ck, n, A, L, T):
  import torch
  from trex_b200.relaxed import soft_sankoff_cost, to_native_leaf_cost
  from trex_b200.sankoff import SankoffEngine

  rng = np.random.default_rng(9)
  adj, cost, _ = _soft_case("random", n, A, L, seed=5, cost_kind=ck)
  n_all = 2 * n - 1
  lc = rng.random((n, L, A)) * 3
  dp64, tot64 = so.soft_sankoff_fwd(adj, cost, lc, n_all, n, T)
  g64, dC64 = so.soft_sankoff_bwd(adj, cost, dp64, n_all, n, T, g=2.5)
  eng = SankoffEngine(n_all, n, A, L)
  C = torch.from_numpy(cost).cuda().requires_grad_(True)
  leaf = to_native_leaf_cost(torch.from_numpy(lc.astype(np.float32)).cuda()).requires_grad_(True)
  total = soft_sankoff_cost(torch.from_numpy(adj), C, leaf, n_all, A, n, T, engine=eng)
  assert abs(float(total) - tot64) <= RTOL * abs(tot64)
  (2.5 * total).backward()
  assert _relerr(C.grad.cpu().numpy(), dC64) <= RTOL
  dleaf = leaf.grad[:, :, :L].permute(2, 0, 1).cpu().numpy()
  assert np.abs(dleaf - g64[:, :n]).max() <= 2.5 * RTOL


def test_soft_limit_T_to_zero_is_hard_score():
  import torch
  from trex_b200.sankoff import SankoffEngine

  adj, cost, states = _soft_case("balanced", 32, 4, 500, seed=2, cost_kind="unit")
  eng = SankoffEngine(63, 32, 4, 500)
  eng.set_adjacency(torch.from_numpy(adj)).set_leaf_states_u8(torch.from_numpy(states[:32]))
  hard = float(eng.hard_forward(torch.from_numpy(cost))[0])
  soft = float(eng.soft_forward(torch.from_numpy(cost), 0.02)[0])
  assert hard - 0.05 * 500 <= soft <= hard + 1e-3


def _tree_inputs(N, L, Q, seed, adj_kind):
  rng = np.random.default_rng(seed)
  n = (N + 1) // 2
  S = to.softmax(rng.normal(size=(N, L, Q))).astype(np.float32)
  S[:n] = np.eye(Q, dtype=np.float32)[rng.integers(0, Q, size=(n, L))]  # one-hot leaves
  if adj_kind == "bernoulli":  # benchmark_tree_cost.py:57
    adj = (rng.random((N, N)) < 0.1).astype(np.float32)
  else:  # row-softmax over the update_tree mask structure (tree.py:81-107)
    adj = to.update_tree(rng.normal(size=(N - 1, N - n)).astype(np.float32),
                         np.zeros((N - 1, N - n), np.float32)).astype(np.float32)
  return S, adj, rng


TREE_CASES = [(127, 1000, 4, "softmax"), (127, 1000, 4, "bernoulli"), (10, 100, 20, "bernoulli"),
              (50, 200, 20, "bernoulli"), (200, 150, 20, "bernoulli"), (7, 3, 5, "softmax"), (65, 33, 3, "softmax")]


@pytest.mark.parametrize("N,L,Q,ak", TREE_CASES)
@pytest.mark.parametrize("ckind", ["none", "unit", "vector", "asym"])
def test_tree_cost_value_and_grads(N, L, Q, ak, ckind):
  import torch
  from trex_b200.tree import tree_cost_value_and_grad

  S, adj, rng = _tree_inputs(N, L, Q, seed=N + L + Q, adj_kind=ak)
  C = {"none": None, "unit": synth.unit_cost(Q), "vector": (rng.random(Q) + 0.5).astype(np.float32),
       "asym": rng.random((Q, Q)).astype(np.float32)}[ckind]
  want, dS64, dA64, dC64 = to.soft_cost_value_and_grad(S, adj, C)
  _, (t1, t2, t3) = to.compute_soft_cost(S, adj, C, dtype=np.float64)
  cost, dS, dA, dC = tree_cost_value_and_grad(
    torch.from_numpy(S).cuda(), torch.from_numpy(adj).cuda(), None if C is None else torch.from_numpy(C).cuda())
  # the expanded quadratic form cancels; bound = 1e-5 relative + 4 eps32 of the terms
  bound = RTOL * abs(want) + 4 * np.finfo(np.float32).eps * (abs(t1) + abs(t2) + abs(t3))
  assert abs(float(cost) - want) <= bound
  assert _relerr(dS.cpu().numpy(), dS64) <= RTOL
  assert _relerr(dA.cpu().numpy(), dA64) <= RTOL
  if C is not None:
    assert _relerr(dC.cpu().numpy(), dC64) <= RTOL
  if ckind != "asym":
    # the direct (N,N,L,Q) form of benchmark_tree_cost.py:10-32 equals tree.py's expanded form only
    # for symmetric C; the library follows tree.py (the shipped function) for asymmetric C
    naive = to.naive_compute_soft_cost(S, adj, C)
    assert abs(float(cost) - naive) <= bound + 1e-9 * abs(naive)


def test_tree_cost_autograd_matches_reference_signatures():
  import torch
  from trex_b200.tree import compute_soft_cost, compute_surrogate_cost

  S, adj, _ = _tree_inputs(15, 20, 4, seed=1, adj_kind="softmax")
  St = torch.from_numpy(S).cuda().requires_grad_(True)
  At = torch.from_numpy(adj).cuda().requires_grad_(True)
  c = compute_surrogate_cost(St, At)
  assert c.shape == () and float(c) >= 0  # tests/test_tree.py:54-61
  (3.0 * c).backward()
  want, dS64, dA64, _ = to.soft_cost_value_and_grad(S, adj, None)
  assert _relerr(St.grad.cpu().numpy(), 3.0 * dS64) <= RTOL
  assert _relerr(At.grad.cpu().numpy(), 3.0 * dA64) <= RTOL
  c2 = compute_soft_cost(St.detach(), At.detach(), torch.eye(4, device="cuda"))
  assert abs(float(c2) - float(c)) <= RTOL * abs(float(c))


TC_CASES = [(5, 7, 3), (127, 1000, 4), (128, 40, 20), (129, 33, 5), (200, 150, 20), (256, 64, 4), (257, 20, 4)]


@pytest.mark.parametrize("N,L,Q", TC_CASES)
def test_tree_cost_tcgen05_and_simt_paths_agree_with_oracle(N, L, Q, monkeypatch):
  """Identity cost, N <= 256 runs on tcgen05 (3xTF32); TREX_TREE_COST_PATH=simt pins the exact-f32
  CUDA-core path; N = 257 falls back to it.  Both must meet the 1e-5 bar against float64
  (tree.py:163-209) and agree with each other."""
  import torch
  from trex_b200.tree import tree_cost_value_and_grad

  S, adj, _ = _tree_inputs(N, L, Q, seed=3 * N + L + Q, adj_kind="bernoulli")
  want, dS64, dA64, _ = to.soft_cost_value_and_grad(S, adj, None)
  _, (t1, t2, t3) = to.compute_soft_cost(S, adj, None, dtype=np.float64)
  bound = RTOL * abs(want) + 4 * np.finfo(np.float32).eps * (abs(t1) + abs(t2) + abs(t3))
  St, At = torch.from_numpy(S).cuda(), torch.from_numpy(adj).cuda()
  got = {}
  for path in ("auto", "simt"):
    monkeypatch.setenv("TREX_TREE_COST_PATH", path)
    cost, dS, dA, _ = tree_cost_value_and_grad(St, At, None)
    torch.cuda.synchronize()
    assert abs(float(cost) - want) <= bound, path
    assert _relerr(dS.cpu().numpy(), dS64) <= RTOL, path
    assert _relerr(dA.cpu().numpy(), dA64) <= RTOL, path
    got[path] = (float(cost), dS.cpu().numpy(), dA.cpu().numpy())
  assert abs(got["auto"][0] - got["simt"][0]) <= 2 * bound
  # value only (no gradients requested)
  monkeypatch.setenv("TREX_TREE_COST_PATH", "auto")
  cost, dS, dA, _ = tree_cost_value_and_grad(St, At, None, want_dseq=False, want_dadj=False)
  assert dS is None and dA is None and abs(float(cost) - want) <= bound


@pytest.mark.parametrize("n,A,L", [(256, 20, 10_000), (128, 4, 5_000)])
def test_full_size_relaxed_properties(n, A, L):
  """Sizes of BASELINE configs c3 / c4 (one tree): size-independent properties of the relaxed path.
  * every node's cotangent row sums to 1 (softmin weights are a distribution, so mass is conserved
    from the root down) and every leaf receives mass 1 per site;
  * sum(dC) = number of edges x sites;
  * the O(A) uniform-cost kernels (C = 1 - I) and the general exp-domain kernels (same C with one entry
    moved by 1e-6, which changes the device-side dispatch and nothing else) agree to 1e-5."""
  import torch
  from trex_b200.sankoff import SankoffEngine

  rng = np.random.default_rng(n * 3 + A)
  par = synth.balanced_parents(n)
  adj = synth.parents_to_adjacency(par)
  states = synth.evolve(par, L, A, 0.1, rng)[:n]
  n_all, T = 2 * n - 1, 1.0
  cost = synth.unit_cost(A)
  cost_g = cost.copy()
  cost_g[0, 1] += 1e-6
  out = []
  for c in (cost, cost_g):
    eng = SankoffEngine(n_all, n, A, L)
    eng.set_adjacency(torch.from_numpy(adj)).set_leaf_states_u8(torch.from_numpy(states))
    tot = float(eng.soft_forward(torch.from_numpy(c), T)[0])
    eng.soft_backward(torch.from_numpy(c), T)
    out.append((tot, eng.soft_dp[..., :L].clone(), eng.gbar[..., :L].clone(), eng.dleaf[..., :L].clone(),
                eng.dcost[0].clone()))
    del eng
  (t0, dp0, g0, dl0, dc0), (t1, dp1, g1, dl1, dc1) = out
  assert torch.allclose(g0.sum(dim=2), torch.ones_like(g0[:, :, 0]), atol=2e-5)
  assert torch.allclose(dl0.sum(dim=2), torch.ones_like(dl0[:, :, 0]), atol=2e-5)
  assert abs(float(dc0.sum()) - 2 * (n - 1) * L) <= 1e-5 * 2 * (n - 1) * L
  assert abs(t0 - t1) <= RTOL * abs(t0)
  assert float((dp0 - dp1).abs().max()) <= RTOL * float(dp0.abs().max())
  assert float((g0 - g1).abs().max()) <= 2 * RTOL
  assert float((dl0 - dl1).abs().max()) <= 2 * RTOL
  assert float((dc0 - dc1).abs().max()) <= RTOL * float(dc0.abs().max())


def _subtree_leaves(par, k, n):
  kids = [v for v in range(len(par)) if par[v] == k]
  out = []
  for v in kids:
    out += [v] if v < n else _subtree_leaves(par, v, n)
  return out


@pytest.mark.parametrize("ck,A,T", [("unit", 20, 1.0), ("int", 20, 1.0), ("unit", 4, 0.5), ("int", 5, 0.01)])
@pytest.mark.parametrize("compact", [False, True])
def test_reverse_pass_on_a_forest_zeroes_the_unreached_rows(ck, A, T, compact):
  """A subtree detached from the root (its parent is left with one child — the -1 fill of sankoff.py:60): the
  cotangents of its nodes and the leaf-gradient rows of its leaves are ZERO (written, not left stale), everything
  reachable matches the oracle."""
  import torch
  from trex_b200.sankoff import SankoffEngine

  n, L = 12, 70
  rng = np.random.default_rng(A + int(10 * T))
  par = synth.random_join_parents(n, rng)
  n_all = 2 * n - 1
  cand = [k for k in range(n, n_all - 1) if len(_subtree_leaves(par, k, n)) >= 2]
  k = cand[len(cand) // 2]
  adj = synth.parents_to_adjacency(par)
  adj[k, par[k]] = 0                                  # detach internal node k
  lost = _subtree_leaves(par, k, n)
  states = synth.evolve(par, L, A, 0.2, rng)
  cost = synth.unit_cost(A) if ck == "unit" else synth.random_int_cost(A, 5, rng)
  lc = so.leaf_cost_from_states(states[:n], A)
  dp64, tot64 = so.soft_sankoff_fwd(adj, cost, lc, n_all, n, T)
  g64, dC64 = so.soft_sankoff_bwd(adj, cost, dp64, n_all, n, T)
  eng = SankoffEngine(n_all, n, A, L)
  eng.set_adjacency(torch.from_numpy(adj)).set_leaf_states_u8(torch.from_numpy(states[:n]))
  eng.soft_forward(torch.from_numpy(cost), T)
  assert abs(float(eng.soft_total[0]) - tot64) <= RTOL * abs(tot64)
  eng.gbar.fill_(float("nan"))
  eng.dleaf.fill_(float("nan"))
  eng.dleaf_obs.fill_(float("nan"))
  eng.soft_backward(torch.from_numpy(cost), T, want_dleaf="compact" if compact else True)
  gbar = eng.gbar[0, :, :, :L].permute(2, 0, 1).cpu().numpy()
  # The parent of the detached node is left with a missing child = the reference's all-1e5 row (sankoff.py:60,152),
  # so every dp row above it is ~1e5: float32 resolves that to 2^-7, and the softmin weights exp(-(x - m)/T) of ANY
  # float32 evaluation move by 2^-24 * max|dp| / T (same clause as tests/test_gpu_bench_shapes.py).
  gtol = max(RTOL, 2.0 ** -24 * np.abs(dp64[:, n:]).max() / T)
  assert not np.isnan(gbar).any()
  assert np.all(gbar[:, k - n] == 0.0)                 # the detached subtree: exact zeros, written by the kernel
  assert np.abs(gbar - g64[:, n:]).max() <= gtol
  assert _relerr(eng.dcost[0].cpu().numpy(), dC64) <= gtol
  if compact:
    obs = eng.dleaf_obs[0, :, :L].cpu().numpy()
    assert np.all(obs[lost] == 0.0)
    assert np.abs(obs - g64[:, :n].sum(axis=2).T).max() <= gtol
  else:
    dleaf = eng.dleaf[0, :, :, :L].permute(2, 0, 1).cpu().numpy()
    assert np.all(dleaf[:, lost] == 0.0)
    assert np.abs(dleaf - g64[:, :n]).max() <= gtol


@pytest.mark.parametrize("ck,A,T", [("unit", 20, 1.0), ("int", 20, 1.0), ("int", 5, 0.01)])
@pytest.mark.parametrize("kind", ["shared_leaf", "shared_node"])
def test_reverse_pass_poisons_dcost_when_the_adjacency_is_not_a_tree(ck, A, T, kind):
  """A leaf or an internal node with two parents: the forward passes still match the reference semantics (and the
  oracle); the reverse pass is defined on trees only and says so in its result — every entry of dC is NaN, with
  no host synchronisation needed to find out (include/trex_b200.h, TREX_PLAN_NOT_A_TREE / TREX_PLAN_SHARED_LEAF)."""
  import torch
  from trex_b200 import _lib
  from trex_b200.sankoff import SankoffEngine

  n, L = 10, 40
  rng = np.random.default_rng(3)
  par = synth.random_join_parents(n, rng)
  n_all = 2 * n - 1
  adj = synth.parents_to_adjacency(par)
  if kind == "shared_leaf":
    victim = 0
    other = next(k for k in range(n, n_all) if k != par[victim] and np.count_nonzero(adj[:, k]) == 2)
    donor = int(np.nonzero(adj[:, other])[0][-1])
    adj[donor, other] = 0                            # `other` now takes leaf 0 as a child as well
    adj[victim, other] = 1
  else:
    victim = next(k for k in range(n, n_all - 2) if par[k] < n_all - 1)
    other = next(k for k in range(victim + 1, n_all) if k != par[victim])
    donor = int(np.nonzero(adj[:, other])[0][-1])
    adj[donor, other] = 0
    adj[victim, other] = 1
  states = synth.evolve(par, L, A, 0.2, rng)
  cost = synth.unit_cost(A) if ck == "unit" else synth.random_int_cost(A, 5, rng)
  eng = SankoffEngine(n_all, n, A, L)
  eng.set_adjacency(torch.from_numpy(adj)).set_leaf_states_u8(torch.from_numpy(states[:n]))
  _, status = eng.plan_status()
  assert status & (_lib.PLAN_NOT_A_TREE | _lib.PLAN_SHARED_LEAF)
  # forward passes: unchanged semantics
  tabs = so.run_sankoff_tables(adj, cost, states[:n], n_all, A, n)
  eng.hard_forward(torch.from_numpy(cost))
  assert np.array_equal(eng.export_dp().cpu().numpy(), tabs["dp"])
  dp64, tot64 = so.soft_sankoff_fwd(adj, cost, so.leaf_cost_from_states(states[:n], A), n_all, n, T)
  eng.soft_forward(torch.from_numpy(cost), T)
  assert abs(float(eng.soft_total[0]) - tot64) <= RTOL * abs(tot64)
  eng.soft_backward(torch.from_numpy(cost), T, want_dleaf="compact")
  assert bool(torch.isnan(eng.dcost[0]).all())


@pytest.mark.parametrize("N,L,Q", [(127, 1000, 4), (5, 7, 4), (128, 40, 20), (64, 500, 4), (100, 3000, 4)])
def test_tree_cost_one_kernel_mode_vs_three_kernel_chain_and_oracle(N, L, Q, monkeypatch):
  """One-kernel mode of the tcgen05 path (tree_cost_tc.cu: tagged flags, per-slice slots, launch counter kept on the
  device) against the three-kernel chain and float64 (tree.py:163-266): one workspace reused over many calls with
  changing inputs, value-only calls in between, and a CUDA-graph replay (same kernel arguments every replay)."""
  import torch
  from trex_b200._lib import load
  from trex_b200.tree import tree_cost_value_and_grad

  lib = load()
  ws = torch.empty(lib.trex_tree_cost_workspace_bytes(N, L * Q, Q), dtype=torch.uint8, device="cuda")
  ws.random_(0, 256)                                             # a fresh workspace holds anything
  launches = {}
  for rep in range(4):
    S, adj, _ = _tree_inputs(N, L, Q, seed=7 * N + L + Q + rep, adj_kind="bernoulli")
    want, dS64, dA64, _ = to.soft_cost_value_and_grad(S, adj, None)
    _, (t1, t2, t3) = to.compute_soft_cost(S, adj, None, dtype=np.float64)
    bound = RTOL * abs(want) + 4 * np.finfo(np.float32).eps * (abs(t1) + abs(t2) + abs(t3))
    St, At = torch.from_numpy(S).cuda(), torch.from_numpy(adj).cuda()
    for one in ("1", "0"):
      monkeypatch.setenv("TREX_TREE_COST_ONE", one)
      before = lib.trex_kernel_launches()
      cost, dS, dA, _ = tree_cost_value_and_grad(St, At, None, workspace=ws)
      launches[one] = lib.trex_kernel_launches() - before
      torch.cuda.synchronize()
      assert abs(float(cost) - want) <= bound, (rep, one)
      assert _relerr(dS.cpu().numpy(), dS64) <= RTOL, (rep, one)
      assert _relerr(dA.cpu().numpy(), dA64) <= RTOL, (rep, one)
      c2, _, _, _ = tree_cost_value_and_grad(St, At, None, want_dseq=False, want_dadj=False, workspace=ws)
      assert abs(float(c2) - want) <= bound, (rep, one)
  assert launches["1"] == 1 and launches["0"] >= 2
  # graph replay: the launch counter advances on the device, so identical launches stay correct
  monkeypatch.setenv("TREX_TREE_COST_ONE", "1")
  out = {"cost": torch.empty((), device="cuda"), "dS": torch.empty_like(St), "dA": torch.empty_like(At)}
  stream = torch.cuda.Stream()
  with torch.cuda.stream(stream):
    tree_cost_value_and_grad(St, At, None, workspace=ws, out=out)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=stream):
      tree_cost_value_and_grad(St, At, None, workspace=ws, out=out)
    for _ in range(5):
      out["cost"].zero_(); out["dS"].zero_(); out["dA"].zero_()
      g.replay()
  torch.cuda.synchronize()
  assert abs(float(out["cost"]) - want) <= bound
  assert _relerr(out["dS"].cpu().numpy(), dS64) <= RTOL and _relerr(out["dA"].cpu().numpy(), dA64) <= RTOL
