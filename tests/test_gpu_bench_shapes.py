"""GPU parity tests AT THE BENCHMARKED SHAPES, against the oracle (never GPU vs GPU).

``bench.py`` times c5 (1024 taxa x 100k sites, A = 20) — at n_leaves >= 64 every walk kernel takes the
two-launch subtree-segment path (``plan.cu:use_segments``), which the small-tree cases of the other test
files never reach.  Sites are independent (sankoff.py:97), so a site slice of the benchmark alignment
exercises exactly the code the bench runs: c5 slice = 1024 taxa x 2000 sites (ragged: not a multiple of
the 64-site warp tile), c3 slice = 256 taxa x 4096 sites, both with the unit cost and with a random
symmetric integer matrix in {0..5}, T in {1.0, 0.1}, TREX_SEGMENTS = 0 and 1.

Checkers: hard path — the plain-C restatement with the reference's own data structures
(``oracle/sankoff_oracle.c``, itself checked against the golden vectors of the unmodified reference) and
the NumPy restatement for the back-pointer table; relaxed path — the float64 instance of the C
restatement, which tests/test_relaxed_oracles.py pins against torch.autograd over a literal transcription
of the reference recurrence.  Bars: bit-exact (hard), 1e-5 relative (relaxed), with two float32-conditioning
clauses that profiles/r02_b_relaxed_err_vs_f32_port.txt backs with measurements (the float32 instance of the
C restatement — the reference's default dtype — shows the same or larger deviations from float64 on these inputs):
* the marginals are softmin weights exp(-(x - m)/T): a half-ulp rounding of a dp entry of magnitude D moves a
  weight by 2^-24 * D / T relative, so their bar is max(1e-5, 2^-24 * max|dp| / T) (1.4e-4 at the c5 slice with
  the integer matrix at T = 0.1, where max|dp| = 578; the float32 port is off by 1.5e-4 there);
* the total is a sum of per-site values of both signs (at T = 1 the entropy term -T log(#near-optimal paths)
  makes many of them negative: -2 247 against sum |.| = 66 596 at the c5 slice with the integer matrix), so its
  bar is 1e-5 of the l1 norm of the per-site values."""

import functools
import os

import numpy as np
import pytest

from oracle import c_oracle
from oracle import sankoff_oracle as so
from trex_b200 import synth

pytestmark = pytest.mark.gpu

RTOL = 1e-5
SHAPES = {"c5slice": (1024, 20, 2000, 5), "c3slice": (256, 20, 4096, 3)}


@functools.lru_cache(maxsize=None)
def _inputs(shape, cost_kind):
  n, A, L, seed = SHAPES[shape]
  rng = np.random.default_rng(seed)
  par = synth.balanced_parents(n)
  adj = synth.parents_to_adjacency(par)
  states = np.ascontiguousarray(synth.evolve(par, L, A, 0.1, rng)[:n])
  cost = synth.unit_cost(A) if cost_kind == "unit" else synth.random_int_cost(A, 5, rng)
  return adj, cost, states


@functools.lru_cache(maxsize=2)
def _hard_oracle(shape, cost_kind):
  n, A, L, _ = SHAPES[shape]
  adj, cost, states = _inputs(shape, cost_kind)
  recon, dp, total, root_min = c_oracle.run_sankoff(adj, cost, states.astype(np.float32), 2 * n - 1, A, n,
                                                    return_path=True, want_dp=True)
  return recon, dp, total, root_min


@functools.lru_cache(maxsize=2)  # (0.65 GB per float64 table at the c5 slice; seg varies fastest)
def _soft_oracle(shape, cost_kind, T):
  n, A, L, _ = SHAPES[shape]
  adj, cost, states = _inputs(shape, cost_kind)
  return c_oracle.soft_sankoff_f64(adj, cost, 2 * n - 1, A, n, T, seqs=states.astype(np.float32))


class _segments:
  """TREX_SEGMENTS is read by the library at every launch (plan.cu:use_segments)."""

  def __init__(self, v):
    self.v = v

  def __enter__(self):
    self.old = os.environ.get("TREX_SEGMENTS")
    os.environ["TREX_SEGMENTS"] = str(self.v)

  def __exit__(self, *a):
    if self.old is None:
      os.environ.pop("TREX_SEGMENTS", None)
    else:
      os.environ["TREX_SEGMENTS"] = self.old


def _engine(shape, cost_kind):
  import torch
  from trex_b200.sankoff import SankoffEngine

  n, A, L, _ = SHAPES[shape]
  adj, cost, states = _inputs(shape, cost_kind)
  eng = SankoffEngine(2 * n - 1, n, A, L)
  eng.set_adjacency(torch.from_numpy(adj)).set_leaf_states_u8(torch.from_numpy(states))
  return eng, torch.from_numpy(cost).cuda()


@pytest.mark.parametrize("seg", [0, 1])
@pytest.mark.parametrize("cost_kind", ["unit", "int"])
@pytest.mark.parametrize("shape", list(SHAPES))
def test_hard_path_bit_exact_at_bench_shapes(shape, cost_kind, seg):
  n, A, L, _ = SHAPES[shape]
  recon, dp_ref, total_ref, root_min_ref = _hard_oracle(shape, cost_kind)
  with _segments(seg):
    eng, cost = _engine(shape, cost_kind)
    eng.hard_forward(cost, want_argmin=True)
    eng.traceback(want_hist=True)
    dp = eng.export_dp().cpu().numpy()
    assert np.array_equal(dp, dp_ref)                                   # whole table incl. the implicit leaf rows
    assert np.array_equal(eng.root_min[0, :L].cpu().numpy(), root_min_ref)
    assert float(eng.total[0]) == float(root_min_ref.astype(np.float64).sum())
    anc = eng.anc[0, :, :L].cpu().numpy()
    assert np.array_equal(anc, recon[n:].astype(np.uint8))              # the reference's stack traceback
    hist = eng.edge_hist[0].cpu().numpy()
    assert hist.sum() == (2 * n - 2) * L
    # the reconstruction realises the score: sum over edges of C[parent state, child state] (tree.py:269-296)
    adj, cost_h, _ = _inputs(shape, cost_kind)
    assert float((hist * cost_h.astype(np.float64)).sum()) == float(eng.total[0])


@pytest.mark.parametrize("seg", [0, 1])
@pytest.mark.parametrize("cost_kind", ["unit", "int"])
def test_back_pointers_bit_exact_on_the_1024_leaf_tree(cost_kind, seg):
  """bt table of run_dp (sankoff.py:80-83) on the first 160 sites of the c5 slice vs the NumPy restatement."""
  import torch
  from trex_b200.sankoff import SankoffEngine

  n, A, _, _ = SHAPES["c5slice"]
  L = 160
  adj, cost, states = _inputs("c5slice", cost_kind)
  states = np.ascontiguousarray(states[:, :L])
  tabs = so.run_sankoff_tables(adj, cost, states, 2 * n - 1, A, n)
  with _segments(seg):
    eng = SankoffEngine(2 * n - 1, n, A, L)
    eng.set_adjacency(torch.from_numpy(adj)).set_leaf_states_u8(torch.from_numpy(states))
    eng.hard_forward(torch.from_numpy(cost), want_argmin=True)
    assert eng.wrote_pointer_words() == (cost_kind == "unit")
    arg = eng.argmin_table()[0, :, :, :L, :].permute(2, 0, 3, 1).cpu().numpy()  # (L, n_int, slot, A)
    assert np.array_equal(arg, tabs["argmin"][:, n:].astype(np.uint8))
    assert np.array_equal(eng.export_dp().cpu().numpy(), tabs["dp"])
    eng.traceback()
    want = so.traceback_topdown(tabs, 2 * n - 1, n)
    assert np.array_equal(eng.anc[0, :, :L].cpu().numpy(), want[n:].astype(np.uint8))


@pytest.mark.parametrize("seg", [0, 1])
@pytest.mark.parametrize("T", [1.0, 0.1])
@pytest.mark.parametrize("cost_kind", ["unit", "int"])
@pytest.mark.parametrize("shape", list(SHAPES))
def test_relaxed_path_vs_float64_oracle_at_bench_shapes(shape, cost_kind, T, seg):
  n, A, L, _ = SHAPES[shape]
  tot64, dC64, dp64, g64 = _soft_oracle(shape, cost_kind, T)
  with _segments(seg):
    eng, cost = _engine(shape, cost_kind)
    eng.soft_forward(cost, T)
    dp = eng.export_dp(soft=True).cpu().numpy()
    max_dp = np.abs(dp64[:, n:]).max()
    assert np.abs(dp[:, n:] - dp64[:, n:]).max() <= RTOL * max_dp
    root = dp64[:, -1, :]
    m = root.min(axis=1)
    per_site = m - T * np.log(np.exp(-(root - m[:, None]) / T).sum(axis=1))
    assert abs(per_site.sum() - tot64) <= 1e-9 * np.abs(per_site).sum()
    assert abs(float(eng.soft_total[0]) - tot64) <= RTOL * np.abs(per_site).sum()
    gtol = max(RTOL, 2.0 ** -24 * max_dp / T)  # float32 conditioning of the softmin weights (module docstring)
    # dense leaf gradient
    eng.soft_backward(cost, T, want_dleaf=True, want_dcost=True)
    gbar = eng.gbar[0, :, :, :L].permute(2, 0, 1).cpu().numpy()           # (L, n_int, A)
    assert np.abs(gbar - g64[:, n:]).max() <= gtol                         # marginals are O(1) probabilities
    dleaf = eng.dleaf[0, :, :, :L].permute(2, 0, 1).cpu().numpy()
    assert np.abs(dleaf - g64[:, :n]).max() <= RTOL
    dC = eng.dcost[0].cpu().numpy()
    assert np.abs(dC - dC64).max() <= RTOL * np.abs(dC64).max()
    # compact leaf gradient (what bench.py asks for: leaves are state codes)
    eng.gbar.zero_()
    eng.dcost.zero_()
    eng.soft_backward(cost, T, want_dleaf="compact", want_dcost=True)
    gbar2 = eng.gbar[0, :, :, :L].permute(2, 0, 1).cpu().numpy()
    assert np.abs(gbar2 - g64[:, n:]).max() <= gtol
    assert np.abs(eng.dcost[0].cpu().numpy() - dC64).max() <= RTOL * np.abs(dC64).max()
    obs = eng.dleaf_obs[0, :, :L].cpu().numpy()                            # (n, L)
    want_obs = g64[:, :n].sum(axis=2).T                                    # the one non-zero entry of each row
    _, _, states = _inputs(shape, cost_kind)
    picked = np.take_along_axis(g64[:, :n], states.T[:, :, None].astype(np.int64), axis=2)[:, :, 0].T
    assert np.abs(want_obs - picked).max() <= 1e-12                        # ... which is the observed-state entry
    assert np.abs(obs - want_obs).max() <= RTOL


@pytest.mark.parametrize("T", [1.0, 0.1])
def test_wide_reverse_kernel_build_vs_float64_oracle(T):
  """The 255-register build of the uniform-cost reverse kernel (sankoff_soft_bwd_uniform.cu: WIDE) is what the
  bench runs at 100k sites; the launcher picks it from ~76k sites on, so it is forced here on the c5 slice."""
  shape, cost_kind = "c5slice", "unit"
  n, A, L, _ = SHAPES[shape]
  tot64, dC64, dp64, g64 = _soft_oracle(shape, cost_kind, T)
  old = os.environ.get("TREX_BWD_WIDE")
  os.environ["TREX_BWD_WIDE"] = "1"
  try:
    eng, cost = _engine(shape, cost_kind)
    eng.soft_forward(cost, T)
    gtol = max(RTOL, 2.0 ** -24 * np.abs(dp64[:, n:]).max() / T)
    for mode in (True, "compact"):
      eng.gbar.zero_()
      eng.soft_backward(cost, T, want_dleaf=mode, want_dcost=True)
      gbar = eng.gbar[0, :, :, :L].permute(2, 0, 1).cpu().numpy()
      assert np.abs(gbar - g64[:, n:]).max() <= gtol
      assert np.abs(eng.dcost[0].cpu().numpy() - dC64).max() <= RTOL * np.abs(dC64).max()
    dleaf = eng.dleaf[0, :, :, :L].permute(2, 0, 1).cpu().numpy()
    assert np.abs(dleaf - g64[:, :n]).max() <= RTOL
    obs = eng.dleaf_obs[0, :, :L].cpu().numpy()
    assert np.abs(obs - g64[:, :n].sum(axis=2).T).max() <= RTOL
  finally:
    if old is None:
      os.environ.pop("TREX_BWD_WIDE", None)
    else:
      os.environ["TREX_BWD_WIDE"] = old
