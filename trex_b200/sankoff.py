"""Host-side mirror of the reference's ``trex.sankoff`` (``/root/reference/src/trex/sankoff.py``)
on top of the C ABI of ``libtrex_b200.so``.  Same names, argument order and return values as
the reference; arrays are torch CUDA tensors (PyTorch is only the owner of device memory and
streams here — every computation below is a call into the library).

Two levels:
* ``run_sankoff`` / ``run_dp`` / ``vectorized_dp`` / ``backtrack_sankoff_jit`` — drop-in
  signatures, reference array layouts (``dp (L, n_all, A)``, ``recon (n_all, L)``).
* ``SankoffEngine`` — the native, allocation-free path (site-contiguous tables resident in HBM)
  used by ``bench.py`` and by callers that do not want the layout exports.
"""

from __future__ import annotations

import torch

from . import _lib
from ._lib import call, ptr, stream_ptr


def pad_sites(n_sites: int) -> int:
  """Row pitch of the native tables: multiple of 16 sites."""
  return (int(n_sites) + 15) // 16 * 16


class SankoffEngine:
  """Buffers + kernels for one (n_all, n_leaves, A, L, batch) problem shape.

  Native layouts (see include/trex_b200.h): ``leaf`` u8 [n_leaves][stride];
  ``dp``/``gbar`` f32 [batch][n_int][A][stride]; ``argmin`` u8 [batch][n_int][A][stride][2] (slot last);
  ``anc`` u8 [batch][n_int][stride].  Buffers are created lazily, once.
  """

  def __init__(self, n_all: int, n_leaves: int, n_states: int, n_sites: int, batch: int = 1,
               device=None):
    if not torch.cuda.is_available():
      raise RuntimeError("trex_b200 needs a CUDA device; there is no CPU fallback")
    _lib.load()
    self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    self.n_all, self.n_leaves, self.A = int(n_all), int(n_leaves), int(n_states)
    self.n_int = self.n_all - self.n_leaves
    self.L, self.batch = int(n_sites), int(batch)
    self.stride = pad_sites(n_sites)
    self.plan_words = _lib.load().trex_sankoff_plan_words(self.n_all, self.n_leaves)
    self._bufs = {}

  # -- buffers ------------------------------------------------------------------------------
  def buf(self, name, shape, dtype, zero=False):
    parent = getattr(self, "_parent", None)
    if parent is not None and name not in self._own:
      return parent.buf(name, shape, dtype, zero)
    t = self._bufs.get(name)
    if t is None:
      t = (torch.zeros if zero else torch.empty)(shape, dtype=dtype, device=self.device)
      self._bufs[name] = t
    return t

  @property
  def plan(self):
    return self.buf("plan", (self.batch, self.plan_words), torch.int32)

  @property
  def leaf(self):
    return self.buf("leaf", (self.n_leaves, self.stride), torch.uint8, zero=True)

  @property
  def dp(self):
    return self.buf("dp", (self.batch, self.n_int, self.A, self.stride), torch.float32)

  @property
  def argmin(self):
    return self.buf("argmin", (self.batch, self.n_int, self.A, self.stride, 2), torch.uint8)

  @property
  def anc(self):
    return self.buf("anc", (self.batch, self.n_int, self.stride), torch.uint8)

  @property
  def root_min(self):
    return self.buf("root_min", (self.batch, self.stride), torch.float32)

  @property
  def total(self):
    return self.buf("total", (self.batch,), torch.float64)

  @property
  def soft_dp(self):
    return self.buf("soft_dp", (self.batch, self.n_int, self.A, self.stride), torch.float32)

  @property
  def soft_total(self):
    return self.buf("soft_total", (self.batch,), torch.float64)

  @property
  def gbar(self):
    return self.buf("gbar", (self.batch, self.n_int, self.A, self.stride), torch.float32)

  @property
  def dleaf(self):
    return self.buf("dleaf", (self.batch, self.n_leaves, self.A, self.stride), torch.float32, zero=True)

  @property
  def dleaf_obs(self):
    return self.buf("dleaf_obs", (self.batch, self.n_leaves, self.stride), torch.float32, zero=True)

  @property
  def dcost(self):
    return self.buf("dcost", (self.batch, self.A, self.A), torch.float64)

  @property
  def edge_hist(self):
    return self.buf("edge_hist", (self.batch, self.A, self.A), torch.int64)

  # -- site blocks ----------------------------------------------------------------------------
  _SITE_TABLES = ("leaf", "dp", "argmin", "anc", "root_min", "soft_dp", "gbar", "dleaf", "dleaf_obs")

  def site_block(self, lo: int, hi: int):
    """An engine over sites [lo, hi) of this one's tables (views, no copies; ``lo`` a multiple of 64).

    Sites are independent under Sankoff (sankoff.py:97), so a long alignment can be streamed in
    blocks: the host->device copy of block b+1 overlaps the kernels of block b (``bench.py`` e2e leg).
    The plan is shared; totals / dC / edge histogram are per block and add up over the blocks."""
    lo, hi = int(lo), int(hi)
    assert 0 <= lo < hi <= self.L and lo % 64 == 0, (lo, hi)
    blk = SankoffEngine.__new__(SankoffEngine)
    blk.__dict__.update({k: v for k, v in self.__dict__.items() if k != "_bufs"})
    blk.L = hi - lo
    blk._pad_owner = hi == self.L  # only the last block may touch the pad columns of the row pitch
    blk._bufs = {"plan": self.plan}
    for name in self._SITE_TABLES:
      t = getattr(self, name)
      blk._bufs[name] = t[..., lo:, :] if name == "argmin" else t[..., lo:]
    return blk

  def twin(self, own=("leaf", "anc")):
    """A second engine of the same shape that owns only the tables named in ``own`` and shares every other
    one (and the plan) with this engine — the second buffer slot of a double-buffered pipeline."""
    t = SankoffEngine.__new__(SankoffEngine)
    t.__dict__.update({k: v for k, v in self.__dict__.items() if k != "_bufs"})
    t._bufs, t._parent, t._own, t._pad_owner = {}, self, set(own), True
    return t

  def site_blocks(self, n_blocks: int):
    per = max(64, ((self.L + n_blocks - 1) // n_blocks + 63) // 64 * 64)
    return [self.site_block(lo, min(self.L, lo + per)) for lo in range(0, self.L, per)]

  # -- inputs -------------------------------------------------------------------------------
  def set_adjacency(self, adj, zero_root_selfloop=True):
    """adj: (n_all, n_all) or (batch, n_all, n_all), any numeric dtype (cast to f32 as
    sankoff.py:144 does)."""
    adj = adj.to(device=self.device, dtype=torch.float32).contiguous()
    if adj.dim() == 2:
      adj = adj.unsqueeze(0).expand(self.batch, -1, -1).contiguous() if self.batch > 1 else adj.unsqueeze(0)
    assert adj.shape == (self.batch, self.n_all, self.n_all), adj.shape
    call("trex_sankoff_plan_for_sites", ptr(adj), self.n_all, self.n_leaves, self.batch,
         1 if zero_root_selfloop else 0, self.L, ptr(self.plan), stream_ptr())
    return self

  def set_leaf_states_u8(self, states):
    """states: (n_leaves, L) uint8 codes (>= A means 'no state')."""
    self.leaf[:, : self.L].copy_(states[: self.n_leaves], non_blocking=True)
    if getattr(self, "_pad_owner", True) and self.leaf.shape[1] > self.L:
      self.leaf[:, self.L:].fill_(255)
    return self

  def set_sequences_f32(self, seqs):
    """seqs: (>= n_leaves, L) float32 as the reference takes them (sankoff.py:145)."""
    seqs = seqs.to(device=self.device, dtype=torch.float32).contiguous()
    call("trex_encode_leaf_states", ptr(seqs), self.n_leaves, self.L, seqs.stride(0), self.A,
         ptr(self.leaf), self.stride, stream_ptr())
    return self

  # -- kernels ------------------------------------------------------------------------------
  def hard_forward(self, cost, want_argmin=True, mode=_lib.MODE_AUTO, compact_arg=True):
    """``compact_arg``: let the library write the back-pointers as pointer words when the cost matrix is uniform
    (``TREX_MODE_COMPACT_ARG``, include/trex_b200.h: 8 instead of 2A bytes per node and site); ``traceback`` reads
    either form, ``export_bt`` needs ``compact_arg=False``."""
    cost = cost.to(device=self.device, dtype=torch.float32).contiguous()
    if want_argmin:
      mode = mode | (_lib.MODE_COMPACT_ARG if compact_arg else 0)
      self._arg_cost, self._arg_mode = cost, mode  # the traceback repeats the forward's decision on the same matrix
    call("trex_sankoff_hard_fwd", ptr(self.plan), self.n_all, self.n_leaves, self.A, self.L,
         self.stride, self.batch, ptr(self.leaf), ptr(cost), ptr(self.dp),
         ptr(self.argmin) if want_argmin else None, ptr(self.root_min), ptr(self.total), mode,
         stream_ptr())
    return self.total

  def traceback(self, want_hist=False):
    cost, mode = getattr(self, "_arg_cost", None), getattr(self, "_arg_mode", _lib.MODE_AUTO)
    call("trex_sankoff_traceback_cost", ptr(self.plan), self.n_all, self.n_leaves, self.A, self.L,
         self.stride, self.batch, ptr(self.leaf), ptr(cost), mode, ptr(self.dp), ptr(self.argmin), ptr(self.anc),
         ptr(self.edge_hist) if want_hist else None, stream_ptr())
    return self.anc

  def soft_forward(self, cost, temperature, leaf_cost=None):
    cost = cost.to(device=self.device, dtype=torch.float32).contiguous()
    call("trex_sankoff_soft_fwd", ptr(self.plan), self.n_all, self.n_leaves, self.A, self.L,
         self.stride, self.batch, None if leaf_cost is not None else ptr(self.leaf),
         ptr(leaf_cost), ptr(cost), float(temperature), ptr(self.soft_dp), ptr(self.soft_total),
         stream_ptr())
    return self.soft_total

  def soft_backward(self, cost, temperature, leaf_cost=None, gscale=None, want_dleaf=True,
                    want_dcost=True):
    """``want_dleaf``: True = dense ``dleaf`` [leaf][A][site]; "compact" = ``dleaf_obs`` [leaf][site]
    (leaves given as state codes only: the entry at the observed state, every other one is exactly 0)."""
    cost = cost.to(device=self.device, dtype=torch.float32).contiguous()
    if want_dleaf == "compact":
      if leaf_cost is not None:
        raise ValueError("the compact leaf gradient is defined for leaves given as state codes")
      call("trex_sankoff_soft_bwd_compact", ptr(self.plan), self.n_all, self.n_leaves, self.A, self.L,
           self.stride, self.batch, ptr(self.leaf), ptr(cost), float(temperature), ptr(self.soft_dp),
           ptr(gscale), ptr(self.gbar), ptr(self.dleaf_obs), ptr(self.dcost) if want_dcost else None,
           stream_ptr())
      return self.gbar
    call("trex_sankoff_soft_bwd", ptr(self.plan), self.n_all, self.n_leaves, self.A, self.L,
         self.stride, self.batch, None if leaf_cost is not None else ptr(self.leaf),
         ptr(leaf_cost), ptr(cost), float(temperature), ptr(self.soft_dp), ptr(gscale),
         ptr(self.gbar), ptr(self.dleaf) if want_dleaf else None,
         ptr(self.dcost) if want_dcost else None, stream_ptr())
    return self.gbar

  # -- exports to the reference layouts -------------------------------------------------------
  def export_dp(self, tree=0, soft=False):
    out = torch.empty((self.L, self.n_all, self.A), dtype=torch.float32, device=self.device)
    src = (self.soft_dp if soft else self.dp)[tree]
    call("trex_sankoff_export_dp", self.n_all, self.n_leaves, self.A, self.L, self.stride,
         ptr(self.leaf), ptr(src), ptr(out), stream_ptr())
    return out

  def export_recon(self, leaf_seqs_f32=None, with_path=True, tree=0):
    out = torch.empty((self.n_all, self.L), dtype=torch.float32, device=self.device)
    call("trex_sankoff_export_recon", self.n_all, self.n_leaves, self.L, self.stride,
         ptr(leaf_seqs_f32), 0 if leaf_seqs_f32 is None else leaf_seqs_f32.stride(0),
         ptr(self.leaf), ptr(self.anc[tree]) if with_path else None, ptr(out), stream_ptr())
    return out

  def wrote_pointer_words(self) -> bool:
    """Whether the last ``hard_forward(want_argmin=True)`` left pointer words in ``argmin`` — the device-side test
    of include/trex_b200.h (TREX_MODE_COMPACT_ARG) repeated on the host; copies the cost matrix back (a sync): for
    exports and tests, not for the hot path."""
    mode = getattr(self, "_arg_mode", 0)
    if not (mode & _lib.MODE_COMPACT_ARG) or (mode & 0xff) != _lib.MODE_AUTO or not 4 <= self.A <= 20:
      return False
    c = self._arg_cost.cpu()
    a = float(c[0, 1])
    want = a * (1.0 - torch.eye(self.A))
    return bool(torch.equal(c, want)) and 0.0 <= a < 65536.0 and a == int(a) and a * self.n_all < 262000.0

  def argmin_table(self):
    """The back-pointers as the byte table u8 [batch][n_int][A][stride][2], whichever form ``hard_forward`` wrote
    (pointer words are decoded with torch ops into a new tensor)."""
    if not self.wrote_pointer_words():
      return self.argmin
    q = self.argmin[:, :, :4].to(torch.int64)                       # rows 0-3: [B][n_int][4][stride][2] bytes
    q16 = q[..., 0] | (q[..., 1] << 8)                              # [B][n_int][4][stride]
    words = torch.stack([q16[:, :, 0] | (q16[:, :, 1] << 16), q16[:, :, 2] | (q16[:, :, 3] << 16)], dim=-1)
    words = words[:, :, None]                                       # [B][n_int][1][stride][2]
    s = torch.arange(self.A, device=self.device).view(1, 1, self.A, 1, 1)
    return torch.where(((words >> (5 + s)) & 1).bool(), s, words & 31).to(torch.uint8)

  def export_bt(self, tree=0):
    out = torch.empty((self.L, self.n_all, self.A, 4), dtype=torch.float32, device=self.device)
    table = self.argmin_table()[tree].contiguous()
    call("trex_sankoff_export_bt", ptr(self.plan[tree]), self.n_all, self.n_leaves, self.A,
         self.L, self.stride, ptr(table), ptr(out), stream_ptr())
    return out

  def plan_status(self, tree=0):
    """(n_reach, status bits) — forces a device sync; for diagnostics and tests."""
    hdr = self.plan[tree, : _lib.PLAN_HDR].cpu()
    return int(hdr[1]), int(hdr[3])


# ---------------------------------------------------------------------------------------------
# Drop-in functions (reference signatures)
# ---------------------------------------------------------------------------------------------


def run_sankoff(adjacency_matrix, cost_matrix, sequences, n_all, n_states, n_leaves, *,
                return_path=False):
  """``trex.sankoff.run_sankoff`` (sankoff.py:114-188): returns
  ``(reconstructed (n_all, L) f32, dp (L, n_all, A) f32, total_cost f32 scalar)``."""
  sequences = torch.as_tensor(sequences)
  adjacency_matrix = torch.as_tensor(adjacency_matrix)
  cost_matrix = torch.as_tensor(cost_matrix)
  L = sequences.shape[1]
  eng = SankoffEngine(n_all, n_leaves, n_states, L)
  seqs = sequences.to(device=eng.device, dtype=torch.float32).contiguous()
  eng.set_adjacency(adjacency_matrix, zero_root_selfloop=True)
  eng.set_sequences_f32(seqs)
  eng.hard_forward(cost_matrix, want_argmin=return_path)
  if return_path:
    eng.traceback()
  recon = eng.export_recon(seqs, with_path=return_path)
  dp = eng.export_dp()
  total = eng.total[0].to(torch.float32)  # the reference's f32 sum (sankoff.py:187)
  return recon, dp, total


def vectorized_dp(adjacency_matrix, dynamic_programming_table, backtracking_table, sequences,
                  cost_matrix):
  """``trex.sankoff.vectorized_dp`` (sankoff.py:97): dp (L, n_all, A), bt (L, n_all, A, 4),
  sequences (rows, L).  The incoming tables must be the reference's initial fill (1e5 / 0)."""
  dpt = torch.as_tensor(dynamic_programming_table)
  L, n_all, A = dpt.shape
  n_leaves = (n_all + 1) // 2  # sankoff.py:45-46
  eng = SankoffEngine(n_all, n_leaves, A, L)
  eng.set_adjacency(torch.as_tensor(adjacency_matrix), zero_root_selfloop=False)
  eng.set_sequences_f32(torch.as_tensor(sequences))
  eng.hard_forward(torch.as_tensor(cost_matrix), want_argmin=True, compact_arg=False)
  return eng.export_dp(), eng.export_bt()


def run_dp(adjacency_matrix, dynamic_programming_table, backtracking_table, sequences,
           cost_matrix):
  """``trex.sankoff.run_dp`` (sankoff.py:24-94) for one site: dp (n_all, A), bt (n_all, A, 4)."""
  dpt = torch.as_tensor(dynamic_programming_table)
  seq = torch.as_tensor(sequences).reshape(-1, 1)
  dp, bt = vectorized_dp(adjacency_matrix, dpt.unsqueeze(0), None, seq, cost_matrix)
  return dp[0], bt[0]


def backtrack_sankoff_jit(root_node, root_state, backtracking_table, n_all, n_leaves):
  """``trex.sankoff.backtrack_sankoff_jit`` (sankoff.py:191-267) for one site (or a leading
  site axis): backtracking_table (n_all, A, 4) f32 -> reconstructed chars (n_all,) int32."""
  _lib.load()
  bt = torch.as_tensor(backtracking_table).to(device="cuda", dtype=torch.float32).contiguous()
  single = bt.dim() == 3
  if single:
    bt = bt.unsqueeze(0)
  n_sites, _, A, _ = bt.shape
  rn = torch.as_tensor(root_node).to(device=bt.device, dtype=torch.int32).reshape(1)
  rs = torch.as_tensor(root_state).to(device=bt.device, dtype=torch.int32).reshape(-1).contiguous()
  out = torch.empty((n_sites, n_all), dtype=torch.int32, device=bt.device)
  call("trex_sankoff_backtrack_ref", ptr(bt), int(n_all), int(n_leaves), int(A), int(n_sites),
       ptr(rn), ptr(rs), ptr(out), stream_ptr())
  return out[0] if single else out
