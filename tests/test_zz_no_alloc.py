"""GPU test of the allocation-free entry point ``trex_compute_cost_noalloc`` (include/trex_b200.h; no counterpart in
the reference).  Collected LAST on purpose (``test_zz``):
a failure here must not hide any other test.

Through the C ABI, two calls in a row (a larger problem first, then a smaller one, so the second call meets the
first one's leftovers in the shared static scratch) are compared with the existing ``trex_compute_cost`` on the
same inputs — exact equality, the bar the compute_cost tests use (integer-valued sums) — and with the oracle."""

import numpy as np
import pytest

from oracle import tree_oracle as to

pytestmark = pytest.mark.gpu


def _inputs(N, L, Q, seed, cost_kind):
  import torch

  rng = np.random.default_rng(seed)
  states = rng.integers(0, Q, size=(N, L))
  S = np.eye(Q, dtype=np.float32)[states]                       # one-hot (N, L, Q)
  parents = np.array([rng.integers(i + 1, N) if i < N - 1 else N - 1 for i in range(N)])
  adj = np.zeros((N, N), dtype=np.float32)
  adj[np.arange(N), parents] = 1.0
  if cost_kind == "int":
    sub = rng.integers(0, 6, size=(Q, Q)).astype(np.float32)
  else:
    sub = (np.ones((Q, Q)) - np.eye(Q)).astype(np.float32)
  return (S, adj, sub), tuple(torch.from_numpy(x).cuda() for x in (S, adj, sub))


def _existing(St, At, Ct):
  """the existing entry point, caller-provided scratch"""
  import torch
  from trex_b200._lib import call, ptr, stream_ptr

  N, L, Q = St.shape
  st = torch.empty((N, L), dtype=torch.int32, device="cuda")
  par = torch.empty(N, dtype=torch.int32, device="cuda")
  acc = torch.empty(1, dtype=torch.float64, device="cuda")
  out = torch.empty((), dtype=torch.float32, device="cuda")
  call("trex_compute_cost", ptr(St), ptr(At), ptr(Ct), N, L, Q, ptr(st), ptr(par), ptr(acc), ptr(out), stream_ptr())
  return out


def _noalloc(St, At, Ct, out):
  from trex_b200._lib import call, ptr, stream_ptr

  N, L, Q = St.shape
  call("trex_compute_cost_noalloc", ptr(St), ptr(At), ptr(Ct), N, L, Q, ptr(out), stream_ptr())
  return out


def test_noalloc_matches_existing_entry_point_on_two_calls_in_a_row():
  import torch
  from trex_b200 import _lib

  lib = _lib.load()
  assert lib.trex_compute_cost_noalloc_scratch_bytes() == 256 + 4 * 4096 + 4 * (1 << 20)
  cases = [(127, 1000, 4, 11, "int"), (31, 100, 20, 12, "unit")]  # big first, then small: stale scratch contents
  prepared = [_inputs(*c) for c in cases]
  want = [_existing(*dev) for _, dev in prepared]
  outs = [torch.full((), -1.0, dtype=torch.float32, device="cuda") for _ in cases]
  _noalloc(*prepared[0][1], torch.empty((), dtype=torch.float32, device="cuda"))  # loads the module, finds the symbol
  torch.cuda.synchronize()
  free0, _ = torch.cuda.mem_get_info()
  alloc0 = torch.cuda.memory_allocated()
  for (_, dev), out in zip(prepared, outs):  # two calls in a row, same stream, nothing in between
    _noalloc(*dev, out)
  torch.cuda.synchronize()
  free1, _ = torch.cuda.mem_get_info()
  assert torch.cuda.memory_allocated() == alloc0   # torch's allocator was not touched
  assert free1 >= free0                            # and the driver handed out no device memory either
  for (host, _), out, w in zip(prepared, outs, want):
    assert float(out) == float(w)                                  # exact: same kernels, integer-valued sum
    S, adj, sub = host
    assert float(out) == float(to.compute_cost(S, adj, sub))      # and the oracle of tree.py:269-296
  # a third call repeating the first problem gives the first answer again (accumulator cleared per call)
  again = _noalloc(*prepared[0][1], torch.empty((), dtype=torch.float32, device="cuda"))
  assert float(again) == float(want[0])


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
  import torch
  from trex_b200._lib import TrexError, call, ptr, stream_ptr

  small = torch.zeros(64, dtype=torch.float32, device="cuda")
  out = torch.empty((), dtype=torch.float32, device="cuda")
  with pytest.raises(TrexError, match="TREX_NOALLOC_MAX_CELLS"):   # 1025 * 1024 cells > 2^20
    call("trex_compute_cost_noalloc", ptr(small), ptr(small), ptr(small), 1025, 1024, 4, ptr(out), stream_ptr())
  with pytest.raises(TrexError, match="TREX_NOALLOC_MAX_NODES"):
    call("trex_compute_cost_noalloc", ptr(small), ptr(small), ptr(small), 4097, 1, 4, ptr(out), stream_ptr())
  torch.cuda.synchronize()  # nothing was launched: the device is still healthy
  (_, dev) = _inputs(9, 17, 3, 5, "int")
  assert float(_noalloc(*dev, out)) == float(_existing(*dev))
