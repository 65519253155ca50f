"""CPU tests: the C-ABI library loads and exports every symbol include/trex_b200.h declares,
argument validation fails loudly without touching a GPU, and the multi-rank host logic works
under gloo with world_size 2."""

import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="session")
def lib():
  from trex_b200 import _lib

  if not os.path.exists(_lib.LIB_PATH):
    subprocess.run(["make", "-j", "8"], cwd=os.path.join(ROOT, "trex_b200", "csrc"), check=True)
  return _lib.load()


def _header_symbols():
  src = open(os.path.join(ROOT, "include", "trex_b200.h")).read()
  return sorted(set(re.findall(r"TREX_API\s+[\w\s\*]+?\b(trex_\w+)\s*\(", src)))


def test_library_exports_every_declared_symbol(lib):
  from trex_b200 import _lib

  syms = _header_symbols()
  assert len(syms) >= 15
  out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True, check=True).stdout
  exported = set(re.findall(r" T (trex_\w+)", out))
  assert set(syms) <= exported, sorted(set(syms) - exported)
  assert set(syms) == set(_lib.SIGNATURES), sorted(set(syms) ^ set(_lib.SIGNATURES))
  for s in syms:
    assert getattr(lib, s) is not None


def test_host_side_queries_and_argument_errors(lib):
  assert lib.trex_version() >= 100
  assert lib.trex_sankoff_plan_words(2047, 1024) == 8 + 9 * 1023 + 2 * 128  # header, 9 per-node arrays, segment table
  assert lib.trex_tree_cost_workspace_bytes(127, 4000, 4) >= 127 * 127 * 8
  # null pointers / bad sizes are rejected on the host, before any CUDA call
  rc = lib.trex_sankoff_plan(None, 5, 3, 1, 1, None, None)
  assert rc == 1 and b"null pointer" in lib.trex_last_error()
  buf = (C.c_char * 64)()
  p = C.cast(buf, C.c_void_p)
  rc = lib.trex_sankoff_hard_fwd(p, 5, 3, 40, 10, 16, 1, p, p, p, None, None, None, 0, None)
  assert rc == 1 and b"n_states" in lib.trex_last_error()
  rc = lib.trex_sankoff_hard_fwd(p, 5, 3, 4, 10, 10, 1, p, p, p, None, None, None, 0, None)
  assert rc == 1 and b"site_stride" in lib.trex_last_error()
  rc = lib.trex_sankoff_soft_fwd(p, 5, 3, 4, 10, 16, 1, p, p, p, C.c_float(1.0), p, None, None)
  assert rc == 1 and b"exactly one" in lib.trex_last_error()
  rc = lib.trex_tree_cost_fwd_bwd(p, p, None, 2, 4, 4, 4, p, None, None, None, p, None)
  assert rc == 1 and b"mismatch" in lib.trex_last_error()


def test_missing_library_fails_loudly(monkeypatch):
  from trex_b200 import _lib

  monkeypatch.setattr(_lib, "_lib", None)
  monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libtrex_b200.so")
  with pytest.raises(ImportError, match="no CPU or PyTorch fallback"):
    _lib.load()


def test_engine_refuses_to_run_without_cuda():
  import torch
  from trex_b200.sankoff import SankoffEngine

  if torch.cuda.is_available():
    pytest.skip("CUDA present")
  with pytest.raises(RuntimeError, match="no CPU fallback"):
    SankoffEngine(5, 3, 2, 10)


def test_shard_range_covers_everything():
  from trex_b200.dist import shard_range

  for n, w in [(100_000, 8), (100_000, 3), (17, 4), (5, 8), (12_500, 1)]:
    got = [shard_range(n, w, r, align=16) for r in range(w)]
    assert got[0][0] == 0 and max(b for _, b in got) == n
    for (a0, b0), (a1, b1) in zip(got, got[1:]):
      assert b0 == a1 or (a1 == n and b1 == n)
    assert all(a % 16 == 0 or a == n for a, _ in got)
    assert sum(b - a for a, b in got) == n


def _worker(rank, world, port, q):
  import torch
  import torch.distributed as dist
  from oracle import sankoff_oracle as so
  from trex_b200 import synth
  from trex_b200.dist import PackedAllReduce, shard_range

  os.environ["MASTER_ADDR"] = "127.0.0.1"
  os.environ["MASTER_PORT"] = str(port)
  dist.init_process_group("gloo", rank=rank, world_size=world)
  rng = np.random.default_rng(0)  # identical inputs on every rank
  n, A, L = 8, 4, 101
  par = synth.random_join_parents(n, rng)
  adj = synth.parents_to_adjacency(par)
  states = synth.evolve(par, L, A, 0.2, rng)
  cost = synth.unit_cost(A)
  a, b = shard_range(L, world, rank, align=16)
  # each rank scores its own site shard (the oracle stands in for the GPU kernels on CPU)
  tabs = so.run_sankoff_tables(adj, cost, states[:n, a:b], 2 * n - 1, A, n)
  part = so.per_site_root_min(tabs["dp"]).astype(np.float64).sum()
  hist = np.full((A, A), float(rank + 1))
  par_red = PackedAllReduce([(), (A, A)], device="cpu")
  tot, h = par_red([torch.tensor(part, dtype=torch.float64), torch.from_numpy(hist)])
  full = so.per_site_root_min(so.run_sankoff_tables(adj, cost, states[:n], 2 * n - 1, A, n)["dp"]).sum()
  q.put((rank, float(tot), float(full), float(h.sum())))
  dist.destroy_process_group()


def test_site_sharded_reduction_world_size_2_gloo():
  import torch.multiprocessing as mp

  ctx = mp.get_context("spawn")
  q = ctx.Queue()
  port = 29500 + os.getpid() % 2000
  procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
  for p in procs:
    p.start()
  res = [q.get(timeout=120) for _ in procs]
  for p in procs:
    p.join(timeout=60)
    assert p.exitcode == 0
  for _, tot, full, hs in res:
    assert tot == full            # sum of shard scores == score of the whole alignment
    assert hs == 16 * (1 + 2)     # packed buffer reduced in the same single all-reduce


def test_xla_ffi_shim_type_checks_and_matches_the_python_binding():
  """The jax.ffi binding cannot run here (no jaxlib on any machine of this project: UNVERIFIED).
  What can be checked: the shim's handler bodies type-check against a minimal mock of xla/ffi/api/ffi.h
  (tests/mock_xla), every FFI target the Python side registers is defined by the shim, no handler returns F64
  (the reference runs with x64 disabled, src/trex/types.py:12) and attributes are NumPy scalars, not jax Arrays."""
  import re
  import shutil
  import subprocess

  root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
  shim = os.path.join(root, "trex_b200", "csrc", "xla_ffi_shim.cc")
  src = open(shim).read()
  py = open(os.path.join(root, "trex_b200", "jax_ffi.py")).read()
  gxx = shutil.which("g++")
  if gxx:
    cuda_inc = "/usr/local/cuda/include"
    cmd = [gxx, "-std=c++17", "-fsyntax-only", "-Wall", "-Werror", "-I", os.path.join(root, "tests", "mock_xla")]
    if os.path.isdir(cuda_inc):
      cmd += ["-I", cuda_inc]
      r = subprocess.run(cmd + [shim], capture_output=True, text=True)
      assert r.returncode == 0, r.stderr[-2000:]
  defined = set(re.findall(r"XLA_FFI_DEFINE_HANDLER_SYMBOL\((\w+),", src))
  block = py[py.index("for _name in"):py.index("jax.ffi.register_ffi_target")]
  registered = set(re.findall(r'"(trex_\w+_ffi)"', block))
  called = set(re.findall(r'ffi_call\(\s*"(\w+)"', py))
  assert registered == defined and called <= registered and registered
  assert "Buffer<ffi::F64>>" not in src and "float64" not in py.split('"""', 2)[2]
  assert not re.search(r"=\s*jnp\.(int|float)\d+\(", py)          # attributes: np.int32 / np.float32 ...
  assert 'vmap_method="expand_dims"' in py
