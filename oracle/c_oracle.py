"""TEST INFRASTRUCTURE ONLY — ctypes loader of the plain-C restatement (``sankoff_oracle.c``).

Used by tests (second checker) and by ``bench.py``'s CPU-baseline / ``--impl reference`` legs.
"""

from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(_HERE, "_build", "liboracle.so")
_lib = None


def load():
  global _lib
  if _lib is None:
    if not os.path.exists(_PATH):
      subprocess.run(["make"], cwd=_HERE, check=True)
    _lib = C.CDLL(_PATH)
    _lib.oracle_num_threads.restype = C.c_int
  return _lib


def num_threads() -> int:
  return int(load().oracle_num_threads())


def _p(a):
  return None if a is None else a.ctypes.data_as(C.c_void_p)


def run_sankoff(adj, cost, seqs, n_all, n_states, n_leaves, return_path=False, want_dp=True):
  """Same contract as the reference's run_sankoff; returns (recon, dp | None, total, root_min)."""
  adj = np.ascontiguousarray(adj, dtype=np.float32)
  cost = np.ascontiguousarray(cost, dtype=np.float32)
  seqs = np.ascontiguousarray(seqs, dtype=np.float32)
  L = seqs.shape[1]
  recon = np.empty((n_all, L), dtype=np.float32)
  dp = np.empty((L, n_all, n_states), dtype=np.float32) if want_dp else None
  root_min = np.empty(L, dtype=np.float32)
  total = C.c_float(0)
  load().oracle_run_sankoff(_p(adj), _p(cost), _p(seqs), C.c_int(n_all), C.c_int(n_states),
                            C.c_int(n_leaves), C.c_long(L), C.c_int(1 if return_path else 0),
                            _p(recon), _p(dp), _p(root_min), C.byref(total))
  return recon, dp, np.float32(total.value), root_min


def soft_sankoff(adj, cost, seqs, n_all, n_states, n_leaves, T, want_gbar=False):
  """Relaxed forward+reverse (restatement, not reference code): (total f64, dC f64, gbar|None)."""
  adj = np.ascontiguousarray(adj, dtype=np.float32)
  cost = np.ascontiguousarray(cost, dtype=np.float32)
  seqs = np.ascontiguousarray(seqs, dtype=np.float32)
  L = seqs.shape[1]
  dC = np.zeros((n_states, n_states), dtype=np.float64)
  gbar = np.empty((L, n_all, n_states), dtype=np.float32) if want_gbar else None
  total = C.c_double(0)
  load().oracle_soft_sankoff(_p(adj), _p(cost), _p(seqs), C.c_int(n_all), C.c_int(n_states),
                             C.c_int(n_leaves), C.c_long(L), C.c_float(T), C.byref(total), _p(dC),
                             _p(gbar))
  return total.value, dC, gbar


def soft_sankoff_f64(adj, cost, n_all, n_states, n_leaves, T, seqs=None, leaf_cost=None, want_dp=True,
                     want_gbar=True):
  """float64 instance of the same restatement — the checker of the 1e-5 parity tests at the benchmark
  shapes (the NumPy version needs minutes there).  Leaves: ``seqs`` (rows, L) state codes or
  ``leaf_cost`` (n_leaves, L, A).  Returns ``(total, dC (A,A), dp (L,n_all,A) | None, gbar (L,n_all,A) | None)``."""
  adj = np.ascontiguousarray(adj, dtype=np.float32)
  cost = np.ascontiguousarray(cost, dtype=np.float32)
  assert (seqs is None) != (leaf_cost is None)
  if seqs is not None:
    seqs = np.ascontiguousarray(seqs, dtype=np.float32)
    L = seqs.shape[1]
  else:
    leaf_cost = np.ascontiguousarray(leaf_cost, dtype=np.float64)
    L = leaf_cost.shape[1]
  dC = np.zeros((n_states, n_states), dtype=np.float64)
  dp = np.empty((L, n_all, n_states), dtype=np.float64) if want_dp else None
  gbar = np.empty((L, n_all, n_states), dtype=np.float64) if want_gbar else None
  total = C.c_double(0)
  load().oracle_soft_sankoff_f64(_p(adj), _p(cost), _p(seqs), _p(leaf_cost), C.c_int(n_all), C.c_int(n_states),
                                 C.c_int(n_leaves), C.c_long(L), C.c_double(T), C.byref(total), _p(dC),
                                 _p(dp), _p(gbar))
  return total.value, dC, dp, gbar


def set_num_threads(n: int) -> int:
  """OpenMP thread count of the C legs (``torch.distributed.run`` exports OMP_NUM_THREADS=1)."""
  load().oracle_set_num_threads(C.c_int(int(n)))
  return num_threads()
