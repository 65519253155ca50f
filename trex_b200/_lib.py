"""ctypes binding of ``libtrex_b200.so`` (C ABI declared in ``include/trex_b200.h``).

There is no fallback: if the shared library is missing or a call fails, an exception is raised.
"""

from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# TREX_B200_LIB: load another build of the same library (kernel-variant experiments); default = the in-tree build
LIB_PATH = os.environ.get("TREX_B200_LIB") or os.path.join(_HERE, "libtrex_b200.so")

MODE_AUTO = 0
MODE_GENERAL = 1
MODE_COMPACT_ARG = 0x100  # OR-ed into mode: pointer words instead of the byte table when the cost is uniform
PLAN_HAS_SENTINEL_CHILD = 1
PLAN_NOT_A_TREE = 2
PLAN_SHARED_LEAF = 4
PLAN_HDR = 8

_vp, _i, _i64, _f = C.c_void_p, C.c_int, C.c_int64, C.c_float

# name -> (restype, argtypes); mirrors include/trex_b200.h one to one
SIGNATURES = {
  "trex_last_error": (C.c_char_p, []),
  "trex_version": (_i, []),
  "trex_kernel_launches": (C.c_ulonglong, []),
  "trex_sankoff_plan_words": (C.c_size_t, [_i, _i]),
  "trex_sankoff_plan": (_i, [_vp, _i, _i, _i, _i, _vp, _vp]),
  "trex_sankoff_plan_for_sites": (_i, [_vp, _i, _i, _i, _i, _i64, _vp, _vp]),
  "trex_encode_leaf_states": (_i, [_vp, _i, _i64, _i64, _i, _vp, _i64, _vp]),
  "trex_sankoff_hard_fwd": (_i, [_vp, _i, _i, _i, _i64, _i64, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp]),
  "trex_sankoff_traceback": (_i, [_vp, _i, _i, _i, _i64, _i64, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
  "trex_sankoff_traceback_cost": (_i, [_vp, _i, _i, _i, _i64, _i64, _i, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp]),
  "trex_sankoff_export_dp": (_i, [_i, _i, _i, _i64, _i64, _vp, _vp, _vp, _vp]),
  "trex_sankoff_export_recon": (_i, [_i, _i, _i64, _i64, _vp, _i64, _vp, _vp, _vp, _vp]),
  "trex_sankoff_export_bt": (_i, [_vp, _i, _i, _i, _i64, _i64, _vp, _vp, _vp]),
  "trex_sankoff_backtrack_ref": (_i, [_vp, _i, _i, _i, _i64, _vp, _vp, _vp]),
  "trex_cast_f64_to_f32": (_i, [_vp, _vp, _i64, _vp]),
  "trex_sankoff_soft_fwd": (_i, [_vp, _i, _i, _i, _i64, _i64, _i, _vp, _vp, _vp, _f, _vp, _vp, _vp]),
  "trex_sankoff_soft_bwd": (_i, [_vp, _i, _i, _i, _i64, _i64, _i, _vp, _vp, _vp, _f, _vp, _vp, _vp, _vp, _vp, _vp]),
  "trex_sankoff_soft_bwd_compact": (_i, [_vp, _i, _i, _i, _i64, _i64, _i, _vp, _vp, _f, _vp, _vp, _vp, _vp, _vp, _vp]),
  "trex_tree_cost_workspace_bytes": (C.c_size_t, [_i, _i64, _i]),
  "trex_tree_cost_fwd_bwd": (_i, [_vp, _vp, _vp, _i, _i, _i64, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
  "trex_update_seq_fwd": (_i, [_vp, _i, _i64, _i, _f, _vp, _vp]),
  "trex_update_seq_bwd": (_i, [_vp, _vp, _i, _i64, _i, _f, _vp, _vp]),
  "trex_update_tree_fwd": (_i, [_vp, _vp, _vp, _f, _i, _i, _vp, _vp]),
  "trex_update_tree_bwd": (_i, [_vp, _vp, _vp, _f, _i, _i, _vp, _vp]),
  "trex_graph_constraint": (_i, [_vp, _i, _f, _vp, _vp, _vp, _vp]),
  "trex_discretize_topology": (_i, [_vp, _i, _vp, _vp, _vp]),
  "trex_compute_cost": (_i, [_vp, _vp, _vp, _i, _i64, _i, _vp, _vp, _vp, _vp, _vp]),
  "trex_compute_cost_noalloc_scratch_bytes": (C.c_size_t, []),
  "trex_compute_cost_noalloc": (_i, [_vp, _vp, _vp, _i, _i64, _i, _vp, _vp]),
  "trex_landscape_loss_fwd_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i64, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
  "trex_tree_adjacency": (_i, [_i, _i, _vp, _vp]),
  "trex_generate_groundtruth_workspace_bytes": (C.c_size_t, [_i, _i, _i64]),
  "trex_generate_groundtruth": (_i, [_i, _i, _i, _i64, _i64, C.c_uint64, _vp, _vp, _vp, _vp, _vp, _vp]),
  "trex_nk_landscape": (_i, [_i64, _i, _i, C.c_uint64, _vp, _vp, _vp]),
  "trex_nk_fitness": (_i, [_vp, _i64, _i, _i64, _i, _i, _vp, _vp, _vp, _vp, _vp]),
  "trex_nk_generate_tree_workspace_bytes": (C.c_size_t, [_i, _i64]),
  "trex_nk_generate_tree": (_i, [_vp, _i, _vp, _i64, _i64, _i, _i, _vp, _vp, _f, _f, _f, _i, C.c_uint64, _vp, _vp, _vp, _vp,
                                 _vp, _vp, _vp, _vp, _vp]),
  "trex_parse_alignment_host": (_i, [C.c_char_p, C.c_size_t, _i, C.c_char_p, _vp, _vp, _vp, _i64, _vp, C.c_size_t]),
  "trex_parse_newick_host": (_i, [C.c_char_p, C.c_size_t, C.c_char_p, _vp, _vp, _vp, _vp, C.c_size_t]),
  "trex_fit_ancestors_workspace_bytes": (C.c_size_t, [_i, _i, _i64, _i]),
  "trex_fit_ancestors": (_i, [_vp, _vp, _i, _i, _i64, _i, _vp, _vp, _vp, _i, _f, _f, _i, _i64, _vp, _vp, _vp, _vp]),
}

_lib = None


class TrexError(RuntimeError):
  """A libtrex_b200 entry point returned a non-zero status."""


def load() -> C.CDLL:
  global _lib
  if _lib is None:
    if not os.path.exists(LIB_PATH):
      raise ImportError(
        f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "or `make -C trex_b200/csrc`. trex_b200 has no CPU or PyTorch fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
      fn = getattr(lib, name)  # AttributeError if the library does not export the symbol
      fn.restype = res
      fn.argtypes = args
    _lib = lib
  return _lib


def call(name: str, *args) -> None:
  lib = load()
  rc = getattr(lib, name)(*args)
  if rc != 0:
    raise TrexError(f"{name} failed ({rc}): {lib.trex_last_error().decode()}")


def ptr(t):
  """Device pointer of a torch tensor (None -> NULL)."""
  return None if t is None else C.c_void_p(t.data_ptr())


def stream_ptr():
  import torch

  return C.c_void_p(torch.cuda.current_stream().cuda_stream)
