"""jax.ffi binding of libtrex_b200 — the drop-in surface ``north_star`` names.

Importable only where JAX (>= 0.4.38, ``jax.ffi``) and ``libtrex_b200_ffi.so`` (built by
``make -C trex_b200/csrc ffi``, needs jaxlib's headers) exist.  Neither exists in the build
container, so this module is UNTESTED there; the tested binding is the ctypes one
(``_lib.py`` + ``sankoff.py`` / ``tree.py``), which calls exactly the same C entry points.

Each public function keeps the reference signature (``src/trex/sankoff.py:114-124``,
``src/trex/tree.py:163-167, 212-217``), is ``jax.jit``/``vmap``/``grad`` compatible
(``vmap_method="sequential"``; integer outputs get no cotangent) and is a thin
``jax.ffi.ffi_call`` under ``jax.custom_vjp``.  There is no fallback: import fails without JAX.
"""

from __future__ import annotations

import ctypes
import functools
import os

import jax
import jax.numpy as jnp

_HERE = os.path.dirname(os.path.abspath(__file__))
_FFI_LIB = os.path.join(_HERE, "libtrex_b200_ffi.so")

_lib = ctypes.CDLL(_FFI_LIB)  # raises OSError when the shim was not built: no fallback
for _name in ("trex_run_sankoff_ffi", "trex_soft_fwd_ffi", "trex_soft_bwd_ffi", "trex_tree_cost_ffi"):
  jax.ffi.register_ffi_target(_name, jax.ffi.pycapsule(getattr(_lib, _name)), platform="CUDA")


def _pad16(n):
  return (n + 15) // 16 * 16


@functools.partial(jax.jit, static_argnames=("return_path", "n_all", "n_states", "n_leaves"))
def run_sankoff(adjacency_matrix, cost_matrix, sequences, n_all, n_states, n_leaves, *, return_path=False):
  """``trex.sankoff.run_sankoff`` — (recon (n_all,L) f32, dp (L,n_all,A) f32, total () f32)."""
  L = sequences.shape[1]
  f32 = jnp.float32
  recon, dp, total = jax.ffi.ffi_call(
    "trex_run_sankoff_ffi",
    (jax.ShapeDtypeStruct((n_all, L), f32), jax.ShapeDtypeStruct((L, n_all, n_states), f32),
     jax.ShapeDtypeStruct((1,), jnp.float64)),
    vmap_method="sequential",
  )(adjacency_matrix.astype(f32), cost_matrix.astype(f32), sequences.astype(f32),
    n_leaves=jnp.int32(n_leaves), return_path=bool(return_path))
  return recon, dp, total[0].astype(f32)


def _tree_cost_call(sequences, adjacency, cost_matrix):
  f32 = jnp.float32
  N, L, Q = sequences.shape
  kind = 0 if cost_matrix is None else (1 if cost_matrix.ndim == 1 else 2)
  c = jnp.zeros((1,), f32) if cost_matrix is None else cost_matrix.astype(f32)
  return jax.ffi.ffi_call(
    "trex_tree_cost_ffi",
    (jax.ShapeDtypeStruct((), f32), jax.ShapeDtypeStruct((N, L, Q), f32), jax.ShapeDtypeStruct((N, N), f32),
     jax.ShapeDtypeStruct(c.shape, f32)),
    vmap_method="sequential",
  )(sequences.astype(f32), adjacency.astype(f32), c, cost_kind=jnp.int32(kind))


@jax.custom_vjp
def compute_soft_cost(sequences, adjacency, cost_matrix=None):
  """``trex.tree.compute_soft_cost`` (tree.py:212-266)."""
  return _tree_cost_call(sequences, adjacency, cost_matrix)[0]


def _csc_fwd(sequences, adjacency, cost_matrix=None):
  cost, dS, dA, dC = _tree_cost_call(sequences, adjacency, cost_matrix)
  return cost, (dS, dA, None if cost_matrix is None else dC)


def _csc_bwd(res, g):
  dS, dA, dC = res
  return dS * g, dA * g, None if dC is None else dC * g


compute_soft_cost.defvjp(_csc_fwd, _csc_bwd)


def compute_surrogate_cost(sequences, adjacency):
  """``trex.tree.compute_surrogate_cost`` (tree.py:163-209)."""
  return compute_soft_cost(sequences, adjacency, None)


@functools.partial(jax.custom_vjp, nondiff_argnums=(3, 4, 5))
def soft_sankoff_cost(adjacency_matrix, cost_matrix, leaf_cost_native, n_leaves, n_sites, temperature):
  """Relaxed parsimony score; ``leaf_cost_native`` is [n_leaves][A][pad16(L)] f32."""
  return _soft_fwd(adjacency_matrix, cost_matrix, leaf_cost_native, n_leaves, n_sites, temperature)[2][0]


def _soft_fwd(adj, cost, leaf_cost, n_leaves, n_sites, temperature):
  n_all, A, stride = adj.shape[0], cost.shape[0], leaf_cost.shape[2]
  n_int = n_all - n_leaves
  return jax.ffi.ffi_call(
    "trex_soft_fwd_ffi",
    (jax.ShapeDtypeStruct((8 + 9 * n_int + 256,), jnp.int32), jax.ShapeDtypeStruct((n_int, A, stride), jnp.float32),
     jax.ShapeDtypeStruct((1,), jnp.float64)),
    vmap_method="sequential",
  )(adj.astype(jnp.float32), cost.astype(jnp.float32), leaf_cost, n_leaves=jnp.int32(n_leaves),
    n_sites=jnp.int64(n_sites), temperature=jnp.float32(temperature))


def _ssc_fwd(adj, cost, leaf_cost, n_leaves, n_sites, temperature):
  plan, dp, total = _soft_fwd(adj, cost, leaf_cost, n_leaves, n_sites, temperature)
  return total[0], (plan, dp, cost, leaf_cost, adj.shape[0])


def _ssc_bwd(n_leaves, n_sites, temperature, res, g):
  plan, dp, cost, leaf_cost, n_all = res
  A, stride = cost.shape[0], leaf_cost.shape[2]
  n_int = n_all - n_leaves
  _, dleaf, dcost = jax.ffi.ffi_call(
    "trex_soft_bwd_ffi",
    (jax.ShapeDtypeStruct((n_int, A, stride), jnp.float32), jax.ShapeDtypeStruct(leaf_cost.shape, jnp.float32),
     jax.ShapeDtypeStruct((A, A), jnp.float64)),
    vmap_method="sequential",
  )(plan, cost.astype(jnp.float32), leaf_cost, dp, jnp.asarray(g, jnp.float32).reshape(1),
    n_all=jnp.int32(n_all), n_leaves=jnp.int32(n_leaves), n_sites=jnp.int64(n_sites),
    temperature=jnp.float32(temperature))
  return jnp.zeros((n_all, n_all), jnp.float32), dcost.astype(cost.dtype), dleaf


soft_sankoff_cost.defvjp(_ssc_fwd, _ssc_bwd)
