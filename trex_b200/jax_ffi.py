"""jax.ffi binding of libtrex_b200 — the drop-in surface ``north_star`` names.

Importable only where JAX (>= 0.4.38, ``jax.ffi``) and ``libtrex_b200_ffi.so`` (built by
``make -C trex_b200/csrc ffi``, needs jaxlib's headers) exist.  Neither exists in the build
container, so this module is UNTESTED there; the tested binding is the ctypes one
(``_lib.py`` + ``sankoff.py`` / ``tree.py``), which calls exactly the same C entry points.

Each public function keeps the reference signature (``src/trex/sankoff.py:114-124``,
``src/trex/tree.py:163-167, 212-217``), is ``jax.jit``/``vmap``/``grad`` compatible and is a thin
``jax.ffi.ffi_call`` under ``jax.custom_vjp``.  There is no fallback: import fails without JAX.

Conventions (see the header of ``csrc/xla_ffi_shim.cc``):
* ``vmap_method="expand_dims"``: mapped operands arrive with a leading batch dimension, unmapped ones with a
  leading 1, so ``vmap`` over candidate trees with a shared alignment (``evals/benchmark.py:92-93``) is ONE
  batched launch, not B sequential calls;
* attributes are hashable NumPy scalars (``np.int32`` ...), never jax Arrays;
* no float64 results — the reference runs with x64 disabled (``src/trex/types.py:12``): totals and dC come back
  as float32 (accumulated in f64 inside the library).
UNVERIFIED: never imported or lowered anywhere (no jaxlib on any machine of this project); where jaxlib exists,
``python -c "import trex_b200.jax_ffi as t, jax; print(jax.jit(t.run_sankoff, static_argnames=(...)).lower(...))"``
is the first check to run.
"""

from __future__ import annotations

import ctypes
import functools
import os

import jax
import jax.numpy as jnp
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_FFI_LIB = os.path.join(_HERE, "libtrex_b200_ffi.so")

_lib = ctypes.CDLL(_FFI_LIB)  # raises OSError when the shim was not built: no fallback
for _name in ("trex_run_sankoff_ffi", "trex_soft_fwd_ffi", "trex_soft_bwd_ffi", "trex_tree_cost_ffi",
              "trex_fit_ancestors_ffi"):
  jax.ffi.register_ffi_target(_name, jax.ffi.pycapsule(getattr(_lib, _name)), platform="CUDA")


def _pad16(n):
  return (n + 15) // 16 * 16


def _plan_words(n_all, n_leaves):
  """``trex_sankoff_plan_words`` (host-side query of the library: header + 9 words per internal node + 2 * 128)."""
  _lib.trex_sankoff_plan_words.restype = ctypes.c_size_t
  return int(_lib.trex_sankoff_plan_words(ctypes.c_int(n_all), ctypes.c_int(n_leaves)))


@functools.partial(jax.jit, static_argnames=("return_path", "n_all", "n_states", "n_leaves"))
def run_sankoff(adjacency_matrix, cost_matrix, sequences, n_all, n_states, n_leaves, *, return_path=False):
  """``trex.sankoff.run_sankoff`` — (recon (n_all,L) f32, dp (L,n_all,A) f32, total () f32)."""
  L = sequences.shape[1]
  f32 = jnp.float32
  recon, dp, total = jax.ffi.ffi_call(
    "trex_run_sankoff_ffi",
    (jax.ShapeDtypeStruct((n_all, L), f32), jax.ShapeDtypeStruct((L, n_all, n_states), f32),
     jax.ShapeDtypeStruct((), f32)),
    vmap_method="expand_dims",
  )(adjacency_matrix.astype(f32), cost_matrix.astype(f32), sequences.astype(f32),
    n_leaves=np.int32(n_leaves), return_path=bool(return_path))
  return recon, dp, total


def _tree_cost_call(sequences, adjacency, cost_matrix):
  f32 = jnp.float32
  N, L, Q = sequences.shape
  kind = 0 if cost_matrix is None else (1 if cost_matrix.ndim == 1 else 2)
  c = jnp.zeros((1,), f32) if cost_matrix is None else cost_matrix.astype(f32)
  return jax.ffi.ffi_call(
    "trex_tree_cost_ffi",
    (jax.ShapeDtypeStruct((), f32), jax.ShapeDtypeStruct((N, L, Q), f32), jax.ShapeDtypeStruct((N, N), f32),
     jax.ShapeDtypeStruct(c.shape, f32)),
    vmap_method="expand_dims",
  )(sequences.astype(f32), adjacency.astype(f32), c, cost_kind=np.int32(kind))


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
  return _soft_fwd(adjacency_matrix, cost_matrix, leaf_cost_native, n_leaves, n_sites, temperature)[2]


def _soft_fwd(adj, cost, leaf_cost, n_leaves, n_sites, temperature):
  n_all, A, stride = adj.shape[0], cost.shape[0], leaf_cost.shape[2]
  n_int = n_all - n_leaves
  return jax.ffi.ffi_call(
    "trex_soft_fwd_ffi",
    (jax.ShapeDtypeStruct((_plan_words(n_all, n_leaves),), jnp.int32),
     jax.ShapeDtypeStruct((n_int, A, stride), jnp.float32), jax.ShapeDtypeStruct((), jnp.float32)),
    vmap_method="expand_dims",
  )(adj.astype(jnp.float32), cost.astype(jnp.float32), leaf_cost, n_leaves=np.int32(n_leaves),
    n_sites=np.int64(n_sites), temperature=np.float32(temperature))


def _ssc_fwd(adj, cost, leaf_cost, n_leaves, n_sites, temperature):
  plan, dp, total = _soft_fwd(adj, cost, leaf_cost, n_leaves, n_sites, temperature)
  return total, (plan, dp, cost, leaf_cost, adj.shape[0])


def _ssc_bwd(n_leaves, n_sites, temperature, res, g):
  plan, dp, cost, leaf_cost, n_all = res
  A, stride = cost.shape[0], leaf_cost.shape[2]
  n_int = n_all - n_leaves
  _, dleaf, dcost = jax.ffi.ffi_call(
    "trex_soft_bwd_ffi",
    (jax.ShapeDtypeStruct((n_int, A, stride), jnp.float32), jax.ShapeDtypeStruct(leaf_cost.shape, jnp.float32),
     jax.ShapeDtypeStruct((A, A), jnp.float32)),
    vmap_method="expand_dims",
  )(plan, cost.astype(jnp.float32), leaf_cost, dp, jnp.asarray(g, jnp.float32).reshape(1),
    n_all=np.int32(n_all), n_leaves=np.int32(n_leaves), n_sites=np.int64(n_sites),
    temperature=np.float32(temperature))
  return jnp.zeros((n_all, n_all), jnp.float32), dcost.astype(cost.dtype), dleaf


soft_sankoff_cost.defvjp(_ssc_fwd, _ssc_bwd)


_OPTIMIZERS = {"adam": 0, "sgd": 1, "rmsprop": 2, "adamw": 3}  # TREX_OPT_* (include/trex_b200.h)


@functools.partial(jax.jit, static_argnames=("optimizer", "learning_rate", "clip_norm", "n_iterations"))
def fit_ancestors(adjacency, leaf_seqs, ancestor_logits, optimizer="adam", learning_rate=1e-3, clip_norm=1.0,
                  n_iterations=10_000):
  """The loop of ``run_trex_optimization_configurable`` (``src/trex/evals/benchmark.py:75-200``) as one call:
  returns ``(logits (n_anc,L,Q) f32, losses (n_iterations,) f32, argmax (n_anc,L) i32)``.  Not vmapped."""
  f32 = jnp.float32
  n_anc, L, Q = ancestor_logits.shape
  out, _, _, losses, argmax = jax.ffi.ffi_call(
    "trex_fit_ancestors_ffi",
    (jax.ShapeDtypeStruct((n_anc, L, Q), f32), jax.ShapeDtypeStruct((n_anc, L, Q), f32),
     jax.ShapeDtypeStruct((n_anc, L, Q), f32), jax.ShapeDtypeStruct((n_iterations,), f32),
     jax.ShapeDtypeStruct((n_anc, L), jnp.int32)),
    vmap_method="sequential",
  )(adjacency.astype(f32), leaf_seqs.astype(f32), ancestor_logits.astype(f32),
    optimizer=np.int32(_OPTIMIZERS[optimizer]), learning_rate=np.float32(learning_rate),
    clip_norm=np.float32(clip_norm), n_iterations=np.int32(n_iterations))
  return out, losses, argmax
