"""NumPy stand-in for the handful of ``jax`` primitives the reference's hot path uses.

JAX cannot be installed in this container (no network), so the golden vectors under
``tests/golden/`` are produced by importing the UNMODIFIED reference modules
(``/root/reference/src/trex/{sankoff,tree}.py``, ``/root/reference/benchmark_tree_cost.py``)
with this module registered as ``jax``.  Only semantics documented by JAX are restated:

* ``x.at[idx].set(v)``     — functional update; negative indices wrap, out-of-bounds dropped
* ``jnp.where(c, size=k, fill_value=f)`` — first k true indices ascending, padded with f
* ``jnp.argmin/argmax``    — first extremum (same as NumPy)
* ``lax.fori_loop/scan/while_loop/cond``, ``vmap`` (Python loops), ``jit`` (identity)
* ``jax.nn.softmax`` (max-subtracted), ``jax.nn.one_hot``
* ``jax.random.*``         — NOT reproducible without JAX: ``gumbel`` draws from NumPy and the
  draw is recorded in ``LAST_GUMBEL`` so the fixture stores the noise it was produced with.

This is test infrastructure; nothing in ``trex_b200`` imports it.
"""

from __future__ import annotations

import sys
import types

import numpy as np


class Arr(np.ndarray):
  """ndarray with jax's ``.at[...]`` functional-update property."""

  @property
  def at(self):
    return _At(self)

  def block_until_ready(self):
    return self


def _wrap(x):
  return np.asarray(x).view(Arr)


class _AtIdx:

  def __init__(self, arr, idx):
    self.arr, self.idx = arr, idx

  def set(self, value):
    out = np.array(self.arr, copy=True).view(Arr)
    idx = self.idx if isinstance(self.idx, tuple) else (self.idx,)
    # jax scatter: integer indices wrap when negative and are DROPPED when out of bounds
    fixed = []
    for ax, ix in enumerate(idx):
      if isinstance(ix, (int, np.integer)) or (isinstance(ix, np.ndarray) and ix.ndim == 0 and ix.dtype.kind in "iu"):
        i = int(ix)
        n = out.shape[ax]
        if i < 0:
          i += n
        if i < 0 or i >= n:
          return out
        fixed.append(i)
      else:
        fixed.append(ix)
    out[tuple(fixed)] = value
    return out


class _At:

  def __init__(self, arr):
    self.arr = arr

  def __getitem__(self, idx):
    return _AtIdx(self.arr, idx)


def _where(cond, x=None, y=None, *, size=None, fill_value=None):
  if x is None and y is None:
    idx = np.nonzero(np.asarray(cond))
    if size is not None:
      out = []
      for a in idx:
        b = np.full(size, 0 if fill_value is None else fill_value, dtype=np.int64)
        b[: min(size, a.size)] = a[:size]
        out.append(_wrap(b))
      return tuple(out)
    return tuple(_wrap(a) for a in idx)
  return _wrap(np.where(cond, x, y))


def _tree_map(f, *trees):
  t0 = trees[0]
  if isinstance(t0, (tuple, list)):
    return type(t0)(_tree_map(f, *[t[i] for t in trees]) for i in range(len(t0)))
  if isinstance(t0, dict):
    return {k: _tree_map(f, *[t[k] for t in trees]) for k in t0}
  return f(*trees)


def _fori_loop(lo, hi, body, init):
  val = init
  for i in range(int(lo), int(hi)):
    val = body(i, val)
  return val


def _scan(f, init, xs, length=None):
  carry, ys = init, []
  n = length if xs is None else len(xs)
  for i in range(n):
    carry, y = f(carry, None if xs is None else _tree_map(lambda a: a[i], xs))
    ys.append(y)
  stacked = _tree_map(lambda *a: _wrap(np.stack([np.asarray(v) for v in a])), *ys) if ys and ys[0] is not None else None
  return carry, stacked


def _while_loop(cond, body, val):
  while bool(cond(val)):
    val = body(val)
  return val


def _cond(pred, t, f, *operands):
  return t(*operands) if bool(pred) else f(*operands)


def _vmap(fn, in_axes=0, out_axes=0):
  def wrapped(*args):
    axes = in_axes if isinstance(in_axes, (tuple, list)) else (in_axes,) * len(args)
    n = next(np.asarray(a).shape[ax] for a, ax in zip(args, axes) if ax is not None)
    outs = []
    for b in range(n):
      sl = [a if ax is None else _wrap(np.take(np.asarray(a), b, axis=ax)) for a, ax in zip(args, axes)]
      outs.append(fn(*sl))
    return _tree_map(lambda *a: _wrap(np.stack([np.asarray(v) for v in a], axis=out_axes)), *outs)
  return wrapped


def _jit(fn=None, **_kw):
  if fn is None:
    return lambda f: f
  return fn


def _softmax(x, axis=-1):
  x = np.asarray(x)
  m = np.max(x, axis=axis, keepdims=True)
  m = np.where(np.isfinite(m), m, 0)
  e = np.exp(x - m)
  return _wrap(e / e.sum(axis=axis, keepdims=True))


def _one_hot(x, num_classes, dtype=np.float32):
  x = np.asarray(x).astype(np.int64)
  return _wrap((x[..., None] == np.arange(num_classes)).astype(dtype))


LAST_GUMBEL = {}
_RNG = np.random.default_rng(20261104)


def _gumbel(key, shape=(), dtype=np.float32):
  g = _RNG.gumbel(size=shape).astype(dtype)
  LAST_GUMBEL["noise"] = g
  return _wrap(g)


# Recording random functions (make_golden_generators.py): jax.random's threefry streams are not reproducible without
# JAX, so each call draws from NumPy and appends (function name, array) to DRAWS in CALL ORDER; the golden script
# maps that log onto the explicit draw arrays the oracle and the CUDA generators take.  bernoulli is `uniform < p`
# (as in jax) and logs the uniform.
DRAWS = []


def _rec(name, arr):
  DRAWS.append((name, np.array(arr)))
  return arr


def _r_choice(key, a, shape=(), replace=True, p=None):
  n = int(np.prod(shape)) if shape != () else 1
  out = _RNG.choice(int(a) if np.ndim(a) == 0 else np.asarray(a), size=n, replace=replace, p=p).reshape(shape)
  return _wrap(_rec("choice", out))


def _r_randint(key, shape, minval, maxval, dtype=np.int32):
  return _wrap(_rec("randint", _RNG.integers(minval, maxval, size=shape).astype(dtype)))


def _r_uniform(key, shape=(), dtype=np.float32, minval=0.0, maxval=1.0):
  return _wrap(_rec("uniform", (_RNG.random(size=shape) * (maxval - minval) + minval).astype(dtype)))


def _r_bernoulli(key, p=0.5, shape=None):
  shape = np.shape(p) if shape is None else shape
  u = _rec("bernoulli_uniform", _RNG.random(size=shape).astype(np.float32))
  r = u < np.asarray(p)
  return _wrap(np.asarray(r)) if np.ndim(r) else bool(r)


def _r_normal(key, shape=(), dtype=np.float32):
  r = _rec("normal", np.asarray(_RNG.normal(size=shape)).astype(dtype))
  return _wrap(r) if np.ndim(r) else np.float32(r)


def install():
  """Register the stand-in as ``jax`` / ``jax.numpy`` / ``jax.lax`` / ``jax.nn`` / ``jax.random``."""
  if "jax" in sys.modules and not getattr(sys.modules["jax"], "_IS_NUMPY_SHIM", False):
    raise RuntimeError("a real jax is importable: generate the fixtures with it instead")
  jax = types.ModuleType("jax")
  jax._IS_NUMPY_SHIM = True
  jnp = types.ModuleType("jax.numpy")
  for name in ["zeros", "ones", "full", "arange", "eye", "stack", "concatenate", "zeros_like", "ones_like",
               "array", "asarray", "isnan", "isclose", "allclose", "log", "exp", "einsum", "broadcast_to",
               "transpose", "sqrt", "abs", "mean", "roll", "append", "dot", "squeeze", "minimum", "maximum"]:
    def mk(fn):
      return lambda *a, **k: _wrap(fn(*a, **k))
    setattr(jnp, name, mk(getattr(np, name)))
  for name in ["min", "max", "argmin", "argmax", "sum", "any", "all"]:
    def mk(fn):
      def g(*a, **k):
        r = fn(*a, **k)
        return _wrap(r) if isinstance(r, np.ndarray) else r
      return g
    setattr(jnp, name, mk(getattr(np, name)))
  jnp.where = _where
  jnp.inf, jnp.newaxis, jnp.ndarray = np.inf, np.newaxis, np.ndarray
  jnp.int32, jnp.int8, jnp.float32, jnp.float64, jnp.bfloat16 = np.int32, np.int8, np.float32, np.float64, np.float32
  jnp.dtype = np.dtype
  jnp.bool_ = np.bool_
  lax = types.ModuleType("jax.lax")
  lax.fori_loop, lax.scan, lax.while_loop, lax.cond = _fori_loop, _scan, _while_loop, _cond
  nn = types.ModuleType("jax.nn")
  nn.softmax, nn.one_hot = _softmax, _one_hot
  nn.log_softmax = lambda x, axis=-1: _wrap(np.log(np.asarray(_softmax(x, axis=axis))))
  # lax.map(f, xs, batch_size=...) == stack of f(x) over the leading axis (utils/memory.py:28)
  lax.map = lambda f, xs, batch_size=None: _wrap(np.stack([np.asarray(f(x)) for x in np.asarray(xs)]))
  random = types.ModuleType("jax.random")
  random.PRNGKey = random.key = lambda seed: _wrap(np.array([0, seed], dtype=np.uint32))
  random.split = lambda key, num=2: _wrap(np.stack([np.asarray(key)] * num))
  random.gumbel = _gumbel
  random.choice, random.randint, random.uniform = _r_choice, _r_randint, _r_uniform
  random.bernoulli, random.normal = _r_bernoulli, _r_normal
  debug = types.ModuleType("jax.debug")
  debug.print = lambda *a, **k: None
  tree_util = types.ModuleType("jax.tree_util")
  tree_util.tree_map = _tree_map
  jax.numpy, jax.lax, jax.nn, jax.random, jax.debug, jax.tree_util = jnp, lax, nn, random, debug, tree_util
  jax.jit, jax.vmap, jax.Array = _jit, _vmap, np.ndarray
  jax.block_until_ready = lambda x: x

  class _CustomVjp:  # jax.custom_vjp: the primal function, defvjp ignored (no autodiff in the stand-in)
    def __init__(self, fn, **_):
      self.fn = fn

    def __call__(self, *a, **k):
      return self.fn(*a, **k)

    def defvjp(self, *a, **k):
      return None

  jax.custom_vjp = _CustomVjp

  def _no_autodiff(fn, *a, **k):  # may be built at import time (evals/benchmark.py:807), must not be called
    def call(*aa, **kk):
      raise NotImplementedError("no autodiff in the NumPy stand-in")
    return call

  jax.value_and_grad = jax.grad = _no_autodiff
  for name, mod in [("jax", jax), ("jax.numpy", jnp), ("jax.lax", lax), ("jax.nn", nn), ("jax.random", random),
                    ("jax.debug", debug), ("jax.tree_util", tree_util)]:
    sys.modules[name] = mod
  return jax
