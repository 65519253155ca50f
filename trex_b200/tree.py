"""Host-side mirror of the reference's ``trex.tree`` (``/root/reference/src/trex/tree.py``) for
the soft-adjacency tree cost, on top of ``trex_tree_cost_fwd_bwd`` (include/trex_b200.h).

``compute_surrogate_cost`` / ``compute_soft_cost`` keep the reference signatures and are
differentiable (torch autograd plays the role of ``jax.custom_vjp``): the single library call
returns the value and every gradient, which the backward pass merely scales by the upstream
cotangent.  Arrays are torch CUDA tensors; there is no CPU fallback.
"""

from __future__ import annotations

import torch

from . import _lib
from ._lib import call, ptr, stream_ptr


def tree_cost_value_and_grad(sequences, adjacency, cost_matrix=None, *, want_dseq=True,
                             want_dadj=True, want_dcost=True, workspace=None, out=None):
  """One pass of value + gradients of ``compute_soft_cost`` (tree.py:212-266).

  Returns ``(cost () f32, d_sequences (N,L,Q)|None, d_adjacency (N,N)|None, d_cost|None)`` —
  what ``jax.value_and_grad`` yields in ``src/trex/evals/benchmark.py:167-173``.
  """
  lib = _lib.load()
  if not sequences.is_cuda:
    raise RuntimeError("trex_b200 needs CUDA tensors; there is no CPU fallback")
  S = sequences.to(torch.float32).contiguous()
  adj = adjacency.to(device=S.device, dtype=torch.float32).contiguous()
  N, L, Q = S.shape
  kind = 0
  C = None
  if cost_matrix is not None:
    C = cost_matrix.to(device=S.device, dtype=torch.float32).contiguous()
    kind = 1 if C.dim() == 1 else 2
  if workspace is None:
    nbytes = lib.trex_tree_cost_workspace_bytes(N, L * Q, Q)
    workspace = torch.empty(nbytes, dtype=torch.uint8, device=S.device)
  if out is None:
    out = {}
  cost = out.get("cost")
  if cost is None:
    cost = torch.empty((), dtype=torch.float32, device=S.device)
  dS = (out.get("dS") if out.get("dS") is not None else torch.empty_like(S)) if want_dseq else None
  dA = (out.get("dA") if out.get("dA") is not None else torch.empty_like(adj)) if want_dadj else None
  dC = None
  if want_dcost and C is not None:
    dC = out.get("dC") if out.get("dC") is not None else torch.empty_like(C)
  call("trex_tree_cost_fwd_bwd", ptr(S), ptr(adj), ptr(C), kind, N, L, Q, ptr(cost), ptr(dS),
       ptr(dA), ptr(dC), ptr(workspace), stream_ptr())
  return cost, dS, dA, dC


class _TreeCost(torch.autograd.Function):

  @staticmethod
  def forward(ctx, sequences, adjacency, cost_matrix):
    need = ctx.needs_input_grad
    cost, dS, dA, dC = tree_cost_value_and_grad(
      sequences, adjacency, cost_matrix, want_dseq=need[0], want_dadj=need[1],
      want_dcost=cost_matrix is not None and need[2])
    ctx.grads = (dS, dA, dC)
    ctx.in_dtypes = (sequences.dtype, adjacency.dtype, None if cost_matrix is None else cost_matrix.dtype)
    return cost

  @staticmethod
  def backward(ctx, g):
    dS, dA, dC = ctx.grads
    dt = ctx.in_dtypes
    return (None if dS is None else (dS * g).to(dt[0]), None if dA is None else (dA * g).to(dt[1]),
            None if dC is None else (dC * g).to(dt[2]))


def compute_surrogate_cost(sequences, adjacency):
  """``trex.tree.compute_surrogate_cost`` (tree.py:163-209)."""
  return _TreeCost.apply(sequences, adjacency, None)


def compute_soft_cost(sequences, adjacency, cost_matrix=None):
  """``trex.tree.compute_soft_cost`` (tree.py:212-266); ``cost_matrix`` (Q,Q), (Q,) or None."""
  return _TreeCost.apply(sequences, adjacency, cost_matrix)


# ---------------------------------------------------------------------------------------------
# T3-T8: update_seq, update_tree, enforce_graph_constraints, compute_cost,
# discretize_tree_topology, compute_loss (reference signatures; autograd where the reference
# is differentiated in tests/test_convergence.py and evals/benchmark.py)
# ---------------------------------------------------------------------------------------------


def _stack_ancestors(params):
  a = params["ancestors"] if isinstance(params, dict) else params
  if isinstance(a, (list, tuple)):
    a = torch.stack(list(a))  # jnp.stack(params["ancestors"]), tree.py:129
  return a


class _UpdateSeq(torch.autograd.Function):

  @staticmethod
  def forward(ctx, logits, sequences, temperature):
    _lib.load()
    lg = logits.to(torch.float32).contiguous()
    n_anc, L, Q = lg.shape
    out = sequences.to(device=lg.device, dtype=torch.float32).clone().contiguous()
    n_leaf = out.shape[0] - n_anc
    call("trex_update_seq_fwd", ptr(lg), n_anc, L, Q, float(temperature), ptr(out[n_leaf:]), stream_ptr())
    ctx.save_for_backward(out)
    ctx.meta = (n_anc, L, Q, n_leaf, float(temperature), logits.dtype)
    return out

  @staticmethod
  def backward(ctx, g):
    (out,) = ctx.saved_tensors
    n_anc, L, Q, n_leaf, temp, dt = ctx.meta
    g = g.to(torch.float32).contiguous()
    dlog = torch.empty((n_anc, L, Q), dtype=torch.float32, device=out.device)
    call("trex_update_seq_bwd", ptr(out[n_leaf:]), ptr(g[n_leaf:]), n_anc, L, Q, temp, ptr(dlog), stream_ptr())
    gseq = g.clone()
    gseq[n_leaf:] = 0  # `.at[n_leaves:].set(...)` overwrites those rows of `sequences`
    return dlog.to(dt), gseq, None


def update_seq(params, sequences, temperature=1.0):
  """``trex.tree.update_seq`` (tree.py:110-130): rows ``(n+1)//2:`` become
  ``softmax(stack(params["ancestors"]) * temperature)``.  ``params`` may also be the stacked
  (n_anc, L, Q) array itself (``_update_seq_stacked``, evals/benchmark.py:210-232)."""
  return _UpdateSeq.apply(_stack_ancestors(params), sequences, temperature)


class _UpdateTree(torch.autograd.Function):

  @staticmethod
  def forward(ctx, tree_params, noise, gates, temperature):
    _lib.load()
    tp = tree_params.to(torch.float32).contiguous()
    n_m1, n_anc = tp.shape
    N = n_m1 + 1
    n_leaves = N - n_anc
    nz = None if noise is None else noise.to(device=tp.device, dtype=torch.float32).contiguous()
    gt = None if gates is None else gates.to(device=tp.device, dtype=torch.float32).expand(n_m1, n_anc).contiguous()
    adj = torch.empty((N, N), dtype=torch.float32, device=tp.device)
    call("trex_update_tree_fwd", ptr(tp), ptr(nz), ptr(gt), float(temperature), N, n_leaves, ptr(adj), stream_ptr())
    ctx.save_for_backward(adj, gt if gt is not None else torch.empty(0, device=tp.device))
    ctx.meta = (N, n_leaves, float(temperature), gt is not None, tree_params.dtype)
    return adj

  @staticmethod
  def backward(ctx, g):
    adj, gt = ctx.saved_tensors
    N, n_leaves, temp, has_gates, dt = ctx.meta
    out = torch.empty((N - 1, N - n_leaves), dtype=torch.float32, device=adj.device)
    call("trex_update_tree_bwd", ptr(adj), ptr(g.to(torch.float32).contiguous()), ptr(gt) if has_gates else None,
         temp, N, n_leaves, ptr(out), stream_ptr())
    return out.to(dt), None, None, None


def gumbel_noise(key, shape, device="cuda"):
  """Stand-in for ``jax.random.gumbel(key, shape)`` (tree.py:71).  ``key`` is a
  ``torch.Generator``, an int seed or ``None``; jax's threefry stream itself is not
  reproducible outside JAX, so parity tests pass explicit noise instead."""
  gen = key if isinstance(key, torch.Generator) else None
  if gen is None and key is not None and not torch.is_tensor(key):
    gen = torch.Generator(device=device)
    gen.manual_seed(int(key))
  u = torch.rand(shape, device=device, generator=gen).clamp_(1e-20, 1.0)
  return -torch.log(-torch.log(u))


def update_tree(key, params, temperature=1.0, gates=None, *, noise=None):
  """``trex.tree.update_tree`` (tree.py:50-107).  ``noise`` (keyword, optional) supplies the
  Gumbel draw explicitly; otherwise it is drawn from ``key``."""
  tp = params["tree_params"]
  n_m1, n_anc = tp.shape
  if n_anc == 0:
    return torch.eye(n_m1 + 1, dtype=tp.dtype, device=tp.device)  # tree.py:68-69
  if noise is None:
    noise = gumbel_noise(key, tuple(tp.shape), device=tp.device)
  return _UpdateTree.apply(tp, noise, gates, temperature)


class _GraphConstraint(torch.autograd.Function):

  @staticmethod
  def forward(ctx, adjacency, scaling_factor):
    _lib.load()
    adj = adjacency.to(torch.float32).contiguous()
    N = adj.shape[0]
    out = torch.empty((), dtype=torch.float32, device=adj.device)
    dadj = torch.empty_like(adj) if ctx.needs_input_grad[0] else None
    scratch = torch.empty(1, dtype=torch.float64, device=adj.device)
    call("trex_graph_constraint", ptr(adj), N, float(scaling_factor), ptr(out), ptr(dadj), ptr(scratch), stream_ptr())
    ctx.dadj, ctx.dt = dadj, adjacency.dtype
    return out

  @staticmethod
  def backward(ctx, g):
    return (ctx.dadj * g).to(ctx.dt), None


def enforce_graph_constraints(adjacency, scaling_factor):
  """``trex.tree.enforce_graph_constraints`` (tree.py:133-160)."""
  return _GraphConstraint.apply(adjacency, scaling_factor)


def discretize_tree_topology(adjacency, n_nodes):
  """``trex.tree.discretize_tree_topology`` (tree.py:31-47)."""
  _lib.load()
  adj = adjacency.to(torch.float32).contiguous()
  assert adj.shape[0] == n_nodes and adj.shape[1] == n_nodes
  out = torch.empty_like(adj)
  scratch = torch.empty(n_nodes, dtype=torch.int32, device=adj.device)
  call("trex_discretize_topology", ptr(adj), n_nodes, ptr(scratch), ptr(out), stream_ptr())
  return out


def compute_cost(sequences, adjacency, substitution_matrix):
  """``trex.tree.compute_cost`` (tree.py:269-296): exact discrete traversal cost."""
  _lib.load()
  S = sequences.to(torch.float32).contiguous()
  N, L, Q = S.shape
  adj = adjacency.to(device=S.device, dtype=torch.float32).contiguous()
  sub = substitution_matrix.to(device=S.device, dtype=torch.float32).contiguous()
  st = torch.empty((N, L), dtype=torch.int32, device=S.device)
  par = torch.empty(N, dtype=torch.int32, device=S.device)
  acc = torch.empty(1, dtype=torch.float64, device=S.device)
  out = torch.empty((), dtype=torch.float32, device=S.device)
  call("trex_compute_cost", ptr(S), ptr(adj), ptr(sub), N, L, Q, ptr(st), ptr(par), ptr(acc), ptr(out), stream_ptr())
  return out


def compute_loss(key, params, sequences, _metadata, temperature, adjacency, *, graph_constraint_scale=10.0,
                 verbose=False, fix_seqs=False, fix_tree=False, noise=None):
  """``trex.tree.compute_loss`` (tree.py:299-361).  ``update_tree`` runs at its default
  temperature 1.0 (:338) while ``update_seq`` gets ``temperature`` (:336)."""
  updated_sequences = sequences if fix_seqs else update_seq(params, sequences, temperature)
  updated_tree = adjacency if fix_tree else update_tree(key, params, noise=noise)
  surrogate = compute_surrogate_cost(updated_sequences, updated_tree)
  constraint = enforce_graph_constraints(updated_tree, graph_constraint_scale)
  total = surrogate + temperature * constraint
  if verbose:
    print("Surrogate cost:", float(surrogate), "Tree constraint loss:", float(constraint), "Total loss:", float(total))
  return total
