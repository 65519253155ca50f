"""Host-side mirror of the reference's input generators on top of the device generators of ``libtrex_b200.so``
(SURVEY 8(f) rank 3): ``trex.ground_truth`` (``/root/reference/src/trex/ground_truth.py``), ``trex.nk_model``
(``/root/reference/src/trex/nk_model.py``) and ``create_balanced_binary_tree``
(``/root/reference/src/trex/evals/benchmark.py:781-791``).  Same names, argument order and return shapes / dtypes;
a JAX PRNG key becomes an integer ``seed`` (the device draws from a counter-based Philox stream — jax.random's
threefry streams cannot be reproduced outside JAX), and every function accepts the draws as explicit arrays
(``draws=`` / ``mut_sites=`` / ``offsets=``), which is how the parity tests compare it with the reference.

Arrays are torch CUDA tensors; there is no CPU fallback."""

from __future__ import annotations

from collections import namedtuple

import torch

from . import _lib
from ._lib import call, ptr, stream_ptr
from .sankoff import pad_sites

# trex.utils.types.PhylogeneticTree (masked_sequences, all_sequences, adjacency) + the native u8 state table
PhylogeneticTree = namedtuple("PhylogeneticTree", ["masked_sequences", "all_sequences", "adjacency", "states_u8"])


def _dev(device=None):
  if not torch.cuda.is_available():
    raise RuntimeError("trex_b200 needs a CUDA device; there is no CPU fallback")
  return torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)


def create_balanced_binary_tree(n_leaves: int, device=None):
  """``create_balanced_binary_tree`` (evals/benchmark.py:781-791): (2n-1, 2n-1) f32, no root self-loop."""
  dev = _dev(device)
  n_all = 2 * n_leaves - 1
  adj = torch.empty((n_all, n_all), dtype=torch.float32, device=dev)
  call("trex_tree_adjacency", int(n_leaves), 1, ptr(adj), stream_ptr())
  return adj


def generate_groundtruth(n_leaves, n_states, n_mutations, seq_length, seed=42, *, mut_sites=None, offsets=None,
                         device=None):
  """``generate_groundtruth`` (ground_truth.py:113-197).  Returns ``PhylogeneticTree`` with ``masked_sequences``,
  ``all_sequences`` (n_all, L) f32 and ``adjacency`` (n_all, n_all) f32 as the reference, plus ``states_u8``
  [n_all][pad_sites(L)] — rows ``:n_leaves`` of it are what ``SankoffEngine.set_leaf_states_u8`` takes, so a
  synthetic alignment never has to leave the device.  ``mut_sites`` (n_all-1, n_mutations) i32 / ``offsets``
  (n_all-1, L) u8: explicit draws, indexed by the child node each ``mutate`` call produces."""
  lib = _lib.load()
  dev = _dev(device)
  if not (n_leaves > 0 and (n_leaves & (n_leaves - 1) == 0)):
    raise ValueError("n_leaves must be a power of 2.")  # ground_truth.py:142-144
  n_all = 2 * n_leaves - 1
  L, stride = int(seq_length), pad_sites(seq_length)
  states = torch.zeros((n_all, stride), dtype=torch.uint8, device=dev)
  adj = torch.empty((n_all, n_all), dtype=torch.float32, device=dev)
  ws = torch.empty(lib.trex_generate_groundtruth_workspace_bytes(n_leaves, n_mutations, L), dtype=torch.uint8, device=dev)
  ms = None if mut_sites is None else torch.as_tensor(mut_sites).to(dev, torch.int32).contiguous()
  of = None if offsets is None else torch.as_tensor(offsets).to(dev, torch.uint8).contiguous()
  call("trex_generate_groundtruth", int(n_leaves), int(n_states), int(n_mutations), L, stride, int(seed) & (2**64 - 1),
       ptr(ms), ptr(of), ptr(states), ptr(adj), ptr(ws), stream_ptr())
  all_seqs = states[:, :L].to(torch.float32)
  masked = all_seqs.clone()
  masked[n_leaves:] = 0  # ground_truth.py:175-177
  return PhylogeneticTree(masked, all_seqs, adj, states)


def create_nk_model_landscape(n, k, seed, n_states=2, device=None):
  """``create_nk_model_landscape`` (nk_model.py:17-43): dict with 'interactions' (n, k) i32, 'fitness_tables'
  (n, n_states^(k+1)) f32, 'n_states', 'k'."""
  dev = _dev(device)
  inter = torch.empty((n, max(k, 0)), dtype=torch.int32, device=dev)
  tables = torch.empty((n, n_states ** (k + 1)), dtype=torch.float32, device=dev)
  call("trex_nk_landscape", int(n), int(k), int(n_states), int(seed) & (2**64 - 1), ptr(inter), ptr(tables), stream_ptr())
  return {"interactions": inter, "fitness_tables": tables, "n_states": int(n_states), "k": int(k)}


def batched_get_fitness(sequences, landscape, seq_mask=None):
  """``batched_get_fitness`` (nk_model.py:113): sequences (B, n) integer states -> (B,) f32."""
  seqs = torch.as_tensor(sequences)
  inter, tables = landscape["interactions"], landscape["fitness_tables"]
  seqs = seqs.to(inter.device, torch.uint8).contiguous()
  B, n = seqs.shape
  out = torch.empty(B, dtype=torch.float32, device=inter.device)
  m = None if seq_mask is None else torch.as_tensor(seq_mask).to(inter.device, torch.float32).contiguous()
  call("trex_nk_fitness", ptr(seqs), n, B, n, int(inter.shape[1]), int(landscape["n_states"]), ptr(inter), ptr(tables),
       ptr(m), ptr(out), stream_ptr())
  return out


def get_fitness(sequence, landscape, seq_mask=None):
  """``get_fitness`` (nk_model.py:46-110): scalar f32 mean fitness over the (masked) sites."""
  return batched_get_fitness(torch.as_tensor(sequence).reshape(1, -1), landscape, seq_mask)[0]


def generate_tree_data(landscape, adjacency, root_sequence, mutation_rate, seed, coupled_mutation_prob=0.5,
                       n_states=20, mutation_rate_noise_std=0.0, branch_length=1, *, draws=None):
  """``generate_tree_data`` (nk_model.py:116-278).  Returns ``PhylogeneticTree`` (``masked_sequences`` all zeros as in
  the reference, :271).  ``draws``: optional dict of explicit draw arrays (noise [N]; u_coupled, site, u_accept
  [N][B]; u_mask f32, new_vals u8 [N][B][L]).  Root rule: see include/trex_b200.h (the reference's degenerate
  fallback for adjacencies without a root self-loop is not reproduced)."""
  lib = _lib.load()
  inter, tables = landscape["interactions"], landscape["fitness_tables"]
  dev = inter.device
  if "n_states" in landscape:
    n_states = int(landscape["n_states"])  # nk_model.py:143-145
  adj = torch.as_tensor(adjacency).to(dev, torch.float32).contiguous()
  N = adj.shape[0]
  root = torch.as_tensor(root_sequence).reshape(-1).to(dev, torch.uint8).contiguous()
  L, stride = root.shape[0], pad_sites(root.shape[0])
  states = torch.zeros((N, stride), dtype=torch.uint8, device=dev)
  ws = torch.empty(lib.trex_nk_generate_tree_workspace_bytes(N, stride), dtype=torch.uint8, device=dev)
  d = draws or {}
  f32, i32, u8 = torch.float32, torch.int32, torch.uint8
  get = lambda k, dt: None if d.get(k) is None else torch.as_tensor(d[k]).to(dev, dt).contiguous()
  held = [get("noise", f32), get("u_coupled", f32), get("site", i32), get("u_mask", f32), get("new_vals", u8),
          get("u_accept", f32)]
  call("trex_nk_generate_tree", ptr(adj), N, ptr(root), L, stride, int(inter.shape[1]), n_states, ptr(inter), ptr(tables),
       float(mutation_rate), float(coupled_mutation_prob), float(mutation_rate_noise_std), int(branch_length),
       int(seed) & (2**64 - 1), *[ptr(x) for x in held], ptr(states), ptr(ws), stream_ptr())
  all_seqs = states[:, :L].to(torch.float32)
  return PhylogeneticTree(torch.zeros_like(all_seqs), all_seqs, adj, states)
