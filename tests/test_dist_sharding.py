"""CPU tests (gloo, world_size 2) of the host logic of the two sharded paths of SURVEY 8e rows 2 and 3:
tree-batch sharded scoring (config c4) and the tree cost sharded over the contraction axis.  The CUDA kernels
cannot run here, so the oracle is injected as each rank's local evaluator (test infrastructure: the product
default is the CUDA path, trex_b200/dist.py); what is under test is the partition, the single collective of
each path and the reassembly.  The 2-GPU NCCL twins are in tests/test_gpu_multi.py."""

import os

import numpy as np
import pytest


def _init(rank, world, port):
  import torch.distributed as dist

  os.environ["MASTER_ADDR"] = "127.0.0.1"
  os.environ["MASTER_PORT"] = str(port)
  dist.init_process_group("gloo", rank=rank, world_size=world)
  return dist


def _trees_worker(rank, world, port, q):
  import torch
  from oracle import sankoff_oracle as so
  from trex_b200 import synth
  from trex_b200.dist import score_trees_sharded

  dist = _init(rank, world, port)
  rng = np.random.default_rng(4)  # identical inputs on every rank
  n, A, L, B = 7, 4, 33, 5        # 5 trees over 2 ranks: ragged shards (3 + 2)
  n_all = 2 * n - 1
  adj = np.stack([synth.parents_to_adjacency(synth.random_join_parents(n, rng)) for _ in range(B)])
  states = synth.evolve(synth.balanced_parents(8)[: 2 * 8 - 1], L, A, 0.3, rng)[:n]
  cost = synth.unit_cost(A)
  calls = []

  def local_fn(adj_slice, cost_t, leaf_t, n_states):
    calls.append(adj_slice.shape[0])
    out = [so.per_site_root_min(so.run_sankoff_tables(a.numpy(), cost_t.numpy(), leaf_t.numpy(), n_all, n_states, n)["dp"])
           .astype(np.float64).sum() for a in adj_slice]
    return torch.tensor(out, dtype=torch.float64)

  scores = score_trees_sharded(torch.from_numpy(adj), torch.from_numpy(cost), torch.from_numpy(states), A,
                               local_fn=local_fn)
  want = [float(so.per_site_root_min(so.run_sankoff_tables(a, cost, states, n_all, A, n)["dp"]).astype(np.float64).sum())
          for a in adj]
  q.put((rank, scores.tolist(), want, calls))
  dist.destroy_process_group()


def _tree_cost_worker(rank, world, port, q):
  import torch
  from oracle import tree_oracle as to
  from trex_b200.dist import shard_range, tree_cost_value_and_grad_sharded

  dist = _init(rank, world, port)
  rng = np.random.default_rng(2)
  N, L, Q = 9, 21, 4
  S = to.softmax(rng.normal(size=(N, L, Q)))
  adj = to.softmax(rng.normal(size=(N, N)))
  C = rng.random((Q, Q))

  def local_fn(S_loc, adj_t, cost_t):
    c, dS, dA, dC = to.soft_cost_value_and_grad(S_loc.numpy(), adj_t.numpy(), None if cost_t is None else cost_t.numpy())
    return (torch.tensor(c, dtype=torch.float64), torch.from_numpy(dS), torch.from_numpy(np.asarray(dA)),
            None if dC is None else torch.from_numpy(np.asarray(dC)))

  out = {}
  for name, cm in (("identity", None), ("matrix", C)):
    lo, hi = shard_range(L, world, rank)
    c, dS, dA, dC = tree_cost_value_and_grad_sharded(torch.from_numpy(S[:, lo:hi].copy()), torch.from_numpy(adj),
                                                     None if cm is None else torch.from_numpy(cm), local_fn=local_fn)
    wc, wdS, wdA, wdC = to.soft_cost_value_and_grad(S, adj, cm)
    out[name] = (float(c), float(wc), float(np.abs(dS.numpy() - wdS[:, lo:hi]).max()),
                 float(np.abs(dA.numpy() - wdA).max() / np.abs(wdA).max()),
                 0.0 if cm is None else float(np.abs(dC.numpy() - wdC).max() / np.abs(wdC).max()))
  q.put((rank, out))
  dist.destroy_process_group()


def _spawn(worker, base_port):
  import torch.multiprocessing as mp

  ctx = mp.get_context("spawn")
  q = ctx.Queue()
  port = base_port + os.getpid() % 2000
  procs = [ctx.Process(target=worker, args=(r, 2, port, q)) for r in range(2)]
  for p in procs:
    p.start()
  res = [q.get(timeout=180) for _ in procs]
  for p in procs:
    p.join(timeout=60)
    assert p.exitcode == 0
  return res


def test_tree_batch_sharded_scores_world_size_2_gloo():
  res = _spawn(_trees_worker, 31500)
  seen = {}
  for rank, scores, want, calls in res:
    assert scores == want            # every rank holds all B exact scores, in tree order
    seen[rank] = calls
  assert seen[0] == [3] and seen[1] == [2]  # each rank evaluated only its own shard, once


def test_tree_cost_sharded_over_sites_world_size_2_gloo():
  res = _spawn(_tree_cost_worker, 33500)
  for _, out in res:
    for name, (c, wc, ds_err, da_rel, dc_rel) in out.items():
      assert abs(c - wc) <= 1e-6 * abs(wc), name   # the f32 cast of the f64-reduced scalar
      assert ds_err <= 1e-12                       # dS stays local to its sites
      assert da_rel <= 1e-6 and dc_rel <= 1e-6     # f32 casts of the f64-reduced blocks
