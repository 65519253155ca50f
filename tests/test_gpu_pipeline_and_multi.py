"""GPU tests of the host-buffer pipeline (bench.py's e2e leg) and, on boxes with >= 2 GPUs, the NCCL twins of
tests/test_dist_sharding.py: tree-batch sharded scoring (c4) and the site-sharded tree cost, each compared with
the single-GPU / oracle result."""

import os

import numpy as np
import pytest

from oracle import c_oracle
from oracle import sankoff_oracle as so
from trex_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("use_graph", [False, True])
@pytest.mark.parametrize("n,A,L", [(64, 20, 1000), (16, 4, 100)])
def test_host_pipeline_returns_the_reference_results_for_every_job(n, A, L, use_graph):
  """Three jobs with DIFFERENT trees / alignments through the double-buffered pipeline: the host results of
  each job (ancestral states, totals, dC, edge histogram) equal the oracle's for that job's inputs."""
  import torch
  from trex_b200.pipeline import SankoffHostPipeline

  n_all = 2 * n - 1
  rng = np.random.default_rng(n + L)
  T = 0.7
  jobs, want = [], []
  for j in range(3):
    par = synth.random_join_parents(n, rng)
    adj = synth.parents_to_adjacency(par)
    states = np.ascontiguousarray(synth.evolve(par, L, A, 0.15, rng)[:n])
    cost = synth.unit_cost(A) if j != 1 else synth.random_int_cost(A, 4, rng)
    jobs.append((torch.from_numpy(adj).pin_memory(), torch.from_numpy(cost).pin_memory(),
                 torch.from_numpy(states).pin_memory()))
    recon, _, _, root_min = c_oracle.run_sankoff(adj, cost, states.astype(np.float32), n_all, A, n,
                                                 return_path=True, want_dp=False)
    tot64, dC64, _, _ = c_oracle.soft_sankoff_f64(adj, cost, n_all, A, n, T, seqs=states.astype(np.float32),
                                                  want_dp=False, want_gbar=False)
    want.append((recon[n:].astype(np.uint8), float(root_min.astype(np.float64).sum()), tot64, dC64))
  pipe = SankoffHostPipeline(n_all, n, A, L, temperature=T, use_graph=use_graph)
  got = []
  # results live in two alternating host slots: consume job k before job k+2 is submitted
  for k in range(0, 3, 2):
    outs = pipe.run(jobs[k:k + 2])
    torch.cuda.synchronize()
    chunk = jobs[k:k + 2]
    for i in range(len(chunk)):
      anc_h, packed_h = outs[len(chunk) - 1 - i]
      got.append((anc_h.numpy().copy(), packed_h.numpy().copy()))
  assert pipe.d2h_bytes_per_job == (n - 1) * L + 8 * (2 + 2 * A * A)
  for (anc, packed), (w_anc, w_tot, w_soft, w_dC) in zip(got, want):
    assert np.array_equal(anc, w_anc)
    assert packed[0] == w_tot
    assert abs(packed[1] - w_soft) <= 1e-5 * abs(w_soft)
    dC = packed[2:2 + A * A].reshape(A, A)
    assert np.abs(dC - w_dC).max() <= 1e-5 * np.abs(w_dC).max()
    assert packed[2 + A * A:].sum() == (2 * n - 2) * L


# ---- 2-GPU NCCL tests -----------------------------------------------------------------------------


def _nccl_worker(rank, world, port, q):
  import torch
  import torch.distributed as dist
  from oracle import tree_oracle as to
  from trex_b200.dist import score_trees_sharded, shard_range, tree_cost_value_and_grad_sharded
  from trex_b200.tree import tree_cost_value_and_grad

  os.environ["MASTER_ADDR"] = "127.0.0.1"
  os.environ["MASTER_PORT"] = str(port)
  torch.cuda.set_device(rank)
  dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
  rng = np.random.default_rng(4)
  n, A, L, B = 32, 20, 500, 9
  n_all = 2 * n - 1
  adj = np.stack([synth.parents_to_adjacency(synth.random_join_parents(n, rng)) for _ in range(B)])
  states = np.ascontiguousarray(synth.evolve(synth.balanced_parents(n), L, A, 0.2, rng)[:n])
  cost = synth.unit_cost(A)
  scores = score_trees_sharded(torch.from_numpy(adj).cuda(), torch.from_numpy(cost).cuda(),
                               torch.from_numpy(states).cuda(), A)
  want = [float(so.per_site_root_min(so.run_sankoff_tables(a, cost, states, n_all, A, n)["dp"]).astype(np.float64).sum())
          for a in adj]
  ok_scores = scores.tolist() == want
  # tree cost sharded over the sites vs the single-GPU call on the whole array
  N, Ls, Q = 127, 1000, 4
  S = to.softmax(rng.normal(size=(N, Ls, Q))).astype(np.float32)
  A2 = (rng.random((N, N)) < 0.1).astype(np.float32)
  C = rng.random((Q, Q)).astype(np.float32)
  errs = {}
  for name, cm in (("identity", None), ("matrix", C)):
    lo, hi = shard_range(Ls, world, rank)
    cmd = None if cm is None else torch.from_numpy(cm).cuda()
    c, dS, dA, dC = tree_cost_value_and_grad_sharded(torch.from_numpy(np.ascontiguousarray(S[:, lo:hi])).cuda(),
                                                     torch.from_numpy(A2).cuda(), cmd)
    w, wdS, wdA, wdC = to.soft_cost_value_and_grad(S, A2, cm)
    errs[name] = (abs(float(c) - w) / abs(w), float(np.abs(dS.cpu().numpy() - wdS[:, lo:hi]).max() / np.abs(wdS).max()),
                  float(np.abs(dA.cpu().numpy() - wdA).max() / np.abs(wdA).max()),
                  0.0 if cm is None else float(np.abs(dC.cpu().numpy() - wdC).max() / np.abs(wdC).max()))
  q.put((rank, ok_scores, errs))
  dist.barrier()
  dist.destroy_process_group()


def test_two_gpu_tree_batch_sharding_and_site_sharded_tree_cost_nccl():
  import torch

  if torch.cuda.device_count() < 2:
    pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
  import torch.multiprocessing as mp

  ctx = mp.get_context("spawn")
  q = ctx.Queue()
  port = 35500 + os.getpid() % 2000
  procs = [ctx.Process(target=_nccl_worker, args=(r, 2, port, q)) for r in range(2)]
  for p in procs:
    p.start()
  res = [q.get(timeout=300) for _ in procs]
  for p in procs:
    p.join(timeout=120)
    assert p.exitcode == 0
  for _, ok_scores, errs in res:
    assert ok_scores                                  # exact integer scores of all 9 trees on both ranks
    for name, e in errs.items():
      assert max(e) <= 2e-5, (name, e)                # 1e-5 bar + the cancellation allowance of SURVEY H8
