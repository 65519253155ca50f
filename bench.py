#!/usr/bin/env python
"""bench.py — headline benchmark of the trex_b200 hot path.

  python bench.py --gpus N --steps K --warmup W            (N>1: launched by torchrun)
  python bench.py --impl reference ...                      (CPU port of the reference, rank 0)

Metric (BASELINE.json): Sankoff fwd+bwd DP cell-updates/s, with the tree-cost grad steps/s of
config c2 reported in the same JSON line under "tree_cost".

Workload at every N: BASELINE config c5 — 1024 taxa x 100k sites, A=20, balanced tree, unit
cost, synthetic edge-mutation alignment.  One step = one pass of the hot path over
the rank's sites: device-built plan, hard forward (dp + argmins + score), traceback (+ edge
histogram), relaxed forward, relaxed backward (marginals, leaf gradient, dC).  One DP
cell-update = one dp[node, site, state] value of an internal node in one pass, so a step is
4 * (n-1) * L * A cell-updates (SURVEY §8d unit definitions).  Sites are sharded over ranks
(weak scaling by default: 100k sites per GPU; --scaling strong splits the 100k), and one packed
NCCL all-reduce per step combines the scalar costs, dC and the edge histogram.
"""

from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
  sys.path.insert(0, ROOT)

N_LEAVES, N_STATES, N_SITES = 1024, 20, 100_000
SOFT_T = 1.0
METRIC = "sankoff_fwd_bwd_dp_cell_updates_per_s"
UNIT = "cell-updates/s"


def peaks():
  p = os.path.join(ROOT, "MEASURED_PEAKS.json")
  if os.path.exists(p):
    with open(p) as f:
      return float(json.load(f)["hbm_gbs"]), "measured"
  return 6650.0, "fallback"


def alg_bytes_per_site(n, A):
  """SURVEY §8(d) algorithmic (compulsory) bytes per site and tree, per pass."""
  return {
    "hard_fwd": n + (2 * n - 3) * 4 * A + (n - 1) * 2 * A,
    "hard_bwd": 3 * (n - 1) + 1,
    "soft_fwd": n + (2 * n - 3) * 4 * A,
    "soft_bwd": n + (4 * n - 5) * 4 * A,
  }


def make_inputs(n, A, L, seed):
  from trex_b200 import synth

  rng = np.random.default_rng(seed)
  par = synth.balanced_parents(n)
  adj = synth.parents_to_adjacency(par)
  states = synth.evolve(par, L, A, 0.1, rng)[:n]
  return adj, synth.unit_cost(A), np.ascontiguousarray(states)


class ClockSampler:
  """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""

  Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
       "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

  def __init__(self, index):
    self.rows, self.proc, self.index = [], None, index

  def start(self):
    try:
      self.proc = subprocess.Popen(
        ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index), "-lms", "100"],
        stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
      threading.Thread(target=self._read, daemon=True).start()
    except OSError:
      self.proc = None

  def _read(self):
    for line in self.proc.stdout:
      self.rows.append(line.strip())

  def stop(self):
    if self.proc is None:
      return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
    self.proc.terminate()
    sm, mx, reasons = [], [], set()
    names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    for r in self.rows:
      f = [x.strip() for x in r.split(",")]
      try:
        sm.append(float(f[0]))
        mx.append(float(f[1]))
      except (ValueError, IndexError):
        continue
      for nm, val in zip(names, f[2:6]):
        if val.lower().startswith("active"):
          reasons.add(nm)
    return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
            "samples": len(sm), "reasons": sorted(reasons)}


# ----------------------------------------------------------------------------------------------
# CPU arm: the plain-C port of the reference (oracle/), all host threads
# ----------------------------------------------------------------------------------------------


def cpu_step(adj, cost, seqs_f32, n, A):
  from oracle import c_oracle

  n_all = 2 * n - 1
  c_oracle.run_sankoff(adj, cost, seqs_f32, n_all, A, n, return_path=True, want_dp=False)
  c_oracle.soft_sankoff(adj, cost, seqs_f32, n_all, A, n, SOFT_T)


def cpu_baseline(sample_sites):
  from oracle import c_oracle

  adj, cost, states = make_inputs(N_LEAVES, N_STATES, sample_sites, seed=5)
  seqs = states.astype(np.float32)
  cpu_step(adj, cost, seqs[:, :256].copy(), N_LEAVES, N_STATES)  # warm the thread pool
  t0 = time.perf_counter()
  cpu_step(adj, cost, seqs, N_LEAVES, N_STATES)
  dt = time.perf_counter() - t0
  cells = 4 * (N_LEAVES - 1) * sample_sites * N_STATES
  return {"value": cells / dt, "unit": UNIT, "cores": c_oracle.num_threads(), "kind": "port",
          "sample": f"{sample_sites} of {N_SITES} sites of the c5 alignment, one pass (hard fwd+traceback, "
                    f"relaxed fwd+bwd), plain-C OpenMP port of sankoff.py (JAX not installable)",
          "seconds": dt}


def run_reference(args):
  rank = int(os.environ.get("RANK", "0"))
  if rank != 0:
    return
  from oracle import c_oracle

  sample = args.ref_sites
  adj, cost, states = make_inputs(N_LEAVES, N_STATES, sample, seed=5)
  seqs = states.astype(np.float32)
  for _ in range(args.warmup):
    cpu_step(adj, cost, seqs, N_LEAVES, N_STATES)
  t0 = time.perf_counter()
  for _ in range(args.steps):
    cpu_step(adj, cost, seqs, N_LEAVES, N_STATES)
  dt = time.perf_counter() - t0
  cells = 4 * (N_LEAVES - 1) * sample * N_STATES * args.steps
  val = cells / dt
  desc = (f"{sample} of {N_SITES} sites per step; plain-C OpenMP port of the reference's sankoff.py "
          f"(+ relaxed restatement); the reference itself is JAX, not installable offline")
  line = {
    "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
    "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
    "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32",
    "data": "synthetic", "config": workload_config(args, sample),
    "cpu_baseline": {"value": val, "unit": UNIT, "cores": c_oracle.num_threads(), "kind": "port", "sample": desc},
    "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    "gpu_launches": 0,
  }
  emit(line)


def workload_config(args, sites_per_rank):
  return {"workload": "c5: 1024 taxa x 100k sites, A=20, balanced tree, unit cost; hard fwd + traceback + "
                      "relaxed fwd + bwd (T=1), site-sharded",
          "n_leaves": N_LEAVES, "n_states": N_STATES, "n_sites_total_c5": N_SITES,
          "sites_per_rank": int(sites_per_rank), "parallelism": f"sites{args.gpus}",
          "l2": "inputs_exceed_l2 (tables are GBs per pass; no explicit flush)",
          "step_streams": 2 if sites_per_rank <= args.overlap_below else 1}


# ----------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------


def tree_cost_bench(torch, iters=200):
  """BASELINE config c2: soft-topology tree_cost + grad, 64 taxa (127 nodes) x 1k sites, DNA, and the
  largest shape of the reference's own script (benchmark_tree_cost.py:41-43: N=200, L=1000, Q=20).
  Identity cost -> tcgen05 path.  Timed twice: replayed from a CUDA graph (device-bound) and
  dispatched call by call from Python through ctypes (what an eager caller sees)."""
  from oracle import tree_oracle as to
  from trex_b200 import _lib
  from trex_b200.tree import tree_cost_value_and_grad

  def one(N, L, Q, n, seed, softmax_adj):
    rng = np.random.default_rng(seed)
    S = to.softmax(rng.normal(size=(N, L, Q))).astype(np.float32)
    if n:
      S[:n] = np.eye(Q, dtype=np.float32)[rng.integers(0, Q, size=(n, L))]
    if softmax_adj:
      adj = to.update_tree(rng.normal(size=(N - 1, N - n)).astype(np.float32),
                           np.zeros((N - 1, N - n), np.float32)).astype(np.float32)
    else:
      adj = (rng.random((N, N)) < 0.1).astype(np.float32)  # benchmark_tree_cost.py:57
    St, At = torch.from_numpy(S).cuda(), torch.from_numpy(adj).cuda()
    ws = torch.empty(_lib.load().trex_tree_cost_workspace_bytes(N, L * Q, Q), dtype=torch.uint8, device="cuda")
    out = {"cost": torch.empty((), device="cuda"), "dS": torch.empty_like(St), "dA": torch.empty_like(At)}
    call = lambda: tree_cost_value_and_grad(St, At, None, workspace=ws, out=out)
    for _ in range(10):
      call()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
      call()
    e1.record()
    torch.cuda.synchronize()
    ms_eager = e0.elapsed_time(e1) / iters
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
      call()
      stream.synchronize()
      g = torch.cuda.CUDAGraph()
      with torch.cuda.graph(g, stream=stream):
        call()
      g.replay()
      stream.synchronize()
      e0.record(stream)
      for _ in range(iters):
        g.replay()
      e1.record(stream)
      stream.synchronize()
    ms = e0.elapsed_time(e1) / iters
    t0 = time.perf_counter()
    reps = 3
    for _ in range(reps):
      to.soft_cost_value_and_grad(S, adj, None, dtype=np.float32)
    cpu_ms = (time.perf_counter() - t0) / reps * 1e3
    flop = 4.0 * N * N * L * Q
    return {"shape": f"{N} nodes x {L} sites x {Q} states, identity cost, value + dS + dAdj",
            "ms_per_step": ms, "grad_steps_per_s": 1e3 / ms, "ms_per_step_eager_python": ms_eager,
            "flop_per_step": flop, "useful_tflops": flop / ms / 1e9, "tf32x3_issued_tflops": 3 * flop / ms / 1e9,
            "cpu_grad_steps_per_s": 1e3 / cpu_ms}

  c2 = one(127, 1000, 4, 64, 2, True)
  big = one(200, 1000, 20, 0, 242, False)
  bf16 = None
  pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
  if os.path.exists(pk):
    with open(pk) as f:
      bf16 = float(json.load(f).get("bf16_tflops", 0)) or None
  tf32_peak = (bf16 or 1590.0) / 2  # no TF32 figure is measured on this pool; dense tf32 = bf16 / 2
  return {"metric": "tree_cost_grad_steps_per_s", "value": c2["grad_steps_per_s"], "unit": "grad steps/s",
          "ms_per_step": c2["ms_per_step"], "config": "c2: " + c2["shape"], "c2": c2, "script_n200": big,
          "path": "tcgen05 kind::tf32, 3xTF32 split, TMA + TMEM (tree_cost_tc.cu); CUDA-graph replay",
          "roofline": {"bound": "tensor", "achieved": big["tf32x3_issued_tflops"], "peak": tf32_peak,
                       "unit": "TFLOP/s", "frac": big["tf32x3_issued_tflops"] / tf32_peak,
                       "note": "issued tf32 flops (3 MMAs per product) at N=200, K=20000 over measured bf16 peak / 2; "
                               "c2 itself is launch/latency bound (SURVEY H7)"},
          "cpu_baseline": {"value": c2["cpu_grad_steps_per_s"], "unit": "grad steps/s", "kind": "port",
                           "sample": "NumPy/BLAS float32 restatement of tree.py value+grad at c2"}}


def fit_loop_bench(torch):
  """the whole optimiser loop of run_trex_optimization_configurable
  (evals/benchmark.py:75-200; adam + clip_by_global_norm(1.0), lr 1e-3, its defaults) as one persistent
  kernel.  One "grad step" = loss value + gradient w.r.t. the ancestor logits + optimiser update.
  device: CUDA events around one library call of `iters` iterations (setup kernels included);
  e2e: host leaf states / adjacency / initial logits in pinned memory -> argmax + loss curve back on the host."""
  from oracle import optimize_oracle as oo
  from trex_b200 import optimize as O
  from trex_b200 import synth

  def one(n, L, Q, seed, iters):
    rng = np.random.default_rng(seed)
    parent = synth.balanced_parents(n)
    N = parent.shape[0]
    leaf = synth.evolve(parent, L, Q, 0.1, rng)[:n]
    adj = synth.parents_to_adjacency(parent).astype(np.float32)
    init = rng.normal(size=(N - n, L, Q)).astype(np.float32)
    opt = O.create_optimizer("adam", 1e-3)
    fit = O.AncestorFit(leaf, N, n, Q, adj, init, opt)
    losses = torch.empty(iters, device="cuda")
    fit.run(iters, losses)  # warm-up (module load, attribute set)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fit.run(iters, losses)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    # end to end through the reference-shaped call, host buffers in and out
    leaf_p = torch.from_numpy(leaf.astype(np.int32)).pin_memory()
    adj_p = torch.from_numpy(adj).pin_memory()
    init_p = torch.from_numpy(init).pin_memory()
    out_h = torch.empty((N - n, L), dtype=torch.int32).pin_memory()
    loss_h = torch.empty(iters, dtype=torch.float32).pin_memory()
    def e2e():
      a, l = O.run_trex_optimization_configurable(
        leaf_p.cuda(non_blocking=True), N, n, Q, adj_p.cuda(non_blocking=True), None, n_iterations=iters,
        return_losses=True, init_ancestors=init_p.cuda(non_blocking=True))
      out_h.copy_(a, non_blocking=True)
      loss_h.copy_(l, non_blocking=True)
    e2e()
    torch.cuda.synchronize()
    e0.record()
    e2e()
    e1.record()
    torch.cuda.synchronize()
    ms_e2e = e0.elapsed_time(e1)
    cpu_iters = 3
    t0 = time.perf_counter()
    oo.run_trex_optimization_configurable(leaf, N, n, Q, adj, init, "adam", 1e-3, cpu_iters, dtype=np.float32)
    cpu_s = (time.perf_counter() - t0) / cpu_iters
    return {"shape": f"{N} nodes ({n} leaves) x {L} sites x {Q} states, balanced tree, adam lr 1e-3 + clip 1.0",
            "iterations": iters, "us_per_step": ms * 1e3 / iters, "grad_steps_per_s": iters / (ms * 1e-3),
            "e2e_grad_steps_per_s": iters / (ms_e2e * 1e-3),
            "e2e_h2d_bytes": int(leaf_p.numel() * 4 + adj_p.numel() * 4 + init_p.numel() * 4),
            "e2e_d2h_bytes": int(out_h.numel() * 4 + loss_h.numel() * 4),
            "final_loss": float(losses[-1]), "first_loss": float(losses[0]),
            "cpu_grad_steps_per_s": 1.0 / cpu_s}

  c2 = one(64, 1000, 4, 2, 10000)
  q20 = one(100, 1000, 20, 242, 2000)
  return {"metric": "fit_ancestors_grad_steps_per_s", "value": c2["grad_steps_per_s"], "unit": "grad steps/s",
          "config": "c2: " + c2["shape"], "c2": c2, "script_q20": q20,
          "path": "fit_ancestors_kernel: persistent cooperative kernel, 1 grid barrier per iteration, state resident in shared memory",
          "cpu_baseline": {"value": c2["cpu_grad_steps_per_s"], "unit": "grad steps/s", "kind": "port",
                           "sample": "3 iterations of the NumPy float32 restatement of the loop at c2 (optax/JAX absent)"}}


def landscape_bench(torch, iters=50):
  """the parental-guidance term of _compute_loss_landscape_aware_stacked
  (evals/benchmark.py:283-303) — value + d/dS in one library call, NK landscape with K = 1, q = 20
  (evals/benchmark.py:922-929) at 255 nodes x 1000 sites; CPU: the NumPy float64 restatement, one call."""
  from oracle import landscape_oracle as lo
  from oracle import tree_oracle as to
  from trex_b200 import landscape as LS
  from trex_b200 import synth

  n, L, Q, k = 128, 1000, 20, 1
  rng = np.random.default_rng(7)
  parent = synth.balanced_parents(n)
  N = parent.shape[0]
  adj = synth.parents_to_adjacency(parent).astype(np.float32)
  S = to.softmax(rng.normal(size=(N, L, Q))).astype(np.float32)
  inter = rng.integers(0, L, size=(L, k)).astype(np.int32)
  tables = rng.random((L, Q ** (k + 1))).astype(np.float32)
  land = {"interactions": torch.from_numpy(inter).cuda(), "fitness_tables": torch.from_numpy(tables).cuda()}
  St, At = torch.from_numpy(S).cuda(), torch.from_numpy(adj).cuda()
  it_, tb, kk = LS._landscape_arrays(land, k, St.device)
  call = lambda: LS._call(St, At, None, it_, tb, None, kk, True, False, True)
  for _ in range(5):
    call()
  e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  torch.cuda.synchronize()
  e0.record()
  for _ in range(iters):
    cost, _, dS = call()
  e1.record()
  torch.cuda.synchronize()
  ms = e0.elapsed_time(e1) / iters
  t0 = time.perf_counter()
  ref = lo.fitness_cost(S, adj, inter, tables, k)
  lo.fitness_cost_grad(S, adj, inter, tables, k)
  cpu_ms = (time.perf_counter() - t0) * 1e3
  return {"shape": f"{N} nodes x {L} sites x {Q} states, K = {k}", "ms_per_value_and_grad": ms,
          "value_and_grads_per_s": 1e3 / ms, "cpu_ms_per_value_and_grad": cpu_ms,
          "rel_err_vs_float64_port": abs(float(cost) - ref) / abs(ref)}


def run_gpu(args):
  import torch
  import torch.distributed as dist

  from trex_b200 import _lib
  from trex_b200.sankoff import SankoffEngine

  rank = int(os.environ.get("RANK", "0"))
  world = int(os.environ.get("WORLD_SIZE", "1"))
  local = int(os.environ.get("LOCAL_RANK", "0"))
  if world != args.gpus:
    if world == 1 and args.gpus > 1:
      raise SystemExit("--gpus N>1 must be launched with torch.distributed.run (one rank per GPU)")
  torch.cuda.set_device(local)
  if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
  lib = _lib.load()

  n, A = N_LEAVES, N_STATES
  n_all = 2 * n - 1
  L = args.sites if args.sites else (N_SITES if args.scaling == "weak" else (N_SITES + world - 1) // world)
  adj_h, cost_h, states_h = make_inputs(n, A, L, seed=5 + rank)
  # pinned host buffers: the e2e leg copies these every step
  adj_p = torch.from_numpy(adj_h).pin_memory()
  cost_p = torch.from_numpy(cost_h).pin_memory()
  states_p = torch.from_numpy(states_h).pin_memory()
  adj_d, cost_d, states_d = adj_p.cuda(), cost_p.cuda(), states_p.cuda()

  eng = SankoffEngine(n_all, n, A, L)
  from trex_b200.dist import PackedAllReduce
  reducer = PackedAllReduce([(1,), (1,), (1, A, A), (1, A, A)], device="cuda")
  packed_h = torch.empty(2 + 2 * A * A, dtype=torch.float64).pin_memory()
  names = ["plan", "hard_fwd", "hard_bwd", "soft_fwd", "soft_bwd", "reduce"]

  # Short alignments (a GPU's share under strong scaling) leave SMs idle in the top part of every walk and
  # in the wave tails: there the hard pass and the relaxed pass, which are independent, run on two streams.
  # (No gain at 100k sites per GPU — 15.13 vs 15.06 ms — so the long-alignment step stays on one stream and
  # keeps its per-kernel event timings; with two streams the per-kernel times come from a serial pass.)
  two_streams = L <= args.overlap_below
  stream_a, stream_b = (torch.cuda.Stream(), torch.cuda.Stream()) if two_streams else (None, None)

  def step_two_streams():
    main = torch.cuda.current_stream()
    eng.set_adjacency(adj_d)
    stream_a.wait_stream(main)
    stream_b.wait_stream(main)
    with torch.cuda.stream(stream_a):
      eng.hard_forward(cost_d, want_argmin=True)
      eng.traceback(want_hist=True)
    with torch.cuda.stream(stream_b):
      eng.soft_forward(cost_d, SOFT_T)
      eng.soft_backward(cost_d, SOFT_T, want_dleaf=True, want_dcost=True)
    main.wait_stream(stream_a)
    main.wait_stream(stream_b)
    reducer([eng.total, eng.soft_total, eng.dcost, eng.edge_hist])

  def step(ev=None):
    if two_streams and ev is None:
      return step_two_streams()

    def mark(i):
      if ev is not None:
        ev[i].record()
    mark(0)
    eng.set_adjacency(adj_d)
    mark(1)
    eng.hard_forward(cost_d, want_argmin=True)
    mark(2)
    eng.traceback(want_hist=True)
    mark(3)
    eng.soft_forward(cost_d, SOFT_T)
    mark(4)
    eng.soft_backward(cost_d, SOFT_T, want_dleaf=True, want_dcost=True)
    mark(5)
    # one packed all-reduce: [hard total, soft total, dC (A*A), edge histogram (A*A)]
    reducer([eng.total, eng.soft_total, eng.dcost, eng.edge_hist])
    mark(6)

  eng.set_leaf_states_u8(states_d)
  for _ in range(args.warmup):
    step()
  torch.cuda.synchronize()

  def barrier():
    if world > 1:
      dist.barrier()
    torch.cuda.synchronize()

  # ---- timed region (device resident inputs) ------------------------------------------------
  evs = [[torch.cuda.Event(enable_timing=True) for _ in range(7)] for _ in range(args.steps)]
  e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  sampler = ClockSampler(local)
  if rank == 0:
    sampler.start()
  barrier()
  launches0 = int(lib.trex_kernel_launches())
  e0.record()
  for k in range(args.steps):
    step(None if two_streams else evs[k])
  e1.record()
  barrier()
  # kernels of libtrex_b200.so launched in the timed region, counted by the library itself (the
  # device-side dispatch launches every candidate kernel of an operator; all but one return at once)
  gpu_launches = int(lib.trex_kernel_launches()) - launches0
  if two_streams:  # per-kernel times of the same work, one stream (outside the timed region)
    for k in range(args.steps):
      step(evs[k])
    torch.cuda.synchronize()
  ms_total = e0.elapsed_time(e1)
  clocks = sampler.stop() if rank == 0 else None
  kern_ms = {nm: float(np.mean([evs[k][i].elapsed_time(evs[k][i + 1]) for k in range(args.steps)]))
             for i, nm in enumerate(names)}

  # ---- end-to-end: host buffers in, host results out, through the public API ------------------
  # Successive steps are independent alignments, so the pinned-host -> device copies of step k+1 (leaf
  # states, adjacency, cost matrix) run on a copy stream under the kernels of step k: two leaf tables and
  # two adjacency / cost buffers alternate; every other table is shared.  All K copies happen inside the
  # timed region (the first one is not overlapped with anything).
  engs = [eng, eng.site_block(0, L)]
  engs[1]._bufs["leaf"] = torch.empty_like(eng.leaf)
  engs[1]._pad_owner = True
  adj_x = [torch.empty_like(adj_d) for _ in range(2)]
  cost_x = [torch.empty_like(cost_d) for _ in range(2)]
  copy_stream = torch.cuda.Stream()

  def upload(k):
    j = k & 1
    adj_x[j].copy_(adj_p, non_blocking=True)
    cost_x[j].copy_(cost_p, non_blocking=True)
    engs[j].set_leaf_states_u8(states_p)

  def e2e_run(n_steps):
    main = torch.cuda.current_stream()
    ready = [torch.cuda.Event() for _ in range(n_steps)]
    done = [torch.cuda.Event() for _ in range(n_steps)]
    copy_stream.wait_stream(main)
    with torch.cuda.stream(copy_stream):
      upload(0)
      ready[0].record(copy_stream)
    for k in range(n_steps):
      j = k & 1
      e = engs[j]
      main.wait_event(ready[k])
      if k + 1 < n_steps:
        with torch.cuda.stream(copy_stream):
          if k >= 1:
            copy_stream.wait_event(done[k - 1])  # step k-1 used the buffers step k+1 uploads into
          upload(k + 1)
          ready[k + 1].record(copy_stream)
      e.set_adjacency(adj_x[j])
      e.hard_forward(cost_x[j], want_argmin=True)
      e.traceback(want_hist=True)
      e.soft_forward(cost_x[j], SOFT_T)
      e.soft_backward(cost_x[j], SOFT_T, want_dleaf=True, want_dcost=True)
      done[k].record(main)
      reducer([e.total, e.soft_total, e.dcost, e.edge_hist])
      packed_h.copy_(reducer.buf, non_blocking=True)

  e2e_run(2)
  barrier()
  f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  f0.record()
  e2e_run(args.steps)
  f1.record()
  barrier()
  ms_e2e = f0.elapsed_time(f1)

  t = torch.tensor([ms_total, ms_e2e], dtype=torch.float64, device="cuda")
  if world > 1:
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
  ms_total, ms_e2e = float(t[0]), float(t[1])
  cells_step = 4 * (n - 1) * L * A * world
  value = cells_step * args.steps / (ms_total * 1e-3)
  e2e_value = cells_step * args.steps / (ms_e2e * 1e-3)

  if rank == 0:
    peak, peak_src = peaks()
    per_site = alg_bytes_per_site(n, A)
    roof_all = {}
    for nm in ["hard_fwd", "hard_bwd", "soft_fwd", "soft_bwd"]:
      gbs = per_site[nm] * L / (kern_ms[nm] * 1e-3) / 1e9
      roof_all[nm] = {"ms": kern_ms[nm], "alg_bytes": per_site[nm] * L, "achieved": gbs, "frac": gbs / peak}
    dom = max(["hard_fwd", "soft_fwd", "soft_bwd"], key=lambda k: kern_ms[k])
    # dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full captures
    # (forward kernels: profiles/r01_n_fwd_uniform_ncu_full.txt, segment + top launches: 12.22 + 0.60 and 8.24 + 0.48 GB;
    # soft_bwd re-captured at HEAD with both segment launches, profiles/r01_k_soft_bwd_uniform_ncu_full.txt:
    # 0.98 + 24.28 GB), valid for the c5 shape at 100k sites per rank only
    ncu_traffic = {"hard_fwd": 12.79e9, "soft_fwd": 8.71e9, "soft_bwd": 25.26e9}
    traffic = ncu_traffic.get(dom) if (L == N_SITES) else None
    line = {
      "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
      "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
      "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
      "config": workload_config(args, L),
      "roofline": {"bound": "hbm", "kernel": dom, "achieved": roof_all[dom]["achieved"], "peak": peak,
                   "peak_source": peak_src, "unit": "GB/s", "frac": roof_all[dom]["frac"], "traffic": traffic,
                   "note": "achieved = SURVEY 8(d) algorithmic bytes / CUDA-event launch time on the launching stream"},
      "roofline_all": roof_all,
      "roofline_step": {"bound": "hbm", "alg_bytes": sum(per_site.values()) * L,
                        "achieved": sum(per_site.values()) * L / (ms_total / args.steps * 1e-3) / 1e9, "peak": peak,
                        "unit": "GB/s", "frac": sum(per_site.values()) * L / (ms_total / args.steps * 1e-3) / 1e9 / peak,
                        "note": "all four passes of the step: SURVEY 8(d) algorithmic bytes / step time"},
      "kernel_ms": kern_ms,
      "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": ms_e2e / args.steps,
              "h2d_bytes_per_step": int(states_p.numel() + adj_p.numel() * 4 + cost_p.numel() * 4),
              "d2h_bytes_per_step": int(packed_h.numel() * 8),
              "overlap": "H2D of step k+1 on a copy stream under the kernels of step k (double-buffered inputs)",
              "results": {"hard_total": float(packed_h[0]), "soft_total": float(packed_h[1])}},
      "gpu_launches": gpu_launches,
      "clocks": clocks,
      "results": {"hard_total": float(packed_h[0]), "soft_total": float(packed_h[1])},
    }
    if world == 1 and not args.no_extras:
      line["tree_cost"] = tree_cost_bench(torch)
      line["fit_loop"] = fit_loop_bench(torch)
      line["landscape_loss"] = landscape_bench(torch)
      line["cpu_baseline"] = cpu_baseline(args.cpu_sites)
    emit(line)
  if world > 1:
    dist.destroy_process_group()


_REAL_STDOUT = None


def quiet_stdout():
  """The contract is ONE JSON line on stdout.  Libraries (NCCL's version banner under NCCL_DEBUG, torch
  warnings) also write to fd 1, so fd 1 is pointed at stderr for the run and the JSON line goes to the
  saved descriptor."""
  global _REAL_STDOUT
  sys.stdout.flush()
  _REAL_STDOUT = os.dup(1)
  os.dup2(2, 1)


def emit(line):
  data = (json.dumps(line) + "\n").encode()
  if _REAL_STDOUT is None:
    sys.stdout.write(data.decode())
    sys.stdout.flush()
  else:
    os.write(_REAL_STDOUT, data)


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument("--gpus", type=int, default=1)
  ap.add_argument("--steps", type=int, default=10)
  ap.add_argument("--warmup", type=int, default=3)
  ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
  ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
  ap.add_argument("--sites", type=int, default=0, help="override sites per rank (testing)")
  ap.add_argument("--cpu-sites", type=int, default=4096, help="sites of the bounded cpu_baseline sample")
  ap.add_argument("--ref-sites", type=int, default=2048, help="sites per step of --impl reference")
  ap.add_argument("--overlap-below", type=int, default=50000,
                  help="sites per rank at or below which the hard and relaxed passes run on two streams")
  ap.add_argument("--no-extras", action="store_true", help="skip tree_cost and cpu_baseline legs")
  args = ap.parse_args()
  quiet_stdout()
  args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
  if args.impl == "reference":
    run_reference(args)
  else:
    run_gpu(args)


if __name__ == "__main__":
  main()
