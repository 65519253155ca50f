#!/usr/bin/env python
"""bench.py — headline benchmark of the trex_b200 hot path.

  python bench.py --gpus N --steps K --warmup W            (N>1: launched by torchrun)
  python bench.py --impl reference ...                      (CPU port of the reference, rank 0)

Metric (BASELINE.json): Sankoff fwd+bwd DP cell-updates/s, with the tree-cost grad steps/s of
config c2 reported in the same JSON line under "tree_cost".

Workload at every N: BASELINE config c5 — ONE alignment of 1024 taxa x 100k sites, A=20, balanced tree, unit
cost, synthetic edge-mutation alignment.  One step = one pass of the hot path over the rank's
sites: device-built plan, hard forward (dp + argmins + score), traceback (+ edge histogram), relaxed forward,
relaxed backward (marginals, compact leaf gradient, dC).  One DP cell-update = one dp[node, site, state] value
of an internal node in one pass, so a step is 4 * (n-1) * L * A cell-updates (SURVEY §8d unit definitions).
Sites are sharded over the ranks — STRONG scaling by default: the 100k sites of the one alignment are split N
ways (the north-star case); `--scaling weak` gives every rank its own 100k sites, and at N > 1 a weak-scaling
record is added to the line as `weak_scaling`.  One packed NCCL all-reduce per step combines the scalar costs,
dC and the edge histogram.  The device-resident leg replays the step from a CUDA graph (the all-reduce is
issued on a communication stream and overlaps the next step's kernels — steps are independent jobs — and every
all-reduce is joined before the closing timing event); the e2e leg goes through trex_b200.pipeline.SankoffHostPipeline
with pinned host inputs and host results (ancestral states + packed scalars) every step.
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


def measured_peaks():
  p = os.path.join(ROOT, "MEASURED_PEAKS.json")
  if os.path.exists(p):
    with open(p) as f:
      return json.load(f), "measured (MEASURED_PEAKS.json)"
  return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback (B200_PROFILING.md)"


def alg_bytes_per_site(n, A, words=False):
  """SURVEY §8(d) algorithmic (compulsory) bytes per site and tree, per pass.  ``words``: the back-pointers are pointer
  words (TREX_MODE_COMPACT_ARG, uniform cost: 8 bytes per node and site, all of them read by the traceback) instead
  of the byte table (2A bytes per node and site written, 2 of them read)."""
  return {
    "hard_fwd": n + (2 * n - 3) * 4 * A + (n - 1) * (8 if words else 2 * A),
    "hard_bwd": ((8 if words else 2) + 1) * (n - 1) + 1,
    "soft_fwd": n + (2 * n - 3) * 4 * A,
    "soft_bwd": n + (4 * n - 5) * 4 * A,
  }


def ncu_traffic():
  """dram__bytes_read.sum + dram__bytes_write.sum per launch, from the ncu --set full captures committed under
  profiles/ (profiles/ncu_traffic.json names the capture each number comes from).  Keyed by kernel leg; valid
  for the c5 shape at 100k sites per rank only."""
  p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
  if os.path.exists(p):
    with open(p) as f:
      return json.load(f)
  return {}


_INPUT_CACHE = {}


def make_inputs(n, A, L, seed):
  from trex_b200 import synth

  key = (n, A, L, seed)
  if key not in _INPUT_CACHE:
    rng = np.random.default_rng(seed)
    par = synth.balanced_parents(n)
    adj = synth.parents_to_adjacency(par)
    states = synth.evolve(par, L, A, 0.1, rng)[:n]
    _INPUT_CACHE[key] = (adj, synth.unit_cost(A), np.ascontiguousarray(states))
  return _INPUT_CACHE[key]


class ClockSampler:
  """SM clock and throttle reasons sampled DURING the timed region (B200_PROFILING.md): an in-process NVML
  thread (about 1 kHz, so that even a 20 ms timed region gets samples), nvidia-smi as the fallback."""

  BITS = {"hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40, "sw_power_cap": 0x4}

  def __init__(self, index):
    self.index, self.rows, self.stop_flag, self.thread, self.proc = index, [], False, None, None
    self.window = None
    self.nvml = None

  def start(self):
    try:
      import pynvml

      pynvml.nvmlInit()
      vis = os.environ.get("CUDA_VISIBLE_DEVICES")
      idx = int(vis.split(",")[self.index]) if vis and vis.split(",")[self.index].isdigit() else self.index
      self.handle = pynvml.nvmlDeviceGetHandleByIndex(idx)
      self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
      self.nvml = pynvml
      self.thread = threading.Thread(target=self._poll, daemon=True)
      self.thread.start()
    except Exception:
      self.nvml = None
      self._start_smi()

  def _poll(self):
    nv = self.nvml
    reasons_fn = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
    while not self.stop_flag:
      try:
        mhz = float(nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM))
        bits = int(reasons_fn(self.handle))
        self.rows.append((time.perf_counter(), mhz, bits))
      except Exception:
        pass
      time.sleep(0.001)

  def _start_smi(self):
    q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    try:
      self.proc = subprocess.Popen(
        ["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i", str(self.index), "-lms", "100"],
        stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
      threading.Thread(target=self._read_smi, daemon=True).start()
    except OSError:
      self.proc = None

  def _read_smi(self):
    names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    for line in self.proc.stdout:
      f = [x.strip() for x in line.split(",")]
      try:
        mhz, self.max_mhz = float(f[0]), float(f[1])
      except (ValueError, IndexError):
        continue
      bits = sum(self.BITS[nm] for nm, val in zip(names, f[2:6]) if val.lower().startswith("active"))
      self.rows.append((time.perf_counter(), mhz, bits))

  def mark(self, t0, t1):
    """host-clock window of the timed region (the samples inside it are reported as 'under load')"""
    self.window = (t0, t1)

  def stop(self):
    self.stop_flag = True
    if self.proc is not None:
      self.proc.terminate()
    if self.thread is not None:
      self.thread.join(timeout=1.0)
    if not self.rows:
      return {"sm_mhz": None, "sm_max_mhz": None, "samples": 0, "reasons": ["no clock samples"]}
    rows = self.rows
    inside = [r for r in rows if self.window and self.window[0] <= r[0] <= self.window[1]]
    use = inside if inside else rows
    bits = 0
    for r in use:
      bits |= r[2]
    return {"sm_mhz": float(np.median([r[1] for r in use])), "sm_max_mhz": getattr(self, "max_mhz", None),
            "samples": len(use), "samples_whole_run": len(rows),
            "sm_mhz_whole_run": float(np.median([r[1] for r in rows])),
            "source": "nvml" if self.nvml is not None else "nvidia-smi",
            "reasons": sorted(nm for nm, b in self.BITS.items() if bits & b)}


# ----------------------------------------------------------------------------------------------
# CPU arm: the plain-C port of the reference (oracle/), all host threads
# ----------------------------------------------------------------------------------------------


def host_threads():
  """torch.distributed.run exports OMP_NUM_THREADS=1; the CPU arm uses every core the process may run on."""
  from oracle import c_oracle

  try:
    n = len(os.sched_getaffinity(0))
  except AttributeError:
    n = os.cpu_count() or 1
  return c_oracle.set_num_threads(max(1, n))


def cpu_step(adj, cost, seqs_f32, n, A):
  from oracle import c_oracle

  n_all = 2 * n - 1
  c_oracle.run_sankoff(adj, cost, seqs_f32, n_all, A, n, return_path=True, want_dp=False)
  c_oracle.soft_sankoff(adj, cost, seqs_f32, n_all, A, n, SOFT_T)


def cpu_baseline(sample_sites):
  cores = host_threads()
  adj, cost, states = make_inputs(N_LEAVES, N_STATES, sample_sites, seed=5)
  seqs = states.astype(np.float32)
  cpu_step(adj, cost, seqs[:, :256].copy(), N_LEAVES, N_STATES)  # warm the thread pool
  t0 = time.perf_counter()
  cpu_step(adj, cost, seqs, N_LEAVES, N_STATES)
  dt = time.perf_counter() - t0
  cells = 4 * (N_LEAVES - 1) * sample_sites * N_STATES
  return {"value": cells / dt, "unit": UNIT, "cores": cores, "kind": "port",
          "sample": f"{sample_sites} of {N_SITES} sites of the c5 alignment, one pass (hard fwd+traceback, "
                    f"relaxed fwd+bwd), plain-C OpenMP port of sankoff.py (JAX not installable)",
          "seconds": dt}


def run_reference(args):
  rank = int(os.environ.get("RANK", "0"))
  if rank != 0:
    return
  cores = host_threads()
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
          f"(+ relaxed restatement) on {cores} host threads; the reference itself is JAX, not installable offline")
  cfg = workload_config(args, sample, world=1)
  cfg["parallelism"] = f"openmp{cores}"
  line = {
    "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
    "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
    "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32",
    "data": "synthetic", "config": cfg,
    "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc},
    "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    "gpu_launches": 0,
  }
  emit(line)


def workload_config(args, sites_per_rank, world=None):
  world = args.gpus if world is None else world
  return {"workload": "c5: 1024 taxa x 100k sites, A=20, balanced tree, unit cost; hard fwd + traceback + "
                      "relaxed fwd + bwd (T=1), site-sharded",
          "n_leaves": N_LEAVES, "n_states": N_STATES, "n_sites_total_c5": N_SITES,
          "sites_per_rank": int(sites_per_rank), "parallelism": f"sites{world}",
          "l2": "inputs_exceed_l2 (tables are GBs per pass at 100k sites; at 12.5k sites per rank one pass still "
                "writes > 1 GB against the 126 MB L2; no explicit flush)"}


# ----------------------------------------------------------------------------------------------
# GPU arm: secondary legs (tree cost c2, optimiser loop, landscape loss)
# ----------------------------------------------------------------------------------------------


def measure_tf32_peak(torch):
  """Dense TF32 tensor throughput measured the way MEASURED_PEAKS.json measures bf16: torch.matmul 8192^3 with
  allow_tf32, best of 10 with CUDA events (cuBLAS; a library GEMM as the yardstick, not on any hot path)."""
  n = 8192
  a = torch.randn(n, n, device="cuda")
  b = torch.randn(n, n, device="cuda")
  old = torch.backends.cuda.matmul.allow_tf32
  torch.backends.cuda.matmul.allow_tf32 = True
  try:
    for _ in range(3):
      torch.matmul(a, b)
    best = 1e9
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(10):
      e0.record()
      torch.matmul(a, b)
      e1.record()
      torch.cuda.synchronize()
      best = min(best, e0.elapsed_time(e1))
  finally:
    torch.backends.cuda.matmul.allow_tf32 = old
  del a, b
  return 2.0 * n ** 3 / (best * 1e-3) / 1e12, "torch.matmul f32 8192^3 with allow_tf32, best of 10, CUDA events"


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
  tf32_peak, tf32_how = measure_tf32_peak(torch)
  return {"metric": "tree_cost_grad_steps_per_s", "value": c2["grad_steps_per_s"], "unit": "grad steps/s",
          "ms_per_step": c2["ms_per_step"], "config": "c2: " + c2["shape"], "c2": c2, "script_n200": big,
          "path": "tcgen05 kind::tf32, 3xTF32 split, TMA + TMEM (tree_cost_tc.cu); CUDA-graph replay",
          "roofline": {"bound": "tensor", "achieved": big["useful_tflops"], "peak": tf32_peak,
                       "unit": "TFLOP/s", "frac": big["useful_tflops"] / tf32_peak,
                       "issued_tflops": big["tf32x3_issued_tflops"],
                       "issued_frac": big["tf32x3_issued_tflops"] / tf32_peak, "peak_how": tf32_how,
                       "note": "N=200, K=20000: achieved/frac = USEFUL flops (4 N^2 K per value+grad); issued = the "
                               "3 tf32 MMAs per product of the error-compensated split; c2 itself is "
                               "launch/latency bound (SURVEY H7)"},
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




def event_ms(torch, fn, reps, warm=2):
  for _ in range(warm):
    fn()
  e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  torch.cuda.synchronize()
  e0.record()
  for _ in range(reps):
    fn()
  e1.record()
  torch.cuda.synchronize()
  return e0.elapsed_time(e1) / reps


def general_cost_bench(torch, eng, peak):
  """The GENERAL (non-uniform cost) kernels at the c5 shape — the A x A-in-shared-memory kernels the north star
  describes (sankoff_hard.cu, sankoff_soft_fwd.cu, sankoff_soft_bwd_warp.cu); the headline step above runs the
  O(A) uniform-cost forms because c5 says unit cost.  Random symmetric integer matrix in {1..5}, zero diagonal
  (SURVEY 8d c3's second matrix), T in {1, 0.1}.  Same algorithmic-byte roofline as the headline kernels."""
  from trex_b200 import synth

  n, A, L = eng.n_leaves, eng.A, eng.L
  rng = np.random.default_rng(33)
  ci = torch.from_numpy(synth.random_int_cost(A, 5, rng)).cuda()
  per_site = alg_bytes_per_site(n, A)
  out = {"cost": "random symmetric integer matrix in {1..5}, zero diagonal", "sites": L}

  def rec(name, leg, ms):
    gbs = per_site[leg] * L / (ms * 1e-3) / 1e9
    out[name] = {"ms": ms, "alg_bytes": per_site[leg] * L, "achieved": gbs, "alg_frac": gbs / peak}

  rec("hard_fwd", "hard_fwd", event_ms(torch, lambda: eng.hard_forward(ci, want_argmin=True), 3))
  rec("hard_bwd", "hard_bwd", event_ms(torch, lambda: eng.traceback(want_hist=True), 3))
  for T in (1.0, 0.1):
    rec(f"soft_fwd_T{T}", "soft_fwd", event_ms(torch, lambda: eng.soft_forward(ci, T), 3))
    rec(f"soft_bwd_T{T}", "soft_bwd",
        event_ms(torch, lambda: eng.soft_backward(ci, T, want_dleaf="compact", want_dcost=True), 3))
  return out


def generators_bench(torch):
  """the c5 inputs produced ON THE DEVICE (no host arrays, no H2D): generate_groundtruth
  (ground_truth.py:113-197; fixed-count mutation tree, 1024 leaves x 100k sites, A = 20, 5000 mutations per edge) and
  the NK-landscape variant c5 names (nk_model.py:17-278 with the parameters of evals/benchmark.py:922-929: K = 1,
  q = 20, mutation_rate 0.1, coupled_mutation_prob 0.5, branch_length 1; fitness table 400 x 100k f32 = 160 MB), then
  scored by the hard Sankoff pass to show the states feed the engine directly.  Philox draws (jax.random's streams
  are not reproducible outside JAX); parity on explicit draws is tests/test_gpu_generators.py."""
  from trex_b200 import generators as G
  from trex_b200 import synth
  from trex_b200.sankoff import SankoffEngine

  n, A, L = N_LEAVES, N_STATES, N_SITES
  out = {}
  gt = G.generate_groundtruth(n, A, 5000, L, seed=5)
  out["groundtruth_ms"] = event_ms(torch, lambda: G.generate_groundtruth(n, A, 5000, L, seed=5), 3, warm=1)
  land = G.create_nk_model_landscape(L, 1, seed=5, n_states=A)
  adj = G.create_balanced_binary_tree(n)
  root = torch.randint(0, A, (L,), device="cuda", dtype=torch.uint8)
  out["nk_landscape_ms"] = event_ms(torch, lambda: G.create_nk_model_landscape(L, 1, seed=5, n_states=A), 3, warm=1)
  nk = G.generate_tree_data(land, adj, root, 0.1, seed=5, coupled_mutation_prob=0.5, n_states=A)
  out["nk_tree_ms"] = event_ms(torch, lambda: G.generate_tree_data(land, adj, root, 0.1, seed=5, coupled_mutation_prob=0.5,
                                                                   n_states=A), 3, warm=1)
  cost = torch.from_numpy(synth.unit_cost(A)).cuda()
  for name, tree, a in (("groundtruth", gt, gt.adjacency), ("nk", nk, adj)):
    eng = SankoffEngine(2 * n - 1, n, A, L)
    eng.set_adjacency(a)
    eng.leaf.copy_(tree.states_u8[:n])          # device-to-device: the generated states are the engine's input format
    if eng.leaf.shape[1] > L:
      eng.leaf[:, L:].fill_(255)
    out[name + "_parsimony_score"] = float(eng.hard_forward(cost, want_argmin=False)[0])
    del eng
    torch.cuda.empty_cache()
  out["shape"] = f"{n} leaves x {L} sites, A = {A}"
  return out


def c4_bench(torch, dist, world, rank, reps=5):
  """BASELINE config c4: batched (vmap-over-trees, evals/benchmark.py:92-93) scoring of 512 candidate trees,
  128 taxa x 5000 sites, A in {4, 20}, unit cost; the tree batch is sharded over the ranks
  (trex_b200.dist.score_trees_sharded: no data-path collective, one all-gather of the 512 f64 scores).
  Timed per rank with CUDA events, max over ranks; a tree subset is checked against the plain-C oracle."""
  from trex_b200 import synth
  from trex_b200.dist import score_trees_sharded, shard_range
  from trex_b200.sankoff import SankoffEngine

  n, L, B = 128, 5000, 512
  n_all = 2 * n - 1
  out = {"workload": "c4: 512 random-join trees x 128 taxa x 5000 sites, hard Sankoff scores, unit cost",
         "parallelism": f"trees{world}"}
  rng = np.random.default_rng(4)
  adj = np.stack([synth.parents_to_adjacency(synth.random_join_parents(n, rng)) for _ in range(B)]).astype(np.float32)
  adj_d = torch.from_numpy(adj).cuda()
  lo, hi = shard_range(B, world, rank)
  for A in (4, 20):
    states = np.ascontiguousarray(synth.evolve(synth.balanced_parents(n), L, A, 0.1, rng)[:n])
    leaf_d = torch.from_numpy(states).cuda()
    cost = torch.from_numpy(synth.unit_cost(A)).cuda()
    eng = SankoffEngine(n_all, n, A, L, batch=hi - lo)
    eng.set_leaf_states_u8(leaf_d)
    mine = torch.zeros(-(-B // world), dtype=torch.float64, device="cuda")
    parts = [torch.empty_like(mine) for _ in range(world)]

    def step():
      eng.set_adjacency(adj_d[lo:hi])
      mine[: hi - lo] = eng.hard_forward(cost, want_argmin=False)
      if world > 1:
        dist.all_gather(parts, mine)

    ms = event_ms(torch, step, reps)
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
      dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t[0])
    scores = score_trees_sharded(adj_d, cost, leaf_d, A)  # the public call (same kernels + the all-gather)
    ok = None
    if rank == 0:
      from oracle import c_oracle

      idx = [0, 1, B // 2, B - 1]
      want = [float(c_oracle.run_sankoff(adj[i], synth.unit_cost(A), states.astype(np.float32), n_all, A, n,
                                         want_dp=False)[3].astype(np.float64).sum()) for i in idx]
      ok = bool([float(scores[i]) for i in idx] == want)
    out[f"A{A}"] = {"ms": ms, "trees_per_s": B / ms * 1e3, "cell_updates_per_s": B * (n - 1) * L * A / ms * 1e3,
                    "trees_per_rank": hi - lo, "exact_vs_oracle_on_4_trees": ok, "score0": float(scores[0])}
    del eng
  return out


def sharded_tree_cost_bench(torch, dist, world, rank, reps=20):
  """tree cost value+grad with the sites (the contraction axis) sharded over the ranks and
  one packed all-reduce of [cost, dAdj, dC] (trex_b200.dist.tree_cost_value_and_grad_sharded), at c2 and at
  the script shape N=200, K=20000.  At these sizes a step is microseconds: the all-reduce latency dominates
  and multi-GPU cannot pay for itself — reported, not optimised for."""
  from oracle import tree_oracle as to
  from trex_b200.dist import shard_range, tree_cost_value_and_grad_sharded
  from trex_b200.tree import tree_cost_value_and_grad

  out = {}
  for name, (N, L, Q, seed) in {"c2": (127, 1000, 4, 2), "script_n200": (200, 1000, 20, 242)}.items():
    rng = np.random.default_rng(seed)
    S = to.softmax(rng.normal(size=(N, L, Q))).astype(np.float32)
    adj = (rng.random((N, N)) < 0.1).astype(np.float32)
    lo, hi = shard_range(L, world, rank)
    S_loc = torch.from_numpy(np.ascontiguousarray(S[:, lo:hi])).cuda()
    adj_d = torch.from_numpy(adj).cuda()
    ms = event_ms(torch, lambda: tree_cost_value_and_grad_sharded(S_loc, adj_d, None), reps, warm=5)
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    cost, _, dA, _ = tree_cost_value_and_grad_sharded(S_loc, adj_d, None)
    rel = None
    if rank == 0:
      c1, _, dA1, _ = tree_cost_value_and_grad(torch.from_numpy(S).cuda(), adj_d, None)
      rel = {"cost": abs(float(cost) - float(c1)) / abs(float(c1)),
             "dAdj": float((dA - dA1).abs().max() / dA1.abs().max())}
    out[name] = {"shape": f"{N} nodes x {L} sites x {Q} states", "ms_per_step": float(t[0]),
                 "grad_steps_per_s": 1e3 / float(t[0]), "sites_per_rank": hi - lo,
                 "rel_diff_vs_single_gpu": rel}
  return out


# ----------------------------------------------------------------------------------------------
# GPU arm: the Sankoff step
# ----------------------------------------------------------------------------------------------

LEGS = ["plan", "hard_fwd", "hard_bwd", "soft_fwd", "soft_bwd", "pack"]


class SankoffLeg:
  """One rank's share of the step on device-resident inputs: eager (per-kernel CUDA events) and replayed from a
  CUDA graph (the timed form)."""

  def __init__(self, torch, args, L, adj_h, cost_h, states_h):
    from trex_b200.dist import PackedAllReduce
    from trex_b200.sankoff import SankoffEngine

    self.torch, self.args, self.L = torch, args, L
    n, A = N_LEAVES, N_STATES
    self.adj_p = torch.from_numpy(adj_h).pin_memory()
    self.cost_p = torch.from_numpy(cost_h).pin_memory()
    self.states_p = torch.from_numpy(states_h).pin_memory()
    self.adj_d, self.cost_d = self.adj_p.cuda(), self.cost_p.cuda()
    self.eng = SankoffEngine(2 * n - 1, n, A, L)
    self.eng.set_leaf_states_u8(self.states_p.cuda())
    # two packed buffers alternate: the all-reduce of step k runs on a communication stream under the kernels of
    # step k+1 (steps are independent jobs); both are joined before the closing timing event
    self.reducers = [PackedAllReduce([(1,), (1,), (1, A, A), (1, A, A)], device="cuda") for _ in range(2)]
    self.reducer = self.reducers[0]
    self.comm = torch.cuda.Stream()
    self.reduced = [None, None]
    self.k = 0
    # Short alignments (a GPU's share under strong scaling) leave SMs idle in the top part of every walk and in
    # the wave tails: there the hard pass and the relaxed pass, which are independent, run on two streams.
    self.two_streams = L <= args.overlap_below
    self.side = (torch.cuda.Stream(), torch.cuda.Stream()) if self.two_streams else None
    self.stream = torch.cuda.Stream()
    self.graph = None
    self.graphs = [None, None]

  def step_eager(self, ev=None):
    """single stream, optional event marks between the kernels"""
    e, c = self.eng, self.cost_d
    mark = (lambda i: ev[i].record()) if ev is not None else (lambda i: None)
    mark(0)
    e.set_adjacency(self.adj_d)
    mark(1)
    e.hard_forward(c, want_argmin=True)
    mark(2)
    e.traceback(want_hist=True)
    mark(3)
    e.soft_forward(c, SOFT_T)
    mark(4)
    e.soft_backward(c, SOFT_T, want_dleaf="compact", want_dcost=True)
    mark(5)
    self.reducer.pack([e.total, e.soft_total, e.dcost, e.edge_hist])
    mark(6)

  def step_core(self, j=0):
    self.reducer = self.reducers[j]
    if not self.two_streams:
      return self.step_eager()
    torch, e, c = self.torch, self.eng, self.cost_d
    main = torch.cuda.current_stream()
    sa, sb = self.side
    e.set_adjacency(self.adj_d)
    sa.wait_stream(main)
    sb.wait_stream(main)
    with torch.cuda.stream(sa):
      e.hard_forward(c, want_argmin=True)
      e.traceback(want_hist=True)
    with torch.cuda.stream(sb):
      e.soft_forward(c, SOFT_T)
      e.soft_backward(c, SOFT_T, want_dleaf="compact", want_dcost=True)
    main.wait_stream(sa)
    main.wait_stream(sb)
    self.reducer.pack([e.total, e.soft_total, e.dcost, e.edge_hist])

  def capture(self):
    torch = self.torch
    with torch.cuda.stream(self.stream):
      for j in range(2):
        self.step_core(j)
        self.stream.synchronize()
        if self.args.graph:
          g = torch.cuda.CUDAGraph()
          with torch.cuda.graph(g, stream=self.stream):
            self.step_core(j)
          self.graphs[j] = g
      self.graph = self.graphs[0]
    torch.cuda.synchronize()

  def step(self):
    """the timed step: graph replay (or eager launches) on self.stream, then the packed all-reduce of ITS results
    on the communication stream (it overlaps the next step's kernels, which do not depend on it)"""
    torch = self.torch
    j = self.k & 1
    self.k += 1
    if self.reduced[j] is not None:
      self.stream.wait_event(self.reduced[j])  # the pack of this step overwrites the buffer reduced two steps ago
    if self.graphs[j] is not None:
      self.graphs[j].replay()
    else:
      self.step_core(j)
    packed = torch.cuda.Event()
    packed.record(self.stream)
    with torch.cuda.stream(self.comm):
      self.comm.wait_event(packed)
      self.reducers[j].reduce()
      done = torch.cuda.Event()
      done.record(self.comm)
    self.reduced[j] = done
    self.last = j

  def join(self):
    """self.stream waits for every outstanding all-reduce (called before the closing timing event)"""
    for ev in self.reduced:
      if ev is not None:
        self.stream.wait_event(ev)

  def timed(self, steps, warmup, barrier, sampler=None):
    torch = self.torch
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(self.stream):
      for _ in range(warmup):
        self.step()
      self.join()
    barrier()
    t0 = time.perf_counter()
    with torch.cuda.stream(self.stream):
      e0.record(self.stream)
      for _ in range(steps):
        self.step()
      self.join()
      e1.record(self.stream)
    barrier()
    if sampler is not None:
      sampler.mark(t0, time.perf_counter())
    return e0.elapsed_time(e1)

  def kernel_ms(self, steps):
    torch = self.torch
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(7)] for _ in range(steps)]
    self.step_eager()
    for k in range(steps):
      self.step_eager(evs[k])
    torch.cuda.synchronize()
    return {nm: float(np.mean([evs[k][i].elapsed_time(evs[k][i + 1]) for k in range(steps)]))
            for i, nm in enumerate(LEGS)}


def run_gpu(args):
  import torch
  import torch.distributed as dist

  from trex_b200 import _lib
  from trex_b200.dist import shard_range
  from trex_b200.pipeline import SankoffHostPipeline

  rank = int(os.environ.get("RANK", "0"))
  world = int(os.environ.get("WORLD_SIZE", "1"))
  local = int(os.environ.get("LOCAL_RANK", "0"))
  if world != args.gpus and world == 1 and args.gpus > 1:
    raise SystemExit("--gpus N>1 must be launched with torch.distributed.run (one rank per GPU)")
  torch.cuda.set_device(local)
  if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
  lib = _lib.load()
  n, A = N_LEAVES, N_STATES

  def barrier():
    if world > 1:
      dist.barrier()
    torch.cuda.synchronize()

  def max_over_ranks(vals):
    t = torch.tensor(vals, dtype=torch.float64, device="cuda")
    if world > 1:
      dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(x) for x in t]

  # ---- inputs: strong = this rank's contiguous block of the ONE c5 alignment; weak = its own 100k sites ----
  if args.sites:
    L, lo = args.sites, 0
    adj_h, cost_h, states_h = make_inputs(n, A, L, seed=5 + rank)
  elif args.scaling == "strong":
    adj_h, cost_h, full = make_inputs(n, A, N_SITES, seed=5)
    lo, hi = shard_range(N_SITES, world, rank, align=64)
    L = hi - lo
    states_h = np.ascontiguousarray(full[:, lo:hi])
  else:
    L, lo = N_SITES, 0
    adj_h, cost_h, states_h = make_inputs(n, A, L, seed=5 + rank)
  sites_total = int(sum(max_over_ranks([0.0])) or 0)  # placeholder (keeps NCCL warm before the timed region)
  t = torch.tensor([float(L)], dtype=torch.float64, device="cuda")
  if world > 1:
    dist.all_reduce(t)
  sites_total = int(t[0])

  leg = SankoffLeg(torch, args, L, adj_h, cost_h, states_h)
  leg.capture()
  sampler = ClockSampler(local)
  if rank == 0:
    sampler.start()
  ms_total = leg.timed(args.steps, args.warmup, barrier, sampler)
  clocks = sampler.stop() if rank == 0 else None
  packed = leg.reducers[leg.last].buf.cpu()
  launches0 = int(lib.trex_kernel_launches())
  leg.step_eager()
  launches_per_step = int(lib.trex_kernel_launches()) - launches0
  kern_ms = leg.kernel_ms(min(args.steps, 5))

  # ---- end to end: pinned host inputs -> host results through the public host pipeline, every step ----
  pipe = SankoffHostPipeline(2 * n - 1, n, A, L, temperature=SOFT_T, leaf_grad="compact", use_graph=bool(args.graph),
                             engine=leg.eng)
  jobs = [(leg.adj_p, leg.cost_p, leg.states_p)]
  pipe.run(jobs * 2)
  barrier()
  f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  f0.record()
  outs = pipe.run(jobs * args.steps)
  f1.record()
  barrier()
  ms_e2e = f0.elapsed_time(f1)
  anc_h, packed_h = outs[0]
  e2e_check = {"hard_total": float(packed_h[0]), "soft_total": float(packed_h[1]),
               "anc_checksum": int(anc_h.to(torch.int64).sum())}

  ms_total, ms_e2e = max_over_ranks([ms_total, ms_e2e])
  cells_step = 4 * (n - 1) * sites_total * A
  value = cells_step * args.steps / (ms_total * 1e-3)
  e2e_value = cells_step * args.steps / (ms_e2e * 1e-3)

  # ---- weak-scaling record next to the strong headline (N > 1) ----
  weak = None
  if world > 1 and args.scaling == "strong" and not args.sites and not args.no_extras:
    w_adj, w_cost, w_states = make_inputs(n, A, N_SITES, seed=5 + rank)
    wleg = SankoffLeg(torch, args, N_SITES, w_adj, w_cost, w_states)
    wleg.capture()
    (w_ms,) = max_over_ranks([wleg.timed(args.steps, args.warmup, barrier)])
    weak = {"scaling": "weak", "sites_per_rank": N_SITES, "ms_per_step": w_ms / args.steps,
            "value": 4 * (n - 1) * N_SITES * world * A * args.steps / (w_ms * 1e-3), "unit": UNIT}
    del wleg
    torch.cuda.empty_cache()

  extras = {}
  if not args.no_extras:
    extras["c4"] = c4_bench(torch, dist, world, rank)
    if world > 1:
      extras["tree_cost_site_sharded"] = sharded_tree_cost_bench(torch, dist, world, rank)

  if rank == 0:
    pk, peak_src = measured_peaks()
    peak = float(pk["hbm_gbs"])
    per_site = alg_bytes_per_site(n, A, words=True)  # c5 says unit cost: the engine writes pointer words
    traffic_tab = ncu_traffic() if L == N_SITES else {}
    roof_all = {}
    for nm in ["hard_fwd", "hard_bwd", "soft_fwd", "soft_bwd"]:
      gbs = per_site[nm] * L / (kern_ms[nm] * 1e-3) / 1e9
      tr = (traffic_tab.get(nm) or {}).get("bytes")
      roof_all[nm] = {"ms": kern_ms[nm], "alg_bytes": per_site[nm] * L, "achieved": gbs, "alg_frac": gbs / peak,
                      "dram_bytes_ncu": tr,
                      "dram_frac": (tr / (kern_ms[nm] * 1e-3) / 1e9 / peak) if tr else None,
                      "ncu_source": (traffic_tab.get(nm) or {}).get("source")}
    dom = max(["hard_fwd", "soft_fwd", "soft_bwd"], key=lambda k: kern_ms[k])
    step_ms = ms_total / args.steps
    step_alg = sum(per_site.values()) * L
    step_tr = sum((roof_all[k]["dram_bytes_ncu"] or 0) for k in roof_all) if all(
      roof_all[k]["dram_bytes_ncu"] for k in roof_all) else None
    line = {
      "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
      "warmup": args.warmup, "ms_per_step": step_ms, "higher_is_better": True,
      "scaling": "weak" if args.scaling == "weak" else "strong", "vs_baseline": None, "dtype": "f32",
      "data": "synthetic", "config": dict(workload_config(args, L, world), sites_total=sites_total,
                                          step_streams=2 if leg.two_streams else 1,
                                          cuda_graph=bool(args.graph)),
      "roofline": {"bound": "hbm", "kernel": dom, "achieved": roof_all[dom]["achieved"], "peak": peak,
                   "peak_source": peak_src, "unit": "GB/s", "frac": roof_all[dom]["alg_frac"],
                   "traffic": roof_all[dom]["dram_bytes_ncu"], "dram_frac": roof_all[dom]["dram_frac"],
                   "note": "achieved / frac = SURVEY 8(d) ALGORITHMIC bytes per launch / CUDA-event launch time on the "
                           "launching stream (child rows served from L2 count although they never reach DRAM, so "
                           "forward fractions can exceed 1); dram_frac = ncu dram bytes of the same launch / the "
                           "same time = the HBM fraction proper"},
      "roofline_all": roof_all,
      "roofline_step": {"bound": "hbm", "alg_bytes": step_alg, "achieved": step_alg / (step_ms * 1e-3) / 1e9,
                        "peak": peak, "unit": "GB/s", "alg_frac": step_alg / (step_ms * 1e-3) / 1e9 / peak,
                        "dram_bytes_ncu": step_tr,
                        "dram_frac": (step_tr / (step_ms * 1e-3) / 1e9 / peak) if step_tr else None,
                        "note": "all four passes of the step over the step time"},
      "kernel_ms": kern_ms,
      "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": ms_e2e / args.steps,
              "h2d_bytes_per_step": pipe.h2d_bytes_per_job, "d2h_bytes_per_step": pipe.d2h_bytes_per_job,
              "api": "trex_b200.pipeline.SankoffHostPipeline.run: run_sankoff(..., return_path=True)-shaped job + "
                     "relaxed fwd/bwd; pinned host adjacency / cost / leaf states in, host ancestral states (u8) + "
                     "[totals, dC, edge histogram] out, every step",
              "overlap": "H2D of step k+1 and D2H of step k on copy streams under the kernels (double-buffered)",
              "results": e2e_check},
      "gpu_launches": launches_per_step * args.steps,
      "gpu_launches_note": f"{launches_per_step} libtrex_b200 kernels per step (counted by the library on an eager "
                           f"pass) x {args.steps} steps; the timed region replays them from a CUDA graph"
                           if args.graph else f"{launches_per_step} per step",
      "clocks": clocks,
      "results": {"hard_total": float(packed[0]), "soft_total": float(packed[1])},
    }
    if weak:
      line["weak_scaling"] = weak
    line.update(extras)
    if world == 1 and not args.no_extras:
      def secondary(name, fn, *a):
        # a failing secondary leg must not take the headline line with it: record the error and go on
        try:
          line[name] = fn(*a)
        except Exception as e:  # noqa: BLE001
          line[name] = {"error": f"{type(e).__name__}: {e}"[:300]}
          print(f"bench: secondary leg {name} failed: {e!r}", file=sys.stderr, flush=True)

      if L == N_SITES:
        secondary("general_cost", general_cost_bench, torch, leg.eng, peak)
      secondary("generators", generators_bench, torch)
      secondary("tree_cost", tree_cost_bench, torch)
      secondary("fit_loop", fit_loop_bench, torch)
      secondary("landscape_loss", landscape_bench, torch)
      secondary("cpu_baseline", cpu_baseline, args.cpu_sites)
    emit(line)
  if world > 1:
    dist.barrier()
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
  ap.add_argument("--scaling", default="strong", choices=["weak", "strong"],
                  help="strong (default): the ONE 100k-site c5 alignment split over the ranks; weak: 100k sites per rank")
  ap.add_argument("--sites", type=int, default=0, help="override sites per rank (testing)")
  ap.add_argument("--cpu-sites", type=int, default=4096, help="sites of the bounded cpu_baseline sample")
  ap.add_argument("--ref-sites", type=int, default=2048, help="sites per step of --impl reference")
  ap.add_argument("--overlap-below", type=int, default=50000,
                  help="sites per rank at or below which the hard and relaxed passes run on two streams")
  ap.add_argument("--graph", type=int, default=1, help="1: replay the step from a CUDA graph (default); 0: eager launches")
  ap.add_argument("--no-extras", action="store_true", help="skip the secondary legs (c4, tree cost, cpu_baseline, ...)")
  args = ap.parse_args()
  quiet_stdout()
  args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
  if args.impl == "reference":
    run_reference(args)
  else:
    run_gpu(args)


if __name__ == "__main__":
  main()
