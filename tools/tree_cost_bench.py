#!/usr/bin/env python
"""Tree-cost value+grad timing (c2 and the benchmark_tree_cost.py shapes), through the C ABI,
replayed from a CUDA graph so host dispatch is out of the timed region.  Checks against the
float64 oracle first.  Usage: python tools/tree_cost_bench.py [--iters N]"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import tree_oracle as to  # noqa: E402
from trex_b200 import _lib  # noqa: E402
from trex_b200.tree import tree_cost_value_and_grad  # noqa: E402


STEPS_PER_GRAPH = int(os.environ.get("STEPS_PER_GRAPH", "1"))


def case(N, L, Q, seed, iters):
  rng = np.random.default_rng(seed)
  S = to.softmax(rng.normal(size=(N, L, Q))).astype(np.float32)
  adj = (rng.random((N, N)) < 0.1).astype(np.float32)  # benchmark_tree_cost.py:57
  St, At = torch.from_numpy(S).cuda(), torch.from_numpy(adj).cuda()
  ws = torch.empty(_lib.load().trex_tree_cost_workspace_bytes(N, L * Q, Q), dtype=torch.uint8, device="cuda")
  out = {"cost": torch.empty((), device="cuda"), "dS": torch.empty_like(St), "dA": torch.empty_like(At)}
  tree_cost_value_and_grad(St, At, None, workspace=ws, out=out)
  want, dS64, dA64, _ = to.soft_cost_value_and_grad(S, adj, None)
  err = {
    "cost": abs(float(out["cost"]) - want) / abs(want),
    "dS": float(np.abs(out["dS"].cpu().numpy() - dS64).max() / np.abs(dS64).max()),
    "dA": float(np.abs(out["dA"].cpu().numpy() - dA64).max() / np.abs(dA64).max()),
  }
  stream = torch.cuda.Stream()
  with torch.cuda.stream(stream):
    for _ in range(3):
      tree_cost_value_and_grad(St, At, None, workspace=ws, out=out)
    g = torch.cuda.CUDAGraph()
    stream.synchronize()
    with torch.cuda.graph(g, stream=stream):
      for _ in range(STEPS_PER_GRAPH):
        tree_cost_value_and_grad(St, At, None, workspace=ws, out=out)
    g.replay()
    stream.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(iters):
      g.replay()
    e1.record(stream)
    stream.synchronize()
  ms = e0.elapsed_time(e1) / iters / STEPS_PER_GRAPH
  K = L * Q
  flop = 4.0 * N * N * K
  return {"N": N, "L": L, "Q": Q, "us_per_step": ms * 1e3, "grad_steps_per_s": 1e3 / ms,
          "useful_tflops": flop / (ms * 1e-3) / 1e12, "tf32x3_issued_tflops": 3 * flop / (ms * 1e-3) / 1e12,
          "rel_err": err}


if __name__ == "__main__":
  ap = argparse.ArgumentParser()
  ap.add_argument("--iters", type=int, default=200)
  ap.add_argument("--only", type=int, default=0, help="run just the case with this N (127 = c2)")
  a = ap.parse_args()
  cases = [(127, 1000, 4, 2)] + [(N, 1000, 20, 42 + N) for N in (10, 50, 100, 200)]
  rows = [case(N, L, Q, seed, a.iters) for (N, L, Q, seed) in cases if not a.only or a.only == N]
  for r in rows:
    print(json.dumps(r))
