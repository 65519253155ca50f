"""Timing probes of the ancestor-fitting kernel: TREX_FIT_DEBUG bit mask switches phases off
(results are then meaningless; only the time per iteration is of interest)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from trex_b200 import optimize as O, synth

def one(n, L, Q, iters, dbg):
  os.environ["TREX_FIT_DEBUG"] = str(dbg)
  rng = np.random.default_rng(0)
  parent = synth.balanced_parents(n); N = parent.shape[0]
  leaf = synth.evolve(parent, L, Q, 0.1, rng)[:n]
  adj = synth.parents_to_adjacency(parent).astype(np.float32)
  init = rng.normal(size=(N - n, L, Q)).astype(np.float32)
  fit = O.AncestorFit(leaf, N, n, Q, adj, init, O.create_optimizer("adam", 1e-3))
  losses = torch.empty(iters, device="cuda")
  fit.run(iters, losses); torch.cuda.synchronize()
  e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  e0.record(); fit.run(iters, losses); e1.record(); torch.cuda.synchronize()
  return e0.elapsed_time(e1) * 1e3 / iters

if __name__ == "__main__":
 for shape in [(64, 1000, 4, 5000), (100, 1000, 20, 500)]:
  for dbg in [0, 16, 1, 15, 14, 31]:
    print(shape, "debug", dbg, "us/iter %.2f" % one(*shape, dbg), flush=True)
