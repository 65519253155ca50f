"""Per-kernel times of the c5 step (unit cost) for the site counts in LS under the current environment
(TREX_BWD_WIDE, TREX_SEG_TARGET, TREX_SEGMENTS ...): one line per L."""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from trex_b200 import synth
from trex_b200.sankoff import SankoffEngine

def t(f, reps=10):
  for _ in range(3): f()
  e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  torch.cuda.synchronize(); e0.record()
  for _ in range(reps): f()
  e1.record(); torch.cuda.synchronize()
  return round(e0.elapsed_time(e1) / reps, 4)

n, A = 1024, 20
tag = {k: os.environ[k] for k in ("TREX_BWD_WIDE", "TREX_SEG_TARGET", "TREX_SEGMENTS") if k in os.environ}
for L in [int(x) for x in os.environ.get("LS", "12500").split(",")]:
  rng = np.random.default_rng(5)
  par = synth.balanced_parents(n); adj = torch.from_numpy(synth.parents_to_adjacency(par)).cuda()
  eng = SankoffEngine(2*n-1, n, A, L)
  eng.set_leaf_states_u8(torch.from_numpy(np.ascontiguousarray(synth.evolve(par, L, A, 0.1, rng)[:n])))
  cu = torch.from_numpy(synth.unit_cost(A)).cuda()
  eng.set_adjacency(adj)
  r = {"hard_fwd": t(lambda: eng.hard_forward(cu, want_argmin=True)), "traceback": t(lambda: eng.traceback(want_hist=True)),
       "soft_fwd": t(lambda: eng.soft_forward(cu, 1.0)),
       "soft_bwd": t(lambda: eng.soft_backward(cu, 1.0, want_dleaf="compact", want_dcost=True))}
  r["sum"] = round(sum(r.values()), 4)
  print(L, json.dumps(tag), json.dumps(r), flush=True)
  del eng; torch.cuda.empty_cache()
