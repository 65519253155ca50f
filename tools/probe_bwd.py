"""soft_bwd variants at the c5 shape (and one GPU's share under strong scaling): dense / compact / no leaf
gradient, with and without dC, unit and integer cost.  Prints one JSON object."""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from trex_b200 import synth
from trex_b200.sankoff import SankoffEngine

def t(f, reps=5):
  for _ in range(2): f()
  e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  torch.cuda.synchronize(); e0.record()
  for _ in range(reps): f()
  e1.record(); torch.cuda.synchronize()
  return e0.elapsed_time(e1) / reps

n, A = 1024, 20
out = {}
for L in [int(x) for x in os.environ.get("LS", "100000,12500").split(",")]:
  rng = np.random.default_rng(5)
  par = synth.balanced_parents(n); adj = synth.parents_to_adjacency(par)
  states = synth.evolve(par, L, A, 0.1, rng)[:n]
  eng = SankoffEngine(2*n-1, n, A, L)
  eng.set_adjacency(torch.from_numpy(adj)).set_leaf_states_u8(torch.from_numpy(np.ascontiguousarray(states)))
  cu = torch.from_numpy(synth.unit_cost(A)).cuda()
  ci = torch.from_numpy(synth.random_int_cost(A, 5, rng)).cuda()
  r = {}
  for name, c in (("unit", cu), ("int", ci)):
    r[f"{name}_hard_fwd"] = t(lambda: eng.hard_forward(c, want_argmin=True))
    r[f"{name}_traceback"] = t(lambda: eng.traceback(want_hist=True))
    r[f"{name}_soft_fwd"] = t(lambda: eng.soft_forward(c, 1.0))
    for dl in (True, "compact", False):
      for dc in (True, False):
        r[f"{name}_soft_bwd_dleaf={dl}_dC={dc}"] = t(lambda: eng.soft_backward(c, 1.0, want_dleaf=dl, want_dcost=dc))
  r["plan"] = t(lambda: eng.set_adjacency(torch.from_numpy(adj).cuda()), reps=20)
  out[str(L)] = r
  del eng
  torch.cuda.empty_cache()
print(json.dumps(out, indent=1))
