import sys, os, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from trex_b200 import synth
from trex_b200.sankoff import SankoffEngine
n, A, L = 1024, 20, int(os.environ.get("L", 100000))
rng = np.random.default_rng(5)
par = synth.balanced_parents(n); adj = synth.parents_to_adjacency(par)
states = synth.evolve(par, L, A, 0.1, rng)[:n]
eng = SankoffEngine(2*n-1, n, A, L)
eng.set_adjacency(torch.from_numpy(adj)).set_leaf_states_u8(torch.from_numpy(np.ascontiguousarray(states)))
cu = torch.from_numpy(synth.unit_cost(A)).cuda()
ci = torch.from_numpy(synth.random_int_cost(A, 5, rng)).cuda()
def t(f, reps=5):
  for _ in range(2): f()
  e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  torch.cuda.synchronize(); e0.record()
  for _ in range(reps): f()
  e1.record(); torch.cuda.synchronize()
  return e0.elapsed_time(e1) / reps
print("hard uniform  +arg", t(lambda: eng.hard_forward(cu, want_argmin=True)))
print("hard uniform  -arg", t(lambda: eng.hard_forward(cu, want_argmin=False)))
print("hard general(unit,mode1) +arg", t(lambda: eng.hard_forward(cu, want_argmin=True, mode=1)))
print("hard packed (int cost) +arg", t(lambda: eng.hard_forward(ci, want_argmin=True)))
print("hard packed (int cost) -arg", t(lambda: eng.hard_forward(ci, want_argmin=False)))
print("soft fwd uniform", t(lambda: eng.soft_forward(cu, 1.0)))
print("soft fwd general (int cost)", t(lambda: eng.soft_forward(ci, 1.0)))
eng.soft_forward(cu, 1.0)
print("soft bwd (unit) +dC", t(lambda: eng.soft_backward(cu, 1.0)))
print("soft bwd (unit) -dC", t(lambda: eng.soft_backward(cu, 1.0, want_dcost=False)))
print("soft bwd (unit) -dC -dleaf", t(lambda: eng.soft_backward(cu, 1.0, want_dcost=False, want_dleaf=False)))
eng.hard_forward(cu, want_argmin=True)
print("traceback +hist", t(lambda: eng.traceback(want_hist=True)))
print("traceback -hist", t(lambda: eng.traceback(want_hist=False)))
