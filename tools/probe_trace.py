"""Traceback timing at c5: pointer words (unit cost) and byte table (integer cost), TREX_TRACE_V = 1 / 2 / 4."""
import os, sys
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
def t(f, reps=10):
  for _ in range(3): f()
  e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  torch.cuda.synchronize(); e0.record()
  for _ in range(reps): f()
  e1.record(); torch.cuda.synchronize()
  return e0.elapsed_time(e1) / reps
for name, c, compact in (("words", cu, True), ("bytes(unit)", cu, False), ("bytes(int)", ci, True)):
  eng.hard_forward(c, want_argmin=True, compact_arg=compact)
  ref = None
  for v in ("1", "2", "4"):
    os.environ["TREX_TRACE_V"] = v
    th, tn = t(lambda: eng.traceback(want_hist=True)), t(lambda: eng.traceback(want_hist=False))
    anc = eng.anc.clone()
    same = True if ref is None else bool(torch.equal(ref, anc))
    ref = anc
    print(f"{name:12s} V={v}  +hist {th:.3f} ms  -hist {tn:.3f} ms  same={same}")
