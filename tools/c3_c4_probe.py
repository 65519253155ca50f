"""BASELINE configs c3 and c4 on one GPU (they are parity-test cases; this records their times).
c4: 512 candidate trees (random joins) x 128 taxa x 5 000 sites, hard scores, A in {4, 20}; all 512 trees as one
    batched launch, and 64 trees = one GPU's share of the 8-way tree-batch sharding.
c3: 256 taxa x 10 000 sites, A = 20: relaxed forward + backward (+ dC, + d leaf) and hard forward + traceback."""
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

out = {}
rng = np.random.default_rng(4)
n, L = 128, 5000
for A in (4, 20):
  par0 = synth.balanced_parents(n)
  states = synth.evolve(par0, L, A, 0.1, rng)[:n]
  cost = torch.from_numpy(synth.unit_cost(A)).cuda()
  for B in (512, 64):
    adj = np.stack([synth.parents_to_adjacency(synth.random_join_parents(n, rng)) for _ in range(B)]).astype(np.float32)
    adj_d = torch.from_numpy(adj).cuda()
    eng = SankoffEngine(2 * n - 1, n, A, L, batch=B)
    eng.set_leaf_states_u8(torch.from_numpy(np.ascontiguousarray(states)))
    def score():
      eng.set_adjacency(adj_d)
      eng.hard_forward(cost, want_argmin=False)
    ms = t(score)
    cells = B * (n - 1) * L * A
    out[f"c4_A{A}_trees{B}"] = {"ms": ms, "trees_per_s": B / ms * 1e3, "cell_updates_per_s": cells / ms * 1e3,
                                "score0": float(eng.total[0])}
    del eng
    torch.cuda.empty_cache()
n, L, A = 256, 10000, 20
rng = np.random.default_rng(3)
par = synth.balanced_parents(n)
states = synth.evolve(par, L, A, 0.1, rng)[:n]
eng = SankoffEngine(2 * n - 1, n, A, L)
eng.set_adjacency(torch.from_numpy(synth.parents_to_adjacency(par))).set_leaf_states_u8(torch.from_numpy(np.ascontiguousarray(states)))
cost = torch.from_numpy(synth.unit_cost(A)).cuda()
def c3():
  eng.hard_forward(cost, want_argmin=True); eng.traceback(want_hist=True)
  eng.soft_forward(cost, 1.0); eng.soft_backward(cost, 1.0, want_dleaf=True, want_dcost=True)
ms = t(c3, reps=20)
out["c3"] = {"ms": ms, "cell_updates_per_s": 4 * (n - 1) * L * A / ms * 1e3}
print(json.dumps(out, indent=1))
