"""Target of the ncu captures: the c5 step (unit cost; GENERAL=1: integer cost -> the general kernels), one
warm-up step + one captured step.  TREE_COST=1 instead runs the tcgen05 tree cost at c2 and at N=200, K=20000."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from trex_b200 import synth
if os.environ.get("TREE_COST"):
  from oracle import tree_oracle as to
  from trex_b200.tree import tree_cost_value_and_grad
  for (N, L, Q, seed) in ((127, 1000, 4, 2), (200, 1000, 20, 242)):
    rng = np.random.default_rng(seed)
    S = torch.from_numpy(to.softmax(rng.normal(size=(N, L, Q))).astype(np.float32)).cuda()
    A = torch.from_numpy((rng.random((N, N)) < 0.1).astype(np.float32)).cuda()
    for _ in range(2):
      tree_cost_value_and_grad(S, A, None)
    torch.cuda.synchronize()
  sys.exit(0)
from trex_b200.sankoff import SankoffEngine
n, A, L = 1024, 20, int(os.environ.get("L", 100000))
rng = np.random.default_rng(5)
par = synth.balanced_parents(n); adj = torch.from_numpy(synth.parents_to_adjacency(par)).cuda()
states = synth.evolve(par, L, A, 0.1, rng)[:n]
eng = SankoffEngine(2*n-1, n, A, L)
eng.set_leaf_states_u8(torch.from_numpy(np.ascontiguousarray(states)))
c = torch.from_numpy(synth.random_int_cost(A, 5, rng) if os.environ.get("GENERAL") else synth.unit_cost(A)).cuda()
for _ in range(2):
  eng.set_adjacency(adj)
  eng.hard_forward(c, want_argmin=True); eng.traceback(want_hist=True)
  eng.soft_forward(c, 1.0); eng.soft_backward(c, 1.0, want_dleaf="compact", want_dcost=True)
  torch.cuda.synchronize()
