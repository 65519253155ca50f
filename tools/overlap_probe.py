import sys, os
sys.path.insert(0, "/root/repo")
import numpy as np, torch
import bench
from trex_b200.sankoff import SankoffEngine
n, A, L = 1024, 20, int(sys.argv[1]) if len(sys.argv) > 1 else 100000
adj_h, cost_h, states_h = bench.make_inputs(n, A, L, seed=5)
adj_d, cost_d, states_d = [torch.from_numpy(x).cuda() for x in (adj_h, cost_h, states_h)]
eng = SankoffEngine(2 * n - 1, n, A, L)
eng.set_leaf_states_u8(states_d)
sa, sb = torch.cuda.Stream(), torch.cuda.Stream()
def serial():
  eng.set_adjacency(adj_d)
  eng.hard_forward(cost_d, want_argmin=True); eng.traceback(want_hist=True)
  eng.soft_forward(cost_d, 1.0); eng.soft_backward(cost_d, 1.0, want_dleaf=True, want_dcost=True)
def overlapped():
  main = torch.cuda.current_stream()
  eng.set_adjacency(adj_d)
  sa.wait_stream(main); sb.wait_stream(main)
  with torch.cuda.stream(sa):
    eng.hard_forward(cost_d, want_argmin=True); eng.traceback(want_hist=True)
  with torch.cuda.stream(sb):
    eng.soft_forward(cost_d, 1.0); eng.soft_backward(cost_d, 1.0, want_dleaf=True, want_dcost=True)
  main.wait_stream(sa); main.wait_stream(sb)
for name, fn in (("serial", serial), ("overlapped", overlapped), ("serial", serial), ("overlapped", overlapped)):
  for _ in range(3): fn()
  torch.cuda.synchronize()
  e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  e0.record()
  for _ in range(10): fn()
  e1.record(); torch.cuda.synchronize()
  print(L, name, e0.elapsed_time(e1) / 10, "ms/step", float(eng.total[0]), float(eng.soft_total[0]))
