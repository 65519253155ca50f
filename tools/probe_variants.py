"""Times soft_bwd (unit cost, compact d-leaf, +dC) at the c5 shape for every library build in tools/_bin/lib_*.so
(kernel-variant experiments; each variant runs in its own process because the library path is fixed at import)."""
import glob, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CODE = r'''
import os, sys, numpy as np, torch
sys.path.insert(0, %r)
from trex_b200 import synth
from trex_b200.sankoff import SankoffEngine
n, A = 1024, 20
for L in (100000, 12500):
  rng = np.random.default_rng(5)
  par = synth.balanced_parents(n); adj = synth.parents_to_adjacency(par)
  states = synth.evolve(par, L, A, 0.1, rng)[:n]
  eng = SankoffEngine(2*n-1, n, A, L)
  eng.set_adjacency(torch.from_numpy(adj)).set_leaf_states_u8(torch.from_numpy(np.ascontiguousarray(states)))
  cu = torch.from_numpy(synth.unit_cost(A)).cuda()
  eng.soft_forward(cu, 1.0)
  def t(f, reps=8):
    for _ in range(3): f()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(reps): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
  a = t(lambda: eng.soft_backward(cu, 1.0, want_dleaf="compact", want_dcost=True))
  b = t(lambda: eng.soft_backward(cu, 1.0, want_dleaf="compact", want_dcost=False))
  print("  L=%%d  +dC %%.3f ms   -dC %%.3f ms   dcost[0,1]=%%.6e" %% (L, a, b, float(eng.dcost[0,0,1])), flush=True)
  del eng; torch.cuda.empty_cache()
''' % ROOT
libs = [""] + sorted(glob.glob(os.path.join(ROOT, "tools", "_bin", "lib_*.so")))
for lib in libs:
  print(os.path.basename(lib) or "in-tree build", flush=True)
  env = dict(os.environ)
  if lib: env["TREX_B200_LIB"] = lib
  subprocess.run([sys.executable, "-c", CODE], env=env)
