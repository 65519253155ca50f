"""Errors of the relaxed path vs the float64 oracle at the benchmark shapes, next to the errors of the
float32 instance of the same C restatement (what an f32 implementation of the recurrence can achieve)."""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import c_oracle
from trex_b200 import synth
from trex_b200.sankoff import SankoffEngine

out = []
for shape, (n, A, L, seed) in {"c5slice": (1024, 20, 2000, 5), "c3slice": (256, 20, 4096, 3)}.items():
  for ck in ("unit", "int"):
    rng = np.random.default_rng(seed)
    par = synth.balanced_parents(n); adj = synth.parents_to_adjacency(par)
    states = np.ascontiguousarray(synth.evolve(par, L, A, 0.1, rng)[:n])
    cost = synth.unit_cost(A) if ck == "unit" else synth.random_int_cost(A, 5, rng)
    for T in (1.0, 0.1):
      tot64, dC64, dp64, g64 = c_oracle.soft_sankoff_f64(adj, cost, 2*n-1, A, n, T, seqs=states.astype(np.float32))
      tot32, dC32, g32 = c_oracle.soft_sankoff(adj, cost, states.astype(np.float32), 2*n-1, A, n, T, want_gbar=True)
      for seg in (0, 1):
        os.environ["TREX_SEGMENTS"] = str(seg)
        eng = SankoffEngine(2*n-1, n, A, L)
        eng.set_adjacency(torch.from_numpy(adj)).set_leaf_states_u8(torch.from_numpy(states))
        c = torch.from_numpy(cost).cuda()
        eng.soft_forward(c, T)
        dp = eng.export_dp(soft=True).cpu().numpy()
        eng.soft_backward(c, T, want_dleaf="compact", want_dcost=True)
        gbar = eng.gbar[0, :, :, :L].permute(2, 0, 1).cpu().numpy()
        dC = eng.dcost[0].cpu().numpy()
        r = {"shape": shape, "cost": ck, "T": T, "seg": seg, "max_dp": float(np.abs(dp64[:, n:]).max()),
             "dp_rel": float(np.abs(dp[:, n:] - dp64[:, n:]).max() / np.abs(dp64[:, n:]).max()),
             "total_rel": abs(float(eng.soft_total[0]) - tot64) / abs(tot64),
             "gbar_abs": float(np.abs(gbar - g64[:, n:]).max()),
             "dC_rel": float(np.abs(dC - dC64).max() / np.abs(dC64).max()),
             "f32port_total_rel": abs(tot32 - tot64) / abs(tot64),
             "f32port_gbar_abs": float(np.abs(g32[:, n:] - g64[:, n:]).max()),
             "f32port_dC_rel": float(np.abs(dC32 - dC64).max() / np.abs(dC64).max())}
        out.append(r)
        print(json.dumps(r), flush=True)
        del eng
