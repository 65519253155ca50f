import sys, os
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import tree_oracle as to
from trex_b200 import _lib
from trex_b200.tree import tree_cost_value_and_grad
for (N, L, Q) in [(127, 1000, 4), (10, 100, 20), (200, 150, 20), (7, 4, 4)]:
  rng = np.random.default_rng(N)
  S = to.softmax(rng.normal(size=(N, L, Q))).astype(np.float32)
  adj = (rng.random((N, N)) < 0.1).astype(np.float32)
  St, At = torch.from_numpy(S).cuda(), torch.from_numpy(adj).cuda()
  ws = torch.zeros(_lib.load().trex_tree_cost_workspace_bytes(N, L * Q, Q), dtype=torch.uint8, device="cuda")
  out = {"cost": torch.empty((), device="cuda"), "dS": torch.zeros_like(St), "dA": torch.empty_like(At)}
  tree_cost_value_and_grad(St, At, None, workspace=ws, out=out)
  torch.cuda.synchronize()
  want, dS64, dA64, _ = to.soft_cost_value_and_grad(S, adj, None)
  dS = out["dS"].cpu().numpy()
  print(N, L, Q, "cost", float(out["cost"]), want, "dA err", np.abs(out["dA"].cpu().numpy() - dA64).max() / np.abs(dA64).max(),
        "dS err", np.abs(dS - dS64).max() / np.abs(dS64).max())
  d = np.abs(dS - dS64).reshape(N, -1)
  bad = np.argwhere(d > 1e-4 * np.abs(dS64).max())
  print("  bad count", len(bad), "of", d.size, "first", bad[:5].tolist(), "rows bad", np.unique(bad[:, 0])[:10], "cols bad", np.unique(bad[:, 1])[:12])
  if len(bad):
    i, j = bad[0]
    print("  got", dS.reshape(N, -1)[i, j:j+4], "want", dS64.reshape(N, -1)[i, j:j+4])
