"""Host-buffer front end of the Sankoff hot path: what a caller with NumPy / pinned host arrays uses.

One *job* is the work of ``run_sankoff(adjacency, cost, sequences, ..., return_path=True)``
(``/root/reference/src/trex/sankoff.py:114-188``: device-built schedule, hard forward, root argmin +
traceback) plus the relaxed forward / reverse pass on the same inputs.  Inputs are HOST
arrays (adjacency f32, cost f32, leaf states u8); the results come back in HOST memory:

* ``anc``      u8 (n_int, L)  — the ancestral reconstruction, rows ``n_leaves:`` of the reference's
  ``reconstructed_seqs`` (``sankoff.py:182-185``; its leaf rows are the caller's own input);
* ``packed``   f64 ``[hard total, relaxed total, dC (A*A), edge histogram (A*A)]`` (summed over ranks when a
  process group is given — the one collective of the site-sharded path).

The reference also returns the dp table (``(L, n_all, A)`` f32 — 16.4 GB at 1024 taxa x 100k sites,
SURVEY F7); here it stays on the device (``engine.dp``, native layout) together with the relaxed marginals
``engine.gbar`` and is exported on request (``SankoffEngine.export_dp``).

Jobs are independent, so the pipeline is double-buffered: the host->device copies of job k+1 run on a copy
stream under the kernels of job k, the device->host copy of job k's reconstruction under the kernels of
job k+1.  With ``use_graph`` the kernels of each buffer slot are captured once into a CUDA graph.
"""

from __future__ import annotations

import torch

from .dist import PackedAllReduce
from .sankoff import SankoffEngine


class SankoffHostPipeline:

  def __init__(self, n_all, n_leaves, n_states, n_sites, temperature=1.0, leaf_grad="compact", group=None,
               use_graph=True, engine: SankoffEngine | None = None):
    self.eng = engine if engine is not None else SankoffEngine(n_all, n_leaves, n_states, n_sites)
    e = self.eng
    self.T, self.leaf_grad, self.group = float(temperature), leaf_grad, group
    A = e.A
    dev = e.device
    # slot 0 uses the engine's own leaf / anc tables, slot 1 a second pair; every other table is shared
    self.engs = [e, e.twin(own=("leaf", "anc"))]
    self.adj_d = [torch.empty((e.n_all, e.n_all), dtype=torch.float32, device=dev) for _ in range(2)]
    self.cost_d = [torch.empty((A, A), dtype=torch.float32, device=dev) for _ in range(2)]
    self.reducers = [PackedAllReduce([(1,), (1,), (1, A, A), (1, A, A)], device=dev) for _ in range(2)]
    self.anc_h = [torch.empty((e.n_int, e.L), dtype=torch.uint8).pin_memory() for _ in range(2)]
    self.packed_h = [torch.empty(2 + 2 * A * A, dtype=torch.float64).pin_memory() for _ in range(2)]
    self.h2d = torch.cuda.Stream(device=dev)
    self.d2h = torch.cuda.Stream(device=dev)
    self.use_graph = use_graph
    self._graphs = [None, None]
    self.h2d_bytes_per_job = int(e.n_leaves * e.L + e.n_all * e.n_all * 4 + A * A * 4)
    self.d2h_bytes_per_job = int(e.n_int * e.L + self.packed_h[0].numel() * 8)

  # -- one job on slot j, all on the current stream -------------------------------------------------
  def _kernels(self, j):
    e = self.engs[j]
    e.set_adjacency(self.adj_d[j])
    e.hard_forward(self.cost_d[j], want_argmin=True)
    e.traceback(want_hist=True)
    e.soft_forward(self.cost_d[j], self.T)
    e.soft_backward(self.cost_d[j], self.T, want_dleaf=self.leaf_grad, want_dcost=True)
    self.reducers[j].pack([e.total, e.soft_total, e.dcost, e.edge_hist])

  def _compute(self, j):
    if not self.use_graph:
      return self._kernels(j)
    if self._graphs[j] is None:
      cur = torch.cuda.current_stream()
      side = torch.cuda.Stream(device=self.eng.device)
      side.wait_stream(cur)
      with torch.cuda.stream(side):
        self._kernels(j)  # allocate every lazily created buffer outside the capture
        side.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
          self._kernels(j)
      cur.wait_stream(side)
      self._graphs[j] = g
    self._graphs[j].replay()

  def _upload(self, j, adj_h, cost_h, states_h):
    self.adj_d[j].copy_(adj_h, non_blocking=True)
    self.cost_d[j].copy_(cost_h, non_blocking=True)
    self.engs[j].set_leaf_states_u8(states_h)

  def run(self, jobs):
    """``jobs``: sequence of ``(adjacency_host f32 (n_all,n_all), cost_host f32 (A,A), leaf_states_host u8
    (n_leaves, L))`` — pinned for asynchronous copies.  Returns the list of ``(anc_host, packed_host)`` of the
    LAST TWO jobs' slots as views (a streaming caller consumes slot k & 1 after ``out_done[k]``); every job's
    results are copied to the host inside this call.  The caller synchronises (``torch.cuda.synchronize``)."""
    jobs = list(jobs)
    n = len(jobs)
    main = torch.cuda.current_stream()
    ready = [torch.cuda.Event() for _ in range(n)]
    computed = [torch.cuda.Event() for _ in range(n)]
    out_done = [torch.cuda.Event() for _ in range(n)]
    self.h2d.wait_stream(main)
    self.d2h.wait_stream(main)
    if self.use_graph:
      for j in range(min(2, n)):
        if self._graphs[j] is None:  # first use: capture (runs the kernels once on whatever the slot holds)
          self._upload(j, *jobs[j])
          self._compute(j)
      main.synchronize()
    with torch.cuda.stream(self.h2d):
      self._upload(0, *jobs[0])
      ready[0].record(self.h2d)
    for k in range(n):
      j = k & 1
      main.wait_event(ready[k])
      if k + 1 < n:
        with torch.cuda.stream(self.h2d):
          if k >= 1:
            self.h2d.wait_event(computed[k - 1])  # job k-1 read the buffers job k+1 uploads into
          self._upload((k + 1) & 1, *jobs[k + 1])
          ready[k + 1].record(self.h2d)
      if k >= 2:
        main.wait_event(out_done[k - 2])  # the traceback of job k overwrites the anc table of job k-2
      self._compute(j)
      self.reducers[j].reduce(self.group)
      computed[k].record(main)
      with torch.cuda.stream(self.d2h):
        self.d2h.wait_event(computed[k])
        e = self.engs[j]
        self.anc_h[j].copy_(e.anc[0, :, : e.L], non_blocking=True)
        self.packed_h[j].copy_(self.reducers[j].buf, non_blocking=True)
        out_done[k].record(self.d2h)
    main.wait_stream(self.d2h)
    main.wait_stream(self.h2d)
    return [(self.anc_h[(n - 1 - i) & 1], self.packed_h[(n - 1 - i) & 1]) for i in range(min(2, n))]
