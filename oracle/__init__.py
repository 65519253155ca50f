"""TEST INFRASTRUCTURE ONLY — CPU restatement of the maraxen/trex tree-scoring hot path.

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl reference`` legs may
import it, and there only as the checker (or as the timed CPU baseline), never as the path
that is shipped or measured as the GPU result.  ``trex_b200`` never imports this package and
fails loudly when ``libtrex_b200.so`` is missing.

Pinning status:
* hard Sankoff / traceback / tree cost: the reference is pure JAX and JAX is not installable
  here, so the restatement is pinned by (a) running the UNMODIFIED reference source files
  under a NumPy stand-in for the handful of jax primitives they use
  (``tests/golden/make_golden.py`` → ``tests/golden/*.npz``), (b) the hand-derived known
  answer for the reference's own test inputs and (c) independent cross-checks
  (brute force, Fitch, ``compute_cost`` consistency, the naive O(N^2 L Q) tree-cost form).
* relaxed (softmin) Sankoff: the reference has no such code — **parity unpinned**; the
  restatement is the specification, checked by float64 finite differences and its T→0 limit.
"""
