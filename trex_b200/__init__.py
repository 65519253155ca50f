"""trex_b200 — B200-native (sm_100a) tree-scoring hot path behind the maraxen/trex API.

Only what the hot path needs lives here: ``csrc/`` (CUDA kernels + the C ABI of
``libtrex_b200.so``), the ctypes binding (``_lib``) and host-side mirrors of the reference
modules (``sankoff``, ``tree``, ``relaxed``, ``optimize``, ``landscape``).  Importing the package does not need a GPU;
calling into it does, and there is no CPU / PyTorch fallback.
"""

__all__ = ["sankoff", "tree", "relaxed", "optimize", "landscape", "synth"]
