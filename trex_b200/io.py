"""File formats -> the input contract of the hot path, on top of the host-side parsers of
``libtrex_b200.so`` (``trex_parse_alignment_host``, ``trex_parse_newick_host``; ``csrc/formats.cu``).

The reference has no on-disk format at all — its inputs exist only as in-memory arrays — so these functions replace
nothing; they produce exactly what ``run_sankoff`` / ``SankoffEngine`` take: ``states`` u8 (n_taxa, L) with 255 =
"no state" (``/root/reference/src/trex/sankoff.py:50``: the leaf row stays all-1e5) and an adjacency (2n-1, 2n-1)
f32 with ``adjacency[i, parent[i]] = 1`` (``utils/types.py:33``), children before parents, root last
(``sankoff.py:45-46, 87-92``).  Host code only: NumPy arrays in and out, no GPU needed."""

from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib

DNA = "ACGT"
PROTEIN = "ARNDCQEGHILKMFPSTWYV"
FASTA, PHYLIP = 0, 1


def _names(buf):
  return [s for s in buf.value.decode().split("\n") if s]


def parse_alignment(text, fmt="fasta", alphabet=DNA):
  """``text`` (str or bytes) in FASTA or PHYLIP -> ``(states u8 (n_taxa, L), names list[str])``."""
  lib = _lib.load()
  data = text.encode() if isinstance(text, str) else bytes(text)
  f = {"fasta": FASTA, "phylip": PHYLIP}[fmt] if isinstance(fmt, str) else int(fmt)
  n_taxa, n_sites = C.c_int(0), C.c_int64(0)
  ab = alphabet.encode()
  _lib.call("trex_parse_alignment_host", data, len(data), f, ab, C.byref(n_taxa), C.byref(n_sites), None, 0, None, 0)
  states = np.empty((n_taxa.value, n_sites.value), dtype=np.uint8)
  names = C.create_string_buffer(len(data) + n_taxa.value + 2)
  _lib.call("trex_parse_alignment_host", data, len(data), f, ab, C.byref(n_taxa), C.byref(n_sites),
            states.ctypes.data_as(C.c_void_p), n_sites.value, names, len(names))
  return states, _names(names)


def read_alignment(path, fmt=None, alphabet=DNA):
  if fmt is None:
    fmt = "phylip" if str(path).lower().endswith((".phy", ".phylip")) else "fasta"
  with open(path, "rb") as fh:
    return parse_alignment(fh.read(), fmt, alphabet)


def parse_newick(text, taxon_names=None):
  """Newick -> ``(adjacency f32 (2n-1, 2n-1), parent i32 (2n-1,), leaf names in leaf-index order)``.  With
  ``taxon_names`` (the alignment's row order) leaf i of the tree is row i of the alignment."""
  data = text.encode() if isinstance(text, str) else bytes(text)
  tn = None if taxon_names is None else ("\n".join(taxon_names) + "\n").encode()
  n = C.c_int(0)
  _lib.call("trex_parse_newick_host", data, len(data), tn, C.byref(n), None, None, None, 0)
  n_all = 2 * n.value - 1
  parent = np.empty(n_all, dtype=np.int32)
  adj = np.empty((n_all, n_all), dtype=np.float32)
  names = C.create_string_buffer(len(data) + n.value + 2)
  _lib.call("trex_parse_newick_host", data, len(data), tn, C.byref(n), parent.ctypes.data_as(C.c_void_p),
            adj.ctypes.data_as(C.c_void_p), names, len(names))
  return adj, parent, _names(names)


def read_newick(path, taxon_names=None):
  with open(path, "rb") as fh:
    return parse_newick(fh.read(), taxon_names)


# ---- writers (plain Python: used by the tests to round-trip, and handy for exporting synthetic data) ---------------


def format_fasta(states, names, alphabet=DNA, width=60):
  out = []
  for name, row in zip(names, np.asarray(states)):
    seq = "".join(alphabet[s] if s < len(alphabet) else "-" for s in row)
    out.append(">" + name)
    out += [seq[i:i + width] for i in range(0, len(seq), width)]
  return "\n".join(out) + "\n"


def format_phylip(states, names, alphabet=DNA, interleaved_width=None):
  states = np.asarray(states)
  seqs = ["".join(alphabet[s] if s < len(alphabet) else "-" for s in row) for row in states]
  head = f" {states.shape[0]} {states.shape[1]}\n"
  if not interleaved_width:
    return head + "".join(f"{n}  {s}\n" for n, s in zip(names, seqs))
  blocks = []
  for i in range(0, states.shape[1], interleaved_width):
    rows = [(f"{n}  " if i == 0 else "") + s[i:i + interleaved_width] for n, s in zip(names, seqs)]
    blocks.append("\n".join(rows))
  return head + "\n\n".join(blocks) + "\n"


def format_newick(parent, names):
  """parent array (root: -1) -> Newick with unit branch lengths; leaves 0..n-1 carry ``names``."""
  parent = np.asarray(parent)
  kids = {}
  for v, p in enumerate(parent):
    if p >= 0:
      kids.setdefault(int(p), []).append(v)
  root = int(np.nonzero(parent < 0)[0][-1])
  order, stack = [], [root]
  while stack:  # iterative post-order
    v = stack.pop()
    order.append(v)
    stack += kids.get(v, [])
  text = {}
  for v in reversed(order):
    text[v] = names[v] if v not in kids else "(" + ",".join(text[c] + ":1" for c in kids[v]) + ")"
  return text[root] + ";"
