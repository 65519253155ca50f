"""File formats: FASTA / PHYLIP / Newick -> the input contract of the hot path.  The reference has
no file format, so there is no parity target; the tests pin the contract (u8 states with 255 = no state; children before
parents, root last, adjacency[i, parent[i]] = 1), round trips, and — on the GPU — that parsed inputs score exactly like
the arrays they were written from.  The parsers are host code: everything but the last test runs without a GPU."""

import numpy as np
import pytest

from trex_b200 import io, synth


def _clades(parent, n):
  """set of leaf sets, one per internal node: the topology, whatever the numbering"""
  parent = np.asarray(parent)
  leaves = {v: frozenset([v]) for v in range(n)}
  kids = {}
  for v, p in enumerate(parent):
    if p >= 0:
      kids.setdefault(int(p), []).append(v)
  def rec(v):
    if v in leaves:
      return leaves[v]
    s = frozenset().union(*[rec(c) for c in kids[v]])
    leaves[v] = s
    return s
  import sys
  sys.setrecursionlimit(100000)
  return {rec(v) for v in kids}


@pytest.mark.parametrize("alphabet,A", [(io.DNA, 4), (io.PROTEIN, 20)])
def test_alignment_round_trips_in_every_layout(alphabet, A):
  rng = np.random.default_rng(A)
  n, L = 9, 137
  st = rng.integers(0, A, size=(n, L)).astype(np.uint8)
  st[rng.random((n, L)) < 0.05] = 255                       # gaps -> "no state"
  names = [f"taxon_{i}" for i in range(n)]
  for text, fmt in [(io.format_fasta(st, names, alphabet, width=60), "fasta"),
                    (io.format_fasta(st, names, alphabet, width=1000), "fasta"),
                    (io.format_phylip(st, names, alphabet), "phylip"),
                    (io.format_phylip(st, names, alphabet, interleaved_width=50), "phylip")]:
    got, got_names = io.parse_alignment(text, fmt, alphabet)
    assert got.dtype == np.uint8 and np.array_equal(got, st) and got_names == names
  # sequential PHYLIP with wrapped sequences (line count not a multiple of the taxon count)
  seqs = ["".join(alphabet[s] if s < A else "-" for s in row) for row in st]
  wrapped = f"{n} {L}\n" + "".join(f"{nm} {s[:100]}\n{s[100:]}\n" for nm, s in zip(names, seqs)) + "\n\n"
  lines = [l for l in wrapped.split("\n") if l.strip()]
  if (len(lines) - 1) % n == 0:                              # make sure the sequential branch is the one exercised
    wrapped = f"{n} {L}\n" + "".join(f"{nm} {s[:50]}\n{s[50:100]}\n{s[100:]}\n" for nm, s in zip(names[:-1], seqs[:-1]))
    wrapped += f"{names[-1]} {seqs[-1][:100]}\n{seqs[-1][100:]}\n"
  got, _ = io.parse_alignment(wrapped, "phylip", alphabet)
  assert np.array_equal(got, st)


def test_alignment_details_and_errors():
  from trex_b200._lib import TrexError

  text = "; old-style comment\n> seq1 some description\nacgtNN-?\r\n>seq2\nAC GT\nacgt\n"
  st, names = io.parse_alignment(text)
  assert names == ["seq1", "seq2"]
  assert st.tolist() == [[0, 1, 2, 3, 255, 255, 255, 255], [0, 1, 2, 3, 0, 1, 2, 3]]   # case-insensitive, unknown -> 255
  with pytest.raises(TrexError, match="must be aligned"):
    io.parse_alignment(">a\nACGT\n>b\nACG\n")
  with pytest.raises(TrexError, match="before the first"):
    io.parse_alignment("ACGT\n>a\nACGT\n")
  with pytest.raises(TrexError, match="no records"):
    io.parse_alignment("\n\n")
  with pytest.raises(TrexError, match="number of taxa"):
    io.parse_alignment("hello\nworld\n", "phylip")
  with pytest.raises(TrexError, match="header says"):
    io.parse_alignment("2 5\na ACGT\nb ACGTA\n", "phylip")


@pytest.mark.parametrize("kind,n", [("balanced", 16), ("cat", 40), ("random", 33), ("random", 2), ("cat", 3000)])
def test_newick_gives_the_node_order_run_dp_needs(kind, n):
  rng = np.random.default_rng(n)
  par = {"balanced": synth.balanced_parents, "cat": synth.caterpillar_parents}.get(kind, lambda k: synth.random_join_parents(k, rng))(n)
  names = [f"L{i}" for i in range(n)]
  text = io.format_newick(par, names)
  shuffled = [names[i] for i in rng.permutation(n)]          # the alignment's row order differs from the tree's
  adj, parent, leaf_names = io.parse_newick(text, shuffled)
  n_all = 2 * n - 1
  assert parent.shape == (n_all,) and parent[-1] == -1 and leaf_names == shuffled
  assert np.all(parent[:-1] > np.arange(n_all - 1))          # every child index below its parent's, root last
  assert np.all(parent[:-1] >= n)                            # parents are internal nodes
  assert np.all(np.bincount(parent[:-1], minlength=n_all)[n:] == 2)   # binary
  want = np.zeros((n_all, n_all), np.float32)
  want[np.arange(n_all - 1), parent[:-1]] = 1
  assert np.array_equal(adj, want) and adj[-1].sum() == 0    # no root self-loop
  # same topology once leaves are matched by NAME
  to_orig = [names.index(nm) for nm in leaf_names]
  got = {frozenset(to_orig[v] for v in c) for c in _clades(parent, n)}
  assert got == _clades(par, n)
  _, parent2, names2 = io.parse_newick(text)                 # no taxon list: order of appearance
  assert sorted(names2) == sorted(names) and np.all(parent2[:-1] > np.arange(n_all - 1))


def test_newick_syntax_and_errors():
  from trex_b200._lib import TrexError

  adj, parent, names = io.parse_newick("[a comment] ((A:0.1,'B b''s':2e-3)inner:1.5[&x=1],((C)):0.2)root;")
  assert names == ["A", "B b's", "C"] and parent.tolist() == [3, 3, 4, 4, -1]      # unary nodes are transparent
  with pytest.raises(TrexError, match="resolve multifurcations"):
    io.parse_newick("(A,B,C);")
  with pytest.raises(TrexError, match="expected"):
    io.parse_newick("((A,B);")
  with pytest.raises(TrexError, match="not among the taxon names"):
    io.parse_newick("(A,B);", ["A", "X"])
  with pytest.raises(TrexError, match="taxon names given"):
    io.parse_newick("(A,B);", ["A", "B", "C"])
  with pytest.raises(TrexError, match="at least two leaves"):
    io.parse_newick("A;")


@pytest.mark.gpu
@pytest.mark.parametrize("A,alphabet", [(4, io.DNA), (20, io.PROTEIN)])
def test_parsed_files_score_like_the_arrays_they_were_written_from(A, alphabet, tmp_path):
  import torch
  from trex_b200.sankoff import run_sankoff

  rng = np.random.default_rng(A)
  n, L = 24, 300
  par = synth.random_join_parents(n, rng)
  states = synth.evolve(par, L, A, 0.2, rng)[:n]
  states[3, 7] = 255
  names = [f"sp{i}" for i in range(n)]
  order = rng.permutation(n)                                  # the FASTA lists the taxa in another order than the tree
  (tmp_path / "aln.fasta").write_text(io.format_fasta(states[order], [names[i] for i in order], alphabet))
  (tmp_path / "tree.nwk").write_text(io.format_newick(par, names))
  st, aln_names = io.read_alignment(tmp_path / "aln.fasta", alphabet=alphabet)
  adj, parent, _ = io.read_newick(tmp_path / "tree.nwk", aln_names)
  cost = synth.unit_cost(A)
  n_all = 2 * n - 1
  got = run_sankoff(torch.from_numpy(adj), torch.from_numpy(cost), torch.from_numpy(st.astype(np.float32)), n_all, A, n)[2]
  want = run_sankoff(torch.from_numpy(synth.parents_to_adjacency(par)), torch.from_numpy(cost),
                     torch.from_numpy(states.astype(np.float32)), n_all, A, n)[2]
  assert float(got) == float(want) and float(got) > 0
