#!/usr/bin/env python
"""Group the SASS page of an .ncu-rep into runs of instructions with the same execution count
(~basic blocks) and print, per run: executions, static length, dynamic instructions, stall
samples and the opcode mix.  Usage: python profiles/ncu_blocks.py rep.ncu-rep [kernel-regex] [min%]"""
import collections
import csv
import io
import subprocess
import sys


def main(rep, kern=None, min_pct=1.0):
  args = ["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"]
  if kern:
    args += ["--kernel-name", "regex:" + kern]
  out = subprocess.run(args, capture_output=True, text=True).stdout
  rows = list(csv.reader(io.StringIO(out)))
  hi = next(i for i, r in enumerate(rows) if "Source" in r and "Instructions Executed" in r)
  h = rows[hi]
  si, ie, ss = h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
  blocks, cur = [], None
  for r in rows[hi + 1:]:
    try:
      n, s = int(r[ie]), int(r[ss])
    except (ValueError, IndexError):
      break
    toks = r[si].split()
    op = (toks[1] if toks[0].startswith("@") and len(toks) > 1 else toks[0]).split(".")[0]
    if cur is None or cur["n"] != n:
      cur = {"n": n, "len": 0, "samples": 0, "ops": collections.Counter(), "addr": r[0]}
      blocks.append(cur)
    cur["len"] += 1
    cur["samples"] += s
    cur["ops"][op] += 1
  tot = sum(b["n"] * b["len"] for b in blocks)
  ts = max(1, sum(b["samples"] for b in blocks))
  print(f"total dynamic warp instructions {tot}, static {sum(b['len'] for b in blocks)}")
  off = 0
  for b in blocks:
    dyn = b["n"] * b["len"]
    if 100.0 * dyn / tot >= min_pct or 100.0 * b["samples"] / ts >= min_pct:
      mix = " ".join(f"{k}:{v}" for k, v in b["ops"].most_common(8))
      print(f"@{off:6d} exec {b['n']:10d} x len {b['len']:5d} = {100.0 * dyn / tot:5.1f}% dyn, "
            f"{100.0 * b['samples'] / ts:5.1f}% samples | {mix}")
    off += b["len"]


if __name__ == "__main__":
  main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else None, float(sys.argv[3]) if len(sys.argv) > 3 else 1.0)
