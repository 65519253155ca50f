#!/usr/bin/env python
"""Dump the judged metrics of an .ncu-rep (ncu --set full) as text: per kernel launch the
duration, DRAM bytes, throughput percentages, occupancy, and the SASS opcode histogram with
stall-sample shares.  Usage: python profiles/summarize_ncu.py gpurun_out/prof.ncu-rep > profiles/x.txt"""
import collections
import csv
import io
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "lts__t_sector_hit_rate.pct", "sm__cycles_elapsed.max"]


def ncu(rep, *args):
  return subprocess.run(["ncu", "-i", rep, *args], capture_output=True, text=True).stdout


def main(rep):
  rows = list(csv.reader(io.StringIO(ncu(rep, "--page", "raw", "--csv"))))
  hdr, units, data = rows[0], rows[1], rows[2:]
  idx = {h: i for i, h in enumerate(hdr)}
  names = []
  for d in data:
    name = d[idx["Kernel Name"]]
    names.append(name)
    print(f"== {name}")
    for w in WANT:
      if w in idx:
        print(f"   {w:72s} {d[idx[w]]:>22s} {units[idx[w]]}")
    # ncu scales each launch's value on its own (Gbyte / Mbyte / Kbyte in the same column): normalise to GB
    scale = {"byte": 1e-9, "Kbyte": 1e-6, "Mbyte": 1e-3, "Gbyte": 1.0, "Tbyte": 1e3}
    def gb(col):
      raw = d[idx[col]]
      return float(raw) * scale.get(units[idx[col]], 1.0)
    print(f"   {'traffic = dram read + write':72s} {gb('dram__bytes_read.sum') + gb('dram__bytes_write.sum'):22.6f} Gbyte")
  for name in dict.fromkeys(n.split("<")[0].split("(")[0].replace("void ", "").strip() for n in names):
    out = ncu(rep, "--page", "source", "--csv", "--kernel-name", "regex:" + name)
    rows = list(csv.reader(io.StringIO(out)))
    hi = next((i for i, r in enumerate(rows) if "Source" in r and "Instructions Executed" in r), None)
    if hi is None:
      continue
    h = rows[hi]
    si, ie, ss = h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
    cnt, smp = collections.Counter(), collections.Counter()
    for r in rows[hi + 1:]:
      try:
        n, s = int(r[ie]), int(r[ss])
      except (ValueError, IndexError):
        continue
      toks = r[si].split()
      if not toks:
        continue
      op = (toks[1] if toks[0].startswith("@") and len(toks) > 1 else toks[0]).split(".")[0]
      cnt[op] += n
      smp[op] += s
    tot, ts = sum(cnt.values()), max(1, sum(smp.values()))
    print(f"-- SASS opcode mix of {name} (first matching launch): {tot} warp instructions")
    for k, v in cnt.most_common(16):
      print(f"   {k:10s} {v:14d} {100 * v / tot:5.1f}%   stall samples {100 * smp[k] / ts:5.1f}%")


if __name__ == "__main__":
  main(sys.argv[1])
