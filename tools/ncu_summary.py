"""Text summary of an `ncu --set full --import-source on` report: per launch the headline metrics, the SASS opcode mix
weighted by executions, the hottest basic blocks (instructions that share an execution count) and the warp-stall
reasons.  usage: python tools/ncu_summary.py REPORT.ncu-rep [kernel-name-substring] > profiles/NAME.txt"""

import collections
import csv
import io
import subprocess
import sys

METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
           "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
           "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "smsp__inst_executed.sum",
           "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
           "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
           "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
           "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
           "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "lts__t_sector_hit_rate.pct",
           "l1tex__t_sector_hit_rate.pct", "sm__cycles_elapsed.max"]


def ncu(rep, *args):
  return subprocess.run(["ncu", "-i", rep, *args], capture_output=True, text=True).stdout


def main():
  rep = sys.argv[1]
  want = sys.argv[2] if len(sys.argv) > 2 else ""
  raw = list(csv.reader(io.StringIO(ncu(rep, "--page", "raw", "--csv"))))
  hdr, units, rows = raw[0], raw[1], raw[2:]
  col = {h: i for i, h in enumerate(hdr)}
  for r in rows:
    name = r[col["Kernel Name"]]
    if want not in name:
      continue
    print("==", name)
    for m in METRICS:
      if m in col:
        print(f"   {m:86s} {r[col[m]]:>16s} {units[col[m]]}")
    stalls = [(float(r[i].replace(",", "") or 0), h) for h, i in col.items()
              if h.startswith("smsp__pcsamp_warps_issue_stalled_") and not h.endswith("_not_issued")]
    tot = sum(v for v, _ in stalls) or 1
    print("   warp-stall samples: " + ", ".join(f"{h.replace('smsp__pcsamp_warps_issue_stalled_', '')} {100 * v / tot:.1f}%"
                                              for v, h in sorted(stalls, reverse=True)[:8]))
  src = list(csv.reader(io.StringIO(ncu(rep, "--page", "source", "--csv", "--print-source", "sass"))))
  kernels, cur = [], None
  for r in src:
    if r and r[0] == "Kernel Name":
      cur = {"name": r[1], "rows": []}
      kernels.append(cur)
    elif cur is not None and r and r[0] != "Address":
      cur["rows"].append(r)
  seen = set()
  for k in kernels:
    total = sum(int(r[5]) for r in k["rows"])
    key = (k["name"], total)
    if want not in k["name"] or key in seen:
      continue
    seen.add(key)
    print(f"-- {k['name']}: {total} warp instructions")
    op_of = lambda r: (r[1].split()[1] if r[1].strip().startswith("@") else r[1].split()[0]).split(".")[0]
    mix, smp = collections.Counter(), collections.Counter()
    for r in k["rows"]:
      mix[op_of(r)] += int(r[5])
      smp[op_of(r)] += int(r[2])
    st = sum(smp.values()) or 1
    print("   opcode mix: " + ", ".join(f"{o} {100 * c / total:.1f}% (stall samples {100 * smp[o] / st:.1f}%)"
                                      for o, c in mix.most_common(12)))
    blocks = collections.defaultdict(list)
    for r in k["rows"]:
      blocks[int(r[5])].append(r)
    for cnt, rs in sorted(blocks.items(), key=lambda kv: -kv[0] * len(kv[1]))[:5]:
      ops = collections.Counter(op_of(r) for r in rs)
      print(f"   block executed {cnt} x {len(rs)} instr = {100 * cnt * len(rs) / total:.1f}% of instructions, "
            f"{100 * sum(int(r[2]) for r in rs) / st:.1f}% of stall samples: "
            + ", ".join(f"{o} {c}" for o, c in ops.most_common(10)))


if __name__ == "__main__":
  main()
