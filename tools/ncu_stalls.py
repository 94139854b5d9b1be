"""Top warp-stall reasons per kernel from an `ncu --set full` raw CSV export (`ncu -i X.ncu-rep --page raw --csv`).

  python tools/ncu_stalls.py gpurun_out/full.csv [top_n]
"""
import csv
import sys

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 6
with open(path) as f:
    rows = list(csv.reader(l for l in f if not l.startswith("==")))
names, data = rows[0], rows[2:]
stall = [(i, n) for i, n in enumerate(names) if "issue_stalled" in n and n.endswith("_per_warp_active.pct")]
if not stall:
    stall = [(i, n) for i, n in enumerate(names) if "issue_stalled" in n and "not_issued" not in n]
kix = names.index("Kernel Name")
extra = [n for n in ("smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__inst_executed_pipe_fma.sum",
                     "smsp__inst_executed_pipe_fmaheavy.sum", "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_lsu.sum",
                     "smsp__inst_executed_pipe_uniform.sum", "sm__inst_executed_pipe_xu.sum") if n in names]
for r in data:
    k = r[kix].split("(")[0].replace("isd::", "")
    vals = []
    for i, n in stall:
        try:
            vals.append((float(r[i].replace(",", "")), n.replace("smsp__average_warps_issue_stalled_", "").replace("smsp__average_warp_latency_issue_stalled_", "")
                         .replace("_per_warp_active.pct", "").replace(".ratio", "")))
        except ValueError:
            pass
    vals.sort(reverse=True)
    ex = " ".join("%s=%s" % (n.split("__")[1].replace("inst_executed_", "").replace(".sum", "").replace(".ratio", ""), r[names.index(n)]) for n in extra)
    print("%-30s %s | %s" % (k[:30], "  ".join("%s %.1f" % (n, v) for v, n in vals[:top]), ex))
