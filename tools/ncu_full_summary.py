"""Summarise an `ncu --set full` report (exported with `ncu -i X.ncu-rep --page raw --csv > X.csv`): one row per
launch with duration, DRAM bytes read/written, DRAM / SM throughput %, issue-slot %, FMA-pipe %, occupancy, registers.

  python tools/ncu_full_summary.py gpurun_out/full.csv [profiles/out.txt] ["header comment"]
"""
import csv
import sys

COLS = [("gpu__time_duration.sum", "us", 1e-3), ("dram__bytes_read.sum", "rdMB", None), ("dram__bytes_write.sum", "wrMB", None),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%", 1), ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%", 1),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%", 1),
        ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma%", 1),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%", 1), ("l1tex__t_sector_hit_rate.pct", "l1hit%", 1),
        ("lts__t_sector_hit_rate.pct", "l2hit%", 1), ("launch__registers_per_thread", "regs", 1), ("launch__grid_size", "grid", 1),
        ("launch__block_size", "blk", 1)]


def to_bytes(v, unit):
    v = float(v.replace(",", ""))
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


def main():
    path = sys.argv[1]
    out = sys.argv[2] if len(sys.argv) > 2 else None
    head = sys.argv[3] if len(sys.argv) > 3 else ""
    with open(path) as f:
        rows = list(csv.reader(l for l in f if not l.startswith("==")))
    names, units, data = rows[0], rows[1], rows[2:]
    ix = {n: i for i, n in enumerate(names)}
    lines = ["# " + head] if head else []
    lines.append("%-30s" % "kernel" + "".join("%9s" % c[1] for c in COLS))
    for r in data:
        k = r[ix["Kernel Name"]].split("(")[0].replace("isd::", "")
        cells = []
        for name, label, scale in COLS:
            if name not in ix or r[ix[name]] == "":
                cells.append("%9s" % "-")
                continue
            if scale is None:
                cells.append("%9.2f" % (to_bytes(r[ix[name]], units[ix[name]]) / 1e6))
            else:
                v = float(r[ix[name]].replace(",", ""))
                if label == "us":
                    v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(units[ix[name]], 1e-3)
                cells.append("%9.1f" % v if label not in ("regs", "grid", "blk") else "%9d" % v)
        lines.append("%-30s" % k[:30] + "".join(cells))
    text = "\n".join(lines)
    print(text)
    if out:
        with open(out, "w") as f:
            f.write(text + "\n")


if __name__ == "__main__":
    main()
