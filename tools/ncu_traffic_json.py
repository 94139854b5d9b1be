"""profiles/rNN_traffic.json from an `ncu --set full` raw-page CSV: DRAM bytes (read + write) per kernel per step, under the
names bench.py's kernel table uses (the kernels' own names; the adjacency kernels summed).

  python tools/ncu_traffic_json.py gpurun_out/full.csv profiles/r02_traffic.json "workload text" "source text"
"""
import csv
import json
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[0]
ix = {n: i for i, n in enumerate(hdr)}
ADJ = ("k_vertex_degree", "k_uscan", "k_umbrella_fill", "k_umbrella_sort", "k_mc_adjacency")
out = {}
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    name = re.sub(r"^void\s+", "", r[ix["Kernel Name"]])
    name = re.sub(r"<.*", "", name).split("(")[0].replace("isd::", "")
    if name in ADJ:
        name = "adjacency"
    def num(col):
        return float(r[ix[col]].replace(",", ""))
    unit_r, unit_w = rows[1][ix["dram__bytes_read.sum"]], rows[1][ix["dram__bytes_write.sum"]]
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    b = num("dram__bytes_read.sum") * scale[unit_r] + num("dram__bytes_write.sum") * scale[unit_w]
    us = num("gpu__time_duration.sum") * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3}[rows[1][ix["gpu__time_duration.sum"]]]
    e = out.setdefault(name, {"dram_bytes": 0.0, "us": 0.0, "launches": 0})
    e["dram_bytes"] += b
    e["us"] += us
    e["launches"] += 1
json.dump({"workload": sys.argv[3], "source": sys.argv[4], "kernels": out}, open(sys.argv[2], "w"), indent=1)
for k, v in out.items():
    print("%-28s %10.2f MB %9.1f us x%d" % (k, v["dram_bytes"] / 1e6, v["us"], v["launches"]))
