"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel time and share of one bench step."""
import csv
import sys
from collections import OrderedDict

path, out = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else None)
with open(path) as f:
    lines = [l for l in f if not l.startswith("==")]
rows = [(r["Kernel Name"].split("(")[0].replace("isd::", ""), float(r["Metric Value"].replace(",", ""))) for r in csv.DictReader(lines)]
rows = [r for r in rows if not r[0].startswith(("k_ffma", "k_fmul", "k_f32x2"))]
names = [r[0] for r in rows]
first = "k_tile_bounds" if "k_tile_bounds" in names else "k_eval_field"
idx = [i for i, n in enumerate(names) if n == first]
s, e = idx[len(idx) // 2], idx[len(idx) // 2 + 1]
agg = OrderedDict()
for n, v in rows[s:e]:
    agg.setdefault(n, [0.0, 0])
    agg[n][0] += v
    agg[n][1] += 1
tot = sum(v[0] for k, v in agg.items() if "at::" not in k)
lines = []
for k, v in agg.items():
    lines.append("%-44s %3d %9.1f us %6.1f%%" % (k[:44], v[1], v[0] / 1000, 100 * v[0] / tot))
lines.append("total (ours) %.1f us" % (tot / 1000))
text = "\n".join(lines)
print(text)
if out:
    with open(out, "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none; one bench.py step (C2 256^3, eval+MC+OB02)\n" + text + "\n")
