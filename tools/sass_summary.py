"""Per-kernel SASS instruction counts of the built library (static): packed f32x2 (FFMA2 / FADD2 / FMUL2), scalar FP32, loads and
stores by space, local-memory traffic, MUFU, branches.  `python tools/sass_summary.py > profiles/rNN_sass_summary.txt`."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "implisolid_b200", "libimplisolid_b200.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
demangle = {}
names = re.findall(r"Function : (\S+)", sass)
if names:
    out = subprocess.run(["cu++filt"] + names, capture_output=True, text=True).stdout.splitlines()
    demangle = dict(zip(names, out))
cols = ["total", "FFMA2", "FADD2", "FMUL2", "FFMA", "FADD", "FMUL", "MUFU", "LDG", "STG", "LDS", "STS", "LDL", "STL", "LDC", "BRA", "ATOM"]
rows = []
cur, cnt = None, None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        if cur:
            rows.append((cur, cnt))
        cur, cnt = m.group(1), collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and cur:
        op = m.group(1)
        cnt["total"] += 1
        key = {"ATOMG": "ATOM", "RED": "ATOM", "ATOMS": "ATOM", "LDCU": "LDC"}.get(op, op)
        if key in cols:
            cnt[key] += 1
if cur:
    rows.append((cur, cnt))
print("# cuobjdump -sass %s: static instruction counts per kernel (sm_100a)" % os.path.relpath(lib, ROOT))
print("%-44s" % "kernel" + "".join("%7s" % c for c in cols))
for name, c in sorted(rows, key=lambda r: -r[1]["total"]):
    d = demangle.get(name, name)
    d = re.sub(r"\(.*", "", d).replace("void ", "").replace("isd::", "")
    print("%-44s" % d[:43] + "".join("%7d" % c[k] for k in cols))
tma = len(re.findall(r"UBLKCP|UTMALDG|SYNCS", sass))
print("# bulk-async / TMA / mbarrier instructions (UBLKCP, UTMALDG, SYNCS) in the library: %d" % tma)
