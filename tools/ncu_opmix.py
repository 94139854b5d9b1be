"""Dynamic instruction mix of one kernel from an ncu report's source page:
   ncu -i X.ncu-rep --page source --csv --kernel-name regex:NAME > src.csv ; python tools/ncu_opmix.py src.csv [top]
Prints warp-instructions executed per opcode (share of the kernel), mean active threads, and the hottest SASS lines."""
import csv
import re
import sys
from collections import defaultdict

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
rows = list(csv.reader(open(path)))
hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hdr_i]
ix = {n: i for i, n in enumerate(hdr)}
ops = defaultdict(lambda: [0, 0, 0])
total = thr = 0
for r in rows[hdr_i + 1:]:
    if len(r) < len(hdr) or not r[ix["Instructions Executed"]].isdigit():
        continue
    src = r[ix["Source"]].strip()
    m = re.match(r"(?:@!?U?P[0-9T]+\s+)?([A-Z0-9_]+)", src)
    if not m:
        continue
    n = int(r[ix["Instructions Executed"]])
    t = int(r[ix["Thread Instructions Executed"]])
    s = int(r[ix["# Samples"]] or 0)
    op = m.group(1)
    ops[op][0] += n
    ops[op][1] += t
    ops[op][2] += s
    total += n
    thr += t
print("%s: %d warp instructions, %.1f active threads/instruction" % (rows[0][1][:90] if len(rows[0]) > 1 else path, total, thr / max(total, 1)))
groups = {"fp32": ("FADD", "FMUL", "FFMA", "FADD2", "FMUL2", "FFMA2"), "fp-cmp/sel": ("FSETP", "FSEL", "FMNMX", "FCHK", "FSET"), "mufu": ("MUFU",),
          "int/logic": ("IMAD", "IADD3", "LOP3", "SHF", "LEA", "ISETP", "VIADD", "IADD", "SEL", "PLOP3", "I2FP", "F2I", "I2F", "F2F", "PRMT", "VIADDMNMX", "IABS"),
          "mov": ("MOV", "HFMA2", "CS2R", "UMOV", "S2R", "R2UR", "S2UR"), "ld/st global": ("LDG", "STG", "ATOMG", "RED", "ATOM"), "local": ("LDL", "STL"),
          "shared": ("LDS", "STS"), "const": ("LDC", "LDCU", "ULDC"), "branch": ("BRA", "BSSY", "BSYNC", "CALL", "RET", "EXIT", "BRX", "WARPSYNC", "BAR", "JMP"),
          "fp64": ("DADD", "DMUL", "DFMA", "DSETP")}
g = defaultdict(int)
for op, (n, t, s) in ops.items():
    name = next((k for k, v in groups.items() if op in v), "other")
    g[name] += n
print("  ".join("%s %.1f%%" % (k, 100.0 * v / total) for k, v in sorted(g.items(), key=lambda kv: -kv[1])))
for op, (n, t, s) in sorted(ops.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%-10s %12d %5.1f%%  thr/inst %4.1f  samples %d" % (op, n, 100.0 * n / total, t / max(n, 1), s))
