"""One-off fuzz of K1's interval bounds / liveness masks on the host (tests/emu): random CSG trees of growing depth, boxes of
several sizes near the surface.  Prints one line per tree; any non-zero violation counter is a soundness bug.
  python tools/fuzz_bounds.py [first_seed] [n_seeds]
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from implisolid_b200 import shapes          # noqa: E402
from tests.oracle_lib import load_emu       # noqa: E402

first, count = (int(sys.argv[1]) if len(sys.argv) > 1 else 100), (int(sys.argv[2]) if len(sys.argv) > 2 else 40)
emu = load_emu()
bad = 0
for seed in range(first, first + count):
    sj = shapes.random_tree(seed, depth=3 + seed % 4, all_prims=True)
    rng = np.random.default_rng(seed)
    pts = rng.uniform(-1.1, 1.1, (20000, 3)).astype(np.float32)
    f, _, _ = emu.eval(sj, pts)
    order = np.argsort(np.abs(f))[:600]
    tot = {}
    for he in (0.004, 0.015, 0.06, 0.2):
        boxes = np.concatenate([pts[order], (he * rng.uniform(0.4, 1.6, (600, 3))).astype(np.float32)], axis=1)
        nodes, st = emu.bounds_check(sj, boxes, samples=8, seed=seed)
        for k, v in st.items():
            tot[k] = tot.get(k, 0) + v
    viol = tot["bounds_violations"] + tot["value_mismatches"] + tot["gradient_mismatches"] + tot["looser_second_level"]
    bad += viol
    print(seed, nodes, tot, "VIOLATION" if viol else "")
print("violations:", bad)
