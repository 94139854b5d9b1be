"""Generates tests/golden/oracle_regression.json: regression hashes of the CPU oracle on the reference's demo specs.

The reference holds no tests or golden vectors for this path and cannot be built here, so these hashes pin
the ORACLE against accidental change (they do not pin it to the reference).  Known answers that do come from the
reference's own source (octahedron adjacency, make_a_cube) live in tests/test_oracle_known_answers.py.
Run:  python tests/golden/make_golden.py
"""
import hashlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from implisolid_b200 import shapes          # noqa: E402
from tests.oracle_lib import load_oracle    # noqa: E402

CASES = {
    "c1_sphere_r28": (shapes.c1_sphere, 28, 0.55, {}),
    "c2_cube_cyl_r40": (shapes.c2_cube_minus_cylinder, 40, 0.8, {}),
    "moon_r40": (shapes.demo_moon, 40, 1.0, {}),
    "cone_r28": (shapes.simple_cone, 28, 0.5, {}),
    "c3_metaballs_r32": (shapes.c3_metaballs, 32, 1.2, {}),
    "c4_chain_r40": (shapes.c4_deep_chain, 40, 1.0, {}),
    # the reference's serial float sum of the edge lengths (exact_reduce 1) ...
    "c1_sphere_r28_ob02": (shapes.c1_sphere, 28, 0.55, dict(iters=1, c=1.0, projection=1, qem=1, exact_reduce=1)),
    "c2_cube_cyl_r40_ob02": (shapes.c2_cube_minus_cylinder, 40, 0.8, dict(iters=1, c=1.0, projection=1, qem=1, exact_reduce=1)),
    # ... and the product's default 64-bit fixed-point sum
    "c1_sphere_r28_ob02_fixed_point_sum": (shapes.c1_sphere, 28, 0.55, dict(iters=1, c=1.0, projection=1, qem=1)),
    "c2_cube_cyl_r40_ob02_fixed_point_sum": (shapes.c2_cube_minus_cylinder, 40, 0.8, dict(iters=1, c=1.0, projection=1, qem=1)),
    "c4_chain_r40_ob02_fixed_point_sum": (shapes.c4_deep_chain, 40, 1.0, dict(iters=1, c=1.0, projection=1, qem=1)),
    "heart_r40": (shapes.n1_heart, 40, 1.0, {}),
}


def digest(a):
    return hashlib.sha256(a.tobytes()).hexdigest()[:16]


def run_case(o, shape, R, half, kw):
    o.build_geometry(shape(), shapes.mc_settings(R, half, **kw))
    v, f = o.get_v(), o.get_f()
    rec = {"nv": int(v.shape[0]), "nf": int(f.shape[0]), "verts": digest(v), "faces": digest(f), "field": digest(o.get_field()),
           "cases": digest(o.get_case_codes())}
    o.finish_geometry()
    return rec


def main():
    o = load_oracle()
    out = {name: run_case(o, *spec) for name, spec in CASES.items()}
    with open(os.path.join(ROOT, "tests", "golden", "oracle_regression.json"), "w") as fh:
        json.dump(out, fh, indent=1, sort_keys=True)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
