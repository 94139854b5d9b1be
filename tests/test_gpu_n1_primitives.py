"""GPU parity for the rest of the factory's analytic vocabulary (SURVEY §8 f-N1): itorus, iheart, tetrahedron,
implicit_double_mushroom, half_plane, top_bottom_lid — through the C ABI against the CPU oracle.

Parity bar: tetrahedron / top_bottom_lid / half_plane are binary32 and bit-exact.  torus / heart / double_mushroom take
double-precision powers (std::pow promotes, torus.hpp:74, heart.hpp:94, double_mushroom.hpp:86): the oracle calls glibc's
pow like the reference, the kernels multiply and take sqrt.  The doubles may differ in the last place, the binary32
results agree except with probability ~2^-29 per node, so the tests allow a handful of 1-ulp field differences but
require identical case codes and topology, and positions within 1e-5 of the box diagonal (north_star's tolerance).
"""
import numpy as np
import pytest

from implisolid_b200 import shapes

pytestmark = pytest.mark.gpu

N1 = [
    ("torus", shapes.n1_torus, 56, 0.7),
    ("heart", shapes.n1_heart, 60, 0.8),
    ("tetrahedron", shapes.n1_tetrahedron, 44, 0.9),
    ("mushroom", shapes.n1_mushroom, 52, 0.8),
    ("half_plane_cut", shapes.n1_half_plane_cut, 48, 0.8),
    ("lid_slab", shapes.n1_lid_slab, 40, 0.8),
    ("medley", shapes.n1_medley, 72, 0.9),
]
IDS = [c[0] for c in N1]


def _ulp_diff(a, b):
    ia, ib = a.view(np.int32).astype(np.int64), b.view(np.int32).astype(np.int64)
    ia = np.where(ia < 0, -(ia & 0x7fffffff), ia)
    ib = np.where(ib < 0, -(ib & 0x7fffffff), ib)
    return np.abs(ia - ib)


@pytest.mark.parametrize("name,shape,R,half", N1, ids=IDS)
def test_field_cases_mesh(gpu_lib, oracle, name, shape, R, half):
    sj = shape()
    oracle.build_geometry(sj, shapes.mc_settings(R, half))
    ov, of, ofield, ocases = oracle.get_v(), oracle.get_f(), oracle.get_field(), oracle.get_case_codes()
    oracle.finish_geometry()
    gpu_lib.build_geometry(sj, shapes.mc_settings(R, half, debug_cases=1, prune=0))
    gv, gf, gfield, gcases = gpu_lib.get_v(), gpu_lib.get_f(), gpu_lib.get_field(), gpu_lib.get_case_codes()
    gpu_lib.finish_geometry()
    assert of.shape[0] > 100
    d = _ulp_diff(gfield.ravel(), ofield.ravel())
    assert d.max() <= 1 and (d != 0).sum() <= 4, "field: %d nodes differ, worst %d ulp" % ((d != 0).sum(), d.max())
    assert np.array_equal(gcases, ocases), "cube case codes differ"
    assert np.array_equal(gf, of), "faces (indices and order) differ"
    tol = 1e-5 * 2 * half * np.sqrt(3.0)
    assert np.abs(gv - ov).max() <= tol
    if (d != 0).sum() == 0:
        assert np.array_equal(gv.view(np.uint32), ov.view(np.uint32))


@pytest.mark.parametrize("name,shape,R,half", N1 + [("medley_r150", shapes.n1_medley, 150, 0.9), ("heart_r130", shapes.n1_heart, 130, 0.8),
                                                ("torus_r140", shapes.n1_torus, 140, 0.7), ("mushroom_r120", shapes.n1_mushroom, 120, 0.8)],
                         ids=IDS + ["medley_r150", "heart_r130", "torus_r140", "mushroom_r120"])
def test_pruning_and_specialisation_change_nothing(gpu_lib, name, shape, R, half):
    """The interval bounds of the new primitives (field_kernel.cuh program_bounds) must never discard a tile or a CSG child
    that matters: dense evaluation and the default pruned + specialised evaluation give the same mesh bit for bit."""
    sj = shape()
    gpu_lib.build_geometry(sj, shapes.mc_settings(R, half, prune=0, specialize=0, debug_cases=1))
    v0, f0, cases0 = gpu_lib.get_v(), gpu_lib.get_f(), gpu_lib.get_case_codes()
    n0 = gpu_lib.last_counters()["nodes_full_eval"]
    gpu_lib.finish_geometry()
    gpu_lib.build_geometry(sj, shapes.mc_settings(R, half, debug_cases=1))
    v1, f1, cases1 = gpu_lib.get_v(), gpu_lib.get_f(), gpu_lib.get_case_codes()
    n1 = gpu_lib.last_counters()["nodes_full_eval"]
    gpu_lib.finish_geometry()
    assert np.array_equal(cases0, cases1)
    assert np.array_equal(f0, f1) and np.array_equal(v0.view(np.uint32), v1.view(np.uint32))
    if R >= 100:
        assert n1 < 0.7 * n0, (n1, n0)


def test_direct_evaluation(gpu_lib, oracle):
    rng = np.random.default_rng(9)
    pts = np.concatenate([rng.uniform(-1, 1, (3000, 3)), rng.normal(0, 0.3, (1097, 3))]).astype(np.float32)
    for name, shape, _, _ in N1:
        sj = shape()
        gf_, of_ = gpu_lib.query_implicit_values(sj, pts), oracle.query_implicit_values(sj, pts)
        assert _ulp_diff(gf_, of_).max() <= 1 and (gf_.view(np.uint32) != of_.view(np.uint32)).sum() <= 2, name
        g_raw, o_raw = gpu_lib.query_normals(sj, pts, normalize_and_invert=False), oracle.query_normals(sj, pts, normalize_and_invert=False)
        assert np.allclose(g_raw, o_raw, rtol=1e-6, atol=1e-7), name
        assert (g_raw.view(np.uint32) != o_raw.view(np.uint32)).sum() <= 6, name


@pytest.mark.parametrize("name,shape,R,half", [N1[0], N1[1], N1[6]], ids=["torus", "heart", "medley"])
def test_ob02_on_new_primitives(gpu_lib, oracle, name, shape, R, half):
    sj = shape()
    mc = shapes.mc_settings(R, half, iters=1, c=1.0, projection=1, qem=1)
    gv, gf = gpu_lib.polygonize(sj, mc)
    ov, of = oracle.polygonize(sj, mc)
    assert np.array_equal(gf, of)
    # OB02 moves vertices by gradient bisection; 1e-5 of the box diagonal is north_star's bar
    assert np.abs(gv - ov).max() <= 1e-5 * 2 * half * np.sqrt(3.0)
