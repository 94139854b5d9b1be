"""GPU parity: field evaluation + marching cubes through the C ABI vs the CPU oracle, same JSON inputs.

Bar (north_star): cube case codes, per-cube triangle counts and face topology bit-exact; here additionally the field,
the vertex positions and the vertex/face ORDER are bit-exact, because the kernels reproduce the reference's scan order.
"""
import json

import numpy as np
import pytest

from implisolid_b200 import shapes

pytestmark = pytest.mark.gpu

CASES = [
    ("c1_sphere", shapes.c1_sphere, 28, 0.55),
    ("c1_sphere_r3", shapes.c1_sphere, 3, 0.55),
    ("c1_sphere_r31", shapes.c1_sphere, 31, 0.6),
    ("c1_sphere_r33", shapes.c1_sphere, 33, 0.6),
    ("c2_cube_cyl", shapes.c2_cube_minus_cylinder, 64, 0.8),
    ("c2_cube_cyl_r100", shapes.c2_cube_minus_cylinder, 100, 0.8),
    ("c3_metaballs", shapes.c3_metaballs, 48, 1.2),
    ("c4_chain", shapes.c4_deep_chain, 72, 1.0),
    ("moon", shapes.demo_moon, 40, 1.0),
    ("cone", shapes.simple_cone, 28, 0.5),
    ("bushy", shapes.bushy_tree, 56, 0.9),
]


def _mc(resolution, half, **gpu):
    return shapes.mc_settings(resolution, half, **gpu)


@pytest.mark.parametrize("name,shape,R,half", CASES, ids=[c[0] for c in CASES])
def test_field_cases_mesh_bit_exact(gpu_lib, oracle, name, shape, R, half):
    sj = shape()
    oracle.build_geometry(sj, _mc(R, half))
    ov, of = oracle.get_v(), oracle.get_f()
    ofield, ocases = oracle.get_field(), oracle.get_case_codes()
    oracle.finish_geometry()

    gpu_lib.build_geometry(sj, _mc(R, half, debug_cases=1, prune=0))
    gv, gf = gpu_lib.get_v(), gpu_lib.get_f()
    gfield, gcases = gpu_lib.get_field(), gpu_lib.get_case_codes()
    # borrowed views agree with the copies (get_v_ptr / get_f_ptr)
    assert np.array_equal(gpu_lib.get_v_view(), gv) and np.array_equal(gpu_lib.get_f_view(), gf)
    gpu_lib.finish_geometry()

    assert gfield.shape == ofield.shape
    assert np.array_equal(gfield.view(np.uint32), ofield.view(np.uint32)), "field differs from the oracle bit pattern"
    assert np.array_equal(gcases, ocases), "cube case codes differ"
    assert gf.shape == of.shape and gv.shape == ov.shape
    assert np.array_equal(gf, of), "faces (indices and order) differ"
    assert np.array_equal(gv.view(np.uint32), ov.view(np.uint32)), "vertex positions differ"
    if name.startswith("c1_sphere") and R >= 28:
        assert abs(np.linalg.norm(gv, axis=1) - 0.5).max() < 2e-3


@pytest.mark.parametrize("name,shape,R,half", CASES + [("c2_r200", shapes.c2_cube_minus_cylinder, 200, 0.8), ("c4_r160", shapes.c4_deep_chain, 160, 1.0),
                                                  ("c3_r120", shapes.c3_metaballs, 120, 1.2), ("moon_r130", shapes.demo_moon, 130, 1.0),
                                                  ("c3_sum_blend_r120", shapes.c3_sum_blend, 120, 1.2)],
                         ids=[c[0] for c in CASES] + ["c2_r200", "c4_r160", "c3_r120", "moon_r130", "c3_sum_blend_r120"])
def test_tile_pruning_changes_nothing(gpu_lib, name, shape, R, half):
    """Interval pruning (the default) must give the same mesh, bit for bit, as evaluating every node; and every tile
    it does evaluate holds the same field bits."""
    sj = shape()
    gpu_lib.build_geometry(sj, _mc(R, half, prune=0, debug_cases=1))
    v0, f0, field0, cases0 = gpu_lib.get_v(), gpu_lib.get_f(), gpu_lib.get_field(), gpu_lib.get_case_codes()
    n0 = gpu_lib.last_counters()["nodes_full_eval"]
    gpu_lib.finish_geometry()
    gpu_lib.build_geometry(sj, _mc(R, half, prune=1, debug_cases=1))
    v1, f1, field1, cases1 = gpu_lib.get_v(), gpu_lib.get_f(), gpu_lib.get_field(), gpu_lib.get_case_codes()
    n1 = gpu_lib.last_counters()["nodes_full_eval"]
    gpu_lib.finish_geometry()
    assert np.array_equal(f0, f1) and np.array_equal(v0.view(np.uint32), v1.view(np.uint32))
    assert np.array_equal(cases0, cases1)
    assert n1 <= n0
    if R >= 100:
        assert n1 < 0.6 * n0, (n1, n0)
    # field values next to the surface (every node of a cut cube) are present and identical
    S = field0.shape[0]
    cut = (cases0 != 0) & (cases0 != 255)
    zz, yy, xx = np.nonzero(cut)
    for dz in (0, 1):
        for dy in (0, 1):
            for dx in (0, 1):
                a = field0[zz + 1 + dz, yy + 1 + dy, xx + 1 + dx]
                b = field1[zz + 1 + dz, yy + 1 + dy, xx + 1 + dx]
                assert np.array_equal(a.view(np.uint32), b.view(np.uint32))


SPEC_CASES = [c for c in CASES if c[0] in ("c2_cube_cyl_r100", "c3_metaballs", "c4_chain", "moon", "bushy")] + [
    ("c4_r160", shapes.c4_deep_chain, 160, 1.0), ("c3_r120", shapes.c3_metaballs, 120, 1.2), ("bushy_r150", shapes.bushy_tree, 150, 0.9),
    ("c2_r200", shapes.c2_cube_minus_cylinder, 200, 0.8), ("c4_r301", shapes.c4_deep_chain, 301, 1.0)]


@pytest.mark.parametrize("name,shape,R,half", SPEC_CASES, ids=[c[0] for c in SPEC_CASES])
def test_tile_specialisation_changes_nothing(gpu_lib, name, shape, R, half):
    """Per-tile liveness masks (the default) skip CSG children that interval bounds prove are never selected in a tile.
    The field must keep the same bits in every node of every tile that holds a cut cube, and the mesh must be identical."""
    sj = shape()
    gpu_lib.build_geometry(sj, _mc(R, half, specialize=0, debug_cases=1))
    v0, f0, field0, cases0 = gpu_lib.get_v(), gpu_lib.get_f(), gpu_lib.get_field(), gpu_lib.get_case_codes()
    gpu_lib.finish_geometry()
    gpu_lib.build_geometry(sj, _mc(R, half, specialize=1, debug_cases=1))
    v1, f1, field1, cases1 = gpu_lib.get_v(), gpu_lib.get_f(), gpu_lib.get_field(), gpu_lib.get_case_codes()
    gpu_lib.finish_geometry()
    assert np.array_equal(cases0, cases1)
    assert np.array_equal(f0, f1) and np.array_equal(v0.view(np.uint32), v1.view(np.uint32))
    S = field0.shape[0]
    T = (S + 7) // 8
    tiles = np.zeros((T, T, T), dtype=bool)
    cut = (cases0 != 0) & (cases0 != 255)
    zz, yy, xx = np.nonzero(cut)
    for dz in (0, 1):
        for dy in (0, 1):
            for dx in (0, 1):
                tiles[(zz + 1 + dz) // 8, (yy + 1 + dy) // 8, (xx + 1 + dx) // 8] = True
    node_mask = np.repeat(np.repeat(np.repeat(tiles, 8, axis=0), 8, axis=1), 8, axis=2)[:S, :S, :S]
    assert node_mask.sum() > 0 or f0.shape[0] == 0
    a, b = field0[node_mask], field1[node_mask]
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), "specialised tiles differ in %d nodes" % int((a.view(np.uint32) != b.view(np.uint32)).sum())


def test_mesh_is_closed_manifold(gpu_lib):
    v, f = gpu_lib.polygonize(shapes.c2_cube_minus_cylinder(), _mc(96, 0.8))
    e = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]])
    e.sort(axis=1)
    key = e[:, 0].astype(np.int64) * v.shape[0] + e[:, 1]
    _, counts = np.unique(key, return_counts=True)
    assert (counts == 2).all()                      # every edge shared by exactly two faces
    assert f.shape[0] % 2 == 0                      # mesh_algorithms.hpp:59
    assert v.shape[0] - len(counts) + f.shape[0] == 2 * 2 - 2 * 1 or True   # genus depends on the shape; Euler checked for the sphere below


def test_sphere_euler_characteristic(gpu_lib):
    v, f = gpu_lib.polygonize(shapes.c1_sphere(), _mc(64, 0.6))
    e = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]])
    e.sort(axis=1)
    n_edges = len(np.unique(e[:, 0].astype(np.int64) * v.shape[0] + e[:, 1]))
    assert v.shape[0] - n_edges + f.shape[0] == 2


def test_empty_and_state_machine(gpu_lib, capfd):
    # a sphere outside the box: no geometry; getters return empty arrays
    far = json.dumps({"type": "ellipsoid", "matrix": [1, 0, 0, 10, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1]})
    gpu_lib.build_geometry(far, _mc(16, 0.5))
    assert gpu_lib.get_v_size() == 0 and gpu_lib.get_f_size() == 0
    assert gpu_lib.get_v().shape == (0, 3)
    # second build without finish_geometry is ignored (mcc2.cpp:316-319)
    gpu_lib.build_geometry(shapes.c1_sphere(), _mc(16, 0.55))
    assert gpu_lib.get_v_size() == 0
    gpu_lib.finish_geometry()
    gpu_lib.build_geometry(shapes.c1_sphere(), _mc(16, 0.55))
    assert gpu_lib.get_v_size() > 0
    gpu_lib.finish_geometry()


def test_ignore_root_matrix(gpu_lib, oracle):
    sj = shapes.c2_cube_minus_cylinder()
    mc = json.loads(_mc(40, 0.8))
    mc["ignore_root_matrix"] = 1
    ov, of = oracle.polygonize(sj, json.dumps(mc))
    gv, gf = gpu_lib.polygonize(sj, json.dumps(mc))
    assert np.array_equal(gf, of) and np.array_equal(gv.view(np.uint32), ov.view(np.uint32))


def test_anisotropic_box(gpu_lib, oracle):
    mc = json.dumps({"resolution": 37, "box": {"xmin": -0.7, "xmax": 0.9, "ymin": -0.6, "ymax": 0.6, "zmin": -0.55, "zmax": 0.8},
                     "subdiv": {"enabled": 0}, "overall_repeats": 0})
    sj = shapes.bushy_tree()
    ov, of = oracle.polygonize(sj, mc)
    gv, gf = gpu_lib.polygonize(sj, mc)
    assert np.array_equal(gf, of) and np.array_equal(gv.view(np.uint32), ov.view(np.uint32))


def test_direct_evaluation_service(gpu_lib, oracle):
    rng = np.random.default_rng(5)
    pts = rng.uniform(-1, 1, (4097, 3)).astype(np.float32)
    for shape in (shapes.c2_cube_minus_cylinder, shapes.c3_metaballs, shapes.c4_deep_chain, shapes.bushy_tree, shapes.simple_cone):
        sj = shape()
        assert np.array_equal(gpu_lib.query_implicit_values(sj, pts).view(np.uint32), oracle.query_implicit_values(sj, pts).view(np.uint32))
        g_raw, o_raw = gpu_lib.query_normals(sj, pts, normalize_and_invert=False), oracle.query_normals(sj, pts, normalize_and_invert=False)
        assert np.array_equal(g_raw.view(np.uint32), o_raw.view(np.uint32))
        g_n, o_n = gpu_lib.query_normals(sj, pts), oracle.query_normals(sj, pts)
        assert np.array_equal(g_n.view(np.uint32), o_n.view(np.uint32))
    # protocol errors (mcc2.cpp:726-812)
    assert gpu_lib.lib.set_x(pts.ctypes.data, 50000) is False
    f, g = gpu_lib.eval_points(shapes.c4_deep_chain(), rng.uniform(-1, 1, (70001, 3)).astype(np.float32), gradients=True)
    assert f.shape == (70001,) and np.isfinite(f).all() and g.shape == (70001, 3)
