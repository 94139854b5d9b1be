"""GPU parity: OB02 stages (adjacency, vertex resampling, centroid projection + bisection, QEM) vs the CPU oracle.

Tolerance (north_star): vertex positions within 1e-5 of the bounding-box diagonal.  Every oracle comparison runs in BOTH
mean-edge modes: the product's default (64-bit fixed-point sum, DESIGN.md §3.3 — the mode bench.py times; the oracle's
`compute_average_edge_length_fixed_point` restates it and the two must agree on the mean bit for bit) and
"gpu":{"exact_reduce":1} (the reference's serial float sum, centroids_projection.cpp:70-82).
"""
import numpy as np
import pytest

from implisolid_b200 import shapes
from tests.oracle_lib import load_oracle

pytestmark = pytest.mark.gpu

OCTA_F = np.array([[0, 2, 4], [2, 1, 4], [1, 3, 4], [3, 0, 4], [2, 0, 5], [1, 2, 5], [3, 1, 5], [0, 3, 5]], dtype=np.int32)
OCTA_V = np.array([[1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1]], dtype=np.float32) * 0.45


def _tol(half):
    return 1e-5 * 2 * half * np.sqrt(3.0)


def test_adjacency_octahedron_and_mc_mesh(gpu_lib, oracle):
    off, items = gpu_lib.umbrella(OCTA_F, 6)
    ooff, oitems = oracle.umbrella(OCTA_F, 6)
    assert np.array_equal(off, ooff) and np.array_equal(items, oitems)
    # known answer from mesh_algorithms.hpp:155-175 style: every vertex of an octahedron has 4 faces, ascending
    assert (np.diff(off) == 4).all()
    assert np.array_equal(gpu_lib.faces_of_faces(OCTA_F, 6), oracle.faces_of_faces(OCTA_F))
    v, f = oracle.polygonize(shapes.c2_cube_minus_cylinder(), shapes.mc_settings(40, 0.8))
    off, items = gpu_lib.umbrella(f, v.shape[0])
    ooff, oitems = oracle.umbrella(f, v.shape[0])
    assert np.array_equal(off, ooff) and np.array_equal(items, oitems)
    assert np.array_equal(gpu_lib.faces_of_faces(f, v.shape[0]), oracle.faces_of_faces(f))


@pytest.mark.parametrize("name,shape,R,half", [
    ("sphere", shapes.c1_sphere, 28, 0.55),
    ("cube_cyl", shapes.c2_cube_minus_cylinder, 48, 0.8),
    ("metaballs", shapes.c3_metaballs, 40, 1.2),
    ("chain", shapes.c4_deep_chain, 56, 1.0),
    ("bushy", shapes.bushy_tree, 44, 0.9),
    ("short_alpha_list", shapes.tiny_sphere, 12, 0.02),      # mean edge 0.0029: six alphas
    ("empty_alpha_list", shapes.tiny_sphere, 40, 0.02),      # mean edge 0.0009: none, every centroid stays where it is
])
@pytest.mark.parametrize("mode", [{}, {"exact_reduce": 1}], ids=["default_fixed_point_sum", "serial_float_sum"])
def test_full_pipeline_vs_oracle(gpu_lib, oracle, name, shape, R, half, mode):
    sj = shape()
    mc = shapes.mc_settings(R, half, iters=1, c=1.0, projection=1, qem=1, **mode)
    oracle.build_geometry(sj, mc)
    ov, of = oracle.get_v(), oracle.get_f()
    oc, on = oracle.projected_centroids()
    ost = oracle.stats()
    oedge = oracle.average_edge()
    oracle.finish_geometry()
    gpu_lib.build_geometry(sj, mc)
    gv, gf = gpu_lib.get_v(), gpu_lib.get_f()
    gc, gn = gpu_lib.projected_centroids()
    gst = gpu_lib.last_counters()
    gedge = gpu_lib.average_edge()
    gpu_lib.finish_geometry()
    assert np.array_equal(gf, of)
    if mode:
        assert abs(gedge - oedge) <= 1e-6 * oedge
    else:
        assert np.float32(gedge).tobytes() == np.float32(oedge).tobytes(), (gedge, oedge)      # integer sum: the same bits
    assert gst["projection_by_type"] == ost["projection_by_type"], (gst, ost)
    assert gst["projection_failed"] == ost["projection_failed"]
    assert gst["bisection_capped"] == ost["bisection_capped"]
    tol = _tol(half)
    assert np.abs(gc - oc).max() <= tol, np.abs(gc - oc).max()
    ok = np.isfinite(on).all(axis=1) & np.isfinite(gn).all(axis=1)
    assert np.abs(gn[ok] - on[ok]).max() <= 1e-4
    assert np.abs(gv - ov).max() <= tol, np.abs(gv - ov).max()


@pytest.mark.parametrize("mode", [{}, {"exact_reduce": 1}], ids=["default_fixed_point_sum", "serial_float_sum"])
def test_resampling_only_and_projection_without_qem(gpu_lib, oracle, mode):
    sj = shapes.c2_cube_minus_cylinder()
    for kw in (dict(iters=2, c=0.4), dict(iters=0, projection=1, qem=0), dict(iters=1, c=1.0, projection=1, qem=1, repeats=2)):
        mc = shapes.mc_settings(36, 0.8, **mode, **kw)
        ov, of = oracle.polygonize(sj, mc)
        gv, gf = gpu_lib.polygonize(sj, mc)
        assert np.array_equal(gf, of)
        assert np.abs(gv - ov).max() <= _tol(0.8), (kw, np.abs(gv - ov).max())
    # projection without QEM leaves the mesh untouched (centroids_projection.cpp:1304-1310)
    mv, _ = gpu_lib.polygonize(sj, shapes.mc_settings(36, 0.8))
    pv, _ = gpu_lib.polygonize(sj, shapes.mc_settings(36, 0.8, projection=1, qem=0))
    assert np.array_equal(mv, pv)


def test_ob02_on_fixture_mesh(gpu_lib, oracle):
    sj = shapes.c1_sphere()
    got = gpu_lib.ob02_on_mesh(sj, OCTA_V, OCTA_F, resample_iters=1, c=1.0, projection=1, qem=1)
    want = oracle.ob02_on_mesh(sj, OCTA_V, OCTA_F, resample_iters=1, c=1.0, projection=1, qem=1)
    assert np.abs(got - want).max() <= 1e-5


def test_projected_centroids_lie_on_surface(gpu_lib):
    sj = shapes.c3_metaballs()
    gpu_lib.build_geometry(sj, shapes.mc_settings(96, 1.2, projection=1, qem=1))
    c, _ = gpu_lib.projected_centroids()
    st = gpu_lib.last_counters()
    gpu_lib.finish_geometry()
    f = gpu_lib.eval_points(sj, c)
    resolved = sum(st["projection_by_type"])
    assert resolved + st["projection_failed"] == c.shape[0]
    # |f| <= ROOT_TOLERANCE for every centroid that found the surface and did not hit the bisection cap
    assert (np.abs(f) > 1.0001e-4).sum() <= st["projection_failed"] + st["bisection_capped"]


@pytest.mark.parametrize("shape,R,half", [(shapes.c4_deep_chain, 96, 1.0), (shapes.bushy_tree, 80, 0.9), (shapes.c3_metaballs, 64, 1.2),
                                          (shapes.n1_medley, 72, 0.9), (shapes.c2_cube_minus_cylinder, 64, 0.8), (shapes.demo_moon, 64, 0.6)],
                         ids=["c4", "bushy", "c3", "medley", "c2", "moon"])
def test_point_masks_change_nothing(gpu_lib, shape, R, half):
    """OB02 evaluations near the surface run per-tile instruction lists made from K1's liveness masks (gpu.point_spec) and so
    skip CSG children that can never be selected there: vertices, projected centroids and their normals must keep every bit."""
    sj = shape()
    out = []
    for ps in (0, 1):
        gpu_lib.build_geometry(sj, shapes.mc_settings(R, half, iters=2, c=1.0, projection=1, qem=1, point_spec=ps))
        v, f = gpu_lib.get_v(), gpu_lib.get_f()
        c, n = gpu_lib.projected_centroids()
        gpu_lib.finish_geometry()
        out.append((v, f, c, n))
    (v0, f0, c0, n0), (v1, f1, c1, n1) = out
    assert np.array_equal(f0, f1)
    assert np.array_equal(v0.view(np.uint32), v1.view(np.uint32))
    assert np.array_equal(c0.view(np.uint32), c1.view(np.uint32)) and np.array_equal(n0.view(np.uint32), n1.view(np.uint32))


@pytest.mark.parametrize("mode", [{}, {"exact_reduce": 1}], ids=["default_fixed_point_sum", "serial_float_sum"])
@pytest.mark.parametrize("seed", [1, 5, 9, 15])
def test_random_trees_full_pipeline_vs_oracle(gpu_lib, oracle, seed, mode):
    """Random CSG trees at R=40 (seed 9 resolves centroids by every one of the seven direction types, thousands fail): faces
    and the per-type projection counts exact, vertices within tolerance; per-tile point lists on and off give the same bits."""
    sj = shapes.random_tree(seed)
    kw = dict(iters=1, c=1.0, projection=1, qem=1, **mode)
    oracle.build_geometry(sj, shapes.mc_settings(40, 1.0, **kw))
    ov, of, ost = oracle.get_v(), oracle.get_f(), oracle.stats()
    oracle.finish_geometry()
    out = []
    for ps in (1, 0):
        gpu_lib.build_geometry(sj, shapes.mc_settings(40, 1.0, point_spec=ps, **kw))
        out.append((gpu_lib.get_v(), gpu_lib.get_f(), gpu_lib.last_counters()))
        gpu_lib.finish_geometry()
    (gv, gf, gst), (gv0, gf0, _) = out
    assert np.array_equal(gf, of) and np.array_equal(gf0, of)
    assert gst["projection_by_type"] == ost["projection_by_type"], (gst, ost)
    assert gst["projection_failed"] == ost["projection_failed"]
    assert np.abs(gv - ov).max() <= _tol(1.0), np.abs(gv - ov).max()
    assert np.array_equal(gv.view(np.uint32), gv0.view(np.uint32))


@pytest.mark.parametrize("shape,R,half", [(shapes.c4_deep_chain, 96, 1.0), (shapes.bushy_tree, 80, 0.9), (shapes.c3_metaballs, 64, 1.2),
                                          (shapes.n1_medley, 72, 0.9), (shapes.c2_cube_minus_cylinder, 64, 0.8), (lambda: shapes.random_tree(9), 48, 1.0)],
                         ids=["c4", "bushy", "c3", "medley", "c2", "random9"])
def test_pair_kernels_change_nothing(gpu_lib, shape, R, half):
    """With per-tile point lists the OB02 kernels take two centroids per thread through the packed f32x2 interpreters
    ("gpu":{"pair":1}, the default): vertices, projected centroids, their normals and every counter must keep the bits of the
    one-centroid-per-thread kernels (pair 0), over two OB02 repeats."""
    sj = shape()
    out = []
    # perm: faces taken grouped by tile / in face order; ss: the search's trial points in their own lane-refill kernel / in lockstep
    for pair, split, perm, ss in ((0, 0, 1, 1), (1, 0, 1, 1), (1, 1, 1, 0), (1, 2, 1, 1), (1, 2, 0, 1), (1, 0, 0, 0), (1, 2, 0, 0)):
        gpu_lib.build_geometry(sj, shapes.mc_settings(R, half, iters=2, c=1.0, projection=1, qem=1, repeats=2, point_spec=1, pair=pair, finish_split=split, perm=perm,
                                                      search_split=ss))
        v, f = gpu_lib.get_v(), gpu_lib.get_f()
        c, n = gpu_lib.projected_centroids()
        st = gpu_lib.last_counters()
        gpu_lib.finish_geometry()
        out.append((v, f, c, n, st))
    v0, f0, c0, n0, s0 = out[0]
    for v1, f1, c1, n1, s1 in out[1:]:
        assert np.array_equal(f0, f1)
        assert np.array_equal(v0.view(np.uint32), v1.view(np.uint32))
        assert np.array_equal(c0.view(np.uint32), c1.view(np.uint32)) and np.array_equal(n0.view(np.uint32), n1.view(np.uint32))
        for k in ("projection_by_type", "projection_failed", "bisection_capped", "bisected"):
            assert s0[k] == s1[k], (k, s0[k], s1[k])


@pytest.mark.parametrize("shape,R,half", [(shapes.c4_deep_chain, 96, 1.0), (shapes.bushy_tree, 80, 0.9), (shapes.c3_metaballs, 64, 1.2),
                                          (shapes.n1_medley, 72, 0.9), (shapes.c2_cube_minus_cylinder, 100, 0.8), (shapes.demo_moon, 90, 0.6),
                                          (lambda: shapes.random_tree(9), 48, 1.0), (shapes.c1_sphere, 21, 0.55)],
                         ids=["c4", "bushy", "c3", "medley", "c2", "moon", "random9", "sphere"])
def test_mc_adjacency_equals_generic(gpu_lib, shape, R, half):
    """Umbrellas and faces-of-faces straight from the cube case codes (csrc/mc_adjacency.cuh, the default for the mesh K4 emits)
    against the generic kernels that build them from the face list ("gpu":{"mc_adjacency":0}) and against the oracle's tables:
    identical arrays, hence identical OB02 results bit for bit."""
    sj = shape()
    out = []
    for topo in (1, 0):
        gpu_lib.build_geometry(sj, shapes.mc_settings(R, half, iters=1, c=1.0, projection=1, qem=1, mc_adjacency=topo))
        v, f = gpu_lib.get_v(), gpu_lib.get_f()
        off, items, fof = gpu_lib.debug_adjacency(v.shape[0], f.shape[0])
        gpu_lib.finish_geometry()
        out.append((v, f, off, items, fof))
    (v1, f1, off1, it1, fof1), (v0, f0, off0, it0, fof0) = out
    assert np.array_equal(f1, f0)
    assert np.array_equal(off1, off0) and np.array_equal(it1, it0)
    assert np.array_equal(fof1, fof0)
    assert np.array_equal(v1.view(np.uint32), v0.view(np.uint32))
    o = load_oracle()
    ooff, oitems = o.umbrella(f1, v1.shape[0])
    assert np.array_equal(off1, ooff) and np.array_equal(it1, oitems)
    assert np.array_equal(fof1.reshape(-1, 3), o.faces_of_faces(f1))


@pytest.mark.parametrize("world", [2, 5])
def test_mc_adjacency_equals_generic_on_halo_slabs(gpu_lib, world):
    """The same on emulated z-slabs in halo mode, where the local mesh has a ghost layer (vertices, no faces) below and ends at
    the slab's top: every rank's exported slab keeps its bits."""
    sj = shapes.c4_deep_chain()
    L = gpu_lib.lib
    try:
        for r in range(world):
            res = []
            for topo in (1, 0):
                L.isd_dist_emulate(r, world)
                L.isd_dist_set_offsets(0, 0)
                L.isd_dist_set_edge_total(3 << 40, 40000)
                gpu_lib.build_geometry(sj, shapes.mc_settings(72, 1.0, iters=1, c=1.0, projection=1, qem=1, mc_adjacency=topo))
                res.append((gpu_lib.get_v(), gpu_lib.get_f()))
                gpu_lib.finish_geometry()
            assert np.array_equal(res[0][1], res[1][1])
            assert np.array_equal(res[0][0].view(np.uint32), res[1][0].view(np.uint32))
    finally:
        L.isd_dist_emulate(0, 1)
