"""GPU parity at BASELINE.json's sizes: the configs' own resolutions where the oracle (threaded build, same bits as
the parity build: tests/test_oracle_threads.py) finishes in seconds, bounded windows of the grid where it does not.

  C2  cube minus cylinder, 256^3: field eval + marching cubes, BOTH cube variants (analytic six planes / HEAD's SDF-table
      lookup, cube.hpp:215-270, with a synthetic table), faces and vertices bit for bit; with OB02 (default settings) within
      tolerance and with equal counters
  C3  64-centre metaballs, OB02 at 256^3 against the whole oracle; at 512^3 windows of the oracle's marching cubes bit for bit
  C4  32-primitive chain, OB02 at 192^3 against the whole oracle; at 1024^3 and 2048^3 (the edge cases that only appear at
      size: 32-bit offsets, > 256 tiles per row, the work-list read-back) windows of the oracle's marching cubes bit for bit,
      mesh sizes, closed 2-manifold
Tolerance: vertex positions within 1e-5 x the bounding-box diagonal (north_star)."""
import numpy as np
import pytest

import bench
from implisolid_b200 import shapes
from tests.oracle_lib import load_oracle_mt

pytestmark = pytest.mark.gpu
OB02 = dict(iters=1, c=1.0, projection=1, qem=1)


def _tol(half):
    return 1e-5 * 2 * half * np.sqrt(3.0)


def _closed_manifold(v, f):
    e = np.sort(np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]]), axis=1).astype(np.int64)
    _, cnt = np.unique(e[:, 0] * (v.shape[0] + 1) + e[:, 1], return_counts=True)
    return bool((cnt == 2).all())


def _box_sdf_table(n=40, h=0.05):
    """A synthetic stand-in for the rabbit table of cube.hpp:65-72: the signed distance to the box [-0.5, 0.5]^3 (positive
    outside, as the lookup negates it), sampled on n^3 points with spacing h from origin -n*h/2."""
    o = -0.5 * n * h
    c = o + h * np.arange(n, dtype=np.float64)
    z, y, x = np.meshgrid(c, c, c, indexing="ij")
    q = np.stack([np.abs(x) - 0.5, np.abs(y) - 0.5, np.abs(z) - 0.5])
    d = np.linalg.norm(np.maximum(q, 0.0), axis=0) + np.minimum(q.max(axis=0), 0.0)
    return d.astype(np.float32), h, (o, o, o)


@pytest.mark.parametrize("cube_mode", [0, 1], ids=["analytic_cube", "sdf_table_cube"])
def test_c2_256_both_cube_variants(gpu_lib, cube_mode):
    o = load_oracle_mt()
    sj = shapes.c2_cube_minus_cylinder()
    if cube_mode:
        table, h, origin = _box_sdf_table()
        gpu_lib.set_sdf_grid(table, h, origin)
        o.set_sdf_grid(table, h, origin)
    mc = shapes.mc_settings(256, 0.8, cube_mode=cube_mode)
    ov, of = o.polygonize(sj, mc)
    gv, gf = gpu_lib.polygonize(sj, mc)
    assert ov.shape[0] > 100000
    assert np.array_equal(gf, of)
    assert np.array_equal(gv.view(np.uint32), ov.view(np.uint32))
    # and the field itself on a smaller grid (dense, pruning off), bit for bit
    small = shapes.mc_settings(72, 0.8, cube_mode=cube_mode, prune=0)
    o.build_geometry(sj, small); ofield = o.get_field(); o.finish_geometry()
    gpu_lib.build_geometry(sj, small); gfield = gpu_lib.get_field(); gpu_lib.finish_geometry()
    assert np.array_equal(gfield.view(np.uint32), ofield.view(np.uint32))


@pytest.mark.parametrize("name,shape,R,half", [("c2_256", shapes.c2_cube_minus_cylinder, 256, 0.8), ("c3_256", shapes.c3_metaballs, 256, 1.2),
                                               ("c4_192", shapes.c4_deep_chain, 192, 1.0),
                                               # C3 perf variant (ii) of SURVEY §8(d): one node, 64 blobs summed ("balls" extension)
                                               ("c3_sum_blend_160", shapes.c3_sum_blend, 160, 1.2)])
def test_full_pipeline_default_settings_at_size(gpu_lib, name, shape, R, half):
    o = load_oracle_mt()
    sj, mc = shape(), shapes.mc_settings(R, half, **OB02)
    o.build_geometry(sj, mc)
    ov, of, ost, oedge = o.get_v(), o.get_f(), o.stats(), o.average_edge()
    o.finish_geometry()
    gpu_lib.build_geometry(sj, mc)
    gv, gf, gst, gedge = gpu_lib.get_v(), gpu_lib.get_f(), gpu_lib.last_counters(), gpu_lib.average_edge()
    gpu_lib.finish_geometry()
    assert np.array_equal(gf, of)
    assert np.float32(gedge).tobytes() == np.float32(oedge).tobytes()
    assert gst["projection_by_type"] == ost["projection_by_type"] and gst["projection_failed"] == ost["projection_failed"]
    assert gst["bisection_capped"] == ost["bisection_capped"]
    assert np.abs(gv - ov).max() <= _tol(half), np.abs(gv - ov).max()
    assert _closed_manifold(gv, gf)


@pytest.mark.parametrize("name,shape,R,half,nv,nf", [
    ("c3_512", shapes.c3_metaballs, 512, 1.2, 798962, 1597904),
    ("c4_1024", shapes.c4_deep_chain, 1024, 1.0, 2606308, 5212532),
    ("c5_2048", shapes.c4_deep_chain, 2048, 1.0, 10434182, 20868276),
])
def test_marching_cubes_windows_at_full_size(gpu_lib, name, shape, R, half, nv, nf):
    """Oracle windows of two cube layers at six heights of the full-size grid: the top layer's faces and their vertices must sit
    bit for bit, as one contiguous run, in the GPU's mesh of the whole grid (bench.match_window_to_mesh)."""
    o = load_oracle_mt()
    sj, mc = shape(), shapes.mc_settings(R, half)
    gv, gf = gpu_lib.polygonize(sj, mc)
    assert (gv.shape[0], gf.shape[0]) == (nv, nf)       # the sizes every earlier run of these configs has produced
    assert gf.shape[0] % 2 == 0 and gf.min() == 0 and gf.max() == nv - 1
    checked = 0
    for z0, z1 in bench.cpu_sample_windows(R, 6, 2):
        wv, wf = o.polygonize_layers(sj, mc, z0, z1)
        _, top = o.polygonize_layers(sj, mc, z1 - 1, z1)
        assert bench.match_window_to_mesh(np, wv, wf, top.shape[0], gv, gf), (z0, z1)
        checked += top.shape[0]
    assert checked > 1000
    if R <= 1024:
        assert _closed_manifold(gv, gf)
    else:       # 63 M edges: the Euler characteristic alone (V - E + F with E = 3F/2 for a closed surface)
        assert (gv.shape[0] - gf.shape[0] // 2) % 2 == 0
