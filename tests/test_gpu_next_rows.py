"""GPU parity for the SURVEY §8 (f) rows N2 (fused per-vertex normals), N3 (progressive results) and N4 (1-to-4 subdivision
with the rand() noise), through the C ABI against the CPU oracle."""
import json

import numpy as np
import pytest

from implisolid_b200 import shapes

pytestmark = pytest.mark.gpu


def _mc(R, half, subdiv=1, noise=0.0, **kw):
    mc = json.loads(shapes.mc_settings(R, half, **kw))
    mc["subdiv"] = {"enabled": subdiv}
    mc["debug"] = {"post_subdiv_noise": noise}
    return json.dumps(mc)


SUB = [("sphere", shapes.c1_sphere, 28, 0.55), ("c2", shapes.c2_cube_minus_cylinder, 64, 0.8), ("c4", shapes.c4_deep_chain, 72, 1.0),
       ("medley", shapes.n1_medley, 60, 0.9)]


@pytest.mark.parametrize("name,shape,R,half", SUB, ids=[c[0] for c in SUB])
def test_subdivision_matches_the_oracle_exactly(gpu_lib, oracle, name, shape, R, half):
    sj = shape()
    ov, of = oracle.polygonize(sj, _mc(R, half, noise=0.0))
    gv, gf = gpu_lib.polygonize(sj, _mc(R, half, noise=0.0))
    assert gf.shape == of.shape and gf.shape[0] % 4 == 0
    assert np.array_equal(gf, of), "subdivided faces (indices and order) differ"
    assert np.array_equal(gv.view(np.uint32), ov.view(np.uint32)), "midpoints differ"


def test_subdivision_noise_is_the_glibc_rand_stream(gpu_lib, oracle):
    sj = shapes.c2_cube_minus_cylinder()
    oracle.srand(1)
    gpu_lib.srand(1)
    for R in (20, 33):                   # the second build continues the stream where the first one stopped
        ov, of = oracle.polygonize(sj, _mc(R, 0.8, noise=0.01))
        gv, gf = gpu_lib.polygonize(sj, _mc(R, 0.8, noise=0.01))
        assert np.array_equal(gf, of)
        assert np.array_equal(gv.view(np.uint32), ov.view(np.uint32))
    # a build with zero amplitude still consumes its draws (rand() is called for every coordinate)
    ov, _ = oracle.polygonize(sj, _mc(16, 0.8, noise=0.0))
    gv, _ = gpu_lib.polygonize(sj, _mc(16, 0.8, noise=0.0))
    ov, _ = oracle.polygonize(sj, _mc(20, 0.8, noise=0.02))
    gv, _ = gpu_lib.polygonize(sj, _mc(20, 0.8, noise=0.02))
    assert np.array_equal(gv.view(np.uint32), ov.view(np.uint32))


def test_full_default_pipeline_with_subdivision(gpu_lib, oracle):
    """MC -> 2 x (resample, project + QEM) -> subdivision on the last repeat only (mcc2.cpp:396-401)."""
    sj = shapes.c2_cube_minus_cylinder()
    mc = _mc(40, 0.8, noise=0.0, iters=1, c=1.0, projection=1, qem=1, repeats=2)
    ov, of = oracle.polygonize(sj, mc)
    gv, gf = gpu_lib.polygonize(sj, mc)
    assert np.array_equal(gf, of)
    assert np.abs(gv - ov).max() <= 1e-5 * 1.6 * np.sqrt(3.0)


def test_progress_reports_match_the_oracle(gpu_lib, oracle):
    sj = shapes.c2_cube_minus_cylinder()
    mc = _mc(48, 0.8, subdiv=0, iters=2, c=1.0, projection=1, qem=1)
    specs = '{"progressCallback_id": 11, "call_id": 4, "shape_id": 9}'
    o_rep, (ov, of) = oracle.polygonize_progressive(sj, mc, call_specs=specs)
    g_rep, (gv, gf) = gpu_lib.polygonize_progressive(sj, mc, call_specs=specs)
    assert len(g_rep) == len(o_rep) == 3
    tol = 1e-5 * 1.6 * np.sqrt(3.0)
    for (a_v, a_f, a_ids), (b_v, b_f, b_ids) in zip(g_rep, o_rep):
        assert a_ids == b_ids == (11, 9, 4)
        assert np.array_equal(a_f, b_f)
        assert np.abs(a_v - b_v).max() <= tol
    assert np.array_equal(g_rep[0][0].view(np.uint32), o_rep[0][0].view(np.uint32))      # the marching-cubes mesh is bit-exact
    assert not np.array_equal(g_rep[0][0], g_rep[1][0]) and not np.array_equal(g_rep[1][0], g_rep[2][0])
    assert np.array_equal(g_rep[2][0], gv) and np.array_equal(gf, of)
    # without a callback nothing is reported and the result is the same
    v2, f2 = gpu_lib.polygonize(sj, mc)
    assert np.array_equal(v2, gv) and np.array_equal(f2, gf)


@pytest.mark.parametrize("eager", [0, 1])
def test_fused_normals_equal_the_evaluation_service(gpu_lib, oracle, eager):
    """isd_get_n == set_object / set_x(final vertices) / calculate_implicit_gradients(true), the flow of
    implisolid_main.js:551-600, on the oracle — with no 50 000-point cap."""
    for shape, R, half in ((shapes.c2_cube_minus_cylinder, 150, 0.8), (shapes.c3_metaballs, 48, 1.2), (shapes.n1_heart, 40, 0.8)):
        sj = shape()
        kw = dict(normals=1) if eager else {}
        gpu_lib.build_geometry(sj, shapes.mc_settings(R, half, iters=1, c=1.0, projection=1, qem=1, **kw))
        gv, gn = gpu_lib.get_v(), gpu_lib.get_n()
        assert np.array_equal(gpu_lib.get_n_view().view(np.uint32), gn.view(np.uint32))
        gpu_lib.finish_geometry()
        assert gn.shape == gv.shape
        if shape is shapes.c2_cube_minus_cylinder:
            assert gv.shape[0] > 50000                                  # beyond set_x's cap (mcc2.cpp:770-773)
        on = np.concatenate([oracle.query_normals(sj, gv[i:i + 40000]) for i in range(0, gv.shape[0], 40000)])
        assert np.array_equal(gn.view(np.uint32), on.view(np.uint32))
        if shape is shapes.c2_cube_minus_cylinder:      # elsewhere |grad f| <= 1e-4 is common and takes the reference's -42 branch
            assert (np.abs(np.linalg.norm(gn, axis=1) - 1.0) < 1e-4).mean() > 0.99


def test_normals_survive_the_evaluation_service(gpu_lib, oracle):
    """The direct-evaluation service uploads its own shape program; lazy normals must re-upload the build's."""
    sj, other = shapes.c2_cube_minus_cylinder(), shapes.c1_sphere()
    gpu_lib.build_geometry(sj, shapes.mc_settings(40, 0.8))
    gv = gpu_lib.get_v()
    pts = np.random.default_rng(1).uniform(-1, 1, (100, 3)).astype(np.float32)
    gpu_lib.query_implicit_values(other, pts)
    gn = gpu_lib.get_n()
    gpu_lib.finish_geometry()
    assert np.array_equal(gn.view(np.uint32), oracle.query_normals(sj, gv).view(np.uint32))
