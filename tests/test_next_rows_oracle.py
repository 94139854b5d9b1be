"""CPU: oracle restatement of the SURVEY §8 (f) rows N3 (progressive results) and N4 (1-to-4 subdivision + rand() noise),
pinned to what the reference source states: subdiv_1to4.hpp's face layout diagram (:305-316), the midpoint rule (:198-214),
the first-appearance numbering of the std::map walk (:118-160), libc's rand() itself, and grand_algorithm's three report
points (mcc2.cpp:351,372,390)."""
import ctypes
import json

import numpy as np

from implisolid_b200 import shapes


def _mc(R, half, subdiv=1, noise=0.0, **kw):
    mc = json.loads(shapes.mc_settings(R, half, **kw))
    mc["subdiv"] = {"enabled": subdiv}
    mc["debug"] = {"post_subdiv_noise": noise}
    return json.dumps(mc)


def test_private_rand_is_glibc_rand(oracle):
    oracle.lib.oracle_rand_check.restype = ctypes.c_int
    for seed in (1, 2, 12345, 0):
        assert oracle.lib.oracle_rand_check(seed, 50000) == 0


def test_subdivision_of_one_triangle_follows_the_diagram(oracle):
    v = np.array([[0, 0, 0], [2, 0, 0], [0, 4, 0]], dtype=np.float32)
    f = np.array([[0, 1, 2]], dtype=np.int32)
    vo, fo, nvo = np.zeros((6, 3), np.float32), np.zeros((4, 3), np.int32), ctypes.c_int()
    oracle.lib.oracle_subdivide(v.ctypes.data_as(ctypes.c_void_p), 3, f.ctypes.data_as(ctypes.c_void_p), 1, vo.ctypes.data_as(ctypes.c_void_p),
                                fo.ctypes.data_as(ctypes.c_void_p), ctypes.byref(nvo))
    assert nvo.value == 6
    # midpoints are created in slot order e01, e12, e20 -> m01 = 3, m12 = 4, m20 = 5
    assert np.array_equal(vo[3:], np.array([[1, 0, 0], [1, 2, 0], [0, 2, 0]], dtype=np.float32))
    # subdiv_1to4.hpp:262-318: the old face becomes (m12, m20, m01); (v0,m01,m20), (v1,m12,m01), (v2,m20,m12) are appended
    assert fo.tolist() == [[4, 5, 3], [0, 3, 5], [1, 4, 3], [2, 5, 4]]


def test_subdivision_invariants_on_a_mesh(oracle):
    v0, f0 = oracle.polygonize(shapes.c2_cube_minus_cylinder(), _mc(24, 0.8, subdiv=0))
    v1, f1 = oracle.polygonize(shapes.c2_cube_minus_cylinder(), _mc(24, 0.8, subdiv=1, noise=0.0))
    nv, nf = v0.shape[0], f0.shape[0]
    e = np.concatenate([f0[:, [0, 1]], f0[:, [1, 2]], f0[:, [2, 0]]])
    e.sort(axis=1)
    n_edges = len(np.unique(e[:, 0].astype(np.int64) * nv + e[:, 1]))
    assert f1.shape[0] == 4 * nf and v1.shape[0] == nv + n_edges
    assert np.array_equal(v1[:nv].view(np.uint32), v0.view(np.uint32))            # old vertices untouched when the noise is 0
    # every new vertex is the midpoint of an old edge, and new indices appear in first-use order over (face, slot)
    m = np.stack([f1[nf::3, 1], f1[nf + 1::3, 1], f1[nf + 2::3, 1]], axis=1)       # m01, m12, m20 of every old face
    a = np.stack([f0[:, 0], f0[:, 1], f0[:, 2]], axis=1)
    b = np.stack([f0[:, 1], f0[:, 2], f0[:, 0]], axis=1)
    assert np.array_equal(v1[m], (v0[a] * np.float32(0.5) + v0[b] * np.float32(0.5)))
    flat = m.reshape(-1)
    _, first = np.unique(flat, return_index=True)
    assert np.array_equal(flat[np.sort(first)], np.arange(nv, nv + n_edges))
    assert np.array_equal(f1[:nf], np.stack([m[:, 1], m[:, 2], m[:, 0]], axis=1))
    # closed manifold in, closed manifold out
    e = np.concatenate([f1[:, [0, 1]], f1[:, [1, 2]], f1[:, [2, 0]]])
    e.sort(axis=1)
    _, counts = np.unique(e[:, 0].astype(np.int64) * v1.shape[0] + e[:, 1], return_counts=True)
    assert (counts == 2).all()


def test_noise_uses_the_rand_stream_in_vertex_order(oracle):
    sj = shapes.c1_sphere()
    oracle.srand(1)
    v0, _ = oracle.polygonize(sj, _mc(12, 0.55, noise=0.0))
    oracle.srand(1)
    v1, _ = oracle.polygonize(sj, _mc(12, 0.55, noise=0.01))      # amplitude = 0.01 * 10 (polygonizer_algorithm_ob02.hpp:122-141)
    libc = ctypes.CDLL("libc.so.6")
    libc.srand(1)
    r = np.array([libc.rand() for _ in range(v0.size)], dtype=np.int64).astype(np.float32) / np.float32(2147483647)
    amp = np.float32(np.float32(0.01) * np.float32(10.0))
    expect = (v0.reshape(-1).astype(np.float64) + (r.astype(np.float64) - 0.5) * np.float64(amp)).astype(np.float32)
    assert np.array_equal(v1.reshape(-1).view(np.uint32), expect.view(np.uint32))
    # the generator is never re-seeded: the next build continues the stream (the reference's behaviour)
    v2, _ = oracle.polygonize(sj, _mc(12, 0.55, noise=0.01))
    assert not np.array_equal(v1, v2)


def test_progress_reports(oracle):
    sj = shapes.c2_cube_minus_cylinder()
    mc = _mc(20, 0.8, subdiv=0, iters=2, c=1.0, projection=1, qem=1)
    reports, (v, f) = oracle.polygonize_progressive(sj, mc, call_specs='{"progressCallback_id": 7, "call_id": 3, "shape_id": 5}')
    assert [r[2] for r in reports] == [(7, 5, 3)] * 3            # after MC, after the resampling loop, after projection
    v_mc, f_mc = oracle.polygonize(sj, _mc(20, 0.8, subdiv=0))
    assert np.array_equal(reports[0][0], v_mc) and np.array_equal(reports[0][1], f_mc)
    v_rs, _ = oracle.polygonize(sj, _mc(20, 0.8, subdiv=0, iters=2, c=1.0))
    assert np.array_equal(reports[1][0], v_rs) and not np.array_equal(v_rs, v_mc)
    assert np.array_equal(reports[2][0], v) and all(np.array_equal(r[1], f) for r in reports)
    # plain build_geometry reports too, with ids -1 (report_back_enabled is always true, polygonizer_algorithm_ob02.hpp:46)
    reports, _ = oracle.polygonize_progressive(sj, _mc(12, 0.8, subdiv=0))
    assert [r[2] for r in reports] == [(-1, -1, -1)] * 2
