"""More pins for the oracle (parity is UNPINNED by the reference: it has no tests and cannot be built here — DESIGN.md §4):
known answers the reference's sources hold, published test vectors of the third-party algorithms it depends on, and
independent implementations (numpy) of the linear algebra.  CPU only."""
import json

import numpy as np
import pytest

from implisolid_b200 import shapes


def test_mt11213b_published_validation_value(oracle):
    """boost::random::mt11213b (make_random_pm1.hpp:15-28 draws from it): Boost's validation value — the 10000th output of a
    default-constructed (seed 5489) engine is 3809585648 (boost/random/mersenne_twister.hpp)."""
    assert int(oracle.mt11213b_raw(5489, 10000)[-1]) == 3809585648


def test_qem_eigen_solver_against_numpy_svd(oracle):
    """qem.hpp:397-406: JacobiSVD of the 3x3 normal matrix A = sum n n^T (symmetric PSD), rank = #{sigma_i >= sigma_0/680}.
    The oracle's cyclic Jacobi must give numpy's singular values, an orthonormal V that reconstructs A, and the same rank."""
    rng = np.random.default_rng(7)
    for trial in range(200):
        k = rng.integers(1, 8)
        n = rng.normal(size=(k, 3))
        if trial % 3 == 0:      # nearly coplanar / collinear normals: the rank-deficient cases QEM exists for
            n[:, 2] *= 10.0 ** -rng.integers(1, 6)
        if trial % 7 == 0:
            n[:, 1] *= 10.0 ** -rng.integers(1, 6)
        n /= np.linalg.norm(n, axis=1, keepdims=True)
        A = (n[:, :, None] * n[:, None, :]).sum(axis=0).astype(np.float32)
        w, V = oracle.sym_eig3(A)
        s = np.linalg.svd(A.astype(np.float64), compute_uv=False)
        assert w[0] >= w[1] >= w[2]
        assert np.abs(w - s).max() <= 2e-6 * max(1.0, s[0]), (w, s)
        assert np.abs(V.astype(np.float64).T @ V - np.eye(3)).max() <= 1e-5
        assert np.abs((V * w) @ V.T - A).max() <= 1e-5 * max(1.0, s[0])
        thr = s[0] / 680.0
        if np.abs(s - thr).min() > 1e-4 * s[0]:      # away from the threshold the rank is unambiguous
            assert int((np.maximum(w, 0) >= max(w[0] / 680.0, np.finfo(np.float32).tiny)).sum()) == int((s >= thr).sum())


FAILED_TETRAHEDRA = [   # implicit_function/tetrahedron.hpp:346-383, "FAILED TESTS" 1-3, and what rejects each
    ([[0, 0, 0], [0, 0, 0], [10, 1, 0], [1, 1, 10]], 1),          # points are too close
    ([[0, 0, 0], [0, 0, 10], [10, 0, 0], [0.1, 0, 0]], 2),        # cannot make plane based at points 1,3,4
    ([[0, 0, 0], [0, 0, 0.01], [0, 10, 0], [5, 5, 0.1]], 3),      # point 1 too close to opposite plane
]


@pytest.mark.parametrize("corners,check", FAILED_TETRAHEDRA)
def test_tetrahedron_negative_fixtures(oracle, corners, check):
    got = oracle.tetrahedron_integrity(corners)
    assert got != 0
    # fixture 2's four points are coplanar as well; the reference names the first check that trips
    assert got == check, got


def test_tetrahedron_planes_known_answer(oracle):
    corners = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1]], dtype=np.float32)
    assert oracle.tetrahedron_integrity(corners) == 0
    planes = oracle.tetrahedron_planes(corners)
    hom = np.concatenate([corners, np.ones((4, 1), np.float32)], axis=1)
    d = planes @ hom.T                # d[k][i] = plane k at corner i
    for k in range(4):
        for i in range(4):
            if i == k:
                assert d[k, i] > 0    # oriented towards the opposite corner: inside is positive (tetrahedron.hpp:73-131)
            else:
                assert abs(d[k, i]) < 1e-6
    # the unit right tetrahedron: x >= 0, y >= 0, z >= 0, x + y + z <= 1
    assert np.allclose(planes[0], [-1, -1, -1, 1]) and np.allclose(planes[1], [1, 0, 0, 0])
    assert np.allclose(planes[2], [0, 1, 0, 0]) and np.allclose(planes[3], [0, 0, 1, 0])


def test_make_a_square_fixture(oracle):
    """mcc2.cpp:275-290: two triangles (0,2,1), (0,3,2) on a unit square — the open-mesh fixture beside make_a_cube."""
    faces = np.array([[0, 2, 1], [0, 3, 2]], dtype=np.int32)
    off, items = oracle.umbrella(faces, 4)
    assert off.tolist() == [0, 2, 3, 5, 6]
    assert items.tolist() == [0, 1, 0, 0, 1, 1]
    assert faces.shape[0] % 2 == 0      # make_edge_lookup's precondition (mesh_algorithms.hpp:59)


@pytest.mark.parametrize("name,shape,R,half", [
    ("heart", shapes.n1_heart, 40, 1.0),               # examples/js/example_objects.js:72-78 (HEART)
    ("simple_cone", shapes.simple_cone, 28, 0.5),      # :79 (SIMPLE_CONE)
    ("moon", shapes.demo_moon, 40, 0.6),               # :22-70 (MOON)
    ("c2", shapes.c2_cube_minus_cylinder, 40, 0.8),
    ("c4", shapes.c4_deep_chain, 40, 1.0),
    ("random9", lambda: shapes.random_tree(9), 40, 1.0),
])
def test_reference_assertions_hold_on_the_demo_shapes(oracle, name, shape, R, half):
    """The invariants the reference asserts under ASSERT_USED, as counters: bisection inputs bracket the surface
    (bisection.hpp:51-95, centroids_projection.cpp:1140-1143), bisection ends within ROOT_TOLERANCE unless it hit our
    iteration cap, every marching-cubes mesh has an even number of faces (mesh_algorithms.hpp:59) and is a closed 2-manifold."""
    sj = shape()
    oracle.build_geometry(sj, shapes.mc_settings(R, half, iters=1, c=1.0, projection=1, qem=1))
    v, f = oracle.get_v(), oracle.get_f()
    inv, st = oracle.invariants(), oracle.stats()
    oracle.finish_geometry()
    assert f.shape[0] % 2 == 0
    e = np.sort(np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]]), axis=1)
    _, cnt = np.unique(e, axis=0, return_counts=True)
    assert (cnt == 2).all()
    assert inv["not_conjugated"] == 0, (inv, st)
    assert inv["unresolved_after_bisection"] == 0, (inv, st)
    assert np.isfinite(v).all()


def test_union_of_many_children_compiles_to_a_flat_stack(emu, oracle):
    """object_factory.hpp:537-580 folds an n-ary Union into a left-deep chain of identity-matrix nodes.  The product's compiler
    must take any n (the reference does): 100 children run in O(1) value-stack and frame depth and give the oracle's bits."""
    rng = np.random.default_rng(3)
    kids = []
    for k in range(100):
        t = rng.uniform(-0.8, 0.8, size=3)
        s = rng.uniform(0.1, 0.25)
        kids.append({"type": ["ellipsoid", "cylinder", "cone", "cube"][k % 4],
                     "matrix": [float(np.float32(x)) for x in [s, 0, 0, t[0], 0, s, 0, t[1], 0, 0, s, t[2], 0, 0, 0, 1]]})
    sj = json.dumps({"type": "Union", "matrix": shapes.EYE, "children": kids})
    _, stats = emu.compile(sj)
    assert stats["nodes"] == 199
    assert stats["value_depth"] <= 4 and stats["frame_depth"] <= 2, stats
    pts = rng.uniform(-1, 1, size=(4000, 3)).astype(np.float32)
    f1, f4, g = emu.eval(sj, pts)
    fo = oracle.query_implicit_values(sj, pts)
    go = oracle.query_normals(sj, pts, normalize_and_invert=False)
    assert np.array_equal(f1.view(np.uint32), fo.view(np.uint32)) and np.array_equal(f4.view(np.uint32), fo.view(np.uint32))
    assert np.array_equal(g.view(np.uint32), go.view(np.uint32))
