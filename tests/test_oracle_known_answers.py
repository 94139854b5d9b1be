"""CPU: pins the oracle against every known answer the reference's own source holds for this path, plus analytic
properties (the reference has no tests and cannot be built here: parity is otherwise unpinned)."""
import json
import os

import numpy as np
import pytest

from implisolid_b200 import shapes
from tests.golden.make_golden import CASES, run_case

HERE = os.path.dirname(os.path.abspath(__file__))

# mesh_algorithms.hpp:155-175 — "It is correct." comment: octahedron faces and the faces around every vertex
REF_OCTA_FACES = np.array([[0, 1, 2], [1, 0, 3], [0, 2, 4], [3, 0, 4], [1, 5, 2], [1, 3, 5], [2, 5, 4], [3, 4, 5]], dtype=np.int32)
REF_OCTA_UMBRELLA = [[0, 1, 2, 3], [0, 1, 4, 5], [0, 2, 4, 6], [1, 3, 5, 7], [2, 3, 6, 7], [4, 5, 6, 7]]
# mcc2.cpp:254-273 make_a_cube
CUBE_V = (np.array([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0], [0, 0, 1], [1, 0, 1], [1, 1, 1], [0, 1, 1]], dtype=np.float32) - 0.5)
CUBE_F = np.array([[0, 2, 1], [0, 3, 2], [0, 5, 4], [0, 1, 5], [1, 6, 5], [1, 2, 6], [2, 3, 6], [6, 3, 7], [4, 3, 0], [4, 7, 3], [4, 5, 6], [4, 6, 7]],
                  dtype=np.int32)


def test_umbrella_known_answer(oracle):
    off, items = oracle.umbrella(REF_OCTA_FACES, 6)
    got = [list(items[off[v]:off[v + 1]]) for v in range(6)]
    assert got == REF_OCTA_UMBRELLA


def test_faces_of_faces_symmetric_on_fixtures(oracle):
    for faces in (REF_OCTA_FACES, CUBE_F):
        fof = oracle.faces_of_faces(faces)
        for f in range(faces.shape[0]):
            for e in range(3):
                g = fof[f, e]
                assert g != f and f in fof[g]                      # mesh_algorithms.hpp:124-130 symmetry invariant
                assert len(set(faces[f]) & set(faces[g])) == 2      # neighbours share exactly one edge


def test_golden_regression(oracle):
    with open(os.path.join(HERE, "golden", "oracle_regression.json")) as fh:
        golden = json.load(fh)
    for name, spec in CASES.items():
        assert run_case(oracle, *spec) == golden[name], name


def test_sphere_analytic(oracle):
    v, f = oracle.polygonize(shapes.c1_sphere(), shapes.mc_settings(28, 0.55))
    r = np.linalg.norm(v, axis=1)
    assert abs(r - 0.5).max() < 1e-3                    # linear interpolation of 1 - 4 r^2: O(h^2) inside
    e = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]])
    e.sort(axis=1)
    key, counts = np.unique(e[:, 0].astype(np.int64) * v.shape[0] + e[:, 1], return_counts=True)
    assert (counts == 2).all() and f.shape[0] % 2 == 0  # closed 2-manifold, even face count (mesh_algorithms.hpp:59)
    assert v.shape[0] - len(key) + f.shape[0] == 2      # genus 0
    # with OB02 (projection + QEM) vertices land on the surface to the root tolerance
    v2, f2 = oracle.polygonize(shapes.c1_sphere(), shapes.mc_settings(28, 0.55, iters=1, c=1.0, projection=1, qem=1))
    assert np.array_equal(f, f2)
    assert abs(np.linalg.norm(v2, axis=1) - 0.5).max() < 1e-3   # QEM puts vertices on the tangent planes: slightly outside


def test_vertex_numbering_is_first_appearance_by_owner_cube(oracle):
    """The property the CUDA kernels rely on: vertex k's edge is first seen by the cube at (-1,-1) across it, so the
    reference numbering is cube scan order (marching_cubes.hpp:1041-1057,1554-1584)."""
    oracle.build_geometry(shapes.c2_cube_minus_cylinder(), shapes.mc_settings(24, 0.8))
    keys = oracle.vertex_keys()
    S = oracle.grid_size()
    oracle.finish_geometry()
    node, axis = keys // 3, keys % 3
    x, y, z = node % S, (node // S) % S, node // (S * S)
    ox = np.where(axis == 0, x, x - 1)
    oy = np.where(axis == 1, y, y - 1)
    oz = np.where(axis == 2, z, z - 1)
    owner = (oz * S + oy) * S + ox
    assert (np.diff(owner) >= 0).all()
    assert ox.min() >= 1 and oy.min() >= 1 and oz.min() >= 1 and max(ox.max(), oy.max(), oz.max()) <= S - 3


def test_mc_tables_self_consistent():
    import re
    src = open(os.path.join(os.path.dirname(HERE), "implisolid_b200", "csrc", "mc_tables.h")).read()

    def grab(name):
        body = re.search(name + r"\[[^\]]*\]\s*=\s*\{(.*?)\};", src, re.S).group(1)
        return [int(t, 0) for t in re.findall(r"-?0x[0-9a-fA-F]+|-?\d+", body)]
    edge, tri, ntri, owned = grab("ISD_MC_EDGE_MASK"), grab("ISD_MC_TRI"), grab("ISD_MC_NTRI"), grab("ISD_MC_OWNED")
    assert len(edge) == 256 and len(tri) == 4096
    for c in range(256):
        row = tri[c * 16:(c + 1) * 16]
        used = [e for e in row if e >= 0]
        assert len(used) == 3 * ntri[c]
        mask = 0
        for e in used:
            mask |= 1 << e
        assert mask == edge[c]                                # edgeTable[c] <=> edges referenced by triTable[c]
        assert (owned[c] & 3) == sum(1 for e in (5, 6, 10) if mask >> e & 1)
        assert edge[c] == edge[255 - c]


def test_alpha_list_and_inverse(oracle):
    a = oracle.alpha_list(0.03)
    # Appendix D of SURVEY.md: +-1/2 | +-1/4, +-3/4 | +-1/8 ... first ten as used by direction types 1-6
    assert np.allclose(a[:10], [0.5, -0.5, 0.25, -0.25, 0.75, -0.75, 0.125, -0.125, 0.375, -0.375])
    rng = np.random.default_rng(3)
    for _ in range(20):
        q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
        m = np.zeros((3, 4), dtype=np.float32)
        m[:, :3] = q * rng.uniform(0.2, 2.0)
        m[:, 3] = rng.uniform(-1, 1, 3)
        inv = oracle.invert_matrix(m.reshape(-1)).reshape(3, 4)
        full = np.vstack([m.astype(np.float64), [0, 0, 0, 1]])
        assert np.abs(np.linalg.inv(full)[:3] - inv).max() < 1e-5


def test_projection_puts_centroids_on_surface(oracle):
    sj = shapes.c2_cube_minus_cylinder()
    oracle.build_geometry(sj, shapes.mc_settings(32, 0.8, projection=1, qem=1))
    c, n = oracle.projected_centroids()
    st = oracle.stats()
    oracle.finish_geometry()
    f = oracle.query_implicit_values(sj, c[:40000])
    assert (np.abs(f) > 1.0001e-4).sum() <= st["projection_failed"] + st["bisection_capped"]   # bisection.hpp post-condition


def test_make_a_cube_fixture_ob02(oracle):
    """OB02 on the reference's 12-triangle cube fixture (mcc2.cpp:254-273) with a sphere as the implicit object: all 12
    centroids find the surface along the gradient, and QEM moves each corner onto the three tangent planes around it
    (outside the sphere, like the corners of a circumscribed polyhedron)."""
    v = oracle.ob02_on_mesh(shapes.c1_sphere(), CUBE_V * 0.6, CUBE_F, resample_iters=0, projection=1, qem=1)
    st = oracle.stats()
    assert st["projection_by_type"][0] == 12 and st["projection_failed"] == 0
    assert np.isfinite(v).all() and (np.linalg.norm(v, axis=1) > 0.5).all()


# sha256 of the reference's two tables (marching_cubes.hpp:1207-1240, :1246-1503), "edge values,...;triangle values,..."
MC_TABLES_SHA256 = "ca355f6a8bfd54741af8032c0fadb6dff3c0cf22b0ffd5fc349ddc21de657ed9"


def _table(src, name):
    import re
    body = re.search(name + r"\[[^\]]*\]\s*=\s*\{(.*?)\};", src, re.S).group(1)
    body = re.sub(r"//[^\n]*", "", body)
    return [int(t, 0) for t in re.findall(r"-?0x[0-9a-fA-F]+|-?\d+", body)]


def test_mc_tables_are_the_references():
    """The two marching-cubes tables (product's mc_tables.h and the oracle's omc_tables.h) are verbatim transcriptions of
    the reference's `mc_edge_lookup_table` and `mc_triangles_table` (marching_cubes.hpp:1207-1240, :1246-1503): compared with
    the reference's source where it is present (the build container), and everywhere with the digest of that source's tables
    recorded from it."""
    import hashlib
    root = os.path.dirname(HERE)
    prod = open(os.path.join(root, "implisolid_b200", "csrc", "mc_tables.h")).read()
    orac = open(os.path.join(root, "oracle", "omc_tables.h")).read()
    edge, tri = _table(prod, "ISD_MC_EDGE_MASK"), _table(prod, "ISD_MC_TRI")
    assert len(edge) == 256 and len(tri) == 4096
    assert _table(orac, "OMC_EDGE_MASK") == edge and _table(orac, "OMC_TRI") == tri
    digest = hashlib.sha256((",".join(map(str, edge)) + ";" + ",".join(map(str, tri))).encode()).hexdigest()
    ref = "/root/reference/js_iteration_2/marching_cubes.hpp"
    if os.path.exists(ref):
        src = open(ref).read()
        assert _table(src, "mc_edge_lookup_table") == edge
        assert _table(src, "mc_triangles_table") == tri
    assert digest == MC_TABLES_SHA256, digest
