"""CPU: the umbrella of a marching-cubes mesh derived from the cube case codes alone (implisolid_b200/mc_adjacency.py, the
host mirror of csrc/mc_adjacency.cuh) equals the oracle's make_neighbour_faces_of_vertex
(mesh_algorithms.hpp:178-209) built from the finished face list — offsets and items, i.e. the ascending order too."""
import numpy as np
import pytest

from implisolid_b200 import mc_adjacency, shapes


@pytest.mark.parametrize("name,R,half", [("c1_sphere", 20, 0.55), ("c2_cube_minus_cylinder", 30, 0.8), ("c3_metaballs", 28, 1.2),
                                         ("c4_deep_chain", 30, 1.0), ("bushy_tree", 26, 0.9), ("demo_moon", 36, 0.6)])
def test_umbrella_from_case_codes(oracle, name, R, half):
    oracle.build_geometry(getattr(shapes, name)(), shapes.mc_settings(R, half))
    v, f, cases = oracle.get_v(), oracle.get_f(), oracle.get_case_codes()
    oracle.finish_geometry()
    assert f.shape[0] > 0
    off, items = mc_adjacency.umbrella_from_cases(cases)
    ooff, oitems = oracle.umbrella(f, v.shape[0])
    assert np.array_equal(off, ooff) and np.array_equal(items, oitems)
    assert np.diff(off).max() <= 20          # four cubes around an edge, at most five triangles of a case on one edge vertex


@pytest.mark.parametrize("name,R,half", [("c1_sphere", 20, 0.55), ("c2_cube_minus_cylinder", 30, 0.8), ("c3_metaballs", 28, 1.2),
                                         ("c4_deep_chain", 30, 1.0), ("bushy_tree", 26, 0.9), ("demo_moon", 36, 0.6), ("n1_medley", 24, 0.9)])
def test_faces_of_faces_from_case_codes(oracle, name, R, half):
    """The face across every side, from the case codes and the cube-face tables of mc_adjacency.cuh, equals the oracle's
    build_faces_of_faces (mesh_algorithms.hpp:50-131) on the finished face list."""
    oracle.build_geometry(getattr(shapes, name)(), shapes.mc_settings(R, half))
    f, cases = oracle.get_f(), oracle.get_case_codes()
    oracle.finish_geometry()
    fof = mc_adjacency.faces_of_faces_from_cases(cases)
    assert np.array_equal(fof, oracle.faces_of_faces(f))
