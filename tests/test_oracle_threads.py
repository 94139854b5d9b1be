"""The threaded -O3 -march=native build of the oracle (bench.py's CPU arm, large parity cases) gives the bits of the
single-threaded parity build, and its cube-layer windows are slices of the whole grid's marching-cubes mesh."""
import numpy as np
import pytest

from implisolid_b200 import shapes
from tests.oracle_lib import load_oracle_mt


@pytest.mark.parametrize("shape,R,half", [(shapes.c4_deep_chain, 48, 1.0), (shapes.c3_metaballs, 40, 1.2), (shapes.c2_cube_minus_cylinder, 44, 0.8)],
                         ids=["c4", "c3", "c2"])
def test_threaded_build_is_bit_identical(oracle, shape, R, half):
    sj = shape()
    mc = shapes.mc_settings(R, half, iters=1, c=1.0, projection=1, qem=1)
    v1, f1 = oracle.polygonize(sj, mc)
    for threads in (1, 3, 8):
        o = load_oracle_mt(threads)
        v, f = o.polygonize(sj, mc)
        assert np.array_equal(f, f1)
        assert np.array_equal(v.view(np.uint32), v1.view(np.uint32)), threads


def test_layer_windows_are_slices_of_the_whole_mesh(oracle):
    """oracle_build_layers (MarchingCubes' window, omc.hpp): the faces of the window's top layer, renamed by vertex position,
    are a contiguous run of the whole grid's faces — what bench.py checks at 1024^3 against the GPU mesh."""
    import bench
    sj = shapes.c4_deep_chain()
    mc = shapes.mc_settings(96, 1.0)
    gv, gf = oracle.polygonize(sj, mc)
    o = load_oracle_mt(4)
    seen = 0
    for z0, z1 in bench.cpu_sample_windows(96, 6, 2):
        wv, wf = o.polygonize_layers(sj, mc, z0, z1)
        _, top = o.polygonize_layers(sj, mc, z1 - 1, z1)
        assert bench.match_window_to_mesh(np, wv, wf, top.shape[0], gv, gf)
        seen += top.shape[0]
    assert seen > 0
    # all layers, one window per layer: the face counts add up to the whole mesh's
    total = sum(o.polygonize_layers(sj, mc, z, z + 1)[1].shape[0] for z in range(1, 96 + 3))
    assert total == gf.shape[0]
