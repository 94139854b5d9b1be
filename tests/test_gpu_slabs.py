"""GPU: z-slab sharding, emulated on one GPU (rank by rank) — the stitched mesh must equal the single-GPU mesh exactly,
indices included.  The real multi-process NCCL path is exercised by bench.py --gpus N and tests/test_gpu_nccl.py."""
import numpy as np
import pytest

from implisolid_b200 import shapes, slabs

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("shape,R,half", [(shapes.c2_cube_minus_cylinder, 70, 0.8), (shapes.c4_deep_chain, 61, 1.0)])
def test_emulated_slabs_stitch_to_the_single_gpu_mesh(gpu_lib, world, shape, R, half):
    sj, mc = shape(), shapes.mc_settings(R, half)
    L = gpu_lib.lib
    L.isd_dist_emulate(0, 1)
    v1, f1 = gpu_lib.polygonize(sj, mc)
    try:
        counts = []
        for r in range(world):
            L.isd_dist_emulate(r, world)
            gpu_lib.build_geometry(sj, mc)
            s = gpu_lib.last_slab()
            assert (s["z_begin"], s["z_end"]) == slabs.slab_layers(R, r, world)
            counts.append((s["nv"], s["nf"]))
            gpu_lib.finish_geometry()
        offsets, totals = slabs.global_offsets(counts)
        assert totals == (v1.shape[0], f1.shape[0])
        vs, fs = [], []
        for r in range(world):
            L.isd_dist_emulate(r, world)
            L.isd_dist_set_offsets(offsets[r][0], offsets[r][1])
            gpu_lib.build_geometry(sj, mc)
            vs.append(gpu_lib.get_v())
            fs.append(gpu_lib.get_f())
            gpu_lib.finish_geometry()
        v, f = np.concatenate(vs), np.concatenate(fs)
    finally:
        L.isd_dist_emulate(0, 1)
    assert np.array_equal(f, f1)
    assert np.array_equal(v.view(np.uint32), v1.view(np.uint32))
