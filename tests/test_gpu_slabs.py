"""GPU: z-slab sharding, emulated on one GPU (rank by rank) — the stitched mesh must equal the single-GPU mesh exactly,
indices included.  The real multi-process NCCL path is exercised by bench.py --gpus N and tests/test_gpu_nccl.py."""
import numpy as np
import pytest

from implisolid_b200 import shapes, slabs

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("balance", [0, 1])
@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("shape,R,half", [(shapes.c2_cube_minus_cylinder, 70, 0.8), (shapes.c4_deep_chain, 61, 1.0)])
def test_emulated_slabs_stitch_to_the_single_gpu_mesh(gpu_lib, world, shape, R, half, balance):
    """balance=0: equal numbers of cube layers (SURVEY §8e's formula, mirrored by implisolid_b200/slabs.py); balance=1 (the
    default): boundaries follow the surface — contiguous, covering, at least one layer each, and the same stitched mesh."""
    sj, mc = shape(), shapes.mc_settings(R, half, balance=balance)
    L = gpu_lib.lib
    L.isd_dist_emulate(0, 1)
    v1, f1 = gpu_lib.polygonize(sj, mc)
    try:
        counts = []
        for r in range(world):
            L.isd_dist_emulate(r, world)
            gpu_lib.build_geometry(sj, mc)
            s = gpu_lib.last_slab()
            if balance == 0:
                assert (s["z_begin"], s["z_end"]) == slabs.slab_layers(R, r, world)
            assert s["z_end"] > s["z_begin"] and s["z_begin"] == (1 if r == 0 else prev_end)
            prev_end = s["z_end"]
            counts.append((s["nv"], s["nf"]))
            gpu_lib.finish_geometry()
        assert prev_end == R + 5 - 2
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


OB02 = dict(iters=1, c=1.0, projection=1, qem=1)


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("shape,R,half,kw", [(shapes.c2_cube_minus_cylinder, 70, 0.8, OB02), (shapes.c4_deep_chain, 61, 1.0, OB02),
                                            (shapes.c2_cube_minus_cylinder, 54, 0.8, dict(iters=2, c=1.0, projection=1, qem=1)),
                                            (shapes.n1_medley, 50, 0.9, dict(iters=1, c=0.5, projection=1, qem=0))],
                         ids=["c2", "c4", "c2_iters2", "medley_noqem"])
def test_halo_slabs_with_ob02_stitch_to_the_single_gpu_mesh(gpu_lib, world, shape, R, half, kw):
    """Halo mode: every (emulated) rank optimises its own slab plus a halo of recomputed cube layers; the only things a rank
    learns from the others are the counts and the all-reduced fixed-point edge sum, which the test hands over.  The
    concatenation of the exported slabs must equal the single-GPU OB02 result bit for bit, indices included.  (One projection
    per build here: the emulation hands over one all-reduce result; overall_repeats > 1 is covered over NCCL, tools/dist_check.py.)"""
    sj, mc = shape(), shapes.mc_settings(R, half, **kw)
    L = gpu_lib.lib
    L.isd_dist_emulate(0, 1)
    v1, f1 = gpu_lib.polygonize(sj, mc)
    total = L.isd_dist_last_edge_sum()
    assert total > 0 and L.isd_dist_last_halo() == 0
    try:
        counts = []
        for r in range(world):                       # counts come from the marching-cubes stage alone
            L.isd_dist_emulate(r, world)
            gpu_lib.build_geometry(sj, shapes.mc_settings(R, half))
            s = gpu_lib.last_slab()
            counts.append((s["nv"], s["nf"]))
            gpu_lib.finish_geometry()
        offsets, totals = slabs.global_offsets(counts)
        assert totals == (v1.shape[0], f1.shape[0])
        vs, fs, part = [], [], 0
        for r in range(world):
            L.isd_dist_emulate(r, world)
            L.isd_dist_set_offsets(offsets[r][0], offsets[r][1])
            L.isd_dist_set_edge_total(total, f1.shape[0])
            gpu_lib.build_geometry(sj, mc)
            assert L.isd_dist_last_halo() == slabs.halo_layers(1, kw["iters"], kw["projection"], kw["qem"])
            s = gpu_lib.last_slab()
            assert (s["nv"], s["nf"]) == counts[r]
            part += L.isd_dist_last_edge_sum()
            vs.append(gpu_lib.get_v())
            fs.append(gpu_lib.get_f())
            gpu_lib.finish_geometry()
        v, f = np.concatenate(vs), np.concatenate(fs)
    finally:
        L.isd_dist_emulate(0, 1)
    assert part == total                             # the shares of the all-reduce add up exactly (integer sum)
    assert np.array_equal(f, f1)
    assert np.array_equal(v.view(np.uint32), v1.view(np.uint32))


def test_halo_zero_falls_back_to_replicated(gpu_lib):
    L = gpu_lib.lib
    try:
        L.isd_dist_emulate(1, 2)
        gpu_lib.build_geometry(shapes.c1_sphere(), shapes.mc_settings(20, 0.55, halo=0, **OB02))
        assert L.isd_dist_last_halo() == 0
        gpu_lib.finish_geometry()
    finally:
        L.isd_dist_emulate(0, 1)
