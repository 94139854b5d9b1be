"""Is a halo of H - 1 layers enough now that the ghost layer's vertices are written?  Emulated ranks, halo forced to
formula - 1 (and - 2 as a control that must fail), stitched mesh against the single-GPU mesh bit for bit."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import implisolid_b200
from implisolid_b200 import shapes, slabs

lib = implisolid_b200.load_library(); L = lib.lib
cases = [("c2", shapes.c2_cube_minus_cylinder, 70, 0.8, dict(iters=1, c=1.0, projection=1, qem=1)),
         ("c4", shapes.c4_deep_chain, 96, 1.0, dict(iters=1, c=1.0, projection=1, qem=1)),
         ("c4_r160", shapes.c4_deep_chain, 160, 1.0, dict(iters=1, c=1.0, projection=1, qem=1)),
         ("c3", shapes.c3_metaballs, 64, 1.2, dict(iters=1, c=1.0, projection=1, qem=1)),
         ("c2_iters2", shapes.c2_cube_minus_cylinder, 54, 0.8, dict(iters=2, c=1.0, projection=1, qem=1)),
         ("medley_noqem", shapes.n1_medley, 50, 0.9, dict(iters=1, c=0.5, projection=1, qem=0))]
for name, shape, R, half, kw in cases:
    sj = shape()
    L.isd_dist_emulate(0, 1)
    v1, f1 = lib.polygonize(sj, shapes.mc_settings(R, half, **kw))
    total = L.isd_dist_last_edge_sum()
    full = slabs.halo_layers(1, kw["iters"], kw["projection"], kw["qem"])
    for world in (2, 3, 5):
        for H in (full, full - 1, full - 2):
            counts = []
            for r in range(world):
                L.isd_dist_emulate(r, world)
                lib.build_geometry(sj, shapes.mc_settings(R, half)); s = lib.last_slab(); counts.append((s["nv"], s["nf"])); lib.finish_geometry()
            offsets, totals = slabs.global_offsets(counts)
            vs, fs = [], []
            for r in range(world):
                L.isd_dist_emulate(r, world)
                L.isd_dist_set_offsets(offsets[r][0], offsets[r][1])
                L.isd_dist_set_edge_total(total, f1.shape[0])
                lib.build_geometry(sj, shapes.mc_settings(R, half, halo=H, **kw))
                vs.append(lib.get_v()); fs.append(lib.get_f()); lib.finish_geometry()
            v, f = np.concatenate(vs), np.concatenate(fs)
            same = np.array_equal(f, f1) and v.shape == v1.shape and np.array_equal(v.view(np.uint32), v1.view(np.uint32))
            ndiff = int((v.view(np.uint32) != v1.view(np.uint32)).any(axis=1).sum()) if v.shape == v1.shape else -1
            print("%-14s world %d halo %d (formula %d): %s  vertices differing %d" % (name, world, H, full, "IDENTICAL" if same else "differs", ndiff), flush=True)
    L.isd_dist_emulate(0, 1)
