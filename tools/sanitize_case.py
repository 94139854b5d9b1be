"""Small builds for compute-sanitizer: C2 96^3 and C4 64^3 with OB02 (pair kernels and point lists forced on), a subdivided build,
and a 2-rank emulated halo build of C4 (slab + halo, point lists)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import implisolid_b200
from implisolid_b200 import shapes

lib = implisolid_b200.load_library()
ob = dict(iters=1, c=1.0, projection=1, qem=1)
for sj, mc in [(shapes.c2_cube_minus_cylinder(), shapes.mc_settings(96, 0.8, **ob)),
               (shapes.c4_deep_chain(), shapes.mc_settings(64, 1.0, point_spec=1, **ob)),
               (shapes.random_tree(9), shapes.mc_settings(40, 1.0, point_spec=1, **ob)),
               (shapes.c3_metaballs(), shapes.mc_settings(48, 1.2, point_spec=1, pair=0, **ob))]:
    v, f = lib.polygonize(sj, mc)
    print("built", v.shape, f.shape, lib.last_counters()["projection_by_type"])
L = lib.lib
sj, mc = shapes.c4_deep_chain(), shapes.mc_settings(61, 1.0, point_spec=1, **ob)
for r in range(2):
    L.isd_dist_emulate(r, 2)
    L.isd_dist_set_offsets(0, 0)
    L.isd_dist_set_edge_total(1 << 40, 10000)
    lib.build_geometry(sj, mc)
    print("rank", r, lib.last_slab())
    lib.finish_geometry()
L.isd_dist_emulate(0, 1)
