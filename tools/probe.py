import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import implisolid_b200
from implisolid_b200 import shapes
lib = implisolid_b200.load_library()
for R in (200, 256, 512):
    for kw in ({}, dict(iters=1, c=1.0, projection=1, qem=1)):
        sj = shapes.c2_cube_minus_cylinder()
        for rep in range(3):
            lib.build_geometry_device(sj, shapes.mc_settings(R, 0.8, **kw)); lib.finish_geometry()
        c = lib.last_counters(); t = lib.last_timings()
        print(R, bool(kw), c["nodes"], c["nodes_full_eval"], c["cut_cubes"], {k: round(v, 4) for k, v in t.items()})
for name, R, half in (("c4_deep_chain", 512, 1.0), ("c3_metaballs", 512, 1.2), ("c4_deep_chain", 1024, 1.0)):
    sj = getattr(shapes, name)()
    for rep in range(2):
        lib.build_geometry_device(sj, shapes.mc_settings(R, half)); lib.finish_geometry()
    c = lib.last_counters(); t = lib.last_timings()
    print(name, R, c["nodes"], c["nodes_full_eval"], c["cut_cubes"], lib.last_slab()["nv"], {k: round(v, 4) for k, v in t.items()})
