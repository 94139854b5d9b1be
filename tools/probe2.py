import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import implisolid_b200
from implisolid_b200 import shapes
lib = implisolid_b200.load_library()
sj = shapes.c2_cube_minus_cylinder()
for gpu in ({}, {"prune": 1}, {"debug_cases": 1}, {"prune": 0}, {"prune": 1}, {}):
    mc = shapes.mc_settings(200, 0.8, **gpu)
    lib.build_geometry_device(sj, mc); lib.finish_geometry()
    c = lib.last_counters()
    print(gpu, c["nodes"], c["nodes_full_eval"], lib.last_timings()["field"])
