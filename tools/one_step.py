"""One polygonization step for profilers: warm builds outside the capture range, then `--reps` builds between
cudaProfilerStart/Stop (use `ncu --profile-from-start off`).

  python tools/one_step.py c2 256 [--ob02] [--reps 1] [--gpu key=val ...]
"""
import argparse
import ctypes
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import implisolid_b200
from implisolid_b200 import shapes

CFG = {"c1": (shapes.c1_sphere, 0.55), "c2": (shapes.c2_cube_minus_cylinder, 0.8), "c3": (shapes.c3_metaballs, 1.2), "c3ii": (shapes.c3_sum_blend, 1.2),
       "c4": (shapes.c4_deep_chain, 1.0)}

ap = argparse.ArgumentParser()
ap.add_argument("cfg")
ap.add_argument("R", type=int)
ap.add_argument("--ob02", action="store_true")
ap.add_argument("--reps", type=int, default=1)
ap.add_argument("--warm", type=int, default=2)
ap.add_argument("--gpu", nargs="*", default=[])
ap.add_argument("--flops", action="store_true", help="one more build with the flop counters on: flops per kernel and per face")
ap.add_argument("--emulate", default="", help="r/w: build rank r's slab of a w-GPU run on this GPU (halo mode, no NCCL)")
a = ap.parse_args()

if os.environ.get("ISD_LIB_OVERRIDE"):      # A/B against another build of the library (profiling only)
    import implisolid_b200.api as _api
    _api._LIB = os.environ["ISD_LIB_OVERRIDE"]
lib = implisolid_b200.load_library()
fn, half = CFG[a.cfg]
kw = dict(iters=1, c=1.0, projection=1, qem=1) if a.ob02 else {}
for kv in a.gpu:
    k, v = kv.split("=")
    kw[k] = int(v)
sj, mc = fn(), shapes.mc_settings(a.R, half, **kw)
if a.emulate:
    r, w = (int(x) for x in a.emulate.split("/"))
    voff = foff = 0      # global offsets of the slab = what the slabs below it hold (marching cubes only, same cuts)
    kw_mc = {k: v for k, v in kw.items() if k not in ("iters", "c", "projection", "qem", "kernel_times")}
    for below in range(r):
        lib.lib.isd_dist_emulate(below, w)
        lib.lib.isd_dist_set_offsets(0, 0)
        lib.build_geometry_device(sj, shapes.mc_settings(a.R, half, **kw_mc)); lib.finish_geometry()
        sl = lib.last_slab(); voff += sl["nv"]; foff += sl["nf"]
    lib.lib.isd_dist_emulate(r, w)
    lib.lib.isd_dist_set_offsets(voff, foff)
    lib.lib.isd_dist_set_edge_total(95371408039445, 5212532)      # about C4 1024^3's mean edge 0.00142 x 3 x faces
for _ in range(a.warm):
    lib.build_geometry_device(sj, mc)
    lib.finish_geometry()
rt = ctypes.CDLL("libcudart.so.12")
rt.cudaDeviceSynchronize()
rt.cudaProfilerStart()
t0 = time.perf_counter()
for _ in range(a.reps):
    lib.build_geometry_device(sj, mc)
    lib.finish_geometry()
rt.cudaDeviceSynchronize()
wall_ms = (time.perf_counter() - t0) * 1e3 / a.reps
rt.cudaProfilerStop()
print("wall per build (host clock, %d builds back to back): %.4f ms" % (a.reps, wall_ms))
c, t, s = lib.last_counters(), lib.last_timings(), lib.last_slab()
kt = lib.last_kernel_times()
if kt:
    print(" ".join("%s=%.3f" % (k.split("(")[0].replace("k_", ""), v[0]) for k, v in sorted(kt.items(), key=lambda kv: -kv[1][0])))
    gp = lib.last_kernel_gaps()
    print("gaps before:", " ".join("%s=%.3f" % (k.split("(")[0].replace("k_", ""), v) for k, v in sorted(gp.items(), key=lambda kv: -kv[1])[:8]), "sum=%.3f" % sum(gp.values()))
print(a.cfg, a.R, "nodes", c["nodes"], "evaluated", c["nodes_full_eval"], "cut", c["cut_cubes"], "nv", s["nv"], "nf", s["nf"],
      {k: round(v, 4) for k, v in t.items()})

if a.flops:
    kw2 = dict(kw); kw2["count_flops"] = 1; kw2.pop("kernel_times", None)
    lib.build_geometry_device(sj, shapes.mc_settings(a.R, half, **kw2)); lib.finish_geometry()
    fl = lib.last_kernel_flops()
    nf = max(lib.last_slab()["nf"], 1)
    print("flops:", " ".join("%s=%.3gG(%.0f/face%s)" % (k.split("(")[0].replace("k_", ""), v / 1e9, v / nf, (", %.1f TF/s" % (v / (kt[k][0] * 1e-3) / 1e12)) if kt and k in kt and kt[k][0] > 0 else "")
                             for k, v in sorted(fl.items(), key=lambda kv: -kv[1]) if v > 0))
