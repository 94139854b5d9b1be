#!/usr/bin/env python
"""bench.py — voxels/sec of the polygonization hot path (field eval + marching cubes + OB02) on N B200s.

A step = one pass of the path over one synthetic shape spec: build_geometry(shape_json, mc_json) -> mesh.

Headline workload, the SAME for every N (strong scaling): BASELINE.json's deep-CSG config at the north_star's size —
C4, a left-deep chain of 32 transformed primitives (seed 1234), 1024^3, with the OB02 steps the metric names
(vertex resampling 1 iteration, centroid projection, QEM).  At N > 1 the grid is cut into z-slabs (halo mode).
`other_configs` in the same line: C2 256^3 and C3 512^3 (N = 1) and C5 = the C4 shape at 2048^3 (every N), a few steps each.

value     : voxels / s, device-resident (mesh stays in HBM), CUDA events on the library's stream, L2 flushed between steps.
e2e       : the same through the reference-facing C API with host buffers: JSON strings in, get_v/get_f views out.
roofline  : the kernel with the largest share of the step; every kernel is timed by its own pair of events in an instrumented
            repeat of the timed steps ("gpu":{"kernel_times":1}: +5 % per step, so `value` comes from the plain steps);
            FP32 operations come from an instrumented, untimed step ("gpu":{"count_flops":1}).
cpu_baseline / --impl reference : the oracle restatement of the reference's CPU algorithm (the reference itself needs
            Boost/Eigen/Emscripten and cannot be built here), built -O3 -march=native, all host threads, on a bounded
            sample of the same workload: windows of cube layers spread evenly over the 1024^3 grid.
"""
import argparse
import hashlib
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "voxels/sec (eval+MC+OB02)"
UNIT = "voxels/s"
OB02 = dict(iters=1, c=1.0, projection=1, qem=1)
OB02_TEXT = "eval+MC+OB02(resample 1, projection, QEM)"


def configs():
    from implisolid_b200 import shapes
    return {
        "c4_1024": dict(shape=shapes.c4_deep_chain(), R=1024, half=1.0,
                        name="C4 deep CSG chain (32 transformed primitives, seed 1234), %s, R=1024, box +-1.0" % OB02_TEXT),
        "c2_256": dict(shape=shapes.c2_cube_minus_cylinder(), R=256, half=0.8,
                       name="C2 cube-minus-cylinder (affine), %s, R=256, box +-0.8" % OB02_TEXT),
        "c3_512": dict(shape=shapes.c3_metaballs(), R=512, half=1.2,
                       name="C3 64-centre metaballs (Union of 16 meta_balls nodes), %s, R=512, box +-1.2" % OB02_TEXT),
        "c3ii_512": dict(shape=shapes.c3_sum_blend(), R=512, half=1.2,
                         name="C3(ii) one meta_balls node, 64 blobs summed (extension, not expressible in the reference), %s, R=512, box +-1.2" % OB02_TEXT),
        "c5_2048": dict(shape=shapes.c4_deep_chain(), R=2048, half=1.0,
                        name="C5 = the C4 shape at R=2048, %s, box +-1.0" % OB02_TEXT),
    }


HEADLINE = "c4_1024"


def settings(cfg, **gpu):
    from implisolid_b200 import shapes
    return shapes.mc_settings(cfg["R"], cfg["half"], **OB02, **gpu)


def run_config(name, world):
    """The `config` object of the JSON line: identical in both arms."""
    c = configs()[name]
    return {"workload": c["name"], "resolution": c["R"], "voxels_per_step": float(c["R"]) ** 3, "gpus": world,
            "parallelism": "z-slabs x%d" % world if world > 1 else "single GPU",
            "l2": "flushed between timed steps (256 MB write)", "fp_mode": "exact (no FMA contraction)"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._read, daemon=True)
        self.thread.start()

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
            except ValueError:
                continue
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        med = sm[len(sm) // 2] if sm else None
        return {"sm_mhz": med, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d.get("hbm_gbs", 6650.0)), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic():
    """DRAM bytes per launch (read + write) per kernel from the committed `ncu --set full` capture of the headline workload."""
    for name in ("r02_traffic.json", "r01_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                d = json.load(f)
            return {k: float(v["dram_bytes"]) for k, v in d["kernels"].items()}, "profiles/" + name + " (" + d.get("workload", "?") + ")"
        except Exception:
            continue
    return {}, None


def fp32_peak_tflops():
    """Non-tensor FP32 peaks of this GPU, measured by tools/fp32_peak: separately rounded FMUL+FADD (scalar and packed f32x2)."""
    exe = os.path.join(ROOT, "tools", "fp32_peak")
    try:
        if not os.path.exists(exe):      # __graft_entry__.build() makes it; a bare checkout builds it here
            subprocess.run(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-o", exe, exe + ".cu"], check=True, capture_output=True, timeout=120)
        out = subprocess.run([exe], capture_output=True, text=True, timeout=60).stdout.strip().splitlines()[-1]
        return json.loads(out)
    except Exception:
        return None


# ---- the CPU arm: bounded sample of the headline workload ----------------------------------------------------------------
def cpu_sample_windows(R, n_windows, layers):
    """`n_windows` windows of `layers` cube layers each, centred on equal fractions of the grid's cube layers [1, R+2)."""
    total = R + 2
    out = []
    for k in range(n_windows):
        mid = 1 + int((k + 0.5) * total / n_windows)
        z0 = max(1, mid - layers // 2)
        out.append((z0, min(R + 3, z0 + layers)))
    return out


def time_cpu_sample(oracle, cfg, windows):
    """Wall time of the oracle on the windows (field evaluation of their node planes, marching cubes, OB02 on the slices'
    meshes) and the voxels they hold.  Returns (seconds, voxels, per-stage seconds, vertices, faces)."""
    mc = settings(cfg)
    R = cfg["R"]
    stage = {}
    nv = nf = 0
    t0 = time.perf_counter()
    for z0, z1 in windows:
        v, f = oracle.polygonize_layers(cfg["shape"], mc, z0, z1)
        nv += v.shape[0]
        nf += f.shape[0]
        for k, s in oracle.stage_seconds().items():
            stage[k] = stage.get(k, 0.0) + s
    dt = time.perf_counter() - t0
    voxels = float(sum(z1 - z0 for z0, z1 in windows)) * (R + 2) ** 2 * (float(R) ** 3 / float(R + 2) ** 3)
    return dt, voxels, stage, nv, nf


def plan_cpu_sample(oracle, cfg, target_s):
    """Probe one 2-layer window in the middle, then spread windows of 2 layers evenly so the sample takes about target_s."""
    R = cfg["R"]
    t_probe, _, _, _, _ = time_cpu_sample(oracle, cfg, cpu_sample_windows(R, 1, 2))
    n = int(max(4, min(64, target_s / max(t_probe, 1e-3))))      # the middle of the grid is denser than the ends: an overestimate
    return cpu_sample_windows(R, n, 2)


def run_reference(args):
    """Reference arm: the reference's CPU algorithm on rank 0's host cores, same config/metric; each step a bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    from tests.oracle_lib import load_oracle_mt, host_threads
    cfg = configs()[HEADLINE]
    threads = host_threads()
    o = load_oracle_mt(threads)
    windows = plan_cpu_sample(o, cfg, target_s=12.0)
    steps = max(1, min(args.steps, 3))
    warm = 1 if args.warmup > 0 else 0
    for _ in range(warm):
        time_cpu_sample(o, cfg, windows[: max(1, len(windows) // 4)])
    tot_t = tot_vox = 0.0
    stage = {}
    for _ in range(steps):
        dt, vox, st, nv, nf = time_cpu_sample(o, cfg, windows)
        tot_t += dt
        tot_vox += vox
        for k, s in st.items():
            stage[k] = stage.get(k, 0.0) + s / steps
    value = tot_vox / tot_t
    sample = ("%d windows of 2 cube layers spread evenly over the %d layers of the grid (%.2f%% of the voxels), %.1f s per step; "
              "oracle restatement built -O3 -march=native (no FMA contraction, no fast-math), %d host threads of %d cores; the "
              "reference is single-threaded C++ that cannot be compiled here (Boost/Eigen/Emscripten absent)"
              % (len(windows), cfg["R"] + 2, 100.0 * 2 * len(windows) / (cfg["R"] + 2), tot_t / steps, threads, os.cpu_count() or 0))
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": steps, "warmup": warm,
            "ms_per_step": tot_t / steps * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": run_config(HEADLINE, world),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample,
                             "stage_s_per_step": stage, "sample_vertices": nv, "sample_faces": nf},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# ---- parity helpers -------------------------------------------------------------------------------------------------------
def digest(a):
    return hashlib.blake2b(memoryview(a).cast("B"), digest_size=8).hexdigest()


def match_window_to_mesh(np, wv, wf, n_tail, gv, gf):
    """A marching-cubes mesh of a window of cube layers (oracle) against the whole grid's mesh (GPU).  The last `n_tail` faces of
    the window (its top layer's: scan order is z-major) only use vertices whose owner cube — the first cube of the scan to see
    the edge, which is the one whose arithmetic places the vertex — lies inside the window.  Those faces must therefore appear
    in the whole mesh as ONE contiguous run with the same corner positions bit for bit, in the same corner order, and the
    window's vertex numbers must map one-to-one onto the whole mesh's along that run (the same topology).  Faces are matched
    by their nine coordinates, not vertex by vertex: a vertex that lands exactly on a grid node (f = 0 there) shares its position
    with the vertices of the other cut edges at that node."""
    if n_tail == 0:
        return True
    tail = np.ascontiguousarray(wf[wf.shape[0] - n_tail:])
    wkeys = np.ascontiguousarray(wv[tail].reshape(n_tail, 9)).view(np.dtype((np.void, 36))).ravel()
    # candidate starts: faces of the whole mesh whose first corner has the first tail face's first corner position
    first = np.ascontiguousarray(wv[tail[0, 0]]).view(np.uint32)
    g0 = gv.view(np.uint32)[gf[:, 0]]
    for s in np.flatnonzero((g0 == first).all(axis=1)):
        if s + n_tail > gf.shape[0]:
            continue
        run = gf[s:s + n_tail]
        gkeys = np.ascontiguousarray(gv[run].reshape(n_tail, 9)).view(np.dtype((np.void, 36))).ravel()
        if not (gkeys == wkeys).all():
            continue
        pairs = np.unique(np.stack([tail.ravel().astype(np.int64), run.ravel().astype(np.int64)], axis=1), axis=0)
        if np.unique(pairs[:, 0]).shape[0] == pairs.shape[0] and np.unique(pairs[:, 1]).shape[0] == pairs.shape[0]:
            return True
    return False


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-other-configs", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    args.warmup = max(args.warmup, 3)

    import numpy as np
    import torch
    import implisolid_b200

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU path")
    torch.cuda.set_device(local_rank)
    lib = implisolid_b200.load_library()
    lib.lib.isd_init(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        uid = torch.zeros(128, dtype=torch.uint8)
        if rank == 0:
            buf = np.zeros(128, dtype=np.uint8)
            lib.lib.isd_dist_unique_id(buf.ctypes.data)
            uid = torch.from_numpy(buf)
        uid = uid.cuda()
        dist.broadcast(uid, 0)
        buf = uid.cpu().numpy().copy()
        lib.lib.isd_dist_init(rank, world, buf.ctypes.data)

    cfgs = configs()
    stream = torch.cuda.ExternalStream(lib.lib.isd_stream())
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")      # > 126 MB L2

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    def reduce_max(x):
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def gather_all(obj):
        if dist is None:
            return [obj]
        out = [None] * world
        dist.all_gather_object(out, obj)
        return out

    def timed_device_steps(cfg, mc, steps, warmup):
        """`steps` device-resident builds, each between its own pair of events on the library's stream, L2 flushed before each.
        Returns (ms per step: max over ranks of the summed times / steps, per-stage ms, per-kernel (ms, launches), launches)."""
        for _ in range(warmup):
            lib.build_geometry_device(cfg["shape"], mc)
            lib.finish_geometry()
        barrier()
        starts = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
        ends = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
        stage_sum, kern_sum, launches = {}, {}, 0
        with torch.cuda.stream(stream):
            for i in range(steps):
                flush.zero_()                      # L2 flush between timed iterations, outside the timed interval
                starts[i].record(stream)
                lib.build_geometry_device(cfg["shape"], mc)
                lib.finish_geometry()
                ends[i].record(stream)
                for k, v in lib.last_timings().items():
                    stage_sum[k] = stage_sum.get(k, 0.0) + v
                for k, (ms, n) in lib.last_kernel_times().items():
                    a = kern_sum.setdefault(k, [0.0, 0])
                    a[0] += ms
                    a[1] += n
                launches += lib.last_counters()["launches"]
        barrier()
        ms = sum(s.elapsed_time(e) for s, e in zip(starts, ends))
        ms_per_step = reduce_max(ms) / steps
        stages = {k: v / steps for k, v in stage_sum.items()}
        kernels = {k: (a[0] / steps, a[1] / steps) for k, a in kern_sum.items()}
        return ms_per_step, stages, kernels, launches

    # ---- headline: device-resident throughput ("value") ----
    cfg = cfgs[HEADLINE]
    R = cfg["R"]
    voxels = float(R) ** 3
    mc_dev = settings(cfg)                      # the timed region: default settings, nothing but the path's own kernels in the stream
    mc_timed = settings(cfg, kernel_times=1)    # the instrumented region: a pair of events around every launch (+4 % per step)
    sampler = ClockSampler(local_rank)
    sampler.start()
    # first builds allocate; at N > 1 the slab boundaries also settle over the first builds of a spec (measured busy times travel
    # in the step's all-gather, at most 8 re-cuts, three builds each: DESIGN.md §5).  Not part of the warm-up count.
    settle_builds = 2 if world == 1 else 30
    for _ in range(settle_builds):
        lib.build_geometry_device(cfg["shape"], mc_dev)
        lib.finish_geometry()
    barrier()
    sampler.rows.clear()
    ms_per_step, stages, _, launches = timed_device_steps(cfg, mc_dev, args.steps, args.warmup)
    # the same steps again with per-kernel events: the kernels' durations are the same, the step is longer by the event records
    ms_instrumented, _, kernels, _ = timed_device_steps(cfg, mc_timed, args.steps, 3)
    # keep the same load up until the region is >= 0.5 s long, so that nvidia-smi (100 ms period) has samples of it; the
    # count comes from the reduced time, so every rank runs the same number of collectives.  Timed as one block, no flush.
    extra = int(min(5000, max(0, math.ceil((600.0 - ms_per_step * args.steps) / max(ms_per_step, 0.05)))))
    sustained_ms = None
    if extra > 0:
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            e0.record(stream)
            for _ in range(extra):
                lib.build_geometry_device(cfg["shape"], mc_dev)
                lib.finish_geometry()
            e1.record(stream)
        barrier()
        sustained_ms = reduce_max(e0.elapsed_time(e1)) / extra
    clocks = sampler.stop()
    value = voxels / (ms_per_step * 1e-3)
    counters = lib.last_counters()
    # The timed builds repeat one spec, so after the first they take their array sizes from the build before (checked at the end of
    # every build) and never wait for the device in mid-step; every kernel still runs on every step.  The same steps with that
    # switched off ("gpu":{"speculate":0}: the host waits for the device's counts three times per step) are reported beside them.
    speculative_builds = counters["speculative"]
    ms_waiting, _, _, _ = timed_device_steps(cfg, settings(cfg, speculate=0), max(3, min(args.steps, 20)), 3)
    prog = lib.last_program_stats()
    slab = lib.last_slab()
    per_rank = gather_all({"rank": rank, "stage_ms": stages, "nv": slab["nv"], "nf": slab["nf"], "z": [slab["z_begin"], slab["z_end"]],
                           "cut_cubes": counters["cut_cubes"], "nodes_full_eval": counters["nodes_full_eval"]})

    # ---- FP32 operations per kernel: one instrumented, untimed step ----
    lib.build_geometry_device(cfg["shape"], settings(cfg, count_flops=1))
    lib.finish_geometry()
    flops = lib.last_kernel_flops()
    barrier()

    # ---- end to end through the reference-facing C API, host buffers in and out ("e2e") ----
    mc_plain = settings(cfg)

    def step_e2e():
        lib.build_geometry(cfg["shape"], mc_plain)      # JSON strings (host) in; mesh copied to pinned host memory inside
        v = lib.get_v_view()                   # get_v_ptr / get_f_ptr: typed views on the library's host buffers, what the
        f = lib.get_f_view()                   # reference's JS layer does (HEAPF32.subarray, implisolid_main.js:230-237)
        if len(v) and len(f):
            _ = float(v[-1, 2]) + int(f[-1, 2])     # touch the last elements: the data is in host memory
        lib.finish_geometry()
        return v, f
    for _ in range(2):
        v, f = step_e2e()
    barrier()
    e2e_steps = max(3, min(args.steps, 20))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        v, f = step_e2e()
    torch.cuda.synchronize()
    e2e_dt = reduce_max(time.perf_counter() - t0)
    e2e_value = voxels / (e2e_dt / e2e_steps)
    h2d = len(cfg["shape"]) + len(mc_plain) + prog["words"] * 4 + 2048
    d2h = int(sum(gather_all(int(v.nbytes + f.nbytes) + 64)))
    e2e_counters = lib.last_counters()
    slab_e2e = lib.last_slab()
    v, f = v.copy(), f.copy()      # the views die with the next build

    # ---- N > 1: the sharded mesh against the single-GPU mesh of the same spec (rank 0 builds it whole) ----
    parity = None
    if world > 1:
        mine = {"rank": rank, "voff": slab_e2e["voff"], "nv": int(v.shape[0]), "foff": slab_e2e["foff"], "nf": int(f.shape[0]),
                "hv": digest(np.ascontiguousarray(v)), "hf": digest(np.ascontiguousarray(f))}
        everyone = gather_all(mine)
        if rank == 0:
            lib.lib.isd_dist_suspend()
            fv, ff = lib.polygonize(cfg["shape"], mc_plain)
            lib.lib.isd_dist_resume()
            ok, at_v, at_f = True, 0, 0
            for m in sorted(everyone, key=lambda m: m["rank"]):
                ok = ok and m["voff"] == at_v and m["foff"] == at_f
                ok = ok and digest(np.ascontiguousarray(fv[at_v:at_v + m["nv"]])) == m["hv"]
                ok = ok and digest(np.ascontiguousarray(ff[at_f:at_f + m["nf"]])) == m["hf"]
                at_v += m["nv"]
                at_f += m["nf"]
            ok = ok and at_v == fv.shape[0] and at_f == ff.shape[0]
            parity = {"parity_ok": bool(ok), "what": "blake2b of every rank's vertex bits and global face indices == the same ranges of the mesh "
                      "rank 0 builds whole on one GPU (same spec, default settings); ranges tile the mesh", "nv": int(fv.shape[0]), "nf": int(ff.shape[0])}
        barrier()

    # ---- the other BASELINE configs, a few steps each ----
    others = {}
    if not args.no_other_configs:
        plan = [("c5_2048", 5, 30)] if world > 1 else [("c2_256", 20, 3), ("c3_512", 10, 3), ("c3ii_512", 10, 3), ("c5_2048", 3, 2)]
        for name, steps, warm in plan:
            c = cfgs[name]
            try:
                ms_o, st_o, _, _ = timed_device_steps(c, settings(c), steps, warm)
            except Exception as e:      # a config that does not fit is reported, not fatal
                others[name] = {"error": str(e)}
                continue
            s_o = lib.last_slab()
            c_o = lib.last_counters()
            tot = gather_all((s_o["nv"], s_o["nf"]))
            others[name] = {"workload": c["name"], "ms_per_step": ms_o, "value": float(c["R"]) ** 3 / (ms_o * 1e-3), "unit": UNIT, "steps": steps,
                            "warmup": warm, "nv": int(sum(t[0] for t in tot)), "nf": int(sum(t[1] for t in tot)), "stage_ms_rank0": st_o,
                            "projection_failed_rank0": c_o["projection_failed"], "bisection_capped_rank0": c_o["bisection_capped"]}

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    # ---- roofline entries: every kernel of the step, measured live (its own events inside the timed region) ----
    hbm_peak, hbm_src = measured_peaks()
    fp = fp32_peak_tflops()
    fp_peak = fp["fmul2_fadd2_tflops"] if fp else None
    traffic, traffic_src = ncu_traffic()
    nf_local, nv_local = slab["nf"], slab["nv"]
    n_work_tiles = counters["nodes_full_eval"] // 512
    cut = counters["cut_cubes"]
    # algorithmic HBM bytes per launch (per-unit figures x the units one launch processes)
    hbm_bytes = {
        "k_mc_classify_tiles": n_work_tiles * (10 * 10 * 4 + 64 * 16),      # 10x10 rows of one sign word per tile in; 64 segments x 16 B out
        "k_row_scan": (R + 2) ** 2 * 24.0,                                   # 12 B row total in, 12 B row prefix out
        "k_seg_apply": n_work_tiles * 64 * 24.0 + cut * 8.0,                 # segment record in, offsets out; 8 B per cut cube out
        "k_mc_emit": cut * 8.0 + 12.0 * (nv_local + nf_local) + 16.0 * nv_local,     # cube list in, vertices + faces out, <= 4 field values per vertex in
        "adjacency(k_vertex_degree+k_uscan+k_umbrella_fill+k_umbrella_sort)": 12.0 * nf_local * 5 + 8.0 * nv_local * 2,
        "adjacency(k_mc_adjacency<0>+k_uscan+k_mc_adjacency<1>)": 24.0 * nf_local + 12.0 * nv_local + 16.0 * cut,      # items + faces-of-faces out, degree / offsets, cube list twice
        "k_faces_of_faces": 24.0 * nf_local,
        "k_face_weights": 40.0 * nf_local,
        "k_resample": 28.0 * nv_local + 20.0 * 3 * nf_local,
        "k_face_prepare": 30.0 * nf_local,
        "k_qem": 16.0 * nv_local + 28.0 * 3 * nf_local,
        "k_build_point_lists": n_work_tiles * 2.0 * 4 * (prog["nodes"] * 2),
    }
    per_kernel = {}
    for k, (ms, n) in sorted(kernels.items(), key=lambda kv: -kv[1][0]):
        e = {"ms_per_step": ms, "launches_per_step": n, "share_of_step": ms / ms_instrumented}
        if k in flops and ms > 0:
            tf = flops[k] / (ms * 1e-3) / 1e12
            e.update({"bound": "fp32", "achieved": tf, "peak": fp_peak, "unit": "TFLOP/s", "frac": tf / fp_peak if fp_peak else None,
                      "flops_per_step": flops[k]})
        elif k in hbm_bytes and ms > 0:
            gbs = hbm_bytes[k] / (ms * 1e-3) / 1e9
            e.update({"bound": "hbm", "achieved": gbs, "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak, "algorithmic_bytes": hbm_bytes[k]})
        short = k.split("(")[0]
        e["traffic"] = traffic.get(short)
        per_kernel[k] = e
    dominant_kernel = max(kernels, key=lambda k: kernels[k][0]) if kernels else None
    roofline = dict(per_kernel.get(dominant_kernel, {}))
    roofline["kernel"] = dominant_kernel
    roofline["peak_source"] = ("tools/fp32_peak on this GPU: separately rounded multiply and add on register pairs (FFMA2(a,b,-0) + FADD2), the exact mode's "
                               "ceiling; the scalar forms give %.1f, and %.1f TFLOP/s of fused FFMA are out of reach without contraction. Packed instructions "
                               "issue at half rate: they save issue slots, not FP32-pipe cycles" % (fp["fmul_fadd_tflops"], fp["ffma_tflops"])) if roofline.get("bound") == "fp32" and fp else hbm_src
    roofline["traffic_source"] = traffic_src
    roofline["how"] = ("achieved = FP32 operations the kernel's CSG evaluations execute in one step (counted by an instrumented untimed step: the per-tile "
                       "instruction lists each evaluation really runs, mul and add separately) / the kernel's mean time per step inside the timed region"
                       if roofline.get("bound") == "fp32" else "achieved = algorithmic bytes per launch (DESIGN.md §3) / the kernel's mean time inside the timed region")
    t_mc = sum(kernels.get(k, (0.0, 0))[0] for k in ("k_mc_classify_tiles", "k_row_scan", "k_seg_apply"))
    roof_mc = {"kernel": "k_mc_classify_tiles + k_row_scan + k_seg_apply", "bound": "hbm", "ms": t_mc,
               "achieved": 4.0 * counters["nodes"] / (t_mc * 1e-3) / 1e9 if t_mc > 0 else None, "peak": hbm_peak, "unit": "GB/s",
               "frac": 4.0 * counters["nodes"] / (t_mc * 1e-3) / 1e9 / hbm_peak if t_mc > 0 else None,
               "note": "CONVENTIONAL figure for comparison with a field-reading classifier (4 B/node of the resident grid / stage time); the kernels read "
                       "1 bit/node over the K1 work list only, so this exceeds the HBM peak by design. The honest per-kernel entries are in `kernels`."}
    dominant_stage = max((k for k in stages if k not in ("total_host", "d2h")), key=lambda k: stages[k])
    config = run_config(HEADLINE, world)
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config, "settle_builds": settle_builds,
            "mesh": {"nv": slab["global_nv"] if world > 1 else slab["nv"], "nf": slab["global_nf"] if world > 1 else slab["nf"],
                     "grid_nodes_per_gpu": counters["nodes"], "nodes_full_eval_rank0": counters["nodes_full_eval"], "cut_cubes_rank0": cut,
                     "tree_nodes": prog["nodes"], "flops_per_node_whole_tree": prog["flops_per_point"]},
            "timed_region_ms": ms_per_step * args.steps,
            "instrumented": {"ms_per_step": ms_instrumented, "steps": args.steps, "what": "the timed steps repeated with a pair of CUDA events around every "
                             "launch (\"gpu\":{\"kernel_times\":1}); `kernels` and `roofline` come from this region, `value` from the plain one"},
            "sustained": {"ms_per_step": sustained_ms, "steps": extra, "l2": "not flushed, back to back"},
            "stage_ms": stages, "dominant_stage": dominant_stage, "roofline": roofline, "kernels": per_kernel, "roofline_mc_conventional": roof_mc,
            "counters": {"projection_by_type": e2e_counters["projection_by_type"], "projection_failed": e2e_counters["projection_failed"],
                         "bisection_capped": e2e_counters["bisection_capped"], "bisected": e2e_counters["bisected"], "scope": "rank 0's slab + halo" if world > 1 else "whole mesh"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": e2e_steps},
            "gpu_launches": launches, "clocks": clocks, "other_configs": others,
            "speculation": {"timed_builds_speculative": bool(speculative_builds), "what": "a build of the spec the previous build had sizes its arrays and "
                            "grids from that build's counts instead of waiting for the device; all kernels run every step and the device's own counts are "
                            "compared after the step's last synchronisation (a mismatch rebuilds); nothing of the result is reused",
                            "ms_per_step_waiting": ms_waiting, "value_waiting": voxels / (ms_waiting * 1e-3), "steps_waiting": max(3, min(args.steps, 20))}}
    if world > 1:
        line["per_rank"] = sorted(per_rank, key=lambda r: r["rank"])
        smin = {k: min(r["stage_ms"][k] for r in per_rank) for k in stages}
        smax = {k: max(r["stage_ms"][k] for r in per_rank) for k in stages}
        line["stage_ms_min_over_ranks"], line["stage_ms_max_over_ranks"] = smin, smax
        work = {k: v for k, v in smax.items() if k not in ("total_host", "d2h")}
        line["limiting_stage"] = max(work, key=lambda k: work[k])
        line["parity"] = parity
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline_leg(np, lib, cfg, v, f)
    print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


def cpu_baseline_leg(np, lib, cfg, gv_full, gf_full):
    """Rank 0, N = 1: the oracle (threaded -O3 build, same bits as the parity build) on a bounded sample of the headline
    workload, plus parity checks of the GPU results against it — vertices, not only faces."""
    from implisolid_b200 import shapes
    from tests.oracle_lib import load_oracle_mt, host_threads
    threads = host_threads()
    o = load_oracle_mt(threads)
    windows = plan_cpu_sample(o, cfg, target_s=15.0)
    dt, vox, stage, nv, nf = time_cpu_sample(o, cfg, windows)
    out = {"value": vox / dt, "unit": UNIT, "cores": threads, "kind": "port",
           "sample": "%d windows of 2 cube layers spread evenly over the grid's %d layers (%.2f%% of the voxels), %.1f s; oracle built -O3 -march=native, "
                     "%d threads" % (len(windows), cfg["R"] + 2, 100.0 * 2 * len(windows) / (cfg["R"] + 2), dt, threads),
           "stage_s": stage, "sample_vertices": nv, "sample_faces": nf}
    # (1) at the headline size: marching cubes of three windows, bit for bit inside the GPU's 1024^3 mesh (MC only: OB02 moves
    #     every vertex with the whole mesh's mean edge, which a window does not know)
    gv, gf = lib.polygonize(cfg["shape"], shapes.mc_settings(cfg["R"], cfg["half"]))
    mc_only = shapes.mc_settings(cfg["R"], cfg["half"])
    ok = True
    checked = 0
    for z0, z1 in (windows[len(windows) // 8], windows[len(windows) // 2], windows[-1 - len(windows) // 8]):
        wv, wf = o.polygonize_layers(cfg["shape"], mc_only, z0, z1)
        _, top = o.polygonize_layers(cfg["shape"], mc_only, z1 - 1, z1)      # the top layer alone: how many of the faces are its
        ok = ok and match_window_to_mesh(np, wv, wf, top.shape[0], gv, gf)
        checked += top.shape[0]
    out["headline_mc_windows_faces_checked"] = int(checked)
    out["headline_mc_windows_bit_exact"] = bool(ok)
    # (2) the whole path, default settings, the same shape at R = 160: faces equal, vertices within 1e-5 x bbox diagonal, counters equal
    small = shapes.mc_settings(160, cfg["half"], **OB02)
    o.build_geometry(cfg["shape"], small)
    ov, of, ost = o.get_v(), o.get_f(), o.stats()
    o.finish_geometry()
    lib.build_geometry(cfg["shape"], small)
    sv, sf, sst = lib.get_v(), lib.get_f(), lib.last_counters()
    lib.finish_geometry()
    tol = 1e-5 * 2 * cfg["half"] * math.sqrt(3.0)
    faces_equal = bool(of.shape == sf.shape and (of == sf).all())
    dv = float(np.abs(ov - sv).max()) if ov.shape == sv.shape else float("inf")
    out["full_path_r160"] = {"faces_equal": faces_equal, "max_abs_dv": dv, "tolerance": tol, "within_tolerance": bool(dv <= tol),
                             "projection_counters_equal": bool(sst["projection_by_type"] == ost["projection_by_type"] and
                                                               sst["projection_failed"] == ost["projection_failed"]),
                             "nv": int(sv.shape[0]), "nf": int(sf.shape[0])}
    # (3) the full-size OB02 result: closed 2-manifold (every edge in exactly two faces)
    e = np.sort(np.concatenate([gf_full[:, [0, 1]], gf_full[:, [1, 2]], gf_full[:, [2, 0]]]), axis=1).astype(np.int64)
    keys = e[:, 0] * (gv_full.shape[0] + 1) + e[:, 1]
    _, cnt = np.unique(keys, return_counts=True)
    out["headline_mesh"] = {"closed_manifold": bool((cnt == 2).all()), "finite": bool(np.isfinite(gv_full).all()),
                            "euler_characteristic": int(gv_full.shape[0] - keys.shape[0] // 2 + gf_full.shape[0])}
    out["mesh_matches_gpu"] = bool(ok and faces_equal and dv <= tol)
    return out


if __name__ == "__main__":
    main()
