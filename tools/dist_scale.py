"""z-slab scaling of the BASELINE configs (run under torchrun, one rank per GPU; also works as a single process):

  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/dist_scale.py c4 1024 [--ob02] [--steps K]

Every rank builds its slab of the shape at resolution R (one ncclAllGather of {nv,nf} inside the library; with --ob02 the
stitched mesh is gathered and OB02 runs as the library does for sharded meshes).  Timing: CUDA events on the library's
stream around each step, max over ranks per step, median over steps.  Rank 0 prints one JSON line.
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import implisolid_b200  # noqa: E402
from implisolid_b200 import shapes  # noqa: E402

SHAPES = {"c2": (shapes.c2_cube_minus_cylinder, 0.8), "c3": (shapes.c3_metaballs, 1.2), "c4": (shapes.c4_deep_chain, 1.0),
          "medley": (shapes.n1_medley, 0.9)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("shape")
    ap.add_argument("resolution", type=int)
    ap.add_argument("--ob02", action="store_true")
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    lib = implisolid_b200.load_library()
    lib.lib.isd_init(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        uid = torch.zeros(128, dtype=torch.uint8)
        if rank == 0:
            buf = np.zeros(128, dtype=np.uint8)
            lib.lib.isd_dist_unique_id(buf.ctypes.data)
            uid = torch.from_numpy(buf)
        uid = uid.cuda()
        dist.broadcast(uid, 0)
        buf = uid.cpu().numpy().copy()
        lib.lib.isd_dist_init(rank, world, buf.ctypes.data)
    shape_fn, half = SHAPES[a.shape]
    sj = shape_fn()
    kw = dict(iters=1, c=1.0, projection=1, qem=1) if a.ob02 else {}
    mc = shapes.mc_settings(a.resolution, half, **kw)
    stream = torch.cuda.ExternalStream(lib.lib.isd_stream())
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(a.warmup):
        lib.build_geometry_device(sj, mc)
        lib.finish_geometry()
    barrier()
    per_step, stages = [], {}
    with torch.cuda.stream(stream):
        for i in range(a.steps):
            flush.zero_()
            barrier()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(stream)
            lib.build_geometry_device(sj, mc)
            lib.finish_geometry()
            e.record(stream)
            torch.cuda.synchronize()
            t = torch.tensor([s.elapsed_time(e)], dtype=torch.float64, device="cuda")
            if dist is not None:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            per_step.append(float(t.item()))
            for k, v in lib.last_timings().items():
                stages.setdefault(k, []).append(v)
    barrier()
    st = torch.tensor([float(np.median(stages[k])) for k in sorted(stages)], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(st, op=dist.ReduceOp.MAX)
    slab, cnt = lib.last_slab(), lib.last_counters()
    mem = torch.tensor([torch.cuda.mem_get_info()[0]], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(mem, op=dist.ReduceOp.MIN)
    if rank == 0:
        ms = float(np.median(per_step))
        R = a.resolution
        print(json.dumps({"shape": a.shape, "resolution": R, "ob02": a.ob02, "n_gpus": world, "ms_per_step_median": ms, "ms_per_step_min": min(per_step),
                          "voxels_per_s": R ** 3 / (ms * 1e-3), "nv": slab["global_nv"], "nf": slab["global_nf"],
                          "stage_ms_max_over_ranks": dict(zip(sorted(stages), [round(x, 4) for x in st.tolist()])),
                          "nodes_rank0": cnt["nodes"], "nodes_full_eval_rank0": cnt["nodes_full_eval"],
                          "min_free_hbm_gb": float(mem.item()) / 1e9, "steps": a.steps}), flush=True)
    if dist is not None:
        lib.lib.isd_dist_finalize()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
