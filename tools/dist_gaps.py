"""Diagnostic (torchrun): per-rank kernel times and the idle stream time before each kernel of sharded builds — the gap before
k_project_search holds the step's one collective (ncclAllGather of 24 bytes per rank) and the wait for the slowest rank.

  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/dist_gaps.py [c4 1024]
"""
import os, sys, time
import numpy as np, torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import implisolid_b200
from implisolid_b200 import shapes
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
lib = implisolid_b200.load_library(); lib.lib.isd_init(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
uid = torch.zeros(128, dtype=torch.uint8)
if rank == 0:
    buf = np.zeros(128, dtype=np.uint8); lib.lib.isd_dist_unique_id(buf.ctypes.data); uid = torch.from_numpy(buf)
uid = uid.cuda(); dist.broadcast(uid, 0); buf = uid.cpu().numpy().copy()
lib.lib.isd_dist_init(rank, world, buf.ctypes.data)
R = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
sj = shapes.c4_deep_chain()
for extra in ({}, {"kernel_times": 1}):
    mc = shapes.mc_settings(R, 1.0, iters=1, c=1.0, projection=1, qem=1, **extra)
    for _ in range(3):
        lib.build_geometry_device(sj, mc); lib.finish_geometry()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    n = 30
    for _ in range(n):
        lib.build_geometry_device(sj, mc); lib.finish_geometry()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) * 1e3 / n
    kt, gp, st = lib.last_kernel_times(), lib.last_kernel_gaps(), lib.last_timings()
    for r in range(world):
        dist.barrier()
        if r == rank:
            print("rank %d %s wall %.3f ms/build  stages %s" % (rank, extra, wall, {k: round(v, 3) for k, v in st.items() if v}), flush=True)
            if kt:
                print("   kernels:", " ".join("%s=%.3f" % (k.split("(")[0].replace("k_", ""), v[0]) for k, v in sorted(kt.items(), key=lambda kv: -kv[1][0])[:10]), "sum=%.3f" % sum(v[0] for v in kt.values()), flush=True)
                print("   gaps before:", " ".join("%s=%.3f" % (k.split("(")[0].replace("k_", ""), v) for k, v in sorted(gp.items(), key=lambda kv: -kv[1])[:6]), "sum=%.3f" % sum(gp.values()), flush=True)
torch.cuda.synchronize(); dist.barrier()
lib.lib.isd_dist_finalize(); dist.destroy_process_group()
