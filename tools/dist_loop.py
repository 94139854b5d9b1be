"""Debug helper (torchrun): repeated sharded builds with OB02, prints progress per rank."""
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
sj = shapes.c2_cube_minus_cylinder(); mc = shapes.mc_settings(int(sys.argv[1]) if len(sys.argv) > 1 else 320, 0.8, iters=1, c=1.0, projection=1, qem=1)
for i in range(8):
    t0 = time.perf_counter()
    lib.build_geometry_device(sj, mc); lib.finish_geometry()
    print("rank", rank, "device step", i, "%.2f ms" % ((time.perf_counter() - t0) * 1e3), lib.last_timings(), flush=True)
for i in range(4):
    t0 = time.perf_counter()
    lib.build_geometry(sj, mc); v = lib.get_v_view(); f = lib.get_f_view(); x = float(v[-1, 2]) + int(f[-1, 2]); lib.finish_geometry()
    print("rank", rank, "e2e step", i, "%.2f ms" % ((time.perf_counter() - t0) * 1e3), v.shape, f.shape, flush=True)
torch.cuda.synchronize(); dist.barrier()
lib.lib.isd_dist_finalize(); dist.destroy_process_group()
print("rank", rank, "done", flush=True)
