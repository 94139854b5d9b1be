"""Multi-process check of the z-slab path over NCCL (run under torchrun, one rank per GPU):
every rank first builds the whole mesh alone, then the ranks build it sharded (one ncclAllGather of {nv,nf} inside the
library), gather the slabs (ncclBroadcast group) and compare.  Prints one JSON line on rank 0; exit code 1 on mismatch."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import implisolid_b200  # noqa: E402
from implisolid_b200 import shapes  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    lib = implisolid_b200.load_library()
    lib.lib.isd_init(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    cases = [("c2", shapes.c2_cube_minus_cylinder, 96, 0.8, {}), ("c4", shapes.c4_deep_chain, 120, 1.0, {}),
             ("c2_ob02", shapes.c2_cube_minus_cylinder, 64, 0.8, dict(iters=1, c=1.0, projection=1, qem=1)),                 # halo mode
             ("c4_ob02", shapes.c4_deep_chain, 90, 1.0, dict(iters=2, c=1.0, projection=1, qem=1, repeats=2)),               # halo mode
             ("c2_ob02_replicated", shapes.c2_cube_minus_cylinder, 64, 0.8, dict(iters=1, c=1.0, projection=1, qem=1, halo=0)),
             # a small off-centre shape with equal-layer slabs: at least one rank's slab (plus halo) holds no faces and must still
             # take part in the step's collective (the mean-edge gather); with and without projection, MC only, and balanced cuts
             ("tiny_empty_slabs_ob02", lambda: shapes.tiny_sphere(0.3, (0.1, -0.2, 0.55)), 64, 1.0, dict(iters=1, c=1.0, projection=1, qem=1, balance=0)),
             ("tiny_empty_slabs_resample_only", lambda: shapes.tiny_sphere(0.3, (0.1, -0.2, 0.55)), 64, 1.0, dict(iters=1, c=1.0, balance=0)),
             ("tiny_empty_slabs_mc", lambda: shapes.tiny_sphere(0.3, (0.1, -0.2, 0.55)), 64, 1.0, dict(balance=0)),
             ("tiny_balanced_ob02", lambda: shapes.tiny_sphere(0.3, (0.1, -0.2, 0.55)), 64, 1.0, dict(iters=1, c=1.0, projection=1, qem=1, repeats=2)),
             ("c4_ob02_point_lists", shapes.c4_deep_chain, 128, 1.0, dict(iters=1, c=1.0, projection=1, qem=1, point_spec=1))]
    singles = [lib.polygonize(shape(), shapes.mc_settings(R, half, **kw)) for _, shape, R, half, kw in cases]
    uid = torch.zeros(128, dtype=torch.uint8)
    if rank == 0:
        buf = np.zeros(128, dtype=np.uint8)
        lib.lib.isd_dist_unique_id(buf.ctypes.data)
        uid = torch.from_numpy(buf)
    uid = uid.cuda()
    dist.broadcast(uid, 0)
    buf = uid.cpu().numpy().copy()
    lib.lib.isd_dist_init(rank, world, buf.ctypes.data)
    results = []
    ok = True
    for (name, shape, R, half, kw), (v1, f1) in zip(cases, singles):
        sj, mc = shape(), shapes.mc_settings(R, half, **kw)
        lib.build_geometry_device(sj, mc)
        slab = lib.last_slab()
        lib.lib.isd_dist_gather_mesh()
        v, f = lib.get_v(), lib.get_f()
        lib.finish_geometry()
        same_f = bool(f.shape == f1.shape and np.array_equal(f, f1))
        dv = float(np.abs(v - v1).max()) if v.shape == v1.shape else float("inf")
        good = same_f and dv == 0.0      # OB02 included: the mean edge is an integer (fixed-point) sum, the same on 1 and N GPUs
        ok = ok and good
        results.append({"case": name, "slab": slab, "halo": lib.lib.isd_dist_last_halo(), "faces_equal": same_f, "max_dv": dv, "ok": good})
    lib.lib.isd_dist_finalize()
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(json.dumps({"world": world, "all_ranks_ok": bool(flag.item()), "rank0": results}))
    dist.destroy_process_group()
    sys.exit(0 if flag.item() else 1)


if __name__ == "__main__":
    main()
