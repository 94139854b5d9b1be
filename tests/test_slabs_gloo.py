"""CPU, world_size 2 over gloo: the z-slab stitching arithmetic (implisolid_b200/slabs.py) — partition, the allgather
of {nv, nf}, exclusive prefix — checked against the oracle's global vertex/face numbering."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from implisolid_b200 import shapes, slabs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _slab_counts_from_oracle(R, rank, world):
    """Vertices/faces whose owner cube lies in this rank's layers, from the oracle's mesh (what K2+K3 count on a GPU)."""
    sys.path.insert(0, ROOT)
    from tests.oracle_lib import load_oracle
    o = load_oracle()
    o.build_geometry(shapes.bushy_tree(), shapes.mc_settings(R, 0.9))
    keys, f, S = o.vertex_keys(), o.get_f(), o.grid_size()
    cases = o.get_case_codes()
    o.finish_geometry()
    node, axis = keys // 3, keys % 3
    z = node // (S * S)
    owner_z = np.where(axis == 2, z, z - 1)
    zb, ze = slabs.slab_layers(R, rank, world)
    mine = (owner_z >= zb) & (owner_z < ze)
    # faces per cube layer from the case codes
    import re
    src = open(os.path.join(ROOT, "implisolid_b200", "csrc", "mc_tables.h")).read()
    body = re.search(r"ISD_MC_NTRI\[[^\]]*\]\s*=\s*\{(.*?)\};", src, re.S).group(1)
    ntri = np.array([int(t) for t in re.findall(r"\d+", body)], dtype=np.int64)
    per_layer = ntri[cases].sum(axis=(1, 2))            # cases[z-1][y-1][x-1]
    nf = int(per_layer[zb - 1:ze - 1].sum())
    idx = np.nonzero(mine)[0]
    return int(mine.sum()), nf, (int(idx.min()) if idx.size else None, int(idx.max()) if idx.size else None), int(keys.shape[0]), int(f.shape[0])


def _worker(rank, world, port, R, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    nv, nf, (lo, hi), gnv, gnf = _slab_counts_from_oracle(R, rank, world)
    counts = slabs.exchange_counts(nv, nf)
    offsets, totals = slabs.global_offsets(counts)
    voff, foff = offsets[rank]
    ok = totals == (gnv, gnf)
    # the slab's vertices are exactly the contiguous global range [voff, voff + nv)
    if nv:
        ok = ok and lo == voff and hi == voff + nv - 1
    out[rank] = (ok, counts, offsets)
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_slab_counts_allgather_gloo(world):
    R = 26
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), R, out), nprocs=world, join=True)
    assert len(out) == world
    for r in range(world):
        ok, counts, offsets = out[r]
        assert ok, (r, counts, offsets)
    assert out[0][1] == out[world - 1][1]


def test_partition_covers_all_layers():
    for R in (3, 28, 256, 1024, 2048):
        for world in (1, 2, 4, 8):
            prev = 1
            for r in range(world):
                zb, ze = slabs.slab_layers(R, r, world)
                assert zb == prev and ze >= zb
                prev = ze
                z0, nz = slabs.slab_node_planes(R, r, world)
                assert z0 >= 0 and z0 + nz <= R + 5
            assert prev == R + 3            # cube layers are z in [1, S-2), S = R + 5


@pytest.mark.parametrize("world", [2, 3, 4, 8])
@pytest.mark.parametrize("R", [16, 61, 256])
def test_balanced_cuts_properties(R, world):
    """slabs.balanced_cuts (mirror of geometry_api.cu: balanced_layers): strictly increasing, covering, >= 1 layer per rank,
    and balanced to within one layer's weight for any histogram."""
    rng = np.random.default_rng(R * 31 + world)
    layers, bins = R + 2, (R + 5 + 7) // 8
    for trial in range(20):
        kind = trial % 4
        if kind == 0:
            hist = np.zeros(bins, dtype=np.int64)                     # empty field: uniform term alone
        elif kind == 1:
            hist = rng.integers(0, 500, bins)
        elif kind == 2:
            hist = np.zeros(bins, dtype=np.int64); hist[rng.integers(0, bins)] = 10000      # all the surface in one bin
        else:
            hist = np.zeros(bins, dtype=np.int64); hist[:max(1, bins // 4)] = 300           # surface at the bottom only
        cuts = slabs.balanced_cuts(hist, R, world)
        assert cuts[0] == 0 and cuts[-1] == layers and len(cuts) == world + 1
        assert all(b > a for a, b in zip(cuts, cuts[1:]))
        uniform = int(hist.sum()) * 8 // (4 * layers) + 1
        w = np.array([int(hist[(l + 1) >> 3]) + uniform for l in range(layers)])
        cum = np.concatenate([[0], np.cumsum(w)])
        for r in range(1, world):
            target = int(cum[-1]) * r // world
            forced_lo = cuts[r] == cuts[r - 1] + 1
            forced_hi = cuts[r] == layers - (world - r)
            if not (forced_lo or forced_hi):
                assert cum[cuts[r] - 1] < target <= cum[cuts[r]]


def test_balanced_cuts_equal_plain_partition_on_a_flat_histogram():
    R, world = 254, 4
    hist = np.full((R + 12) // 8, 64)
    cuts = slabs.balanced_cuts(hist, R, world)
    plain = [slabs.slab_layers(R, r, world)[0] - 1 for r in range(world)] + [R + 2]
    assert max(abs(a - b) for a, b in zip(cuts, plain)) <= 1


def test_halo_radius_and_clamping():
    assert slabs.halo_layers(1, 1, True, True) == 3
    assert slabs.halo_layers(2, 1, True, True) == 6
    assert slabs.halo_layers(1, 0, True, False) == 1
    assert slabs.halo_layers(3, 2, False, False) == 12
    assert slabs.halo_layers(2, 1, True, True, override=5) == 5
    R, world = 64, 4
    S = R + 5
    for r in range(world):
        zb, ze = slabs.slab_layers(R, r, world)
        hb, he = slabs.halo_slab(R, r, world, 7)
        assert 1 <= hb <= zb < ze <= he <= S - 2
        assert hb == max(1, zb - 7) and he == min(S - 2, ze + 7)
