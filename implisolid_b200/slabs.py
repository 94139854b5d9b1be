"""z-slab partition of the sampling grid across the GPUs of one box — host-side arithmetic only.

The reference has no sharding; this is the partition `csrc/geometry_api.cu: slab_layers` applies, kept
here in Python so launchers and tests can reason about it without a GPU.

Cubes are visited z-major (marching_cubes.hpp:1041-1057) and both vertices and faces are numbered in that order, so
slab g's vertices/faces are one contiguous global range after those of slabs 0..g-1.  Exchanging the per-slab counts
(one allgather of two int64 per rank) and taking an exclusive prefix is all the stitching needs.
"""


def slab_layers(resolution, rank, world):
    """Cube layers [z_begin, z_end) owned by `rank`; cube layers are z in [1, S-2), S = resolution + 5."""
    layers = resolution + 2
    return 1 + rank * layers // world, 1 + (rank + 1) * layers // world


def slab_node_planes(resolution, rank, world):
    """Node planes [z0, z0+nz) a rank evaluates: its layers, one plane above, and one halo plane below
    (the cubes of layer z_begin-1 introduced the vertices on the slab's bottom plane; they are re-classified,
    not exchanged)."""
    S = resolution + 5
    if world == 1:
        return 0, S
    zb, ze = slab_layers(resolution, rank, world)
    return zb - 1, (ze - zb) + 2


def global_offsets(counts):
    """counts: [(nv, nf)] per rank, as gathered.  Returns [(voff, foff)] per rank and the totals."""
    out, v, f = [], 0, 0
    for nv, nf in counts:
        out.append((v, f))
        v += int(nv)
        f += int(nf)
    return out, (v, f)


def exchange_counts(local_nv, local_nf, group=None):
    """All-gather of {nv, nf} with torch.distributed (whatever backend the group has: nccl on GPUs, gloo in CPU tests).
    Used by launchers that drive the two-phase slab API from Python; the library's own path calls ncclAllGather."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    mine = torch.tensor([int(local_nv), int(local_nf)], dtype=torch.int64, device=dev)
    gathered = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(gathered, mine, group=group)
    return [(int(t[0]), int(t[1])) for t in gathered]


def balanced_cuts(tile_hist, resolution, world):
    """Cube-layer cuts that follow the surface (`gpu.balance`), the mirror of csrc/geometry_api.cu: balanced_layers.

    tile_hist[b] = surface tiles whose z-range is node layers [8b, 8b+8) (k_balance_hist over the whole grid).  Cube layer l
    (z = l + 1) weighs its bin's count plus a uniform term worth a quarter of the mean; rank r gets the layers whose
    cumulative weight falls in [r, r+1)/world of the total, never fewer than one.  Integer arithmetic throughout, so every
    rank computes the same cuts.  Returns world+1 layer indices; rank r owns cube layers [1 + cuts[r], 1 + cuts[r+1])."""
    layers = resolution + 2
    total_bins = sum(int(h) for h in tile_hist)
    uniform = total_bins * 8 // (4 * layers) + 1
    cum = [0] * (layers + 1)
    for l in range(layers):
        cum[l + 1] = cum[l] + int(tile_hist[(l + 1) >> 3]) + uniform
    cuts = [0] * (world + 1)
    cuts[world] = layers
    for r in range(1, world):
        target = cum[layers] * r // world
        l = cuts[r - 1] + 1
        while l < layers - (world - r) and cum[l] < target:
            l += 1
        cuts[r] = l
    return cuts


def halo_layers(overall_repeats, vresampl_iters, projection, qem, override=-1):
    """Cube layers a slab recomputes on each side in halo mode (csrc/geometry_api.cu: setup_grid): a
    resampling iteration reads two layers around a vertex, projection + QEM one; at least one."""
    if override > 0:
        return override
    per_repeat = 2 * vresampl_iters + (1 if (projection and qem) else 0)
    return max(1, overall_repeats * per_repeat)


def halo_slab(resolution, rank, world, halo, cuts=None):
    """Cube layers [zb, ze) a rank computes in halo mode: its own layers widened by `halo` and clamped to [1, S-2)."""
    S = resolution + 5
    if cuts is None:
        zb, ze = slab_layers(resolution, rank, world)
    else:
        zb, ze = 1 + cuts[rank], 1 + cuts[rank + 1]
    return max(zb - halo, 1), min(ze + halo, S - 2)
