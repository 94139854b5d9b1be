"""Umbrella (faces around every vertex) and faces-of-faces of a marching-cubes mesh straight from the cube case codes — host-side
mirror of csrc/mc_adjacency.cuh (k_mc_adjacency), kept so that the construction and its tables are pinned by a CPU test
against the oracle (tests/test_mc_adjacency.py); the kernel itself is held to the generic kernels on the GPU.

The reference builds the umbrella from the finished face list (mesh_algorithms.hpp:178-209: faces around a vertex in
ascending face index).  For a mesh that comes from marching cubes none of that is needed: a vertex sits on one grid edge,
introduced by the first cube in scan order that sees the edge (as its local edge 5, 6 or 10, mc_kernels.cuh), and the faces
around it are the triangles of the four cubes sharing that edge that use it.  Cubes are scanned z, y, x and faces are
numbered cube by cube in table order, so walking the four cubes in scan order and each cube's triangles in table order yields
the umbrella already sorted.  No atomics, no sort, degree <= 4 cubes x 5 triangles.
"""
import os
import re

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

# owned local edge -> the four cubes around that grid edge as (dx, dy, dz, local edge seen from that cube)
_AROUND = {
    5: [(0, 0, 0, 5), (1, 0, 0, 7), (0, 0, 1, 1), (1, 0, 1, 3)],       # y-edge at (x+1, y, z+1)
    6: [(0, 0, 0, 6), (0, 1, 0, 4), (0, 0, 1, 2), (0, 1, 1, 0)],       # x-edge at (x, y+1, z+1)
    10: [(0, 0, 0, 10), (1, 0, 0, 11), (0, 1, 0, 9), (1, 1, 0, 8)],    # z-edge at (x+1, y+1, z)
}
_OWNED_EDGE = (5, 6, 10)      # ISD_MC_OWNED's 2-bit edge codes 0, 1, 2


def load_tables():
    src = open(os.path.join(_HERE, "csrc", "mc_tables.h")).read()

    def grab(name):
        body = re.search(name + r"\[[^\]]*\]\s*=\s*\{(.*?)\};", src, re.S).group(1)
        return np.array([int(t, 0) for t in re.findall(r"-?0x[0-9a-fA-F]+|-?\d+", body)], dtype=np.int64)
    return grab("ISD_MC_TRI").reshape(256, 16), grab("ISD_MC_NTRI"), grab("ISD_MC_OWNED")


def umbrella_from_cases(cases):
    """cases[z][y][x]: case code of the visited cubes (x, y, z each over [1, S-2), index 0 = cube 1).  Returns the CSR
    (offsets[nv + 1], items) of faces around each vertex, vertices and faces numbered as marching cubes numbers them."""
    tri, ntri, owned = load_tables()
    cases = np.asarray(cases, dtype=np.int64)
    nz, ny, nx = cases.shape
    flat = cases.reshape(-1)
    fbase = np.concatenate([[0], np.cumsum(ntri[flat])])                      # first face of every cube, scan order
    nown = owned[flat] & 3
    offsets, items = [0], []
    for c in np.nonzero(nown)[0]:                                               # vertices appear cube by cube
        z, r = divmod(int(c), ny * nx)
        y, x = divmod(r, nx)
        own = int(owned[flat[c]])
        for j in range(own & 3):                                                # first-appearance order inside the cube
            e = _OWNED_EDGE[(own >> (2 + 2 * j)) & 3]
            around = []
            for dx, dy, dz, le in _AROUND[e]:
                xx, yy, zz = x + dx, y + dy, z + dz
                if xx >= nx or yy >= ny or zz >= nz:                            # beyond the visited range: sealed, no triangles
                    continue
                around.append(((zz * ny + yy) * nx + xx, le))
            for cube, le in sorted(around):                                     # scan order = ascending face index
                row = tri[flat[cube]]
                for t in range(int(ntri[flat[cube]])):
                    if le in row[3 * t:3 * t + 3]:
                        items.append(int(fbase[cube]) + t)
            offsets.append(len(items))
    return np.array(offsets, dtype=np.int64), np.array(items, dtype=np.int64)


# cube faces: 0 -x, 1 +x, 2 -y, 3 +y, 4 -z, 5 +z (mc_adjacency.cuh d_edge_faces / d_across)
_EDGE_FACES = [{2, 4}, {1, 4}, {3, 4}, {0, 4}, {2, 5}, {1, 5}, {3, 5}, {0, 5}, {0, 2}, {1, 2}, {1, 3}, {0, 3}]
_ACROSS = {0: {3: 1, 7: 5, 8: 9, 11: 10}, 1: {1: 3, 5: 7, 9: 8, 10: 11}, 2: {0: 2, 4: 6, 8: 11, 9: 10},
           3: {2: 0, 6: 4, 10: 9, 11: 8}, 4: {0: 4, 1: 5, 2: 6, 3: 7}, 5: {4: 0, 5: 1, 6: 2, 7: 3}}
_STEP = {0: (-1, 0, 0), 1: (1, 0, 0), 2: (0, -1, 0), 3: (0, 1, 0), 4: (0, 0, -1), 5: (0, 0, 1)}


def faces_of_faces_from_cases(cases):
    """The face across every side of every face (build_faces_of_faces, mesh_algorithms.hpp:50-131) from the case codes alone: the
    faces that hold both end vertices of a side are triangles of the face's own cube and — when both grid edges lie on one cube
    face — of the neighbour across it; fewer than two -> 0, else the first in face order unless that is the face itself, then
    the last.  Returns an (nf, 3) array."""
    tri, ntri, _ = load_tables()
    cases = np.asarray(cases, dtype=np.int64)
    nz, ny, nx = cases.shape
    flat = cases.reshape(-1)
    fbase = np.concatenate([[0], np.cumsum(ntri[flat])])
    out = np.zeros((int(fbase[-1]), 3), dtype=np.int64)

    def holders(cube, ea, eb):
        row = tri[flat[cube]]
        return [int(fbase[cube]) + t for t in range(int(ntri[flat[cube]])) if ea in row[3 * t:3 * t + 3] and eb in row[3 * t:3 * t + 3]]

    for c in np.nonzero(ntri[flat])[0]:
        z, r = divmod(int(c), ny * nx)
        y, x = divmod(r, nx)
        row = tri[flat[c]]
        for t in range(int(ntri[flat[c]])):
            f = int(fbase[c]) + t
            for j in range(3):
                ea, eb = int(row[3 * t + j]), int(row[3 * t + (j + 1) % 3])
                if ea == eb:
                    continue
                found = holders(c, ea, eb)
                common = _EDGE_FACES[ea] & _EDGE_FACES[eb]
                if common:
                    d = min(common)
                    dx, dy, dz = _STEP[d]
                    xx, yy, zz = x + dx, y + dy, z + dz
                    if 0 <= xx < nx and 0 <= yy < ny and 0 <= zz < nz:
                        found += holders((zz * ny + yy) * nx + xx, _ACROSS[d][ea], _ACROSS[d][eb])
                if len(found) > 1:
                    found.sort()
                    out[f, j] = found[0] if found[0] != f else found[-1]
    return out
