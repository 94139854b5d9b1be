// Mesh adjacency of a marching-cubes mesh straight from the cube case codes (no atomics, no sort, no walk over the face list).
//
// The reference builds both tables from the finished face list: make_neighbour_faces_of_vertex (mesh_algorithms.hpp:178-209: the
// faces around a vertex in ascending face index, "umbrella") and make_edge_lookup + build_faces_of_faces (:50-131: per face and
// side, the face across that side).  ob02_kernels.cuh does the same for any mesh (degree count -> scan -> fill by atomics -> sort;
// then a walk over one end vertex's umbrella per side).  For the mesh K4 has just emitted none of that is needed:
//   * a vertex sits on ONE grid edge, introduced by the first cube in scan order that sees the edge (as its local edge 5, 6 or
//     10, mc_kernels.cuh); the faces around it are the triangles of the four cubes sharing that edge that use it.  Cubes are
//     scanned z, y, x and faces are numbered cube by cube in table order, so walking the four cubes in scan order and each cube's
//     triangles in table order yields the umbrella already sorted (construction pinned on the CPU against the oracle's umbrella:
//     implisolid_b200/mc_adjacency.py, tests/test_mc_adjacency.py);
//   * a triangle side joins the vertices of two grid edges of one cube.  Two grid edges lie together in that cube only, or — when
//     both are on one cube face — in that cube and the neighbour across the face: the faces holding both end vertices are
//     triangles of those one or two cubes whose table rows name both (local) edges.  The reference's rule is then applied as
//     written: fewer than two such faces -> 0 (boundary side); else the first in ascending face order unless that is the face
//     itself, in which case the last.
// Passes (one thread per cut cube, the list K3 wrote, ghost layer of a z-slab included: its cubes own local vertices):
//   k_mc_adjacency<false> : degree of every vertex the cube owns  -> deg[]           (then one exclusive scan, k_uscan)
//   k_mc_adjacency<true>  : the umbrella items of those vertices  -> items[offsets[v] ..]; and, when asked, faces-of-faces of
//                           the cube's own triangles               -> fof[3 f ..]
// Only faces this GPU holds count (cube layers [emit_z0, zc_end)): that is the local mesh the generic kernels would see.
// Both results are held equal, bit for bit, to the generic kernels' (tests/test_gpu_ob02.py::test_mc_adjacency_equals_generic).
#pragma once
#include "mc_kernels.cuh"

namespace isd {

// bit t of d_edge_tris[case * 12 + e]: triangle t of the case's table row uses local edge e (filled by the host from ISD_MC_TRI)
__device__ uint8_t d_edge_tris[256 * 12];

// sign bits of nodes x0 .. x0+4 of one row (x0 >= 0; the row is padded with a whole word)
__device__ __forceinline__ unsigned sign_bits5(const uint32_t* __restrict__ row, int x0) {
    const int w = x0 >> 5;
    return __funnelshift_r(__ldg(row + w), __ldg(row + w + 1), x0 & 31) & 31u;
}

// first face (local numbering) of the cube at (x, y, z): its segment's offset + the triangles of the cubes before it
__device__ __forceinline__ long long cube_face_base(const McDesc& m, const McBuffers& mb, const EmitArgs& ea, int x, int y, int z) {
    const int zl = z - m.g.z0;
    const long long tile = ((long long)(zl >> 3) * m.g.ty + (y >> 3)) * m.g.tx + (x >> 3);
    const long long seg = (long long)__ldg(mb.tile_slot + tile) * 64 + ((zl & 7) * 8 + (y & 7));
    return (long long)mb.seg_off[seg].y + planes_t(mb.seg_bits[seg], (1u << (x & 7)) - 1u) - ea.f_local0;
}

// cube faces: 0 -x, 1 +x, 2 -y, 3 +y, 4 -z, 5 +z.  d_edge_faces[e]: bit d set if local edge e lies on face d; d_across[d][e]: the
// local id of edge e (on face d) seen from the neighbour across face d.  Global memory like the other tables (lane-divergent
// look-ups, L1-resident).
__device__ uint8_t d_edge_faces[12] = {
    (1 << 2) | (1 << 4), (1 << 1) | (1 << 4), (1 << 3) | (1 << 4), (1 << 0) | (1 << 4),      // 0..3: bottom ring
    (1 << 2) | (1 << 5), (1 << 1) | (1 << 5), (1 << 3) | (1 << 5), (1 << 0) | (1 << 5),      // 4..7: top ring
    (1 << 0) | (1 << 2), (1 << 1) | (1 << 2), (1 << 1) | (1 << 3), (1 << 0) | (1 << 3)};     // 8..11: verticals
__device__ int8_t d_across[6][12] = {
    {-1, -1, -1, 1, -1, -1, -1, 5, 9, -1, -1, 10},      // -x: 3->1 7->5 8->9 11->10
    {-1, 3, -1, -1, -1, 7, -1, -1, -1, 8, 11, -1},      // +x: 1->3 5->7 9->8 10->11
    {2, -1, -1, -1, 6, -1, -1, -1, 11, 10, -1, -1},     // -y: 0->2 4->6 8->11 9->10
    {-1, -1, 0, -1, -1, -1, 4, -1, -1, -1, 9, 8},       // +y: 2->0 6->4 10->9 11->8
    {4, 5, 6, 7, -1, -1, -1, -1, -1, -1, -1, -1},       // -z: 0->4 1->5 2->6 3->7
    {-1, -1, -1, -1, 0, 1, 2, 3, -1, -1, -1, -1}};      // +z: 4->0 5->1 6->2 7->3

#ifndef ISD_ADJ_THREADS
#define ISD_ADJ_THREADS 128
#endif
constexpr int kAdjThreads = ISD_ADJ_THREADS;
template <bool FILL>
__global__ void __launch_bounds__(kAdjThreads) k_mc_adjacency(McDesc m, McBuffers mb, EmitArgs ea, const uint32_t* __restrict__ signs,
                                                      const unsigned long long* __restrict__ cubes, const unsigned* __restrict__ n_cubes_ptr,
                                                      unsigned* __restrict__ deg, const unsigned* __restrict__ offsets, int* __restrict__ items,
                                                      int* __restrict__ fof) {
    const unsigned n_cubes = *n_cubes_ptr;
    const int plane_lo = m.g.z0, plane_hi = m.g.z0 + m.g.nz - 1;      // resident node planes
    for (unsigned i = blockIdx.x * kAdjThreads + threadIdx.x; i < n_cubes; i += gridDim.x * kAdjThreads) {
        const unsigned long long id = cubes[i];
        const long long seg = (long long)(id >> 3);
        const int l = (int)(id & 7ull), r = (int)(seg & 63);
        const int t = (int)mb.list[seg >> 6];
        const int tx = t % m.g.tx, ty = (t / m.g.tx) % m.g.ty, tzl = t / (m.g.tx * m.g.ty);
        const int x = 8 * tx + l, y = 8 * ty + (r & 7), z = m.g.z0 + 8 * tzl + (r >> 3);
        // sign rows (dy, dz) in {-1, 0, 1, 2}^2: bit j <-> node x - 1 + j.  Rows of planes this GPU does not hold read as "outside";
        // cubes that would need them hold no local faces and are never looked at (local() below).
        unsigned R[4][4];
#pragma unroll
        for (int dz = 0; dz < 4; ++dz)
#pragma unroll
            for (int dy = 0; dy < 4; ++dy) {
                const int zz = z + dz - 1, yy = y + dy - 1;
                R[dz][dy] = (zz >= plane_lo && zz <= plane_hi && yy < m.g.S) ? sign_bits5(sign_row(m, signs, yy, zz), x - 1) : 31u;
            }
        // The ten cubes the two tables look at, in a fixed order: 0 self, 1 +x, 2 +y, 3 +z, 4 +x+z, 5 +y+z, 6 +x+y, 7 -x, 8 -y, 9 -z.
        // Their case codes and (second pass) first faces are computed ONCE, by straight-line code every lane runs in step (the
        // look-ups of the ten cubes are independent loads in flight together); the divergent parts below only index these arrays.
        // A cube that holds no faces of the local mesh — outside the visited range or this GPU's cube layers — gets code 0: no triangles.
        constexpr int NB[10][3] = {{0, 0, 0}, {1, 0, 0}, {0, 1, 0}, {0, 0, 1}, {1, 0, 1}, {0, 1, 1}, {1, 1, 0}, {-1, 0, 0}, {0, -1, 0}, {0, 0, -1}};
        unsigned cdv[10];
        int fbv[10];
#pragma unroll
        for (int q = 0; q < 10; ++q) {
            if (!FILL && q >= 7) { cdv[q] = 0u; continue; }      // the degree pass only needs the cubes around the owned edges
            const int dx = NB[q][0], dy = NB[q][1], dz = NB[q][2];
            const int xx = x + dx, yy = y + dy, zz = z + dz;
            const bool local = xx >= 1 && xx <= m.g.S - 3 && yy >= 1 && yy <= m.g.S - 3 && zz >= m.emit_z0 && zz < m.zc_end;
            const unsigned cd = case_code(R[dz + 1][dy + 1], R[dz + 1][dy + 2], R[dz + 2][dy + 1], R[dz + 2][dy + 2], dx + 1);
            cdv[q] = local ? cd : 0u;
        }
        // the cube's own code as marching cubes sees it (a ghost cube owns vertices although its faces are not local)
        const unsigned code_self = case_code(R[1][1], R[1][2], R[2][1], R[2][2], 1);
        if (FILL) {
#pragma unroll
            for (int q = 0; q < 10; ++q) {
                fbv[q] = 0;
                if (cdv[q] != 0u && cdv[q] != 255u) fbv[q] = (int)cube_face_base(m, mb, ea, x + NB[q][0], y + NB[q][1], z + NB[q][2]);
            }
        }
        const unsigned own = d_case_info[code_self] >> 8, nv = own & 3u;
        // ---- umbrellas of the vertices this cube owns -------------------------------------------------------------------------
        if (nv) {
            const unsigned long long sb_self = mb.seg_bits[seg];
            const long long vloc = (long long)mb.seg_off[seg].x + planes_v(sb_self, (1u << l) - 1u) - ea.v_local0;
            // the four cubes around an owned edge in scan order, as indices into the list above (4 bits each), and the edge's local id
            // seen from each (4 bits each):  code 0 = edge 5: y-edge at (x+1, y, z+1): self 5, +x 7, +z 1, +x+z 3
            //                               code 1 = edge 6: x-edge at (x, y+1, z+1): self 6, +y 4, +z 2, +y+z 0
            //                               code 2 = edge 10: z-edge at (x+1, y+1, z): self 10, +x 11, +y 9, +x+y 8
            for (unsigned j = 0; j < nv; ++j) {
                const unsigned e = (own >> (2 + 2 * j)) & 3u;
                const unsigned idx4 = e == 0u ? 0x4310u : (e == 1u ? 0x5320u : 0x6210u);
                const unsigned le4 = e == 0u ? 0x3175u : (e == 1u ? 0x0246u : 0x89BAu);
                unsigned n = 0;
                unsigned at = FILL ? offsets[vloc + j] : 0u;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const unsigned q = (idx4 >> (4 * k)) & 15u, le = (le4 >> (4 * k)) & 15u;
                    unsigned mask = d_edge_tris[cdv[q] * 12 + le];
                    if (FILL) {
                        const int fb = fbv[q];
                        while (mask) {
                            const int tt = __ffs(mask) - 1;
                            mask &= mask - 1;
                            items[at++] = fb + tt;
                        }
                    } else {
                        n += __popc(mask);
                    }
                }
                if (!FILL) deg[vloc + j] = n;
            }
        }
        // ---- faces across the sides of this cube's own triangles -----------------------------------------------------------------
        if (FILL && fof != nullptr) {
            const unsigned cd0 = cdv[0];      // 0 for a ghost cube: its faces are not local
            const unsigned nt = d_case_info[cd0] & 0xffu;
            const int fb_self = fbv[0];
            const int8_t* tri = d_tri_table + cd0 * 16;
            for (unsigned tt = 0; tt < nt; ++tt) {
                const int e3[3] = {tri[tt * 3], tri[tt * 3 + 1], tri[tt * 3 + 2]};
                const int f = fb_self + (int)tt;
#pragma unroll
                for (int j = 0; j < 3; ++j) {
                    const int e_a = e3[j], e_b = e3[(j + 1) % 3];
                    int res = 0;
                    if (e_a != e_b) {
                        const unsigned m_own = d_edge_tris[cd0 * 12 + e_a] & d_edge_tris[cd0 * 12 + e_b];
                        int first = fb_self + (__ffs(m_own) - 1), last = fb_self + (31 - __clz(m_own));
                        int count = __popc(m_own);
                        const unsigned common = (unsigned)d_edge_faces[e_a] & (unsigned)d_edge_faces[e_b];
                        if (common) {      // the side lies on a cube face: the neighbour across it may hold it too
                            const int d = __ffs(common) - 1;                  // 0 -x, 1 +x, 2 -y, 3 +y, 4 -z, 5 +z
                            const unsigned q = (0x392817u >> (4 * d)) & 15u;      // -> 7, 1, 8, 2, 9, 3 in the list above
                            const unsigned cd = cdv[q];
                            const unsigned m_nb = d_edge_tris[cd * 12 + d_across[d][e_a]] & d_edge_tris[cd * 12 + d_across[d][e_b]];
                            if (m_nb) {
                                const int lo = fbv[q] + (__ffs(m_nb) - 1), hi = fbv[q] + (31 - __clz(m_nb));
                                first = lo < first ? lo : first;
                                last = hi > last ? hi : last;
                                count += __popc(m_nb);
                            }
                        }
                        if (count > 1) res = (first != f) ? first : last;
                    }
                    fof[(long long)f * 3 + j] = res;
                }
            }
        }
    }
}

}  // namespace isd
