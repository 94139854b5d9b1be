// K2-K4 — marching cubes on the resident sign bits and field, reproducing the reference's output ORDER as well as its
// values.
//
// Reference (marching_cubes.hpp): render_geometry :1019-1072 visits cubes z,y,x in [1,S-2); polygonize_single_cube
// :518-718 forms the case code (:553-560), interpolates the cut edges (VIntX/Y/Z :401-495) and walks the triangle
// table; posnormtriv/flush_geometry_queue :727-821,1520-1658 number vertices by first appearance through a
// std::map keyed by edge id.
//
// First appearance without a map.  An edge is shared by four cubes; the first of them in scan order is the one that
// sees it as local edge 6 (x-edges), 5 (y-edges) or 10 (z-edges) — the cube at (-1,-1) across the edge.  Because the
// two outer node layers are sealed to -1e7 (:895-963) every cut edge has that cube inside the visited range, so
//   global vertex order = cube scan order, and inside a cube the first-appearance order of {5,6,10} in its triangle
//   list (table ISD_MC_OWNED, derived from the reference's own tables by tools/gen_mc_tables.py),
//   global face order   = cube scan order x table order,
// and the vertex position is the one that cube computes (its fx/fy/fz, fx2 = fx + w ..., :570-576).  Vertex and face
// indices therefore match the reference exactly, not only after canonical re-ordering.
//
// Passes.  The case code is a function of eight sign bits, and K1 already packed those (1 bit/node), so classification
// never touches the 4 B/node field:
//   K2 mc_classify : one thread per chunk (32 x-adjacent cubes) reads the 8 sign words around it; a chunk whose 4 rows
//                    are uniformly 0 or 1 is empty (the common case).  Cut chunks are finished by the whole warp, one
//                    lane per cube: case code -> {new vertices, triangles}, 2 B / chunk.
//   K3 scan        : exclusive prefix of the chunk counts in scan order + ordered list of non-empty chunks.
//   K4 mc_emit     : one warp per non-empty chunk; case codes of its own row and of the three rows that introduced its
//                    other edges come from the sign words, warp prefix sums become vertex ranks; field values are read
//                    only at the <= 4 corners each new vertex interpolates between.
// Algorithmic HBM bytes: 1 bit/node read (K2) [4 B/node if the field itself were read] + 12 B per vertex + 12 B per
// triangle written (K4).
#pragma once
#include <cuda_runtime.h>
#include "field_kernel.cuh"
#include "mc_tables.h"

namespace isd {

// per case: bits 0-7 triangles, bits 8-15 ISD_MC_OWNED (count + first-appearance order of edges 5,6,10).
// Global memory (L1-resident): look-ups are lane-divergent, which constant memory would serialise.
__device__ uint16_t d_case_info[256];
__device__ int8_t d_tri_table[256 * 16];
__device__ uint16_t d_edge_mask[256];
#define ISD_EDGE_MASK(code) ((unsigned)d_edge_mask[code])

struct McDesc {
    GridDesc g;
    int scan_z0;      // first cube layer classified (ghost layer of a slab, or emit_z0)
    int emit_z0;      // first cube layer whose geometry this GPU emits
    int zc_end;       // one past the last cube layer
    int NX;           // S - 3 cubes per row / rows per layer (x, y in [1, S-2))
    int K;            // chunks per row: chunk k holds cubes x in [32k, 32k+32) (aligned with the sign words)
    long long n_chunks;
    long long ghost_chunks;     // chunks of the ghost layer (0 or NX * K)
};

__device__ __forceinline__ long long chunk_index(const McDesc& m, int z, int y, int k) {
    return ((long long)(z - m.scan_z0) * m.NX + (y - 1)) * m.K + k;
}
__device__ __forceinline__ const uint32_t* sign_row(const McDesc& m, const uint32_t* signs, int y, int z) {
    return signs + ((long long)(z - m.g.z0) * m.g.S + y) * m.g.SW;
}
// sign bits of nodes x = 32k .. 32k+32 (33 bits) of one row
__device__ __forceinline__ unsigned long long row_bits(const uint32_t* __restrict__ row, int k) {
    return (unsigned long long)__ldg(row + k) | ((unsigned long long)(__ldg(row + k + 1) & 1u) << 32);
}
// case code of the cube at bit position `b` given the four 33-bit rows (y,z) (y+1,z) (y,z+1) (y+1,z+1); :553-560
__device__ __forceinline__ unsigned case_code(unsigned long long r00, unsigned long long r10, unsigned long long r01, unsigned long long r11, int b) {
    const unsigned a = (unsigned)(r00 >> b) & 3u, c = (unsigned)(r10 >> b) & 3u, d = (unsigned)(r01 >> b) & 3u, e = (unsigned)(r11 >> b) & 3u;
    // bits: f0 (x,y,z)->1  f1 (x+1,y,z)->2  f3 (x+1,y+1,z)->4  f2 (x,y+1,z)->8  f4->16 f5->32 f7->64 f6->128
    return (a & 1u) | (a & 2u) | ((c & 2u) << 1) | ((c & 1u) << 3) | ((d & 1u) << 4) | ((d & 2u) << 4) | ((e & 2u) << 5) | ((e & 1u) << 7);
}

// ---- K2 ---------------------------------------------------------------------------------------------------------
constexpr unsigned long long kRowMask = 0x1ffffffffull;

// counts[c] = new vertices | triangles << 8 | cut cubes << 16.  For a cut chunk, lane_bits[5c .. 5c+4] hold, bit per
// lane, the two bits of each cube's vertex count and the three bits of its triangle count, so that K4 can turn any
// (chunk, lane) into a vertex / face rank with popcounts.
__global__ void __launch_bounds__(256) k_mc_classify(McDesc m, const uint32_t* __restrict__ signs, uint32_t* __restrict__ counts,
                                                     uint32_t* __restrict__ lane_bits, uint8_t* __restrict__ dbg_cases) {
    const long long c = (long long)blockIdx.x * 256 + threadIdx.x;
    const bool valid = c < m.n_chunks;
    const int lane = threadIdx.x & 31;
    int k = 0, y = 1, z = m.scan_z0;
    unsigned long long r00 = 0, r10 = 0, r01 = 0, r11 = 0;
    bool cut = false;
    if (valid) {
        k = (int)(c % m.K);
        const long long row = c / m.K;
        y = 1 + (int)(row % m.NX);
        z = m.scan_z0 + (int)(row / m.NX);
        r00 = row_bits(sign_row(m, signs, y, z), k);
        r10 = row_bits(sign_row(m, signs, y + 1, z), k);
        r01 = row_bits(sign_row(m, signs, y, z + 1), k);
        r11 = row_bits(sign_row(m, signs, y + 1, z + 1), k);
        const unsigned long long any = r00 | r10 | r01 | r11, all = r00 & r10 & r01 & r11;
        cut = !(any == 0ull || all == kRowMask) || dbg_cases != nullptr;
    }
    unsigned my_count = 0;
    unsigned todo = __ballot_sync(0xffffffffu, cut);
    while (todo) {                         // the warp finishes one cut chunk at a time, one lane per cube
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        const unsigned long long s00 = __shfl_sync(0xffffffffu, r00, src), s10 = __shfl_sync(0xffffffffu, r10, src);
        const unsigned long long s01 = __shfl_sync(0xffffffffu, r01, src), s11 = __shfl_sync(0xffffffffu, r11, src);
        const int sk = __shfl_sync(0xffffffffu, k, src), sy = __shfl_sync(0xffffffffu, y, src), sz = __shfl_sync(0xffffffffu, z, src);
        const int x = 32 * sk + lane;
        const bool xin = x >= 1 && x <= m.g.S - 3;
        const unsigned code = xin ? case_code(s00, s10, s01, s11, lane) : 0u;
        unsigned nv = 0, nt = 0;
        if (code != 0u && code != 255u) {
            const unsigned info = d_case_info[code];
            nv = (info >> 8) & 3u;
            nt = info & 0xffu;
        }
        const unsigned v0 = __ballot_sync(0xffffffffu, nv & 1u), v1 = __ballot_sync(0xffffffffu, nv & 2u);
        const unsigned t0 = __ballot_sync(0xffffffffu, nt & 1u), t1 = __ballot_sync(0xffffffffu, nt & 2u), t2 = __ballot_sync(0xffffffffu, nt & 4u);
        if (lane == src) {
            my_count = (__popc(v0) + 2u * __popc(v1)) | ((__popc(t0) + 2u * __popc(t1) + 4u * __popc(t2)) << 8) | (__popc(t0 | t1 | t2) << 16);
            if (my_count) {
                uint32_t* lb = lane_bits + c * 5;
                lb[0] = v0; lb[1] = v1; lb[2] = t0; lb[3] = t1; lb[4] = t2;
            }
        }
        if (dbg_cases != nullptr && xin)
            dbg_cases[((long long)(sz - m.scan_z0) * m.NX + (sy - 1)) * m.NX + (x - 1)] = (uint8_t)code;
    }
    if (valid) counts[c] = my_count;
}

// ---- K3: exclusive scan of the chunk counts -------------------------------------------------------------------------
constexpr int kScanThreads = 256, kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;

struct Triple { unsigned v, f, a; };
__device__ __forceinline__ Triple operator+(Triple p, Triple q) { return Triple{p.v + q.v, p.f + q.f, p.a + q.a}; }

__device__ __forceinline__ Triple warp_inclusive(Triple t, int lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned v = __shfl_up_sync(0xffffffffu, t.v, d), f = __shfl_up_sync(0xffffffffu, t.f, d), a = __shfl_up_sync(0xffffffffu, t.a, d);
        if (lane >= d) { t.v += v; t.f += f; t.a += a; }
    }
    return t;
}
// exclusive prefix of `mine` over the block (blockDim = kScanThreads); *total gets the block sum
__device__ __forceinline__ Triple block_exclusive(Triple mine, Triple* total) {
    __shared__ Triple s_warp[kScanThreads / 32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const Triple inc = warp_inclusive(mine, lane);
    if (lane == 31) s_warp[w] = inc;
    __syncthreads();
    Triple base{0, 0, 0}, sum{0, 0, 0};
#pragma unroll
    for (int i = 0; i < kScanThreads / 32; ++i) {
        if (i < w) base = base + s_warp[i];
        sum = sum + s_warp[i];
    }
    __syncthreads();
    *total = sum;
    return Triple{base.v + inc.v - mine.v, base.f + inc.f - mine.f, base.a + inc.a - mine.a};
}

__device__ __forceinline__ void load_counts8(const uint32_t* counts, long long first, long long n, unsigned (&c)[kScanItems]) {
    if (first + kScanItems <= n) {
        const uint4 a = *reinterpret_cast<const uint4*>(counts + first), b = *reinterpret_cast<const uint4*>(counts + first + 4);
        c[0] = a.x; c[1] = a.y; c[2] = a.z; c[3] = a.w; c[4] = b.x; c[5] = b.y; c[6] = b.z; c[7] = b.w;
    } else {
#pragma unroll
        for (int i = 0; i < kScanItems; ++i) c[i] = (first + i < n) ? counts[first + i] : 0u;
    }
}
__device__ __forceinline__ Triple unpack_count(unsigned c) { return Triple{c & 0xffu, (c >> 8) & 0xffu, c >> 16}; }

__global__ void __launch_bounds__(kScanThreads) k_scan_reduce(const uint32_t* __restrict__ counts, long long n, Triple* __restrict__ partials) {
    const long long first = ((long long)blockIdx.x * kScanThreads + threadIdx.x) * kScanItems;
    unsigned c[kScanItems];
    load_counts8(counts, first, n, c);
    Triple t{0, 0, 0};
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) t = t + unpack_count(c[i]);
    Triple total;
    block_exclusive(t, &total);
    if (threadIdx.x == 0) partials[blockIdx.x] = total;
}

// one block: partials[i] <- exclusive prefix; totals[0..2] <- grand totals {vertices, triangles, cut cubes}
__global__ void __launch_bounds__(kScanThreads) k_scan_partials(Triple* __restrict__ partials, int nblocks, unsigned* __restrict__ totals) {
    Triple carry{0, 0, 0};
    for (int base = 0; base < nblocks; base += kScanThreads) {
        const int i = base + threadIdx.x;
        const Triple mine = (i < nblocks) ? partials[i] : Triple{0, 0, 0};
        Triple total;
        const Triple ex = block_exclusive(mine, &total);
        if (i < nblocks) partials[i] = carry + ex;
        carry = carry + total;
    }
    if (threadIdx.x == 0) { totals[0] = carry.v; totals[1] = carry.f; totals[2] = carry.a; }
}

// offsets[c] = {first vertex, first triangle} of chunk c; cubes[] = the cut cubes in scan order, as chunk * 32 + lane
__global__ void __launch_bounds__(kScanThreads) k_scan_apply(const uint32_t* __restrict__ counts, long long n, const Triple* __restrict__ partials,
                                                            const uint32_t* __restrict__ lane_bits, uint2* __restrict__ offsets,
                                                            unsigned* __restrict__ cubes) {
    const long long first = ((long long)blockIdx.x * kScanThreads + threadIdx.x) * kScanItems;
    unsigned c[kScanItems];
    load_counts8(counts, first, n, c);
    Triple t{0, 0, 0};
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) t = t + unpack_count(c[i]);
    Triple total;
    Triple run = partials[blockIdx.x] + block_exclusive(t, &total);
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) {
        if (first + i < n) {
            offsets[first + i] = make_uint2(run.v, run.f);
            if (c[i] != 0u) {
                const uint32_t* lb = lane_bits + (first + i) * 5;
                unsigned act = lb[2] | lb[3] | lb[4];
                unsigned slot = run.a;
                while (act) {
                    const int l = __ffs(act) - 1;
                    act &= act - 1;
                    cubes[slot++] = (unsigned)((first + i) * 32 + l);
                }
            }
        }
        run = run + unpack_count(c[i]);
    }
}

// ---- K4 ---------------------------------------------------------------------------------------------------------
// rank of local edge `code` (0: edge 5, 1: edge 6, 2: edge 10) among the vertices a cube with OWNED byte `own`
// introduces, or -1 if that edge is not cut.
__device__ __forceinline__ int owned_rank(unsigned own, unsigned code) {
    const unsigned n = own & 3u;
    int r = -1;
    if (n > 0 && ((own >> 2) & 3u) == code) r = 0;
    if (n > 1 && ((own >> 4) & 3u) == code) r = 1;
    if (n > 2 && ((own >> 6) & 3u) == code) r = 2;
    return r;
}

struct EmitArgs {
    long long vbase;     // global index of the first vertex counted by the scan (ghost layer included)
    long long fbase;
    long long v_local0;  // scan offset of the first vertex this GPU stores (= ghost vertex count)
    long long f_local0;
};

// 34 sign bits of one row around chunk k: bit j <-> node x = 32k - 1 + j  (j = 0 is the halo node of the previous word)
__device__ __forceinline__ unsigned long long row_bits34(const uint32_t* __restrict__ row, int k) {
    const unsigned prev = (k > 0) ? (__ldg(row + k - 1) >> 31) : 1u;
    return (unsigned long long)prev | ((unsigned long long)__ldg(row + k) << 1) | ((unsigned long long)(__ldg(row + k + 1) & 1u) << 33);
}

// first vertex (scan numbering) of the cube at (chunk c, lane l): chunk prefix + vertices of the lanes before it
__device__ __forceinline__ unsigned cube_vertex_base(const uint2* __restrict__ offsets, const uint32_t* __restrict__ lane_bits, long long c, int l) {
    const unsigned lt = (1u << l) - 1u;
    return offsets[c].x + __popc(__ldg(lane_bits + c * 5) & lt) + 2u * __popc(__ldg(lane_bits + c * 5 + 1) & lt);
}

constexpr int kEmitThreads = 128;

// One thread per cut cube.  Its own new vertices (edges 5, 6, 10) are interpolated from the field; the other nine
// edges' vertices were introduced by six neighbouring cubes, whose vertex ranks follow from their chunk's prefix and
// lane bits and their own case code (for the order of their new vertices).
__global__ void __launch_bounds__(kEmitThreads) k_mc_emit(McDesc m, EmitArgs ea, const float* __restrict__ field, const uint32_t* __restrict__ signs,
                                                          const uint2* __restrict__ offsets, const uint32_t* __restrict__ lane_bits,
                                                          const unsigned* __restrict__ cubes, const unsigned* __restrict__ n_cubes_ptr,
                                                          float* __restrict__ verts, int* __restrict__ faces) {
    const unsigned n_cubes = *n_cubes_ptr;
    const int S = m.g.S;
    for (unsigned i = blockIdx.x * kEmitThreads + threadIdx.x; i < n_cubes; i += gridDim.x * kEmitThreads) {
        const unsigned id = cubes[i];
        const long long chunk = id >> 5;
        const int lane = (int)(id & 31u);
        const int k = (int)(chunk % m.K);
        const long long row = chunk / m.K;
        const int y = 1 + (int)(row % m.NX);
        const int z = m.scan_z0 + (int)(row / m.NX);
        if (z < m.emit_z0) continue;          // ghost layer: counted here, emitted by the slab below
        const int x = 32 * k + lane;
        unsigned long long R[3][3];           // sign rows (dy, dz) in {-1,0,1}^2, 34 bits each
#pragma unroll
        for (int dz = 0; dz < 3; ++dz)
#pragma unroll
            for (int dy = 0; dy < 3; ++dy) R[dz][dy] = row_bits34(sign_row(m, signs, y + dy - 1, z + dz - 1), k);
        // rows: 0 own (y,z)   1 B (y,z-1)   2 C (y-1,z)   3 A (y-1,z-1); the cube at x sits at bit position lane + 1
        const int rdy[4] = {1, 1, 0, 0}, rdz[4] = {1, 0, 1, 0};
        unsigned code[4], codep[3];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const unsigned long long a = R[rdz[r]][rdy[r]], b = R[rdz[r]][rdy[r] + 1], c = R[rdz[r] + 1][rdy[r]], d = R[rdz[r] + 1][rdy[r] + 1];
            code[r] = case_code(a, b, c, d, lane + 1);
            if (r < 3) codep[r] = (x > 1) ? case_code(a, b, c, d, lane) : 0u;      // the cube at x - 1 (x = 0 is never visited)
        }
        const unsigned code_self = code[0];
        const unsigned info_self = d_case_info[code_self];
        const unsigned nt = info_self & 0xffu, own_self = info_self >> 8, nv = own_self & 3u;
        const unsigned emask = ISD_EDGE_MASK(code_self);
        const long long vself = (long long)cube_vertex_base(offsets, lane_bits, chunk, lane);
        // ---- vertices introduced by this cube -----------------------------------------------------------------------
        if (nv) {
            const float fx = ((float)x + m.g.xi0) * m.g.wx, fy = ((float)y + m.g.yi0) * m.g.wy, fz = ((float)z + m.g.zi0) * m.g.wz;   // :1041-1053
            const float fx2 = fx + m.g.wx, fy2 = fy + m.g.wy, fz2 = fz + m.g.wz;                                                    // :570-576
            const float* p = field + (long long)(z - m.g.z0) * m.g.plane + (long long)y * m.g.pitch + x;
            const float f7 = __ldg(p + m.g.plane + m.g.pitch + 1);
            const long long vloc = vself - ea.v_local0;
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                if ((unsigned)j < nv) {
                    const unsigned e = (own_self >> (2 + 2 * j)) & 3u;
                    float px, py, pz;
                    if (e == 0u) {          // edge 5: VIntY(fx2, fy, fz2, field5, field7)  :627-634, :452-457
                        const float f5 = __ldg(p + m.g.plane + 1);
                        const float mu = (0.0f - f5) / (f7 - f5);
                        px = fx2; py = fy + mu * m.g.wy; pz = fz2;
                    } else if (e == 1u) {   // edge 6: VIntX(fx, fy2, fz2, field6, field7)  :636-643, :416-421
                        const float f6 = __ldg(p + m.g.plane + m.g.pitch);
                        const float mu = (0.0f - f6) / (f7 - f6);
                        px = fx + mu * m.g.wx; py = fy2; pz = fz2;
                    } else {                // edge 10: VIntZ(fx2, fy2, fz, field3, field7) :674-681, :478-483
                        const float f3 = __ldg(p + m.g.pitch + 1);
                        const float mu = (0.0f - f3) / (f7 - f3);
                        px = fx2; py = fy2; pz = fz + mu * m.g.wz;
                    }
                    float* o = verts + (vloc + j) * 3;
                    o[0] = px; o[1] = py; o[2] = pz;
                }
            }
        }
        if (nt == 0) continue;
        // ---- vertex index of every cut edge (SURVEY Appendix A + first-appearance cube) --------------------------------
        // neighbour cubes: chunk of row r at the same k; the cube at x-1 is lane-1 of it, or lane 31 of chunk k-1
        const long long crow[3] = {chunk, chunk_index(m, z - 1, y, k), chunk_index(m, z, y - 1, k)};      // own, B, C
        const long long cA = chunk_index(m, z - 1, y - 1, k);
        const int lp = (lane > 0) ? lane - 1 : 31;
        long long ev[12];
#pragma unroll
        for (int q = 0; q < 12; ++q) ev[q] = 0;
        auto vidx = [&](long long c, int l, unsigned cd, unsigned edge_code) -> long long {
            const unsigned own = d_case_info[cd] >> 8;
            return ea.vbase + (long long)cube_vertex_base(offsets, lane_bits, c, l) + owned_rank(own, edge_code);
        };
        ev[5] = ea.vbase + vself + owned_rank(own_self, 0u);
        ev[6] = ea.vbase + vself + owned_rank(own_self, 1u);
        ev[10] = ea.vbase + vself + owned_rank(own_self, 2u);
        if (emask & 0x001u) ev[0] = vidx(cA, lane, code[3], 1u);                    // x-edge at (x,y,z)    <- cube (x, y-1, z-1) edge 6
        if (emask & 0x006u) {                                                       // cube (x, y, z-1): edges 5 and 6
            const long long b = ea.vbase + (long long)cube_vertex_base(offsets, lane_bits, crow[1], lane);
            const unsigned own = d_case_info[code[1]] >> 8;
            ev[1] = b + owned_rank(own, 0u);                                        // y-edge at (x+1,y,z)
            ev[2] = b + owned_rank(own, 1u);                                        // x-edge at (x,y+1,z)
        }
        if (emask & 0x210u) {                                                       // cube (x, y-1, z): edges 6 and 10
            const long long b = ea.vbase + (long long)cube_vertex_base(offsets, lane_bits, crow[2], lane);
            const unsigned own = d_case_info[code[2]] >> 8;
            ev[4] = b + owned_rank(own, 1u);                                        // x-edge at (x,y,z+1)
            ev[9] = b + owned_rank(own, 2u);                                        // z-edge at (x+1,y,z)
        }
        const long long dprev = (lane > 0) ? 0 : -1;                                // chunk k-1 of the same row
        if (emask & 0x008u) ev[3] = vidx(crow[1] + dprev, lp, codep[1], 0u);        // y-edge at (x,y,z)    <- cube (x-1, y, z-1) edge 5
        if (emask & 0x100u) ev[8] = vidx(crow[2] + dprev, lp, codep[2], 2u);        // z-edge at (x,y,z)    <- cube (x-1, y-1, z) edge 10
        if (emask & 0x880u) {                                                       // cube (x-1, y, z): edges 5 and 10
            const long long b = ea.vbase + (long long)cube_vertex_base(offsets, lane_bits, crow[0] + dprev, lp);
            const unsigned own = d_case_info[codep[0]] >> 8;
            ev[7] = b + owned_rank(own, 0u);                                        // y-edge at (x,y,z+1)
            ev[11] = b + owned_rank(own, 2u);                                       // z-edge at (x,y+1,z)
        }
        // ---- triangles ---------------------------------------------------------------------------------------------------
        const unsigned ltm = (1u << lane) - 1u;
        const uint32_t* lb = lane_bits + chunk * 5;
        const unsigned tbefore = __popc(__ldg(lb + 2) & ltm) + 2u * __popc(__ldg(lb + 3) & ltm) + 4u * __popc(__ldg(lb + 4) & ltm);
        const long long floc = (long long)offsets[chunk].y + tbefore - ea.f_local0;
        const int8_t* tri = d_tri_table + code_self * 16;
        for (unsigned t = 0; t < nt; ++t) {
            int* o = faces + (floc + t) * 3;
#pragma unroll
            for (int c3 = 0; c3 < 3; ++c3) {
                const int e = tri[t * 3 + c3];
                long long idx = 0;
#pragma unroll
                for (int q = 0; q < 12; ++q)
                    if (e == q) idx = ev[q];
                o[c3] = (int)idx;
            }
        }
    }
}

}  // namespace isd
