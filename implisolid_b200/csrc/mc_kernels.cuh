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

__global__ void __launch_bounds__(256) k_mc_classify(McDesc m, const uint32_t* __restrict__ signs, uint16_t* __restrict__ counts,
                                                     uint8_t* __restrict__ dbg_cases) {
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
        unsigned packed = 0;
        if (code != 0u && code != 255u) {
            const unsigned info = d_case_info[code];
            packed = ((info >> 8) & 3u) | ((info & 0xffu) << 8);
        }
        const unsigned total = __reduce_add_sync(0xffffffffu, packed);
        if (lane == src) my_count = total;
        if (dbg_cases != nullptr && xin)
            dbg_cases[((long long)(sz - m.scan_z0) * m.NX + (sy - 1)) * m.NX + (x - 1)] = (uint8_t)code;
    }
    if (valid) counts[c] = (uint16_t)my_count;
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

__device__ __forceinline__ void load_counts8(const uint16_t* counts, long long first, long long n, unsigned (&c)[kScanItems]) {
    if (first + kScanItems <= n) {
        const uint4 raw = *reinterpret_cast<const uint4*>(counts + first);    // first is a multiple of 8: 16-byte aligned
        c[0] = raw.x & 0xffffu; c[1] = raw.x >> 16; c[2] = raw.y & 0xffffu; c[3] = raw.y >> 16;
        c[4] = raw.z & 0xffffu; c[5] = raw.z >> 16; c[6] = raw.w & 0xffffu; c[7] = raw.w >> 16;
    } else {
#pragma unroll
        for (int i = 0; i < kScanItems; ++i) c[i] = (first + i < n) ? counts[first + i] : 0u;
    }
}

__global__ void __launch_bounds__(kScanThreads) k_scan_reduce(const uint16_t* __restrict__ counts, long long n, Triple* __restrict__ partials) {
    const long long first = ((long long)blockIdx.x * kScanThreads + threadIdx.x) * kScanItems;
    unsigned c[kScanItems];
    load_counts8(counts, first, n, c);
    Triple t{0, 0, 0};
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) { t.v += c[i] & 0xffu; t.f += c[i] >> 8; t.a += (c[i] != 0u); }
    Triple total;
    block_exclusive(t, &total);
    if (threadIdx.x == 0) partials[blockIdx.x] = total;
}

// one block: partials[i] <- exclusive prefix; totals[0..2] <- grand totals
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

__global__ void __launch_bounds__(kScanThreads) k_scan_apply(const uint16_t* __restrict__ counts, long long n, const Triple* __restrict__ partials,
                                                            uint2* __restrict__ offsets, unsigned* __restrict__ active) {
    const long long first = ((long long)blockIdx.x * kScanThreads + threadIdx.x) * kScanItems;
    unsigned c[kScanItems];
    load_counts8(counts, first, n, c);
    Triple t{0, 0, 0};
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) { t.v += c[i] & 0xffu; t.f += c[i] >> 8; t.a += (c[i] != 0u); }
    Triple total;
    Triple run = partials[blockIdx.x] + block_exclusive(t, &total);
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) {
        if (first + i < n) {
            offsets[first + i] = make_uint2(run.v, run.f);
            if (c[i] != 0u) active[run.a] = (unsigned)(first + i);
        }
        run.v += c[i] & 0xffu; run.f += c[i] >> 8; run.a += (c[i] != 0u);
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

constexpr int kEmitWarps = 8;

// 34 sign bits of one row around chunk k: bit j <-> node x = 32k - 1 + j  (j = 0 is the halo node of the previous word)
__device__ __forceinline__ unsigned long long row_bits34(const uint32_t* __restrict__ row, int k) {
    const unsigned prev = (k > 0) ? (__ldg(row + k - 1) >> 31) : 1u;
    return (unsigned long long)prev | ((unsigned long long)__ldg(row + k) << 1) | ((unsigned long long)(__ldg(row + k + 1) & 1u) << 33);
}

__global__ void __launch_bounds__(kEmitWarps * 32) k_mc_emit(McDesc m, EmitArgs ea, const float* __restrict__ field, const uint32_t* __restrict__ signs,
                                                             const uint2* __restrict__ offsets, const unsigned* __restrict__ active,
                                                             const unsigned* __restrict__ n_active_ptr, float* __restrict__ verts, int* __restrict__ faces) {
    const unsigned n_active = *n_active_ptr;
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    const int S = m.g.S;
    for (unsigned slot = blockIdx.x * kEmitWarps + w; slot < n_active; slot += gridDim.x * kEmitWarps) {
        const long long chunk = active[slot];
        const int k = (int)(chunk % m.K);
        const long long row = chunk / m.K;
        const int y = 1 + (int)(row % m.NX);
        const int z = m.scan_z0 + (int)(row / m.NX);
        if (z < m.emit_z0) continue;          // ghost layer: counted here, emitted by the slab below
        const int x = 32 * k + lane;
        // sign rows (dy, dz) in {-1,0,1}^2, 34 bits each (warp-uniform loads)
        unsigned long long R[3][3];
#pragma unroll
        for (int dz = 0; dz < 3; ++dz)
#pragma unroll
            for (int dy = 0; dy < 3; ++dy) R[dz][dy] = row_bits34(sign_row(m, signs, y + dy - 1, z + dz - 1), k);
        const bool xin = (x >= 1 && x <= S - 3);
        // rows: 0 own (y,z)   1 B (y,z-1)   2 C (y-1,z)   3 A (y-1,z-1); cube at lane -> bit position lane + 1
        const int rdy[4] = {1, 1, 0, 0}, rdz[4] = {1, 0, 1, 0};
        unsigned own[4], own_halo[3];
        long long base[4];
        unsigned code_self = 0;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int yy = y + rdy[r] - 1, zz = z + rdz[r] - 1;
            const unsigned long long a = R[rdz[r]][rdy[r]], b = R[rdz[r]][rdy[r] + 1], c = R[rdz[r] + 1][rdy[r]], d = R[rdz[r] + 1][rdy[r] + 1];
            const unsigned code = xin ? case_code(a, b, c, d, lane + 1) : 0u;
            if (r == 0) code_self = code;
            const unsigned info = (code != 0u && code != 255u) ? d_case_info[code] : 0u;
            own[r] = info >> 8;
            const unsigned n = own[r] & 3u;
            const unsigned b0 = __ballot_sync(0xffffffffu, n & 1u), b1 = __ballot_sync(0xffffffffu, n & 2u);
            const unsigned before = __popc(b0 & lt) + 2u * __popc(b1 & lt);
            long long coff = 0;
            if (yy >= 1 && zz >= m.scan_z0) coff = offsets[chunk_index(m, zz, yy, k)].x;
            base[r] = ea.vbase + coff + before;
            if (r < 3) {
                // halo cube x = 32k - 1 of the same row (introducer of this chunk's lane-0 edges 3/7/8/11)
                const unsigned hcode = case_code(a, b, c, d, 0);
                const unsigned hinfo = (hcode != 0u && hcode != 255u && k > 0) ? d_case_info[hcode] : 0u;
                own_halo[r] = hinfo >> 8;
            }
        }
        long long i5[3], i6[4], i10[3], p5[3], p10[3];      // p*: the cube at x-1
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            i6[r] = base[r] + owned_rank(own[r], 1u);
            if (r < 3) {
                i5[r] = base[r] + owned_rank(own[r], 0u);
                i10[r] = base[r] + owned_rank(own[r], 2u);
                long long q5 = __shfl_up_sync(0xffffffffu, i5[r], 1), q10 = __shfl_up_sync(0xffffffffu, i10[r], 1);
                if (lane == 0) {
                    const long long hbase = base[r] - (own_halo[r] & 3u);     // the halo cube's vertices end where this chunk's begin
                    q5 = hbase + owned_rank(own_halo[r], 0u);
                    q10 = hbase + owned_rank(own_halo[r], 2u);
                }
                p5[r] = q5; p10[r] = q10;
            }
        }
        const unsigned info_self = (code_self != 0u && code_self != 255u) ? d_case_info[code_self] : 0u;
        const unsigned nt = info_self & 0xffu;
        // ---- vertices introduced by this cube ---------------------------------------------------------------------
        const unsigned nv = own[0] & 3u;
        if (nv) {
            const float fx = ((float)x + m.g.xi0) * m.g.wx, fy = ((float)y + m.g.yi0) * m.g.wy, fz = ((float)z + m.g.zi0) * m.g.wz;   // :1041-1053
            const float fx2 = fx + m.g.wx, fy2 = fy + m.g.wy, fz2 = fz + m.g.wz;                                                    // :570-576
            const float* p = field + (long long)(z - m.g.z0) * m.g.plane + (long long)y * m.g.pitch + x;
            const float f7 = __ldg(p + m.g.plane + m.g.pitch + 1);
            const long long vloc = base[0] - ea.vbase - ea.v_local0;
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                if ((unsigned)j < nv) {
                    const unsigned e = (own[0] >> (2 + 2 * j)) & 3u;
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
        // ---- triangles ---------------------------------------------------------------------------------------------------
        const unsigned t0 = __ballot_sync(0xffffffffu, nt & 1u), t1 = __ballot_sync(0xffffffffu, nt & 2u), t2 = __ballot_sync(0xffffffffu, nt & 4u);
        if (nt) {
            const unsigned tbefore = __popc(t0 & lt) + 2u * __popc(t1 & lt) + 4u * __popc(t2 & lt);
            const long long floc = (long long)offsets[chunk].y + tbefore - ea.f_local0;
            long long ev[12];            // vertex index per local edge (SURVEY Appendix A + first-appearance cube)
            ev[0] = i6[3];   // x-edge at (x,y,z)      <- cube (x, y-1, z-1) edge 6
            ev[1] = i5[1];   // y-edge at (x+1,y,z)    <- cube (x, y, z-1)   edge 5
            ev[2] = i6[1];   // x-edge at (x,y+1,z)    <- cube (x, y, z-1)   edge 6
            ev[3] = p5[1];   // y-edge at (x,y,z)      <- cube (x-1, y, z-1) edge 5
            ev[4] = i6[2];   // x-edge at (x,y,z+1)    <- cube (x, y-1, z)   edge 6
            ev[5] = i5[0];
            ev[6] = i6[0];
            ev[7] = p5[0];   // y-edge at (x,y,z+1)    <- cube (x-1, y, z)   edge 5
            ev[8] = p10[2];  // z-edge at (x,y,z)      <- cube (x-1, y-1, z) edge 10
            ev[9] = i10[2];  // z-edge at (x+1,y,z)    <- cube (x, y-1, z)   edge 10
            ev[10] = i10[0];
            ev[11] = p10[0]; // z-edge at (x,y+1,z)    <- cube (x-1, y, z)   edge 10
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
}

}  // namespace isd
