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
// never touches the 4 B/node field; and a cube can only be cut if its home node (x,y,z) lies in a work-list tile of K1
// (a tile proven inside or outside over its box grown by one cell has no cut cube), so every pass below walks the work
// list, never the grid — work and memory are proportional to the surface:
//   K2 mc_classify_tiles : one 64-thread block per row of tiles (ty,tz) that has work-list tiles, one thread per row of
//                    cubes (y,z) in it.  The block walks its work-list tiles in x order; a thread reads the 4x9 sign bits
//                    around its 8-cube segment, turns the 8 case codes into counts and bit planes (new vertices: 2 bits,
//                    triangles: 3 bits per cube) and keeps a running total along its row, which is the segment's offset
//                    in the row.  Row totals need no atomics: one thread owns a row.
//   K3 row_scan    : exclusive prefix of the row totals in scan order (single pass, decoupled look-back); seg_apply makes
//                    the segment offsets absolute and writes the ordered list of cut cubes.
//   K4 mc_emit     : one thread per cut cube; case codes of its own cube and of the six cubes that introduced its other
//                    edges come from the sign bits, popcounts of their segments' bit planes become vertex ranks; field
//                    values are read only at the <= 4 corners each new vertex interpolates between.
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
    long long n_rows;       // (zc_end - scan_z0) * NX rows of cubes, numbered in scan order
    long long ghost_rows;   // rows of the ghost layer (0 or NX)
};

__device__ __forceinline__ long long row_index(const McDesc& m, int y, int z) { return (long long)(z - m.scan_z0) * m.NX + (y - 1); }
__device__ __forceinline__ const uint32_t* sign_row(const McDesc& m, const uint32_t* signs, int y, int z) {
    return signs + ((long long)(z - m.g.z0) * m.g.S + y) * m.g.SW;
}
// sign bits of nodes x0 .. x0+3 of one row (x0 >= 0; the row is padded with a whole word)
__device__ __forceinline__ unsigned sign_bits4(const uint32_t* __restrict__ row, int x0) {
    const int w = x0 >> 5;
    return __funnelshift_r(__ldg(row + w), __ldg(row + w + 1), x0 & 31) & 15u;
}
// sign bits of nodes 8*tx .. 8*tx+8 (9 bits) of one row
__device__ __forceinline__ unsigned sign_bits9(const uint32_t* __restrict__ row, int tx) {
    const uint8_t* b = reinterpret_cast<const uint8_t*>(row);
    return (unsigned)__ldg(b + tx) | (((unsigned)__ldg(b + tx + 1) & 1u) << 8);
}
// case code of the cube at bit position `b` given the four sign rows (y,z) (y+1,z) (y,z+1) (y+1,z+1); :553-560
__device__ __forceinline__ unsigned case_code(unsigned r00, unsigned r10, unsigned r01, unsigned r11, int b) {
    const unsigned a = (r00 >> b) & 3u, c = (r10 >> b) & 3u, d = (r01 >> b) & 3u, e = (r11 >> b) & 3u;
    // bits: f0 (x,y,z)->1  f1 (x+1,y,z)->2  f3 (x+1,y+1,z)->4  f2 (x,y+1,z)->8  f4->16 f5->32 f7->64 f6->128
    return (a & 1u) | (a & 2u) | ((c & 2u) << 1) | ((c & 1u) << 3) | ((d & 1u) << 4) | ((d & 2u) << 4) | ((e & 2u) << 5) | ((e & 1u) << 7);
}

struct Triple { unsigned v, f, a; };
__device__ __forceinline__ Triple operator+(Triple p, Triple q) { return Triple{p.v + q.v, p.f + q.f, p.a + q.a}; }

// What K1 leaves for marching cubes (field_kernel.cuh, k_tile_bounds) and what the passes share.  A SEGMENT is the 8 cubes
// of one row (y,z) whose home nodes lie in one work-list tile; segment index = slot * 64 + (lz * 8 + ly).
//   seg_bits : bits 0-7 / 8-15 the two bit planes of the cubes' new-vertex counts, 16-23 / 24-31 / 32-39 the three bit planes
//              of their triangle counts, bits 40-63 the cut cubes before the segment in its row
//   seg_off  : {vertices, triangles} before the segment — in its row after K2, absolute after seg_apply
struct McBuffers {
    const unsigned* list;          // work-list: tile index per slot
    const unsigned* counters;      // [0] work-list length
    const unsigned* tile_slot;     // tile index -> slot (valid for work-list tiles)
    const uint8_t* tile_any;       // [tz][ty]: the row of tiles has work-list tiles
    const uint32_t* tile_bits;     // [tz][ty][TW]: bit per tile
    int TW;
    unsigned long long* seg_bits;
    uint2* seg_off;
    Triple* row_tot;               // per row of cubes (written for rows inside rows of tiles that have work)
    Triple* row_base;              // exclusive prefix of row_tot in scan order
};

__device__ __forceinline__ unsigned planes_v(unsigned long long sb, unsigned below) {      // new vertices of the cubes selected by `below`
    return __popc((unsigned)sb & 0xffu & below) + 2u * __popc((unsigned)(sb >> 8) & 0xffu & below);
}
__device__ __forceinline__ unsigned planes_t(unsigned long long sb, unsigned below) {
    return __popc((unsigned)(sb >> 16) & 0xffu & below) + 2u * __popc((unsigned)(sb >> 24) & 0xffu & below) +
           4u * __popc((unsigned)(sb >> 32) & 0xffu & below);
}
__device__ __forceinline__ unsigned planes_cut(unsigned long long sb) {      // bit per cube: has triangles
    return ((unsigned)(sb >> 16) | (unsigned)(sb >> 24) | (unsigned)(sb >> 32)) & 0xffu;
}

// ---- K2 ---------------------------------------------------------------------------------------------------------
// Block = one row of tiles (ty,tz); 64 * kClassifyTiles threads = kClassifyTiles tiles x 64 rows of cubes.  The block lists its work-list tiles in x
// order, then takes them four at a time: every thread classifies one 8-cube segment, the segment totals of the group meet in
// shared memory, and each row's running total (kept by all four threads of that row) advances by the group.
// (Small grids have few rows of tiles and many tiles in each: four tiles at a time; large grids the opposite: one.)
constexpr int kMaxTilesPerRow = 1024;

template <int kClassifyTiles>
__global__ void __launch_bounds__(64 * kClassifyTiles) k_mc_classify_tiles(McDesc m, McBuffers mb, const uint32_t* __restrict__ signs) {
    const int trow = blockIdx.x;
    if (!mb.tile_any[trow]) return;
    __shared__ uint16_t s_tiles[kMaxTilesPerRow];
    __shared__ int s_ntiles;
    __shared__ Triple s_tot[kClassifyTiles][64];
    if (threadIdx.x < 32) {      // ordered list of the row's work-list tiles (one lane per 32 tiles, popcount ranks)
        int base = 0;
        for (int w0 = 0; w0 < mb.TW; w0 += 32) {
            const int w = w0 + (int)threadIdx.x;
            unsigned bits = (w < mb.TW) ? __ldg(mb.tile_bits + (long long)trow * mb.TW + w) : 0u;
            int cnt = __popc(bits), inc = cnt;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int o = __shfl_up_sync(0xffffffffu, inc, d);
                if ((int)threadIdx.x >= d) inc += o;
            }
            int at = base + inc - cnt;
            while (bits) {
                s_tiles[at++] = (uint16_t)(32 * w + __ffs(bits) - 1);
                bits &= bits - 1;
            }
            base += __shfl_sync(0xffffffffu, inc, 31);
        }
        if (threadIdx.x == 0) s_ntiles = base;
    }
    __syncthreads();
    const int ntiles = s_ntiles;
    const int ty = trow % m.g.ty, tz = trow / m.g.ty;
    const int r = threadIdx.x & 63, j = threadIdx.x >> 6;
    const int ly = r & 7, lz = r >> 3;
    const int y = 8 * ty + ly, z = m.g.z0 + 8 * tz + lz;
    const bool row_valid = y >= 1 && y <= m.g.S - 3 && z >= m.scan_z0 && z < m.zc_end;
    const uint32_t *s00 = nullptr, *s10 = nullptr, *s01 = nullptr, *s11 = nullptr;
    if (row_valid) {
        s00 = sign_row(m, signs, y, z); s10 = sign_row(m, signs, y + 1, z);
        s01 = sign_row(m, signs, y, z + 1); s11 = sign_row(m, signs, y + 1, z + 1);
    }
    Triple run{0, 0, 0};
    for (int i0 = 0; i0 < ntiles; i0 += kClassifyTiles) {
        const int i = i0 + j;
        unsigned long long sb = 0ull;
        long long seg = -1;
        Triple tot{0, 0, 0};
        if (i < ntiles) {
            const int tx = s_tiles[i];
            seg = (long long)__ldg(mb.tile_slot + (long long)trow * m.g.tx + tx) * 64 + r;
            if (row_valid) {
                const unsigned r00 = sign_bits9(s00, tx), r10 = sign_bits9(s10, tx), r01 = sign_bits9(s01, tx), r11 = sign_bits9(s11, tx);
                if ((r00 | r10 | r01 | r11) != 0u && (r00 & r10 & r01 & r11) != 0x1ffu) {
                    unsigned v0 = 0, v1 = 0, t0 = 0, t1 = 0, t2 = 0;
#pragma unroll
                    for (int l = 0; l < 8; ++l) {
                        const int x = 8 * tx + l;
                        const bool xin = x >= 1 && x <= m.g.S - 3;
                        const unsigned code = xin ? case_code(r00, r10, r01, r11, l) : 0u;
                        if (code != 0u && code != 255u) {
                            const unsigned info = d_case_info[code];
                            const unsigned nv = (info >> 8) & 3u, nt = info & 0xffu;
                            v0 |= (nv & 1u) << l; v1 |= ((nv >> 1) & 1u) << l;
                            t0 |= (nt & 1u) << l; t1 |= ((nt >> 1) & 1u) << l; t2 |= ((nt >> 2) & 1u) << l;
                        }
                    }
                    sb = (unsigned long long)(v0 | (v1 << 8) | (t0 << 16) | (t1 << 24)) | ((unsigned long long)t2 << 32);
                    tot = Triple{planes_v(sb, 0xffu), planes_t(sb, 0xffu), (unsigned)__popc(planes_cut(sb))};
                }
            }
        }
        s_tot[j][r] = tot;
        __syncthreads();
        Triple before = run, all = run;
#pragma unroll
        for (int q = 0; q < kClassifyTiles; ++q) {
            const Triple t = s_tot[q][r];
            if (q < j) before = before + t;
            all = all + t;
        }
        run = all;
        if (seg >= 0) {
            if (sb) {
                mb.seg_off[seg] = make_uint2(before.v, before.f);
                sb |= (unsigned long long)before.a << 40;
            }
            mb.seg_bits[seg] = sb;
        }
        __syncthreads();
    }
    if (row_valid && j == 0) mb.row_tot[row_index(m, y, z)] = run;
}

// Debug only (isd_get_case_codes): the dense array of case codes, straight from the sign bits.
__global__ void __launch_bounds__(256) k_debug_cases(McDesc m, const uint32_t* __restrict__ signs, uint8_t* __restrict__ dbg_cases) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i >= m.n_rows * m.NX) return;
    const int x = 1 + (int)(i % m.NX);
    const long long row = i / m.NX;
    const int y = 1 + (int)(row % m.NX), z = m.scan_z0 + (int)(row / m.NX);
    const unsigned r00 = sign_bits4(sign_row(m, signs, y, z), x), r10 = sign_bits4(sign_row(m, signs, y + 1, z), x);
    const unsigned r01 = sign_bits4(sign_row(m, signs, y, z + 1), x), r11 = sign_bits4(sign_row(m, signs, y + 1, z + 1), x);
    dbg_cases[i] = (uint8_t)case_code(r00, r10, r01, r11, 0);
}

// ---- K3 ---------------------------------------------------------------------------------------------------------
// Exclusive prefix of the row totals in scan order, one launch of at most kRowScanMaxBlocks blocks: a block takes a ticket,
// sums its contiguous range of rows, publishes the sum, waits for the sums of the tickets before it (all of them at once, one
// thread per predecessor — an earlier ticket is always running or done, so the wait cannot deadlock), then scans its range
// again from that prefix.  `scan_state` = {ticket, pad...} followed by {ready, v, f, a} per block, zeroed before the launch.
// totals[0..2] <- grand totals, totals[4..5] <- vertices / triangles before row `ghost_rows`.
constexpr int kRowScanThreads = 512, kRowScanMaxBlocks = 296;

__device__ __forceinline__ Triple warp_inclusive(Triple t, int lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned v = __shfl_up_sync(0xffffffffu, t.v, d), f = __shfl_up_sync(0xffffffffu, t.f, d), a = __shfl_up_sync(0xffffffffu, t.a, d);
        if (lane >= d) { t.v += v; t.f += f; t.a += a; }
    }
    return t;
}
// exclusive prefix of `mine` over the block (kRowScanThreads threads); *total gets the block sum
__device__ __forceinline__ Triple block_exclusive(Triple mine, Triple* total, Triple* s_warp) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const Triple inc = warp_inclusive(mine, lane);
    if (lane == 31) s_warp[w] = inc;
    __syncthreads();
    Triple base{0, 0, 0}, sum{0, 0, 0};
#pragma unroll
    for (int i = 0; i < kRowScanThreads / 32; ++i) {
        if (i < w) base = base + s_warp[i];
        sum = sum + s_warp[i];
    }
    __syncthreads();
    *total = sum;
    return Triple{base.v + inc.v - mine.v, base.f + inc.f - mine.f, base.a + inc.a - mine.a};
}
__device__ __forceinline__ Triple row_total(const McDesc& m, const McBuffers& mb, long long row) {
    if (row >= m.n_rows) return Triple{0, 0, 0};
    const int y = 1 + (int)(row % m.NX), z = m.scan_z0 + (int)(row / m.NX);
    if (!mb.tile_any[((z - m.g.z0) >> 3) * m.g.ty + (y >> 3)]) return Triple{0, 0, 0};      // never written: no work-list tile there
    return mb.row_tot[row];
}

__global__ void __launch_bounds__(kRowScanThreads) k_row_scan(McDesc m, McBuffers mb, long long rows_per_block, unsigned* __restrict__ scan_state,
                                                              unsigned* __restrict__ totals) {
    __shared__ Triple s_warp[kRowScanThreads / 32];
    __shared__ unsigned s_ticket;
    if (threadIdx.x == 0) s_ticket = atomicAdd(scan_state, 1u);
    __syncthreads();
    const unsigned blk = s_ticket;
    const long long first = (long long)blk * rows_per_block, last = first + rows_per_block;      // rows_per_block % kRowScanThreads == 0
    Triple mine{0, 0, 0}, sum;
    for (long long r0 = first; r0 < last; r0 += kRowScanThreads) mine = mine + row_total(m, mb, r0 + threadIdx.x);
    block_exclusive(mine, &sum, s_warp);
    volatile unsigned* state = scan_state + 4;
    if (threadIdx.x == 0) {
        state[blk * 4 + 1] = sum.v; state[blk * 4 + 2] = sum.f; state[blk * 4 + 3] = sum.a;
        __threadfence();
        state[blk * 4] = 1u;
    }
    Triple before{0, 0, 0};
    for (unsigned j = threadIdx.x; j < blk; j += kRowScanThreads)
        while (state[j * 4] == 0u) {}
    __threadfence();      // one fence after all the flags have been seen, then the sums
    for (unsigned j = threadIdx.x; j < blk; j += kRowScanThreads) before = before + Triple{state[j * 4 + 1], state[j * 4 + 2], state[j * 4 + 3]};
    Triple carry;
    block_exclusive(before, &carry, s_warp);
    for (long long r0 = first; r0 < last; r0 += kRowScanThreads) {
        const long long row = r0 + threadIdx.x;
        const Triple t = row_total(m, mb, row);
        Triple step;
        const Triple ex = block_exclusive(t, &step, s_warp);
        const Triple base = carry + ex;
        if (row < m.n_rows) mb.row_base[row] = base;
        if (m.ghost_rows > 0 && row == m.ghost_rows) { totals[4] = base.v; totals[5] = base.f; }
        if (row == m.n_rows - 1) { totals[0] = base.v + t.v; totals[1] = base.f + t.f; totals[2] = base.a + t.a; }
        carry = carry + step;
    }
}

// One thread per segment of the work list: offsets become absolute, and the cut cubes go, in scan order, into cubes[]
// as segment * 8 + position.
//
// Faces grouped by tile (`perm`, optional).  The OB02 kernels evaluate the tree at the faces' centroids through per-tile
// instruction lists, two faces per thread; in face order (scan order: z, y, x) a warp's 64 faces lie in four or five tiles and its
// lanes run as many different lists.  A segment's faces are one contiguous run of face indices, and a tile is 64 consecutive
// segments = 64 consecutive threads here: the block adds up each tile's face count, one thread reserves the tile's (even) number
// of slots with ONE atomic, and every segment writes its run of local face indices behind those of the tile's earlier segments.
// perm then lists every face exactly once, tile by tile (tiles in no particular order — results are per face and do not depend
// on it), padded with -1 to an even count per tile so that the pairs (2t, 2t + 1) the kernels take never straddle two tiles.
struct PermArgs {
    int* perm;                         // nullptr: not wanted
    unsigned* cursor;                  // slots handed out so far (zeroed before the launch); the final value is perm's length
    const unsigned* ghost_faces;       // triangles of the ghost layer (counted by the scan, not emitted), or nullptr
};
__global__ void __launch_bounds__(256) k_seg_apply(McDesc m, McBuffers mb, unsigned long long* __restrict__ cubes, PermArgs pa) {
    __shared__ unsigned s_cnt[8];
    __shared__ unsigned s_base[4];
    const long long n_seg = (long long)mb.counters[0] * 64;
    const unsigned ghost_f = (pa.perm != nullptr && pa.ghost_faces != nullptr) ? *pa.ghost_faces : 0u;
    for (long long seg0 = (long long)blockIdx.x * 256; seg0 < n_seg; seg0 += (long long)gridDim.x * 256) {
        const long long seg = seg0 + threadIdx.x;      // n_seg is a multiple of 64: a tile's segments are all in or all out
        unsigned n_faces = 0u, face0 = 0u;
        if (seg < n_seg) {
            const unsigned long long sb = mb.seg_bits[seg];
            unsigned cut = planes_cut(sb);
            if ((sb & 0xffffffffffull) != 0ull) {
                const int t = (int)mb.list[seg >> 6], r = (int)(seg & 63);
                const int ty = (t / m.g.tx) % m.g.ty, tzl = t / (m.g.tx * m.g.ty);
                const int y = 8 * ty + (r & 7), z = m.g.z0 + 8 * tzl + (r >> 3);
                const Triple base = mb.row_base[row_index(m, y, z)];
                const uint2 rel = mb.seg_off[seg];
                mb.seg_off[seg] = make_uint2(base.v + rel.x, base.f + rel.y);
                unsigned long long slot = (unsigned long long)base.a + (sb >> 40);
                while (cut) {
                    const int l = __ffs(cut) - 1;
                    cut &= cut - 1;
                    cubes[slot++] = (unsigned long long)seg * 8ull + (unsigned)l;
                }
                if (pa.perm != nullptr && z >= m.emit_z0) { n_faces = planes_t(sb, 0xffu); face0 = base.f + rel.y - ghost_f; }
            }
        }
        if (pa.perm != nullptr) {      // uniform branch
            const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, tile = w >> 1;      // a tile = two warps
            unsigned inc = n_faces;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const unsigned o = __shfl_up_sync(0xffffffffu, inc, d);
                if (lane >= d) inc += o;
            }
            if (lane == 31) s_cnt[w] = inc;
            __syncthreads();
            const unsigned before = inc - n_faces + ((w & 1) ? s_cnt[w - 1] : 0u);
            if ((threadIdx.x & 63) == 0) {
                const unsigned total = s_cnt[2 * tile] + s_cnt[2 * tile + 1];
                s_base[tile] = total ? atomicAdd(pa.cursor, (total + 1u) & ~1u) : 0u;
            }
            __syncthreads();
            if (n_faces) {
                int* dst = pa.perm + s_base[tile] + before;
                for (unsigned j = 0; j < n_faces; ++j) dst[j] = (int)(face0 + j);
            }
            __syncthreads();
        }
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
    int ghost_vertices;  // halo mode: the ghost layer's cubes write their vertices too (positions only; their faces belong to the slab below)
};

// first vertex (scan numbering) of the cube at (x, y, z): its segment's offset + the new vertices of the cubes before it
__device__ __forceinline__ long long cube_vertex_base(const McDesc& m, const McBuffers& mb, int x, int y, int z) {
    const int zl = z - m.g.z0;
    const long long tile = ((long long)(zl >> 3) * m.g.ty + (y >> 3)) * m.g.tx + (x >> 3);
    const long long seg = (long long)__ldg(mb.tile_slot + tile) * 64 + ((zl & 7) * 8 + (y & 7));
    return (long long)mb.seg_off[seg].x + planes_v(mb.seg_bits[seg], (1u << (x & 7)) - 1u);
}

constexpr int kEmitThreads = 128;

// One thread per cut cube.  Its own new vertices (edges 5, 6, 10) are interpolated from the field; the other nine
// edges' vertices were introduced by six neighbouring cubes, whose vertex ranks follow from their segment's offset and
// bit planes and their own case code (for the order of their new vertices).
__global__ void __launch_bounds__(kEmitThreads) k_mc_emit(McDesc m, McBuffers mb, EmitArgs ea, const float* __restrict__ field,
                                                          const uint32_t* __restrict__ signs, const unsigned long long* __restrict__ cubes,
                                                          const unsigned* __restrict__ n_cubes_ptr, float* __restrict__ verts, int* __restrict__ faces) {
    const unsigned n_cubes = *n_cubes_ptr;
    for (unsigned i = blockIdx.x * kEmitThreads + threadIdx.x; i < n_cubes; i += gridDim.x * kEmitThreads) {
        const unsigned long long id = cubes[i];
        const long long seg = (long long)(id >> 3);
        const int l = (int)(id & 7ull), r = (int)(seg & 63);
        const int t = (int)mb.list[seg >> 6];
        const int tx = t % m.g.tx, ty = (t / m.g.tx) % m.g.ty, tzl = t / (m.g.tx * m.g.ty);
        const int x = 8 * tx + l, y = 8 * ty + (r & 7), z = m.g.z0 + 8 * tzl + (r >> 3);
        // Ghost layer: counted here, emitted by the slab below.  In halo mode its VERTICES are still written (the local mesh keeps
        // them in its first slots): the lowest halo layer's faces hang on them, and although nothing the slab exports depends on
        // those faces, the OB02 kernels evaluate the tree at their centroids — with the slots left at zero those were giant
        // triangles to the origin whose every evaluation ran the whole tree outside the work-list tiles and failed every search
        // direction (100 us per evaluation in a warp of its own: it dominated the sharded runs' projection).
        const bool ghost = z < m.emit_z0;
        if (ghost && !ea.ghost_vertices) continue;
        unsigned R[3][3];                     // sign rows (dy, dz) in {-1,0,1}^2: bit j <-> node x - 1 + j
#pragma unroll
        for (int dz = 0; dz < 3; ++dz)
#pragma unroll
            for (int dy = 0; dy < 3; ++dy)      // a ghost cube only needs its own case code: the plane below it is not resident
                R[dz][dy] = (ghost && dz == 0) ? 0u : sign_bits4(sign_row(m, signs, y + dy - 1, z + dz - 1), x - 1);
        // rows: 0 own (y,z)   1 B (y,z-1)   2 C (y-1,z)   3 A (y-1,z-1); the cube at x sits at bit position 1
        const int rdy[4] = {1, 1, 0, 0}, rdz[4] = {1, 0, 1, 0};
        unsigned code[4], codep[3];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const unsigned a = R[rdz[q]][rdy[q]], b = R[rdz[q]][rdy[q] + 1], c = R[rdz[q] + 1][rdy[q]], d = R[rdz[q] + 1][rdy[q] + 1];
            code[q] = case_code(a, b, c, d, 1);
            if (q < 3) codep[q] = (x > 1) ? case_code(a, b, c, d, 0) : 0u;      // the cube at x - 1 (x = 0 is never visited)
        }
        const unsigned code_self = code[0];
        const unsigned info_self = d_case_info[code_self];
        const unsigned nt = info_self & 0xffu, own_self = info_self >> 8, nv = own_self & 3u;
        const unsigned emask = ISD_EDGE_MASK(code_self);
        const unsigned long long sb_self = mb.seg_bits[seg];
        const uint2 off_self = mb.seg_off[seg];
        const long long vself = (long long)off_self.x + planes_v(sb_self, (1u << l) - 1u);
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
        if (nt == 0 || ghost) continue;
        // ---- vertex index of every cut edge (SURVEY Appendix A + first-appearance cube) --------------------------------
        long long ev[12];
#pragma unroll
        for (int q = 0; q < 12; ++q) ev[q] = 0;
        auto vidx = [&](int cx, int cy, int cz, unsigned cd, unsigned edge_code) -> long long {
            const unsigned own = d_case_info[cd] >> 8;
            return ea.vbase + cube_vertex_base(m, mb, cx, cy, cz) + owned_rank(own, edge_code);
        };
        ev[5] = ea.vbase + vself + owned_rank(own_self, 0u);
        ev[6] = ea.vbase + vself + owned_rank(own_self, 1u);
        ev[10] = ea.vbase + vself + owned_rank(own_self, 2u);
        if (emask & 0x001u) ev[0] = vidx(x, y - 1, z - 1, code[3], 1u);             // x-edge at (x,y,z)    <- cube (x, y-1, z-1) edge 6
        if (emask & 0x006u) {                                                       // cube (x, y, z-1): edges 5 and 6
            const long long b = ea.vbase + cube_vertex_base(m, mb, x, y, z - 1);
            const unsigned own = d_case_info[code[1]] >> 8;
            ev[1] = b + owned_rank(own, 0u);                                        // y-edge at (x+1,y,z)
            ev[2] = b + owned_rank(own, 1u);                                        // x-edge at (x,y+1,z)
        }
        if (emask & 0x210u) {                                                       // cube (x, y-1, z): edges 6 and 10
            const long long b = ea.vbase + cube_vertex_base(m, mb, x, y - 1, z);
            const unsigned own = d_case_info[code[2]] >> 8;
            ev[4] = b + owned_rank(own, 1u);                                        // x-edge at (x,y,z+1)
            ev[9] = b + owned_rank(own, 2u);                                        // z-edge at (x+1,y,z)
        }
        if (emask & 0x008u) ev[3] = vidx(x - 1, y, z - 1, codep[1], 0u);            // y-edge at (x,y,z)    <- cube (x-1, y, z-1) edge 5
        if (emask & 0x100u) ev[8] = vidx(x - 1, y - 1, z, codep[2], 2u);            // z-edge at (x,y,z)    <- cube (x-1, y-1, z) edge 10
        if (emask & 0x880u) {                                                       // cube (x-1, y, z): edges 5 and 10
            const long long b = ea.vbase + cube_vertex_base(m, mb, x - 1, y, z);
            const unsigned own = d_case_info[codep[0]] >> 8;
            ev[7] = b + owned_rank(own, 0u);                                        // y-edge at (x,y,z+1)
            ev[11] = b + owned_rank(own, 2u);                                       // z-edge at (x,y+1,z)
        }
        // ---- triangles ---------------------------------------------------------------------------------------------------
        const long long floc = (long long)off_self.y + planes_t(sb_self, (1u << l) - 1u) - ea.f_local0;
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
