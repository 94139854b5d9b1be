// K1 — field evaluation over the sampling grid, tile by tile, with the sign bits marching cubes needs.
//
// Replaces MarchingCubes::prepare_grid + eval_shape + seal_exterior (marching_cubes.hpp:1662-1738, :895-963): the
// reference materialises a 12 B/node coordinate array and runs the CSG tree over it with two N-sized temporaries per
// tree node; here a node's coordinate is computed from its index in registers, the whole flattened program
// (csg_eval.cuh) runs per thread, and the two outermost node layers get the seal value without being evaluated.
//
// Two kernels:
//   k_tile_bounds : one thread per 8x8x8-node tile.  Interval evaluation of the tree over the tile's box grown by one
//                   cell (composite transforms, outward margins) proves most tiles entirely inside or entirely outside;
//                   no cut edge can touch such a tile, so its field values are never read and it only gets sign bits.
//                   The rest go on a work list.  With pruning off every tile goes on the list (dense field, parity mode).
//   k_eval_tiles  : persistent blocks walk the work list; 128 threads = 64 rows x 2 quads evaluate 4 x-adjacent nodes
//                   each (bit-exact arithmetic), store them as one aligned float4, and pack the signs (f < 0,
//                   marching_cubes.hpp:553-560) into one byte per 8 nodes.
// Layout in HBM: field[zl * plane + y * pitch + x] (pitch = S rounded up to 8), sign bits as 32-bit words
// signs[(zl * S + y) * SW + x / 32], bit x % 32; padding bits are 1 ("outside", like the seal).
// Work per evaluated node: 4 B + 1 bit written, 0 read; flops = Program::flops_per_point.
#pragma once
#include <cuda_runtime.h>
#include "csg_eval.cuh"
#include "csg_bounds.cuh"

namespace isd {

// The program lives in global memory; every thread of a block walks it in step, so its 128-bit fetches are broadcasts
// served by L1.  (It used to sit in __constant__ memory: the interpreter's program counter is not provably uniform,
// so every word became an indexed LDC through the address-divergence unit, which ncu showed as the busiest pipe of
// k_eval_tiles — 12 of them per matrix against 3 LDG.128 now.)
__device__ uint4 d_program[kMaxProgramWords / 4];

struct ConstProgram {
    __device__ __forceinline__ W4 load4(int i) const {
        const uint4 q = __ldg(&d_program[i >> 2]);
        return W4{q.x, q.y, q.z, q.w};
    }
};

// Flop counting ("gpu":{"count_flops":1}, an untimed instrumented pass of bench.py): every evaluation adds the flops of the
// instruction list it ran (list_flops / program_flops, csg_eval.cuh) to d_flops_now; the host moves the sum to the launching
// kernel's slot after each launch.  The timed kernels are the same binaries with the flag off (one uniform branch per evaluation).
__constant__ int c_count_flops;
__device__ unsigned long long d_flops_now;
__device__ unsigned long long d_flops_by_kernel[32];
__device__ __forceinline__ void add_flops(unsigned long long n) {
    const unsigned m = __activemask();
    const unsigned sum = __reduce_add_sync(m, (unsigned)n);      // one evaluation is far below 2^27 flops
    if ((threadIdx.x & 31) == (unsigned)(__ffs(m) - 1)) atomicAdd(&d_flops_now, (unsigned long long)sum);
}
__global__ void k_flops_move(int id) { d_flops_by_kernel[id] += d_flops_now; d_flops_now = 0ull; }

struct GridDesc {
    int S;             // nodes per axis = R + 5
    int pitch;         // floats per x-row in HBM
    int SW;            // sign words per row = ceil(S / 32) + 1 (the last one is padding, all ones)
    int z0, nz;        // node planes [z0, z0 + nz) are resident on this GPU
    int tx, ty, tz;    // 8x8x8 tiles covering S x S x nz
    float xmin, ymin, zmin;
    float wx, wy, wz;  // (max - min) / R   (marching_cubes.hpp:231-243)
    float two_wx, two_wy, two_wz;   // skip_count_l * width (:1691)
    float xi0, yi0, zi0;            // min / width - 2  (:1026-1028)
    long long plane;   // pitch * S
};

constexpr float kExteriorValue = -10000000.0f;     // marching_cubes.hpp:1734

// node coordinate exactly as prepare_grid forms it (:1679-1693): ((REAL)i * factor + min) - 2 * width
__device__ __forceinline__ float node_coord(int i, float w, float mn, float two_w) { return ((float)i * w + mn) - two_w; }
__device__ __forceinline__ bool sealed_index(int i, int S) { return i <= 1 || i >= S - 2; }

// Box of node index ranges [x0,x1] x [y0,y1] x [z0,z1] grown by one cell on every side (+ a hair): class and, with SPEC,
// the decisions behind it.
template <bool SPEC, bool MASKED, class InMask>
__device__ inline uint8_t classify_box(const GridDesc& g, const uint32_t* __restrict__ bounds_prog, int x0, int y0, int z0, int x1, int y1, int z1,
                                       uint32_t* decisions, const InMask& in_mask) {
    const bool touches_seal = x0 <= 1 || y0 <= 1 || z0 <= 1 || x1 + 1 >= g.S - 2 || y1 + 1 >= g.S - 2 || z1 + 1 >= g.S - 2;   // the grown box must stay clear of the seal too
    const float ax = node_coord(x0 - 1, g.wx, g.xmin, g.two_wx), bx = node_coord(x1 + 1, g.wx, g.xmin, g.two_wx);
    const float ay = node_coord(y0 - 1, g.wy, g.ymin, g.two_wy), by = node_coord(y1 + 1, g.wy, g.ymin, g.two_wy);
    const float az = node_coord(z0 - 1, g.wz, g.zmin, g.two_wz), bz = node_coord(z1 + 1, g.wz, g.zmin, g.two_wz);
    const Ival b = program_bounds<SPEC, MASKED>(bounds_prog, 0.5f * (ax + bx), 0.5f * (ay + by), 0.5f * (az + bz), 0.5f * (bx - ax) * 1.001f,
                                                0.5f * (by - ay) * 1.001f, 0.5f * (bz - az) * 1.001f, decisions, in_mask);
    if (b.hi < 0.0f) return TILE_OUTSIDE;
    if (b.lo > 0.0f && !touches_seal) return TILE_INSIDE;
    return TILE_MIXED;
}

// Level 1: one thread per supertile (4x4x4 tiles = 32^3 nodes).  Outside supertiles need nothing (the sign array is pre-set
// to ones); inside and mixed ones go on the supertile list (inside flagged in bit 31), mixed ones with their liveness mask.
// counters[2] receives the list length.
constexpr unsigned kSuperInside = 0x80000000u;
template <bool SPEC>
__global__ void __launch_bounds__(128) k_super_bounds(GridDesc g, const uint32_t* __restrict__ bounds_prog, int prune, SpecDesc spec,
                                                      unsigned* __restrict__ super_list, unsigned* __restrict__ counters) {
    const int stx = (g.tx + 3) / 4, sty = (g.ty + 3) / 4, stz = (g.tz + 3) / 4;
    const int t = blockIdx.x * 128 + threadIdx.x;
    const bool valid = t < stx * sty * stz;
    uint8_t cls = TILE_MIXED;
    uint32_t decisions[SPEC ? kDecisionWords : 1];
    if (valid && prune) {
        const int sx = t % stx, r = t / stx, sy = r % sty, sz = r / sty;
        const int x0 = sx * 32, y0 = sy * 32, z0 = g.z0 + sz * 32;
        const int x1 = min(x0 + 31, g.S - 1), y1 = min(y0 + 31, g.S - 1), z1 = min(z0 + 31, g.z0 + g.nz - 1);
        cls = classify_box<SPEC, false>(g, bounds_prog, x0, y0, z0, x1, y1, z1, decisions, NoMask());
    }
    const bool listed = valid && cls != TILE_OUTSIDE;
    const unsigned m = __ballot_sync(0xffffffffu, listed);
    if (m) {
        const int lane = threadIdx.x & 31;
        unsigned base = 0;
        if (lane == 0) base = atomicAdd(counters + 2, (unsigned)__popc(m));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (listed) {
            const unsigned at = base + __popc(m & ((1u << lane) - 1u));
            super_list[at] = (unsigned)t | (cls == TILE_INSIDE ? kSuperInside : 0u);
            if (SPEC && cls == TILE_MIXED) {
                uint32_t* out = spec.super_masks + (size_t)at * spec.mask_words;
                if (prune) write_live_mask(spec, decisions, out);
                else for (int w = 0; w < spec.mask_words; ++w) out[w] = 0xffffffffu;
            }
        }
    }
}

// z-slab load balance: how much surface each layer of tiles holds, over the WHOLE grid (g.z0 = 0, g.nz = S).  One thread per
// supertile; a mixed supertile is looked at again in its four 32x32x8 slices and every mixed slice adds one to its tile
// layer's bin.  Integer counts from interval bounds: every rank computes the same histogram, so all ranks derive the same
// slab boundaries without talking to each other.
__global__ void __launch_bounds__(128) k_balance_hist(GridDesc g, const uint32_t* __restrict__ bounds_prog, unsigned* __restrict__ hist) {
    const int stx = (g.tx + 3) / 4, sty = (g.ty + 3) / 4, stz = (g.tz + 3) / 4;
    const int t = blockIdx.x * 128 + threadIdx.x;
    if (t >= stx * sty * stz) return;
    const int sx = t % stx, r = t / stx, sy = r % sty, sz = r / sty;
    const int x0 = sx * 32, y0 = sy * 32, z0 = sz * 32;
    const int x1 = min(x0 + 31, g.S - 1), y1 = min(y0 + 31, g.S - 1), z1 = min(z0 + 31, g.S - 1);
    uint32_t none[1];
    if (classify_box<false, false>(g, bounds_prog, x0, y0, z0, x1, y1, z1, none, NoMask()) != TILE_MIXED) return;
    for (int j = 0; j < 4; ++j) {
        const int a = z0 + 8 * j, b = min(a + 7, g.S - 1);
        if (a > g.S - 1) break;
        if (classify_box<false, false>(g, bounds_prog, x0, y0, a, x1, y1, b, none, NoMask()) == TILE_MIXED) atomicAdd(hist + (a >> 3), 1u);
    }
}

// Where the work-list tiles are, for marching cubes (mc_kernels.cuh walks the work list, row of tiles by row of tiles):
// tile_slot[tile] = work-list position; tile_any[tz * ty_count + ty] = 1 if some tile of that row of tiles is listed;
// tile_bits = one bit per tile, TW words per row of tiles.
struct TileHints {
    unsigned* tile_slot;
    uint8_t* tile_any;
    uint32_t* tile_bits;
    int TW;
};

// Level 2: one 64-thread block per listed supertile (grid-stride), one thread per tile.  Tiles of an inside supertile clear
// their sign bytes; tiles of a mixed one are bounded again over their own box — only the primitives still live in the
// supertile are looked at — and the mixed ones go on the work list with their own liveness mask.
// counters[0] receives the work-list length.
template <bool SPEC>
__global__ void __launch_bounds__(64) k_tile_bounds(GridDesc g, const uint32_t* __restrict__ bounds_prog, int prune, SpecDesc spec,
                                                    const unsigned* __restrict__ super_list, uint32_t* __restrict__ signs,
                                                    unsigned* __restrict__ list, unsigned* __restrict__ counters, TileHints hints) {
    const unsigned n_super = counters[2];
    const int stx = (g.tx + 3) / 4, sty = (g.ty + 3) / 4;
    uint32_t decisions[SPEC ? kDecisionWords : 1];
    for (unsigned item = blockIdx.x; item < n_super; item += gridDim.x) {
        const unsigned entry = super_list[item];
        const int st = (int)(entry & ~kSuperInside);
        const int sx = st % stx, r = st / stx, sy = r % sty, sz = r / sty;
        const int tx = sx * 4 + (threadIdx.x & 3), ty = sy * 4 + ((threadIdx.x >> 2) & 3), tzl = sz * 4 + (threadIdx.x >> 4);
        const bool valid = tx < g.tx && ty < g.ty && tzl < g.tz;
        uint8_t cls = (entry & kSuperInside) ? TILE_INSIDE : TILE_MIXED;
        if (valid && cls == TILE_MIXED && prune) {
            const int x0 = tx * 8, y0 = ty * 8, z0 = g.z0 + tzl * 8;
            const int x1 = min(x0 + 7, g.S - 1), y1 = min(y0 + 7, g.S - 1), z1 = min(z0 + 7, g.z0 + g.nz - 1);
            if (SPEC) cls = classify_box<SPEC, true>(g, bounds_prog, x0, y0, z0, x1, y1, z1, decisions, WordMask{spec.super_masks + (size_t)item * spec.mask_words});
            else cls = classify_box<SPEC, false>(g, bounds_prog, x0, y0, z0, x1, y1, z1, decisions, NoMask());
        }
        // inside tiles: clear their sign bytes (the array is pre-set to ones = outside)
        if (valid && cls == TILE_INSIDE) {
            uint8_t* sb = reinterpret_cast<uint8_t*>(signs);
            for (int dz = 0; dz < 8; ++dz) {
                const int zl = tzl * 8 + dz;
                if (zl >= g.nz) break;
                for (int dy = 0; dy < 8; ++dy) {
                    const int y = ty * 8 + dy;
                    if (y >= g.S) break;
                    sb[((long long)zl * g.S + y) * (g.SW * 4) + tx] = 0;
                }
            }
        }
        // mixed tiles -> work list (warp-aggregated append)
        const bool mixed = valid && cls == TILE_MIXED;
        const unsigned m = __ballot_sync(0xffffffffu, mixed);
        if (m) {
            const int lane = threadIdx.x & 31;
            unsigned base = 0;
            if (lane == 0) base = atomicAdd(counters, (unsigned)__popc(m));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (mixed) {
                const unsigned at = base + __popc(m & ((1u << lane) - 1u));
                const int t = (tzl * g.ty + ty) * g.tx + tx;
                list[at] = (unsigned)t;
                hints.tile_slot[t] = at;
                hints.tile_any[tzl * g.ty + ty] = 1;
                atomicOr(hints.tile_bits + (long long)(tzl * g.ty + ty) * hints.TW + (tx >> 5), 1u << (tx & 31));
                if (SPEC) {
                    uint32_t* out = spec.masks + (size_t)at * spec.mask_words;
                    if (prune) write_live_mask(spec, decisions, out);
                    else for (int w = 0; w < spec.mask_words; ++w) out[w] = 0xffffffffu;
                }
            }
        }
    }
}

#ifndef ISD_TILE_NODES
#define ISD_TILE_NODES 8      // measured on C4 1024^3: 8 nodes/thread, 64-thread blocks, >= 10 blocks/SM: 0.63 ms; 4 nodes/thread: 0.77 ms
#endif
constexpr int kNodesPerThread = ISD_TILE_NODES;      // 4 or 8 x-adjacent nodes per thread (2 or 4 packed pairs)
constexpr int kTileThreads = 512 / kNodesPerThread;
#ifndef ISD_EVAL_MINBLOCKS
#define ISD_EVAL_MINBLOCKS 10
#endif

struct TileLive {             // liveness mask of the tile being evaluated (uniform across the block)
    const uint32_t* mask;
    unsigned long long first64;       // ids 0..63 in a register pair: no load on the interpreter's critical path
    __device__ __forceinline__ TileLive(const uint32_t* m, int words) : mask(m) {
        const uint32_t lo = __ldg(m), hi = words > 1 ? __ldg(m + 1) : 0u;
        first64 = ((unsigned long long)hi << 32) | lo;
    }
    __device__ __forceinline__ bool operator()(uint32_t id) const {
        if (id < 64u) return (first64 >> id) & 1ull;
        return (__ldg(mask + (id >> 5)) >> (id & 31)) & 1u;
    }
};

// Per-tile instruction lists for evaluations at arbitrary points (the OB02 kernels: centroids, search points, bisection
// midpoints).  A work-list tile's liveness mask was proven over the tile's box grown by one cell, so it holds at every point
// whose containing node cell lies in the tile; k_build_point_lists turns every such mask into the tile's value list
// (spec_entry) and gradient list (spec_entry_gradient) once per build, and an evaluation near the surface runs its tile's
// list: no liveness tests, no dead-node jumps, no operand bookkeeping.  Elsewhere (tiles proven inside / outside, points off
// the resident planes) there is no list and the whole tree runs.  Same bits either way (tests/emu checks the interpreters
// against each other on the host).
struct PointSpec {
    GridDesc g;
    const uint32_t* tile_bits;
    const unsigned* tile_slot;
    const uint32_t* lists;        // per work-list slot two records of `stride` words, value list then gradient list:
                                  // word 0 = number of entries, entries from word 1 (count and first entries share a sector)
    int TW, stride, enabled;
};
__constant__ PointSpec c_point_spec;      // set by the host after K1 of every build; disabled whenever another program is uploaded
// work-list slot of the tile holding the point's node cell, or -1
__device__ __forceinline__ int point_slot(const PointSpec& ps, float x, float y, float z) {
    if (!ps.enabled) return -1;
    // node_coord(i) = i * w + min - 2w  =>  containing cell i = floor((p - min) / w + 2)
    const float fx = (x - ps.g.xmin) / ps.g.wx + 2.0f, fy = (y - ps.g.ymin) / ps.g.wy + 2.0f, fz = (z - ps.g.zmin) / ps.g.wz + 2.0f;
    if (!(fx >= 0.0f && fy >= 0.0f && fz >= (float)ps.g.z0)) return -1;      // also rejects NaN
    if (!(fx < (float)ps.g.S && fy < (float)ps.g.S && fz < (float)(ps.g.z0 + ps.g.nz))) return -1;
    const int tx = (int)fx >> 3, ty = (int)fy >> 3, tz = ((int)fz - ps.g.z0) >> 3;
    const int row = tz * ps.g.ty + ty;
    if (!((__ldg(ps.tile_bits + (size_t)row * ps.TW + (tx >> 5)) >> (tx & 31)) & 1u)) return -1;
    return (int)__ldg(ps.tile_slot + (size_t)row * ps.g.tx + tx);
}
struct GlobalList {
    const uint32_t* p;
    __device__ __forceinline__ uint32_t operator[](int k) const { return __ldg(p + k); }
};

// One warp per work-list tile: the tile's two instruction lists from its liveness mask, program order kept by ballot ranks.
__global__ void __launch_bounds__(128) k_build_point_lists(const unsigned* __restrict__ list_count, const uint32_t* __restrict__ masks, int mask_words,
                                                           const uint4* __restrict__ itab, int n_instr, int stride, uint32_t* __restrict__ lists) {
    const unsigned n = list_count[0];
    const int lane = threadIdx.x & 31;
    for (unsigned slot = blockIdx.x * 4 + (threadIdx.x >> 5); slot < n; slot += gridDim.x * 4) {
        const TileLive live(masks + (size_t)slot * mask_words, mask_words);
        uint32_t* lv = lists + (size_t)slot * 2 * stride + 1;
        uint32_t* lg = lv + stride;
        int nv = 0, ng = 0;
        for (int i0 = 0; i0 < n_instr; i0 += 32) {
            const int i = i0 + lane;
            uint32_t eg = 0u;
            if (i < n_instr) {
                const uint4 it = __ldg(itab + i);
                eg = spec_entry_gradient(W4{it.x, it.y, it.z, it.w}, live);
            }
            const uint32_t ev = (eg & LE_PASS) ? 0u : eg;      // spec_entry's answer: in value mode a pass-through does nothing
            const unsigned mg = __ballot_sync(0xffffffffu, eg != 0u), mv = __ballot_sync(0xffffffffu, ev != 0u);
            const unsigned below = (1u << lane) - 1u;
            if (eg) lg[ng + __popc(mg & below)] = eg;
            if (ev) lv[nv + __popc(mv & below)] = ev;
            ng += __popc(mg); nv += __popc(mv);
        }
        if (lane == 0) { lv[-1] = (uint32_t)nv; lg[-1] = (uint32_t)ng; }
    }
}

// SPEC: before evaluating a tile the block turns its liveness mask into the list of instructions the tile really runs
// (spec_entry, one instruction per thread, order kept by ballot ranks) in shared memory; eval_values_list then has no
// liveness tests, no dead-node jumps and no operand bookkeeping left.
template <bool SPEC>
__global__ void __launch_bounds__(kTileThreads, ISD_EVAL_MINBLOCKS) k_eval_tiles(GridDesc g, SdfGridView sdf, const unsigned* __restrict__ list,
                                                            const unsigned* __restrict__ list_count, const uint32_t* __restrict__ masks,
                                                            int mask_words, const uint4* __restrict__ itab, int n_instr,
                                                            float* __restrict__ field, uint32_t* __restrict__ signs) {
    __shared__ uint32_t s_list[SPEC ? kMaxListEntries : 1];
    __shared__ int s_warp_count[kTileThreads / 32];
    __shared__ int s_n_list;
    const unsigned n = list_count[0];
    unsigned* next_item = const_cast<unsigned*>(list_count) + 1;      // dynamic queue: specialised tiles differ a lot in cost
    constexpr int NPT = kNodesPerThread, QPR = 8 / NPT;      // threads per row of 8 nodes
    const int q = threadIdx.x % QPR, row = threadIdx.x / QPR;
    uint8_t* sb = reinterpret_cast<uint8_t*>(signs);
    __shared__ unsigned s_item;
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) s_item = atomicAdd(next_item, 1u);
        __syncthreads();
        const unsigned item = s_item;
        if (item >= n) break;
        if (SPEC) {
            const TileLive live(masks + (size_t)item * mask_words, mask_words);
            const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
            int base = 0;
            for (int i0 = 0; i0 < n_instr; i0 += kTileThreads) {       // uniform trip count
                const int i = i0 + (int)threadIdx.x;
                uint32_t e = 0u;
                if (i < n_instr) {
                    const uint4 it = __ldg(itab + i);
                    e = spec_entry(W4{it.x, it.y, it.z, it.w}, live);
                }
                const unsigned m = __ballot_sync(0xffffffffu, e != 0u);
                if (lane == 0) s_warp_count[warp] = __popc(m);
                __syncthreads();
                int before = base;
                for (int w = 0; w < warp; ++w) before += s_warp_count[w];
                if (e) s_list[before + __popc(m & ((1u << lane) - 1u))] = e;
                for (int w = 0; w < kTileThreads / 32; ++w) base += s_warp_count[w];
                __syncthreads();
            }
            if (threadIdx.x == 0) s_n_list = base;
            __syncthreads();
        }
        const int t = (int)list[item];
        const int tx = t % g.tx;
        const int r = t / g.tx;
        const int ty = r % g.ty, tzl = r / g.ty;
        const int x0 = tx * 8 + q * NPT, y = ty * 8 + (row & 7), zl = tzl * 8 + (row >> 3);
        const int z = g.z0 + zl;
        unsigned nib = (1u << NPT) - 1u;
        const bool in_grid = (y < g.S) && (zl < g.nz) && (x0 < g.pitch);
        if (in_grid) {
            float4* dst = reinterpret_cast<float4*>(field + (long long)zl * g.plane + (long long)y * g.pitch + x0);
            const bool row_sealed = sealed_index(y, g.S) || sealed_index(z, g.S);
            if (row_sealed || x0 + NPT - 1 <= 1 || x0 >= g.S - 2) {
#pragma unroll
                for (int h = 0; h < NPT / 4; ++h) dst[h] = make_float4(kExteriorValue, kExteriorValue, kExteriorValue, kExteriorValue);
            } else {
                float f[NPT];
                const float yy = node_coord(y, g.wy, g.ymin, g.two_wy);
                const float zz = node_coord(z, g.wz, g.zmin, g.two_wz);
                if (SPEC) {      // packed pairs (f32x2): half the floating-point instructions, the same bits
                    F2 px[NPT / 2], py[NPT / 2], pz[NPT / 2], pf[NPT / 2];
#pragma unroll
                    for (int p = 0; p < NPT / 2; ++p) {
                        px[p] = F2{node_coord(x0 + 2 * p, g.wx, g.xmin, g.two_wx), node_coord(x0 + 2 * p + 1, g.wx, g.xmin, g.two_wx)};
                        py[p] = splat(yy); pz[p] = splat(zz);
                    }
                    eval_values_list2<NPT / 2>(ConstProgram(), sdf, s_list, s_n_list, px, py, pz, pf);
#pragma unroll
                    for (int p = 0; p < NPT / 2; ++p) { f[2 * p] = pf[p].x; f[2 * p + 1] = pf[p].y; }
                    if (c_count_flops) add_flops((unsigned long long)NPT * list_flops(ConstProgram(), s_list, s_n_list, false));
                } else {
#pragma unroll
                    for (int h = 0; h < NPT / 4; ++h) {
                        float cx[4], cy[4], cz[4], f4[4];
#pragma unroll
                        for (int p = 0; p < 4; ++p) { cx[p] = node_coord(x0 + 4 * h + p, g.wx, g.xmin, g.two_wx); cy[p] = yy; cz[p] = zz; }
                        eval_values<4>(ConstProgram(), sdf, AllLive(), cx, cy, cz, f4);
#pragma unroll
                        for (int p = 0; p < 4; ++p) f[4 * h + p] = f4[p];
                    }
                    if (c_count_flops) add_flops((unsigned long long)NPT * program_flops(ConstProgram(), false));
                }
                nib = 0;
#pragma unroll
                for (int p = 0; p < NPT; ++p) {
                    if (sealed_index(x0 + p, g.S)) f[p] = kExteriorValue;
                    nib |= (f[p] < 0.0f ? 1u : 0u) << p;
                }
#pragma unroll
                for (int h = 0; h < NPT / 4; ++h) dst[h] = make_float4(f[4 * h], f[4 * h + 1], f[4 * h + 2], f[4 * h + 3]);
            }
        }
        if (NPT == 4) {      // two quads -> one sign byte (lanes 2j, 2j+1 share a row)
            const unsigned other = __shfl_xor_sync(0xffffffffu, nib, 1);
            if (q == 0 && y < g.S && zl < g.nz) sb[((long long)zl * g.S + y) * (g.SW * 4) + tx] = (uint8_t)(nib | (other << 4));
        } else {
            if (y < g.S && zl < g.nz) sb[((long long)zl * g.S + y) * (g.SW * 4) + tx] = (uint8_t)nib;
        }
    }
}

// K10 — f and grad f on an arbitrary point list (mcc2.cpp:815-889 calculate_implicit_values/gradients).
// mode bit 0: write f; bit 1: write gradient; bit 2: normalize_and_invert (g <- -g/|g|, or -42 g when |g| <= 1e-4,
// mcc2.cpp:852-870).
__global__ void __launch_bounds__(128) k_eval_points(SdfGridView sdf, const float* __restrict__ xyz, long long n,
                                                     float* __restrict__ f_out, float* __restrict__ g_out, int mode) {
    const long long i = (long long)blockIdx.x * 128 + threadIdx.x;
    if (i >= n) return;
    const float x = xyz[i * 3], y = xyz[i * 3 + 1], z = xyz[i * 3 + 2];
    float f, g0, g1, g2;
    eval_value_gradient(ConstProgram(), sdf, x, y, z, f, g0, g1, g2);
    if (c_count_flops) add_flops(program_flops(ConstProgram(), true));
    if (mode & 1) f_out[i] = f;
    if (mode & 2) {
        if (mode & 4) {
            const float norm = sqrtf(g0 * g0 + g1 * g1 + g2 * g2);
            // norm_factor is formed in double in the reference (-1.0/norm) and rounded once: same value as the float division
            const float k = (norm > 0.0001f) ? (-1.0f / norm) : -42.0f;
            g0 = g0 * k; g1 = g1 * k; g2 = g2 * k;
        }
        g_out[i * 3] = g0; g_out[i * 3 + 1] = g1; g_out[i * 3 + 2] = g2;
    }
}

}  // namespace isd
