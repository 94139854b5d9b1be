// K5-K9 — the OB02 mesh-optimisation steps on the device-resident mesh.
//
// Reference (paths relative to /root/reference/js_iteration_2/):
//   adjacency            mesh_algorithms.hpp:50-209 (make_edge_lookup, build_faces_of_faces, make_neighbour_faces_of_vertex)
//   vertex resampling    vertex_resampling.hpp:47-225, apply_v_s_to_mc_buffers.hpp:280-326
//   centroid projection  centroids_projection.cpp:70-82,144-362,421-1311, polygoniser/bisection.hpp:117-459
//   QEM                  qem.hpp:256-599
// The reference runs every step as whole-array passes over shrinking index sets (std::map / vector<vector>
// adjacency, one eval_implicit call per alpha over all still-active centroids).  Every centroid's search and every
// vertex's update is independent of the others, so here one thread owns one centroid (or vertex) and walks the whole
// sequence — direction types, alphas, bisection — in registers, calling the flattened CSG program inline.  Field
// values are bit-identical to the reference's, so the first alpha that flips the sign is the same one.
#pragma once
#include <cuda_runtime.h>
#include "field_kernel.cuh"

namespace isd {

constexpr float kRootTolerance = (float)(0.001 / 10.0);     // configs.hpp:33
constexpr float kMinGradientLen = 0.000001f;                 // configs.hpp:57
constexpr int kBisectionMaxIter = 100;                       // ours: the reference has no cap (bisection.hpp:204,415-437)
constexpr int kMaxAlphas = 512;

// make_alpha_list (centroids_projection.cpp:144-194), built on the device from the mean edge length: the host never has to
// wait for that scalar.  Direction types 1-6 use the first ten alphas (:211); below a mean edge of 0.004 the list is shorter
// than that and the reference slices out of range (undefined behaviour) — here they use what exists.
struct AlphaList {
    float alpha[kMaxAlphas];
    int n_full, n_short;
    float avg_edge;
    unsigned ticket;          // last-block-done counter of k_face_prepare
    unsigned long long edge_sum;      // fixed-point (2^-32) sum of the edge lengths k_face_prepare was asked to sum
    unsigned long long edge_total;    // z-slab runs: the all-reduced sum
};
__device__ inline void make_alpha_list(float initial_step_size, float min_step_size, float max_dist, int max_iter, AlphaList* out) {
    const float unit_of_length = max_dist;
    int n = 0;
    float step_size = initial_step_size;
    while (step_size > min_step_size) {
        step_size = (float)(step_size * 0.5);
        const int total_steps = (int)(floor((double)(max_dist / fabsf(step_size)) + 0.001));
        const int max_steps = min(max_iter, total_steps);
        for (int i = 1; i < max_steps + 1; i += 2) {
            const float alpha = ((float)i) * step_size;
            if (n + 1 < kMaxAlphas) { out->alpha[n] = alpha / unit_of_length; out->alpha[n + 1] = -alpha / unit_of_length; }
            n += 2;
        }
    }
    if (n > kMaxAlphas) n = kMaxAlphas;
    out->n_full = n;
    out->n_short = n < 10 ? n : 10;
}

// counters: [0..6] centroids resolved by direction type, [7] failed, [8] bisection capped, [9] bisected,
// [10] centroids still unresolved after phase A, [11] range ticket of k_finish_bisect_refill, [12] of k_search_trials_refill
struct Ob02Counters { unsigned long long c[16]; };

// one atomic per warp instead of one per thread: all counters are bumped by up to every thread of the grid
__device__ __forceinline__ void count_if(bool pred, unsigned long long* ctr) {
    const unsigned m = __ballot_sync(__activemask(), pred);
    if (pred && (threadIdx.x & 31) == (__ffs(m) - 1)) atomicAdd(ctr, (unsigned long long)__popc(m));
}

__device__ __forceinline__ float norm2f(float x, float y, float z) { return sqrtf(x * x + y * y + z * z); }
__device__ __forceinline__ float sign_tol(float v) { return (v > kRootTolerance) ? 1.0f : (v < -kRootTolerance) ? -1.0f : 0.0f; }

// f at a point / f and grad f at a point.  Near the surface the point lies in a work-list tile of K1, whose instruction
// lists hold only what can matter there (same bits, a fraction of the work for deep trees); elsewhere the whole tree runs.
// MASKED is a compile-time switch: shallow trees (and meshes that did not come from K1) run the plain instantiation, which
// carries neither the tile look-up nor a second copy of the interpreter.
template <bool MASKED>
__device__ __forceinline__ float eval_f(const SdfGridView& sdf, float x, float y, float z) {
    float cx[1] = {x}, cy[1] = {y}, cz[1] = {z}, out[1];
    const int slot = MASKED ? point_slot(c_point_spec, x, y, z) : -1;
    if (MASKED && slot >= 0) {
        const uint32_t* rec = c_point_spec.lists + (size_t)slot * 2 * c_point_spec.stride;
        eval_values_list<1>(ConstProgram(), sdf, GlobalList{rec + 1}, (int)__ldg(rec), cx, cy, cz, out);
        if (c_count_flops) add_flops(list_flops(ConstProgram(), GlobalList{rec + 1}, (int)__ldg(rec), false));
    } else {
        eval_values<1>(ConstProgram(), sdf, AllLive(), cx, cy, cz, out);
        if (c_count_flops) add_flops(program_flops(ConstProgram(), false));
    }
    return out[0];
}
template <bool MASKED>
__device__ __forceinline__ void eval_fg(const SdfGridView& sdf, float x, float y, float z, float& f, float& g0, float& g1, float& g2) {
    const int slot = MASKED ? point_slot(c_point_spec, x, y, z) : -1;
    if (MASKED && slot >= 0) {
        const uint32_t* rec = c_point_spec.lists + ((size_t)slot * 2 + 1) * c_point_spec.stride;
        eval_value_gradient_list(ConstProgram(), sdf, GlobalList{rec + 1}, (int)__ldg(rec), x, y, z, f, g0, g1, g2);
        if (c_count_flops) add_flops(list_flops(ConstProgram(), GlobalList{rec + 1}, (int)__ldg(rec), true));
    } else {
        eval_value_gradient(ConstProgram(), sdf, x, y, z, f, g0, g1, g2);
        if (c_count_flops) add_flops(program_flops(ConstProgram(), true));
    }
}

// ---- generic exclusive scan of a u32 array (three launches) ----------------------------------------------------------
constexpr int kUScanThreads = 256, kUScanItems = 4, kUScanTile = kUScanThreads * kUScanItems;

__device__ __forceinline__ unsigned block_exclusive_u32(unsigned mine, unsigned* total) {
    __shared__ unsigned s_w[kUScanThreads / 32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    unsigned inc = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned v = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += v;
    }
    if (lane == 31) s_w[w] = inc;
    __syncthreads();
    unsigned base = 0, sum = 0;
#pragma unroll
    for (int i = 0; i < kUScanThreads / 32; ++i) {
        if (i < w) base += s_w[i];
        sum += s_w[i];
    }
    __syncthreads();
    *total = sum;
    return base + inc - mine;
}
// out[i] = exclusive prefix of in[0..n), out[n] = total; one launch of at most kUScanMaxBlocks blocks over contiguous ranges
// (same scheme as k_row_scan, mc_kernels.cuh): ticket, range sum, publish, wait for the earlier tickets, scan the range.
// `state` = {ticket, pad, pad, pad} + {ready, sum} per block, zeroed before the launch.
constexpr int kUScanMaxBlocks = 296;
__global__ void __launch_bounds__(kUScanThreads) k_uscan(const unsigned* __restrict__ in, long long n, long long per_block,
                                                       unsigned* __restrict__ state, unsigned* __restrict__ out) {
    __shared__ unsigned s_ticket;
    if (threadIdx.x == 0) s_ticket = atomicAdd(state, 1u);
    __syncthreads();
    const unsigned blk = s_ticket;
    const long long first = (long long)blk * per_block, last = first + per_block;      // per_block % kUScanThreads == 0
    unsigned mine = 0, sum;
    for (long long i = first + threadIdx.x; i < last; i += kUScanThreads) mine += (i < n) ? in[i] : 0u;
    block_exclusive_u32(mine, &sum);
    volatile unsigned* st = state + 4;
    if (threadIdx.x == 0) {
        st[blk * 2 + 1] = sum;
        __threadfence();
        st[blk * 2] = 1u;
    }
    unsigned before = 0;
    for (unsigned j = threadIdx.x; j < blk; j += kUScanThreads)
        while (st[j * 2] == 0u) {}
    __threadfence();      // one fence after all the flags have been seen, then the sums
    for (unsigned j = threadIdx.x; j < blk; j += kUScanThreads) before += st[j * 2 + 1];
    unsigned carry;
    block_exclusive_u32(before, &carry);
    for (long long i0 = first; i0 < last; i0 += kUScanThreads) {
        const long long i = i0 + threadIdx.x;
        const unsigned v = (i < n) ? in[i] : 0u;
        unsigned step;
        const unsigned ex = block_exclusive_u32(v, &step);
        if (i < n) out[i] = carry + ex;
        if (i == n - 1) out[n] = carry + ex + v;
        carry += step;
    }
}

// ---- K6 adjacency ------------------------------------------------------------------------------------------------
// make_neighbour_faces_of_vertex (mesh_algorithms.hpp:178-209): faces around a vertex in ascending face index
__global__ void k_vertex_degree(const int* __restrict__ faces, int nf, unsigned* __restrict__ deg) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nf * 3) return;
    atomicAdd(&deg[faces[i]], 1u);
}
__global__ void k_umbrella_fill(const int* __restrict__ faces, int nf, const unsigned* __restrict__ offsets, unsigned* __restrict__ cursor,
                                int* __restrict__ items) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nf * 3) return;
    const int v = faces[i];
    const unsigned slot = atomicAdd(&cursor[v], 1u);
    items[offsets[v] + slot] = i / 3;
}
__global__ void k_umbrella_sort(int nv, const unsigned* __restrict__ offsets, int* __restrict__ items) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= nv) return;
    const unsigned b = offsets[v], e = offsets[v + 1];
    if (e - b <= 16u) {                             // umbrellas hold ~6 faces: sort them in registers (rank by counting; ties cannot
        int key[16];                                // be told apart, a face listed twice is twice the same value)
#pragma unroll
        for (int i = 0; i < 16; ++i) key[i] = (b + i < e) ? items[b + i] : 0x7fffffff;
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            if (b + i < e) {
                int rank = 0;
#pragma unroll
                for (int j = 0; j < 16; ++j) rank += (key[j] < key[i] || (key[j] == key[i] && j < i)) ? 1 : 0;
                items[b + rank] = key[i];
            }
        }
        return;
    }
    for (unsigned i = b + 1; i < e; ++i) {          // insertion sort for the rare big umbrella
        const int key = items[i];
        unsigned j = i;
        while (j > b && items[j - 1] > key) { items[j] = items[j - 1]; --j; }
        items[j] = key;
    }
}
// make_edge_lookup + build_faces_of_faces (mesh_algorithms.hpp:50-131).  The reference's map keeps, per undirected
// edge, the first face that inserted it and the last face that found it; a face's neighbour across the edge is the
// first one unless that is the face itself.  Both are read here from the (ascending) umbrella of one end vertex.
__global__ void k_faces_of_faces(const int* __restrict__ faces, int nf, const unsigned* __restrict__ offsets, const int* __restrict__ items,
                                 int* __restrict__ fof) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nf * 3) return;
    const int f = i / 3, j = i - f * 3;
    const int a = faces[f * 3 + j], b = faces[f * 3 + (j + 1) % 3];
    int first = -1, last = -1, count = 0, prev = -1;
    for (unsigned k = offsets[a], e = offsets[a + 1]; k < e; ++k) {
        const int g = items[k];
        if (g == prev) continue;        // a face listed twice (repeated vertex) is still one face
        prev = g;
        const int g0 = faces[g * 3], g1 = faces[g * 3 + 1], g2 = faces[g * 3 + 2];
        // does g have the undirected edge {a,b}?  (consecutive corners, cyclically — always true for a triangle holding both)
        const bool has = (a != b) ? ((g0 == b) || (g1 == b) || (g2 == b)) : false;
        if (has) {
            if (first < 0) first = g;
            last = g;
            ++count;
        }
    }
    int r;
    if (count <= 1) r = 0;              // boundary edge: faces_of_edges[e][1] keeps its zero initialisation (:66, :100)
    else r = (first != f) ? first : last;
    fof[i] = r;
}

// ---- K5 centroids and their unit gradients ---------------------------------------------------------------------------
// compute_centroids (implicit_vectorised_algorithms.hpp:161-171) + eval_gradient + normalize_1111
// (vectorised_algorithms/normalise_inplace.hpp:60-70)
__global__ void __launch_bounds__(128) k_centroids(const float* __restrict__ verts, const int* __restrict__ faces, int nf, float* __restrict__ cen) {
    const int f = blockIdx.x * 128 + threadIdx.x;
    if (f >= nf) return;
    const int f0 = faces[f * 3], f1 = faces[f * 3 + 1], f2 = faces[f * 3 + 2];
#pragma unroll
    for (int d = 0; d < 3; ++d) cen[f * 3 + d] = (verts[f0 * 3 + d] + verts[f1 * 3 + d] + verts[f2 * 3 + d]) / 3.0f;
}
// both at once (vertex resampling: the gradient is taken at the centroid just stored, same bits)
template <bool MASKED>
__global__ void __launch_bounds__(128) k_centroid_gradients(SdfGridView sdf, const float* __restrict__ verts, const int* __restrict__ faces, int nf,
                                                            float* __restrict__ cen, float* __restrict__ nrm) {
    const int f = blockIdx.x * 128 + threadIdx.x;
    if (f >= nf) return;
    const int f0 = faces[f * 3], f1 = faces[f * 3 + 1], f2 = faces[f * 3 + 2];
    float c[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        c[d] = (verts[f0 * 3 + d] + verts[f1 * 3 + d] + verts[f2 * 3 + d]) / 3.0f;
        cen[f * 3 + d] = c[d];
    }
    float fv, g0, g1, g2;
    eval_fg<MASKED>(sdf, c[0], c[1], c[2], fv, g0, g1, g2);
    const float norm = norm2f(g0, g1, g2);
    nrm[f * 3] = g0 / norm; nrm[f * 3 + 1] = g1 / norm; nrm[f * 3 + 2] = g2 / norm;
}
__global__ void __launch_bounds__(128) k_unit_gradients(SdfGridView sdf, const float* __restrict__ pts, int n, float* __restrict__ nrm) {
    const int i = blockIdx.x * 128 + threadIdx.x;
    if (i >= n) return;
    float f, g0, g1, g2;
    eval_fg<false>(sdf, pts[i * 3], pts[i * 3 + 1], pts[i * 3 + 2], f, g0, g1, g2);
    const float norm = norm2f(g0, g1, g2);
    nrm[i * 3] = g0 / norm; nrm[i * 3 + 1] = g1 / norm; nrm[i * 3 + 2] = g2 / norm;
}

// ---- K7 vertex resampling -------------------------------------------------------------------------------------------
// kij / wi (vertex_resampling.hpp:47-91).  std::acos(float) is glibc's acosf in the reference; here acos is taken in
// double and rounded, which agrees with a correctly rounded acosf except in rare double-rounding cases.
__global__ void __launch_bounds__(128) k_face_weights(const float* __restrict__ cen, const float* __restrict__ nrm, const int* __restrict__ fof,
                                                      int nf, float c, float* __restrict__ w) {
    const int i = blockIdx.x * 128 + threadIdx.x;
    if (i >= nf) return;
    float ki = 0.0f;
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        const int o = fof[i * 3 + j];
        float mimj = nrm[i * 3] * nrm[o * 3] + nrm[i * 3 + 1] * nrm[o * 3 + 1] + nrm[i * 3 + 2] * nrm[o * 3 + 2];
        if (mimj > 1.0f) mimj = 1.0f;
        if (mimj < -1.0f) mimj = -1.0f;
        const float pipj = norm2f(cen[i * 3] - cen[o * 3], cen[i * 3 + 1] - cen[o * 3 + 1], cen[i * 3 + 2] - cen[o * 3 + 2]);
        float kij = 0.0f;
        if (pipj != 0.0f) kij = (float)acos((double)mimj) / pipj;
        ki += kij;
    }
    w[i] = (float)(1.0 + (double)(c * ki));     // REAL w = 1.0 + c*ki: the sum is formed in double
}
// vertex_resampling_VV1 (vertex_resampling.hpp:93-141)
__global__ void __launch_bounds__(128) k_resample(int nv, const unsigned* __restrict__ offsets, const int* __restrict__ items, const float* __restrict__ w,
                                                  const float* __restrict__ cen, float* __restrict__ verts) {
    const int v = blockIdx.x * 128 + threadIdx.x;
    if (v >= nv) return;
    const unsigned b = offsets[v], e = offsets[v + 1];
    float sum_w = 0.0f;
    for (unsigned k = b; k < e; ++k) sum_w += w[items[k]];
    float nx = 0.0f, ny = 0.0f, nz = 0.0f;
    for (unsigned k = b; k < e; ++k) {
        const int f = items[k];
        const float wj = w[f] / sum_w;
        nx += wj * cen[f * 3]; ny += wj * cen[f * 3 + 1]; nz += wj * cen[f * 3 + 2];
    }
    verts[v * 3] = nx; verts[v * 3 + 1] = ny; verts[v * 3 + 2] = nz;
}

// ---- K8 centroid projection -----------------------------------------------------------------------------------------
// compute_average_edge_length (centroids_projection.cpp:70-82): three lengths per face
__global__ void __launch_bounds__(128) k_edge_lengths(const float* __restrict__ verts, const int* __restrict__ faces, int nf, float* __restrict__ len3) {
    const int j = blockIdx.x * 128 + threadIdx.x;
    if (j >= nf) return;
    const int f0 = faces[j * 3], f1 = faces[j * 3 + 1], f2 = faces[j * 3 + 2];
    const float* a = verts + f0 * 3; const float* b = verts + f1 * 3; const float* c = verts + f2 * 3;
    len3[j * 3] = norm2f(a[0] - b[0], a[1] - b[1], a[2] - b[2]);
    len3[j * 3 + 1] = norm2f(a[0] - c[0], a[1] - c[1], a[2] - c[2]);
    len3[j * 3 + 2] = norm2f(c[0] - b[0], c[1] - b[1], c[2] - b[2]);
}
// The reference accumulates the 3*Nf lengths into one float, in order.  "exact" mode repeats that serial sum bit for
// bit (one warp: coalesced loads, lane 0 adds in order); the default is a pairwise tree in double, which is the
// better-conditioned value and what large meshes should use (the serial float sum stagnates beyond ~10^7 terms).
__global__ void __launch_bounds__(32) k_sum_serial_float(const float* __restrict__ v, long long n, int nf, float* __restrict__ out) {
    const int lane = threadIdx.x;
    float acc = 0.0f;
    for (long long base = 0; base < n; base += 32) {
        const float mine = (base + lane < n) ? v[base + lane] : 0.0f;
        const int cnt = (n - base < 32) ? (int)(n - base) : 32;
        for (int l = 0; l < cnt; ++l) {
            const float t = __shfl_sync(0xffffffffu, mine, l);
            acc += t;
        }
    }
    if (lane == 0) *out = (float)(acc / (3. * nf));     // edge_length/(3.*nfaces): double division, rounded to REAL
}
__global__ void __launch_bounds__(256) k_sum_double(const float* __restrict__ v, long long n, double* __restrict__ acc) {
    __shared__ double s[8];
    double t = 0.0;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) t += (double)v[i];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) t += __shfl_xor_sync(0xffffffffu, t, d);
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = t;
    __syncthreads();
    if (threadIdx.x == 0) {
        double b = 0.0;
        for (int i = 0; i < 8; ++i) b += s[i];
        atomicAdd(acc, b);
    }
}
__global__ void k_finish_average(const double* __restrict__ acc, int nf, float* __restrict__ out) { *out = (float)(*acc / (3. * nf)); }

// Projection prologue in one launch: centroids, the three edge lengths per face (compute_average_edge_length,
// centroids_projection.cpp:70-82), the mean edge and the alpha list.  The lengths are summed in 64-bit fixed point (quantum
// 2^-32): integer addition is associative, so the mean does not depend on block scheduling NOR on how a z-slab run splits
// the faces between GPUs — one GPU and eight give the same bits.  Only faces in [sum_f0, sum_f1) are summed (a slab's own
// faces; its halo faces belong to the neighbours' sums).  With `finish` the last block turns the sum into the mean over `nf`
// faces and builds the list; without, the sum is left in al->edge_sum for the all-reduce and k_alpha_from_sum.
// `len3` (optional) keeps the lengths for the reference-order serial float sum (exact_reduce), which then overwrites the mean.
__global__ void __launch_bounds__(256) k_face_prepare(const float* __restrict__ verts, const int* __restrict__ faces, int nf, float* __restrict__ cen,
                                                      float* __restrict__ len3, unsigned long long* __restrict__ partials, AlphaList* __restrict__ al,
                                                      int sum_f0, int sum_f1, int finish) {
    __shared__ unsigned long long s[8];
    __shared__ bool s_last;
    const int j = blockIdx.x * 256 + threadIdx.x;
    unsigned long long t = 0ull;
    if (j < nf) {
        const int f0 = faces[j * 3], f1 = faces[j * 3 + 1], f2 = faces[j * 3 + 2];
        const float* a = verts + f0 * 3; const float* b = verts + f1 * 3; const float* c = verts + f2 * 3;
#pragma unroll
        for (int d = 0; d < 3; ++d) cen[j * 3 + d] = (a[d] + b[d] + c[d]) / 3.0f;
        const float l0 = norm2f(a[0] - b[0], a[1] - b[1], a[2] - b[2]);
        const float l1 = norm2f(a[0] - c[0], a[1] - c[1], a[2] - c[2]);
        const float l2 = norm2f(c[0] - b[0], c[1] - b[1], c[2] - b[2]);
        if (len3 != nullptr) { len3[j * 3] = l0; len3[j * 3 + 1] = l1; len3[j * 3 + 2] = l2; }
        const double sum3 = (double)l0 + (double)l1 + (double)l2;
        if (j >= sum_f0 && j < sum_f1 && sum3 < 2147483648.0) t = (unsigned long long)__double2ll_rn(sum3 * 4294967296.0);
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) t += __shfl_xor_sync(0xffffffffu, t, d);
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = t;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long b = 0ull;
        for (int i = 0; i < 8; ++i) b += s[i];
        partials[blockIdx.x] = b;
        __threadfence();
        s_last = atomicAdd(&al->ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    unsigned long long acc = 0ull;
    for (unsigned i = threadIdx.x; i < gridDim.x; i += 256) acc += ((volatile unsigned long long*)partials)[i];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long b = 0ull;
        for (int i = 0; i < 8; ++i) b += s[i];
        al->edge_sum = b;
        al->ticket = 0u;
        if (finish) {
            const float avg = (float)(((double)b * (1.0 / 4294967296.0)) / (3. * nf));
            al->avg_edge = avg;
            make_alpha_list(avg, 0.001f, avg, 20, al);
        }
    }
}
// z-slab runs: the mean over the whole mesh from the all-reduced sum; the face count is the sum of every slab's count
// (`counts` = the all-gathered {nv, nf} pairs) or `nf_fixed` when emulating ranks on one GPU
__global__ void k_alpha_from_sum(AlphaList* __restrict__ al, const unsigned long long* __restrict__ total, const long long* __restrict__ counts,
                                 int world, long long nf_fixed) {
    long long nf = nf_fixed;
    if (counts != nullptr) {
        nf = 0;
        for (int r = 0; r < world; ++r) nf += counts[2 * r + 1];
    }
    const float avg = (float)(((double)*total * (1.0 / 4294967296.0)) / (3. * (double)nf));
    al->avg_edge = avg;
    make_alpha_list(avg, 0.001f, avg, 20, al);
}
// real z-slab runs: the gathered {nv, nf, edge sum} triples of all ranks (geometry_api.cu enqueue_slab_gather), summed in rank
// order — integer additions, the same total on every rank and the same as one GPU's sum over the whole mesh
__global__ void k_alpha_from_gather(AlphaList* __restrict__ al, const long long* __restrict__ triples, int world) {
    long long nf = 0;
    unsigned long long total = 0ull;
    for (int r = 0; r < world; ++r) { nf += triples[4 * r + 1]; total += (unsigned long long)triples[4 * r + 2]; }      // 4 = kGatherWords
    al->edge_total = total;
    const float avg = (float)(((double)total * (1.0 / 4294967296.0)) / (3. * (double)nf));
    al->avg_edge = avg;
    make_alpha_list(avg, 0.001f, avg, 20, al);
}
// exact_reduce: the alpha list again, from the serial float mean
__global__ void k_alpha_from_mean(const float* __restrict__ mean, AlphaList* __restrict__ al) {
    al->avg_edge = *mean;
    make_alpha_list(*mean, 0.001f, *mean, 20, al);
}

// The order in which the pair kernels take the faces: face order (perm == nullptr: unit u = faces 2u, 2u + 1), or grouped by
// tile (mc_kernels.cuh, k_seg_apply: perm lists every face once, each tile's run padded with -1 to an even length, so the two
// members of a unit always come from one tile).  Results are per face and do not depend on the order.
struct FaceOrder {
    const int* perm;
    const unsigned* len;      // perm's length (device; even)
};
__device__ __forceinline__ long long unit_count(const FaceOrder& fo, int n) { return fo.perm ? (long long)(*fo.len / 2u) : ((long long)n + 1) / 2; }
// the two faces of unit u (u < unit_count): fa >= 0 always, fb = -1 when the unit has one member
__device__ __forceinline__ void unit_faces(const FaceOrder& fo, long long u, int n, int& fa, int& fb) {
    if (fo.perm) { fa = fo.perm[2 * u]; fb = fo.perm[2 * u + 1]; }
    else { fa = (int)(2 * u); fb = 2 * u + 1 < (long long)n ? (int)(2 * u + 1) : -1; }
}

struct ProjectArgs {
    FaceOrder order;
    const float* verts;
    const int* faces;
    const float* cen;        // centroids (input)
    const float* noise;      // make_random_pm1(n, 1e-6) for direction type 2 (phase B only)
    const AlphaList* al;     // mean edge length + alpha list (device)
    float4* best;            // per centroid: candidate on the other side of the surface, and f there
    float* fc;               // f(centroid)
    int* leftover;           // centroids phase A left over, in no particular order (count: counters->c[10])
    unsigned char* state;    // bit 0: already success; bit 1: still searching after phase A
    float* out;              // projected centroids
    float* nrm_out;          // unit gradient at projected centroids
    Ob02Counters* counters;
    int n;
    int range;               // lane-refill kernels: units per ticket (a multiple of 32; the host shrinks it for small meshes so that the grid fills the GPU)
    int speculative;         // k_project_finish: do nothing if phase A left centroids over (the host then runs phase B and calls again)
};

// produce_facet_normals(force_normalisation = true)  implicit_vectorised_algorithms.hpp:52-133
__device__ __forceinline__ void facet_normal(const float* __restrict__ verts, const int* __restrict__ faces, int fi, float& x, float& y, float& z) {
    const float min_area = (30.0f * ((1.0f / 1000.0f) / 1000.0f)) * (30.0f * ((1.0f / 1000.0f) / 1000.0f));   // configs.hpp:38-48
    const float min_norm_sq = min_area * min_area;
    const int f0 = faces[fi * 3], f1 = faces[fi * 3 + 1], f2 = faces[fi * 3 + 2];
    const float x1 = verts[f1 * 3] - verts[f0 * 3], y1 = verts[f1 * 3 + 1] - verts[f0 * 3 + 1], z1 = verts[f1 * 3 + 2] - verts[f0 * 3 + 2];
    const float x2 = verts[f2 * 3] - verts[f0 * 3], y2 = verts[f2 * 3 + 1] - verts[f0 * 3 + 1], z2 = verts[f2 * 3 + 2] - verts[f0 * 3 + 2];
    x = y1 * z2 - z1 * y2;
    y = z1 * x2 - x1 * z2;
    z = x1 * y2 - y1 * x2;
    const float n2 = x * x + y * y + z * z;
    if (n2 < min_norm_sq) {
        const float o = 1.0f / sqrtf(3.0f);
        x = o; y = o; z = o;
    } else {
        const float n = sqrtf(n2);
        x = x / n; y = y / n; z = z / n;
    }
}

// tries the alphas of one direction; returns true (and the trial point) at the first sign change  (:604-892)
template <bool MASKED>
__device__ __forceinline__ bool try_direction(const SdfGridView& sdf, float cx, float cy, float cz, float dx, float dy, float dz, float edge,
                                              const float* __restrict__ alphas, int n_alpha, float s_c, float& bx, float& by, float& bz, float& bf) {
    for (int ai = 0; ai < n_alpha; ++ai) {
        const float c = edge * __ldg(alphas + ai);          // (length_factor * alpha) * 4.0 / 4.0
        const float x = cx + c * dx, y = cy + c * dy, z = cz + c * dz;
        const float f = eval_f<MASKED>(sdf, x, y, z);
        if (sign_tol(f) * s_c <= 0.0f) { bx = x; by = y; bz = z; bf = f; return true; }
    }
    return false;
}

// set_centers_on_surface, search part, phase A: direction types 0 (gradient) and 1 (facet normal).  The value at the candidate
// is kept beside it (k_project_finish starts from it), and f(centroid) is the value the gradient pass returns: the gradient
// interpreter selects and computes f with the same operations as the value interpreter (tests/test_program_emulation.py holds
// the two equal bit for bit on every shape).  Centroids left over go to a list; phase B works from it.
template <bool MASKED>
__global__ void __launch_bounds__(128) k_project_search(SdfGridView sdf, ProjectArgs a) {
    const int i = blockIdx.x * 128 + threadIdx.x;
    if (i >= a.n) return;
    const float cx = a.cen[i * 3], cy = a.cen[i * 3 + 1], cz = a.cen[i * 3 + 2];
    const float edge = a.al->avg_edge;
    const float* alphas = a.al->alpha;
    const int n_alpha_full = a.al->n_full, n_alpha_short = a.al->n_short;
    float bx = -10.0f, by = -10.0f, bz = -10.0f, bf = 0.0f;
    float fc, g0, g1, g2;
    eval_fg<MASKED>(sdf, cx, cy, cz, fc, g0, g1, g2);
    a.fc[i] = fc;
    // normalise_inplace(g, min_gradient_len)  normalise_inplace.hpp:19-58
    float norm = norm2f(g0, g1, g2);
    const float norm_raw = norm, graw0 = g0, graw1 = g1, graw2 = g2;
    norm = (norm < kMinGradientLen) ? 1.0f : norm;
    const float factor = 1.0f / norm;
    g0 = g0 * factor; g1 = g1 * factor; g2 = g2 * factor;
    const float s_c = sign_tol(fc);
    if (s_c < 0.0f) { g0 = -g0; g1 = -g1; g2 = -g2; }
    const float d0 = -g0 * s_c, d1 = -g1 * s_c, d2 = -g2 * s_c;
    bool ok;
    if (s_c == 0.0f && n_alpha_full > 0) {
        // |f(centroid)| <= ROOT_TOLERANCE: the first trial point passes the reference's test whatever f is there (sign * 0 <= 0; the
        // primitives are NaN-free on finite points), and k_project_finish returns the centroid itself without looking at the
        // candidate or its value — so the trial evaluation is skipped.  The unit gradient the finish would take at its result is
        // the one of this very point: it is written here (same operations on the same evaluation) and state bit 2 says so.
        const float c = edge * __ldg(alphas);
        bx = cx + c * d0; by = cy + c * d1; bz = cz + c * d2; bf = 0.0f;
        ok = true;
        a.nrm_out[i * 3] = graw0 / norm_raw; a.nrm_out[i * 3 + 1] = graw1 / norm_raw; a.nrm_out[i * 3 + 2] = graw2 / norm_raw;
    } else {
        ok = try_direction<MASKED>(sdf, cx, cy, cz, d0, d1, d2, edge, alphas, n_alpha_full, s_c, bx, by, bz, bf);
    }
    int type = 0;
    if (!ok) {
        float nx, ny, nz;
        facet_normal(a.verts, a.faces, i, nx, ny, nz);
        ok = try_direction<MASKED>(sdf, cx, cy, cz, nx, ny, nz, edge, alphas, n_alpha_short, s_c, bx, by, bz, bf);
        type = 1;
    }
    if (ok) {
        a.best[i] = make_float4(bx, by, bz, bf);
        a.state[i] = (s_c == 0.0f && n_alpha_full > 0) ? 5 : 1;
    } else {
        a.state[i] = 2;
    }
    {   // left-overs: one atomic per warp hands out list slots
        const unsigned m = __ballot_sync(__activemask(), !ok);
        if (!ok) {
            const int leader = __ffs(m) - 1, lane = threadIdx.x & 31;
            unsigned long long base = 0ull;
            if (lane == leader) base = atomicAdd(&a.counters->c[10], (unsigned long long)__popc(m));
            base = __shfl_sync(m, base, leader);
            a.leftover[base + __popc(m & ((1u << lane) - 1u))] = i;
        }
    }
    count_if(ok && type == 0, &a.counters->c[0]);
    count_if(ok && type == 1, &a.counters->c[1]);
}

// Phase B: direction types 2..6 for the centroids phase A left over (rare; needs the host-generated noise).  A group of lanes per
// centroid: the candidates (five directions x the short alpha list, in the reference's order of trial) are independent
// evaluations, so the lanes take one each and the first success in that order wins, as in the reference's sequential walk.
// The group is the smallest of 8 / 16 / 32 lanes that holds all candidates at once (10 candidates at C4 1024^3, where the short
// list has two alphas: two centroids per warp), else the whole warp in rounds of 32.
template <bool MASKED>
__global__ void __launch_bounds__(128) k_project_leftovers(SdfGridView sdf, ProjectArgs a, int n_left) {
    const int lane = threadIdx.x & 31;
    const float edge = a.al->avg_edge;
    const float* alphas = a.al->alpha;
    const int n_short = a.al->n_short;
    if (n_short == 0) return;      // mean edge below the reference's smallest step: no candidates, every left-over stays failed
    {   // n_left may be a previous build's count (speculative builds): never walk past what this step listed
        const int listed = (int)a.counters->c[10];
        n_left = n_left < listed ? n_left : listed;
    }
    const int n_cand = 5 * n_short;
    const int G = n_cand > 16 ? 32 : (n_cand > 8 ? 16 : 8), per_warp = 32 / G;
    const int sub = lane / G, gl = lane - sub * G;
    const unsigned gmask = G == 32 ? 0xffffffffu : ((1u << G) - 1u);
    for (int l0 = (blockIdx.x * 4 + (threadIdx.x >> 5)) * per_warp; l0 < n_left; l0 += gridDim.x * 4 * per_warp) {
        const int li = l0 + sub;
        const bool valid = li < n_left;
        const int i = valid ? a.leftover[li] : 0;
        const float cx = a.cen[i * 3], cy = a.cen[i * 3 + 1], cz = a.cen[i * 3 + 2];
        const float s_c = sign_tol(a.fc[i]);
        float dir[5][3];
        {
            float nx, ny, nz;
            facet_normal(a.verts, a.faces, i, nx, ny, nz);
            // type 2: facet normal x noise, normalize_1111   (create_directions_bundle :230-262; add_inplace works on copies)
            const float px = a.noise[i * 3], py = a.noise[i * 3 + 1], pz = a.noise[i * 3 + 2];
            float zx = ny * pz - nz * py, zy = nz * px - nx * pz, zz = nx * py - ny * px;
            {
                const float n = norm2f(zx, zy, zz);
                zx = zx / n; zy = zy / n; zz = zz / n;
            }
            // type 3: facet normal x previous direction, normalise_inplace
            float wx = ny * zz - nz * zy, wy = nz * zx - nx * zz, wz = nx * zy - ny * zx;
            float n = norm2f(wx, wy, wz);
            n = (n < kMinGradientLen) ? 1.0f : n;
            const float factor = 1.0f / n;
            wx = wx * factor; wy = wy * factor; wz = wz * factor;
            dir[0][0] = zx; dir[0][1] = zy; dir[0][2] = zz;
            dir[1][0] = wx; dir[1][1] = wy; dir[1][2] = wz;
    #pragma unroll
            for (int d = 0; d < 3; ++d) { dir[2 + d][0] = d == 0 ? 1.0f : 0.0f; dir[2 + d][1] = d == 1 ? 1.0f : 0.0f; dir[2 + d][2] = d == 2 ? 1.0f : 0.0f; }
        }
        bool ok = false;
        int type = -1;
        float bx = -10.0f, by = -10.0f, bz = -10.0f, bf = 0.0f;
        for (int c0 = 0; __any_sync(0xffffffffu, valid && !ok && c0 < n_cand); c0 += G) {
            const int c = c0 + gl;
            bool mine = false;
            float x = 0.0f, y = 0.0f, z = 0.0f, f = 0.0f;
            if (valid && !ok && c < n_cand) {
                const int d = c / n_short, ai = c - d * n_short;
                float dx = 0.0f, dy = 0.0f, dz = 0.0f;
    #pragma unroll
                for (int k = 0; k < 5; ++k)
                    if (d == k) { dx = dir[k][0]; dy = dir[k][1]; dz = dir[k][2]; }
                const float cc = edge * __ldg(alphas + ai);
                x = cx + cc * dx; y = cy + cc * dy; z = cz + cc * dz;
                f = eval_f<MASKED>(sdf, x, y, z);
                mine = sign_tol(f) * s_c <= 0.0f;
            }
            const unsigned m = (__ballot_sync(0xffffffffu, mine) >> (sub * G)) & gmask;
            const int w = m ? __ffs(m) - 1 : 0, src = sub * G + w;
            const float sx = __shfl_sync(0xffffffffu, x, src), sy = __shfl_sync(0xffffffffu, y, src), sz = __shfl_sync(0xffffffffu, z, src);
            const float sf = __shfl_sync(0xffffffffu, f, src);
            if (m && !ok) {
                bx = sx; by = sy; bz = sz; bf = sf;
                type = 2 + (c0 + w) / n_short;
                ok = true;
            }
        }
        if (gl == 0 && valid) {
            if (ok) {
                a.best[i] = make_float4(bx, by, bz, bf);
                a.state[i] = 1;
                atomicAdd(&a.counters->c[type], 1ull);
            } else {
                a.state[i] = 2;
            }
        }
    }
}

// set_centers_on_surface, second half (:893-1214): zero tests, bisection between the centroid and the candidate,
// final assembly; then the unit gradient at the projected centroid (compute_centroid_gradient, :1262).
template <bool MASKED>
__global__ void __launch_bounds__(128) k_project_finish(SdfGridView sdf, ProjectArgs a) {
    const int i = blockIdx.x * 128 + threadIdx.x;
    if (i >= a.n) return;
    if (a.speculative && a.counters->c[10] != 0ull) return;
    const float cx = a.cen[i * 3], cy = a.cen[i * 3 + 1], cz = a.cen[i * 3 + 2];
    const bool already = (a.state[i] & 1) != 0;
    const float fc = a.fc[i];
    float bx = cx, by = cy, bz = cz, f2 = fc;      // no candidate: the reference evaluates the centroid again
    if (already) {
        const float4 b = a.best[i];
        bx = b.x; by = b.y; bz = b.z; f2 = b.w;    // f at the candidate, from the search that found it
    }
    count_if(!already, &a.counters->c[7]);
    const bool zeros2 = fabsf(f2) <= kRootTolerance;
    const bool zeros1 = fabsf(fc) <= kRootTolerance;
    if (zeros1) { bx = cx; by = cy; bz = cz; }
    const bool zeros12 = zeros2 || zeros1;
    float ox = cx, oy = cy, oz = cz;
    count_if(already && !zeros12, &a.counters->c[9]);
    if (already && !zeros12) {
        float x1x = cx, x1y = cy, x1z = cz, x2x = bx, x2y = by, x2z = bz;
        if (f2 < -kRootTolerance) { x1x = bx; x1y = by; x1z = bz; x2x = cx; x2y = cy; x2z = cz; }
        for (int it = 1;; ++it) {
            const float mx = (x1x + x2x) / 2.0f, my = (x1y + x2y) / 2.0f, mz = (x1z + x2z) / 2.0f;
            const float v = eval_f<MASKED>(sdf, mx, my, mz);
            if (fabsf(v) <= kRootTolerance || it >= kBisectionMaxIter) {
                if (fabsf(v) > kRootTolerance) atomicAdd(&a.counters->c[8], 1ull);
                ox = mx; oy = my; oz = mz;
                break;
            }
            if (v > kRootTolerance) { x2x = mx; x2y = my; x2z = mz; }
            if (v < -kRootTolerance) { x1x = mx; x1y = my; x1z = mz; }
        }
    }
    if (zeros12) { ox = bx; oy = by; oz = bz; }
    a.out[i * 3] = ox; a.out[i * 3 + 1] = oy; a.out[i * 3 + 2] = oz;
    if (a.state[i] & 4) return;      // the result is the centroid and the search has written the unit gradient there
    float f, g0, g1, g2;
    eval_fg<MASKED>(sdf, ox, oy, oz, f, g0, g1, g2);
    const float norm = norm2f(g0, g1, g2);
    a.nrm_out[i * 3] = g0 / norm; a.nrm_out[i * 3 + 1] = g1 / norm; a.nrm_out[i * 3 + 2] = g2 / norm;
}

// ---- two centroids per thread (f32x2 packed pairs, csg_eval.cuh) ----------------------------------------------------------------
// The evaluations above are bound by instruction issue, and more than half of what they issue is not floating point: list
// entries, matrix loads, flag decoding, branches.  Faces 2t and 2t+1 of a marching-cubes mesh come from the same or from
// neighbouring cubes, so their centroids — and the search and bisection points around them — almost always lie in one tile and
// run the SAME instruction list: one thread takes both, one instruction stream serves two points (eval_values_list2<1>,
// eval_value_gradient_list2), and the floating-point work itself goes through FMUL2/FADD2 register pairs.  Every lane performs
// the scalar kernels' operations in their order: the results keep every bit (tests: test_pair_kernels_change_nothing).
//
// A tile's liveness mask was proven over the tile's box grown by one cell; in node-cell terms the box covers the tile's own
// cells and the one layer of cells below them on each axis (classify_box, field_kernel.cuh).  Two points in different tiles can
// therefore still share a list: that of the tile with the larger index on every axis, provided neither point's cell lies more
// than one cell below it.  Otherwise the two points are evaluated one after the other by the scalar interpreters.
__device__ __forceinline__ int pair_slot(const PointSpec& ps, float ax, float ay, float az, float bx, float by, float bz) {
    if (!ps.enabled) return -1;
    const float fax = (ax - ps.g.xmin) / ps.g.wx + 2.0f, fay = (ay - ps.g.ymin) / ps.g.wy + 2.0f, faz = (az - ps.g.zmin) / ps.g.wz + 2.0f;
    const float fbx = (bx - ps.g.xmin) / ps.g.wx + 2.0f, fby = (by - ps.g.ymin) / ps.g.wy + 2.0f, fbz = (bz - ps.g.zmin) / ps.g.wz + 2.0f;
    const float S = (float)ps.g.S, zlo = (float)ps.g.z0, zhi = (float)(ps.g.z0 + ps.g.nz);
    if (!(fax >= 0.0f && fay >= 0.0f && faz >= zlo && fbx >= 0.0f && fby >= 0.0f && fbz >= zlo)) return -1;      // also rejects NaN
    if (!(fax < S && fay < S && faz < zhi && fbx < S && fby < S && fbz < zhi)) return -1;
    const int cax = (int)fax, cay = (int)fay, caz = (int)faz - ps.g.z0, cbx = (int)fbx, cby = (int)fby, cbz = (int)fbz - ps.g.z0;
    const int tx = max(cax, cbx) >> 3, ty = max(cay, cby) >> 3, tz = max(caz, cbz) >> 3;
    if (min(cax, cbx) < tx * 8 - 1 || min(cay, cby) < ty * 8 - 1 || min(caz, cbz) < tz * 8 - 1) return -1;
    const int row = tz * ps.g.ty + ty;
    if (!((__ldg(ps.tile_bits + (size_t)row * ps.TW + (tx >> 5)) >> (tx & 31)) & 1u)) return -1;
    return (int)__ldg(ps.tile_slot + (size_t)row * ps.g.tx + tx);
}
#ifndef ISD_PAIR_INLINE
#define ISD_PAIR_INLINE 1
#endif
#if ISD_PAIR_INLINE
#define ISD_PAIR_FN __device__ __forceinline__
#else
#define ISD_PAIR_FN __device__ __noinline__      // one copy per kernel, but every load inside pays two R2UR for its descriptor
#endif
#ifndef ISD_PAIR_MINBLOCKS
#define ISD_PAIR_MINBLOCKS 6      // 80 registers: measured best for search / finish (5: 96 regs, 7: 72 regs and spills in the search)
#endif
#ifndef ISD_PAIR_MINBLOCKS_CG
#define ISD_PAIR_MINBLOCKS_CG 7
#endif
#ifndef ISD_PAIR_MINBLOCKS_G      // the search's gradient pass (trials elsewhere) and the final unit gradients
#define ISD_PAIR_MINBLOCKS_G 7      // 72 registers: 3.955 against 3.991 ms per C4 1024^3 build with 6 (80), 4.065 with 5 (96)
#endif
// f at two points / f and grad f at two points (.x = first point, .y = second)
ISD_PAIR_FN F2 eval_f_pair(SdfGridView sdf, float ax, float ay, float az, float bx, float by, float bz) {
    const int slot = pair_slot(c_point_spec, ax, ay, az, bx, by, bz);
    if (slot >= 0) {
        const uint32_t* rec = c_point_spec.lists + (size_t)slot * 2 * c_point_spec.stride;
        F2 cx[1] = {F2{ax, bx}}, cy[1] = {F2{ay, by}}, cz[1] = {F2{az, bz}}, out[1];
        eval_values_list2<1>(ConstProgram(), sdf, GlobalList{rec + 1}, (int)__ldg(rec), cx, cy, cz, out);
        if (c_count_flops) add_flops(2ull * list_flops(ConstProgram(), GlobalList{rec + 1}, (int)__ldg(rec), false));
        return out[0];
    }
    return F2{eval_f<true>(sdf, ax, ay, az), eval_f<true>(sdf, bx, by, bz)};
}
struct FG2 { F2 f, g0, g1, g2; };
ISD_PAIR_FN FG2 eval_fg_pair(SdfGridView sdf, float ax, float ay, float az, float bx, float by, float bz) {
    FG2 r;
    const int slot = pair_slot(c_point_spec, ax, ay, az, bx, by, bz);
    if (slot >= 0) {
        const uint32_t* rec = c_point_spec.lists + ((size_t)slot * 2 + 1) * c_point_spec.stride;
        eval_value_gradient_list2(ConstProgram(), sdf, GlobalList{rec + 1}, (int)__ldg(rec), F2{ax, bx}, F2{ay, by}, F2{az, bz}, r.f, r.g0, r.g1, r.g2);
        if (c_count_flops) add_flops(2ull * list_flops(ConstProgram(), GlobalList{rec + 1}, (int)__ldg(rec), true));
        return r;
    }
    eval_fg<true>(sdf, ax, ay, az, r.f.x, r.g0.x, r.g1.x, r.g2.x);
    eval_fg<true>(sdf, bx, by, bz, r.f.y, r.g0.y, r.g1.y, r.g2.y);
    return r;
}

// k_centroid_gradients, faces 2t and 2t+1 per thread
__global__ void __launch_bounds__(128, ISD_PAIR_MINBLOCKS_CG) k_centroid_gradients_pair(SdfGridView sdf, const float* __restrict__ verts, const int* __restrict__ faces, int nf,
                                                                 float* __restrict__ cen, float* __restrict__ nrm, FaceOrder order) {
    const long long u = (long long)blockIdx.x * 128 + threadIdx.x;
    if (u >= unit_count(order, nf)) return;
    int fa, fb;
    unit_faces(order, u, nf, fa, fb);
    const bool two = fb >= 0;
    float c[2][3];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const int f = (k == 1 && two) ? fb : fa;      // a unit of one: the thread evaluates its face twice
        const int f0 = faces[f * 3], f1 = faces[f * 3 + 1], f2 = faces[f * 3 + 2];
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            c[k][d] = (verts[f0 * 3 + d] + verts[f1 * 3 + d] + verts[f2 * 3 + d]) / 3.0f;
            cen[f * 3 + d] = c[k][d];
        }
    }
    const FG2 r = eval_fg_pair(sdf, c[0][0], c[0][1], c[0][2], c[1][0], c[1][1], c[1][2]);
    {
        const float norm = norm2f(r.g0.x, r.g1.x, r.g2.x);
        nrm[fa * 3] = r.g0.x / norm; nrm[fa * 3 + 1] = r.g1.x / norm; nrm[fa * 3 + 2] = r.g2.x / norm;
    }
    if (two) {
        const float norm = norm2f(r.g0.y, r.g1.y, r.g2.y);
        nrm[fb * 3] = r.g0.y / norm; nrm[fb * 3 + 1] = r.g1.y / norm; nrm[fb * 3 + 2] = r.g2.y / norm;
    }
}

// k_project_search, centroids 2t and 2t+1 per thread.  Each centroid's result is the first sign change along its own sequence of
// trial points (direction type 0 over the whole alpha list, then type 1 over the short list), exactly as in the scalar kernel;
// the two sequences advance in lockstep and a lane that has finished a phase rides along on its partner's point.
// TRIALS_LATER: the kernel stops after the gradient pass; a centroid that needs trial points is left pending (state 0, its search
// direction and the sign of f in `best`) for k_search_trials_refill.
template <bool TRIALS_LATER>
__global__ void __launch_bounds__(128, TRIALS_LATER ? ISD_PAIR_MINBLOCKS_G : ISD_PAIR_MINBLOCKS) k_project_search_pair(SdfGridView sdf, ProjectArgs a) {
    const long long u = (long long)blockIdx.x * 128 + threadIdx.x;
    if (u >= unit_count(a.order, a.n)) return;
    int ia, ib;
    unit_faces(a.order, u, a.n, ia, ib);
    const bool two = ib >= 0;
    const int idx[2] = {ia, two ? ib : ia};
    const float edge = a.al->avg_edge;
    const float* alphas = a.al->alpha;
    const int n_alpha_full = a.al->n_full, n_alpha_short = a.al->n_short;
    float c[2][3], d[2][3], s_c[2], best[2][4];
    bool ok[2] = {false, false}, settled[2] = {false, false};
    int type[2] = {0, 0};
#pragma unroll
    for (int k = 0; k < 2; ++k) {
#pragma unroll
        for (int q = 0; q < 3; ++q) c[k][q] = a.cen[idx[k] * 3 + q];
        best[k][0] = -10.0f; best[k][1] = -10.0f; best[k][2] = -10.0f; best[k][3] = 0.0f;
    }
    {
        const FG2 r = eval_fg_pair(sdf, c[0][0], c[0][1], c[0][2], c[1][0], c[1][1], c[1][2]);
        const float fc[2] = {r.f.x, r.f.y};
        float g[2][3] = {{r.g0.x, r.g1.x, r.g2.x}, {r.g0.y, r.g1.y, r.g2.y}};
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            if (k == 0 || two) a.fc[idx[k]] = fc[k];
            float norm = norm2f(g[k][0], g[k][1], g[k][2]);      // normalise_inplace(g, min_gradient_len)  normalise_inplace.hpp:19-58
            norm = (norm < kMinGradientLen) ? 1.0f : norm;
            const float factor = 1.0f / norm;
            float g0 = g[k][0] * factor, g1 = g[k][1] * factor, g2 = g[k][2] * factor;
            s_c[k] = sign_tol(fc[k]);
            if (s_c[k] < 0.0f) { g0 = -g0; g1 = -g1; g2 = -g2; }
            d[k][0] = -g0 * s_c[k]; d[k][1] = -g1 * s_c[k]; d[k][2] = -g2 * s_c[k];
            if (s_c[k] == 0.0f && n_alpha_full > 0) {      // see k_project_search: no trial evaluation, the unit gradient is final
                const float cc = edge * __ldg(alphas);
                best[k][0] = c[k][0] + cc * d[k][0]; best[k][1] = c[k][1] + cc * d[k][1]; best[k][2] = c[k][2] + cc * d[k][2]; best[k][3] = 0.0f;
                ok[k] = true; settled[k] = true;
                if (k == 0 || two) {
                    const float nr = norm2f(g[k][0], g[k][1], g[k][2]);
                    a.nrm_out[idx[k] * 3] = g[k][0] / nr; a.nrm_out[idx[k] * 3 + 1] = g[k][1] / nr; a.nrm_out[idx[k] * 3 + 2] = g[k][2] / nr;
                }
            }
        }
    }
    const bool pending_mode = TRIALS_LATER && (n_alpha_full > 0 || n_alpha_short > 0);
    for (int phase = 0; phase < 2 && !pending_mode; ++phase) {
        if (phase == 1) {
            if (ok[0] && ok[1]) break;
#pragma unroll
            for (int k = 0; k < 2; ++k)
                if (!ok[k]) { facet_normal(a.verts, a.faces, idx[k], d[k][0], d[k][1], d[k][2]); type[k] = 1; }
        }
        const int n_alpha = phase == 0 ? n_alpha_full : n_alpha_short;
        bool done[2] = {ok[0], ok[1]};      // finished with this phase
        for (int ai = 0; ai < n_alpha && !(done[0] && done[1]); ++ai) {
            const float cc = edge * __ldg(alphas + ai);          // (length_factor * alpha) * 4.0 / 4.0
            float p[2][3];
#pragma unroll
            for (int k = 0; k < 2; ++k)
#pragma unroll
                for (int q = 0; q < 3; ++q) p[k][q] = c[k][q] + cc * d[k][q];
            if (done[0]) { p[0][0] = p[1][0]; p[0][1] = p[1][1]; p[0][2] = p[1][2]; }
            if (done[1]) { p[1][0] = p[0][0]; p[1][1] = p[0][1]; p[1][2] = p[0][2]; }
            const F2 f = eval_f_pair(sdf, p[0][0], p[0][1], p[0][2], p[1][0], p[1][1], p[1][2]);
            const float fv[2] = {f.x, f.y};
#pragma unroll
            for (int k = 0; k < 2; ++k)
                if (!done[k] && sign_tol(fv[k]) * s_c[k] <= 0.0f) {
                    best[k][0] = p[k][0]; best[k][1] = p[k][1]; best[k][2] = p[k][2]; best[k][3] = fv[k];
                    ok[k] = true; done[k] = true;
                }
        }
    }
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        if (k == 1 && !two) break;
        if (ok[k]) {
            a.best[idx[k]] = make_float4(best[k][0], best[k][1], best[k][2], best[k][3]);
            a.state[idx[k]] = settled[k] ? 5 : 1;
        } else if (pending_mode) {
            a.best[idx[k]] = make_float4(d[k][0], d[k][1], d[k][2], s_c[k]);
            a.state[idx[k]] = 0;
        } else {
            a.state[idx[k]] = 2;
        }
    }
#pragma unroll
    for (int k = 0; k < 2; ++k) {   // left-overs: one atomic per warp hands out list slots
        const bool left = !ok[k] && !pending_mode && (k == 0 || two);
        const unsigned m = __ballot_sync(__activemask(), left);
        if (left) {
            const int leader = __ffs(m) - 1, lane = threadIdx.x & 31;
            unsigned long long base = 0ull;
            if (lane == leader) base = atomicAdd(&a.counters->c[10], (unsigned long long)__popc(m));
            base = __shfl_sync(m, base, leader);
            a.leftover[base + __popc(m & ((1u << lane) - 1u))] = idx[k];
        }
        const bool mine = ok[k] && (k == 0 || two);
        count_if(mine && type[k] == 0, &a.counters->c[0]);
        count_if(mine && type[k] == 1, &a.counters->c[1]);
    }
}

// k_project_finish, centroids 2t and 2t+1 per thread: the two bisections advance in lockstep, each with its own interval and
// its own stopping test; a lane that has finished rides along on its partner's midpoint.
__global__ void __launch_bounds__(128, ISD_PAIR_MINBLOCKS) k_project_finish_pair(SdfGridView sdf, ProjectArgs a) {
    const long long u = (long long)blockIdx.x * 128 + threadIdx.x;
    if (u >= unit_count(a.order, a.n)) return;
    if (a.speculative && a.counters->c[10] != 0ull) return;
    int ia, ib;
    unit_faces(a.order, u, a.n, ia, ib);
    const bool two = ib >= 0;
    const int idx[2] = {ia, two ? ib : ia};
    float o[2][3], x1[2][3], x2[2][3];
    bool zeros12[2], active[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const int i = idx[k];
        const bool mine = k == 0 || two;
        const float cx = a.cen[i * 3], cy = a.cen[i * 3 + 1], cz = a.cen[i * 3 + 2];
        const bool already = (a.state[i] & 1) != 0;
        const float fc = a.fc[i];
        float bx = cx, by = cy, bz = cz, f2 = fc;      // no candidate: the reference evaluates the centroid again
        if (already) {
            const float4 b = a.best[i];
            bx = b.x; by = b.y; bz = b.z; f2 = b.w;    // f at the candidate, from the search that found it
        }
        count_if(mine && !already, &a.counters->c[7]);
        const bool zeros2 = fabsf(f2) <= kRootTolerance;
        const bool zeros1 = fabsf(fc) <= kRootTolerance;
        if (zeros1) { bx = cx; by = cy; bz = cz; }
        zeros12[k] = zeros2 || zeros1;
        o[k][0] = cx; o[k][1] = cy; o[k][2] = cz;
        active[k] = already && !zeros12[k];
        count_if(mine && active[k], &a.counters->c[9]);
        x1[k][0] = cx; x1[k][1] = cy; x1[k][2] = cz; x2[k][0] = bx; x2[k][1] = by; x2[k][2] = bz;
        if (f2 < -kRootTolerance) { x1[k][0] = bx; x1[k][1] = by; x1[k][2] = bz; x2[k][0] = cx; x2[k][1] = cy; x2[k][2] = cz; }
        if (zeros12[k]) { o[k][0] = bx; o[k][1] = by; o[k][2] = bz; }
    }
    for (int it = 1; active[0] || active[1]; ++it) {
        float m[2][3];
#pragma unroll
        for (int k = 0; k < 2; ++k)
#pragma unroll
            for (int q = 0; q < 3; ++q) m[k][q] = (x1[k][q] + x2[k][q]) / 2.0f;
        if (!active[0]) { m[0][0] = m[1][0]; m[0][1] = m[1][1]; m[0][2] = m[1][2]; }
        if (!active[1]) { m[1][0] = m[0][0]; m[1][1] = m[0][1]; m[1][2] = m[0][2]; }
        const F2 f = eval_f_pair(sdf, m[0][0], m[0][1], m[0][2], m[1][0], m[1][1], m[1][2]);
        const float v[2] = {f.x, f.y};
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            if (!active[k]) continue;
            if (fabsf(v[k]) <= kRootTolerance || it >= kBisectionMaxIter) {
                if (fabsf(v[k]) > kRootTolerance && (k == 0 || two)) atomicAdd(&a.counters->c[8], 1ull);
                o[k][0] = m[k][0]; o[k][1] = m[k][1]; o[k][2] = m[k][2];
                active[k] = false;
                continue;
            }
            if (v[k] > kRootTolerance) { x2[k][0] = m[k][0]; x2[k][1] = m[k][1]; x2[k][2] = m[k][2]; }
            if (v[k] < -kRootTolerance) { x1[k][0] = m[k][0]; x1[k][1] = m[k][1]; x1[k][2] = m[k][2]; }
        }
    }
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        if (k == 1 && !two) break;
        a.out[idx[k] * 3] = o[k][0]; a.out[idx[k] * 3 + 1] = o[k][1]; a.out[idx[k] * 3 + 2] = o[k][2];
    }
    if ((a.state[idx[0]] & 4) && (a.state[idx[1]] & 4)) return;      // both unit gradients were written by the search
    const FG2 r = eval_fg_pair(sdf, o[0][0], o[0][1], o[0][2], o[1][0], o[1][1], o[1][2]);
    const float g[2][3] = {{r.g0.x, r.g1.x, r.g2.x}, {r.g0.y, r.g1.y, r.g2.y}};
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        if (k == 1 && !two) break;
        const int i = idx[k];
        const float norm = norm2f(g[k][0], g[k][1], g[k][2]);
        a.nrm_out[i * 3] = g[k][0] / norm; a.nrm_out[i * 3 + 1] = g[k][1] / norm; a.nrm_out[i * 3 + 2] = g[k][2] / norm;
    }
}

// k_project_finish in two passes (pair path, "gpu":{"finish_split":1}).  Only a third of the centroids of a typical mesh
// bisect at all (the others already sit within ROOT_TOLERANCE at the centroid or at the candidate), so inside the one-kernel form
// about half of the lanes — those whose two centroids both need nothing — idle through every bisection step.  Pass 1
// (k_finish_bisect_units): a warp walks a contiguous range of UNITS (faces 2u and 2u+1: the pair the other pair kernels take,
// neighbours in space that share a tile list) 32 at a time, settles the centroids that need no bisection on the spot and queues
// the units with at least one that does (ballot ranks: mesh order is kept); whenever 32 units are waiting, every lane takes one
// and bisects its two centroids in lockstep, a member that needs nothing riding along as in k_project_finish_pair.  Pass 2
// (k_unit_gradients_pair) takes the unit gradient at all projected centroids.  Per centroid the operations are those of
// k_project_finish, in the same order.
constexpr int kFinishRange = 512;      // units per warp
__device__ __forceinline__ void bisect_unit(const SdfGridView& sdf, const ProjectArgs& a, int entry) {
    const int u = entry >> 2;
    const int idx[2] = {2 * u, 2 * u + 1};
    bool active[2] = {(entry & 1) != 0, (entry & 2) != 0};
    const bool need[2] = {active[0], active[1]};
    float o[2][3], x1[2][3], x2[2][3];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
#pragma unroll
        for (int q = 0; q < 3; ++q) { o[k][q] = 0.0f; x1[k][q] = 0.0f; x2[k][q] = 0.0f; }
        if (!need[k]) continue;
        const int i = idx[k];
        const float cx = a.cen[i * 3], cy = a.cen[i * 3 + 1], cz = a.cen[i * 3 + 2];
        const float4 b = a.best[i];
        x1[k][0] = cx; x1[k][1] = cy; x1[k][2] = cz; x2[k][0] = b.x; x2[k][1] = b.y; x2[k][2] = b.z;
        if (b.w < -kRootTolerance) { x1[k][0] = b.x; x1[k][1] = b.y; x1[k][2] = b.z; x2[k][0] = cx; x2[k][1] = cy; x2[k][2] = cz; }
    }
    for (int it = 1; active[0] || active[1]; ++it) {
        float m[2][3];
#pragma unroll
        for (int k = 0; k < 2; ++k)
#pragma unroll
            for (int q = 0; q < 3; ++q) m[k][q] = (x1[k][q] + x2[k][q]) / 2.0f;
        if (!active[0]) { m[0][0] = m[1][0]; m[0][1] = m[1][1]; m[0][2] = m[1][2]; }
        if (!active[1]) { m[1][0] = m[0][0]; m[1][1] = m[0][1]; m[1][2] = m[0][2]; }
        const F2 f = eval_f_pair(sdf, m[0][0], m[0][1], m[0][2], m[1][0], m[1][1], m[1][2]);
        const float v[2] = {f.x, f.y};
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            if (!active[k]) continue;
            if (fabsf(v[k]) <= kRootTolerance || it >= kBisectionMaxIter) {
                if (fabsf(v[k]) > kRootTolerance) atomicAdd(&a.counters->c[8], 1ull);
                o[k][0] = m[k][0]; o[k][1] = m[k][1]; o[k][2] = m[k][2];
                active[k] = false;
                continue;
            }
            if (v[k] > kRootTolerance) { x2[k][0] = m[k][0]; x2[k][1] = m[k][1]; x2[k][2] = m[k][2]; }
            if (v[k] < -kRootTolerance) { x1[k][0] = m[k][0]; x1[k][1] = m[k][1]; x1[k][2] = m[k][2]; }
        }
    }
#pragma unroll
    for (int k = 0; k < 2; ++k)
        if (need[k]) { a.out[idx[k] * 3] = o[k][0]; a.out[idx[k] * 3 + 1] = o[k][1]; a.out[idx[k] * 3 + 2] = o[k][2]; }
}
#ifndef ISD_FINISH_UNITS_MINBLOCKS
#define ISD_FINISH_UNITS_MINBLOCKS 5
#endif
__global__ void __launch_bounds__(128, ISD_FINISH_UNITS_MINBLOCKS) k_finish_bisect_units(SdfGridView sdf, ProjectArgs a) {
    if (a.speculative && a.counters->c[10] != 0ull) return;
    __shared__ int s_queue[4][64];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int* queue = s_queue[warp];
    const int n_units = (a.n + 1) / 2;
    const long long first = ((long long)blockIdx.x * 4 + warp) * kFinishRange;
    const long long last = first + kFinishRange < (long long)n_units ? first + kFinishRange : (long long)n_units;
    int waiting = 0;
    unsigned n_failed = 0u, n_bisect = 0u;
    for (long long base = first; base < last; base += 32) {
        const long long u = base + lane;
        int needs = 0;
        if (u < last) {
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const long long i = 2 * u + k;
                if (i >= (long long)a.n) break;
                const float cx = a.cen[i * 3], cy = a.cen[i * 3 + 1], cz = a.cen[i * 3 + 2];
                const bool already = (a.state[i] & 1) != 0;
                const float fc = a.fc[i];
                float bx = cx, by = cy, bz = cz, f2 = fc;      // no candidate: the reference evaluates the centroid again
                if (already) {
                    const float4 b = a.best[i];
                    bx = b.x; by = b.y; bz = b.z; f2 = b.w;    // f at the candidate, from the search that found it
                }
                const bool zeros2 = fabsf(f2) <= kRootTolerance;
                const bool zeros1 = fabsf(fc) <= kRootTolerance;
                if (zeros1) { bx = cx; by = cy; bz = cz; }
                const bool zeros12 = zeros2 || zeros1;
                const bool bisect = already && !zeros12;
                if (!bisect) {
                    float ox = cx, oy = cy, oz = cz;
                    if (zeros12) { ox = bx; oy = by; oz = bz; }
                    a.out[i * 3] = ox; a.out[i * 3 + 1] = oy; a.out[i * 3 + 2] = oz;
                }
                if (!already) ++n_failed;
                if (bisect) { ++n_bisect; needs |= 1 << k; }
            }
        }
        const unsigned m = __ballot_sync(0xffffffffu, needs != 0);
        if (needs) queue[waiting + __popc(m & ((1u << lane) - 1u))] = (int)(u << 2) | needs;
        waiting += __popc(m);
        __syncwarp();
        if (waiting >= 32) {
            bisect_unit(sdf, a, queue[lane]);
            __syncwarp();
            const int rest = waiting - 32;      // < 32: move the tail to the front
            const int moved = lane < rest ? queue[32 + lane] : 0;
            __syncwarp();
            if (lane < rest) queue[lane] = moved;
            waiting = rest;
            __syncwarp();
        }
    }
    if (lane < waiting) bisect_unit(sdf, a, queue[lane]);      // what is left at the end of the range
    // counters: one atomic per warp each
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        n_failed += __shfl_xor_sync(0xffffffffu, n_failed, d);
        n_bisect += __shfl_xor_sync(0xffffffffu, n_bisect, d);
    }
    if (lane == 0) {
        if (n_failed) atomicAdd(&a.counters->c[7], (unsigned long long)n_failed);
        if (n_bisect) atomicAdd(&a.counters->c[9], (unsigned long long)n_bisect);
    }
}
// Trial points of the projection's search with LANE REFILL ("gpu":{"search_split":1}, default on the pair path).  After its
// gradient pass k_project_search_pair walked the two centroids of a thread through their trial points in lockstep: a settled member
// rode along, and a warp ran as long as its slowest lane (ncu, C4 1024^3: 13 of 32 lanes active in that part).  Here the pending
// centroids are queued per warp exactly as in k_finish_bisect_refill below (persistent grid, ranges from a ticket, counters->c[12])
// and a lane that has resolved its unit takes the next one.  Per centroid the sequence of trial points is the reference's:
// direction type 0 over the whole alpha list, then the facet normal over the short list; the first sign change wins.
constexpr int kRefillRange = 64;       // units per ticket at most (both lane-refill kernels; ProjectArgs::range)
#ifndef ISD_TRIALS_REFILL_MINBLOCKS
#define ISD_TRIALS_REFILL_MINBLOCKS 6
#endif
__global__ void __launch_bounds__(128, ISD_TRIALS_REFILL_MINBLOCKS) k_search_trials_refill(SdfGridView sdf, ProjectArgs a) {
    __shared__ int s_queue[4][128];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int* queue = s_queue[warp];
    const long long n_units = unit_count(a.order, a.n);
    const float edge = a.al->avg_edge;
    const float* alphas = a.al->alpha;
    const int n_alpha_full = a.al->n_full, n_alpha_short = a.al->n_short;
    long long next = 0, last = 0;
    bool exhausted = false;
    int head = 0, waiting = 0;
    unsigned n_type0 = 0u, n_type1 = 0u;
#ifdef ISD_TRIALS_DIAG
    unsigned long long diag_steps = 0ull, diag_max_cyc = 0ull, diag_t0 = clock64();
#endif
    int entry = -1;      // this lane's unit (u << 2 | members that search), or -1: idle
    bool active[2] = {false, false};
    int face[2] = {0, 0}, ai[2] = {0, 0}, phase[2] = {0, 0};
    float c[2][3], d[2][3], s_c[2] = {0.0f, 0.0f};
#pragma unroll
    for (int k = 0; k < 2; ++k)
#pragma unroll
        for (int q = 0; q < 3; ++q) { c[k][q] = 0.0f; d[k][q] = 0.0f; }
    for (;;) {
        const unsigned idle = __ballot_sync(0xffffffffu, entry < 0);
        const int n_idle = __popc(idle);
        while (waiting < n_idle && !exhausted) {      // classify the next 32 units
            if (next >= last) {      // draw another range
                unsigned long long t = 0ull;
                if (lane == 0) t = atomicAdd(&a.counters->c[12], 1ull);
                t = __shfl_sync(0xffffffffu, t, 0);
                next = (long long)t * a.range;
                last = next + a.range < n_units ? next + a.range : n_units;
                if (next >= n_units) { exhausted = true; break; }
            }
            const long long u = next + lane;
            int needs = 0;
            if (u < last) {
                int uf[2];
                unit_faces(a.order, u, a.n, uf[0], uf[1]);
#pragma unroll
                for (int k = 0; k < 2; ++k)
                    if (uf[k] >= 0 && a.state[uf[k]] == 0) needs |= 1 << k;
            }
            const unsigned m = __ballot_sync(0xffffffffu, needs != 0);
            if (needs) queue[(head + waiting + __popc(m & ((1u << lane) - 1u))) & 127] = (int)(u << 2) | needs;
            waiting += __popc(m);
            next += 32;
            __syncwarp();
        }
        if (entry < 0) {      // idle lanes take the waiting units in lane order
            const int r = __popc(idle & ((1u << lane) - 1u));
            if (r < waiting) {
                entry = queue[(head + r) & 127];
                unit_faces(a.order, entry >> 2, a.n, face[0], face[1]);
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    active[k] = ((entry >> k) & 1) != 0;
                    if (!active[k]) continue;
                    const long long i = face[k];
                    c[k][0] = a.cen[i * 3]; c[k][1] = a.cen[i * 3 + 1]; c[k][2] = a.cen[i * 3 + 2];
                    const float4 b = a.best[i];
                    d[k][0] = b.x; d[k][1] = b.y; d[k][2] = b.z; s_c[k] = b.w;
                    ai[k] = 0; phase[k] = 0;
                }
            }
        }
        {
            const int taken = waiting < n_idle ? waiting : n_idle;
            head = (head + taken) & 127;
            waiting -= taken;
        }
        __syncwarp();
        if (__ballot_sync(0xffffffffu, entry >= 0) == 0u) break;      // no unit in flight, none waiting, range exhausted
        if (entry >= 0) {
            // a member at the end of a direction's alphas moves on: type 0 -> the facet normal (type 1) -> left over
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                if (!active[k]) continue;
                if (phase[k] == 0 && ai[k] >= n_alpha_full) {
                    phase[k] = 1; ai[k] = 0;
                    facet_normal(a.verts, a.faces, face[k], d[k][0], d[k][1], d[k][2]);
                }
                const bool failed = phase[k] == 1 && ai[k] >= n_alpha_short;
                if (failed) { a.state[face[k]] = 2; active[k] = false; }
                {   // left over: one atomic per warp hands out the list slots of the lanes that fail in this step
                    const unsigned fm = __ballot_sync(__activemask(), failed);
                    if (failed) {
                        const int leader = __ffs(fm) - 1;
                        unsigned long long base = 0ull;
                        if (lane == leader) base = atomicAdd(&a.counters->c[10], (unsigned long long)__popc(fm));
                        base = __shfl_sync(fm, base, leader);
                        a.leftover[base + __popc(fm & ((1u << lane) - 1u))] = face[k];
                    }
                }
            }
            if (active[0] || active[1]) {      // one trial point per searching member
                float p[2][3];
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    const float cc = edge * __ldg(alphas + (active[k] ? ai[k] : 0));          // (length_factor * alpha) * 4.0 / 4.0
#pragma unroll
                    for (int q = 0; q < 3; ++q) p[k][q] = c[k][q] + cc * d[k][q];
                }
                if (!active[0]) { p[0][0] = p[1][0]; p[0][1] = p[1][1]; p[0][2] = p[1][2]; }
                if (!active[1]) { p[1][0] = p[0][0]; p[1][1] = p[0][1]; p[1][2] = p[0][2]; }
#ifdef ISD_TRIALS_DIAG
                const unsigned long long diag_a = clock64();
#endif
                const F2 f = eval_f_pair(sdf, p[0][0], p[0][1], p[0][2], p[1][0], p[1][1], p[1][2]);
                const float fv[2] = {f.x, f.y};
#ifdef ISD_TRIALS_DIAG
                { const unsigned long long dc = clock64() - diag_a; diag_max_cyc = dc > diag_max_cyc ? dc : diag_max_cyc; ++diag_steps; }
#endif
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    if (!active[k]) continue;
                    if (sign_tol(fv[k]) * s_c[k] <= 0.0f) {
                        a.best[face[k]] = make_float4(p[k][0], p[k][1], p[k][2], fv[k]);
                        a.state[face[k]] = 1;
                        if (phase[k] == 0) ++n_type0; else ++n_type1;
                        active[k] = false;
                    } else {
                        ++ai[k];
                    }
                }
            }
            if (!active[0] && !active[1]) entry = -1;
        }
    }
#pragma unroll
    for (int dd = 16; dd > 0; dd >>= 1) {
        n_type0 += __shfl_xor_sync(0xffffffffu, n_type0, dd);
        n_type1 += __shfl_xor_sync(0xffffffffu, n_type1, dd);
    }
#ifdef ISD_TRIALS_DIAG
    {      // diagnostic build only: the warps that lived longest
        const unsigned long long life = clock64() - diag_t0;
        if (lane == 0 && life > 600000ull)
            printf("trials warp %d.%d: life %llu cyc, steps %llu, slowest step %llu cyc, last ticket start %lld of %lld units\n", (int)blockIdx.x, warp, life, diag_steps,
                   diag_max_cyc, next, n_units);
    }
#endif
    if (lane == 0) {
        if (n_type0) atomicAdd(&a.counters->c[0], (unsigned long long)n_type0);
        if (n_type1) atomicAdd(&a.counters->c[1], (unsigned long long)n_type1);
    }
}

// The same pass with LANE REFILL ("gpu":{"finish_split":2}).  In k_project_finish_pair a lane bisects for as long as the slowest
// centroid of its warp (ncu, C4 1024^3: 13.7 of 32 lanes active in the bisection part); the batch form above fills all lanes at the
// start of a batch but still runs every batch to its slowest member.  Here a lane that has finished its unit takes the next one
// from the warp's queue at once, so every evaluation runs with (nearly) all lanes busy; the queue is topped up from the warp's
// range of units, 32 classified at a time, whenever fewer units wait than lanes are idle.  Per centroid the sequence of
// midpoints, tests and the iteration cap are those of k_project_finish.
#ifndef ISD_FINISH_REFILL_MINBLOCKS
#define ISD_FINISH_REFILL_MINBLOCKS 6
#endif
// Persistent grid (one block per resident slot); a warp draws ranges of ProjectArgs::range units (64, or 32 for small meshes) from a ticket (counters->c[11], zeroed
// with the other counters at the start of the projection step) for as long as it has idle lanes, so lanes stay busy across range
// boundaries and only the very end of the kernel runs with part of the lanes.
__global__ void __launch_bounds__(128, ISD_FINISH_REFILL_MINBLOCKS) k_finish_bisect_refill(SdfGridView sdf, ProjectArgs a) {
    if (a.speculative && a.counters->c[10] != 0ull) return;
    __shared__ int s_queue[4][128];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int* queue = s_queue[warp];
    const long long n_units = unit_count(a.order, a.n);
    long long next = 0, last = 0;
    bool exhausted = false;
    int head = 0, waiting = 0;
    unsigned n_failed = 0u, n_bisect = 0u;
    int entry = -1, it = 0;      // this lane's unit (u << 2 | members that bisect), or -1: idle
    bool active[2] = {false, false};
    int face[2] = {0, 0};        // the unit's faces
    float x1[2][3], x2[2][3];
#pragma unroll
    for (int k = 0; k < 2; ++k)
#pragma unroll
        for (int q = 0; q < 3; ++q) { x1[k][q] = 0.0f; x2[k][q] = 0.0f; }
    for (;;) {
        const unsigned idle = __ballot_sync(0xffffffffu, entry < 0);
        const int n_idle = __popc(idle);
        while (waiting < n_idle && !exhausted) {      // classify the next 32 units
            if (next >= last) {      // draw another range
                unsigned long long t = 0ull;
                if (lane == 0) t = atomicAdd(&a.counters->c[11], 1ull);
                t = __shfl_sync(0xffffffffu, t, 0);
                next = (long long)t * a.range;
                last = next + a.range < n_units ? next + a.range : n_units;
                if (next >= n_units) { exhausted = true; break; }
            }
            const long long u = next + lane;
            int needs = 0;
            if (u < last) {
                int uf[2];
                unit_faces(a.order, u, a.n, uf[0], uf[1]);
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    const long long i = uf[k];
                    if (i < 0) break;
                    const float cx = a.cen[i * 3], cy = a.cen[i * 3 + 1], cz = a.cen[i * 3 + 2];
                    const bool already = (a.state[i] & 1) != 0;
                    const float fc = a.fc[i];
                    float bx = cx, by = cy, bz = cz, f2 = fc;      // no candidate: the reference evaluates the centroid again
                    if (already) {
                        const float4 b = a.best[i];
                        bx = b.x; by = b.y; bz = b.z; f2 = b.w;    // f at the candidate, from the search that found it
                    }
                    const bool zeros2 = fabsf(f2) <= kRootTolerance;
                    const bool zeros1 = fabsf(fc) <= kRootTolerance;
                    if (zeros1) { bx = cx; by = cy; bz = cz; }
                    const bool zeros12 = zeros2 || zeros1;
                    const bool bisect = already && !zeros12;
                    if (!bisect) {
                        float ox = cx, oy = cy, oz = cz;
                        if (zeros12) { ox = bx; oy = by; oz = bz; }
                        a.out[i * 3] = ox; a.out[i * 3 + 1] = oy; a.out[i * 3 + 2] = oz;
                    }
                    if (!already) ++n_failed;
                    if (bisect) { ++n_bisect; needs |= 1 << k; }
                }
            }
            const unsigned m = __ballot_sync(0xffffffffu, needs != 0);
            if (needs) queue[(head + waiting + __popc(m & ((1u << lane) - 1u))) & 127] = (int)(u << 2) | needs;
            waiting += __popc(m);
            next += 32;
            __syncwarp();
        }
        if (entry < 0) {      // idle lanes take the waiting units in lane order
            const int r = __popc(idle & ((1u << lane) - 1u));
            if (r < waiting) {
                entry = queue[(head + r) & 127];
                it = 1;
                unit_faces(a.order, entry >> 2, a.n, face[0], face[1]);
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    active[k] = ((entry >> k) & 1) != 0;
                    if (!active[k]) continue;
                    const long long i = face[k];
                    const float cx = a.cen[i * 3], cy = a.cen[i * 3 + 1], cz = a.cen[i * 3 + 2];
                    const float4 b = a.best[i];
                    x1[k][0] = cx; x1[k][1] = cy; x1[k][2] = cz; x2[k][0] = b.x; x2[k][1] = b.y; x2[k][2] = b.z;
                    if (b.w < -kRootTolerance) { x1[k][0] = b.x; x1[k][1] = b.y; x1[k][2] = b.z; x2[k][0] = cx; x2[k][1] = cy; x2[k][2] = cz; }
                }
            }
        }
        {
            const int taken = waiting < n_idle ? waiting : n_idle;
            head = (head + taken) & 127;
            waiting -= taken;
        }
        __syncwarp();
        if (__ballot_sync(0xffffffffu, entry >= 0) == 0u) break;      // no unit in flight, none waiting, range exhausted
        if (entry >= 0) {      // one bisection step of this lane's unit
            float m[2][3];
#pragma unroll
            for (int k = 0; k < 2; ++k)
#pragma unroll
                for (int q = 0; q < 3; ++q) m[k][q] = (x1[k][q] + x2[k][q]) / 2.0f;
            if (!active[0]) { m[0][0] = m[1][0]; m[0][1] = m[1][1]; m[0][2] = m[1][2]; }
            if (!active[1]) { m[1][0] = m[0][0]; m[1][1] = m[0][1]; m[1][2] = m[0][2]; }
            const F2 f = eval_f_pair(sdf, m[0][0], m[0][1], m[0][2], m[1][0], m[1][1], m[1][2]);
            const float v[2] = {f.x, f.y};
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                if (!active[k]) continue;
                if (fabsf(v[k]) <= kRootTolerance || it >= kBisectionMaxIter) {
                    if (fabsf(v[k]) > kRootTolerance) atomicAdd(&a.counters->c[8], 1ull);
                    const long long i = face[k];
                    a.out[i * 3] = m[k][0]; a.out[i * 3 + 1] = m[k][1]; a.out[i * 3 + 2] = m[k][2];
                    active[k] = false;
                    continue;
                }
                if (v[k] > kRootTolerance) { x2[k][0] = m[k][0]; x2[k][1] = m[k][1]; x2[k][2] = m[k][2]; }
                if (v[k] < -kRootTolerance) { x1[k][0] = m[k][0]; x1[k][1] = m[k][1]; x1[k][2] = m[k][2]; }
            }
            ++it;
            if (!active[0] && !active[1]) entry = -1;
        }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        n_failed += __shfl_xor_sync(0xffffffffu, n_failed, d);
        n_bisect += __shfl_xor_sync(0xffffffffu, n_bisect, d);
    }
    if (lane == 0) {
        if (n_failed) atomicAdd(&a.counters->c[7], (unsigned long long)n_failed);
        if (n_bisect) atomicAdd(&a.counters->c[9], (unsigned long long)n_bisect);
    }
}
// unit gradients at points 2t and 2t+1 (compute_centroid_gradient at the projected centroids, centroids_projection.cpp:1262)
__global__ void __launch_bounds__(128, ISD_PAIR_MINBLOCKS_G) k_unit_gradients_pair(SdfGridView sdf, const float* __restrict__ pts, int n, float* __restrict__ nrm,
                                                                                 const Ob02Counters* __restrict__ skip_if_leftovers,
                                                                                 const unsigned char* __restrict__ state, FaceOrder order) {
    const long long u = (long long)blockIdx.x * 128 + threadIdx.x;
    if (u >= unit_count(order, n)) return;
    if (skip_if_leftovers != nullptr && skip_if_leftovers->c[10] != 0ull) return;
    int ia, ib;
    unit_faces(order, u, n, ia, ib);
    const bool two = ib >= 0;
    if (!two) ib = ia;
    if (state != nullptr && (state[ia] & 4) && (state[ib] & 4)) return;      // both written by the search (state bit 2, k_project_search)
    const FG2 r = eval_fg_pair(sdf, pts[ia * 3], pts[ia * 3 + 1], pts[ia * 3 + 2], pts[ib * 3], pts[ib * 3 + 1], pts[ib * 3 + 2]);
    {
        const float norm = norm2f(r.g0.x, r.g1.x, r.g2.x);
        nrm[ia * 3] = r.g0.x / norm; nrm[ia * 3 + 1] = r.g1.x / norm; nrm[ia * 3 + 2] = r.g2.x / norm;
    }
    if (two) {
        const float norm = norm2f(r.g0.y, r.g1.y, r.g2.y);
        nrm[ib * 3] = r.g0.y / norm; nrm[ib * 3 + 1] = r.g1.y / norm; nrm[ib * 3 + 2] = r.g2.y / norm;
    }
}

// ---- K9 QEM vertex update ---------------------------------------------------------------------------------------------
// Eigen's JacobiSVD of the symmetric PSD 3x3 (qem.hpp:397-405) as a cyclic Jacobi eigen-decomposition, operation for
// operation the same as the oracle's restatement; eigenvalues sorted descending.
__device__ __forceinline__ void sym_eig3(const float (&Ain)[3][3], float (&w)[3], float (&V)[3][3]) {
    float A[3][3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) { A[i][j] = Ain[i][j]; V[i][j] = (i == j) ? 1.0f : 0.0f; }
    // stopping rule of Eigen's JacobiSVD (Eigen/src/SVD/JacobiSVD.h): rotate a pair only while an off-diagonal entry exceeds
    // max(FLT_MIN, 2*epsilon*maxDiagEntry); stop after a sweep that rotated nothing
    const float precision = 2.0f * 1.1920929e-07f, considerAsZero = 1.17549435e-38f;
    float maxDiagEntry = fmaxf(fabsf(A[0][0]), fmaxf(fabsf(A[1][1]), fabsf(A[2][2])));
    for (int sweep = 0; sweep < 12; ++sweep) {
        bool finished = true;
#pragma unroll
        for (int p = 0; p < 2; ++p)
#pragma unroll
            for (int q = p + 1; q < 3; ++q) {
                const float threshold = fmaxf(considerAsZero, precision * maxDiagEntry);
                if (!(fabsf(A[p][q]) > threshold || fabsf(A[q][p]) > threshold)) continue;
                finished = false;
                if (A[p][q] == 0.0f) continue;
                const float theta = (A[q][q] - A[p][p]) / (2.0f * A[p][q]);
                const float t = (theta >= 0 ? 1.0f : -1.0f) / (fabsf(theta) + sqrtf(theta * theta + 1.0f));
                const float c = 1.0f / sqrtf(t * t + 1.0f), s = t * c;
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const float akp = A[k][p], akq = A[k][q];
                    A[k][p] = c * akp - s * akq;
                    A[k][q] = s * akp + c * akq;
                }
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const float apk = A[p][k], aqk = A[q][k];
                    A[p][k] = c * apk - s * aqk;
                    A[q][k] = s * apk + c * aqk;
                }
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const float vkp = V[k][p], vkq = V[k][q];
                    V[k][p] = c * vkp - s * vkq;
                    V[k][q] = s * vkp + c * vkq;
                }
                maxDiagEntry = fmaxf(maxDiagEntry, fmaxf(fabsf(A[p][p]), fabsf(A[q][q])));
            }
        if (finished) break;
    }
#pragma unroll
    for (int i = 0; i < 3; ++i) w[i] = A[i][i];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = i + 1; j < 3; ++j)
            if (w[j] > w[i]) {
                const float t = w[i]; w[i] = w[j]; w[j] = t;
#pragma unroll
                for (int k = 0; k < 3; ++k) { const float u = V[k][i]; V[k][i] = V[k][j]; V[k][j] = u; }
            }
}

// vertex_apply_qem + get_A_b (qem.hpp:256-599)
__global__ void __launch_bounds__(128) k_qem(int nv, const unsigned* __restrict__ offsets, const int* __restrict__ items, const float* __restrict__ cen,
                                             const float* __restrict__ nrm, const AlphaList* __restrict__ al,
                                             const Ob02Counters* __restrict__ skip_if_leftovers, float* __restrict__ verts) {
    const int vi = blockIdx.x * 128 + threadIdx.x;
    if (vi >= nv) return;
    if (skip_if_leftovers != nullptr && skip_if_leftovers->c[10] != 0ull) return;      // speculative launch, see projection_step
    const float maxd = al->avg_edge;
    const float o0 = verts[vi * 3], o1 = verts[vi * 3 + 1], o2 = verts[vi * 3 + 2];
    float A[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}}, b[3] = {0, 0, 0};
    for (unsigned k = offsets[vi], e = offsets[vi + 1]; k < e; ++k) {
        const int ni = items[k];
        const float Nx = nrm[ni * 3], Ny = nrm[ni * 3 + 1], Nz = nrm[ni * 3 + 2];
        const float Px = cen[ni * 3] - o0, Py = cen[ni * 3 + 1] - o1, Pz = cen[ni * 3 + 2] - o2;
        const float nn00 = Nx * Nx, nn01 = Nx * Ny, nn02 = Nx * Nz, nn11 = Ny * Ny, nn12 = Ny * Nz, nn22 = Nz * Nz;
        A[0][0] += nn00; A[0][1] += nn01; A[0][2] += nn02;
        A[1][0] += nn01; A[1][1] += nn11; A[1][2] += nn12;
        A[2][2] += nn22; A[2][0] += nn02; A[2][1] += nn12;
        b[0] -= nn00 * Px + nn01 * Py + nn02 * Pz;
        b[1] -= nn01 * Px + nn11 * Py + nn12 * Pz;
        b[2] -= nn02 * Px + nn12 * Py + nn22 * Pz;
    }
    float Sg[3], V[3][3];
    sym_eig3(A, Sg, V);
#pragma unroll
    for (int i = 0; i < 3; ++i) if (Sg[i] < 0) Sg[i] = 0;
    const float tau = 680.0f;
    const float svd_threshold = 1.0f / tau;
    const float thr = fmaxf(Sg[0] * svd_threshold, 1.17549435e-38f);
    int rank = 0;
    while (rank < 3 && !(Sg[rank] < thr)) rank++;
    float acc[3] = {0, 0, 0};
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        if (i < rank) {
            const float utb = -(V[0][i] * b[0] + V[1][i] * b[1] + V[2][i] * b[2]);
            const float y = utb / Sg[i];
            acc[0] += V[0][i] * y; acc[1] += V[1][i] * y; acc[2] += V[2][i] * y;
        }
    }
    const float n0 = acc[0] + o0, n1 = acc[1] + o1, n2 = acc[2] + o2;
    if (maxd > 0) {
        const float dx = n0 - o0, dy = n1 - o1, dz = n2 - o2;
        const float dist2 = dx * dx + dy * dy + dz * dz;
        if (dist2 <= maxd * maxd) {
            verts[vi * 3] = n0; verts[vi * 3 + 1] = n1; verts[vi * 3 + 2] = n2;
        } else {
            const float dist = sqrtf(dist2);
            float dl = maxd * 1.5f;
            if (dl > dist) dl = dist;
            verts[vi * 3] = o0 + dx / dist * dl;
            verts[vi * 3 + 1] = o1 + dy / dist * dl;
            verts[vi * 3 + 2] = o2 + dz / dist * dl;
        }
    } else {
        verts[vi * 3] = n0; verts[vi * 3 + 1] = n1; verts[vi * 3 + 2] = n2;
    }
}

}  // namespace isd
