// K1 — field evaluation over the sampling grid.
//
// Replaces MarchingCubes::prepare_grid + eval_shape + seal_exterior (marching_cubes.hpp:1662-1738, :895-963):
// the reference materialises a 12 B/node coordinate array and runs the CSG tree over it with two N-sized temporaries
// per tree node; here the coordinate of a node is computed from its index in registers, the whole flattened program
// (csg_eval.cuh) runs per thread, and the two outermost node layers get the seal value without being evaluated.
//
// Layout in HBM: field[(z - z0) * plane + y * pitch + x], pitch = S rounded up to 4 floats so that every thread's four
// x-adjacent nodes are one aligned 16-byte store (st.global.v4.f32); plane = pitch * S.
// Work per node: 4 B written, 0 read; flops = Program::flops_per_point (no FMA contraction in exact mode).
#pragma once
#include <cuda_runtime.h>
#include "csg_eval.cuh"

namespace isd {

__constant__ uint32_t c_program[kMaxProgramWords];

struct ConstProgram {
    __device__ __forceinline__ uint32_t operator[](int i) const { return c_program[i]; }
};
struct GlobalProgram {     // per-tile shortened programs live in global memory (read through L1, warp-uniform)
    const uint32_t* w;
    __device__ __forceinline__ uint32_t operator[](int i) const { return __ldg(w + i); }
};

struct GridDesc {
    int S;             // nodes per axis = R + 5
    int pitch;         // floats per x-row in HBM
    int z0, nz;        // node planes [z0, z0 + nz) are resident on this GPU
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

constexpr int kFieldThreads = 128;

__global__ void __launch_bounds__(kFieldThreads) k_eval_field(GridDesc g, SdfGridView sdf, float* __restrict__ field) {
    const int quads_per_row = g.pitch >> 2;
    const long long q = (long long)blockIdx.x * kFieldThreads + threadIdx.x;
    const long long rows = (long long)g.S * g.nz;
    const long long row = q / quads_per_row;
    if (row >= rows) return;
    const int xq = (int)(q - row * quads_per_row);
    const int zl = (int)(row / g.S);
    const int y = (int)(row - (long long)zl * g.S);
    const int z = g.z0 + zl;
    const int x0 = xq << 2;
    float4* dst = reinterpret_cast<float4*>(field + (long long)zl * g.plane + (long long)y * g.pitch + x0);
    const bool row_sealed = sealed_index(y, g.S) || sealed_index(z, g.S);
    if (row_sealed || x0 + 3 <= 1 || x0 >= g.S - 2) {
        *dst = make_float4(kExteriorValue, kExteriorValue, kExteriorValue, kExteriorValue);
        return;
    }
    float cx[4], cy[4], cz[4], f[4];
    const float yy = node_coord(y, g.wy, g.ymin, g.two_wy);
    const float zz = node_coord(z, g.wz, g.zmin, g.two_wz);
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        cx[p] = node_coord(x0 + p, g.wx, g.xmin, g.two_wx);
        cy[p] = yy;
        cz[p] = zz;
    }
    eval_values<4>(ConstProgram(), sdf, cx, cy, cz, f);
#pragma unroll
    for (int p = 0; p < 4; ++p)
        if (sealed_index(x0 + p, g.S)) f[p] = kExteriorValue;
    *dst = make_float4(f[0], f[1], f[2], f[3]);
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
