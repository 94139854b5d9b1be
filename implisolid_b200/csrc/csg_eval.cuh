// Evaluator for the flattened CSG program (see program.hpp): value mode (f only, P points per thread) and
// gradient mode (f and grad f, one pass).
//
// Arithmetic follows the reference's implicit_function classes operation by operation in binary32, with no FMA
// contraction (this file must be compiled with -fmad=false; on the host with -ffp-contract=off), so field values are
// bit-identical to the reference's eval_implicit / eval_gradient (paths relative to
// /root/reference/js_iteration_2/implicit_function/):
//   affine        basic_functions.hpp:140-177  ((m0*x + m1*y) + m2*z) + m3
//   egg           egg.hpp:93-128               radii 0.5, centre 0 (object_factory.hpp:134-143 only builds this one)
//   scylinder     scylinder.hpp:97-176         w=(0,0,1), base z=-0.5, length 1, radius 0.5
//   scone         scone.hpp:81-166             h=1, r1=0, r2=0.5, apex z0=+0.5
//   cube          cube.hpp:199-210 (six planes, analytic), :291-315 (gradient); HEAD's SDF-grid lookup :215-270
//   metaballs     meta_balls_Rydgard.hpp:89-144,292-432
//   union / intersection / subtract   transformed_union.hpp:47,76  transformed_intersection.hpp:49,78
//                                     transformed_subtract.hpp:52,76-85
// Where the factory's fixed parameters make an operation exact (x-0, x/0.5, x*0+..) it is folded; every folded form
// is bit-equal to the written one for finite inputs (checked against the oracle in tests/test_program_emulation.py).
//
// The same code compiles for the host (ISD_HD) so the flattening and the interpreter can be checked against the
// oracle without a GPU; the host build is a test fixture, never a fallback of the product.
#pragma once
#include <stdint.h>
#include <math.h>
#include "program.hpp"

#if defined(__CUDACC__)
#define ISD_HD __host__ __device__ __forceinline__
#define ISD_COLD __host__ __device__ __noinline__      // rarely used primitives with double-precision bodies: kept out of the
#else                                                   // interpreters' hot loops (registers, instruction cache)
#define ISD_HD inline
#define ISD_COLD inline
#endif

namespace isd {

// Optional (-DISD_LIST_PREFETCH=1): the interpreters read list entry k + 1 while they execute entry k (one register), so that
// the loads of an instruction's matrix depend on a value that is already there.  Measured slower (C4 1024^3: 4.32 against 4.22 ms
// per build, k_centroid_gradients_pair 0.417 against 0.371 ms — profiles/r02_ab_prefetch.txt): the extra live register and the
// second load per iteration cost more than the shorter dependency chain saves at 24-28 warps per SM.  Off.
#ifndef ISD_LIST_PREFETCH
#define ISD_LIST_PREFETCH 0
#endif
#if ISD_LIST_PREFETCH
#define ISD_LIST_FIRST(list, n) uint32_t e_next__ = (n) > 0 ? (list)[0] : 0u
#define ISD_LIST_ENTRY(list, n, k) const uint32_t e = e_next__; if ((k) + 1 < (n)) e_next__ = (list)[(k) + 1]
#else
#define ISD_LIST_FIRST(list, n) do {} while (0)
#define ISD_LIST_ENTRY(list, n, k) const uint32_t e = (list)[k]
#endif
// A list entry carries the opcode of the instruction it points at (bits 24-28) and XFORM's SAVE flag (bit 29), so the list
// interpreters do not fetch the 16-byte instruction header; only METABALLS keeps something there (its ball count).
#define ISD_ENTRY_DECODE(prog, e, pc)                                                         \
    const uint32_t op = ((e) >> LE_OPC_SHIFT) & LE_OPC_MASK;                                  \
    const bool save = ((e) & LE_SAVE) != 0;                                                   \
    W4 hdr = W4{0u, 0u, 0u, 0u};                                                              \
    if (op == OP_METABALLS) hdr = (prog).load4(pc)
struct SdfGridView {          // HEAD-literal cube (cube.hpp:65-72): table lives in global memory
    const float* v;
    int nx, ny, nz;
    float h, ox, oy, oz;
};

ISD_HD float as_float(uint32_t w) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(w);
#else
    float f; memcpy(&f, &w, 4); return f;
#endif
}

// std::min as the reference calls it: (b < a) ? b : a
ISD_HD float ref_min(float a, float b) { return (b < a) ? b : a; }

struct W4 { uint32_t x, y, z, w; };      // one 128-bit fetch of the program (Prog::load4(word index), index % 4 == 0)

template <class Prog>
struct Mat12 {
    float m[12];
    ISD_HD void load(const Prog& prog, int at) {
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const W4 q = prog.load4(at + 4 * r);
            m[4 * r] = as_float(q.x); m[4 * r + 1] = as_float(q.y); m[4 * r + 2] = as_float(q.z); m[4 * r + 3] = as_float(q.w);
        }
    }
    ISD_HD void apply(float x, float y, float z, float& ox, float& oy, float& oz) const {
        ox = m[0] * x + m[1] * y + m[2] * z + m[3];
        oy = m[4] * x + m[5] * y + m[6] * z + m[7];
        oz = m[8] * x + m[9] * y + m[10] * z + m[11];
    }
    // (M^-1)^T_{3x3} g   e.g. egg.hpp:124-126
    ISD_HD void back(float g0, float g1, float g2, float& o0, float& o1, float& o2) const {
        o0 = m[0] * g0 + m[4] * g1 + m[8] * g2;
        o1 = m[1] * g0 + m[5] * g1 + m[9] * g2;
        o2 = m[2] * g0 + m[6] * g1 + m[10] * g2;
    }
};

// ---- primitives, local coordinates (x,y,z) already in the primitive's frame -----------------------------------
ISD_HD float prim_ellipsoid(float x, float y, float z) {
    // 1 - ((x/0.5)^2 + (y/0.5)^2 + (z/0.5)^2); scaling by 2 and 4 is exact
    const float X = x + x, Y = y + y, Z = z + z;
    return 1 - (X * X + Y * Y + Z * Z);
}
ISD_HD void grad_ellipsoid(float x, float y, float z, float& g0, float& g1, float& g2) {
    g0 = -8.0f * x; g1 = -8.0f * y; g2 = -8.0f * z;      // (-2*x)/0.25, exact
}
ISD_HD void cyl_terms(float i0, float i1, float i2, float& t0, float& t1, float& r_) {
    // t0 = (i0-0)*0 + (i1-0)*0 + (i2+0.5)*1 == i2+0.5 ;  i - w*t0 - c  with w=(0,0,1), c=(0,0,-0.5)
    t0 = i2 + 0.5f;
    t1 = 1.0f - t0;
    const float d2 = i2 - t0 + 0.5f;
    r_ = 0.5f - sqrtf(i0 * i0 + i1 * i1 + d2 * d2);
}
ISD_HD float prim_cylinder(float x, float y, float z) {
    float t0, t1, r_;
    cyl_terms(x, y, z, t0, t1, r_);
    return ref_min(t0, ref_min(t1, r_));
}
ISD_HD void grad_cylinder(float i0, float i1, float i2, float& g0, float& g1, float& g2) {
    float t0, t1, r_;
    cyl_terms(i0, i1, i2, t0, t1, r_);
    const float c_t0 = (t0 <= t1 && t0 <= r_) ? 1.0f : 0.0f;
    const float c_t1 = (t1 <= t0 && t1 <= r_) ? 1.0f : 0.0f;
    const float c_r = (r_ <= t0 && r_ <= t1) ? 1.0f : 0.0f;
    const float w0 = 0.0f, w1 = 0.0f, w2 = 1.0f, cx = 0.0f, cy = 0.0f, cz = -0.5f;
    g0 = c_t0 * w0 + c_t1 * (-w0) + c_r * (w0 * t0 + cx - i0);
    g1 = c_t0 * w1 + c_t1 * (-w1) + c_r * (w1 * t0 + cy - i1);
    g2 = c_t0 * w2 + c_t1 * (-w2) + c_r * (w2 * t0 + cz - i2);
}
ISD_HD float prim_cone(float x, float y, float z) {
    const float dz = z - 0.5f;
    const float f = -sqrtf(x * x + y * y) + sqrtf(dz * dz * 0.25f);
    const float uperside = -dz - 0.0f;
    const float lowerside = dz + 1.0f;
    return ref_min(f, ref_min(uperside, lowerside));
}
ISD_HD void grad_cone(float x, float y, float z, float& gx, float& gy, float& gz) {
    const float dz = z - 0.5f;
    // -(x*x)/a2 - (y*y)/a2 + dz*dz with a2 = 0.25: the divisions are exact scalings by 4
    const float f = -(x * x) * 4.0f - (y * y) * 4.0f + dz * dz;
    const float uperside = -dz - 0.0f;
    const float lowerside = dz + 1.0f;
    if (uperside < f && uperside < lowerside) { gx = 0.f; gy = 0.f; gz = -1.f; }
    else if (lowerside < f && lowerside < uperside) { gx = 0.f; gy = 0.f; gz = +1.f; }
    else { gx = -8.0f * x; gy = -8.0f * y; gz = 2.0f * dz; }
}
ISD_HD float prim_cube(float x, float y, float z) {
    // the six plane terms ((X-C-p).p)(-2) are exactly 0.5 -/+ coordinate
    const float t0 = 0.5f - x, t1 = x + 0.5f, t2 = 0.5f - y, t3 = y + 0.5f, t4 = 0.5f - z, t5 = z + 0.5f;
    return ref_min(t0, ref_min(t1, ref_min(t2, ref_min(t3, ref_min(t4, t5)))));
}
ISD_HD void grad_cube(float x, float y, float z, float& g0, float& g1, float& g2) {
    const float t[6] = {0.5f - x, x + 0.5f, 0.5f - y, y + 0.5f, 0.5f - z, z + 0.5f};
    int index = 0;
    float mn = t[0];
#pragma unroll
    for (int k = 1; k < 6; ++k)
        if (t[k] < mn) { index = k; mn = t[k]; }
    // -p[index], including the signed zeros of the other two components
    g0 = (index == 0) ? -0.5f : (index == 1) ? 0.5f : -0.0f;
    g1 = (index == 2) ? -0.5f : (index == 3) ? 0.5f : -0.0f;
    g2 = (index == 4) ? -0.5f : (index == 5) ? 0.5f : -0.0f;
}
ISD_HD float prim_cube_sdf(const SdfGridView& g, float px, float py, float pz) {
    float res = 10000;
    if (g.ox + g.h * g.nx < px || px < g.ox) {
    } else if (g.oy + g.h * g.ny < py || py < g.oy) {
    } else if (g.oz + g.h * g.nz < pz || pz < g.oz) {
    } else {
        const int xc = (int)((px - g.ox) / g.h), yc = (int)((py - g.oy) / g.h), zc = (int)((pz - g.oz) / g.h);
        const float xl = g.ox + xc * g.h, yl = g.oy + yc * g.h, zl = g.oz + zc * g.h;
        const float xd = (px - xl) / g.h, yd = (py - yl) / g.h, zd = (pz - zl) / g.h;
        const long total = (long)g.nx * g.ny * g.nz;
        // the reference indexes one past the table at the upper faces (UB); out-of-table reads give 0 here
#define ISD_AT(a, b, c) (((long)(a) + (long)(b) * g.nx + (long)(c) * g.nx * g.ny) < total ? g.v[(long)(a) + (long)(b) * g.nx + (long)(c) * g.nx * g.ny] : 0.0f)
        const float c00 = ISD_AT(xc, yc, zc) * (1 - xd) + ISD_AT(xc + 1, yc, zc) * xd;
        const float c01 = ISD_AT(xc, yc, zc + 1) * (1 - xd) + ISD_AT(xc + 1, yc, zc + 1) * xd;
        const float c10 = ISD_AT(xc, yc + 1, zc) * (1 - xd) + ISD_AT(xc + 1, yc + 1, zc) * xd;
        const float c11 = ISD_AT(xc, yc + 1, zc + 1) * (1 - xd) + ISD_AT(xc + 1, yc + 1, zc + 1) * xd;
#undef ISD_AT
        const float c0 = c00 * (1 - yd) + c10 * yd;
        const float c1 = c01 * (1 - yd) + c11 * yd;
        res = c0 * (1 - zd) + c1 * (zd);
    }
    return -res;
}

// ---- the rest of the factory's analytic vocabulary (SURVEY §8 f-N1) ------------------------------------------------------
// These follow C++'s promotion rules as the reference's expressions are written: std::pow(float, int|double) is a double
// power, a double literal makes its sub-expression double, and the store rounds once to binary32.  Squares and cubes are
// formed by multiplication and pow(.,0.5) by sqrt (IEEE-exact here; glibc's pow is within 0.52 ulp of the same double, so
// after the final rounding to binary32 the two agree except with probability ~2^-29 per point: tolerance-level parity).
//   torus            torus.hpp:55-118            r = 4, rx = ry = rz = 0.2 (the matrix12 constructor, :42-53)
//   heart            heart.hpp:78-126
//   double_mushroom  double_mushroom.hpp:68-134  (0.9, 0.2, 0.2, 1/0.2): r = 0.45, a = b = 0.2, c = 1/5; its own identity
//                    matrix product is carried out as the reference does (signs of zero)
//   top_bottom_lid   top_bottom_lid.hpp:51-76    on untransformed points
//   tetrahedron      tetrahedron.hpp:145-193     four host-built planes, untransformed points
//   half_plane       half_plane.hpp:109-150
ISD_COLD float prim_torus(float x, float y, float z) {
    const float rx = 0.2f, r = 4.0f;
    const double qx = (double)(x / rx), qy = (double)(y / rx), qz = (double)(z / rx);
    const double a = sqrt(qx * qx + qy * qy);
    const double d = (double)r - a;
    return (float)(1 - d * d - qz * qz);
}
ISD_COLD void grad_torus(float x, float y, float z, float& g0, float& g1, float& g2) {
    const float rx = 0.2f, r = 4.0f;
    const float a = (float)sqrt((double)(x * x / (rx * rx) + y * y / (rx * rx)));
    g0 = (2 * x / (rx * rx * a)) * (r - a);
    g1 = (2 * y / (rx * rx * a)) * (r - a);
    g2 = -2 * z / (rx * rx);
}
ISD_HD double heart_core(float i1, float i2, float i3) {
    return (double)(i1 * i1) + (9. / 4.) * (double)i2 * (double)i2 + (double)(i3 * i3);
}
ISD_COLD float prim_heart(float i1, float i2, float i3) {
    const double u = heart_core(i1, i2, i3) - 1.;
    const double d2 = (double)i2, d3 = (double)i3;
    return (float)(-(u * u * u - (double)(i1 * i1 * i3 * i3 * i3) - (9. / 200.) * d2 * d2 * d3 * d3 * d3));
}
ISD_COLD void grad_heart(float i1, float i2, float i3, float& g0, float& g1, float& g2) {
    const double u = heart_core(i1, i2, i3) - 1;
    const float a = (float)(u * u);
    const double d1 = (double)i1, d2 = (double)i2, d3 = (double)i3, da = (double)a;
    g0 = (float)(-6. * d1 * da + 2. * d1 * d3 * d3 * d3);
    g1 = (float)(-(27. / 2) * d2 * da + (9. / 100.) * d2 * d3 * d3 * d3);
    g2 = (float)(-6. * d3 * da + 3. * d1 * d1 * d3 * d3 + (27. / 200.) * d2 * d2 * d3 * d3);
}
ISD_HD void identity_product(float& x, float& y, float& z) {      // matrix_vector_product with the identity, as written
    const float one = 1.0f, zero = 0.0f;
    const float nx = one * x + zero * y + zero * z + zero;
    const float ny = zero * x + one * y + zero * z + zero;
    const float nz = zero * x + zero * y + one * z + zero;
    x = nx; y = ny; z = nz;
}
ISD_COLD float prim_mushroom(float x, float y, float z) {
    identity_product(x, y, z);
    const float r = 0.9f / 2, a = (float)(0.4 / 2), c = 1 / (float)(1 / 0.2);
    const float a2 = a * a, c2 = c * c;
    if (z > r) return r - z;
    if (z < -r) return r + z;
    const double dx = (double)(x - 0.0f), dy = (double)(y - 0.0f), dz = (double)(z - 0.0f);
    return (float)(-(dx * dx / (double)a2 + dy * dy / (double)a2 - dz * dz / (double)c2 - 1));
}
ISD_COLD void grad_mushroom(float x, float y, float z, float& o0, float& o1, float& o2) {
    identity_product(x, y, z);
    const float r = 0.9f / 2, a = (float)(0.4 / 2), c = 1 / (float)(1 / 0.2);
    const float a2 = a * a, c2 = c * c;
    float g0, g1, g2;
    if (z < -r) { g2 = 1; g1 = 0; g0 = 0; }
    else if (z > r) { g2 = -1; g1 = 0; g0 = 0; }
    else {
        g0 = -2 * (x - 0.0f) / a2;
        g1 = -2 * (y - 0.0f) / a2;
        g2 = 2 * (z - 0.0f) / c2;
    }
    // double_mushroom's own (identity) inverse, :128-130; the wrapper's follows in the caller
    const float one = 1.0f, zero = 0.0f;
    o0 = one * g0 + zero * g1 + zero * g2;
    o1 = zero * g0 + one * g1 + zero * g2;
    o2 = zero * g0 + zero * g1 + one * g2;
}
ISD_HD float prim_lid(float z) {
    const float p = z - 0.5f, q = (z + 0.5f) * -1;
    return (p < q) ? q : p;
}
template <class Prog>
ISD_HD float tetra_plane(const Prog& prog, int at, float x, float y, float z) {
    const W4 q = prog.load4(at);
    return as_float(q.x) * x + as_float(q.y) * y + as_float(q.z) * z + as_float(q.w);
}
template <class Prog>
ISD_HD float prim_tetra(const Prog& prog, int at, float x, float y, float z) {
    return ref_min(tetra_plane(prog, at, x, y, z),
                   ref_min(tetra_plane(prog, at + 4, x, y, z), ref_min(tetra_plane(prog, at + 8, x, y, z), tetra_plane(prog, at + 12, x, y, z))));
}
template <class Prog>
ISD_HD void grad_tetra(const Prog& prog, int at, float x, float y, float z, float& g0, float& g1, float& g2) {
    int index = 0;
    float mn = tetra_plane(prog, at, x, y, z);
    for (int k = 1; k < 4; ++k) {
        const float possible_min = tetra_plane(prog, at + 4 * k, x, y, z);
        if (possible_min < mn) { index = k; mn = possible_min; }
    }
    const W4 q = prog.load4(at + 4 * index);
    g0 = as_float(q.x); g1 = as_float(q.y); g2 = as_float(q.z);
}
template <class Prog>
ISD_HD float prim_half_plane(const Prog& prog, int at, float x, float y, float z) {
    const W4 n = prog.load4(at), p = prog.load4(at + 4);
    return (x - as_float(p.x)) * as_float(n.x) + (y - as_float(p.y)) * as_float(n.y) + (z - as_float(p.z)) * as_float(n.z);
}

// One primitive other than METABALLS at one point: (cx,cy,cz) are the coordinates handed down by the parent, M the
// instruction's matrix, `at` the word index of its parameters.
template <class Prog>
ISD_HD float prim_value(uint32_t op, const Prog& prog, int at, const Mat12<Prog>& M, const SdfGridView& sdf, float cx, float cy, float cz) {
    if (op == OP_TETRA) return prim_tetra(prog, at, cx, cy, cz);
    if (op == OP_LID) return prim_lid(cz);
    float x, y, z;
    M.apply(cx, cy, cz, x, y, z);
    if (op == OP_ELLIPSOID) return prim_ellipsoid(x, y, z);
    if (op == OP_CYLINDER) return prim_cylinder(x, y, z);
    if (op == OP_CONE) return prim_cone(x, y, z);
    if (op == OP_CUBE) return prim_cube(x, y, z);
    if (op == OP_CUBE_SDF) return prim_cube_sdf(sdf, x, y, z);
    if (op == OP_TORUS) return prim_torus(x, y, z);
    if (op == OP_HEART) return prim_heart(x, y, z);
    if (op == OP_MUSHROOM) return prim_mushroom(x, y, z);
    return prim_half_plane(prog, at, x, y, z);
}

// ---- packed pairs: two points per thread on the sm_100 f32x2 pipe ------------------------------------------------------------------
// Blackwell issues FMUL2 / FADD2 / FFMA2 on register pairs: two binary32 operations per instruction, each rounded exactly like
// its scalar form, with a scalar register broadcast to both lanes where an operand is `splat(s)`.  The interpreters below
// run TWO points through one instruction stream with them; every lane performs the reference's operations in the reference's
// order, so the results keep every bit of the scalar interpreters (held on the host by tests/emu, where F2 is two plain floats,
// and on the GPU by the parity tests).
// ptxas (12.9) contracts mul.rn.f32x2 followed by add.rn.f32x2 into one FFMA2 even with explicit .rn and -fmad=false, which would
// round once instead of twice.  The product is therefore written fma(a, b, nz) with nz = -0.0 read from constant memory — a value
// the optimiser cannot see — which is fl(a*b) exactly (x + (-0) = x for every x, signed zeros included) and cannot be merged with
// the addition that follows.  Sums and differences are add.rn.f32x2 / sub.rn.f32x2.
struct alignas(8) F2 { float x, y; };
#if defined(__CUDACC__)
__constant__ float c_negzero = -0.0f;
#endif
ISD_HD F2 splat(float s) { return F2{s, s}; }
ISD_HD F2 neg2(F2 a) { return F2{-a.x, -a.y}; }
#if defined(__CUDA_ARCH__)
__device__ __forceinline__ unsigned long long f2_bits(F2 a) { return *reinterpret_cast<unsigned long long*>(&a); }
__device__ __forceinline__ F2 f2_from(unsigned long long v) { return *reinterpret_cast<F2*>(&v); }
#endif
ISD_HD F2 mul2(F2 a, F2 b) {
#if defined(__CUDA_ARCH__)
    unsigned long long r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(f2_bits(a)), "l"(f2_bits(b)), "l"(f2_bits(F2{c_negzero, c_negzero})));
    return f2_from(r);
#else
    return F2{a.x * b.x, a.y * b.y};
#endif
}
ISD_HD F2 add2(F2 a, F2 b) {
#if defined(__CUDA_ARCH__)
    unsigned long long r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(f2_bits(a)), "l"(f2_bits(b)));
    return f2_from(r);
#else
    return F2{a.x + b.x, a.y + b.y};
#endif
}
ISD_HD F2 sub2(F2 a, F2 b) {
#if defined(__CUDA_ARCH__)
    unsigned long long r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(f2_bits(a)), "l"(f2_bits(b)));
    return f2_from(r);
#else
    return F2{a.x - b.x, a.y - b.y};
#endif
}
ISD_HD F2 sqrt2(F2 a) { return F2{sqrtf(a.x), sqrtf(a.y)}; }
ISD_HD F2 min2(F2 a, F2 b) { return F2{ref_min(a.x, b.x), ref_min(a.y, b.y)}; }

// Mat12::apply / Mat12::back on a pair, operation for operation
template <class Prog>
ISD_HD void apply2(const Mat12<Prog>& M, F2 x, F2 y, F2 z, F2& ox, F2& oy, F2& oz) {
    const float* m = M.m;
    ox = add2(add2(add2(mul2(splat(m[0]), x), mul2(splat(m[1]), y)), mul2(splat(m[2]), z)), splat(m[3]));
    oy = add2(add2(add2(mul2(splat(m[4]), x), mul2(splat(m[5]), y)), mul2(splat(m[6]), z)), splat(m[7]));
    oz = add2(add2(add2(mul2(splat(m[8]), x), mul2(splat(m[9]), y)), mul2(splat(m[10]), z)), splat(m[11]));
}
template <class Prog>
ISD_HD void back2(const Mat12<Prog>& M, F2 g0, F2 g1, F2 g2, F2& o0, F2& o1, F2& o2) {
    const float* m = M.m;
    o0 = add2(add2(mul2(splat(m[0]), g0), mul2(splat(m[4]), g1)), mul2(splat(m[8]), g2));
    o1 = add2(add2(mul2(splat(m[1]), g0), mul2(splat(m[5]), g1)), mul2(splat(m[9]), g2));
    o2 = add2(add2(mul2(splat(m[2]), g0), mul2(splat(m[6]), g1)), mul2(splat(m[10]), g2));
}
// the factory's four everyday primitives on a pair (same expressions as prim_ellipsoid / cyl_terms / prim_cone / prim_cube)
ISD_HD F2 prim_ellipsoid2(F2 x, F2 y, F2 z) {
    const F2 X = add2(x, x), Y = add2(y, y), Z = add2(z, z);
    return sub2(splat(1.0f), add2(add2(mul2(X, X), mul2(Y, Y)), mul2(Z, Z)));
}
ISD_HD void cyl_terms2(F2 i0, F2 i1, F2 i2, F2& t0, F2& t1, F2& r_) {
    t0 = add2(i2, splat(0.5f));
    t1 = sub2(splat(1.0f), t0);
    const F2 d2 = add2(sub2(i2, t0), splat(0.5f));
    r_ = sub2(splat(0.5f), sqrt2(add2(add2(mul2(i0, i0), mul2(i1, i1)), mul2(d2, d2))));
}
ISD_HD F2 prim_cylinder2(F2 x, F2 y, F2 z) {
    F2 t0, t1, r_;
    cyl_terms2(x, y, z, t0, t1, r_);
    return min2(t0, min2(t1, r_));
}
ISD_HD F2 prim_cone2(F2 x, F2 y, F2 z) {
    const F2 dz = sub2(z, splat(0.5f));
    const F2 f = add2(neg2(sqrt2(add2(mul2(x, x), mul2(y, y)))), sqrt2(mul2(mul2(dz, dz), splat(0.25f))));
    const F2 uperside = sub2(neg2(dz), splat(0.0f));
    const F2 lowerside = add2(dz, splat(1.0f));
    return min2(f, min2(uperside, lowerside));
}
ISD_HD F2 prim_cube2(F2 x, F2 y, F2 z) {
    const F2 h = splat(0.5f);
    const F2 t0 = sub2(h, x), t1 = add2(x, h), t2 = sub2(h, y), t3 = add2(y, h), t4 = sub2(h, z), t5 = add2(z, h);
    return min2(t0, min2(t1, min2(t2, min2(t3, min2(t4, t5)))));
}
ISD_HD bool packed_primitive(uint32_t op) { return op == OP_ELLIPSOID || op == OP_CYLINDER || op == OP_CONE || op == OP_CUBE; }
// one primitive (not METABALLS) on a pair: packed where it pays, lane by lane otherwise
template <class Prog>
ISD_HD F2 prim_value2(uint32_t op, const Prog& prog, int at, const Mat12<Prog>& M, const SdfGridView& sdf, F2 cx, F2 cy, F2 cz) {
    if (packed_primitive(op)) {
        F2 x, y, z;
        apply2(M, cx, cy, cz, x, y, z);
        if (op == OP_ELLIPSOID) return prim_ellipsoid2(x, y, z);
        if (op == OP_CYLINDER) return prim_cylinder2(x, y, z);
        if (op == OP_CONE) return prim_cone2(x, y, z);
        return prim_cube2(x, y, z);
    }
    return F2{prim_value(op, prog, at, M, sdf, cx.x, cy.x, cz.x), prim_value(op, prog, at, M, sdf, cx.y, cy.y, cz.y)};
}
ISD_HD float csg_select(uint32_t op_minus_union, float f1, float f2) {
    if (op_minus_union == 0u) return (f1 > f2) ? f1 : f2;
    if (op_minus_union == 1u) return (f1 > f2) ? f2 : f1;
    return (f1 < -f2) ? f1 : -f2;
}

// eval_values_list for NP pairs of points (2 * NP points per thread)
template <int NP, class Prog, class List>
ISD_HD void eval_values_list2(const Prog& prog, const SdfGridView& sdf, const List& list, int n_list, F2 (&cx)[NP], F2 (&cy)[NP], F2 (&cz)[NP],
                              F2 (&out)[NP]) {
    F2 spill[kMaxValueStack][NP];
    F2 fs[kMaxFrameStack][3][NP];
    F2 t0[NP], t1[NP];
#pragma unroll
    for (int p = 0; p < NP; ++p) { t0[p] = splat(0.0f); t1[p] = splat(0.0f); }
    int sp = 0, fp = 0;
    ISD_LIST_FIRST(list, n_list);
    for (int k = 0; k < n_list; ++k) {
        ISD_LIST_ENTRY(list, n_list, k);
        int pc = (int)(e & IT_PC_MASK) * 4;
        if (e & LE_CLOSING) {
            if (e & LE_COMBINE) {
                const bool swapped = (e & LE_SWAPPED) != 0;
                const uint32_t op = (e >> LE_OP_SHIFT) & 3u;
#pragma unroll
                for (int p = 0; p < NP; ++p) {
                    const F2 top = t0[p], below = t1[p];
                    const F2 f1 = swapped ? top : below, f2 = swapped ? below : top;
                    t0[p] = F2{csg_select(op, f1.x, f2.x), csg_select(op, f1.y, f2.y)};
                }
                --sp;
                if (sp >= 2) {
#pragma unroll
                    for (int p = 0; p < NP; ++p) t1[p] = spill[sp - 2][p];
                }
            } else if (e & LE_NEGATE) {
#pragma unroll
                for (int p = 0; p < NP; ++p) t0[p] = neg2(t0[p]);
            }
            if (e & LE_RESTORE) {
                --fp;
#pragma unroll
                for (int p = 0; p < NP; ++p) { cx[p] = fs[fp][0][p]; cy[p] = fs[fp][1][p]; cz[p] = fs[fp][2][p]; }
            }
            continue;
        }
        ISD_ENTRY_DECODE(prog, e, pc);
        Mat12<Prog> M;
        M.load(prog, pc + 4);
        pc += kInstrWords;
        if (op == OP_XFORM) {
            if (save) {
#pragma unroll
                for (int p = 0; p < NP; ++p) { fs[fp][0][p] = cx[p]; fs[fp][1][p] = cy[p]; fs[fp][2][p] = cz[p]; }
                ++fp;
            }
#pragma unroll
            for (int p = 0; p < NP; ++p) {
                F2 x, y, z;
                apply2(M, cx[p], cy[p], cz[p], x, y, z);
                cx[p] = x; cy[p] = y; cz[p] = z;
            }
            continue;
        }
        F2 f[NP];
        if (op == OP_METABALLS) {
            const int nb = (int)hdr.z;
            F2 lx[NP], ly[NP], lz[NP];
#pragma unroll
            for (int p = 0; p < NP; ++p) { apply2(M, cx[p], cy[p], cz[p], lx[p], ly[p], lz[p]); f[p] = splat(0.0f); }
            for (int b = 0; b < nb; ++b) {
                const W4 q = prog.load4(pc), q2 = prog.load4(pc + 4);
                const float bx = as_float(q.x), by = as_float(q.y), bz = as_float(q.z);
                const float strength = as_float(q.w), subtract = as_float(q2.x);
                pc += 8;
#pragma unroll
                for (int p = 0; p < NP; ++p) {
                    const F2 fx = sub2(lx[p], splat(bx)), fx2 = mul2(fx, fx);
                    const F2 fy = sub2(ly[p], splat(by)), fy2 = mul2(fy, fy);
                    const F2 fz = sub2(lz[p], splat(bz)), fz2 = mul2(fz, fz);
                    const F2 h = add2(add2(add2(splat(0.000001f), fx2), fy2), fz2);
                    const F2 hinv = F2{1.0f / h.x, 1.0f / h.y};
                    const F2 val = sub2(mul2(splat(strength), hinv), splat(subtract));
                    f[p] = add2(f[p], F2{val.x / 100.0f, val.y / 100.0f});
                }
            }
        } else {
#pragma unroll
            for (int p = 0; p < NP; ++p) f[p] = prim_value2(op, prog, pc, M, sdf, cx[p], cy[p], cz[p]);
        }
        if (sp >= 2) {
#pragma unroll
            for (int p = 0; p < NP; ++p) spill[sp - 2][p] = t1[p];
        }
#pragma unroll
        for (int p = 0; p < NP; ++p) { t1[p] = t0[p]; t0[p] = f[p]; }
        ++sp;
    }
#pragma unroll
    for (int p = 0; p < NP; ++p) out[p] = t0[p];
}

// ---- flop accounting (bench.py's FP32 roofline entries) ---------------------------------------------------------------
// Floating-point operations one instruction executes per point, multiplications and additions counted separately (exact mode
// has no FMA), comparisons / selects / sqrt / division one each.  Value-mode numbers are the ones ShapeCompiler adds up into
// Program::flops_per_point (program.hpp); gradient mode adds the primitive's gradient body and the 3x3 transposed product
// (M.back, 15) on the way up.  A closing CSG instruction: 1 select (+1 negation for a subtraction) in value mode, + M.back in
// gradient mode.
ISD_HD uint32_t op_flops(uint32_t op, uint32_t aux, bool gradient) {
    uint32_t v = 18, g = 15;
    switch (op) {
        case OP_XFORM: return 18;
        case OP_ELLIPSOID: v += 9; g += 3; break;
        case OP_CYLINDER: v += 14; g += 15; break;
        case OP_CONE: v += 14; g += 10; break;
        case OP_CUBE: v += 11; g += 5; break;
        case OP_CUBE_SDF: v += 40; g += 5; break;
        case OP_METABALLS: v += 12 * aux; g += 21 * aux; break;
        case OP_TORUS: v += 13; g += 14; break;
        case OP_HEART: v += 24; g += 30; break;
        case OP_MUSHROOM: v += 18 + 11; g += 18 + 9; break;
        case OP_LID: v = 3; g = 0; break;
        case OP_TETRA: v = 24 + 3; g = 0; break;
        case OP_HALF_PLANE: v += 8; g += 4; break;
        default: return 0;
    }
    return gradient ? v + g : v;
}
ISD_HD uint32_t closing_flops(uint32_t op_minus_union, bool combine, bool negate, bool gradient) {
    uint32_t v = combine ? (op_minus_union == 2u ? 2u : 1u) : (negate ? 1u : 0u);
    if (gradient) v = (combine ? v + 3u : (negate ? 4u : 0u)) + 15u;
    return v;
}
// the whole program at one point (what eval_values / eval_value_gradient run with AllLive)
template <class Prog>
ISD_HD uint32_t program_flops(const Prog& prog, bool gradient) {
    uint32_t total = 0;
    for (int pc = 0;;) {
        const W4 hdr = prog.load4(pc);
        const uint32_t op = hdr.x & OP_MASK;
        if (op == OP_END) break;
        if (op >= OP_UNION) { total += closing_flops(op - OP_UNION, true, false, gradient); pc += kInstrWords; }
        else if (op == OP_XFORM) { total += 18; pc += kInstrWords; }
        else { total += op_flops(op, hdr.z, gradient); pc = (int)hdr.y; }
    }
    return total;
}
// a per-tile instruction list (entries from spec_entry / spec_entry_gradient)
template <class Prog, class List>
ISD_HD uint32_t list_flops(const Prog& prog, const List& list, int n_list, bool gradient) {
    uint32_t total = 0;
    for (int k = 0; k < n_list; ++k) {
        const uint32_t e = list[k];
        if (e & LE_CLOSING) { total += closing_flops((e >> LE_OP_SHIFT) & 3u, (e & LE_COMBINE) != 0, (e & LE_NEGATE) != 0, gradient); continue; }
        const W4 hdr = prog.load4((int)(e & IT_PC_MASK) * 4);
        total += op_flops(hdr.x & OP_MASK, hdr.z, gradient);
    }
    return total;
}

// ---- value mode -------------------------------------------------------------------------------------------------
// Evaluates the program at P points held in registers; cur* are clobbered.
// The two newest values of the stack live in registers (t0 = top, t1 = below); older ones spill to a per-thread
// array (local memory, L1-resident).
//
// `live(id)` says whether node `id` can influence the result at these points (tile specialisation, field_kernel.cuh:
// a child of a min/max node that interval bounds prove is never selected over the whole tile is dead).  A dead node is
// jumped over and pushes nothing; `present` is a bit stack recording, for every node finished so far, whether it left a
// value, so that a closing CSG instruction knows if it has two operands (select as the reference does), or one (pass it
// through; a lone second operand of a subtraction is negated: -f2 is what the reference selects there).  A live CSG
// node always has a live child.  Selecting between the same two floats as the reference, or passing through the one it
// would select, gives the same bits; AllLive reproduces the reference's full evaluation.
struct AllLive {
    ISD_HD bool operator()(uint32_t) const { return true; }
};
template <class Live> struct IsAllLive { static constexpr bool value = false; };
template <> struct IsAllLive<AllLive> { static constexpr bool value = true; };

template <int P, class Prog, class Live>
ISD_HD void eval_values(const Prog& prog, const SdfGridView& sdf, const Live& live, float (&cx)[P], float (&cy)[P], float (&cz)[P],
                        float (&out)[P]) {
    float spill[kMaxValueStack][P];
    float fs[kMaxFrameStack][3][P];
    float t0[P], t1[P];
#pragma unroll
    for (int p = 0; p < P; ++p) { t0[p] = 0.0f; t1[p] = 0.0f; }
    int sp = 0, fp = 0, pc = 0;
    unsigned long long present = 0ull;
    for (;;) {
        const W4 hdr = prog.load4(pc);
        const uint32_t word = hdr.x;
        const uint32_t op = word & OP_MASK;
        if (op == OP_END) break;
        if (op >= OP_UNION) {
            const bool top_there = (present & 1ull) != 0, below_there = (present & 2ull) != 0;
            present = ((present >> 2) << 1) | 1ull;
            pc += kInstrWords;
            const bool swapped = (word & OP_FLAG_SWAPPED) != 0;
            if (top_there && below_there) {
#pragma unroll
                for (int p = 0; p < P; ++p) {
                    const float top = t0[p], below = t1[p];
                    const float f1 = swapped ? top : below, f2 = swapped ? below : top;
                    float r;
                    if (op == OP_UNION) r = (f1 > f2) ? f1 : f2;
                    else if (op == OP_INTERSECT) r = (f1 > f2) ? f2 : f1;
                    else r = (f1 < -f2) ? f1 : -f2;
                    t0[p] = r;
                }
                --sp;
                if (sp >= 2) {
#pragma unroll
                    for (int p = 0; p < P; ++p) t1[p] = spill[sp - 2][p];
                }
            } else if (op == OP_SUBTRACT) {
                // the lone operand sits on top; it is f2 when (top is the second-evaluated child) xor swapped
                const bool lone_is_f2 = top_there != swapped;
                if (lone_is_f2) {
#pragma unroll
                    for (int p = 0; p < P; ++p) t0[p] = -t0[p];
                }
            }
            if (word & OP_FLAG_RESTORE) {
                --fp;
#pragma unroll
                for (int p = 0; p < P; ++p) { cx[p] = fs[fp][0][p]; cy[p] = fs[fp][1][p]; cz[p] = fs[fp][2][p]; }
            }
            continue;
        }
        // node-opening instruction: XFORM or a primitive
        if (!live(word >> OP_ID_SHIFT)) {
            pc = (int)hdr.y;
            present <<= 1;
            continue;
        }
        Mat12<Prog> M;
        M.load(prog, pc + 4);
        pc += kInstrWords;
        if (op == OP_XFORM) {
            if (word & OP_FLAG_SAVE) {
#pragma unroll
                for (int p = 0; p < P; ++p) { fs[fp][0][p] = cx[p]; fs[fp][1][p] = cy[p]; fs[fp][2][p] = cz[p]; }
                ++fp;
            }
#pragma unroll
            for (int p = 0; p < P; ++p) {
                float x, y, z;
                M.apply(cx[p], cy[p], cz[p], x, y, z);
                cx[p] = x; cy[p] = y; cz[p] = z;
            }
        } else {
            float f[P];
            if (op == OP_METABALLS) {
                const int nb = (int)hdr.z;
                float lx[P], ly[P], lz[P];
#pragma unroll
                for (int p = 0; p < P; ++p) { M.apply(cx[p], cy[p], cz[p], lx[p], ly[p], lz[p]); f[p] = 0.0f; }
                for (int b = 0; b < nb; ++b) {
                    const W4 q = prog.load4(pc), q2 = prog.load4(pc + 4);
                    const float bx = as_float(q.x), by = as_float(q.y), bz = as_float(q.z);
                    const float strength = as_float(q.w), subtract = as_float(q2.x);
                    pc += 8;
#pragma unroll
                    for (int p = 0; p < P; ++p) {
                        const float fx = lx[p] - bx, fx2 = fx * fx;
                        const float fy = ly[p] - by, fy2 = fy * fy;
                        const float fz = lz[p] - bz, fz2 = fz * fz;
                        const float h = 0.000001f + fx2 + fy2 + fz2;
                        const float hinv = 1.0f / h;
                        const float val = strength * hinv - subtract;
                        f[p] += val / 100.0f;
                    }
                }
            } else {
#pragma unroll
                for (int p = 0; p < P; ++p) f[p] = prim_value(op, prog, pc, M, sdf, cx[p], cy[p], cz[p]);
                pc = (int)hdr.y;      // past the parameters, if any
            }
            // push
            if (sp >= 2) {
#pragma unroll
                for (int p = 0; p < P; ++p) spill[sp - 2][p] = t1[p];
            }
#pragma unroll
            for (int p = 0; p < P; ++p) { t1[p] = t0[p]; t0[p] = f[p]; }
            ++sp;
            present = (present << 1) | 1ull;
        }
    }
#pragma unroll
    for (int p = 0; p < P; ++p) out[p] = t0[p];
}

// ---- value mode over a per-tile instruction list (tile specialisation) ---------------------------------------------
// spec_entry decides what one program instruction does in a tile whose live nodes are given by `live`:
//   node-opening instruction : runs iff its node is live;
//   closing instruction      : runs iff its CSG node is live and it has something to do — select between two live children
//                              (COMBINE), negate the lone second child of a subtraction (NEGATE: -f2 is what the reference
//                              selects there), or pop the coordinate frame (RESTORE).
// Returns 0 when the instruction is skipped.  The list in program order is then run by eval_values_list, which does
// exactly what eval_values does with the same `live`, minus all the bookkeeping.
template <class Live>
ISD_HD uint32_t spec_entry(const W4& it, const Live& live) {
    if (!live(it.y)) return 0u;
    if (!(it.x & IT_CLOSING)) return (it.x & (IT_PC_MASK | (LE_OPC_MASK << LE_OPC_SHIFT) | LE_SAVE)) | LE_VALID;
    const bool la = live(it.z), lb = live(it.w);
    const uint32_t op = (it.x >> IT_OP_SHIFT) & 3u;
    uint32_t e = (it.x & (IT_PC_MASK | IT_CLOSING | IT_RESTORE | IT_SWAPPED)) | (op << LE_OP_SHIFT);      // same bit positions as LE_*
    if (la && lb) e |= LE_COMBINE;
    else if (op == (uint32_t)OP_SUBTRACT - (uint32_t)OP_UNION && lb) e |= LE_NEGATE;
    if (!(e & (LE_COMBINE | LE_NEGATE | LE_RESTORE))) return 0u;
    return e | LE_VALID;
}

// The same for gradient mode, where a live CSG node with a single live child still has work to do (the child's gradient goes
// through the node's own matrix): such closing instructions stay in the list, marked LE_PASS.
template <class Live>
ISD_HD uint32_t spec_entry_gradient(const W4& it, const Live& live) {
    if (!live(it.y)) return 0u;
    if (!(it.x & IT_CLOSING)) return (it.x & (IT_PC_MASK | (LE_OPC_MASK << LE_OPC_SHIFT) | LE_SAVE)) | LE_VALID;
    const bool la = live(it.z), lb = live(it.w);
    const uint32_t op = (it.x >> IT_OP_SHIFT) & 3u;
    uint32_t e = (it.x & (IT_PC_MASK | IT_CLOSING | IT_RESTORE | IT_SWAPPED)) | (op << LE_OP_SHIFT);
    if (la && lb) e |= LE_COMBINE;
    else if (op == (uint32_t)OP_SUBTRACT - (uint32_t)OP_UNION && lb) e |= LE_NEGATE;
    if (!(e & (LE_COMBINE | LE_NEGATE | LE_RESTORE))) e |= LE_PASS;
    return e | LE_VALID;
}

template <int P, class Prog, class List>
ISD_HD void eval_values_list(const Prog& prog, const SdfGridView& sdf, const List& list, int n_list, float (&cx)[P], float (&cy)[P],
                             float (&cz)[P], float (&out)[P]) {
    float spill[kMaxValueStack][P];
    float fs[kMaxFrameStack][3][P];
    float t0[P], t1[P];
#pragma unroll
    for (int p = 0; p < P; ++p) { t0[p] = 0.0f; t1[p] = 0.0f; }
    int sp = 0, fp = 0;
    ISD_LIST_FIRST(list, n_list);
    for (int k = 0; k < n_list; ++k) {
        ISD_LIST_ENTRY(list, n_list, k);
        int pc = (int)(e & IT_PC_MASK) * 4;
        if (e & LE_CLOSING) {
            if (e & LE_COMBINE) {
                const bool swapped = (e & LE_SWAPPED) != 0;
                const uint32_t op = (e >> LE_OP_SHIFT) & 3u;
#pragma unroll
                for (int p = 0; p < P; ++p) {
                    const float top = t0[p], below = t1[p];
                    const float f1 = swapped ? top : below, f2 = swapped ? below : top;
                    float r;
                    if (op == 0u) r = (f1 > f2) ? f1 : f2;
                    else if (op == 1u) r = (f1 > f2) ? f2 : f1;
                    else r = (f1 < -f2) ? f1 : -f2;
                    t0[p] = r;
                }
                --sp;
                if (sp >= 2) {
#pragma unroll
                    for (int p = 0; p < P; ++p) t1[p] = spill[sp - 2][p];
                }
            } else if (e & LE_NEGATE) {
#pragma unroll
                for (int p = 0; p < P; ++p) t0[p] = -t0[p];
            }
            if (e & LE_RESTORE) {
                --fp;
#pragma unroll
                for (int p = 0; p < P; ++p) { cx[p] = fs[fp][0][p]; cy[p] = fs[fp][1][p]; cz[p] = fs[fp][2][p]; }
            }
            continue;
        }
        ISD_ENTRY_DECODE(prog, e, pc);
        Mat12<Prog> M;
        M.load(prog, pc + 4);
        pc += kInstrWords;
        if (op == OP_XFORM) {
            if (save) {
#pragma unroll
                for (int p = 0; p < P; ++p) { fs[fp][0][p] = cx[p]; fs[fp][1][p] = cy[p]; fs[fp][2][p] = cz[p]; }
                ++fp;
            }
#pragma unroll
            for (int p = 0; p < P; ++p) {
                float x, y, z;
                M.apply(cx[p], cy[p], cz[p], x, y, z);
                cx[p] = x; cy[p] = y; cz[p] = z;
            }
            continue;
        }
        float f[P];
        if (op == OP_METABALLS) {
            const int nb = (int)hdr.z;
            float lx[P], ly[P], lz[P];
#pragma unroll
            for (int p = 0; p < P; ++p) { M.apply(cx[p], cy[p], cz[p], lx[p], ly[p], lz[p]); f[p] = 0.0f; }
            for (int b = 0; b < nb; ++b) {
                const W4 q = prog.load4(pc), q2 = prog.load4(pc + 4);
                const float bx = as_float(q.x), by = as_float(q.y), bz = as_float(q.z);
                const float strength = as_float(q.w), subtract = as_float(q2.x);
                pc += 8;
#pragma unroll
                for (int p = 0; p < P; ++p) {
                    const float fx = lx[p] - bx, fx2 = fx * fx;
                    const float fy = ly[p] - by, fy2 = fy * fy;
                    const float fz = lz[p] - bz, fz2 = fz * fz;
                    const float h = 0.000001f + fx2 + fy2 + fz2;
                    const float hinv = 1.0f / h;
                    const float val = strength * hinv - subtract;
                    f[p] += val / 100.0f;
                }
            }
        } else {
#pragma unroll
            for (int p = 0; p < P; ++p) f[p] = prim_value(op, prog, pc, M, sdf, cx[p], cy[p], cz[p]);
        }
        if (sp >= 2) {
#pragma unroll
            for (int p = 0; p < P; ++p) spill[sp - 2][p] = t1[p];
        }
#pragma unroll
        for (int p = 0; p < P; ++p) { t1[p] = t0[p]; t0[p] = f[p]; }
        ++sp;
    }
#pragma unroll
    for (int p = 0; p < P; ++p) out[p] = t0[p];
}

// One live primitive in gradient mode: f and the gradient in the parent's frame (shared by the two gradient interpreters'
// list form; `pc` is the word index of the primitive's parameters, (cx,cy,cz) the coordinates handed down by the parent).
template <class Prog>
ISD_HD void prim_value_gradient(uint32_t op, const W4& hdr, const Prog& prog, int pc, const Mat12<Prog>& M, const SdfGridView& sdf, float cx,
                                float cy, float cz, float& f, float& o0, float& o1, float& o2) {
    float x, y, z;
    M.apply(cx, cy, cz, x, y, z);
    float g0, g1, g2;
    if (op == OP_METABALLS) {
        const int nb = (int)hdr.z;
        f = 0.0f; g0 = 0.0f; g1 = 0.0f; g2 = 0.0f;
        for (int b = 0; b < nb; ++b) {
            const W4 q = prog.load4(pc), q2 = prog.load4(pc + 4);
            const float bx = as_float(q.x), by = as_float(q.y), bz = as_float(q.z);
            const float strength = as_float(q.w), subtract = as_float(q2.x);
            pc += 8;
            const float fx = x - bx, fx2 = fx * fx;
            const float fy = y - by, fy2 = fy * fy;
            const float fz = z - bz, fz2 = fz * fz;
            {   // implicit_ball :89-100
                const float h = 0.000001f + fx2 + fy2 + fz2;
                const float hinv = 1.0f / h;
                const float val = strength * hinv - subtract;
                f += val / 100.0f;
            }
            {   // gradient_ball :118-136
                const float h = (float)(0.000001 + (double)fx2 + (double)fy2 + (double)fz2);
                const float hinv = 1.0f / h;
                g0 += strength * (-2.0f * fx * hinv * hinv) / 100.0f;
                g1 += strength * (-2.0f * fy * hinv * hinv) / 100.0f;
                g2 += strength * (-2.0f * fz * hinv * hinv) / 100.0f;
            }
        }
    } else if (op == OP_ELLIPSOID) { f = prim_ellipsoid(x, y, z); grad_ellipsoid(x, y, z, g0, g1, g2); }
    else if (op == OP_CYLINDER) { f = prim_cylinder(x, y, z); grad_cylinder(x, y, z, g0, g1, g2); }
    else if (op == OP_CONE) { f = prim_cone(x, y, z); grad_cone(x, y, z, g0, g1, g2); }
    else if (op == OP_CUBE) { f = prim_cube(x, y, z); grad_cube(x, y, z, g0, g1, g2); }
    else if (op == OP_CUBE_SDF) { f = prim_cube_sdf(sdf, x, y, z); grad_cube(x, y, z, g0, g1, g2); }
    else if (op == OP_TORUS) { f = prim_torus(x, y, z); grad_torus(x, y, z, g0, g1, g2); }
    else if (op == OP_HEART) { f = prim_heart(x, y, z); grad_heart(x, y, z, g0, g1, g2); }
    else if (op == OP_MUSHROOM) { f = prim_mushroom(x, y, z); grad_mushroom(x, y, z, g0, g1, g2); }
    else if (op == OP_HALF_PLANE) {
        f = prim_half_plane(prog, pc, x, y, z);
        const W4 n = prog.load4(pc);
        const float s = (f >= 0) ? -1.0f : 1.0f;      // half_plane.hpp:129-138, as written
        g0 = s * as_float(n.x); g1 = s * as_float(n.y); g2 = s * as_float(n.z);
    } else if (op == OP_TETRA) {      // planes in the parent's frame; no matrix on the way back (tetrahedron.hpp:163-193)
        f = prim_tetra(prog, pc, cx, cy, cz);
        grad_tetra(prog, pc, cx, cy, cz, g0, g1, g2);
    } else {                          // OP_LID: (0, 0, 1) everywhere (top_bottom_lid.hpp:62-76)
        f = prim_lid(cz);
        g0 = 0; g1 = 0; g2 = 1;
    }
    if (op == OP_TETRA || op == OP_LID) { o0 = g0; o1 = g1; o2 = g2; }
    else M.back(g0, g1, g2, o0, o1, o2);
}

// ---- gradient mode: f and grad f in one pass (identical to the reference's O(depth^2) recursion) ----
// `live` as in eval_values: a dead node is jumped over; a CSG node with one operand passes it on (value and gradient; the
// lone second operand of a subtraction negated, transformed_subtract.hpp:76-85) through its own matrix, which is exactly
// what the reference does with the child it selects.
template <class Prog, class Live>
ISD_HD void eval_value_gradient(const Prog& prog, const SdfGridView& sdf, const Live& live, float cx, float cy, float cz, float& f_out,
                                float& g0_out, float& g1_out, float& g2_out) {
    float vs[kMaxValueStack][4];      // all but the newest entry; the newest lives in registers (t*), valid when sp > 0
    float fs[kMaxFrameStack][3];
    float tf = 0.0f, t0 = 0.0f, t1 = 0.0f, t2 = 0.0f;
    int sp = 0, fp = 0, pc = 0;
    unsigned long long present = 0ull;
    for (;;) {
        const W4 hdr = prog.load4(pc);
        const uint32_t word = hdr.x;
        const uint32_t op = word & OP_MASK;
        if (op == OP_END) break;
        if (op < OP_UNION && !live(word >> OP_ID_SHIFT)) {      // dead node-opening instruction: skip the node's whole code
            pc = (int)hdr.y;
            present <<= 1;
            continue;
        }
        Mat12<Prog> M;
        M.load(prog, pc + 4);
        pc += kInstrWords;
        if (op == OP_XFORM) {
            if (word & OP_FLAG_SAVE) { fs[fp][0] = cx; fs[fp][1] = cy; fs[fp][2] = cz; ++fp; }
            float x, y, z;
            M.apply(cx, cy, cz, x, y, z);
            cx = x; cy = y; cz = z;
        } else if (op >= OP_UNION) {
            const bool top_there = IsAllLive<Live>::value || (present & 1ull) != 0, below_there = IsAllLive<Live>::value || (present & 2ull) != 0;
            present = ((present >> 2) << 1) | 1ull;
            const bool swapped = (word & OP_FLAG_SWAPPED) != 0;
            float r, a0, a1, a2;
            if (top_there && below_there) {
                --sp;
                const float bf = vs[sp - 1][0], b0 = vs[sp - 1][1], b1 = vs[sp - 1][2], b2 = vs[sp - 1][3];      // the entry below the top
                const float f1 = swapped ? tf : bf, f2 = swapped ? bf : tf;
                bool first;
                if (op == OP_UNION) { first = (f1 > f2); r = first ? f1 : f2; }
                else if (op == OP_INTERSECT) { first = !(f1 > f2); r = (f1 > f2) ? f2 : f1; }
                else { first = (f1 < -f2); r = first ? f1 : -f2; }
                const bool take_top = (first == swapped);      // operand 1 is the top entry when swapped
                a0 = take_top ? t0 : b0; a1 = take_top ? t1 : b1; a2 = take_top ? t2 : b2;
                if (!first && op == OP_SUBTRACT) { a0 = -a0; a1 = -a1; a2 = -a2; }
            } else {
                const bool negate = (op == OP_SUBTRACT) && (top_there != swapped);      // the lone operand is f2
                r = tf; a0 = t0; a1 = t1; a2 = t2;
                if (negate) { r = -r; a0 = -a0; a1 = -a1; a2 = -a2; }
            }
            tf = r;
            M.back(a0, a1, a2, t0, t1, t2);
            if (word & OP_FLAG_RESTORE) { --fp; cx = fs[fp][0]; cy = fs[fp][1]; cz = fs[fp][2]; }
        } else {
            float f, o0, o1, o2;
            prim_value_gradient(op, hdr, prog, pc, M, sdf, cx, cy, cz, f, o0, o1, o2);
            pc = (int)hdr.y;
            if (sp > 0) { vs[sp - 1][0] = tf; vs[sp - 1][1] = t0; vs[sp - 1][2] = t1; vs[sp - 1][3] = t2; }
            tf = f; t0 = o0; t1 = o1; t2 = o2;
            ++sp;
            present = (present << 1) | 1ull;
        }
    }
    f_out = tf; g0_out = t0; g1_out = t1; g2_out = t2;
}
template <class Prog>
ISD_HD void eval_value_gradient(const Prog& prog, const SdfGridView& sdf, float cx, float cy, float cz, float& f_out, float& g0_out,
                                float& g1_out, float& g2_out) {
    eval_value_gradient(prog, sdf, AllLive(), cx, cy, cz, f_out, g0_out, g1_out, g2_out);
}

// Gradient mode over a per-tile instruction list (entries from spec_entry_gradient, program order): what eval_value_gradient
// does with the same liveness, without the dead-node jumps and the operand bookkeeping.
template <class Prog, class List>
ISD_HD void eval_value_gradient_list(const Prog& prog, const SdfGridView& sdf, const List& list, int n_list, float cx, float cy, float cz,
                                     float& f_out, float& g0_out, float& g1_out, float& g2_out) {
    float vs[kMaxValueStack][4];      // all but the newest entry; the newest lives in registers (t*), valid when sp > 0
    float fs[kMaxFrameStack][3];
    float tf = 0.0f, t0 = 0.0f, t1 = 0.0f, t2 = 0.0f;
    int sp = 0, fp = 0;
    ISD_LIST_FIRST(list, n_list);
    for (int k = 0; k < n_list; ++k) {
        ISD_LIST_ENTRY(list, n_list, k);
        const int pc = (int)(e & IT_PC_MASK) * 4;
        Mat12<Prog> M;
        M.load(prog, pc + 4);
        if (e & LE_CLOSING) {
            float r = tf, a0 = t0, a1 = t1, a2 = t2;      // one operand: it passes through the node's matrix (LE_PASS, LE_NEGATE)
            if (e & LE_COMBINE) {
                const bool swapped = (e & LE_SWAPPED) != 0;
                const uint32_t op = (e >> LE_OP_SHIFT) & 3u;
                --sp;
                const float bf = vs[sp - 1][0], b0 = vs[sp - 1][1], b1 = vs[sp - 1][2], b2 = vs[sp - 1][3];      // the entry below the top
                const float f1 = swapped ? tf : bf, f2 = swapped ? bf : tf;
                bool first;
                if (op == 0u) { first = (f1 > f2); r = first ? f1 : f2; }
                else if (op == 1u) { first = !(f1 > f2); r = (f1 > f2) ? f2 : f1; }
                else { first = (f1 < -f2); r = first ? f1 : -f2; }
                const bool take_top = (first == swapped);      // operand 1 is the top entry when swapped
                a0 = take_top ? t0 : b0; a1 = take_top ? t1 : b1; a2 = take_top ? t2 : b2;
                if (!first && op == 2u) { a0 = -a0; a1 = -a1; a2 = -a2; }
            } else if (e & LE_NEGATE) {
                r = -r; a0 = -a0; a1 = -a1; a2 = -a2;
            }
            tf = r;
            M.back(a0, a1, a2, t0, t1, t2);
            if (e & LE_RESTORE) { --fp; cx = fs[fp][0]; cy = fs[fp][1]; cz = fs[fp][2]; }
            continue;
        }
        ISD_ENTRY_DECODE(prog, e, pc);
        if (op == OP_XFORM) {
            if (save) { fs[fp][0] = cx; fs[fp][1] = cy; fs[fp][2] = cz; ++fp; }
            float x, y, z;
            M.apply(cx, cy, cz, x, y, z);
            cx = x; cy = y; cz = z;
            continue;
        }
        float f, o0, o1, o2;
        prim_value_gradient(op, hdr, prog, pc + kInstrWords, M, sdf, cx, cy, cz, f, o0, o1, o2);
        if (sp > 0) { vs[sp - 1][0] = tf; vs[sp - 1][1] = t0; vs[sp - 1][2] = t1; vs[sp - 1][3] = t2; }
        tf = f; t0 = o0; t1 = o1; t2 = o2;
        ++sp;
    }
    f_out = tf; g0_out = t0; g1_out = t1; g2_out = t2;
}

// ---- gradient mode on a pair of points (same list for both) ------------------------------------------------------------------------------
// grad_cylinder from the terms cyl_terms2 has already produced for this lane (grad_cylinder recomputes them: same operations, same bits)
ISD_HD void grad_cylinder_from_terms(float t0, float t1, float r_, float i0, float i1, float i2, float& g0, float& g1, float& g2) {
    const float c_t0 = (t0 <= t1 && t0 <= r_) ? 1.0f : 0.0f;
    const float c_t1 = (t1 <= t0 && t1 <= r_) ? 1.0f : 0.0f;
    const float c_r = (r_ <= t0 && r_ <= t1) ? 1.0f : 0.0f;
    const float w0 = 0.0f, w1 = 0.0f, w2 = 1.0f, cx = 0.0f, cy = 0.0f, cz = -0.5f;
    g0 = c_t0 * w0 + c_t1 * (-w0) + c_r * (w0 * t0 + cx - i0);
    g1 = c_t0 * w1 + c_t1 * (-w1) + c_r * (w1 * t0 + cy - i1);
    g2 = c_t0 * w2 + c_t1 * (-w2) + c_r * (w2 * t0 + cz - i2);
}
template <class Prog>
ISD_HD void prim_value_gradient2(uint32_t op, const W4& hdr, const Prog& prog, int pc, const Mat12<Prog>& M, const SdfGridView& sdf, F2 cx, F2 cy,
                                 F2 cz, F2& f, F2& o0, F2& o1, F2& o2) {
    if (!packed_primitive(op)) {      // lane by lane through the scalar form (transform and back-transform included)
        prim_value_gradient(op, hdr, prog, pc, M, sdf, cx.x, cy.x, cz.x, f.x, o0.x, o1.x, o2.x);
        prim_value_gradient(op, hdr, prog, pc, M, sdf, cx.y, cy.y, cz.y, f.y, o0.y, o1.y, o2.y);
        return;
    }
    F2 x, y, z, g0, g1, g2;
    apply2(M, cx, cy, cz, x, y, z);
    if (op == OP_ELLIPSOID) {
        f = prim_ellipsoid2(x, y, z);
        g0 = mul2(splat(-8.0f), x); g1 = mul2(splat(-8.0f), y); g2 = mul2(splat(-8.0f), z);
    } else if (op == OP_CYLINDER) {
        F2 t0, t1, r_;
        cyl_terms2(x, y, z, t0, t1, r_);
        f = min2(t0, min2(t1, r_));
        grad_cylinder_from_terms(t0.x, t1.x, r_.x, x.x, y.x, z.x, g0.x, g1.x, g2.x);
        grad_cylinder_from_terms(t0.y, t1.y, r_.y, x.y, y.y, z.y, g0.y, g1.y, g2.y);
    } else if (op == OP_CONE) {
        f = prim_cone2(x, y, z);
        grad_cone(x.x, y.x, z.x, g0.x, g1.x, g2.x);
        grad_cone(x.y, y.y, z.y, g0.y, g1.y, g2.y);
    } else {
        f = prim_cube2(x, y, z);
        grad_cube(x.x, y.x, z.x, g0.x, g1.x, g2.x);
        grad_cube(x.y, y.y, z.y, g0.y, g1.y, g2.y);
    }
    back2(M, g0, g1, g2, o0, o1, o2);
}
// one lane of a closing CSG instruction with both operands: the selected value and which operand's gradient goes up
ISD_HD void csg_select_gradient(uint32_t op, bool swapped, float tf, float bf, float t0, float t1, float t2, float b0, float b1, float b2, float& r,
                                float& a0, float& a1, float& a2) {
    const float f1 = swapped ? tf : bf, f2 = swapped ? bf : tf;
    bool first;
    if (op == 0u) { first = (f1 > f2); r = first ? f1 : f2; }
    else if (op == 1u) { first = !(f1 > f2); r = (f1 > f2) ? f2 : f1; }
    else { first = (f1 < -f2); r = first ? f1 : -f2; }
    const bool take_top = (first == swapped);
    a0 = take_top ? t0 : b0; a1 = take_top ? t1 : b1; a2 = take_top ? t2 : b2;
    if (!first && op == 2u) { a0 = -a0; a1 = -a1; a2 = -a2; }
}
template <class Prog, class List>
ISD_HD void eval_value_gradient_list2(const Prog& prog, const SdfGridView& sdf, const List& list, int n_list, F2 cx, F2 cy, F2 cz, F2& f_out,
                                      F2& g0_out, F2& g1_out, F2& g2_out) {
    F2 vs[kMaxValueStack][4];
    F2 fs[kMaxFrameStack][3];
    F2 tf = splat(0.0f), t0 = splat(0.0f), t1 = splat(0.0f), t2 = splat(0.0f);
    int sp = 0, fp = 0;
    ISD_LIST_FIRST(list, n_list);
    for (int k = 0; k < n_list; ++k) {
        ISD_LIST_ENTRY(list, n_list, k);
        const int pc = (int)(e & IT_PC_MASK) * 4;
        Mat12<Prog> M;
        M.load(prog, pc + 4);
        if (e & LE_CLOSING) {
            F2 r = tf, a0 = t0, a1 = t1, a2 = t2;
            if (e & LE_COMBINE) {
                const bool swapped = (e & LE_SWAPPED) != 0;
                const uint32_t op = (e >> LE_OP_SHIFT) & 3u;
                --sp;
                const F2 bf = vs[sp - 1][0], b0 = vs[sp - 1][1], b1 = vs[sp - 1][2], b2 = vs[sp - 1][3];
                csg_select_gradient(op, swapped, tf.x, bf.x, t0.x, t1.x, t2.x, b0.x, b1.x, b2.x, r.x, a0.x, a1.x, a2.x);
                csg_select_gradient(op, swapped, tf.y, bf.y, t0.y, t1.y, t2.y, b0.y, b1.y, b2.y, r.y, a0.y, a1.y, a2.y);
            } else if (e & LE_NEGATE) {
                r = neg2(r); a0 = neg2(a0); a1 = neg2(a1); a2 = neg2(a2);
            }
            tf = r;
            back2(M, a0, a1, a2, t0, t1, t2);
            if (e & LE_RESTORE) { --fp; cx = fs[fp][0]; cy = fs[fp][1]; cz = fs[fp][2]; }
            continue;
        }
        ISD_ENTRY_DECODE(prog, e, pc);
        if (op == OP_XFORM) {
            if (save) { fs[fp][0] = cx; fs[fp][1] = cy; fs[fp][2] = cz; ++fp; }
            F2 x, y, z;
            apply2(M, cx, cy, cz, x, y, z);
            cx = x; cy = y; cz = z;
            continue;
        }
        F2 f, o0, o1, o2;
        prim_value_gradient2(op, hdr, prog, pc + kInstrWords, M, sdf, cx, cy, cz, f, o0, o1, o2);
        if (sp > 0) { vs[sp - 1][0] = tf; vs[sp - 1][1] = t0; vs[sp - 1][2] = t1; vs[sp - 1][3] = t2; }
        tf = f; t0 = o0; t1 = o1; t2 = o2;
        ++sp;
    }
    f_out = tf; g0_out = t0; g1_out = t1; g2_out = t2;
}

}  // namespace isd
