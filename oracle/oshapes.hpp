// ORACLE — TEST INFRASTRUCTURE ONLY (see ojson.hpp header).
//
// CPU restatement of the reference's implicit-function tree, statement by statement,
// in binary32 with the reference's evaluation order (compile with -ffp-contract=off).
// Paths are relative to /root/reference/js_iteration_2/.
//   invert_matrix            basic_functions.hpp:77-128   (uBLAS lu_factorize/lu_substitute restated;
//                            Boost is not in this image, so this elimination order DEFINES the oracle)
//   matrix_vector_product    basic_functions.hpp:140-177
//   prepare_inner_vectors    basic_functions.hpp:361-379
//   egg                      implicit_function/egg.hpp:93-128
//   scylinder                implicit_function/scylinder.hpp:97-176
//   cube                     implicit_function/cube.hpp:199-210 (analytic), :215-270 (HEAD SDF grid), :291-315 (gradient)
//   scone                    implicit_function/scone.hpp:81-166
//   meta_ball_Rydgard        implicit_function/meta_balls_Rydgard.hpp:49-59,89-144,292-432
//   transformed_union/intersection/subtract  implicit_function/transformed_*.hpp
//   object_factory           object_factory.hpp:21-30,56-173,537-673,741-758
#pragma once
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <string>
#include <thread>
#include <vector>
#include "ojson.hpp"

namespace oracle {

typedef float REAL;

struct vects {            // N x 3, AoS (boost::multi_array<REAL,2> in the reference)
    std::vector<REAL> d;
    vects() {}
    explicit vects(size_t n) : d(n * 3, 0.0f) {}
    size_t n() const { return d.size() / 3; }
    REAL* operator[](size_t i) { return &d[i * 3]; }
    const REAL* operator[](size_t i) const { return &d[i * 3]; }
};
typedef std::vector<REAL> scalars;

// ---- basic_functions.hpp:77-128 -------------------------------------------------------------
// Returns false (and leaves inv untouched) when a zero pivot is met, like lu_factorize != 0.
inline bool invert_matrix(const REAL in12[12], REAL inv12[12]) {
    REAL A[4][4], B[4][4];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 4; j++) A[i][j] = in12[i * 4 + j];
    A[3][0] = 0.f; A[3][1] = 0.f; A[3][2] = 0.f; A[3][3] = 1.f;
    int pm[4] = {0, 1, 2, 3};
    int singular = 0;
    // lu_factorize: partial pivoting, reciprocal scaling, rank-1 trailing update
    for (int i = 0; i < 4; i++) {
        int p = i;
        REAL best = std::fabs(A[i][i]);
        for (int k = i + 1; k < 4; k++) {             // index_norm_inf: first maximum
            REAL a = std::fabs(A[k][i]);
            if (a > best) { best = a; p = k; }
        }
        if (A[p][i] != 0.0f) {
            if (p != i) {
                pm[i] = p;
                for (int j = 0; j < 4; j++) std::swap(A[i][j], A[p][j]);
            }
            REAL r = 1.0f / A[i][i];
            for (int k = i + 1; k < 4; k++) A[k][i] *= r;
        } else if (singular == 0) {
            singular = i + 1;
        }
        for (int k = i + 1; k < 4; k++)
            for (int j = i + 1; j < 4; j++) A[k][j] -= A[k][i] * A[i][j];
    }
    if (singular != 0) return false;
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) B[i][j] = (i == j) ? 1.0f : 0.0f;
    // lu_substitute: swap_rows(pm, B)
    for (int i = 0; i < 4; i++)
        if (pm[i] != i)
            for (int j = 0; j < 4; j++) std::swap(B[i][j], B[pm[i]][j]);
    // inplace_solve(unit_lower): column-oriented forward substitution
    for (int n = 0; n < 4; n++)
        for (int l = 0; l < 4; l++) {
            REAL t = B[n][l];
            if (t != 0.0f)
                for (int m = n + 1; m < 4; m++) B[m][l] -= A[m][n] * t;
        }
    // inplace_solve(upper): column-oriented back substitution
    for (int n = 3; n >= 0; n--)
        for (int l = 3; l >= 0; l--) {
            REAL t = B[n][l] /= A[n][n];
            if (t != 0.0f)
                for (int m = n - 1; m >= 0; m--) B[m][l] -= A[m][n] * t;
        }
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 4; j++) inv12[i * 4 + j] = B[i][j];
    return true;
}

// ---- basic_functions.hpp:140-177 ------------------------------------------------------------
inline void matrix_vector_product(const REAL m[12], vects& v) {
    for (size_t i = 0, n = v.n(); i < n; ++i) {
        const REAL x = v[i][0], y = v[i][1], z = v[i][2];
        v[i][0] = m[0] * x + m[1] * y + m[2] * z + m[3];
        v[i][1] = m[4] * x + m[5] * y + m[6] * z + m[7];
        v[i][2] = m[8] * x + m[9] * y + m[10] * z + m[11];
    }
}
inline vects prepare_inner_vectors(const REAL inv[12], const vects& x) {
    vects c = x;
    matrix_vector_product(inv, c);
    return c;
}

// how "cube" evaluates f: HEAD looks up a hard-coded SDF grid (cube.hpp:215-270), the
// analytic six-plane form is what the class name, its gradient and its bounding box mean
// (cube.hpp:199-210, commented out at HEAD).  SURVEY F3.
enum CubeMode { CUBE_ANALYTIC = 0, CUBE_HEAD_SDFGRID = 1 };

struct SdfGrid {   // supplied by the test harness from a fixture (never compiled in)
    int nx = 0, ny = 0, nz = 0;
    float h = 0, ox = 0, oy = 0, oz = 0;
    std::vector<float> v;
};

struct Env {
    CubeMode cube_mode = CUBE_ANALYTIC;
    const SdfGrid* sdf = nullptr;
};

struct implicit_function {
    REAL M[12], Minv[12];
    virtual ~implicit_function() {}
    virtual void eval_implicit(const vects& x, scalars* f) const = 0;
    virtual void eval_gradient(const vects& x, vects* g) const = 0;
    void set_matrix(const REAL m[12]) {
        for (int i = 0; i < 12; i++) { M[i] = m[i]; Minv[i] = 0.0f; }   // reference: uninitialised heap if singular
        invert_matrix(M, Minv);
    }
    // (M^-1)^T_{3x3} g, e.g. egg.hpp:124-126
    void back(REAL g0, REAL g1, REAL g2, REAL* out) const {
        out[0] = Minv[0] * g0 + Minv[4] * g1 + Minv[8] * g2;
        out[1] = Minv[1] * g0 + Minv[5] * g1 + Minv[9] * g2;
        out[2] = Minv[2] * g0 + Minv[6] * g1 + Minv[10] * g2;
    }
};

inline REAL squared(REAL x) { return x * x; }
inline REAL norm_squared(REAL x, REAL y, REAL z) { return x * x + y * y + z * z; }

// ---- egg.hpp ----------------------------------------------------------------------------------
struct egg : implicit_function {
    REAL a = 0.5f, b = 0.5f, c = 0.5f, x0 = 0.f, y0 = 0.f, z0 = 0.f;
    explicit egg(const REAL m[12]) { set_matrix(m); }
    void eval_implicit(const vects& x, scalars* f) const override {
        vects lx = x;
        matrix_vector_product(Minv, lx);
        for (size_t i = 0; i < lx.n(); i++)
            (*f)[i] = 1 - norm_squared((lx[i][0] - x0) / a, (lx[i][1] - y0) / b, (lx[i][2] - z0) / c);
    }
    void eval_gradient(const vects& x, vects* out) const override {
        vects lx = x;
        matrix_vector_product(Minv, lx);
        const REAL a2 = squared(a), b2 = squared(b), c2 = squared(c);
        for (size_t i = 0; i < lx.n(); i++) {
            REAL g0 = -2. * (lx[i][0] - x0) / a2;      // double intermediate, single rounding (Appendix E)
            REAL g1 = -2. * (lx[i][1] - y0) / b2;
            REAL g2 = -2. * (lx[i][2] - z0) / c2;
            back(g0, g1, g2, (*out)[i]);
        }
    }
};

// ---- scylinder.hpp ----------------------------------------------------------------------------
struct scylinder : implicit_function {
    REAL x = 0.f, y = 0.f, z = -0.5f, w[3] = {0.f, 0.f, 1.f}, c_len = 1.0f, radius_u = 0.5f;
    explicit scylinder(const REAL m[12]) { set_matrix(m); }
    void eval_implicit(const vects& xx, scalars* f) const override {
        vects lx = xx;
        matrix_vector_product(Minv, lx);
        for (size_t i = 0; i < lx.n(); i++) {
            REAL i0 = lx[i][0], i1 = lx[i][1], i2 = lx[i][2];
            REAL w0 = w[0], w1 = w[1], w2 = w[2];
            REAL t0 = (i0 - x) * w0 + (i1 - y) * w1 + (i2 - z) * w2;
            REAL t1 = c_len - t0;
            REAL r_ = radius_u - std::sqrt((i0 - w0 * t0 - x) * (i0 - w0 * t0 - x) +
                                           (i1 - w1 * t0 - y) * (i1 - w1 * t0 - y) + (i2 - w2 * t0 - z) * (i2 - w2 * t0 - z));
            (*f)[i] = std::min(t0, std::min(t1, r_));
        }
    }
    void eval_gradient(const vects& xx, vects* out) const override {
        vects lx = xx;
        matrix_vector_product(Minv, lx);
        for (size_t i = 0; i < lx.n(); i++) {
            REAL i0 = lx[i][0], i1 = lx[i][1], i2 = lx[i][2];
            REAL w0 = w[0], w1 = w[1], w2 = w[2];
            REAL t0 = (i0 - x) * w0 + (i1 - y) * w1 + (i2 - z) * w2;
            REAL t1 = c_len - t0;
            REAL r_ = radius_u - std::sqrt((i0 - w0 * t0 - x) * (i0 - w0 * t0 - x) +
                                           (i1 - w1 * t0 - y) * (i1 - w1 * t0 - y) + (i2 - w2 * t0 - z) * (i2 - w2 * t0 - z));
            bool c_t0 = 0, c_t1 = 0, c_r = 0;
            if (t0 <= t1 && t0 <= r_) c_t0 = 1;
            if (t1 <= t0 && t1 <= r_) c_t1 = 1;
            if (r_ <= t0 && r_ <= t1) c_r = 1;
            REAL g0 = c_t0 * w0 + c_t1 * (-w0) + c_r * (w0 * t0 + x - i0);
            REAL g1 = c_t0 * w1 + c_t1 * (-w1) + c_r * (w1 * t0 + y - i1);
            REAL g2 = c_t0 * w2 + c_t1 * (-w2) + c_r * (w2 * t0 + z - i2);
            back(g0, g1, g2, (*out)[i]);
        }
    }
};

// ---- cube.hpp ---------------------------------------------------------------------------------
struct cube : implicit_function {
    REAL p[18];
    REAL x = 0.f, y = 0.f, z = 0.f;
    Env env;
    cube(const REAL m[12], const Env& e) : env(e) {
        for (int i = 0; i < 18; i++) p[i] = 0;
        p[0] = 0.5f; p[3] = -0.5f; p[7] = 0.5f; p[10] = -0.5f; p[14] = 0.5f; p[17] = -0.5f;
        set_matrix(m);
    }
    // one plane term of cube.hpp:204-209 / :296-300, formed in double like the source
    double plane(int k, REAL i1, REAL i2, REAL i3) const {
        return (i1 - x - p[0 + k * 3]) * p[0 + k * 3] * (-2.) + (i2 - y - p[1 + k * 3]) * p[1 + k * 3] * (-2.) +
               (i3 - z - p[2 + k * 3]) * p[2 + k * 3] * (-2.);
    }
    void eval_implicit(const vects& xx, scalars* f) const override {
        vects lx = xx;
        matrix_vector_product(Minv, lx);
        for (size_t i = 0; i < lx.n(); i++) {
            if (env.cube_mode == CUBE_ANALYTIC) {
                REAL i1 = lx[i][0], i2 = lx[i][1], i3 = lx[i][2];
                // min(...) of the six plane terms; every term is exactly representable in binary32
                REAL m5 = (REAL)plane(5, i1, i2, i3);
                REAL m4 = std::min((REAL)plane(4, i1, i2, i3), m5);
                REAL m3 = std::min((REAL)plane(3, i1, i2, i3), m4);
                REAL m2 = std::min((REAL)plane(2, i1, i2, i3), m3);
                REAL m1 = std::min((REAL)plane(1, i1, i2, i3), m2);
                (*f)[i] = std::min((REAL)plane(0, i1, i2, i3), m1);
            } else {
                const SdfGrid& g = *env.sdf;
                const float px = lx[i][0], py = lx[i][1], pz = lx[i][2];
                float res = 10000;
                if (g.ox + g.h * g.nx < px || px < g.ox) {
                } else if (g.oy + g.h * g.ny < py || py < g.oy) {
                } else if (g.oz + g.h * g.nz < pz || pz < g.oz) {
                } else {
                    int xc = (px - g.ox) / g.h, yc = (py - g.oy) / g.h, zc = (pz - g.oz) / g.h;
                    float xl = g.ox + xc * g.h, yl = g.oy + yc * g.h, zl = g.oz + zc * g.h;
                    float xd = (px - xl) / g.h, yd = (py - yl) / g.h, zd = (pz - zl) / g.h;
                    auto at = [&](int a, int b, int c) {
                        size_t idx = (size_t)a + (size_t)b * g.nx + (size_t)c * g.nx * g.ny;
                        return idx < g.v.size() ? g.v[idx] : 0.0f;   // reference reads past the table here (UB)
                    };
                    float c00 = at(xc, yc, zc) * (1 - xd) + at(xc + 1, yc, zc) * xd;
                    float c01 = at(xc, yc, zc + 1) * (1 - xd) + at(xc + 1, yc, zc + 1) * xd;
                    float c10 = at(xc, yc + 1, zc) * (1 - xd) + at(xc + 1, yc + 1, zc) * xd;
                    float c11 = at(xc, yc + 1, zc + 1) * (1 - xd) + at(xc + 1, yc + 1, zc + 1) * xd;
                    float c0 = c00 * (1 - yd) + c10 * yd;
                    float c1 = c01 * (1 - yd) + c11 * yd;
                    res = c0 * (1 - zd) + c1 * (zd);
                }
                (*f)[i] = -res;
            }
        }
    }
    void eval_gradient(const vects& xx, vects* out) const override {
        vects lx = xx;
        matrix_vector_product(Minv, lx);
        for (size_t i = 0; i < lx.n(); i++) {
            REAL i1 = lx[i][0], i2 = lx[i][1], i3 = lx[i][2];
            int index = 0;
            REAL mn = plane(0, i1, i2, i3);
            for (int k = 0; k < 6; k++)
                if (plane(k, i1, i2, i3) < mn) { index = k; mn = plane(k, i1, i2, i3); }
            back(-p[index * 3 + 0], -p[index * 3 + 1], -p[index * 3 + 2], (*out)[i]);
        }
    }
};

// ---- scone.hpp --------------------------------------------------------------------------------
struct scone : implicit_function {
    REAL h = 1.f, r1 = 0.0f, r2 = 0.5f, x0 = 0.f, y0 = 0.f, z0 = 0.5f;
    explicit scone(const REAL m[12]) { set_matrix(m); }
    void eval_implicit(const vects& xx, scalars* f) const override {
        const vects lx = prepare_inner_vectors(Minv, xx);
        const REAL a2 = squared(r2 / h);
        for (size_t i = 0; i < lx.n(); i++) {
            REAL x = lx[i][0], y = lx[i][1], z = lx[i][2];
            REAL fv = -std::sqrt((x - x0) * (x - x0) + (y - y0) * (y - y0)) + std::sqrt((z - z0) * (z - z0) * a2);
            REAL uperside = -(z - z0) - r1;
            REAL lowerside = (z - z0) + h;
            (*f)[i] = std::min(fv, std::min(uperside, lowerside));
        }
    }
    void eval_gradient(const vects& xx, vects* out) const override {
        const vects lx = prepare_inner_vectors(Minv, xx);
        const REAL a2 = squared(r2 / h);
        for (size_t i = 0; i < lx.n(); i++) {
            REAL x = lx[i][0], y = lx[i][1], z = lx[i][2];
            REAL fv = -(x - x0) * (x - x0) / a2 - (y - y0) * (y - y0) / a2 + (z - z0) * (z - z0);
            REAL uperside = -(z - z0) - r1;
            REAL lowerside = (z - z0) + h;
            REAL gx, gy, gz;
            if (uperside < fv && uperside < lowerside) { gx = 0.; gy = 0.; gz = -1.; }
            else if (lowerside < fv && lowerside < uperside) { gx = 0.; gy = 0.; gz = +1.; }
            else { gx = -2 * (x - x0) / a2; gy = -2 * (y - y0) / a2; gz = 2 * (z - z0); }
            back(gx, gy, gz, (*out)[i]);
        }
    }
};

// ---- meta_balls_Rydgard.hpp -------------------------------------------------------------------
struct meta_balls : implicit_function {
    struct ball_t { REAL ballx, bally, ballz, strength, subtract, scale; };
    std::vector<ball_t> balls;
    meta_balls(const REAL m[12], int num_blobs, REAL time, REAL scale) {
        int numblobs = num_blobs;
        for (int ball_i = 0; ball_i < numblobs; ball_i++) {          // :49-59, host double math
            REAL x0 = 0.5, y0 = 0.5, z0 = 0.5;
            REAL D = 1;
            REAL ballx = sin(ball_i + 1.26 * time * (1.03 + 0.5 * cos(0.21 * ball_i))) * 0.27 * D + 0.5 - x0;
            REAL bally = std::abs(cos(ball_i + 1.12 * time * cos(1.22 + 0.1424 * ball_i))) * 0.77 * D - y0;
            REAL ballz = cos(ball_i + 1.32 * time * 0.1 * sin((0.92 + 0.53 * ball_i))) * 0.27 * D + 0.5 - z0;
            REAL subtract = 12;
            REAL strength = 1.2 / ((sqrt(numblobs) - 1) / 4 + 1);
            balls.push_back(ball_t{ballx, bally, ballz, strength, subtract, scale});
        }
        set_matrix(m);
    }
    // extension ("balls" key; SURVEY §8(d) C3 variant (ii), oracle-only semantics: the reference's factory cannot express it):
    // the blobs given explicitly, evaluated by the same loops
    meta_balls(const REAL m[12], const std::vector<REAL>& flat5, REAL scale) {
        for (size_t i = 0; i + 4 < flat5.size(); i += 5) balls.push_back(ball_t{flat5[i], flat5[i + 1], flat5[i + 2], flat5[i + 3], flat5[i + 4], scale});
        set_matrix(m);
    }
    void eval_implicit(const vects& x, scalars* f) const override {
        for (size_t i = 0; i < x.n(); i++) (*f)[i] = 0.0;
        const vects lx = prepare_inner_vectors(Minv, x);
        for (const ball_t& b : balls)
            for (size_t i = 0; i < lx.n(); i++) {
                REAL fx = lx[i][0] - b.ballx; REAL fx2 = fx * fx;
                REAL fy = lx[i][1] - b.bally; REAL fy2 = fy * fy;
                REAL fz = lx[i][2] - b.ballz; REAL fz2 = fz * fz;
                REAL h = ((REAL)0.000001 + fx2 + fy2 + fz2);
                REAL hinv = 1.0 / h;
                REAL val = b.strength * hinv - b.subtract;
                (*f)[i] += val / 100;
            }
    }
    void eval_gradient(const vects& x, vects* out) const override {
        const vects lx = prepare_inner_vectors(Minv, x);
        for (size_t i = 0; i < lx.n(); i++) {
            REAL gx = 0, gy = 0, gz = 0;
            for (const ball_t& b : balls) {
                REAL fz = lx[i][2] - b.ballz; REAL fz2 = fz * fz;
                REAL fy = lx[i][1] - b.bally; REAL fy2 = fy * fy;
                REAL fx = lx[i][0] - b.ballx; REAL fx2 = fx * fx;
                REAL h = 0.000001 + fx2 + fy2 + fz2;          // double literal: summed in double (:118-136)
                REAL hinv = 1.0 / h;
                REAL gradx = b.strength * (-2 * fx * hinv * hinv) / 100;
                REAL grady = b.strength * (-2 * fy * hinv * hinv) / 100;
                REAL gradz = b.strength * (-2 * fz * hinv * hinv) / 100;
                gx += gradx; gy += grady; gz += gradz;
            }
            back(gx, gy, gz, (*out)[i]);
        }
    }
};

// ---- torus.hpp:42-118 -------------------------------------------------------------------------
// std::pow(float, int) and std::pow(float, double) promote to double (C++11 <cmath>), so f is one double
// expression rounded once to REAL; the gradient's `a` is a double power rounded to REAL, the rest binary32.
struct torus : implicit_function {
    REAL r = 4.f, rx = 0.2f, ry = 0.2f, rz = 0.2f;
    explicit torus(const REAL m[12]) { set_matrix(m); }
    void eval_implicit(const vects& xx, scalars* out) const override {
        const vects lx = prepare_inner_vectors(Minv, xx);
        for (size_t i = 0; i < lx.n(); i++) {
            REAL x = lx[i][0], y = lx[i][1], z = lx[i][2];
            REAL f = (1 - std::pow(r - std::pow((std::pow(x / rx, 2) + std::pow(y / ry, 2)), 0.5), 2) - std::pow(z / rz, 2));
            (*out)[i] = f;
        }
    }
    void eval_gradient(const vects& xx, vects* out) const override {
        const vects lx = prepare_inner_vectors(Minv, xx);
        for (size_t i = 0; i < lx.n(); i++) {
            REAL x = lx[i][0], y = lx[i][1], z = lx[i][2];
            REAL a = std::pow(x * x / (rx * rx) + y * y / (ry * ry), 0.5);
            REAL gx = (2 * x / (rx * rx * a)) * (r - a);
            REAL gy = (2 * y / (ry * ry * a)) * (r - a);
            REAL gz = -2 * z / (rz * rz);
            back(gx, gy, gz, (*out)[i]);
        }
    }
};

// ---- heart.hpp:33-52,78-126 -------------------------------------------------------------------
// Mixed precision as written: i1*i1 and i3*i3 are binary32 products, (9./4.)*i2*i2 and everything they are
// added to is double; the i1*i1*i3*i3*i3 chain is binary32; one rounding to REAL at the store.
struct heart : implicit_function {
    explicit heart(const REAL m[12]) { set_matrix(m); }
    void eval_implicit(const vects& xx, scalars* out) const override {
        vects lx = xx;
        matrix_vector_product(Minv, lx);
        for (size_t i = 0; i < lx.n(); i++) {
            REAL i1 = lx[i][0], i2 = lx[i][1], i3 = lx[i][2];
            (*out)[i] = -(std::pow(i1 * i1 + (9. / 4.) * i2 * i2 + i3 * i3 - 1., 3) - i1 * i1 * i3 * i3 * i3 - (9. / 200.) * i2 * i2 * i3 * i3 * i3);
        }
    }
    void eval_gradient(const vects& xx, vects* out) const override {
        vects lx = xx;
        matrix_vector_product(Minv, lx);
        for (size_t i = 0; i < lx.n(); i++) {
            REAL i1 = lx[i][0], i2 = lx[i][1], i3 = lx[i][2];
            REAL a = pow(i1 * i1 + (9. / 4.) * i2 * i2 + i3 * i3 - 1, 2);
            REAL g0 = -6. * i1 * a + 2. * i1 * i3 * i3 * i3;
            REAL g1 = -(27. / 2) * i2 * a + (9. / 100.) * i2 * i3 * i3 * i3;
            REAL g2 = -6. * i3 * a + 3. * i1 * i1 * i3 * i3 + (27. / 200.) * i2 * i2 * i3 * i3;
            back(g0, g1, g2, (*out)[i]);
        }
    }
};

// ---- basic_functions.hpp:22-31, configs.hpp:33 ------------------------------------------------
inline REAL sign(REAL v, REAL tol) { return (v > +tol) ? (+1) : (v < -tol) ? (-1) : (0.0); }
const REAL ROOT_TOLERANCE = 0.001 / 10.0;

// ---- tetrahedron.hpp:18-134,136-193 -----------------------------------------------------------
// The corners are moved by the FORWARD matrix once (constructor :136-143); points are never transformed.
struct tetrahedron : implicit_function {
    REAL p[4][3];
    REAL planes[4][4];
    static void calculatePlaneCoefficients(REAL x1, REAL y1, REAL z1, REAL x2, REAL y2, REAL z2, REAL x3, REAL y3, REAL z3, REAL& a, REAL& b,
                                           REAL& c, REAL& d) {
        a = y1 * z2 - y1 * z3 - y2 * z1 + y2 * z3 + y3 * z1 - y3 * z2;
        b = x1 * z3 - x1 * z2 + x2 * z1 - x2 * z3 - x3 * z1 + x3 * z2;
        c = x1 * y2 - x1 * y3 - x2 * y1 + x2 * y3 + x3 * y1 - x3 * y2;
        d = x1 * y3 * z2 - x1 * y2 * z3 + x2 * y1 * z3 - x2 * y3 * z1 - x3 * y1 * z2 + x3 * y2 * z1;
    }
    void getPlanes() {
        static const int others[4][3] = {{1, 2, 3}, {0, 2, 3}, {0, 1, 3}, {0, 1, 2}};
        for (int k = 0; k < 4; k++) {
            const int* o = others[k];
            calculatePlaneCoefficients(p[o[0]][0], p[o[0]][1], p[o[0]][2], p[o[1]][0], p[o[1]][1], p[o[1]][2], p[o[2]][0], p[o[2]][1],
                                       p[o[2]][2], planes[k][0], planes[k][1], planes[k][2], planes[k][3]);
        }
        for (int k = 0; k < 4; k++) {
            REAL s = sign(planes[k][0] * p[k][0] + planes[k][1] * p[k][1] + planes[k][2] * p[k][2] + planes[k][3], ROOT_TOLERANCE);
            for (int j = 0; j < 4; j++) planes[k][j] *= s;
        }
    }
    // integrity_invariant (tetrahedron.hpp:212-321): the reference asserts it in eval_implicit; its own three FAILED TESTS
    // (:346-383) are the fixtures tests/test_oracle_pins.py holds this restatement — and through it getPlanes — to.
    // Returns 0 when the tetrahedron is sound, else the number of the check that rejects it (1 points too close, 2 three
    // points on a line, 3 a corner too close to the opposite plane).
    int integrity_failure() const {
        const REAL MIN_PRINTABLE_LENGTH = 0.01 / 2.0, MIN_THICKNESS = 0.2;      // configs.hpp:45-46
        for (int i = 0; i < 3; i++)
            for (int j = i + 1; j < 4; j++) {
                REAL d = std::sqrt((p[i][0] - p[j][0]) * (p[i][0] - p[j][0]) + (p[i][1] - p[j][1]) * (p[i][1] - p[j][1]) +
                                   (p[i][2] - p[j][2]) * (p[i][2] - p[j][2]));
                if (d < MIN_PRINTABLE_LENGTH) return 1;
            }
        for (int i = 0; i < 4; i++) {
            const int i_1 = (i + 1) % 4, i_2 = (i + 2) % 4, i_3 = (i + 3) % 4;
            REAL u[3], v[3];
            for (int k = 0; k < 3; k++) { u[k] = p[i_2][k] - p[i_1][k]; v[k] = p[i_3][k] - p[i_1][k]; }
            REAL d = std::sqrt((u[1] * v[2] - u[2] * v[1]) * (u[1] * v[2] - u[2] * v[1]) + (u[2] * v[0] - u[0] * v[2]) * (u[2] * v[0] - u[0] * v[2]) +
                               (u[0] * v[1] - u[1] * v[0]) * (u[0] * v[1] - u[1] * v[0]));
            if (d < MIN_PRINTABLE_LENGTH) return 2;
        }
        for (int i = 0; i < 4; i++) {
            REAL d = planes[i][0] * p[i][0] + planes[i][1] * p[i][1] + planes[i][2] * p[i][2] + planes[i][3];
            d /= std::sqrt(planes[i][0] * planes[i][0] + planes[i][1] * planes[i][1] + planes[i][2] * planes[i][2]);
            d = d < 0 ? -d : d;
            if (!(d >= MIN_THICKNESS)) return 3;      // NaN (a degenerate plane) fails like a small distance
        }
        return 0;
    }
    tetrahedron(const REAL corners[4][3], const REAL m[12]) {
        for (int i = 0; i < 12; i++) { M[i] = m[i]; Minv[i] = (i % 5 == 0) ? 1.0f : 0.0f; }
        vects c(4);
        for (int i = 0; i < 4; i++)
            for (int j = 0; j < 3; j++) c[i][j] = corners[i][j];
        matrix_vector_product(m, c);
        for (int i = 0; i < 4; i++)
            for (int j = 0; j < 3; j++) p[i][j] = c[i][j];
        getPlanes();
    }
    REAL plane(int k, REAL i1, REAL i2, REAL i3) const { return planes[k][0] * i1 + planes[k][1] * i2 + planes[k][2] * i3 + planes[k][3]; }
    void eval_implicit(const vects& x, scalars* out) const override {
        for (size_t i = 0; i < x.n(); i++) {
            REAL i1 = x[i][0], i2 = x[i][1], i3 = x[i][2];
            (*out)[i] = std::min(plane(0, i1, i2, i3), std::min(plane(1, i1, i2, i3), std::min(plane(2, i1, i2, i3), plane(3, i1, i2, i3))));
        }
    }
    void eval_gradient(const vects& x, vects* out) const override {
        for (size_t i = 0; i < x.n(); i++) {
            REAL i1 = x[i][0], i2 = x[i][1], i3 = x[i][2];
            int index = 0;
            REAL mn = plane(0, i1, i2, i3);
            for (int k = 1; k < 4; k++) {
                REAL possible_min = plane(k, i1, i2, i3);
                if (possible_min < mn) { index = k; mn = possible_min; }
            }
            (*out)[i][0] = planes[index][0];
            (*out)[i][1] = planes[index][1];
            (*out)[i][2] = planes[index][2];
        }
    }
};

// ---- double_mushroom.hpp:21-43,68-134 ---------------------------------------------------------
// Own matrices are the identity (the factory wraps it in linearly_transformed, object_factory.hpp:86-99), but the
// identity product is still carried out.  std::pow(float, int) is a double power: f is one double expression.
struct double_mushroom : implicit_function {
    REAL r, a, b, c, x = 0, y = 0, z = 0;
    double_mushroom(REAL height_mm, REAL center_radius_x_mm, REAL center_radius_y_mm, REAL curvature) {
        r = height_mm / 2;
        a = center_radius_x_mm;
        b = center_radius_y_mm;
        c = 1 / curvature;
        for (int i = 0; i < 12; i++) M[i] = Minv[i] = (i % 5 == 0) ? 1.0f : 0.0f;
    }
    void eval_implicit(const vects& xx, scalars* out) const override {
        vects lx = xx;
        matrix_vector_product(Minv, lx);
        const REAL a2 = squared(a), b2 = squared(b), c2 = squared(c);
        for (size_t i = 0; i < lx.n(); i++) {
            if (lx[i][2] > r) (*out)[i] = r - lx[i][2];
            else if (lx[i][2] < -r) (*out)[i] = r + lx[i][2];
            else (*out)[i] = -(std::pow(lx[i][0] - x, 2) / a2 + std::pow(lx[i][1] - y, 2) / b2 - pow(lx[i][2] - z, 2) / c2 - 1);
        }
    }
    void eval_gradient(const vects& xx, vects* out) const override {
        vects lx = xx;
        matrix_vector_product(Minv, lx);
        const REAL a2 = squared(a), b2 = squared(b), c2 = squared(c);
        for (size_t i = 0; i < lx.n(); i++) {
            REAL g0, g1, g2;
            if (lx[i][2] < -r) { g2 = 1; g1 = 0; g0 = 0; }
            else if (lx[i][2] > r) { g2 = -1; g1 = 0; g0 = 0; }
            else {
                g0 = -2 * (lx[i][0] - x) / a2;
                g1 = -2 * (lx[i][1] - y) / b2;
                g2 = 2 * (lx[i][2] - z) / c2;
            }
            back(g0, g1, g2, (*out)[i]);
        }
    }
};

// ---- linearly_transformed.hpp:24-59 -----------------------------------------------------------
struct linearly_transformed : implicit_function {
    const implicit_function* child;
    linearly_transformed(const implicit_function* c, const REAL m[12]) : child(c) { set_matrix(m); }
    void eval_implicit(const vects& x, scalars* f) const override {
        const vects lx = prepare_inner_vectors(Minv, x);
        child->eval_implicit(lx, f);
    }
    void eval_gradient(const vects& x, vects* out) const override {
        const vects lx = prepare_inner_vectors(Minv, x);
        child->eval_gradient(lx, out);
        for (size_t i = 0; i < lx.n(); i++) {
            REAL gx = (*out)[i][0], gy = (*out)[i][1], gz = (*out)[i][2];
            back(gx, gy, gz, (*out)[i]);
        }
    }
};

// ---- top_bottom_lid.hpp:51-76 -----------------------------------------------------------------
// No transform at all (the matrix is only checked); max(z - 0.5, -(z + 0.5)); the gradient's first branch can never
// be taken (z >= 0.5 && 0 > z), so it is (0, 0, 1) everywhere.
struct top_bottom_lid : implicit_function {
    explicit top_bottom_lid(const REAL m[12]) { for (int i = 0; i < 12; i++) { M[i] = m[i]; Minv[i] = (i % 5 == 0) ? 1.0f : 0.0f; } }
    void eval_implicit(const vects& x, scalars* out) const override {
        for (size_t i = 0; i < x.n(); i++) {
            REAL p = x[i][2] - 0.5f, q = (x[i][2] + 0.5f) * -1;
            (*out)[i] = (p < q) ? q : p;          // Eigen's array max
        }
    }
    void eval_gradient(const vects& x, vects* out) const override {
        for (size_t i = 0; i < x.n(); i++) {
            const REAL z = x[i][2];
            (*out)[i][0] = 0;
            (*out)[i][1] = 0;
            (*out)[i][2] = (z >= 0.5 && (0.0 > z && z >= -0.5)) ? -1 : 1;
        }
    }
};

// ---- half_plane.hpp:85-150 --------------------------------------------------------------------
// The reference inverts the 4x4 with Eigen (Matrix4f::inverse(): cofactors / determinant) and evaluates through
// Eigen products.  Eigen is not in this image and the reference pins no version: restated as the plain cofactor
// inverse and left-to-right sums in binary32 — tolerance-level parity only for this primitive.
inline void cofactor_inverse4(const REAL m12[12], REAL inv16[16]) {
    REAL a[16];
    for (int i = 0; i < 12; i++) a[i] = m12[i];
    a[12] = 0; a[13] = 0; a[14] = 0; a[15] = 1;
    REAL cof[16];
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++) {
            REAL s[9];
            int k = 0;
            for (int i = 0; i < 4; i++)
                for (int j = 0; j < 4; j++)
                    if (i != r && j != c) s[k++] = a[i * 4 + j];
            REAL det3 = s[0] * (s[4] * s[8] - s[5] * s[7]) - s[1] * (s[3] * s[8] - s[5] * s[6]) + s[2] * (s[3] * s[7] - s[4] * s[6]);
            cof[r * 4 + c] = ((r + c) & 1) ? -det3 : det3;
        }
    REAL det = a[0] * cof[0] + a[1] * cof[1] + a[2] * cof[2] + a[3] * cof[3];
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++) inv16[r * 4 + c] = cof[c * 4 + r] / det;
}
struct half_plane : implicit_function {
    REAL n[3], pp[3];
    half_plane(const REAL m[12], const REAL plane_vector[3], const REAL plane_point[3]) {
        REAL norm = std::sqrt(plane_vector[0] * plane_vector[0] + plane_vector[1] * plane_vector[1] + plane_vector[2] * plane_vector[2]);
        for (int i = 0; i < 3; i++) { n[i] = plane_vector[i] / norm; pp[i] = plane_point[i]; }
        REAL inv[16];
        cofactor_inverse4(m, inv);
        for (int i = 0; i < 12; i++) { M[i] = m[i]; Minv[i] = inv[i]; }
    }
    REAL value(const REAL* x) const {
        REAL lx = Minv[0] * x[0] + Minv[1] * x[1] + Minv[2] * x[2];
        REAL ly = Minv[4] * x[0] + Minv[5] * x[1] + Minv[6] * x[2];
        REAL lz = Minv[8] * x[0] + Minv[9] * x[1] + Minv[10] * x[2];
        lx = lx + Minv[3]; ly = ly + Minv[7]; lz = lz + Minv[11];
        return (lx - pp[0]) * n[0] + (ly - pp[1]) * n[1] + (lz - pp[2]) * n[2];
    }
    void eval_implicit(const vects& x, scalars* out) const override {
        for (size_t i = 0; i < x.n(); i++) (*out)[i] = value(x[i]);
    }
    void eval_gradient(const vects& x, vects* out) const override {
        for (size_t i = 0; i < x.n(); i++) {
            const REAL s = (value(x[i]) >= 0) ? -1.0f : 1.0f;
            back(s * n[0], s * n[1], s * n[2], (*out)[i]);
        }
    }
};

// ---- transformed_union / intersection / subtract ----------------------------------------------
enum CsgOp { OP_UNION, OP_INTERSECTION, OP_SUBTRACT };
struct csg_node : implicit_function {
    CsgOp op;
    const implicit_function* a;
    const implicit_function* b;
    csg_node(CsgOp op_, const implicit_function* a_, const implicit_function* b_, const REAL m[12]) : op(op_), a(a_), b(b_) {
        set_matrix(m);
    }
    void eval_implicit(const vects& x, scalars* f) const override {
        const size_t n = x.n();
        const vects lx = prepare_inner_vectors(Minv, x);
        scalars f1(n), f2(n);
        a->eval_implicit(lx, &f1);
        b->eval_implicit(lx, &f2);
        for (size_t i = 0; i < n; i++) {
            switch (op) {
                case OP_UNION:        (*f)[i] = (f1[i] > f2[i]) ? f1[i] : f2[i]; break;     // transformed_union.hpp:47
                case OP_INTERSECTION: (*f)[i] = (f1[i] > f2[i]) ? f2[i] : f1[i]; break;     // transformed_intersection.hpp:49
                case OP_SUBTRACT:     (*f)[i] = (f1[i] < -f2[i]) ? f1[i] : -f2[i]; break;   // transformed_subtract.hpp:52
            }
        }
    }
    void eval_gradient(const vects& x, vects* out) const override {
        const size_t n = x.n();
        scalars f1(n), f2(n);
        vects g1(n), g2(n);
        const vects lx = prepare_inner_vectors(Minv, x);
        a->eval_implicit(lx, &f1);
        b->eval_implicit(lx, &f2);
        a->eval_gradient(lx, &g1);
        b->eval_gradient(lx, &g2);
        if (op == OP_SUBTRACT)
            for (size_t i = 0; i < n * 3; i++) g2.d[i] = -g2.d[i];
        for (size_t i = 0; i < n; i++) {
            bool first;
            switch (op) {
                case OP_UNION:        first = (f1[i] > f2[i]); break;
                case OP_INTERSECTION: first = !(f1[i] > f2[i]); break;
                default:              first = (f1[i] < -f2[i]); break;
            }
            const REAL* g = first ? g1[i] : g2[i];
            back(g[0], g[1], g[2], (*out)[i]);
        }
    }
};

// ---- object_factory.hpp -----------------------------------------------------------------------
struct Shape {                       // owns every node (object_collector.hpp:10-32)
    std::vector<std::unique_ptr<implicit_function>> pool;
    const implicit_function* root = nullptr;
    template <class T> T* keep(T* p) { pool.emplace_back(p); return p; }
};

inline void getMatrix12(REAL m[12], const ojson::Node& node) {        // :21-30 (reads 16, UB; we stop at 12)
    const ojson::Node* mat = node.child("matrix");
    if (!mat) { fprintf(stderr, "oracle: node without \"matrix\"\n"); abort(); }
    int i = 0;
    for (auto& kv : mat->kids) {
        if (i >= 12) break;
        float v = 0;
        if (!kv.second.as_float(&v)) { fprintf(stderr, "oracle: bad matrix entry\n"); abort(); }
        m[i++] = v;
    }
    if (i < 12) { fprintf(stderr, "oracle: matrix needs 12 entries\n"); abort(); }
}
inline void copy_eye(REAL m[12]) {
    const REAL eye[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
    for (int j = 0; j < 12; j++) m[j] = eye[j];
}

inline const implicit_function* object_factory(Shape& S, const ojson::Node& node, bool ignore_root_matrix, const Env& env) {
    const ojson::Node* t = node.child("type");
    std::string name = t ? t->text : std::string();
    REAL m[12];
    auto matrix = [&]() { getMatrix12(m, node); if (ignore_root_matrix) copy_eye(m); };
    if (name == "icube" || name == "cube") { matrix(); return S.keep(new cube(m, env)); }
    if (name == "icylinder" || name == "cylinder") { matrix(); return S.keep(new scylinder(m)); }
    if (name == "iellipsoid" || name == "ellipsoid") { matrix(); return S.keep(new egg(m)); }
    if (name == "icone" || name == "cone") { matrix(); return S.keep(new scone(m)); }
    if (name == "iheart") { matrix(); return S.keep(new heart(m)); }                     // :153-162
    if (name == "itorus") { matrix(); return S.keep(new torus(m)); }                     // :163-173
    if (name == "implicit_double_mushroom") {                                            // :86-99
        const implicit_function* dm = S.keep(new double_mushroom(0.9, 0.4 / 2, 0.4 / 2, 1 / 0.2));
        matrix();
        return S.keep(new linearly_transformed(dm, m));
    }
    if (name == "tetrahedron") {                                                         // :175-188, getCorners :32-45
        getMatrix12(m, node);
        REAL corners[4][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
        const ojson::Node* cs = node.child("corners");
        if (!cs || cs->kids.size() != 4) { fprintf(stderr, "oracle: tetrahedron needs 4 corners\n"); abort(); }
        for (int i = 0; i < 4; i++) {
            const ojson::Node& row = cs->kids[i].second;
            if (row.kids.size() != 3) { fprintf(stderr, "oracle: a corner needs 3 coordinates\n"); abort(); }
            for (int j = 0; j < 3; j++)
                if (!row.kids[j].second.as_float(&corners[i][j])) { fprintf(stderr, "oracle: bad corner\n"); abort(); }
        }
        if (ignore_root_matrix) copy_eye(m);
        return S.keep(new tetrahedron(corners, m));
    }
    if (name == "half_plane") {                                                          // :396-434
        matrix();
        REAL pv[3] = {0, 0, 0}, ppt[3] = {0, 0, 0};
        const ojson::Node* a = node.child("plane_vector");
        const ojson::Node* b = node.child("plane_point");
        if (!a || !b || a->kids.size() != 3 || b->kids.size() != 3) { fprintf(stderr, "oracle: half_plane needs plane_vector and plane_point\n"); abort(); }
        for (int j = 0; j < 3; j++) { a->kids[j].second.as_float(&pv[j]); b->kids[j].second.as_float(&ppt[j]); }
        return S.keep(new half_plane(m, pv, ppt));
    }
    if (name == "top_bottom_lid") { matrix(); return S.keep(new top_bottom_lid(m)); }    // :480-505
    if (name == "meta_balls") {                                                          // :654-673
        matrix();
        REAL time = node.get_float("time", 0.1);
        if (const ojson::Node* bl = node.child("balls")) {
            std::vector<REAL> flat;
            for (auto& kv : bl->kids) { float v = 0; if (!kv.second.as_float(&v)) { fprintf(stderr, "oracle: bad balls entry\n"); abort(); } flat.push_back(v); }
            if (flat.empty() || flat.size() % 5 != 0) { fprintf(stderr, "oracle: balls needs 5 numbers per ball\n"); abort(); }
            return S.keep(new meta_balls(m, flat, 1.0));
        }
        return S.keep(new meta_balls(m, 4, time, 1.0));
    }
    if (name == "Union" || name == "Intersection" || name == "Difference") {
        matrix();
        const ojson::Node* ch = node.child("children");
        size_t nch = ch ? ch->kids.size() : 0;
        if (name == "Union") {                                                           // :537-580 left fold
            if (nch < 2) { fprintf(stderr, "oracle: Union needs >= 2 children (reference returns an uninitialised pointer)\n"); abort(); }
            const REAL eye[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
            const implicit_function* a = object_factory(S, ch->kids[0].second, false, env);
            for (size_t i = 1; i < nch; i++) {
                const implicit_function* b = object_factory(S, ch->kids[i].second, false, env);
                bool last = (i + 1 == nch);
                a = S.keep(new csg_node(OP_UNION, a, b, last ? m : eye));
            }
            return a;
        }
        if (nch != 2) { fprintf(stderr, "oracle: %s needs exactly 2 children\n", name.c_str()); abort(); }
        const implicit_function* a = object_factory(S, ch->kids[0].second, false, env);
        const implicit_function* b = object_factory(S, ch->kids[1].second, false, env);
        return S.keep(new csg_node(name == "Intersection" ? OP_INTERSECTION : OP_SUBTRACT, a, b, m));
    }
    fprintf(stderr, "Invalid object you asked for: \"%s\"\n", name.c_str());              // :731-733
    abort();
}

// ---- host threads (bench baselines and large parity cases only) ---------------------------------------------------
// The reference is single-threaded.  Its whole-array passes are pointwise (point i of the output depends on point i of
// the input only), so splitting an array into contiguous chunks and evaluating the chunks on different threads gives
// the same bits as the one call the reference makes.  threads() == 1 (the default) makes the reference's single call.
inline int& threads() { static int t = 1; return t; }

// fn(k) for k in [0, n), work handed out dynamically to threads() std::threads (no OpenMP runtime needed)
template <class F>
inline void parallel_for(long n, const F& fn) {
    const int t = (int)std::min<long>(threads(), n);
    if (t <= 1) { for (long k = 0; k < n; k++) fn(k); return; }
    std::atomic<long> next(0);
    auto worker = [&]() { for (long k; (k = next.fetch_add(1)) < n;) fn(k); };
    std::vector<std::thread> pool;
    for (int i = 1; i < t; i++) pool.emplace_back(worker);
    worker();
    for (std::thread& th : pool) th.join();
}

struct chunked_function : implicit_function {
    const implicit_function* inner;
    explicit chunked_function(const implicit_function* f) : inner(f) {}
    static size_t chunk_of(size_t n) {
        size_t c = (n + (size_t)threads() * 4 - 1) / ((size_t)threads() * 4);
        return c < 4096 ? 4096 : c;
    }
    void eval_implicit(const vects& x, scalars* f) const override {
        const size_t n = x.n();
        if (threads() <= 1 || n < 8192) { inner->eval_implicit(x, f); return; }
        const size_t c = chunk_of(n), nchunks = (n + c - 1) / c;
        parallel_for((long)nchunks, [&](long k) {
            const size_t b = (size_t)k * c, e = std::min(n, b + c);
            vects xs(e - b);
            std::copy(x.d.begin() + b * 3, x.d.begin() + e * 3, xs.d.begin());
            scalars fs(e - b);
            inner->eval_implicit(xs, &fs);
            std::copy(fs.begin(), fs.end(), f->begin() + b);
        });
    }
    void eval_gradient(const vects& x, vects* g) const override {
        const size_t n = x.n();
        if (threads() <= 1 || n < 8192) { inner->eval_gradient(x, g); return; }
        const size_t c = chunk_of(n), nchunks = (n + c - 1) / c;
        parallel_for((long)nchunks, [&](long k) {
            const size_t b = (size_t)k * c, e = std::min(n, b + c);
            vects xs(e - b);
            std::copy(x.d.begin() + b * 3, x.d.begin() + e * 3, xs.d.begin());
            vects gs(e - b);
            inner->eval_gradient(xs, &gs);
            std::copy(gs.d.begin(), gs.d.end(), g->d.begin() + b * 3);
        });
    }
};

}  // namespace oracle
