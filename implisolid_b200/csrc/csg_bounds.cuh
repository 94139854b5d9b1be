// Interval evaluation of the flattened CSG program over a box, the pruning / specialisation decisions it supports and the
// liveness masks made from them (used by K1's k_super_bounds / k_tile_bounds / k_balance_hist, field_kernel.cuh).
// Like csg_eval.cuh this compiles for the host as well, so that tests/emu can check on the CPU that the bounds really
// contain the exact evaluator's values and that a tile's liveness mask never changes a value inside the tile's box; the host
// build is a test fixture, never a fallback of the product.
#pragma once
#include "csg_eval.cuh"

#if defined(__CUDACC__)
#define ISD_HDI __host__ __device__ inline
#else
#define ISD_HDI inline
#endif

namespace isd {

template <class T>
ISD_HD T isd_ldg(const T* p) {
#if defined(__CUDA_ARCH__)
    return __ldg(p);
#else
    return *p;
#endif
}

// ---- interval evaluation (bounds only) -------------------------------------------------------------------------------
struct Ival { float lo, hi; };
ISD_HD Ival iv_sq(float c, float r) {      // range of t^2 for t in [c - r, c + r]
    const float a = fabsf(c);
    const float mn = fmaxf(a - r, 0.0f), mx = a + r;
    return Ival{mn * mn, mx * mx};
}
ISD_HD float fmin3(float a, float b, float c) { return fminf(a, fminf(b, c)); }

// Bounds of the CSG function over the axis-aligned box centre (bx,by,bz), half extents (hx,hy,hz), in grid coordinates.
// Everything is widened outwards: local boxes by a relative+absolute margin that covers the difference between the
// composite transform used here and the nested float transforms of the exact evaluator; f bounds by a final margin.
//
// With DECIDE, every CSG instruction also records (2 bits, in post-order of the CSG nodes) whether one child can never be
// selected anywhere in the box: a is the reference's first child (f1), b its second (f2), and the tests mirror the
// selectors of transformed_union.hpp:47, transformed_intersection.hpp:49, transformed_subtract.hpp:52 with a margin for
// the rounding of both the bounds and the exact evaluation.
enum : uint32_t { KEEP_BOTH = 0, KILL_A = 1, KILL_B = 2 };
ISD_HD bool surely_greater(float x, float y) {      // x > y by more than the bounds' uncertainty
    return x - y > 1e-4f * (1.0f + fabsf(x) + fabsf(y));
}

//
// With MASKED, `in_mask` holds liveness bits proven for an enclosing box (the supertile): dead primitives are not bounded
// at all and push nothing, `present` tracks which operands exist, and a CSG node with one operand passes it through and
// records the other as killed — so the masks derived from these decisions are never looser than the enclosing one.
struct NoMask {
    ISD_HD bool operator()(uint32_t) const { return true; }
};
struct WordMask {
    const uint32_t* m;
    ISD_HD bool operator()(uint32_t id) const { return (isd_ldg(m + (id >> 5)) >> (id & 31)) & 1u; }
};

template <bool DECIDE, bool MASKED, class InMask>
ISD_HDI Ival program_bounds(const uint32_t* __restrict__ bp, float bx, float by, float bz, float hx, float hy, float hz,
                                      uint32_t* __restrict__ decisions, const InMask& in_mask) {
    Ival st[kMaxValueStack];
    int sp = 0, pc = 0, opj = 0;
    unsigned long long present = 0ull;
    for (;;) {
        const uint32_t word = isd_ldg(bp + pc);
        const uint32_t op = word & OP_MASK;
        if (op == OP_END) break;
        if (op >= OP_UNION) {
            ++pc;
            if (MASKED) {
                const bool b_there = (present & 1ull) != 0, a_there = (present & 2ull) != 0;
                present = ((present >> 2) << 1) | ((a_there || b_there) ? 1ull : 0ull);
                if (!(a_there && b_there)) {
                    uint32_t d = KEEP_BOTH;
                    if (b_there) {
                        d = KILL_A;
                        if (op == OP_SUBTRACT) { const Ival b = st[sp - 1]; st[sp - 1] = Ival{-b.hi, -b.lo}; }
                    } else if (a_there) {
                        d = KILL_B;
                    }
                    if (DECIDE) {
                        if ((opj & 15) == 0) decisions[opj >> 4] = 0u;
                        decisions[opj >> 4] |= d << ((opj & 15) * 2);
                        ++opj;
                    }
                    continue;
                }
            }
            const Ival b = st[--sp], a = st[sp - 1];
            Ival r;
            uint32_t d = KEEP_BOTH;
            if (op == OP_UNION) {
                r = Ival{fmaxf(a.lo, b.lo), fmaxf(a.hi, b.hi)};
                if (DECIDE) d = surely_greater(a.lo, b.hi) ? KILL_B : (surely_greater(b.lo, a.hi) ? KILL_A : KEEP_BOTH);
            } else if (op == OP_INTERSECT) {
                r = Ival{fminf(a.lo, b.lo), fminf(a.hi, b.hi)};
                if (DECIDE) d = surely_greater(a.lo, b.hi) ? KILL_A : (surely_greater(b.lo, a.hi) ? KILL_B : KEEP_BOTH);
            } else {
                r = Ival{fminf(a.lo, -b.hi), fminf(a.hi, -b.lo)};
                if (DECIDE) d = surely_greater(-b.hi, a.hi) ? KILL_B : (surely_greater(a.lo, -b.lo) ? KILL_A : KEEP_BOTH);
            }
            if (DECIDE) {
                if (d == KILL_A) r = (op == OP_SUBTRACT) ? Ival{-b.hi, -b.lo} : b;      // tighter: the result IS that child
                else if (d == KILL_B) r = a;
                if ((opj & 15) == 0) decisions[opj >> 4] = 0u;
                decisions[opj >> 4] |= d << ((opj & 15) * 2);
                ++opj;
            }
            st[sp - 1] = r;
            continue;
        }
        if (MASKED && !in_mask(word >> OP_ID_SHIFT)) {      // dead primitive: no bound needed, nothing pushed
            pc += 13;
            if (op == OP_METABALLS) pc += 1 + 5 * (int)isd_ldg(bp + pc);
            else if (op == OP_TETRA) pc += 16;
            else if (op == OP_HALF_PLANE) pc += 8;
            present <<= 1;
            continue;
        }
        present = (present << 1) | 1ull;
        float m[12];
#pragma unroll
        for (int i = 0; i < 12; ++i) m[i] = as_float(isd_ldg(bp + pc + 1 + i));
        pc += 13;
        float cx = m[0] * bx + m[1] * by + m[2] * bz + m[3];
        float cy = m[4] * bx + m[5] * by + m[6] * bz + m[7];
        float cz = m[8] * bx + m[9] * by + m[10] * bz + m[11];
        float rx = fabsf(m[0]) * hx + fabsf(m[1]) * hy + fabsf(m[2]) * hz;
        float ry = fabsf(m[4]) * hx + fabsf(m[5]) * hy + fabsf(m[6]) * hz;
        float rz = fabsf(m[8]) * hx + fabsf(m[9]) * hy + fabsf(m[10]) * hz;
        const float mag = fabsf(cx) + fabsf(cy) + fabsf(cz) + rx + ry + rz + fabsf(m[3]) + fabsf(m[7]) + fabsf(m[11]) + 1.0f;
        const float pad = 1e-3f * mag;
        rx += pad; ry += pad; rz += pad;
        Ival r;
        if (op == OP_ELLIPSOID) {
            const Ival X = iv_sq(cx, rx), Y = iv_sq(cy, ry), Z = iv_sq(cz, rz);
            r = Ival{1.0f - 4.0f * (X.hi + Y.hi + Z.hi), 1.0f - 4.0f * (X.lo + Y.lo + Z.lo)};
        } else if (op == OP_CYLINDER) {
            const Ival X = iv_sq(cx, rx), Y = iv_sq(cy, ry);
            const float rho_lo = sqrtf(X.lo + Y.lo), rho_hi = sqrtf(X.hi + Y.hi);
            const float zl = cz - rz, zh = cz + rz;
            r = Ival{fmin3(zl + 0.5f, 0.5f - zh, 0.5f - rho_hi), fmin3(zh + 0.5f, 0.5f - zl, 0.5f - rho_lo)};
        } else if (op == OP_CONE) {
            const Ival X = iv_sq(cx, rx), Y = iv_sq(cy, ry);
            const float rho_lo = sqrtf(X.lo + Y.lo), rho_hi = sqrtf(X.hi + Y.hi);
            const float dl = (cz - 0.5f) - rz, dh = (cz - 0.5f) + rz;
            const float al = (dl > 0.0f) ? dl : ((dh < 0.0f) ? -dh : 0.0f), ah = fmaxf(fabsf(dl), fabsf(dh));
            r = Ival{fmin3(-rho_hi + 0.5f * al, -dh, dl + 1.0f), fmin3(-rho_lo + 0.5f * ah, -dl, dh + 1.0f)};
        } else if (op == OP_CUBE) {
            const float xl = cx - rx, xh = cx + rx, yl = cy - ry, yh = cy + ry, zl = cz - rz, zh = cz + rz;
            r = Ival{fminf(fmin3(0.5f - xh, xl + 0.5f, 0.5f - yh), fmin3(yl + 0.5f, 0.5f - zh, zl + 0.5f)),
                     fminf(fmin3(0.5f - xl, xh + 0.5f, 0.5f - yl), fmin3(yh + 0.5f, 0.5f - zl, zh + 0.5f))};
        } else if (op == OP_METABALLS) {
            const int nb = (int)isd_ldg(bp + pc);
            ++pc;
            float lo = 0.0f, hi = 0.0f;
            for (int b = 0; b < nb; ++b) {
                const float px = as_float(isd_ldg(bp + pc)), py = as_float(isd_ldg(bp + pc + 1)), pz = as_float(isd_ldg(bp + pc + 2));
                const float strength = as_float(isd_ldg(bp + pc + 3)), subtract = as_float(isd_ldg(bp + pc + 4));
                pc += 5;
                const Ival X = iv_sq(cx - px, rx), Y = iv_sq(cy - py, ry), Z = iv_sq(cz - pz, rz);
                const float dmin = 0.000001f + X.lo + Y.lo + Z.lo, dmax = 0.000001f + X.hi + Y.hi + Z.hi;
                lo += (strength / dmax - subtract) / 100.0f;
                hi += (strength / dmin - subtract) / 100.0f;
            }
            const float slack = 1e-4f * (fabsf(lo) + fabsf(hi) + (float)nb);
            r = Ival{lo - slack, hi + slack};
        } else if (op == OP_TORUS) {
            // 1 - (4 - sqrt(25 x^2 + 25 y^2))^2 - 25 z^2
            const Ival X = iv_sq(cx, rx), Y = iv_sq(cy, ry), Z = iv_sq(cz, rz);
            const float a_lo = 5.0f * sqrtf(X.lo + Y.lo), a_hi = 5.0f * sqrtf(X.hi + Y.hi);
            const float d_lo = 4.0f - a_hi, d_hi = 4.0f - a_lo;
            const float q_hi = fmaxf(d_lo * d_lo, d_hi * d_hi), q_lo = (d_lo > 0.0f || d_hi < 0.0f) ? fminf(d_lo * d_lo, d_hi * d_hi) : 0.0f;
            r = Ival{1.0f - q_hi - 25.0f * Z.hi, 1.0f - q_lo - 25.0f * Z.lo};
            const float slack = 1e-3f * (fabsf(r.lo) + fabsf(r.hi) + 16.0f);
            r.lo -= slack; r.hi += slack;
        } else if (op == OP_HEART) {
            // -((x^2 + 2.25 y^2 + z^2 - 1)^3 - x^2 z^3 - 0.045 y^2 z^3)
            const Ival X = iv_sq(cx, rx), Y = iv_sq(cy, ry), Z = iv_sq(cz, rz);
            const float u_lo = X.lo + 2.25f * Y.lo + Z.lo - 1.0f, u_hi = X.hi + 2.25f * Y.hi + Z.hi - 1.0f;
            const float zl = cz - rz, zh = cz + rz;
            const float c_lo = zl * zl * zl, c_hi = zh * zh * zh;                     // z^3 is increasing
            const float w_lo = X.lo + 0.045f * Y.lo, w_hi = X.hi + 0.045f * Y.hi;    // >= 0
            const float t_lo = fminf(w_lo * c_lo, w_hi * c_lo), t_hi = fmaxf(w_lo * c_hi, w_hi * c_hi);
            const float p_lo = u_lo * u_lo * u_lo, p_hi = u_hi * u_hi * u_hi;
            r = Ival{-(p_hi - t_lo), -(p_lo - t_hi)};
            const float slack = 1e-3f * (fabsf(p_lo) + fabsf(p_hi) + fabsf(t_lo) + fabsf(t_hi) + 1.0f);
            r.lo -= slack; r.hi += slack;
        } else if (op == OP_MUSHROOM) {
            // z > 0.45: 0.45 - z;  z < -0.45: 0.45 + z;  else 1 - 25 x^2 - 25 y^2 + 25 z^2  (hull of the pieces the box meets)
            const float R = 0.45f, zl = cz - rz, zh = cz + rz;
            float lo = 3.0e38f, hi = -3.0e38f;
            if (zh > R - pad) { lo = fminf(lo, R - zh); hi = fmaxf(hi, R - fmaxf(zl, R - pad)); }
            if (zl < -R + pad) { lo = fminf(lo, R + zl); hi = fmaxf(hi, R + fminf(zh, -R + pad)); }
            if (zh >= -R - pad && zl <= R + pad) {
                const Ival X = iv_sq(cx, rx), Y = iv_sq(cy, ry);
                const float a = fmaxf(zl, -R - pad), b = fminf(zh, R + pad);
                const float m = fmaxf(fabsf(a), fabsf(b)), n = (a > 0.0f) ? a : ((b < 0.0f) ? -b : 0.0f);
                lo = fminf(lo, 1.0f - 25.0f * (X.hi + Y.hi) + 25.0f * n * n);
                hi = fmaxf(hi, 1.0f - 25.0f * (X.lo + Y.lo) + 25.0f * m * m);
            }
            const float slack = 1e-3f * (fabsf(lo) + fabsf(hi) + 1.0f);
            r = Ival{lo - slack, hi + slack};
        } else if (op == OP_LID) {
            // max(z - 0.5, -z - 0.5) = |z| - 0.5 on the parent's coordinates (the composite here has no own inverse)
            const float zl = cz - rz, zh = cz + rz;
            const float m = fmaxf(fabsf(zl), fabsf(zh)), n = (zl > 0.0f) ? zl : ((zh < 0.0f) ? -zh : 0.0f);
            r = Ival{n - 0.5f, m - 0.5f};
        } else if (op == OP_TETRA) {
            float lo = 3.0e38f, hi = 3.0e38f;
            for (int k = 0; k < 4; ++k) {
                const float a = as_float(isd_ldg(bp + pc + 4 * k)), b = as_float(isd_ldg(bp + pc + 4 * k + 1));
                const float c = as_float(isd_ldg(bp + pc + 4 * k + 2)), d = as_float(isd_ldg(bp + pc + 4 * k + 3));
                const float mid = a * cx + b * cy + c * cz + d, rad = fabsf(a) * rx + fabsf(b) * ry + fabsf(c) * rz;
                const float slack = 1e-4f * (fabsf(a * cx) + fabsf(b * cy) + fabsf(c * cz) + fabsf(d) + rad);
                lo = fminf(lo, mid - rad - slack);
                hi = fminf(hi, mid + rad + slack);
            }
            pc += 16;
            r = Ival{lo, hi};
        } else if (op == OP_HALF_PLANE) {
            const float n0 = as_float(isd_ldg(bp + pc)), n1 = as_float(isd_ldg(bp + pc + 1)), n2 = as_float(isd_ldg(bp + pc + 2));
            const float p0 = as_float(isd_ldg(bp + pc + 4)), p1 = as_float(isd_ldg(bp + pc + 5)), p2 = as_float(isd_ldg(bp + pc + 6));
            pc += 8;
            const float mid = (cx - p0) * n0 + (cy - p1) * n1 + (cz - p2) * n2, rad = fabsf(n0) * rx + fabsf(n1) * ry + fabsf(n2) * rz;
            const float slack = 1e-4f * (fabsf(cx - p0) + fabsf(cy - p1) + fabsf(cz - p2) + rad);
            r = Ival{mid - rad - slack, mid + rad + slack};
        } else {   // OP_CUBE_SDF: a table look-up, no bound
            r = Ival{-3.0e38f, 3.0e38f};
        }
        st[sp++] = r;
    }
    Ival out = (MASKED && !(present & 1ull)) ? Ival{-1.0f, -1.0f} : st[0];      // (a live root always has a live operand)
    const float slack = 1e-4f * (1.0f + fabsf(out.lo) + fabsf(out.hi));
    out.lo -= slack; out.hi += slack;
    return out;
}

// tile classes
enum : uint8_t { TILE_OUTSIDE = 0, TILE_INSIDE = 1, TILE_MIXED = 2 };

// Tile specialisation (SpecDesc::enabled): every work-list tile also gets a liveness mask, one bit per tree node (bit = node id,
// reverse post-order).  The root is live; a live CSG node passes liveness to the children its decision did not kill.
struct SpecDesc {
    int enabled;
    int num_nodes, num_ops, mask_words;
    const uint32_t* topo;         // bit i: node id i is a CSG node
    uint32_t* masks;              // [work-list index][mask_words]
    uint32_t* super_masks;        // [supertile-list index][mask_words]
};
constexpr int kDecisionWords = kMaxSpecNodes / 2 / 16 + 1;

// Liveness masks from the decisions of one program_bounds<DECIDE> call: walk the nodes in reverse post-order (= id order),
// a stack of pending liveness flags in one register (depth checked on the host).
ISD_HDI void write_live_mask(const SpecDesc& spec, const uint32_t* decisions, uint32_t* __restrict__ out) {
    unsigned long long pending = 1ull;       // the root is live; top of the stack in bit 0
    int opj = spec.num_ops - 1;
    uint32_t word = 0u, topo_word = 0u;
    for (int i = 0; i < spec.num_nodes; ++i) {
        if ((i & 31) == 0) topo_word = isd_ldg(spec.topo + (i >> 5));
        const unsigned long long alive = pending & 1ull;
        pending >>= 1;
        word |= (uint32_t)alive << (i & 31);
        if ((topo_word >> (i & 31)) & 1u) {
            const uint32_t d = (decisions[opj >> 4] >> ((opj & 15) * 2)) & 3u;
            --opj;
            const unsigned long long la = (alive && d != KILL_A) ? 1ull : 0ull, lb = (alive && d != KILL_B) ? 1ull : 0ull;
            pending = (pending << 2) | (la << 1) | lb;      // b's subtree comes first in reverse post-order
        }
        if ((i & 31) == 31 || i + 1 == spec.num_nodes) { out[i >> 5] = word; word = 0u; }
    }
}

}  // namespace isd
