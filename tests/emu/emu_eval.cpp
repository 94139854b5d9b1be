// TEST FIXTURE (CPU): compiles the product's shape compiler (program.hpp) and the __host__ __device__ program
// evaluator (csg_eval.cuh) for the host, so the flattening and the interpreter arithmetic can be checked against the
// oracle without a GPU.  Built by tests/conftest.py into tests/emu/libemu.so; never linked into the product.
#include <cstring>
#include <vector>
#include "../../implisolid_b200/csrc/csg_eval.cuh"
#include "../../implisolid_b200/csrc/slab_balance.hpp"

using namespace isd;

struct HostProg {
    const uint32_t* w;
    W4 load4(int i) const { return W4{w[i], w[i + 1], w[i + 2], w[i + 3]}; }
};

extern "C" {

// returns number of program words; fills stats[0..3] = value depth, frame depth, nodes, flops/point
int emu_compile(const char* shape_json, int ignore_root_matrix, uint32_t* words_out, int cap, double* stats) {
    ShapeOptions o;
    ShapeCompiler c(o);
    auto tree = c.parse(shape_json, ignore_root_matrix != 0);
    ShapeCompiler::Program p = c.compile(*tree);
    if (stats) { stats[0] = p.value_depth; stats[1] = p.frame_depth; stats[2] = p.num_nodes; stats[3] = p.flops_per_point; }
    int n = (int)p.words.size();
    for (int i = 0; i < n && i < cap; i++) words_out[i] = p.words[i];
    return n;
}

void emu_eval(const char* shape_json, int ignore_root_matrix, const float* pts, int n, float* f_out, float* f4_out, float* g_out) {
    ShapeOptions o;
    ShapeCompiler c(o);
    auto tree = c.parse(shape_json, ignore_root_matrix != 0);
    ShapeCompiler::Program p = c.compile(*tree);
    HostProg prog{p.words.data()};
    SdfGridView sdf{nullptr, 0, 0, 0, 0, 0, 0, 0};
    for (int i = 0; i < n; i++) {
        float f, g0, g1, g2;
        eval_value_gradient(prog, sdf, pts[i * 3], pts[i * 3 + 1], pts[i * 3 + 2], f, g0, g1, g2);
        f_out[i] = f;
        g_out[i * 3] = g0; g_out[i * 3 + 1] = g1; g_out[i * 3 + 2] = g2;
    }
    // value mode, 4 points per call like the field kernel
    for (int i = 0; i < n; i += 4) {
        float x[4], y[4], z[4], out[4];
        for (int k = 0; k < 4; k++) {
            int j = (i + k < n) ? i + k : n - 1;
            x[k] = pts[j * 3]; y[k] = pts[j * 3 + 1]; z[k] = pts[j * 3 + 2];
        }
        eval_values<4>(prog, sdf, AllLive(), x, y, z, out);
        for (int k = 0; k < 4 && i + k < n; k++) f4_out[i + k] = out[k];
    }
}

void emu_affine_inverse(const float* in12, float* out12) { affine_inverse(in12, out12); }

}

// ---- tile specialisation: the interpreter's handling of dead nodes ------------------------------------------------------
// Recursive evaluation of the tree (same primitives, nested transforms) that also records, per point, a liveness mask in
// which every child NOT selected by its parent's min/max is dead.  eval_values with that mask must return the same bits
// as the full evaluation: this is the host check of skip offsets, the `present` bit stack and the lone-operand rules.
namespace {
struct MaskLive {
    const uint32_t* m;
    bool operator()(uint32_t id) const { return (m[id >> 5] >> (id & 31)) & 1u; }
};

float eval_tree(const ShapeCompiler& c, const ShapeNode& n, float x, float y, float z, bool alive, uint32_t* mask) {
    if (alive) { uint32_t id = c.id_of(n); mask[id >> 5] |= 1u << (id & 31); }
    const float* m = n.minv;
    const float lx = m[0] * x + m[1] * y + m[2] * z + m[3];
    const float ly = m[4] * x + m[5] * y + m[6] * z + m[7];
    const float lz = m[8] * x + m[9] * y + m[10] * z + m[11];
    SdfGridView sdf{nullptr, 0, 0, 0, 0, 0, 0, 0};
    switch (n.op) {
        case OP_ELLIPSOID: return prim_ellipsoid(lx, ly, lz);
        case OP_CYLINDER: return prim_cylinder(lx, ly, lz);
        case OP_CONE: return prim_cone(lx, ly, lz);
        case OP_CUBE: return prim_cube(lx, ly, lz);
        case OP_CUBE_SDF: return prim_cube_sdf(sdf, lx, ly, lz);
        case OP_METABALLS: {
            float f = 0.0f;
            for (size_t b = 0; b + 4 < n.params.size(); b += 5) {
                const float fx = lx - n.params[b], fy = ly - n.params[b + 1], fz = lz - n.params[b + 2];
                const float h = 0.000001f + fx * fx + fy * fy + fz * fz;
                const float hinv = 1.0f / h;
                const float val = n.params[b + 3] * hinv - n.params[b + 4];
                f += val / 100.0f;
            }
            return f;
        }
        case OP_TORUS: return prim_torus(lx, ly, lz);
        case OP_HEART: return prim_heart(lx, ly, lz);
        case OP_MUSHROOM: return prim_mushroom(lx, ly, lz);
        case OP_LID: return prim_lid(z);
        case OP_TETRA: return prim_tetra(HostProg{reinterpret_cast<const uint32_t*>(n.params.data())}, 0, x, y, z);
        case OP_HALF_PLANE: return prim_half_plane(HostProg{reinterpret_cast<const uint32_t*>(n.params.data())}, 0, lx, ly, lz);
        default: break;
    }
    // CSG: evaluate both children unmarked first to learn the selection, then mark the selected one
    uint32_t scratch[kMaxSpecNodes / 32 + 1];
    const float f1 = eval_tree(c, *n.a, lx, ly, lz, false, scratch);
    const float f2 = eval_tree(c, *n.b, lx, ly, lz, false, scratch);
    bool first;
    float r;
    if (n.op == OP_UNION) { first = f1 > f2; r = first ? f1 : f2; }
    else if (n.op == OP_INTERSECT) { first = !(f1 > f2); r = first ? f1 : f2; }
    else { first = f1 < -f2; r = first ? f1 : -f2; }
    if (alive) eval_tree(c, first ? *n.a : *n.b, lx, ly, lz, true, mask);
    return r;
}
}  // namespace

extern "C" int emu_eval_masked(const char* shape_json, const float* pts, int n, float* f_tree, float* f_masked, int* live_nodes) {
    ShapeOptions o;
    ShapeCompiler c(o);
    auto tree = c.parse(shape_json, false);
    ShapeCompiler::Program p = c.compile(*tree);
    HostProg prog{p.words.data()};
    SdfGridView sdf{nullptr, 0, 0, 0, 0, 0, 0, 0};
    for (int i = 0; i < n; i++) {
        uint32_t mask[kMaxSpecNodes / 32 + 1];
        memset(mask, 0, sizeof(mask));
        f_tree[i] = eval_tree(c, *tree, pts[i * 3], pts[i * 3 + 1], pts[i * 3 + 2], true, mask);
        float x[1] = {pts[i * 3]}, y[1] = {pts[i * 3 + 1]}, z[1] = {pts[i * 3 + 2]}, out[1];
        eval_values<1>(prog, sdf, MaskLive{mask}, x, y, z, out);
        f_masked[i] = out[0];
        // the same through the per-tile instruction list (what k_eval_tiles runs); any disagreement poisons the result
        {
            std::vector<uint32_t> list;
            for (size_t k = 0; k + 3 < p.itab.size(); k += 4) {
                const uint32_t e = spec_entry(W4{p.itab[k], p.itab[k + 1], p.itab[k + 2], p.itab[k + 3]}, MaskLive{mask});
                if (e) list.push_back(e);
            }
            float x2[1] = {pts[i * 3]}, y2[1] = {pts[i * 3 + 1]}, z2[1] = {pts[i * 3 + 2]}, out2[1];
            eval_values_list<1>(prog, sdf, list.data(), (int)list.size(), x2, y2, z2, out2);
            if (memcmp(out2, out, 4) != 0) f_masked[i] = __builtin_nanf("");
        }
        // gradient mode with the same mask (what the OB02 kernels run near the surface): value and gradient must keep their bits
        {
            float fa, ga[3], fb, gb[3];
            eval_value_gradient(prog, sdf, pts[i * 3], pts[i * 3 + 1], pts[i * 3 + 2], fa, ga[0], ga[1], ga[2]);
            eval_value_gradient(prog, sdf, MaskLive{mask}, pts[i * 3], pts[i * 3 + 1], pts[i * 3 + 2], fb, gb[0], gb[1], gb[2]);
            if (memcmp(&fa, &fb, 4) != 0 || memcmp(ga, gb, 12) != 0) f_masked[i] = __builtin_nanf("");
            // and through the per-tile gradient list (what the OB02 kernels run from the lists k_build_point_lists leaves);
            // the value interpreter must also accept that list (its LE_PASS entries do nothing there)
            std::vector<uint32_t> glist;
            for (size_t k = 0; k + 3 < p.itab.size(); k += 4) {
                const uint32_t e = spec_entry_gradient(W4{p.itab[k], p.itab[k + 1], p.itab[k + 2], p.itab[k + 3]}, MaskLive{mask});
                if (e) glist.push_back(e);
            }
            float fc, gc[3];
            eval_value_gradient_list(prog, sdf, glist.data(), (int)glist.size(), pts[i * 3], pts[i * 3 + 1], pts[i * 3 + 2], fc, gc[0], gc[1], gc[2]);
            if (memcmp(&fa, &fc, 4) != 0 || memcmp(ga, gc, 12) != 0) f_masked[i] = __builtin_nanf("");
            float x3[1] = {pts[i * 3]}, y3[1] = {pts[i * 3 + 1]}, z3[1] = {pts[i * 3 + 2]}, out3[1];
            eval_values_list<1>(prog, sdf, glist.data(), (int)glist.size(), x3, y3, z3, out3);
            if (memcmp(out3, out, 4) != 0) f_masked[i] = __builtin_nanf("");
        }
        int cnt = 0;
        for (int k = 0; k < p.num_nodes; k++) cnt += (mask[k >> 5] >> (k & 31)) & 1;
        live_nodes[i] = cnt;
    }
    return p.can_specialize ? p.num_nodes : -p.num_nodes;
}

// ---- interval bounds, pruning decisions and liveness masks (csg_bounds.cuh) ----------------------------------------------
// For every box: bounds of the tree over the box, the liveness mask made from the decisions, then a second level the way K1
// does it (a sub-box bounded under the first mask).  At random points of the box the exact value must lie inside the bounds,
// and every interpreter run under the mask (walk, value list, gradient list) must return the bits of the full evaluation;
// the second-level mask must never be looser than the first.
// out[0] bounds violations, out[1] value mismatches under a mask, out[2] gradient mismatches, out[3] second-level masks with a
// bit the first level lacks, out[4] sum of live nodes over the boxes, out[5] boxes proven entirely inside or outside.
#include "../../implisolid_b200/csrc/csg_bounds.cuh"

extern "C" int emu_bounds_check(const char* shape_json, const float* boxes, int n_boxes, int samples, unsigned seed, long long* out) {
    ShapeOptions o;
    ShapeCompiler c(o);
    auto tree = c.parse(shape_json, false);
    ShapeCompiler::Program p = c.compile(*tree);
    for (int k = 0; k < 6; k++) out[k] = 0;
    if (!p.can_specialize) return -p.num_nodes;
    HostProg prog{p.words.data()};
    SdfGridView sdf{nullptr, 0, 0, 0, 0, 0, 0, 0};
    SpecDesc spec{};
    spec.enabled = 1; spec.num_nodes = p.num_nodes; spec.num_ops = (p.num_nodes - 1) / 2; spec.mask_words = (p.num_nodes + 31) / 32;
    spec.topo = p.topo.data();
    unsigned long long rng = 0x9E3779B97F4A7C15ull ^ seed;
    auto uni = [&rng]() {      // xorshift64*, uniform in [0, 1)
        rng ^= rng >> 12; rng ^= rng << 25; rng ^= rng >> 27;
        return (float)(((rng * 0x2545F4914F6CDD1Dull) >> 40) * (1.0 / 16777216.0));
    };
    std::vector<uint32_t> mask(spec.mask_words), mask2(spec.mask_words);
    uint32_t decisions[kDecisionWords];
    auto check_points = [&](const float* bx, const Ival& b, const std::vector<uint32_t>& m, bool check_bounds) {
        std::vector<uint32_t> vlist, glist;
        for (size_t k = 0; k + 3 < p.itab.size(); k += 4) {
            const W4 it{p.itab[k], p.itab[k + 1], p.itab[k + 2], p.itab[k + 3]};
            const uint32_t ev = spec_entry(it, MaskLive{m.data()}), eg = spec_entry_gradient(it, MaskLive{m.data()});
            if (ev) vlist.push_back(ev);
            if (eg) glist.push_back(eg);
        }
        for (int s = 0; s < samples; s++) {
            const float x = bx[0] + (2.0f * uni() - 1.0f) * bx[3], y = bx[1] + (2.0f * uni() - 1.0f) * bx[4], z = bx[2] + (2.0f * uni() - 1.0f) * bx[5];
            float fa, ga[3];
            eval_value_gradient(prog, sdf, x, y, z, fa, ga[0], ga[1], ga[2]);
            if (check_bounds && !(fa >= b.lo && fa <= b.hi)) out[0]++;
            float X[1] = {x}, Y[1] = {y}, Z[1] = {z}, fm[1];
            eval_values<1>(prog, sdf, MaskLive{m.data()}, X, Y, Z, fm);
            float X2[1] = {x}, Y2[1] = {y}, Z2[1] = {z}, fl[1];
            eval_values_list<1>(prog, sdf, vlist.data(), (int)vlist.size(), X2, Y2, Z2, fl);
            if (memcmp(&fa, fm, 4) != 0 || memcmp(&fa, fl, 4) != 0) out[1]++;
            float fb, gb[3], fc, gc[3];
            eval_value_gradient(prog, sdf, MaskLive{m.data()}, x, y, z, fb, gb[0], gb[1], gb[2]);
            eval_value_gradient_list(prog, sdf, glist.data(), (int)glist.size(), x, y, z, fc, gc[0], gc[1], gc[2]);
            if (memcmp(&fa, &fb, 4) != 0 || memcmp(ga, gb, 12) != 0 || memcmp(&fa, &fc, 4) != 0 || memcmp(ga, gc, 12) != 0) out[2]++;
        }
    };
    for (int i = 0; i < n_boxes; i++) {
        const float* bx = boxes + i * 6;
        const Ival b = program_bounds<true, false>(p.bounds.data(), bx[0], bx[1], bx[2], bx[3], bx[4], bx[5], decisions, NoMask());
        write_live_mask(spec, decisions, mask.data());
        for (int k = 0; k < p.num_nodes; k++) out[4] += (mask[k >> 5] >> (k & 31)) & 1;
        if (b.hi < 0.0f || b.lo > 0.0f) out[5]++;
        check_points(bx, b, mask, true);
        // second level: a sub-box (one octant, chosen at random) bounded under the first level's mask
        float sub[6];
        for (int d = 0; d < 3; d++) { sub[3 + d] = 0.5f * bx[3 + d]; sub[d] = bx[d] + (uni() < 0.5f ? -0.5f : 0.5f) * bx[3 + d]; }
        const Ival b2 = program_bounds<true, true>(p.bounds.data(), sub[0], sub[1], sub[2], sub[3], sub[4], sub[5], decisions, WordMask{mask.data()});
        write_live_mask(spec, decisions, mask2.data());
        for (int w = 0; w < spec.mask_words; w++) if (mask2[w] & ~mask[w]) out[3]++;
        check_points(sub, b2, mask2, true);
    }
    return p.num_nodes;
}

// ---- packed-pair interpreters (csg_eval.cuh F2): two points through one instruction list -------------------------------------
// Points are taken four at a time; the lists come from the liveness mask interval bounds give for the box that holds the four
// (what K1 does per tile, csg_bounds.cuh).  Packed value (1 and 2 pairs) and packed gradient results must carry the bits of the
// scalar list interpreters and of the full tree.  Returns the number of mismatching results.
extern "C" int emu_eval_packed(const char* shape_json, const float* pts, int n) {
    ShapeOptions o;
    ShapeCompiler c(o);
    auto tree = c.parse(shape_json, false);
    ShapeCompiler::Program p = c.compile(*tree);
    if (!p.can_specialize) return -1;
    HostProg prog{p.words.data()};
    SdfGridView sdf{nullptr, 0, 0, 0, 0, 0, 0, 0};
    SpecDesc spec{};
    spec.enabled = 1; spec.num_nodes = p.num_nodes; spec.num_ops = (p.num_nodes - 1) / 2; spec.mask_words = (p.num_nodes + 31) / 32;
    spec.topo = p.topo.data();
    uint32_t decisions[kDecisionWords];
    int bad = 0;
    for (int i = 0; i + 3 < n; i += 4) {
        uint32_t mask[kMaxSpecNodes / 32 + 1], scratch[kMaxSpecNodes / 32 + 1];
        memset(mask, 0, sizeof(mask));
        float ft[4], lo[3] = {1e30f, 1e30f, 1e30f}, hi[3] = {-1e30f, -1e30f, -1e30f};
        for (int j = 0; j < 4; j++) {
            ft[j] = eval_tree(c, *tree, pts[(i + j) * 3], pts[(i + j) * 3 + 1], pts[(i + j) * 3 + 2], false, scratch);
            for (int d = 0; d < 3; d++) { lo[d] = std::min(lo[d], pts[(i + j) * 3 + d]); hi[d] = std::max(hi[d], pts[(i + j) * 3 + d]); }
        }
        program_bounds<true, false>(p.bounds.data(), 0.5f * (lo[0] + hi[0]), 0.5f * (lo[1] + hi[1]), 0.5f * (lo[2] + hi[2]), 0.5f * (hi[0] - lo[0]) * 1.001f + 1e-6f,
                                    0.5f * (hi[1] - lo[1]) * 1.001f + 1e-6f, 0.5f * (hi[2] - lo[2]) * 1.001f + 1e-6f, decisions, NoMask());
        write_live_mask(spec, decisions, mask);
        std::vector<uint32_t> vlist, glist;
        for (size_t k = 0; k + 3 < p.itab.size(); k += 4) {
            const W4 it{p.itab[k], p.itab[k + 1], p.itab[k + 2], p.itab[k + 3]};
            const uint32_t ev = spec_entry(it, MaskLive{mask}), eg = spec_entry_gradient(it, MaskLive{mask});
            if (ev) vlist.push_back(ev);
            if (eg) glist.push_back(eg);
        }
        const float* q = pts + i * 3;
        // scalar references under the same lists
        float fs[4], gs[4][3];
        for (int j = 0; j < 4; j++) {
            float x[1] = {q[j * 3]}, y[1] = {q[j * 3 + 1]}, z[1] = {q[j * 3 + 2]}, out[1];
            eval_values_list<1>(prog, sdf, vlist.data(), (int)vlist.size(), x, y, z, out);
            fs[j] = out[0];
            float fg;
            eval_value_gradient_list(prog, sdf, glist.data(), (int)glist.size(), q[j * 3], q[j * 3 + 1], q[j * 3 + 2], fg, gs[j][0], gs[j][1], gs[j][2]);
            if (memcmp(&fg, &fs[j], 4) != 0 || memcmp(&fs[j], &ft[j], 4) != 0) bad++;
        }
        {   // two pairs at once (what k_eval_tiles runs)
            F2 x[2] = {{q[0], q[3]}, {q[6], q[9]}}, y[2] = {{q[1], q[4]}, {q[7], q[10]}}, z[2] = {{q[2], q[5]}, {q[8], q[11]}}, out[2];
            eval_values_list2<2>(prog, sdf, vlist.data(), (int)vlist.size(), x, y, z, out);
            const float got[4] = {out[0].x, out[0].y, out[1].x, out[1].y};
            if (memcmp(got, fs, 16) != 0) bad++;
        }
        for (int h = 0; h < 2; h++) {   // one pair: value and gradient (what the OB02 kernels run)
            const float* a = q + h * 6;
            F2 x[1] = {{a[0], a[3]}}, y[1] = {{a[1], a[4]}}, z[1] = {{a[2], a[5]}}, out[1];
            eval_values_list2<1>(prog, sdf, vlist.data(), (int)vlist.size(), x, y, z, out);
            if (memcmp(&out[0].x, &fs[h * 2], 4) != 0 || memcmp(&out[0].y, &fs[h * 2 + 1], 4) != 0) bad++;
            F2 f, g0, g1, g2;
            eval_value_gradient_list2(prog, sdf, glist.data(), (int)glist.size(), F2{a[0], a[3]}, F2{a[1], a[4]}, F2{a[2], a[5]}, f, g0, g1, g2);
            const float ga[3] = {g0.x, g1.x, g2.x}, gb[3] = {g0.y, g1.y, g2.y};
            if (memcmp(&f.x, &fs[h * 2], 4) != 0 || memcmp(&f.y, &fs[h * 2 + 1], 4) != 0) bad++;
            if (memcmp(ga, gs[h * 2], 12) != 0 || memcmp(gb, gs[h * 2 + 1], 12) != 0) bad++;
            // the value interpreter must also accept the gradient list
            F2 x3[1] = {{a[0], a[3]}}, y3[1] = {{a[1], a[4]}}, z3[1] = {{a[2], a[5]}}, out3[1];
            eval_values_list2<1>(prog, sdf, glist.data(), (int)glist.size(), x3, y3, z3, out3);
            if (memcmp(&out3[0].x, &fs[h * 2], 4) != 0 || memcmp(&out3[0].y, &fs[h * 2 + 1], 4) != 0) bad++;
        }
    }
    return bad;
}


// z-slab balance (slab_balance.hpp, the code geometry_api.cu runs): first cuts from a tile histogram, then `rounds` feedback steps
// against a synthetic cost model — rank r's busy time = fixed_us + the sum of cost_us[l] over its layers.  cuts_out receives
// (rounds + 1) x (world + 1) entries (the cuts before every round and the final ones), busy_out (rounds + 1) x world times.
// Returns the number of re-cuts made.
extern "C" int emu_slab_balance(const unsigned* hist, int bins, int layers, int world, const long long* cost_us, long long fixed_us, int rounds,
                                int threshold_permille, int* cuts_out, long long* busy_out) {
    SlabBalance b;
    b.init(hist, bins, layers, world);
    std::vector<long long> busy((size_t)world);
    for (int it = 0; it <= rounds; ++it) {
        for (int r = 0; r < world; ++r) {
            long long t = fixed_us;
            for (int l = b.cuts[(size_t)r]; l < b.cuts[(size_t)r + 1]; ++l) t += cost_us[l];
            busy[(size_t)r] = t;
            busy_out[(size_t)it * world + r] = t;
        }
        for (int r = 0; r <= world; ++r) cuts_out[(size_t)it * (world + 1) + r] = b.cuts[(size_t)r];
        if (it < rounds) b.feedback(busy.data(), world, threshold_permille);
    }
    return b.recuts;
}
