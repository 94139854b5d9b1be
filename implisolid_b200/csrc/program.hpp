// Shape spec (JSON) -> CSG tree -> flat device bytecode ("program").
//
// Host-side replacement for the reference's object_factory (object_factory.hpp:56-758) and the
// per-node matrix inversion done in every implicit_function constructor
// (implicit_function/transformed.hpp:16-26, egg.hpp:38-57, ... -> invert_matrix,
// basic_functions.hpp:77-128).  The reference nests transforms (every node applies its OWN inverse
// matrix to the points handed down by its parent, basic_functions.hpp:361-379) and never composes
// them, so the program carries one 3x4 inverse per node and the device applies them in the same
// order — required for bit-identical field values.
//
// Program = array of 32-bit words in global memory, every instruction aligned to 16 bytes so that the interpreter
// fetches it with 128-bit loads (one for the header, three for a 3x4 matrix):
//   header = {opcode | flags | node id << 12, skip, aux, 0}
//   XFORM(+SAVE flag) hdr m[12]     entering a CSG node: (push cur on the frame stack when the node is a CSG child that is
//                                   not evaluated last), cur <- m * cur
//   ELLIPSOID|CYLINDER|CONE|CUBE|CUBE_SDF hdr m[12]     push f(m * cur)   (primitives apply their own inverse)
//   METABALLS hdr(aux = nb) m[12] {bx,by,bz,strength, subtract,0,0,0}*nb
//   TORUS|HEART|MUSHROOM|LID hdr m[12]                  (LID ignores m: top_bottom_lid never transforms its points)
//   TETRA hdr m[12] planes[4][4]                        (m unused: the corners were moved by the forward matrix on the host)
//   HALF_PLANE hdr m[12] {n0,n1,n2,0, p0,p1,p2,0}       (m = cofactor inverse, see cofactor_inverse)
//   UNION|INTERSECT|SUBTRACT (+SWAPPED, +RESTORE flags) hdr m[12]   pop two values, push the selected one; m is the node's
//                                   own inverse, used only in gradient mode to map the selected child's gradient back;
//                                   RESTORE: then pop cur (closes a node entered with the SAVE flag)
//   END hdr
// `skip` is the program counter just past the node's whole code; the opcode word of a node-opening instruction carries
// the node id in bits 12..31.  Both exist for the tile-specialised field kernel (field_kernel.cuh): per grid tile, interval
// bounds prove which children of min/max nodes can never be selected, and the evaluator jumps over those nodes.  A node
// id is the node's index in REVERSE post-order (root = 0), the order in which the liveness pass produces its bits.
// Children are ordered so the deeper child is evaluated last (no frame to keep while descending), which keeps the
// frame stack empty and the value stack at one float per level for the left-deep chains the factory builds.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <vector>
#include "json_min.hpp"

namespace isd {

enum Opcode : uint32_t {
    OP_END = 0,
    OP_XFORM = 1,
    OP_ELLIPSOID = 4,
    OP_CYLINDER = 5,
    OP_CONE = 6,
    OP_CUBE = 7,
    OP_CUBE_SDF = 8,
    OP_METABALLS = 9,
    OP_TORUS = 10,               // itorus                    object_factory.hpp:163-173
    OP_HEART = 11,               // iheart                    :153-162
    OP_TETRA = 12,               // tetrahedron               :175-188   params: planes[4][4]; points are NOT transformed
    OP_MUSHROOM = 13,            // implicit_double_mushroom  :86-99     linearly_transformed(double_mushroom(0.9, 0.2, 0.2, 5))
    OP_HALF_PLANE = 14,          // half_plane                :396-434   params: {n0,n1,n2,0, p0,p1,p2,0}
    OP_LID = 15,                 // top_bottom_lid            :480-505   points are NOT transformed
    OP_UNION = 24,
    OP_INTERSECT = 25,
    OP_SUBTRACT = 26,
    OP_MASK = 0xff,
    OP_FLAG_SWAPPED = 0x100,     // value stack holds (second child, first child)
    OP_FLAG_SAVE = 0x200,        // XFORM: push cur on the frame stack first
    OP_FLAG_RESTORE = 0x400,     // closing CSG instruction: pop cur from the frame stack afterwards
    OP_ID_SHIFT = 12,
};

constexpr int kMaxValueStack = 64;     // device limits (csg_eval.cuh)
constexpr int kMaxFrameStack = 16;
constexpr int kMaxProgramWords = 32768;   // 128 KB of global memory (a __device__ array, field_kernel.cuh)
constexpr int kInstrWords = 16;           // header (4 words) + 3x4 matrix
constexpr int kMaxListEntries = 1024;     // tile specialisation: longest per-tile instruction list kept in shared memory

// Instruction table entry (one uint4 per program instruction, END excluded) and per-tile list entry, tile specialisation:
//   itab.x = pc / 4 | IT_* flags, itab.y = id of the instruction's node, itab.z / itab.w = ids of a CSG node's children a / b
//   list entry = pc / 4 | LE_* flags: what the instruction does in THIS tile (see spec_entry, csg_eval.cuh)
enum : uint32_t {
    IT_PC_MASK = 0xffffu,
    IT_CLOSING = 1u << 16, IT_RESTORE = 1u << 17, IT_SWAPPED = 1u << 18, IT_OP_SHIFT = 19,      // op - OP_UNION in 2 bits
    LE_CLOSING = 1u << 16, LE_RESTORE = 1u << 17, LE_SWAPPED = 1u << 18, LE_OP_SHIFT = 19,
    LE_COMBINE = 1u << 21, LE_NEGATE = 1u << 22, LE_VALID = 1u << 31,
    LE_PASS = 1u << 23,       // gradient lists only: live closing instruction with one operand and nothing to do in value mode
    LE_OPC_SHIFT = 24, LE_OPC_MASK = 31u, LE_SAVE = 1u << 29,      // opening instructions (in itab.x too): the opcode and XFORM's SAVE flag
};
constexpr int kMaxSpecNodes = 2048;       // tile specialisation: per-tile liveness mask of at most this many bits
constexpr int kMaxPendingDepth = 64;      // liveness pass: pending-sibling stack held in one 64-bit register

// ---- 3x4 affine inverse -------------------------------------------------------------------------------------
// Restates what invert_matrix gets from uBLAS lu_factorize + lu_substitute on the 4x4 [M; 0 0 0 1] in binary32:
// partial pivoting by largest |a_ki| (first maximum), column scaled by the reciprocal of the pivot, rank-1 trailing
// update, then row swaps on the identity, unit-lower forward and upper backward substitution, column by column.
// Boost is not available here and the reference pins no version, so this order defines the result.
// Returns false on a zero pivot, leaving `inv` untouched like the reference.
inline bool affine_inverse(const float m12[12], float inv[12]) {
    float lu[16];
    memcpy(lu, m12, 12 * sizeof(float));
    lu[12] = 0.f; lu[13] = 0.f; lu[14] = 0.f; lu[15] = 1.f;
    int swap_with[4];
    for (int c = 0; c < 4; ++c) {
        int piv = c;
        float big = fabsf(lu[c * 4 + c]);
        for (int r = c + 1; r < 4; ++r) {
            float v = fabsf(lu[r * 4 + c]);
            if (v > big) { big = v; piv = r; }
        }
        swap_with[c] = piv;
        if (lu[piv * 4 + c] == 0.0f) return false;
        if (piv != c)
            for (int k = 0; k < 4; ++k) { float t = lu[c * 4 + k]; lu[c * 4 + k] = lu[piv * 4 + k]; lu[piv * 4 + k] = t; }
        const float recip = 1.0f / lu[c * 4 + c];
        for (int r = c + 1; r < 4; ++r) lu[r * 4 + c] *= recip;
        for (int r = c + 1; r < 4; ++r)
            for (int k = c + 1; k < 4; ++k) lu[r * 4 + k] -= lu[r * 4 + c] * lu[c * 4 + k];
    }
    float x[16];
    for (int i = 0; i < 16; ++i) x[i] = (i % 5 == 0) ? 1.0f : 0.0f;
    for (int c = 0; c < 4; ++c)
        if (swap_with[c] != c)
            for (int k = 0; k < 4; ++k) { float t = x[c * 4 + k]; x[c * 4 + k] = x[swap_with[c] * 4 + k]; x[swap_with[c] * 4 + k] = t; }
    for (int n = 0; n < 4; ++n)                 // L y = P b   (unit diagonal)
        for (int col = 0; col < 4; ++col) {
            const float t = x[n * 4 + col];
            if (t != 0.0f)
                for (int r = n + 1; r < 4; ++r) x[r * 4 + col] -= lu[r * 4 + n] * t;
        }
    for (int n = 3; n >= 0; --n)                // U x = y
        for (int col = 3; col >= 0; --col) {
            const float t = (x[n * 4 + col] /= lu[n * 4 + n]);
            if (t != 0.0f)
                for (int r = n - 1; r >= 0; --r) x[r * 4 + col] -= lu[r * 4 + n] * t;
        }
    memcpy(inv, x, 12 * sizeof(float));
    return true;
}

// ---- tree -----------------------------------------------------------------------------------------------------
struct ShapeNode {
    Opcode op = OP_END;
    float m[12], minv[12];
    std::vector<float> params;                       // metaballs: nb*5
    std::unique_ptr<ShapeNode> a, b;                 // CSG children (reference order)
    bool is_csg() const { return op >= OP_UNION; }
    int depth() const { return is_csg() ? 1 + std::max(a->depth(), b->depth()) : 0; }
};

struct ShapeOptions {
    int cube_mode = 0;
};

class ShapeCompiler {
public:
    explicit ShapeCompiler(const ShapeOptions& o) : opt_(o) {}

    // object_factory(const char*, bool)  object_factory.hpp:741-758
    std::unique_ptr<ShapeNode> parse(const char* shape_json, bool ignore_root_matrix) {
        JsonDoc doc(shape_json);
        return build(doc, doc.root(), ignore_root_matrix);
    }

    struct Program {
        std::vector<uint32_t> words;
        // Bounds program (tile pruning, field_kernel.cuh k_tile_bounds): the same tree in plain post-order, every primitive
        // carrying the COMPOSITE grid->local matrix (product of the nested inverses along its path, formed in double).
        // Only used to prove a tile is entirely inside or outside; never to produce field values.
        std::vector<uint32_t> bounds;
        // Tree topology for the liveness pass: bit i = node with id i (reverse post-order) is a CSG node.
        std::vector<uint32_t> topo;
        // Instruction table (4 words per instruction, program order): lets a thread block decide in parallel which
        // instructions a tile executes.
        std::vector<uint32_t> itab;
        bool can_specialize = false;         // nodes <= kMaxSpecNodes and pending depth <= kMaxPendingDepth
        int value_depth = 0, frame_depth = 0, num_nodes = 0;
        bool uses_sdf_grid = false;
        double flops_per_point = 0;          // mul and add counted separately
    };

    Program compile(const ShapeNode& root) {
        Program p;
        vs_ = fs_ = 0;
        ids_.clear();
        int total = count_nodes(root), po = 0;
        assign_ids(root, total, po);
        emit_node(root, /*preserve=*/false, p);
        put_header(p, OP_END, 0u);
        {
            double eye[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
            emit_bounds(root, eye, p);
            p.bounds.push_back(OP_END);
        }
        build_topology(root, total, p);
        if (p.value_depth > kMaxValueStack || p.frame_depth > kMaxFrameStack) {
            fprintf(stderr, "implisolid_b200: shape tree too deep for the device evaluator (value stack %d > %d or frame stack %d > %d)\n",
                    p.value_depth, kMaxValueStack, p.frame_depth, kMaxFrameStack);
            abort();
        }
        if ((int)p.words.size() > kMaxProgramWords) {
            fprintf(stderr, "implisolid_b200: shape program too large (%zu words > %d)\n", p.words.size(), kMaxProgramWords);
            abort();
        }
        return p;
    }

    // id of a node of the tree last compiled (reverse post-order index, root = 0)
    uint32_t id_of(const ShapeNode& n) const {
        for (const auto& e : ids_) if (e.first == &n) return (uint32_t)e.second;
        return 0;
    }

private:
    ShapeOptions opt_;
    int vs_ = 0, fs_ = 0;
    std::vector<std::pair<const ShapeNode*, int>> ids_;      // node -> id (reverse post-order index)

    static int count_nodes(const ShapeNode& n) { return n.is_csg() ? 1 + count_nodes(*n.a) + count_nodes(*n.b) : 1; }
    void assign_ids(const ShapeNode& n, int total, int& po) {
        if (n.is_csg()) { assign_ids(*n.a, total, po); assign_ids(*n.b, total, po); }
        ids_.push_back({&n, total - 1 - po});
        ++po;
    }
    void build_topology(const ShapeNode& root, int total, Program& p) {
        p.topo.assign((size_t)(total + 31) / 32, 0u);
        for (const auto& e : ids_)
            if (e.first->is_csg()) p.topo[(size_t)e.second >> 5] |= 1u << (e.second & 31);
        // pending-sibling depth of the liveness pass (walks ids 0..total-1: pop one, a CSG node pushes two)
        int depth = 1, max_depth = 1;
        for (int i = 0; i < total; ++i) {
            --depth;
            if (p.topo[(size_t)i >> 5] >> (i & 31) & 1u) depth += 2;
            if (depth > max_depth) max_depth = depth;
        }
        p.can_specialize = total <= kMaxSpecNodes && max_depth <= kMaxPendingDepth && total < (1 << (32 - OP_ID_SHIFT)) &&
                           p.itab.size() / 4 <= (size_t)kMaxListEntries && p.words.size() / 4 <= (size_t)IT_PC_MASK;
    }

    static void read_matrix12(const JsonDoc& d, int node, bool ignore_root_matrix, float m[12]) {
        // getMatrix12 (object_factory.hpp:21-30) copies every entry it finds (16 -> overruns REAL[12]); we stop at 12.
        int mat = d.child(node, "matrix");
        if (mat < 0) { fprintf(stderr, "implisolid_b200: shape node without \"matrix\"\n"); abort(); }
        int i = 0;
        for (int c = d.first_child(mat); c >= 0 && i < 12; c = d.next(c)) {
            float v;
            if (!d.to_float(c, &v)) { fprintf(stderr, "implisolid_b200: bad matrix entry\n"); abort(); }
            m[i++] = v;
        }
        if (i < 12) { fprintf(stderr, "implisolid_b200: \"matrix\" needs at least 12 entries\n"); abort(); }
        if (ignore_root_matrix) {   // copy_eye, object_factory.hpp:47
            static const float eye[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
            memcpy(m, eye, sizeof(eye));
        }
    }
    static void set_matrix(ShapeNode& n, const float m[12]) {
        memcpy(n.m, m, sizeof(n.m));
        memset(n.minv, 0, sizeof(n.minv));     // the reference leaves heap garbage when the matrix is singular
        if (!affine_inverse(n.m, n.minv)) fprintf(stderr, "implisolid_b200: warning: singular shape matrix\n");
    }
    static void set_identity_inverse(ShapeNode& n) {
        for (int i = 0; i < 12; ++i) n.minv[i] = (i % 5 == 0) ? 1.0f : 0.0f;
    }
    // `rows` rows of three numbers: "corners":[[x,y,z] x4], or one flat triple ("plane_vector":[x,y,z])
    static void read_rows(const JsonDoc& d, int node, const char* key, int rows, float* out) {
        int arr = d.child(node, key);
        if (arr < 0) { fprintf(stderr, "implisolid_b200: shape node without \"%s\"\n", key); abort(); }
        int k = 0;
        if (rows == 1) {
            for (int c = d.first_child(arr); c >= 0 && k < 3; c = d.next(c))
                if (!d.to_float(c, &out[k++])) { fprintf(stderr, "implisolid_b200: bad \"%s\" entry\n", key); abort(); }
        } else {
            for (int r = d.first_child(arr); r >= 0 && k < rows * 3; r = d.next(r)) {
                int j = 0;
                for (int c = d.first_child(r); c >= 0 && j < 3; c = d.next(c), ++j)
                    if (!d.to_float(c, &out[k++])) { fprintf(stderr, "implisolid_b200: bad \"%s\" entry\n", key); abort(); }
                if (j != 3) { fprintf(stderr, "implisolid_b200: a row of \"%s\" needs 3 numbers\n", key); abort(); }
            }
        }
        if (k != rows * 3) { fprintf(stderr, "implisolid_b200: \"%s\" needs %d x 3 numbers\n", key, rows); abort(); }
    }
    // tetrahedron: constructor (tetrahedron.hpp:136-143: corners moved by the FORWARD matrix, basic_functions.hpp:140-177),
    // calculatePlaneCoefficients :18-28, getPlanes :30-134 (each plane oriented so that the opposite corner is on the
    // positive side, sign() with ROOT_TOLERANCE = 0.001/10, basic_functions.hpp:22-31, configs.hpp:33), all binary32.
    static void plane_through(const float* p1, const float* p2, const float* p3, float* pl) {
        const float x1 = p1[0], y1 = p1[1], z1 = p1[2], x2 = p2[0], y2 = p2[1], z2 = p2[2], x3 = p3[0], y3 = p3[1], z3 = p3[2];
        pl[0] = y1 * z2 - y1 * z3 - y2 * z1 + y2 * z3 + y3 * z1 - y3 * z2;
        pl[1] = x1 * z3 - x1 * z2 + x2 * z1 - x2 * z3 - x3 * z1 + x3 * z2;
        pl[2] = x1 * y2 - x1 * y3 - x2 * y1 + x2 * y3 + x3 * y1 - x3 * y2;
        pl[3] = x1 * y3 * z2 - x1 * y2 * z3 + x2 * y1 * z3 - x2 * y3 * z1 - x3 * y1 * z2 + x3 * y2 * z1;
    }
    static void tetra_planes(const float corners[4][3], const float m[12], std::vector<float>& out) {
        float p[4][3];
        for (int i = 0; i < 4; ++i) {
            const float x = corners[i][0], y = corners[i][1], z = corners[i][2];
            p[i][0] = m[0] * x + m[1] * y + m[2] * z + m[3];
            p[i][1] = m[4] * x + m[5] * y + m[6] * z + m[7];
            p[i][2] = m[8] * x + m[9] * y + m[10] * z + m[11];
        }
        float pl[4][4];
        plane_through(p[1], p[2], p[3], pl[0]);
        plane_through(p[0], p[2], p[3], pl[1]);
        plane_through(p[0], p[1], p[3], pl[2]);
        plane_through(p[0], p[1], p[2], pl[3]);
        const float tol = 0.001 / 10.0;
        for (int k = 0; k < 4; ++k) {
            const float v = pl[k][0] * p[k][0] + pl[k][1] * p[k][1] + pl[k][2] * p[k][2] + pl[k][3];
            const float s = (v > +tol) ? (+1) : (v < -tol) ? (-1) : (0.0);
            for (int j = 0; j < 4; ++j) pl[k][j] *= s;
        }
        out.assign(&pl[0][0], &pl[0][0] + 16);
    }
    // half_plane.hpp:100-107 inverts [M; 0 0 0 1] with Eigen's Matrix4f::inverse() (cofactors over the determinant).  Eigen
    // is absent and unpinned: plain binary32 cofactor expansion, so this primitive has tolerance-level parity only.
    static void cofactor_inverse(const float m12[12], float inv12[12]) {
        float a[16];
        memcpy(a, m12, 12 * sizeof(float));
        a[12] = 0.f; a[13] = 0.f; a[14] = 0.f; a[15] = 1.f;
        float cof[16];
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c) {
                float s[9];
                int k = 0;
                for (int i = 0; i < 4; ++i)
                    for (int j = 0; j < 4; ++j)
                        if (i != r && j != c) s[k++] = a[i * 4 + j];
                const float det3 = s[0] * (s[4] * s[8] - s[5] * s[7]) - s[1] * (s[3] * s[8] - s[5] * s[6]) + s[2] * (s[3] * s[7] - s[4] * s[6]);
                cof[r * 4 + c] = ((r + c) & 1) ? -det3 : det3;
            }
        const float det = a[0] * cof[0] + a[1] * cof[1] + a[2] * cof[2] + a[3] * cof[3];
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 4; ++c) inv12[r * 4 + c] = cof[c * 4 + r] / det;
    }
    static std::unique_ptr<ShapeNode> leaf(Opcode op, const float m[12]) {
        std::unique_ptr<ShapeNode> n(new ShapeNode());
        n->op = op;
        set_matrix(*n, m);
        return n;
    }
    static std::unique_ptr<ShapeNode> csg(Opcode op, std::unique_ptr<ShapeNode> a, std::unique_ptr<ShapeNode> b, const float m[12]) {
        std::unique_ptr<ShapeNode> n(new ShapeNode());
        n->op = op;
        set_matrix(*n, m);
        n->a = std::move(a);
        n->b = std::move(b);
        return n;
    }

    std::unique_ptr<ShapeNode> build(const JsonDoc& d, int node, bool ignore_root_matrix) {
        int t = d.child(node, "type");
        std::string name = t >= 0 ? d.text(t) : std::string();
        float m[12];
        if (name == "icube" || name == "cube") {                                      // object_factory.hpp:110-120
            read_matrix12(d, node, ignore_root_matrix, m);
            return leaf(opt_.cube_mode ? OP_CUBE_SDF : OP_CUBE, m);
        }
        if (name == "icylinder" || name == "cylinder") {                              // :121-133
            read_matrix12(d, node, ignore_root_matrix, m);
            return leaf(OP_CYLINDER, m);
        }
        if (name == "iellipsoid" || name == "ellipsoid") {                            // :134-143
            read_matrix12(d, node, ignore_root_matrix, m);
            return leaf(OP_ELLIPSOID, m);
        }
        if (name == "icone" || name == "cone") {                                      // :144-152
            read_matrix12(d, node, ignore_root_matrix, m);
            return leaf(OP_CONE, m);
        }
        if (name == "meta_balls") {                                                   // :654-673
            read_matrix12(d, node, ignore_root_matrix, m);
            float time = 0.1f;
            d.to_float(d.child(node, "time"), &time);
            std::unique_ptr<ShapeNode> n = leaf(OP_METABALLS, m);
            // Extension (SURVEY §8(d) C3 variant (ii), not expressible in the reference, whose factory always builds four blobs
            // from `time`): "balls": [x, y, z, strength, subtract, ...] gives the node's blobs explicitly — a single-node SUM blend
            // of any number of centres, evaluated by the very loop of meta_balls_Rydgard.hpp:89-144.
            const int balls = d.child(node, "balls");
            if (balls >= 0) {
                for (int c = d.first_child(balls); c >= 0; c = d.next(c)) {
                    float v;
                    if (!d.to_float(c, &v)) { fprintf(stderr, "implisolid_b200: bad \"balls\" entry\n"); abort(); }
                    n->params.push_back(v);
                }
                if (n->params.empty() || n->params.size() % 5 != 0 || n->params.size() / 5 > 512) {
                    fprintf(stderr, "implisolid_b200: \"balls\" needs 5 numbers per ball (x, y, z, strength, subtract), 1 to 512 balls\n");
                    abort();
                }
                return n;
            }
            metaball_centres(4, time, n->params);
            return n;
        }
        if (name == "iheart") {                                                       // :153-162
            read_matrix12(d, node, ignore_root_matrix, m);
            return leaf(OP_HEART, m);
        }
        if (name == "itorus") {                                                       // :163-173
            read_matrix12(d, node, ignore_root_matrix, m);
            return leaf(OP_TORUS, m);
        }
        if (name == "implicit_double_mushroom") {                                     // :86-99
            read_matrix12(d, node, ignore_root_matrix, m);
            return leaf(OP_MUSHROOM, m);
        }
        if (name == "top_bottom_lid") {                                               // :480-505 (matrix read, never applied)
            read_matrix12(d, node, ignore_root_matrix, m);
            std::unique_ptr<ShapeNode> n = leaf(OP_LID, m);
            set_identity_inverse(*n);
            return n;
        }
        if (name == "tetrahedron") {                                                  // :175-188, getCorners :32-45
            read_matrix12(d, node, ignore_root_matrix, m);
            float corners[4][3];
            read_rows(d, node, "corners", 4, &corners[0][0]);
            std::unique_ptr<ShapeNode> n = leaf(OP_TETRA, m);
            set_identity_inverse(*n);
            tetra_planes(corners, m, n->params);
            return n;
        }
        if (name == "half_plane") {                                                   // :396-434
            read_matrix12(d, node, ignore_root_matrix, m);
            float nv[3], pp[3];
            read_rows(d, node, "plane_vector", 1, nv);
            read_rows(d, node, "plane_point", 1, pp);
            std::unique_ptr<ShapeNode> n = leaf(OP_HALF_PLANE, m);
            cofactor_inverse(m, n->minv);
            // half_plane.hpp:88-89: the vector is normalised once in the constructor
            const float norm = sqrtf(nv[0] * nv[0] + nv[1] * nv[1] + nv[2] * nv[2]);
            n->params = {nv[0] / norm, nv[1] / norm, nv[2] / norm, 0.0f, pp[0], pp[1], pp[2], 0.0f};
            return n;
        }
        if (name == "Union" || name == "Intersection" || name == "Difference") {
            read_matrix12(d, node, ignore_root_matrix, m);
            int kids = d.child(node, "children");
            int nkids = kids >= 0 ? d.count_children(kids) : 0;
            if (name == "Union") {
                // left fold of binary unions; inner ones carry the identity, the last one the node matrix
                // (object_factory.hpp:537-580).  One child leaves the reference with an uninitialised pointer: rejected.
                if (nkids < 2) { fprintf(stderr, "implisolid_b200: Union needs at least two children\n"); abort(); }
                static const float eye[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
                int c = d.first_child(kids);
                std::unique_ptr<ShapeNode> acc = build(d, c, false);
                int idx = 1;
                for (c = d.next(c); c >= 0; c = d.next(c), ++idx) {
                    std::unique_ptr<ShapeNode> rhs = build(d, c, false);
                    acc = csg(OP_UNION, std::move(acc), std::move(rhs), (idx + 1 == nkids) ? m : eye);
                }
                return acc;
            }
            if (nkids != 2) {   // :581-653 mis-builds with != 2 children
                fprintf(stderr, "implisolid_b200: %s needs exactly two children\n", name.c_str());
                abort();
            }
            int c0 = d.first_child(kids), c1 = d.next(c0);
            std::unique_ptr<ShapeNode> a = build(d, c0, false);
            std::unique_ptr<ShapeNode> b = build(d, c1, false);
            return csg(name == "Intersection" ? OP_INTERSECT : OP_SUBTRACT, std::move(a), std::move(b), m);
        }
        fprintf(stderr, "Invalid object you asked for: \"%s\"\n", name.c_str());      // :731-733
        abort();
    }

    // meta_ball_Rydgard constructor, implicit_function/meta_balls_Rydgard.hpp:49-59 (double math on the host,
    // each result rounded once to binary32)
    static void metaball_centres(int numblobs, float time, std::vector<float>& out) {
        for (int ball_i = 0; ball_i < numblobs; ++ball_i) {
            float x0 = 0.5, y0 = 0.5, z0 = 0.5;
            float D = 1;
            float ballx = sin(ball_i + 1.26 * time * (1.03 + 0.5 * cos(0.21 * ball_i))) * 0.27 * D + 0.5 - x0;
            float bally = std::abs(cos(ball_i + 1.12 * time * cos(1.22 + 0.1424 * ball_i))) * 0.77 * D - y0;
            float ballz = cos(ball_i + 1.32 * time * 0.1 * sin((0.92 + 0.53 * ball_i))) * 0.27 * D + 0.5 - z0;
            float subtract = 12;
            float strength = 1.2 / ((sqrt((double)numblobs) - 1) / 4 + 1);
            out.push_back(ballx); out.push_back(bally); out.push_back(ballz);
            out.push_back(strength); out.push_back(subtract);
        }
    }

    // returns the index of the header's skip word
    static size_t put_header(Program& p, uint32_t word0, uint32_t aux) {
        p.words.push_back(word0);
        p.words.push_back(0u);
        p.words.push_back(aux);
        p.words.push_back(0u);
        return p.words.size() - 3;
    }
    static void put_itab(Program& p, uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
        p.itab.push_back(x); p.itab.push_back(y); p.itab.push_back(z); p.itab.push_back(w);
    }
    static void put_matrix(Program& p, const float m[12]) {
        for (int i = 0; i < 12; ++i) { uint32_t w; memcpy(&w, &m[i], 4); p.words.push_back(w); }
    }
    void push_value(Program& p) { if (++vs_ > p.value_depth) p.value_depth = vs_; }

    // comp = (own inverse) x parent_comp, 4x4 row-major in double
    static void compose(const float minv[12], const double parent[16], double out[16]) {
        double a[16];
        for (int i = 0; i < 12; ++i) a[i] = minv[i];
        a[12] = 0; a[13] = 0; a[14] = 0; a[15] = 1;
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c) {
                double t = 0;
                for (int k = 0; k < 4; ++k) t += a[r * 4 + k] * parent[k * 4 + c];
                out[r * 4 + c] = t;
            }
    }
    void emit_bounds(const ShapeNode& n, const double parent[16], Program& p) {
        double comp[16];
        compose(n.minv, parent, comp);
        if (!n.is_csg()) {
            p.bounds.push_back((uint32_t)n.op | (id_of(n) << OP_ID_SHIFT));
            for (int i = 0; i < 12; ++i) { float f = (float)comp[i]; uint32_t w; memcpy(&w, &f, 4); p.bounds.push_back(w); }
            if (n.op == OP_METABALLS) {
                p.bounds.push_back((uint32_t)(n.params.size() / 5));
                for (float v : n.params) { uint32_t w; memcpy(&w, &v, 4); p.bounds.push_back(w); }
            } else if (n.op == OP_TETRA || n.op == OP_HALF_PLANE) {      // 16 / 8 words, as in the value program
                for (float v : n.params) { uint32_t w; memcpy(&w, &v, 4); p.bounds.push_back(w); }
            }
            return;
        }
        emit_bounds(*n.a, comp, p);
        emit_bounds(*n.b, comp, p);
        p.bounds.push_back((uint32_t)n.op | (id_of(n) << OP_ID_SHIFT));
    }

    void emit_node(const ShapeNode& n, bool preserve, Program& p, bool parent_identity = false) {
        p.num_nodes++;
        if (!n.is_csg()) {
            put_itab(p, (uint32_t)(p.words.size() / 4) | ((uint32_t)n.op << LE_OPC_SHIFT), id_of(n), 0u, 0u);
            const size_t skip_at = put_header(p, (uint32_t)n.op | (id_of(n) << OP_ID_SHIFT), n.op == OP_METABALLS ? (uint32_t)(n.params.size() / 5) : 0u);
            put_matrix(p, n.minv);
            p.flops_per_point += 18;
            switch (n.op) {
                case OP_ELLIPSOID: p.flops_per_point += 9; break;
                case OP_CYLINDER: p.flops_per_point += 14; break;
                case OP_CONE: p.flops_per_point += 14; break;
                case OP_CUBE: p.flops_per_point += 11; break;
                case OP_CUBE_SDF: p.flops_per_point += 40; p.uses_sdf_grid = true; break;
                case OP_METABALLS: {
                    uint32_t nb = (uint32_t)(n.params.size() / 5);
                    for (uint32_t b = 0; b < nb; ++b) {
                        for (int k = 0; k < 5; ++k) { uint32_t w; memcpy(&w, &n.params[b * 5 + k], 4); p.words.push_back(w); }
                        for (int k = 5; k < 8; ++k) p.words.push_back(0u);
                    }
                    p.flops_per_point += 12.0 * nb;
                    break;
                }
                // double-precision operations counted as one flop each, like the binary32 ones
                case OP_TORUS: p.flops_per_point += 13; break;
                case OP_HEART: p.flops_per_point += 24; break;
                case OP_MUSHROOM: p.flops_per_point += 18 + 11; break;
                case OP_LID: p.flops_per_point += 3 - 18; break;           // no transform
                case OP_TETRA:
                case OP_HALF_PLANE:
                    for (float v : n.params) { uint32_t w; memcpy(&w, &v, 4); p.words.push_back(w); }
                    p.flops_per_point += (n.op == OP_TETRA) ? 24 + 3 - 18 : 8;
                    break;
                default: break;
            }
            p.words[skip_at] = (uint32_t)p.words.size();
            push_value(p);
            return;
        }
        // A node asked to preserve `cur` normally saves it on the frame stack.  A node whose own inverse is the identity and
        // whose parent's is too (the inner nodes of the factory's n-ary Union fold, object_factory.hpp:537-580) leaves cur as
        // it found it: I*(I*x) == I*x bit for bit (the product's final "+ 0" has already turned every -0 into +0), so it
        // needs no frame as long as BOTH its children preserve cur.
        const bool soft = preserve && parent_identity && is_identity(n.minv);
        const bool save = preserve && !soft;
        if (save && ++fs_ > p.frame_depth) p.frame_depth = fs_;
        put_itab(p, (uint32_t)(p.words.size() / 4) | ((uint32_t)OP_XFORM << LE_OPC_SHIFT) | (save ? (uint32_t)LE_SAVE : 0u), id_of(n), 0u, 0u);
        const size_t skip_at = put_header(p, (uint32_t)OP_XFORM | (save ? (uint32_t)OP_FLAG_SAVE : 0u) | (id_of(n) << OP_ID_SHIFT), 0u);
        put_matrix(p, n.minv);
        p.flops_per_point += 18 + 1;
        // Deeper child last (no frame to keep while descending; ties keep the reference order) — except down a run of
        // identity nodes, where the deeper child goes FIRST: it preserves cur for free (soft, above), so a Union of N children
        // runs in two value-stack slots and one frame instead of N slots.
        const bool identity_run = is_identity(n.minv) && n.a->is_csg() && is_identity(n.a->minv) && n.a->depth() >= n.b->depth();
        const bool b_first = !identity_run && n.b->depth() < n.a->depth();
        const ShapeNode& first = b_first ? *n.b : *n.a;
        const ShapeNode& second = b_first ? *n.a : *n.b;
        const bool here_identity = is_identity(n.minv);
        emit_node(first, /*preserve=*/first.is_csg(), p, here_identity);
        emit_node(second, /*preserve=*/soft && second.is_csg(), p, here_identity);
        put_itab(p, (uint32_t)(p.words.size() / 4) | IT_CLOSING | (save ? IT_RESTORE : 0u) | (b_first ? IT_SWAPPED : 0u) |
                        (((uint32_t)n.op - (uint32_t)OP_UNION) << IT_OP_SHIFT), id_of(n), id_of(*n.a), id_of(*n.b));
        put_header(p, (uint32_t)n.op | (b_first ? (uint32_t)OP_FLAG_SWAPPED : 0u) | (save ? (uint32_t)OP_FLAG_RESTORE : 0u), 0u);
        put_matrix(p, n.minv);
        vs_ -= 1;   // pops two, pushes one
        if (save) --fs_;
        p.words[skip_at] = (uint32_t)p.words.size();
    }
    static bool is_identity(const float minv[12]) {      // bitwise: ones on the diagonal, +0 elsewhere
        for (int i = 0; i < 12; ++i) {
            uint32_t w; memcpy(&w, &minv[i], 4);
            if (w != ((i % 5 == 0) ? 0x3f800000u : 0u)) return false;
        }
        return true;
    }
};

}  // namespace isd
