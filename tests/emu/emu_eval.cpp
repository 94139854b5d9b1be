// TEST FIXTURE (CPU): compiles the product's shape compiler (program.hpp) and the __host__ __device__ program
// evaluator (csg_eval.cuh) for the host, so the flattening and the interpreter arithmetic can be checked against the
// oracle without a GPU.  Built by tests/conftest.py into tests/emu/libemu.so; never linked into the product.
#include <cstring>
#include <vector>
#include "../../implisolid_b200/csrc/csg_eval.cuh"

using namespace isd;

struct HostProg {
    const uint32_t* w;
    uint32_t operator[](int i) const { return w[i]; }
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
        eval_values<4>(prog, sdf, x, y, z, out);
        for (int k = 0; k < 4 && i + k < n; k++) f4_out[i + k] = out[k];
    }
}

void emu_affine_inverse(const float* in12, float* out12) { affine_inverse(in12, out12); }

}
