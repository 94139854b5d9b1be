// ORACLE — TEST INFRASTRUCTURE ONLY (see ojson.hpp header).  PARITY UNPINNED: the reference holds no
// tests or golden vectors for this path and cannot be compiled here (Boost/Eigen/Emscripten absent,
// SURVEY §0 F1-F2), so this restatement is pinned only by the reference's known-answer comments and
// analytic properties (tests/test_oracle_*.py).
//
// Exports the reference's extern "C" geometry API (mcc2.cpp:88-134) so one harness drives the oracle and the
// CUDA build alike, restating mcc2.cpp:165-193 (state_t), :222-249 (mc_start), :309-444 (grand_algorithm),
// :446-552 (build_geometry, getters, finish_geometry), :699-911 (direct evaluation service),
// polygoniser_settings.hpp:70-84,101-141,147-305 (settings) and polygonizer_algorithm_ob02.hpp:74-176.
// Extra oracle_* entry points expose intermediates (field, case codes, stats) to the parity tests.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <limits>
#include <string>
#include <vector>
#include "omc.hpp"
#include "oob02.hpp"
#include "osubdiv.hpp"

using namespace oracle;

namespace {

struct mc_settings {
    bounding_box box;
    int resolution;
    bool ignore_root_matrix;
    int overall_repeats;
    struct { int iters; REAL c; } vresampl;
    struct { bool enabled; } qem, projection, subdiv;
    struct { REAL post_subdiv_noise; } debug;
    int cube_mode;     // ours: "gpu.cube_mode" 0 analytic (default) / 1 HEAD SDF grid
    int exact_reduce;  // ours: "gpu.exact_reduce" 1 = the reference's serial float sum of the edge lengths (centroids_projection.cpp:70-82);
                       // 0 (default, like the product) = the product's 64-bit fixed-point sum, a documented deviation
};

bool read_bool_from_json(const ojson::Node& d, const std::string& field, bool dflt, bool* needs_abort,
                         const std::vector<std::string>& forbidden) {
    const int MAGIC = -1;
    int v = d.get_int(field, MAGIC);
    bool val = !!v;
    if (v == MAGIC) val = dflt;
    for (const std::string& bad : forbidden) {
        const int MAGIC2 = 9123456;
        if (d.get_int(bad, MAGIC2) != MAGIC2) {
            std::cerr << "Error: Use " << field << " instead of using" << bad << ". Aborting! " << std::endl;
            *needs_abort = true;
        }
    }
    return val;
}

mc_settings parse_mc_properties_json(const char* json) {
    mc_settings s;
    bool needs_abort = false;
    ojson::Node d = ojson::parse(json);
    const REAL NaN = std::numeric_limits<REAL>::quiet_NaN();
    REAL xmin = d.get_float("box.xmin", NaN), xmax = d.get_float("box.xmax", NaN);
    REAL ymin = d.get_float("box.ymin", NaN), ymax = d.get_float("box.ymax", NaN);
    REAL zmin = d.get_float("box.zmin", NaN), zmax = d.get_float("box.zmax", NaN);
    if (xmin != xmin || xmax != xmax || ymin != ymin || ymax != ymax || zmin != zmin || zmax != zmax) {
        std::cerr << "Error: missing or incorrect values in mc_parameters_json" << std::endl;
        needs_abort = true;
    }
    s.box = bounding_box{xmin, xmax, ymin, ymax, zmin, zmax};
    {
        REAL resolution_real = d.get_float("resolution", -1);
        int resolution_int = static_cast<int>(resolution_real);
        if (resolution_int == -1) resolution_int = 28;
        if (static_cast<REAL>(resolution_int) != resolution_real) {   // NB: a missing "resolution" lands here too (28 != -1)
            std::cerr << "Error: resolution must be integer" << std::endl;
            needs_abort = true;
        }
        s.resolution = resolution_int;
    }
    if (s.resolution <= 2) {
        std::cerr << "Error: missing or incorrect values in mc_parameters_json" << std::endl;
        needs_abort = true;
    }
    s.vresampl.c = d.get_float("vresampl.c", 1.0);
    s.vresampl.iters = d.get_int("vresampl.iters", 0);
    s.projection.enabled = read_bool_from_json(d, "projection.enabled", false, &needs_abort, {"projection.enable"});
    s.qem.enabled = read_bool_from_json(d, "qem.enabled", false, &needs_abort, {"qem.enable"});
    s.subdiv.enabled = read_bool_from_json(d, "subdiv.enabled", true, &needs_abort, {"subdiv.enable"});
    s.debug.post_subdiv_noise = d.get_float("debug.post_subdiv_noise", 0.01);
    s.overall_repeats = d.get_int("overall_repeats", 1);
    s.ignore_root_matrix = d.get_bool("ignore_root_matrix", false);
    s.cube_mode = d.get_int("gpu.cube_mode", 0);
    s.exact_reduce = d.get_int("gpu.exact_reduce", 0);
    if (s.qem.enabled && !s.projection.enabled) {
        std::clog << "Invalid settings: QEM will not be applied if centroid projection is disabled" << std::endl;
        needs_abort = true;
    }
    if (needs_abort) {
        std::cout << "Aborting because of a problem in mc_settings_from_json." << std::endl;
        abort();
    }
    return s;
}

struct state_t {
    bool active = false;
    std::vector<REAL> mc_result_verts;
    std::vector<int> mc_result_faces;
} _state;

// intermediates kept for the parity tests
std::vector<REAL> g_field;
std::vector<uint8_t> g_cases;
std::vector<int64_t> g_keys;
int g_S = 0;
Stats g_stats;
ProjectionDebug g_proj;
SdfGrid g_sdf;
bool g_quiet = true;

Env make_env(int cube_mode) {
    Env e;
    e.cube_mode = cube_mode ? CUBE_HEAD_SDFGRID : CUBE_ANALYTIC;
    e.sdf = &g_sdf;
    if (cube_mode && g_sdf.v.empty()) { std::cerr << "oracle: cube_mode 1 needs oracle_set_sdf_grid()" << std::endl; abort(); }
    return e;
}

// send_mesh_back_to_client (polygonizer_algorithm_ob02.hpp:180-220): the EM_ASM block hands the JS client typed-array views on
// _state.mc_result_verts / faces plus the three ids of worker_call_sepcs_t; here a C callback receives the same.
typedef void (*progress_fn_t)(const REAL*, int, const int*, int, int, int, int, void*);
progress_fn_t g_progress = nullptr;
void* g_progress_user = nullptr;
int g_call_specs[3] = {-1, -1, -1};      // progress_callback_id, shape_id, call_id
void send_mesh_back_to_client() {
    if (!g_progress) return;
    g_progress(_state.mc_result_verts.data(), (int)(_state.mc_result_verts.size() / 3), _state.mc_result_faces.data(),
               (int)(_state.mc_result_faces.size() / 3), g_call_specs[0], g_call_specs[1], g_call_specs[2], g_progress_user);
}

double g_stage_s[4] = {0, 0, 0, 0};      // eval, marching cubes, resampling, projection+qem (seconds, last build)
double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// layer0 < 0: the whole grid (the reference).  Otherwise cube layers [layer0, layer1) only (MarchingCubes' window).
void grand_algorithm(const char* shape_json, const mc_settings& s, int layer0 = -1, int layer1 = -1, bool keep_debug = true) {
    if (_state.active) {
        std::clog << "Error: There are non-deallocated geometry resources. Call finit_geometry() first. Not doing anything.";
        std::clog << "build_geometry() called in a bad state.";
        return;
    }
    Shape shape;
    ojson::Node doc = ojson::parse(shape_json);
    const implicit_function* root = object_factory(shape, doc, s.ignore_root_matrix, make_env(s.cube_mode));
    chunked_function chunked(root);      // threads() == 1: passes every call straight through
    const implicit_function* object = &chunked;
    g_stats = Stats();
    for (double& t : g_stage_s) t = 0;
    {   // polygonize_step_0 / mc_start
        MarchingCubes mc(s.resolution, s.box, layer0, layer1);
        mc.keep_cases = keep_debug;
        double t0 = now_s();
        mc.eval_shape(*root);            // chunks of z-planes, concurrent when threads() > 1
        mc.seal_exterior(-10000000.0);
        double t1 = now_s();
        mc.render_geometry();
        g_stage_s[0] = t1 - t0; g_stage_s[1] = now_s() - t1;
        _state.mc_result_verts = std::move(mc.result_verts);
        _state.mc_result_faces = std::move(mc.result_faces);
        g_field = std::move(mc.field);
        g_cases = std::move(mc.case_codes);
        g_keys = std::move(mc.result_keys);
        g_S = mc.S;
    }
    send_mesh_back_to_client();                                                                  // mcc2.cpp:351
    for (int overall_iter = 0; overall_iter < s.overall_repeats; ++overall_iter) {
        double t0 = now_s();
        for (int i = 0; i < s.vresampl.iters; i++)
            vertex_resampling(*object, s.vresampl.c, _state.mc_result_verts, _state.mc_result_faces);   // polygonize_step_1
        double t1 = now_s();
        g_stage_s[2] += t1 - t0;
        send_mesh_back_to_client();                                                              // mcc2.cpp:372
        if (s.projection.enabled) {
            centroids_projection(*object, _state.mc_result_verts, _state.mc_result_faces, s.qem.enabled, &g_stats, &g_proj,
                                 s.exact_reduce != 0);
            g_stage_s[3] += now_s() - t1;
            send_mesh_back_to_client();                                                          // mcc2.cpp:390
        }
        const bool is_last = overall_iter == s.overall_repeats - 1;
        if (s.subdiv.enabled && (s.overall_repeats <= 1 || is_last))                             // mcc2.cpp:396-410
            polygonize_step_3(_state.mc_result_verts, _state.mc_result_faces, s.debug.post_subdiv_noise, is_last);
    }
    _state.active = true;   // polygonize_terminate
}

struct {
    Shape* shape = nullptr;
    const implicit_function* current_object = nullptr;
    vects* current_x = nullptr;
    vects* current_grad = nullptr;
    scalars* current_f = nullptr;
} svc;

}  // namespace

extern "C" {

void build_geometry(const char* shape_parameters_json, const char* mc_parameters_json) {
    mc_settings s = parse_mc_properties_json(mc_parameters_json);
    grand_algorithm(shape_parameters_json, s);
}
void build_geometry_u(const char* shape_parameters_json, const char* mc_parameters_json, const char* call_specs) {
    if (call_specs && *call_specs) {                 // worker_call_specs.hpp:20-34
        ojson::Node d = ojson::parse(call_specs);
        g_call_specs[0] = d.get_int("progressCallback_id", -1);
        g_call_specs[2] = d.get_int("call_id", -1);
        g_call_specs[1] = d.get_int("shape_id", -1);
    }
    build_geometry(shape_parameters_json, mc_parameters_json);
    g_call_specs[0] = g_call_specs[1] = g_call_specs[2] = -1;
}
void oracle_set_progress_callback(progress_fn_t fn, void* user) { g_progress = fn; g_progress_user = user; }
void oracle_srand(unsigned seed) { subdiv_rand().seed(seed); }
// the private generator against libc's own rand() (test_oracle_known_answers.py): number of mismatches in n draws after srand(seed)
int oracle_rand_check(unsigned seed, int n) {
    GlibcRand g;
    g.seed(seed);
    srand(seed);
    int bad = 0;
    for (int i = 0; i < n; i++) bad += (g.next() != rand());
    return bad;
}
void oracle_subdivide(float* verts_in, int nv, const int* faces_in, int nf, float* verts_out, int* faces_out, int* nv_out) {
    std::vector<REAL> v(verts_in, verts_in + (size_t)nv * 3);
    std::vector<int> f(faces_in, faces_in + (size_t)nf * 3);
    subdivide_all_1to4(v, f);
    *nv_out = (int)(v.size() / 3);
    std::copy(v.begin(), v.end(), verts_out);
    std::copy(f.begin(), f.end(), faces_out);
}
int get_f_size() { return (int)(_state.mc_result_faces.size() / 3); }
int get_v_size() { return (int)(_state.mc_result_verts.size() / 3); }
void get_v(REAL* v_out, int vcount) {
    size_t n = _state.mc_result_verts.size();
    for (size_t i = 0; i < n; i++) v_out[i] = _state.mc_result_verts[i];
    if ((size_t)vcount * 3 != n) std::clog << "sizes dont match: " << n / 3. << " " << vcount << std::endl;
}
void get_f(int* f_out, int fcount) {
    size_t n = _state.mc_result_faces.size();
    for (size_t i = 0; i < n; i++) f_out[i] = _state.mc_result_faces[i];
    if ((size_t)fcount * 3 != n) std::clog << "sizes dont match: " << n / 3. << " " << fcount << std::endl;
}
void* get_v_ptr() { return _state.mc_result_verts.data(); }
void* get_f_ptr() { return _state.mc_result_faces.data(); }
void finish_geometry() { _state.active = false; }

int set_object(const char* shape_parameters_json, bool ignore_root_matrix) {
    if (svc.current_object != nullptr) {
        std::clog << "Error: You cannot unset() the object before a set_object(json)." << std::endl;
        return 0;
    }
    svc.shape = new Shape();
    ojson::Node doc = ojson::parse(shape_parameters_json);
    svc.current_object = object_factory(*svc.shape, doc, ignore_root_matrix, make_env(0));
    return 1;
}
bool unset_object(int id) {
    if (svc.current_object == nullptr) {
        std::clog << "Error: You cannot unset() the object before a set_object(json)." << std::endl;
        return false;
    }
    if (id <= 0) { std::clog << "Incorrect ID: use the same id returned by set_object(json)." << std::endl; return false; }
    if (id != 1) { std::clog << "Incorrect ID. For now, The only id is 1" << std::endl; return false; }
    delete svc.shape;
    svc.shape = nullptr;
    svc.current_object = nullptr;
    return true;
}
bool set_x(void* verts, int n) {
    if (svc.current_x != nullptr) {
        std::clog << "Error: You set() before unset()ing the previous set()." << std::endl;
        return false;
    }
    if (n < 0 || n >= 10000 * 5) {
        std::clog << "Error: n is outside [0, 50000]." << std::endl;
        return false;
    }
    svc.current_x = new vects(n);
    const REAL* rv = reinterpret_cast<const REAL*>(verts);
    for (int i = 0; i < n * 3; i++) svc.current_x->d[i] = rv[i];
    svc.current_f = new scalars(n);
    svc.current_grad = new vects(n);
    return true;
}
void unset_x() {
    if (svc.current_x == nullptr) {
        std::clog << "Error: You cannot unset() before a set()." << std::endl;
        return;
    }
    delete svc.current_x; delete svc.current_f; delete svc.current_grad;
    svc.current_x = nullptr; svc.current_f = nullptr; svc.current_grad = nullptr;
}
void calculate_implicit_values() {
    if (!svc.current_x || !svc.current_f || !svc.current_object) {
        std::clog << "Error: You need to set_x() and set_object() first." << std::endl;
        return;
    }
    if (svc.current_x->n()) svc.current_object->eval_implicit(*svc.current_x, svc.current_f);
}
void* get_values_ptr() { return svc.current_f->data(); }
int get_values_size() { return (int)svc.current_f->size(); }
void calculate_implicit_gradients(bool normalize_and_invert) {
    if (!svc.current_x || !svc.current_grad || !svc.current_object) {
        std::clog << "Error: You need to set_x() and set_object() first." << std::endl;
        return;
    }
    if (svc.current_x->n()) svc.current_object->eval_gradient(*svc.current_x, svc.current_grad);
    if (normalize_and_invert) {
        for (size_t i = 0; i < svc.current_grad->n(); i++) {
            REAL* g = (*svc.current_grad)[i];
            REAL x = g[0], y = g[1], z = g[2];
            REAL norm = std::sqrt(x * x + y * y + z * z);
            REAL norm_factor;
            if (norm > 0.0001) norm_factor = -1.0 / norm;
            else norm_factor = -42.0;
            g[0] = x * norm_factor; g[1] = y * norm_factor; g[2] = z * norm_factor;
        }
    }
}
void* get_gradients_ptr() {
    if (!svc.current_x || !svc.current_f || !svc.current_object) {
        std::clog << "Error: You need to set_x() and set_object() first." << std::endl;
        return nullptr;
    }
    return svc.current_grad->d.data();
}
int get_gradients_size() {
    if (!svc.current_x || !svc.current_f || !svc.current_object) {
        std::clog << "Error: You need to set_x() and set_object() first." << std::endl;
        return 0;
    }
    return (int)svc.current_grad->n() * 3;
}
void* get_pointset_ptr(char*) { return nullptr; }
int get_pointset_size(char*) { return 0; }
void about() { std::cout << "implisolid oracle (CPU restatement; test infrastructure)" << std::endl; }

// ---- oracle-only: host threads, bounded samples of a large grid ---------------------------------
void oracle_set_threads(int n) { threads() = n > 1 ? n : 1; }
int oracle_get_threads() { return threads(); }
// build_geometry restricted to cube layers [layer0, layer1) of the grid the settings describe: field evaluation of the node
// planes those layers touch, marching cubes over them, then the OB02 steps on the mesh that results (its open rims included).
// For timing a bounded sample of a workload too large to run whole, and for checking a slab of a large GPU mesh.
void oracle_build_layers(const char* shape_parameters_json, const char* mc_parameters_json, int layer0, int layer1) {
    mc_settings s = parse_mc_properties_json(mc_parameters_json);
    grand_algorithm(shape_parameters_json, s, layer0, layer1, false);
}
void oracle_stage_seconds(double* out4) { for (int i = 0; i < 4; i++) out4[i] = g_stage_s[i]; }

// ---- oracle-only: pins against the reference's own known answers and independent implementations ----------------------------
int oracle_tetrahedron_integrity(const float* corners12, const float* m12) {
    REAL c[4][3];
    for (int i = 0; i < 4; i++) for (int j = 0; j < 3; j++) c[i][j] = corners12[i * 3 + j];
    tetrahedron t(c, m12);
    return t.integrity_failure();
}
void oracle_tetrahedron_planes(const float* corners12, const float* m12, float* planes16) {
    REAL c[4][3];
    for (int i = 0; i < 4; i++) for (int j = 0; j < 3; j++) c[i][j] = corners12[i * 3 + j];
    tetrahedron t(c, m12);
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) planes16[i * 4 + j] = t.planes[i][j];
}
void oracle_sym_eig3(const float* a9, float* w3, float* v9) {
    REAL A[3][3], w[3], V[3][3];
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) A[i][j] = a9[i * 3 + j];
    sym_eig3(A, w, V);
    for (int i = 0; i < 3; i++) { w3[i] = w[i]; for (int j = 0; j < 3; j++) v9[i * 3 + j] = V[i][j]; }
}
void oracle_mt11213b_raw(unsigned seed, int n, unsigned* out) {
    mt11213b rng(seed);
    for (int i = 0; i < n; i++) out[i] = rng();
}

// ---- oracle-only introspection ----------------------------------------------------------------
int oracle_grid_size() { return g_S; }
const float* oracle_field_ptr() { return g_field.data(); }
long oracle_num_cubes() { return (long)g_cases.size(); }
const uint8_t* oracle_case_codes_ptr() { return g_cases.data(); }
const int64_t* oracle_vertex_keys_ptr() { return g_keys.data(); }
void oracle_get_stats(long* out9) {     // [0..6] by type, [7] failed, [8] capped
    for (int i = 0; i < 7; i++) out9[i] = g_stats.projection_by_type[i];
    out9[7] = g_stats.projection_failed;
    out9[8] = g_stats.bisection_capped;
}
void oracle_get_invariants(long* out2) { out2[0] = g_stats.not_conjugated; out2[1] = g_stats.unresolved_after_bisection; }
int oracle_projection_count() { return (int)g_proj.centroids_after.n(); }
const float* oracle_projected_centroids_ptr() { return g_proj.centroids_after.d.data(); }
const float* oracle_centroid_normals_ptr() { return g_proj.centroid_normals.d.data(); }
float oracle_average_edge() { return g_proj.average_edge; }
void oracle_set_sdf_grid(int nx, int ny, int nz, float h, float ox, float oy, float oz, const float* v) {
    g_sdf.nx = nx; g_sdf.ny = ny; g_sdf.nz = nz; g_sdf.h = h; g_sdf.ox = ox; g_sdf.oy = oy; g_sdf.oz = oz;
    g_sdf.v.assign(v, v + (size_t)nx * ny * nz);
}
// OB02 pieces on caller-supplied meshes (fixtures: octahedron, make_a_cube)
void oracle_faces_of_faces(const int* faces, int nf, int* out) {
    std::vector<int> f(faces, faces + (size_t)nf * 3);
    std::vector<int> fof = build_faces_of_faces(f);
    for (size_t i = 0; i < fof.size(); i++) out[i] = fof[i];
}
// umbrella as CSR: offsets[nv+1], items[3*nf]
void oracle_umbrella(const int* faces, int nf, int nv, int* offsets, int* items) {
    std::vector<int> f(faces, faces + (size_t)nf * 3);
    umbrella_t nb = make_neighbour_faces_of_vertex(f, nv);
    int k = 0;
    for (int v = 0; v < nv; v++) {
        offsets[v] = k;
        for (int fi : nb[v]) items[k++] = fi;
    }
    offsets[nv] = k;
}
int oracle_alpha_list(float average_edge, float* out, int cap) {
    std::vector<REAL> a = make_alpha_list(average_edge, 0.001, average_edge, 20);
    for (size_t i = 0; i < a.size() && (int)i < cap; i++) out[i] = a[i];
    return (int)a.size();
}
void oracle_invert_matrix(const float* in12, float* out12) { invert_matrix(in12, out12); }
void oracle_random_pm1(int n, float amplitude, float* out) {
    vects r = make_random_pm1(n, amplitude);
    for (size_t i = 0; i < r.d.size(); i++) out[i] = r.d[i];
}
// run vertex resampling / projection(+qem) on a caller mesh with a caller shape
void oracle_ob02_on_mesh(const char* shape_json, float* verts, int nv, const int* faces, int nf, int resample_iters, float c,
                         int projection, int qem) {
    Shape shape;
    ojson::Node doc = ojson::parse(shape_json);
    const implicit_function* object = object_factory(shape, doc, false, make_env(0));
    std::vector<REAL> v(verts, verts + (size_t)nv * 3);
    std::vector<int> f(faces, faces + (size_t)nf * 3);
    g_stats = Stats();
    for (int i = 0; i < resample_iters; i++) vertex_resampling(*object, c, v, f);
    if (projection) centroids_projection(*object, v, f, qem != 0, &g_stats, &g_proj);
    for (size_t i = 0; i < v.size(); i++) verts[i] = v[i];
}

}  // extern "C"
