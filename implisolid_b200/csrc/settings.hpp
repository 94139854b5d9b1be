// Polygonizer settings: the mc_parameters_json string of build_geometry().
//
// Same keys, defaults and abort conditions as the reference's mc_settings /
// parse_mc_properties_json (polygoniser_settings.hpp:8-34 struct, :70-84 defaults, :101-141
// read_bool_from_json, :147-305 parser).  Booleans are JSON integers 0/1 there (get<int>); a JSON
// true/false is not an int and silently yields the default — kept.  Extra keys under "gpu" are ours
// and are ignored by the reference parser.
#pragma once
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <limits>
#include "json_min.hpp"

namespace isd {

struct Box { float xmin, xmax, ymin, ymax, zmin, zmax; };

struct McSettings {
    Box box;
    int resolution;
    bool ignore_root_matrix;
    int overall_repeats;
    int vresampl_iters;
    float vresampl_c;
    bool qem_enabled, projection_enabled, subdiv_enabled;
    float post_subdiv_noise;
    // ---- ours ("gpu": {...}) ----
    int cube_mode;        // 0: analytic six-plane cube (default), 1: HEAD-literal SDF grid (needs isd_set_sdf_grid)
    int verbose;
    int kernel_times;     // 1: time every kernel of the build with its own pair of events (isd_last_kernel_times)
    int count_flops;      // 1: the interpreters count the FP32 operations they execute, per kernel (isd_last_kernel_flops); an
                          // instrumented pass for bench.py's roofline entries, never a timed one
    int debug_cases;      // 1: keep the dense per-cube case codes for isd_get_case_codes (parity tests)
    int exact_reduce;     // 1: mean edge length as the reference's serial float sum, bit for bit (slow); 0: 64-bit fixed-point sum
    int prune;            // 1 (default): skip grid tiles proven off-surface by interval bounds (mesh identical; field values
                          // exist only in tiles near the surface); 0: evaluate every node (dense field, parity mode)
    int halo;             // z-slab runs with OB02: -1 (default) every GPU optimises its own slab plus a halo of recomputed cube layers
                          // (no mesh exchange); 0: gather the stitched mesh on every GPU and optimise it replicated
    int balance;          // z-slab runs: 2 (default) slab boundaries follow the surface (equal counts of surface tiles per GPU, from a
                          // bounds-only pass every rank repeats identically) and, over repeated builds of one spec, the measured busy
                          // time of every rank (sent along in the step's one all-gather); 1: surface tiles only; 0: equal numbers of
                          // cube layers (SURVEY §8e's formula)
    int pair;             // 1 (default): with per-tile point lists the OB02 kernels take two centroids per thread through the packed f32x2
                          // interpreters (same bits); 0: one centroid per thread
    int finish_split;     // pair path only, the projection's finish: 2 (default) bisection with lane refill from per-warp queues, then the
                          // gradients of all; 1: bisection in batches of 32 units, then gradients; 0: one kernel (two centroids per thread in lockstep)
    int perm;             // the pair kernels take the faces grouped by tile (a warp's lanes then run one instruction list) instead of in face
                          // order: 1 always, 0 never, -1 (default) when the tree's evaluations are heavy enough to pay for the scattered
                          // access (mean flops per tree node >= 40: metaball blends; measured C3 512^3 -11 %, C4 1024^3 +3.6 %)
    int search_split;     // 1 (default): pair path only — the search's trial points run in a kernel of their own, lanes refilled from per-warp
                          // queues (k_search_trials_refill); 0: in lockstep inside k_project_search_pair
    int refill_range;     // units per ticket of the lane-refill kernels; 0 (default): 128, fewer (>= 32) for meshes too small to fill the GPU
    int mc_adjacency;     // 1 (default): umbrellas and faces-of-faces of the mesh K4 emitted come straight from the cube case codes
                          // (mc_adjacency.cuh); 0: from the face list, as for any other mesh (same tables, bit for bit)
    int speculate;        // 1 (default): a build of the spec the previous build had takes that build's counts (mesh size, cut cubes, left-over
                          // centroids) instead of waiting for the device in mid-step; the counts are checked when the build ends and a
                          // mismatch drops the result and builds again the waiting way; 0: always wait
    int point_spec;       // -1 (default): the OB02 kernels run per-tile instruction lists (from K1's liveness masks) at their evaluation
                          // points when the tree has at least 8 nodes; 0: never; 1: always (when masks exist)
    int normals;          // 1: per-vertex normals (-grad f/|grad f|) computed right after the last step (isd_get_n_ptr); 0: on first use
    int specialize;       // 1 (default): per-tile liveness masks let the evaluator skip CSG children that interval bounds prove
                          // are never selected in that tile (field values bit-identical); 0: every tile runs the whole tree
};

inline bool settings_bool(const JsonDoc& d, const char* field, bool dflt, const char* forbidden, bool* needs_abort) {
    const int kUnset = -1;
    int v = d.get_int(field, kUnset);
    bool val = (v == kUnset) ? dflt : (v != 0);
    const int kMagic = 9123456;
    if (d.get_int(forbidden, kMagic) != kMagic) {
        fprintf(stderr, "Error: Use %s instead of using%s. Aborting! \n", field, forbidden);
        *needs_abort = true;
    }
    return val;
}

inline McSettings parse_mc_settings(const char* json) {
    JsonDoc d(json);
    McSettings s;
    bool needs_abort = false;
    const float nan = std::numeric_limits<float>::quiet_NaN();
    float b[6] = {d.get_float("box.xmin", nan), d.get_float("box.xmax", nan), d.get_float("box.ymin", nan),
                  d.get_float("box.ymax", nan), d.get_float("box.zmin", nan), d.get_float("box.zmax", nan)};
    for (float v : b)
        if (std::isnan(v)) needs_abort = true;
    if (needs_abort) fprintf(stderr, "Error: missing or incorrect values in mc_parameters_json\n");
    s.box = Box{b[0], b[1], b[2], b[3], b[4], b[5]};

    float res_real = d.get_float("resolution", -1.0f);
    int res_int = (int)res_real;
    if (res_int == -1) res_int = 28;
    if ((float)res_int != res_real) {      // also hit when "resolution" is absent (28 != -1), as in the reference
        fprintf(stderr, "Error: resolution must be integer: %g != %g\n", (double)res_int, (double)res_real);
        needs_abort = true;
    }
    s.resolution = res_int;
    if (s.resolution <= 2) {
        fprintf(stderr, "Error: missing or incorrect values in mc_parameters_json\n");
        needs_abort = true;
    }
    if (s.resolution > 65535 - 5) {        // dim_t is uint16 in the reference (mcc2.cpp:68)
        fprintf(stderr, "Error: resolution too large\n");
        needs_abort = true;
    }
    s.vresampl_c = d.get_float("vresampl.c", 1.0f);
    s.vresampl_iters = d.get_int("vresampl.iters", 0);
    s.projection_enabled = settings_bool(d, "projection.enabled", false, "projection.enable", &needs_abort);
    s.qem_enabled = settings_bool(d, "qem.enabled", false, "qem.enable", &needs_abort);
    s.subdiv_enabled = settings_bool(d, "subdiv.enabled", true, "subdiv.enable", &needs_abort);
    s.post_subdiv_noise = d.get_float("debug.post_subdiv_noise", 0.01f);
    s.overall_repeats = d.get_int("overall_repeats", 1);
    s.ignore_root_matrix = d.get_bool("ignore_root_matrix", false);
    s.cube_mode = d.get_int("gpu.cube_mode", 0);
    s.verbose = d.get_int("gpu.verbose", 0);
    s.debug_cases = d.get_int("gpu.debug_cases", 0);
    s.kernel_times = d.get_int("gpu.kernel_times", 0);
    s.count_flops = d.get_int("gpu.count_flops", 0);
    s.exact_reduce = d.get_int("gpu.exact_reduce", 0);
    s.prune = d.get_int("gpu.prune", 1);
    s.specialize = d.get_int("gpu.specialize", 1);
    s.normals = d.get_int("gpu.normals", 0);
    s.halo = d.get_int("gpu.halo", -1);
    s.point_spec = d.get_int("gpu.point_spec", -1);
    s.balance = d.get_int("gpu.balance", 2);
    s.pair = d.get_int("gpu.pair", 1);
    s.finish_split = d.get_int("gpu.finish_split", 2);
    s.speculate = d.get_int("gpu.speculate", 1);
    s.mc_adjacency = d.get_int("gpu.mc_adjacency", 1);
    s.refill_range = d.get_int("gpu.refill_range", 0);
    s.search_split = d.get_int("gpu.search_split", 1);
    s.perm = d.get_int("gpu.perm", -1);
    if (s.qem_enabled && !s.projection_enabled) {
        fprintf(stderr, "Invalid settings: QEM will not be applied if centroid projection is disabled\n");
        needs_abort = true;
    }
    if (needs_abort) {
        fprintf(stderr, "Aborting because of a problem in mc_settings_from_json.\n");
        abort();
    }
    return s;
}

}  // namespace isd
