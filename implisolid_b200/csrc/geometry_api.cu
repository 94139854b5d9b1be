// libimplisolid_b200.so — the reference's extern "C" geometry API (mcc2.cpp:88-134) on one B200, plus isd_* extensions.
//
// Host side of the hot path: what mcc2.cpp's grand_algorithm (:309-444), mc_start (:222-249), the polygonizer state
// machine (polygonizer_algorithm_ob02.hpp:14-176) and the direct-evaluation service (mcc2.cpp:699-911) do, restated
// as: parse the two JSON strings (settings.hpp, program.hpp), flatten the CSG tree to a device program, then launch
//   K1 eval_field -> K2 mc_classify -> K3 scan -> [z-slab: ncclAllGather of {nv,nf}] -> K4 mc_emit
//   -> overall_repeats x { vresampl.iters x (K5,K6,K7) ; projection (K5,K8) ; qem (K9) }
// on one stream.  Only the three scan totals cross back to the host before the mesh itself.  There is no CPU
// fallback: every compute entry point aborts if no CUDA device is usable.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/implisolid_b200.h"
#include "device_util.hpp"
#include "field_kernel.cuh"
#include "mc_kernels.cuh"
#include "mc_adjacency.cuh"
#include "ob02_kernels.cuh"
#include "subdiv_kernels.cuh"
#include "settings.hpp"
#include "slab_balance.hpp"

using namespace isd;

namespace {

// ---------------------------------------------------------------------------------------------------------------------
// NCCL, bound at run time (the library has no link-time dependency on it)
// ---------------------------------------------------------------------------------------------------------------------
struct NcclApi {
    void* lib = nullptr;
    typedef struct { char internal[128]; } unique_id;
    int (*GetUniqueId)(unique_id*) = nullptr;
    int (*CommInitRank)(void**, int, unique_id, int) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
    int (*Broadcast)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    int (*CommDestroy)(void*) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    bool load() {
        if (lib) return true;
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* n : names) {
            lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
            if (lib) break;
        }
        if (!lib) { fprintf(stderr, "implisolid_b200: cannot load libnccl.so.2: %s\n", dlerror()); return false; }
#define ISD_NCCL_SYM(field, name)                                             \
    *(void**)(&field) = dlsym(lib, name);                                     \
    if (!field) { fprintf(stderr, "implisolid_b200: NCCL symbol %s missing\n", name); return false; }
        ISD_NCCL_SYM(GetUniqueId, "ncclGetUniqueId");
        ISD_NCCL_SYM(CommInitRank, "ncclCommInitRank");
        ISD_NCCL_SYM(AllGather, "ncclAllGather");
        ISD_NCCL_SYM(Broadcast, "ncclBroadcast");
        ISD_NCCL_SYM(AllReduce, "ncclAllReduce");
        ISD_NCCL_SYM(GroupStart, "ncclGroupStart");
        ISD_NCCL_SYM(GroupEnd, "ncclGroupEnd");
        ISD_NCCL_SYM(CommDestroy, "ncclCommDestroy");
        ISD_NCCL_SYM(GetErrorString, "ncclGetErrorString");
#undef ISD_NCCL_SYM
        return true;
    }
    void check(int rc, const char* what) {
        if (rc != 0) {
            fprintf(stderr, "implisolid_b200: NCCL error in %s: %s\n", what, GetErrorString ? GetErrorString(rc) : "?");
            abort();
        }
    }
};
constexpr int kNcclInt64 = 4, kNcclUint64 = 5, kNcclFloat32 = 7, kNcclInt32 = 2;   // ncclDataType_t values (nccl.h)
constexpr int kNcclSum = 0;                                                        // ncclRedOp_t

// ---------------------------------------------------------------------------------------------------------------------
// process-wide state (the reference keeps one global state_t _state, mcc2.cpp:165-193)
// ---------------------------------------------------------------------------------------------------------------------
enum { T_FIELD = 0, T_CLASSIFY, T_EMIT, T_RESAMPLE, T_PROJECT, T_QEM, T_COPY, T_TOTAL, T_SUBDIV, T_GATHER, T_COUNT };

struct Context {
    bool ready = false;
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;   // device->host copy of the faces overlaps the OB02 stages (faces are final after K4)
    cudaEvent_t ev_faces;
    bool faces_in_flight = false;
    cudaEvent_t ev[16];

    // geometry state
    bool active = false;          // state_t::active
    long long nv = 0, nf = 0;     // mesh held on this GPU (a slab's part in sharded mode until gathered)
    bool host_valid = false;      // pinned mirrors hold the current mesh
    DeviceBuffer d_verts, d_faces;
    long long local_voff = 0, local_foff = 0;   // sharded, before the gather: the slab sits at its global offset inside buffers sized for the whole mesh
    PinnedBuffer h_verts, h_faces;

    // field + marching cubes scratch
    DeviceBuffer d_field, d_signs, d_tile_list, d_tile_masks, d_super_list, d_super_masks, d_bounds, d_row_tot, d_row_base, d_scan_state, d_tile_hints, d_seg_bits, d_seg_off, d_cubes, d_totals, d_cases;
    PinnedBuffer h_small;         // totals and other small read-backs
    PinnedBuffer h_gather;        // {nv,nf} of this slab + of every rank (sized from world in isd_dist_init)
    PinnedBuffer h_program;       // staging of the bounds program + topology
    GridDesc grid{};
    McDesc mc{};
    bool keep_field = false, have_cases = false;
    int S = 0;

    // OB02 scratch
    DeviceBuffer d_cen, d_nrm, d_deg, d_uoff, d_uitems, d_cursor, d_fof, d_w, d_len3, d_best, d_fc, d_state, d_proj, d_pnrm, d_point_lists,
        d_noise, d_scal, d_counters, d_upartials;
    bool adjacency_valid = false, fof_valid_ = false;
    // adjacency from the marching-cubes topology (mc_adjacency.cuh): what K2-K4 left behind, valid while the mesh is K4's own
    McDesc topo_m{}; McBuffers topo_mb{}; EmitArgs topo_ea{};
    bool topo_valid = false, want_topo = true;      // "gpu":{"mc_adjacency":1}
    DeviceBuffer d_perm;                    // faces grouped by tile for the pair kernels (mc_kernels.cuh k_seg_apply) + its length counter in front
    bool perm_valid = false, want_perm = true;      // "gpu":{"perm":1}
    long long perm_units_bound = 0;         // >= perm length / 2 (the length itself stays on the device)
    float average_edge = 0.f;
    bool alpha_ticket_zeroed = false;
    long long proj_n = 0;

    // polygonize_step_3 (subdivision) scratch, per-vertex normals, progress reports
    DeviceBuffer d_sub_owner, d_sub_flag, d_sub_rank, d_verts2, d_faces2, d_ratio, d_normals;
    PinnedBuffer h_normals;
    bool normals_valid = false, normals_host_valid = false;
    bool point_spec_on = false;             // the OB02 kernels may use K1's tile masks (c_point_spec describes the current program)
    int finish_split = 0;                   // "gpu":{"finish_split":..}: the pair path's projection finish; 0 one kernel, 1 bisect batches of 32 units + gradients, 2 bisect with lane refill + gradients
    int refill_range = 0;                   // "gpu":{"refill_range":..}: units per ticket of the lane-refill kernels, 0 = by mesh size
    bool search_split = true;               // "gpu":{"search_split":1}: pair path, the search's trial points run in a lane-refill kernel of their own
    bool pair_kernels = true;               // "gpu":{"pair":1}: with point lists, two centroids per thread through the f32x2 interpreters
    std::vector<uint32_t> last_program;     // the shape program of the last build (lazy normals re-upload it)
    isd_progress_fn progress_fn = nullptr;
    void* progress_user = nullptr;
    int call_specs[3] = {-1, -1, -1};       // progressCallback_id, shape_id, call_id (worker_call_specs.hpp:20-34)

    // HEAD-literal cube table
    DeviceBuffer d_sdf;
    SdfGridView sdf{nullptr, 0, 0, 0, 0.f, 0.f, 0.f, 0.f};

    // stats
    float timings[T_COUNT] = {0};
    int64_t counters[24] = {0};
    int launches = 0;
    double prog_stats[4] = {0, 0, 0, 0};   // flops/point, nodes, words, value-stack depth

    // z-slab sharding
    int rank = 0, world = 1;
    int comm_rank = 0, comm_world = 1;      // what the communicator was made for (isd_dist_suspend / isd_dist_resume)
    bool emulate = false;
    long long emu_voff = 0, emu_foff = 0;
    NcclApi nccl;
    void* comm = nullptr;
    DeviceBuffer d_gather, d_balance;
    int64_t slab[8] = {0};
    bool gathered = false;
    // halo mode (sharded OB02 without a mesh exchange): the GPU holds a LOCAL mesh of its slab plus `halo` recomputed cube
    // layers on either side, numbered from 0 in scan order (ghost-layer vertices first, never written); the slab's own part is
    // the contiguous range [own_v0, own_v1) x [own_f0, own_f1) and global index = local index + (global offset - own_*0).
    bool halo_mode = false, halo_pending = false, halo_gather_enqueued = false;
    long long halo_counts[2] = {0, 0};       // this slab's {nv, nf}, what it contributes to the gather
    int halo = 0, own_zb = 0, own_ze = 0;
    long long own_v0 = 0, own_v1 = 0, own_f0 = 0, own_f1 = 0;
    unsigned long long emu_edge_total = 0;      // emulated ranks: the global fixed-point edge sum and face count
    long long emu_gnf = 0;
    unsigned long long last_edge_sum = 0;       // this GPU's share of the fixed-point edge sum (own faces only)

    // direct evaluation service (mcc2.cpp:699-705)
    bool svc_object = false;
    std::vector<uint32_t> svc_program;
    bool svc_uses_sdf = false;
    bool svc_x = false;
    int svc_n = 0;
    DeviceBuffer d_svc_x, d_svc_f, d_svc_g;
    PinnedBuffer h_svc_f, h_svc_g;
} g;

void ensure_init(int device = -1) {
    if (g.ready) return;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        fprintf(stderr, "implisolid_b200: no CUDA device available (%s); this library has no CPU path\n",
                e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
        abort();
    }
    if (device < 0) {
        const char* s = getenv("ISD_DEVICE");
        if (!s) s = getenv("LOCAL_RANK");
        device = s ? atoi(s) : 0;
        if (device >= count) device = device % count;
    }
    ISD_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    ISD_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) {
        fprintf(stderr, "implisolid_b200: device %d is sm_%d%d; this build is sm_100a only\n", device, prop.major, prop.minor);
        abort();
    }
    g.device = device;
    g.sm_count = prop.multiProcessorCount;
    ISD_CUDA(cudaStreamCreateWithFlags(&g.stream, cudaStreamNonBlocking));
    ISD_CUDA(cudaStreamCreateWithFlags(&g.copy_stream, cudaStreamNonBlocking));
    ISD_CUDA(cudaEventCreateWithFlags(&g.ev_faces, cudaEventDisableTiming));
    for (auto& ev : g.ev) ISD_CUDA(cudaEventCreate(&ev));
    // tables derived from the reference's own (mc_tables.h)
    uint16_t info[256];
    for (int i = 0; i < 256; ++i) info[i] = (uint16_t)(ISD_MC_NTRI[i] | (ISD_MC_OWNED[i] << 8));
    ISD_CUDA(cudaMemcpyToSymbol(d_case_info, info, sizeof(info)));
    ISD_CUDA(cudaMemcpyToSymbol(d_tri_table, ISD_MC_TRI, sizeof(ISD_MC_TRI)));
    ISD_CUDA(cudaMemcpyToSymbol(d_edge_mask, ISD_MC_EDGE_MASK, sizeof(ISD_MC_EDGE_MASK)));
    {   // which triangles of a case use which local edge (mc_adjacency.cuh), from the reference's own triangle table
        static uint8_t edge_tris[256 * 12];
        memset(edge_tris, 0, sizeof(edge_tris));
        for (int c = 0; c < 256; ++c)
            for (int t = 0; t < (int)ISD_MC_NTRI[c]; ++t)
                for (int k = 0; k < 3; ++k) edge_tris[c * 12 + ISD_MC_TRI[c * 16 + 3 * t + k]] |= (uint8_t)(1u << t);
        ISD_CUDA(cudaMemcpyToSymbol(d_edge_tris, edge_tris, sizeof(edge_tris)));
    }
    g.h_small.reserve(4096);
    g.ready = true;
}

inline void launched() { g.launches++; }
#define ISD_LAUNCH_CHECK() do { ISD_CUDA(cudaGetLastError()); launched(); } while (0)

// Per-kernel device times ("gpu":{"kernel_times":1}; bench.py's roofline entries): a pair of events around every launch,
// read after the build's last synchronisation.  Off by default: two event records per launch are a few per cent of a 0.4 ms step.
enum { KT_SUPER_BOUNDS = 0, KT_TILE_BOUNDS, KT_EVAL_TILES, KT_CLASSIFY, KT_ROW_SCAN, KT_POINT_LISTS, KT_SEG_APPLY, KT_EMIT, KT_ADJACENCY, KT_FOF,
       KT_CENTROID_GRAD, KT_FACE_WEIGHTS, KT_RESAMPLE, KT_FACE_PREPARE, KT_SEARCH, KT_FINISH, KT_LEFTOVERS, KT_QEM, KT_OTHER, KT_UNIT_GRAD, KT_TRIALS, KT_N };
// the names of the kernels a build actually launched under each id (the pair path and the finish's form change them)
const char* g_kernel_names[KT_N] = {"k_super_bounds", "k_tile_bounds", "k_eval_tiles", "k_mc_classify_tiles", "k_row_scan", "k_build_point_lists",
                                    "k_seg_apply", "k_mc_emit", "adjacency(k_vertex_degree+k_uscan+k_umbrella_fill+k_umbrella_sort)", "k_faces_of_faces",
                                    "k_centroid_gradients", "k_face_weights", "k_resample", "k_face_prepare", "k_project_search", "k_project_finish",
                                    "k_project_leftovers", "k_qem", "other(subdivision scan)", "k_unit_gradients_pair", "k_search_trials_refill"};
struct KernelTimes {
    bool on = false;
    std::vector<cudaEvent_t> pool;
    size_t used = 0;
    struct Span { int id; cudaEvent_t a, b; };
    std::vector<Span> spans;
    float ms[KT_N] = {0};
    float gap_ms[KT_N] = {0};      // idle stream time BEFORE each kernel (memsets, copies, host waits, collectives), same indices
    int launches[KT_N] = {0};
    bool counting = false, flag_on_device = false;      // "gpu":{"count_flops":1}: executed FP32 operations per kernel
    double flops[KT_N] = {0};
    cudaEvent_t get() {
        if (used == pool.size()) { cudaEvent_t e; ISD_CUDA(cudaEventCreate(&e)); pool.push_back(e); }
        return pool[used++];
    }
    void reset(bool enable) { on = enable; used = 0; spans.clear(); for (float& v : ms) v = 0; for (float& v : gap_ms) v = 0; for (int& v : launches) v = 0; }
    void start_counting(bool count);
    void finish_counting();
    void collect() {      // the stream has been synchronised
        for (size_t k = 0; k < spans.size(); ++k) {
            const Span& sp = spans[k];
            float t = 0;
            ISD_CUDA(cudaEventElapsedTime(&t, sp.a, sp.b));
            ms[sp.id] += t;
            launches[sp.id]++;
            if (k > 0) { ISD_CUDA(cudaEventElapsedTime(&t, spans[k - 1].b, sp.a)); gap_ms[sp.id] += t; }
        }
        spans.clear(); used = 0;
    }
} g_kt;
void KernelTimes::start_counting(bool count) {
    for (double& v : flops) v = 0;
    if (count != flag_on_device) {
        const int flag = count ? 1 : 0;
        ISD_CUDA(cudaMemcpyToSymbolAsync(c_count_flops, &flag, sizeof(int), 0, cudaMemcpyHostToDevice, g.stream));
        ISD_CUDA(cudaStreamSynchronize(g.stream));      // `flag` is a stack variable
        flag_on_device = count;
    }
    counting = count;
    if (count) {
        static const unsigned long long zeros[33] = {0};
        ISD_CUDA(cudaMemcpyToSymbolAsync(d_flops_by_kernel, zeros, 32 * sizeof(unsigned long long), 0, cudaMemcpyHostToDevice, g.stream));
        ISD_CUDA(cudaMemcpyToSymbolAsync(d_flops_now, zeros, sizeof(unsigned long long), 0, cudaMemcpyHostToDevice, g.stream));
    }
}
void KernelTimes::finish_counting() {
    if (!counting) return;
    unsigned long long h[32];
    ISD_CUDA(cudaMemcpyFromSymbolAsync(h, d_flops_by_kernel, sizeof(h), 0, cudaMemcpyDeviceToHost, g.stream));
    ISD_CUDA(cudaStreamSynchronize(g.stream));
    for (int i = 0; i < KT_N; ++i) flops[i] = (double)h[i];
    counting = false;
}
struct KtScope {
    int id; cudaEvent_t a = nullptr;
    explicit KtScope(int i) : id(i) { if (g_kt.on) { a = g_kt.get(); ISD_CUDA(cudaEventRecord(a, g.stream)); } }
    ~KtScope() { if (a) { cudaEvent_t b = g_kt.get(); ISD_CUDA(cudaEventRecord(b, g.stream)); g_kt.spans.push_back({id, a, b}); } }
};
#define KT_CAT2(a, b) a##b
#define KT_CAT(a, b) KT_CAT2(a, b)
#define KT(id) KtScope KT_CAT(kt_scope_, __LINE__)(id)
#define ISD_TIMED(id, ...) do { { KT(id); __VA_ARGS__; ISD_LAUNCH_CHECK(); } if (g_kt.counting) k_flops_move<<<1, 1, 0, g.stream>>>(id); } while (0)

// per-tile instruction lists for evaluations at arbitrary points (field_kernel.cuh PointSpec): valid only for the program
// and the grid K1 has just bounded
void set_point_spec(const PointSpec* ps) {
    if (ps == nullptr) { g.point_spec_on = false; return; }      // the plain kernel instantiations never read c_point_spec
    ISD_CUDA(cudaMemcpyToSymbolAsync(c_point_spec, ps, sizeof(PointSpec), 0, cudaMemcpyHostToDevice, g.stream));
    g.point_spec_on = true;
}

void upload_program(const std::vector<uint32_t>& words) {
    set_point_spec(nullptr);
    ISD_CUDA(cudaMemcpyToSymbolAsync(d_program, words.data(), words.size() * sizeof(uint32_t), 0, cudaMemcpyHostToDevice, g.stream));
}

// Parsing the shape JSON and flattening the tree is a pure function of (text, ignore_root_matrix, cube_mode) and costs 0.3-0.5 ms
// of host time for a 60-node tree, during which the GPU would sit idle: the last few programs are kept (a caller that
// re-polygonizes a shape at another resolution, as the reference's UI does, compiles it once).
const ShapeCompiler::Program& compile_shape(const char* shape_json, bool ignore_root_matrix, int cube_mode) {
    struct Entry { std::string text; bool ignore_root; int cube_mode; ShapeCompiler::Program prog; unsigned long long used; };
    static std::vector<Entry> cache;
    static unsigned long long clock = 0;
    ++clock;
    for (Entry& e : cache)
        if (e.ignore_root == ignore_root_matrix && e.cube_mode == cube_mode && e.text == shape_json) {
            e.used = clock;
            if (e.prog.uses_sdf_grid && g.sdf.v == nullptr) {
                fprintf(stderr, "implisolid_b200: \"gpu.cube_mode\":1 needs isd_set_sdf_grid() first\n");
                abort();
            }
            return e.prog;
        }
    ShapeOptions opt;
    opt.cube_mode = cube_mode;
    ShapeCompiler compiler(opt);
    std::unique_ptr<ShapeNode> tree = compiler.parse(shape_json, ignore_root_matrix);
    ShapeCompiler::Program p = compiler.compile(*tree);
    if (p.uses_sdf_grid && g.sdf.v == nullptr) {
        fprintf(stderr, "implisolid_b200: \"gpu.cube_mode\":1 needs isd_set_sdf_grid() first\n");
        abort();
    }
    if (cache.size() >= 8) {      // evict the least recently used
        size_t victim = 0;
        for (size_t i = 1; i < cache.size(); ++i) if (cache[i].used < cache[victim].used) victim = i;
        cache.erase(cache.begin() + (long)victim);
    }
    cache.push_back(Entry{std::string(shape_json), ignore_root_matrix, cube_mode, std::move(p), clock});
    return cache.back().prog;
}

// ---------------------------------------------------------------------------------------------------------------------
// marching cubes stage
// ---------------------------------------------------------------------------------------------------------------------
void slab_layers(int S, int rank, int world, int* z_begin, int* z_end) {
    const long long layers = S - 3;                       // cube layers z in [1, S-2)
    *z_begin = 1 + (int)((long long)rank * layers / world);
    *z_end = 1 + (int)((long long)(rank + 1) * layers / world);
}

// Slab boundaries that follow the surface (gpu.balance): cube layers weighted by the number of surface tiles in their layer of
// tiles (k_balance_hist over the whole grid) plus a small uniform term for the per-layer fixed costs; rank r gets the layers
// whose cumulative weight falls in [r, r+1) / world of the total.  Deterministic and identical on every rank.
//
// Feedback (real NCCL runs in halo mode, repeated builds of one spec).  Equal surface is not equal work: where the projection's
// first direction fails, warps walk the whole alpha list and phase B runs (C4 1024^3 on two GPUs: 2.28 against 2.83 ms of kernels
// per step with equal surface, the lighter rank idling in the step's collective).  Every build measures its own busy time (device
// time of the build minus the time spent inside the collective) and sends the previous build's along in the step's one
// all-gather (a fourth word beside {nv, nf, edge sum}), tagged with the generation of the cuts it was measured under.  Once every
// rank has reported a time for the current cuts and the slowest is more than 3 % above the mean, each layer's weight is multiplied by
// its owner's measured cost per unit of weight and the cuts are drawn again — the same table on every rank, hence the same cuts.
// At most kMaxRecuts times per (shape, grid, world); a build after a re-cut has a new slab and does not speculate.
constexpr int kGatherWords = 4;            // per rank: nv, nf, fixed-point edge sum, busy word
constexpr int kMaxRecuts = 8;
constexpr int kRecutThresholdPermille = 1030;
struct BalanceState {
    std::vector<uint32_t> key;
    SlabBalance b;                               // weights, cuts, generation (slab_balance.hpp)
    long long my_busy_us = 0; int my_busy_gen = -1;      // this rank's last completed build
    std::vector<long long> table_busy; std::vector<int> table_gen;      // what the last gather said about every rank
} g_bal;

void balanced_layers(const McSettings& s, const ShapeCompiler::Program& prog, int* z_begin, int* z_end) {
    const int S = g.grid.S, layers = S - 3, world = g.world, bins = (S + 7) / 8;
    std::vector<uint32_t> key(prog.bounds);
    const float boxv[6] = {s.box.xmin, s.box.xmax, s.box.ymin, s.box.ymax, s.box.zmin, s.box.zmax};
    for (float v : boxv) { uint32_t w; memcpy(&w, &v, 4); key.push_back(w); }
    key.push_back((uint32_t)S); key.push_back((uint32_t)world);
    if (key != g_bal.key) {
        GridDesc full = g.grid;
        full.z0 = 0; full.nz = S;
        full.tx = full.pitch / 8; full.ty = (S + 7) / 8; full.tz = (S + 7) / 8;
        const size_t prog_bytes = prog.bounds.size() * sizeof(uint32_t), hist_bytes = (size_t)bins * sizeof(unsigned);
        g.d_balance.reserve(prog_bytes + hist_bytes + 64);
        uint32_t* d_prog = g.d_balance.as<uint32_t>();
        unsigned* d_hist = reinterpret_cast<unsigned*>(g.d_balance.as<char>() + ((prog_bytes + 15) & ~(size_t)15));
        ISD_CUDA(cudaMemcpyAsync(d_prog, prog.bounds.data(), prog_bytes, cudaMemcpyHostToDevice, g.stream));
        ISD_CUDA(cudaMemsetAsync(d_hist, 0, hist_bytes, g.stream));
        const int nsuper = ((full.tx + 3) / 4) * ((full.ty + 3) / 4) * ((full.tz + 3) / 4);
        k_balance_hist<<<(nsuper + 127) / 128, 128, 0, g.stream>>>(full, d_prog, d_hist);
        ISD_LAUNCH_CHECK();
        std::vector<unsigned> hist((size_t)bins);
        ISD_CUDA(cudaMemcpyAsync(hist.data(), d_hist, hist_bytes, cudaMemcpyDeviceToHost, g.stream));
        ISD_CUDA(cudaStreamSynchronize(g.stream));
        g_bal.b.init(hist.data(), bins, layers, world);
        g_bal.key = key;
        g_bal.my_busy_gen = -1;
        g_bal.table_busy.assign((size_t)world, 0); g_bal.table_gen.assign((size_t)world, -1);
    } else if (s.balance >= 2 && !g.emulate && g_bal.b.recuts < kMaxRecuts && (int)g_bal.table_busy.size() == world) {
        bool complete = true;
        for (int r = 0; r < world; ++r)
            if (g_bal.table_gen[(size_t)r] != g_bal.b.gen || g_bal.table_busy[(size_t)r] <= 0) complete = false;
        if (complete) g_bal.b.feedback(g_bal.table_busy.data(), world, kRecutThresholdPermille);
    }
    *z_begin = 1 + g_bal.b.cuts[(size_t)g.rank];
    *z_end = 1 + g_bal.b.cuts[(size_t)g.rank + 1];
}

void setup_grid(const McSettings& s, const ShapeCompiler::Program& prog) {
    const int S = s.resolution + 5;                       // skip_count_l + skip_count_h, marching_cubes.hpp:178-181
    GridDesc& gd = g.grid;
    gd.S = S;
    gd.pitch = (S + 7) & ~7;
    gd.SW = (S + 31) / 32 + 1;
    gd.plane = (long long)gd.pitch * S;
    gd.xmin = s.box.xmin; gd.ymin = s.box.ymin; gd.zmin = s.box.zmin;
    gd.wx = (s.box.xmax - s.box.xmin) / (float)s.resolution;
    gd.wy = (s.box.ymax - s.box.ymin) / (float)s.resolution;
    gd.wz = (s.box.zmax - s.box.zmin) / (float)s.resolution;
    gd.two_wx = 2 * gd.wx; gd.two_wy = 2 * gd.wy; gd.two_wz = 2 * gd.wz;
    gd.xi0 = s.box.xmin / gd.wx - 2; gd.yi0 = s.box.ymin / gd.wy - 2; gd.zi0 = s.box.zmin / gd.wz - 2;
    int zb, ze;
    if (g.world > 1 && s.balance && S - 3 >= 2 * g.world) balanced_layers(s, prog, &zb, &ze);
    else slab_layers(S, g.rank, g.world, &zb, &ze);
    g.own_zb = zb; g.own_ze = ze;
    g.halo_mode = false;
    g.halo = 0;
    if (g.world > 1 && s.halo != 0 && s.overall_repeats > 0 && (s.vresampl_iters > 0 || s.projection_enabled) && !s.exact_reduce &&
        !s.subdiv_enabled && !g.progress_fn) {
        // Dependency radius in cube layers: a resampling iteration moves a vertex from the centroids, normals and
        // neighbour-face curvatures of its umbrella (2 layers of input around it), projection + QEM from the faces around it
        // (1 layer).  (Round 1 added one more layer because the lowest halo layer's faces hung on ghost vertices that were never
        // written; k_mc_emit writes them now and tools/halo_check.py holds the stitched meshes bit-identical without that layer —
        // and different with one layer less still.)
        const int per_repeat = 2 * s.vresampl_iters + ((s.projection_enabled && s.qem_enabled) ? 1 : 0);
        g.halo = s.halo > 0 ? s.halo : (s.overall_repeats * per_repeat > 1 ? s.overall_repeats * per_repeat : 1);
        g.halo_mode = true;
        zb = (zb - g.halo > 1) ? zb - g.halo : 1;
        ze = (ze + g.halo < S - 2) ? ze + g.halo : S - 2;
    }
    McDesc& m = g.mc;
    m.emit_z0 = zb;
    m.zc_end = ze;
    m.scan_z0 = (zb > 1) ? zb - 1 : zb;
    if (g.world == 1) { gd.z0 = 0; gd.nz = S; }
    else { gd.z0 = zb - 1; gd.nz = (ze - zb) + 2; }
    gd.tx = gd.pitch / 8; gd.ty = (S + 7) / 8; gd.tz = (gd.nz + 7) / 8;
    m.NX = S - 3;
    m.n_rows = (long long)(ze - m.scan_z0) * m.NX;
    m.ghost_rows = (long long)(zb - m.scan_z0) * m.NX;
    m.g = gd;
    g.S = S;
}

// The one collective of a z-slab step: every rank contributes {nv, nf, fixed-point edge-length sum of its own faces} (24 bytes);
// afterwards d_gather holds the triples of all ranks in rank order and their copy is on its way to h_gather + 3.  `edge_sum`
// (device, may be null = 0) is read in the stream, so the call can follow the kernel that produces it without a host wait.
__global__ void k_pack_slab_triple(long long* __restrict__ dst, long long nv, long long nf, const unsigned long long* __restrict__ edge_sum, long long busy_word) {
    dst[0] = nv; dst[1] = nf; dst[2] = edge_sum ? (long long)*edge_sum : 0ll; dst[3] = busy_word;
}
// events around the collectives of the running build: the time between them is waiting for the other ranks, not work
std::vector<std::pair<cudaEvent_t, cudaEvent_t>> g_gather_spans;
cudaEvent_t g_ev_build_start = nullptr, g_ev_build_end = nullptr;
void enqueue_slab_gather(const unsigned long long* d_edge_sum) {
    g.d_gather.reserve(((size_t)g.world + 1) * kGatherWords * sizeof(int64_t) + 64);
    int64_t* dg = g.d_gather.as<int64_t>();
    int64_t* hs = g.h_gather.as<int64_t>();
    // this rank's busy time of its previous build, tagged with the generation of the cuts it ran under (balanced_layers)
    const long long busy_word = g_bal.my_busy_gen >= 0 ? (g_bal.my_busy_us & ((1ll << 40) - 1)) | ((long long)(g_bal.my_busy_gen + 1) << 40) : 0ll;
    cudaEvent_t ea = g_kt.get(), eb = g_kt.get();
    ISD_CUDA(cudaEventRecord(ea, g.stream));
    k_pack_slab_triple<<<1, 1, 0, g.stream>>>(reinterpret_cast<long long*>(dg + kGatherWords * g.world), g.halo_counts[0], g.halo_counts[1], d_edge_sum, busy_word);
    ISD_LAUNCH_CHECK();
    g.nccl.check(g.nccl.AllGather(dg + kGatherWords * g.world, dg, kGatherWords, kNcclInt64, g.comm, g.stream), "ncclAllGather");
    ISD_CUDA(cudaEventRecord(eb, g.stream));
    g_gather_spans.push_back({ea, eb});
    ISD_CUDA(cudaMemcpyAsync(hs + 3, dg, (size_t)g.world * kGatherWords * sizeof(int64_t), cudaMemcpyDeviceToHost, g.stream));
    g.halo_gather_enqueued = true;
}

// Sizes a build learns from the device half-way (work-list length, vertex / face / cut-cube totals, a slab's own ranges) are a
// deterministic function of (program, grid, slab, settings).  A build whose key equals the previous build's takes them from
// there and never waits for the device in mid-step ("gpu":{"speculate":1}, default); the device's own values still come back
// with the final synchronisation and are compared — a mismatch (never seen: the path is integer-deterministic) discards the
// result and rebuilds without speculation.  A caller that re-polygonizes the same spec pays for one host wait per build, not four.
struct SpecCache {
    bool valid = false;
    std::vector<uint32_t> key;
    unsigned h[24] = {0};
    // projection: centroids left over after phase A, per overall repeat (-1 unknown)
    long long leftovers[8] = {-1, -1, -1, -1, -1, -1, -1, -1};
    long long face_base[8] = {0, 0, 0, 0, 0, 0, 0, 0};      // global index of local face 0 (noise stream offset), per repeat
} g_spec;
bool g_spec_used = false;

void run_marching_cubes(const McSettings& s, const ShapeCompiler::Program& prog) {
    setup_grid(s, prog);
    const GridDesc& gd = g.grid;
    McDesc& m = g.mc;
    const size_t field_bytes = (size_t)gd.plane * gd.nz * sizeof(float);
    g.d_field.reserve(field_bytes);
    const long long n_rows = m.n_rows;
    int scan_blocks = (int)((n_rows + kRowScanThreads * 2 - 1) / (kRowScanThreads * 2));      // >= 2 rounds per block: fewer blocks to wait for
    if (scan_blocks > kRowScanMaxBlocks) scan_blocks = kRowScanMaxBlocks;
    const long long rows_per_block = ((n_rows + scan_blocks - 1) / scan_blocks + kRowScanThreads - 1) / kRowScanThreads * kRowScanThreads;
    g.d_row_tot.reserve((size_t)n_rows * sizeof(Triple));
    g.d_row_base.reserve((size_t)n_rows * sizeof(Triple));
    g.d_scan_state.reserve(((size_t)scan_blocks + 1) * 8 * sizeof(unsigned));
    g.d_totals.reserve(64);
    uint8_t* dbg_cases = nullptr;
    g.have_cases = false;
    if (s.debug_cases) {
        g.d_cases.reserve((size_t)(m.zc_end - m.scan_z0) * m.NX * m.NX);
        dbg_cases = g.d_cases.as<uint8_t>();
        g.have_cases = true;
    }

    const size_t sign_bytes = (size_t)gd.S * gd.nz * gd.SW * sizeof(uint32_t);
    const int ntiles = gd.tx * gd.ty * gd.tz;
    g.d_signs.reserve(sign_bytes);
    g.d_tile_list.reserve((size_t)ntiles * sizeof(unsigned) + 64);
    // bounds program and tree topology share one upload
    const size_t itab_words = prog.itab.size(), bounds_words = prog.bounds.size(), topo_words = prog.topo.size();
    g.d_bounds.reserve((itab_words + bounds_words + topo_words) * sizeof(uint32_t));
    SpecDesc spec{};
    spec.enabled = (s.specialize && prog.can_specialize && prog.num_nodes > 1) ? 1 : 0;
    spec.num_nodes = prog.num_nodes;
    spec.num_ops = (prog.num_nodes - 1) / 2;
    spec.mask_words = (prog.num_nodes + 31) / 32;
    const uint4* d_itab = g.d_bounds.as<uint4>();                       // [itab | bounds program | topology]
    const uint32_t* d_bounds_prog = g.d_bounds.as<uint32_t>() + itab_words;
    spec.topo = d_bounds_prog + bounds_words;
    const int nsuper = ((gd.tx + 3) / 4) * ((gd.ty + 3) / 4) * ((gd.tz + 3) / 4);      // supertiles: 4x4x4 tiles
    g.d_super_list.reserve((size_t)nsuper * sizeof(unsigned) + 64);
    if (spec.enabled) {
        g.d_tile_masks.reserve((size_t)ntiles * spec.mask_words * sizeof(uint32_t));
        g.d_super_masks.reserve((size_t)nsuper * spec.mask_words * sizeof(uint32_t));
    }
    spec.masks = g.d_tile_masks.as<uint32_t>();
    spec.super_masks = g.d_super_masks.as<uint32_t>();
    TileHints hints;
    hints.TW = (gd.tx + 31) / 32;
    const size_t hint_bits_bytes = (size_t)gd.ty * gd.tz * hints.TW * sizeof(uint32_t), hint_any_bytes = ((size_t)gd.ty * gd.tz + 3) & ~(size_t)3;
    g.d_tile_hints.reserve(hint_bits_bytes + hint_any_bytes + (size_t)ntiles * sizeof(unsigned));
    hints.tile_bits = g.d_tile_hints.as<uint32_t>();
    hints.tile_any = g.d_tile_hints.as<uint8_t>() + hint_bits_bytes;
    hints.tile_slot = reinterpret_cast<unsigned*>(g.d_tile_hints.as<uint8_t>() + hint_bits_bytes + hint_any_bytes);
    unsigned* tile_count = g.d_totals.as<unsigned>() + 8;
    ISD_CUDA(cudaEventRecord(g.ev[0], g.stream));
    {   // K1: bounds -> work list -> tiles
        g.h_program.reserve((itab_words + bounds_words + topo_words) * sizeof(uint32_t));
        uint32_t* staged = g.h_program.as<uint32_t>();      // pinned staging, one copy
        memcpy(staged, prog.itab.data(), itab_words * sizeof(uint32_t));
        memcpy(staged + itab_words, prog.bounds.data(), bounds_words * sizeof(uint32_t));
        memcpy(staged + itab_words + bounds_words, prog.topo.data(), topo_words * sizeof(uint32_t));
        ISD_CUDA(cudaMemcpyAsync(g.d_bounds.ptr, staged, (itab_words + bounds_words + topo_words) * sizeof(uint32_t), cudaMemcpyHostToDevice, g.stream));
        ISD_CUDA(cudaMemsetAsync(g.d_signs.ptr, 0xff, sign_bytes, g.stream));
        ISD_CUDA(cudaMemsetAsync(g.d_tile_hints.ptr, 0, hint_bits_bytes + hint_any_bytes, g.stream));
        ISD_CUDA(cudaMemsetAsync(g.d_scan_state.ptr, 0, ((size_t)scan_blocks + 1) * 8 * sizeof(unsigned), g.stream));
        ISD_CUDA(cudaMemsetAsync(tile_count, 0, 4 * sizeof(unsigned), g.stream));      // work-list length, queue head, supertile-list length
        static int resident[2] = {0, 0};      // blocks per SM of the two instantiations (persistent grid = what fits at once)
        if (!resident[0]) {
            ISD_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&resident[0], k_eval_tiles<false>, kTileThreads, 0));
            ISD_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&resident[1], k_eval_tiles<true>, kTileThreads, 0));
        }
        const int level2_blocks = nsuper < g.sm_count * 32 ? nsuper : g.sm_count * 32;
        const int want = g.sm_count * resident[spec.enabled ? 1 : 0];
        const int blocks = ntiles < want ? ntiles : want;
        if (spec.enabled) {
            ISD_TIMED(KT_SUPER_BOUNDS, k_super_bounds<true><<<(nsuper + 127) / 128, 128, 0, g.stream>>>(gd, d_bounds_prog, s.prune, spec, g.d_super_list.as<unsigned>(), tile_count));
            ISD_TIMED(KT_TILE_BOUNDS, k_tile_bounds<true><<<level2_blocks, 64, 0, g.stream>>>(gd, d_bounds_prog, s.prune, spec, g.d_super_list.as<unsigned>(),
                                                                   g.d_signs.as<uint32_t>(), g.d_tile_list.as<unsigned>(), tile_count, hints));
            ISD_TIMED(KT_EVAL_TILES, k_eval_tiles<true><<<blocks, kTileThreads, 0, g.stream>>>(gd, g.sdf, g.d_tile_list.as<unsigned>(), tile_count, spec.masks,
                                                                     spec.mask_words, d_itab, (int)(itab_words / 4), g.d_field.as<float>(),
                                                                     g.d_signs.as<uint32_t>()));
        } else {
            ISD_TIMED(KT_SUPER_BOUNDS, k_super_bounds<false><<<(nsuper + 127) / 128, 128, 0, g.stream>>>(gd, d_bounds_prog, s.prune, spec, g.d_super_list.as<unsigned>(), tile_count));
            ISD_TIMED(KT_TILE_BOUNDS, k_tile_bounds<false><<<level2_blocks, 64, 0, g.stream>>>(gd, d_bounds_prog, s.prune, spec, g.d_super_list.as<unsigned>(),
                                                                    g.d_signs.as<uint32_t>(), g.d_tile_list.as<unsigned>(), tile_count, hints));
            ISD_TIMED(KT_EVAL_TILES, k_eval_tiles<false><<<blocks, kTileThreads, 0, g.stream>>>(gd, g.sdf, g.d_tile_list.as<unsigned>(), tile_count, nullptr, 0,
                                                                      nullptr, 0, g.d_field.as<float>(), g.d_signs.as<uint32_t>()));
        }
    }
    ISD_CUDA(cudaEventRecord(g.ev[1], g.stream));
    // OB02 evaluates at points near the surface: they run per-tile instruction lists, built below once the work-list length is known
    const bool want_point_lists = spec.enabled && s.point_spec != 0 && (s.point_spec > 0 || prog.num_nodes >= 8) && s.overall_repeats > 0 &&
                                  (s.vresampl_iters > 0 || s.projection_enabled);
    // Per-segment arrays: sized for every tile when that is small (no read-back), else for the work list (one extra
    // read-back, negligible at the grid sizes where it happens).
    unsigned* h_actual = g.h_small.as<unsigned>();      // what the device reports (pinned)
    std::vector<uint32_t> key;
    key.reserve(prog.words.size() + prog.bounds.size() + 128);
    key.insert(key.end(), prog.words.begin(), prog.words.end());
    key.insert(key.end(), prog.bounds.begin(), prog.bounds.end());
    {
        uint32_t raw[(sizeof(McDesc) + 3) / 4] = {0};
        memcpy(raw, &m, sizeof(McDesc));      // grid, box, slab layers (McDesc holds the GridDesc)
        key.insert(key.end(), raw, raw + (sizeof(McDesc) + 3) / 4);
        const uint32_t extra[] = {(uint32_t)s.prune, (uint32_t)s.specialize, (uint32_t)g.rank, (uint32_t)g.world, (uint32_t)g.emulate, (uint32_t)g.halo,
                                  (uint32_t)g.own_zb, (uint32_t)g.own_ze, (uint32_t)(uintptr_t)g.sdf.v, (uint32_t)((uintptr_t)g.sdf.v >> 32), (uint32_t)g.sdf.nx,
                                  (uint32_t)s.overall_repeats, (uint32_t)s.vresampl_iters, (uint32_t)s.projection_enabled, (uint32_t)s.qem_enabled,
                                  (uint32_t)s.exact_reduce, (uint32_t)s.subdiv_enabled, (uint32_t)s.point_spec};
        key.insert(key.end(), extra, extra + sizeof(extra) / 4);
        uint32_t cbits;
        memcpy(&cbits, &s.vresampl_c, 4);
        key.push_back(cbits);
    }
    const bool spec_hit = s.speculate != 0 && g_spec.valid && g_spec.key == key;
    g_spec_used = spec_hit;
    unsigned h[24];
    if (spec_hit) memcpy(h, g_spec.h, sizeof(h));
    unsigned n_work = (unsigned)ntiles;
    if ((size_t)ntiles * 64 * 16 > ((size_t)512 << 20)) {
        if (spec_hit) {
            n_work = h[8];
        } else {
            ISD_CUDA(cudaMemcpyAsync(h_actual + 8, tile_count, sizeof(unsigned), cudaMemcpyDeviceToHost, g.stream));
            ISD_CUDA(cudaStreamSynchronize(g.stream));
            n_work = h_actual[8];
        }
    }
    const size_t n_seg = (size_t)(n_work > 0 ? n_work : 1) * 64;
    g.d_seg_bits.reserve(n_seg * sizeof(unsigned long long));
    g.d_seg_off.reserve(n_seg * sizeof(uint2));
    McBuffers mb;
    mb.list = g.d_tile_list.as<unsigned>();
    mb.counters = tile_count;
    mb.tile_slot = hints.tile_slot; mb.tile_any = hints.tile_any; mb.tile_bits = hints.tile_bits; mb.TW = hints.TW;
    mb.seg_bits = g.d_seg_bits.as<unsigned long long>();
    mb.seg_off = g.d_seg_off.as<uint2>();
    mb.row_tot = g.d_row_tot.as<Triple>();
    mb.row_base = g.d_row_base.as<Triple>();
    {   // K2, K3 first half
        if (dbg_cases) {
            k_debug_cases<<<(unsigned)((n_rows * m.NX + 255) / 256), 256, 0, g.stream>>>(m, g.d_signs.as<uint32_t>(), dbg_cases);
            ISD_LAUNCH_CHECK();
        }
        if (gd.ty * gd.tz <= 8192) ISD_TIMED(KT_CLASSIFY, k_mc_classify_tiles<4><<<gd.ty * gd.tz, 256, 0, g.stream>>>(m, mb, g.d_signs.as<uint32_t>()));
        else ISD_TIMED(KT_CLASSIFY, k_mc_classify_tiles<1><<<gd.ty * gd.tz, 64, 0, g.stream>>>(m, mb, g.d_signs.as<uint32_t>()));
        ISD_TIMED(KT_ROW_SCAN, k_row_scan<<<scan_blocks, kRowScanThreads, 0, g.stream>>>(m, mb, rows_per_block, g.d_scan_state.as<unsigned>(), g.d_totals.as<unsigned>()));
    }
    // totals back to the host: sizes of the vertex, face and cut-cube arrays (+ the ghost layer's share, z-slabs only)
    memset(h_actual, 0, 24 * sizeof(unsigned));
    ISD_CUDA(cudaMemcpyAsync(h_actual, g.d_totals.ptr, 10 * sizeof(unsigned), cudaMemcpyDeviceToHost, g.stream));      // [8] = work-list length
    const long long own_row0 = (long long)(g.own_zb - m.scan_z0) * m.NX, own_row1 = (long long)(g.own_ze - m.scan_z0) * m.NX;
    if (g.halo_mode) {      // vertices / faces counted before the slab's first own row and before the first row above it
        ISD_CUDA(cudaMemcpyAsync(h_actual + 16, mb.row_base + own_row0, sizeof(Triple), cudaMemcpyDeviceToHost, g.stream));
        if (own_row1 < n_rows) ISD_CUDA(cudaMemcpyAsync(h_actual + 20, mb.row_base + own_row1, sizeof(Triple), cudaMemcpyDeviceToHost, g.stream));
    }
    if (!spec_hit) {
        ISD_CUDA(cudaStreamSynchronize(g.stream));
        memcpy(h, h_actual, sizeof(h));
        g_spec.valid = true;
        g_spec.key = key;
        memcpy(g_spec.h, h, sizeof(h));
        for (long long& l : g_spec.leftovers) l = -1;
    }
    if (want_point_lists && h[8] > 0) {
        const int n_instr = (int)(itab_words / 4), stride = n_instr + 1;
        const size_t list_bytes = (size_t)h[8] * 2 * stride * sizeof(uint32_t);
        if (list_bytes <= ((size_t)4 << 30)) {      // enormous trees on enormous grids: the whole tree runs instead
            g.d_point_lists.reserve(list_bytes);
            PointSpec ps;
            ps.g = gd; ps.tile_bits = hints.tile_bits; ps.tile_slot = hints.tile_slot;
            ps.lists = g.d_point_lists.as<uint32_t>();
            ps.TW = hints.TW; ps.stride = stride; ps.enabled = 1;
            const unsigned want_blocks = (h[8] + 3) / 4, cap = (unsigned)g.sm_count * 16;
            ISD_TIMED(KT_POINT_LISTS, k_build_point_lists<<<want_blocks < cap ? want_blocks : cap, 128, 0, g.stream>>>(tile_count, spec.masks, spec.mask_words, d_itab, n_instr, stride,
                                                                                             g.d_point_lists.as<uint32_t>()));
            set_point_spec(&ps);
        }
    }
    const unsigned n_cubes = h[2];
    if (m.ghost_rows == 0) { h[4] = 0; h[5] = 0; }
    g.d_cubes.reserve((size_t)(n_cubes > 0 ? n_cubes : 1) * sizeof(unsigned long long));
    if (n_work > 0) {   // K3, second half: absolute segment offsets + the ordered list of cut cubes
        unsigned blocks = (unsigned)((n_seg + 255) / 256);
        if (blocks > (unsigned)g.sm_count * 32u) blocks = (unsigned)g.sm_count * 32u;
        // faces grouped by tile, for the OB02 pair kernels: only when they will run on this very mesh (per-tile lists on, the
        // local mesh is what K4 emits)
        PermArgs pa{nullptr, nullptr, nullptr};
        g.perm_valid = false;
        const bool local_mesh = g.world == 1 || g.emulate || g.halo_mode;
        if (g.want_perm && g.point_spec_on && g.pair_kernels && local_mesh && s.overall_repeats > 0 && (s.vresampl_iters > 0 || s.projection_enabled)) {
            const size_t slots = (size_t)h[1] + (size_t)h[8] + 2;
            g.d_perm.reserve(16 + slots * sizeof(int));
            ISD_CUDA(cudaMemsetAsync(g.d_perm.ptr, 0, 16, g.stream));
            ISD_CUDA(cudaMemsetAsync(g.d_perm.as<char>() + 16, 0xff, slots * sizeof(int), g.stream));
            pa.cursor = g.d_perm.as<unsigned>();
            pa.perm = reinterpret_cast<int*>(g.d_perm.as<char>() + 16);
            pa.ghost_faces = m.ghost_rows > 0 ? g.d_totals.as<unsigned>() + 5 : nullptr;
            g.perm_valid = true;
            g.perm_units_bound = (long long)(slots / 2) + 1;
        }
        ISD_TIMED(KT_SEG_APPLY, k_seg_apply<<<blocks, 256, 0, g.stream>>>(m, mb, g.d_cubes.as<unsigned long long>(), pa));
    }
    ISD_CUDA(cudaEventRecord(g.ev[2], g.stream));
    const long long scan_v = h[0], scan_f = h[1];
    const long long ghost_v = h[4], ghost_f = h[5];
    long long nv = scan_v - ghost_v, nf = scan_f - ghost_f;
    if (g.halo_mode) {      // what the slab owns, and what it tells the others
        g.own_v0 = h[16]; g.own_f0 = h[17];
        g.own_v1 = (own_row1 < n_rows) ? h[20] : scan_v;
        g.own_f1 = (own_row1 < n_rows) ? h[21] : scan_f;
        nv = g.own_v1 - g.own_v0; nf = g.own_f1 - g.own_f0;
    }

    long long voff = 0, foff = 0, gnv = nv, gnf = nf;
    g.halo_pending = false;
    if (g.world > 1) {
        if (g.emulate) {
            if (g.halo_mode) { gnv = -1; gnf = g.emu_gnf; }
            voff = g.emu_voff; foff = g.emu_foff;
        } else if (g.halo_mode) {
            // Nothing is exchanged here.  The slab's {nv, nf} travel with its edge-length sum in the ONE collective of the step (after
            // k_face_prepare, projection_step) — or, when no projection runs, in a gather of their own when the slab is exported.
            g.halo_counts[0] = nv; g.halo_counts[1] = nf;
            g.halo_gather_enqueued = false;
            g.halo_pending = true;
            gnv = -1; gnf = -1;
        } else {
            // the one collective of the path: {nv, nf} of every slab (16 bytes per rank)
            g.halo_counts[0] = nv; g.halo_counts[1] = nf;
            enqueue_slab_gather(nullptr);
            ISD_CUDA(cudaStreamSynchronize(g.stream));
            const int64_t* hs = g.h_gather.as<int64_t>();
            gnv = 0; gnf = 0;
            for (int r = 0; r < g.world; ++r) {
                if (r == g.rank) { voff = gnv; foff = gnf; }
                gnv += hs[3 + kGatherWords * r]; gnf += hs[4 + kGatherWords * r];
            }
        }
    }
    if (gnv >= (1ll << 31) || gnf >= (1ll << 31)) {
        fprintf(stderr, "implisolid_b200: mesh too large for 32-bit indices (%lld vertices, %lld faces)\n", gnv, gnf);
        abort();
    }
    const bool in_place = g.world > 1 && !g.emulate && !g.halo_mode;      // room for every slab: the gather is then a set of in-place broadcasts
    g.local_voff = in_place ? voff : 0;
    g.local_foff = in_place ? foff : 0;
    const long long store_v = g.halo_mode ? scan_v : (in_place ? gnv : nv), store_f = g.halo_mode ? scan_f - ghost_f : (in_place ? gnf : nf);
    g.d_verts.reserve((size_t)(store_v > 0 ? store_v : 1) * 3 * sizeof(float));
    g.d_faces.reserve((size_t)(store_f > 0 ? store_f : 1) * 3 * sizeof(int));
    if (n_cubes > 0) {   // K4
        EmitArgs ea;
        ea.vbase = voff - ghost_v; ea.fbase = foff - ghost_f;
        ea.v_local0 = ghost_v; ea.f_local0 = ghost_f;
        ea.ghost_vertices = g.halo_mode ? 1 : 0;
        if (g.halo_mode) { ea.vbase = 0; ea.fbase = 0; ea.v_local0 = 0; }      // local numbering = scan order, ghost vertices first
        const unsigned blocks = (n_cubes + kEmitThreads - 1) / kEmitThreads;      // one thread per cut cube (a cap of 16 blocks per SM measured 3 % slower)
        ISD_TIMED(KT_EMIT, k_mc_emit<<<blocks, kEmitThreads, 0, g.stream>>>(m, mb, ea, g.d_field.as<float>(), g.d_signs.as<uint32_t>(),
                                                         g.d_cubes.as<unsigned long long>(), g.d_totals.as<unsigned>() + 2,
                                                         g.d_verts.as<float>() + g.local_voff * 3, g.d_faces.as<int>() + g.local_foff * 3));
    }
    ISD_CUDA(cudaEventRecord(g.ev[3], g.stream));
    g.nv = nv; g.nf = nf;
    if (g.halo_mode) {      // the local mesh the OB02 stages work on; export_own_slab() narrows it to the slab afterwards
        g.nv = scan_v; g.nf = scan_f - ghost_f;
        g.own_f0 -= ghost_f; g.own_f1 -= ghost_f;
    }
    g.gathered = false;
    g.topo_valid = false;
    if (n_cubes > 0 && (g.world == 1 || g.halo_mode)) {      // the local mesh is what K4 emitted, numbered in scan order
        g.topo_m = m; g.topo_mb = mb;
        g.topo_ea.vbase = 0; g.topo_ea.fbase = 0; g.topo_ea.v_local0 = 0; g.topo_ea.f_local0 = g.halo_mode ? ghost_f : 0; g.topo_ea.ghost_vertices = 0;
        g.topo_valid = true;
    }
    g.slab[0] = g.own_zb; g.slab[1] = g.own_ze; g.slab[2] = voff; g.slab[3] = foff;
    g.slab[4] = nv; g.slab[5] = nf; g.slab[6] = gnv; g.slab[7] = gnf;
    g.counters[0] = (int64_t)gd.S * gd.S * gd.nz;
    g.counters[1] = (int64_t)(m.zc_end - m.emit_z0) * m.NX * m.NX;
    g.counters[2] = n_cubes;
    g.counters[14] = (int64_t)field_bytes;
    g.counters[15] = (int64_t)h[8] * 512;
    (void)prog;
}

// all ranks end up with the whole stitched mesh (vertices of rank r at [voff_r, voff_r + nv_r))
void gather_mesh() {
    if (g.world == 1 || g.gathered || g.emulate) return;
    int64_t* hs = g.h_gather.as<int64_t>();
    const long long gnv = g.slab[6], gnf = g.slab[7];
    ISD_CUDA(cudaEventRecord(g.ev[13], g.stream));
    if (g.halo_mode) {      // the exported slab sits inside local arrays: move it to its global offset in arrays sized for the whole mesh
        g.d_verts2.reserve((size_t)(gnv > 0 ? gnv : 1) * 3 * sizeof(float));
        g.d_faces2.reserve((size_t)(gnf > 0 ? gnf : 1) * 3 * sizeof(int));
        if (g.nv) ISD_CUDA(cudaMemcpyAsync(g.d_verts2.as<float>() + g.slab[2] * 3, g.d_verts.as<float>() + g.local_voff * 3, (size_t)g.nv * 3 * sizeof(float), cudaMemcpyDeviceToDevice, g.stream));
        if (g.nf) ISD_CUDA(cudaMemcpyAsync(g.d_faces2.as<int>() + g.slab[3] * 3, g.d_faces.as<int>() + g.local_foff * 3, (size_t)g.nf * 3 * sizeof(int), cudaMemcpyDeviceToDevice, g.stream));
        std::swap(g.d_verts, g.d_verts2);
        std::swap(g.d_faces, g.d_faces2);
    }
    g.nccl.check(g.nccl.GroupStart(), "ncclGroupStart");
    long long vo = 0, fo = 0;
    for (int r = 0; r < g.world; ++r) {      // in place: every rank's slab already sits at its global offset
        const long long rv = hs[3 + kGatherWords * r], rf = hs[4 + kGatherWords * r];
        float* pv = g.d_verts.as<float>() + vo * 3;
        int* pf = g.d_faces.as<int>() + fo * 3;
        if (rv > 0) g.nccl.check(g.nccl.Broadcast(pv, pv, (size_t)rv * 3, kNcclFloat32, r, g.comm, g.stream), "ncclBroadcast");
        if (rf > 0) g.nccl.check(g.nccl.Broadcast(pf, pf, (size_t)rf * 3, kNcclInt32, r, g.comm, g.stream), "ncclBroadcast");
        vo += rv; fo += rf;
    }
    g.nccl.check(g.nccl.GroupEnd(), "ncclGroupEnd");
    ISD_CUDA(cudaEventRecord(g.ev[14], g.stream));
    if (getenv("ISD_GATHER_SYNC")) ISD_CUDA(cudaStreamSynchronize(g.stream));
    g.local_voff = 0; g.local_foff = 0;
    g.nv = gnv; g.nf = gnf;
    g.gathered = true;
    g.perm_valid = false;
    g.topo_valid = false;
    g.host_valid = false;
}

// ---------------------------------------------------------------------------------------------------------------------
// OB02 stages
// ---------------------------------------------------------------------------------------------------------------------
void exclusive_scan_u32(const unsigned* in, long long n, unsigned* out /* n + 1 */, int kt_id = KT_ADJACENCY) {
    if (n == 0) { ISD_CUDA(cudaMemsetAsync(out, 0, sizeof(unsigned), g.stream)); return; }
    int blocks = (int)((n + kUScanThreads * 4 - 1) / (kUScanThreads * 4));      // >= 4 rounds per block: fewer blocks to wait for
    if (blocks > kUScanMaxBlocks) blocks = kUScanMaxBlocks;
    const long long per_block = ((n + blocks - 1) / blocks + kUScanThreads - 1) / kUScanThreads * kUScanThreads;
    const size_t state_bytes = (4 + 2 * (size_t)kUScanMaxBlocks) * sizeof(unsigned);
    g.d_upartials.reserve(state_bytes);
    ISD_CUDA(cudaMemsetAsync(g.d_upartials.ptr, 0, state_bytes, g.stream));
    ISD_TIMED(kt_id, k_uscan<<<blocks, kUScanThreads, 0, g.stream>>>(in, n, per_block, g.d_upartials.as<unsigned>(), out));
}

inline unsigned blocks_for(long long n, int threads) { return (unsigned)((n + threads - 1) / threads); }
// grid of k_project_leftovers (a warp takes 1-4 centroids per round, grid-stride)
#ifndef ISD_LEFT_CAP
#define ISD_LEFT_CAP 64      // blocks per SM at most: 0.105 ms at C4 1024^3 against 0.111 with 16 and 0.137 with 8 (uneven work: finer blocks balance better)
#endif
inline unsigned left_grid(long long n_left) {
    const unsigned want = blocks_for(n_left, 4), cap = (unsigned)g.sm_count * ISD_LEFT_CAP;
    return want < cap ? want : cap;
}

// the order in which the pair kernels take the faces (ob02_kernels.cuh FaceOrder) and the number of units to launch for
FaceOrder face_order() {
    if (!g.perm_valid) return FaceOrder{nullptr, nullptr};
    return FaceOrder{reinterpret_cast<const int*>(g.d_perm.as<char>() + 16), g.d_perm.as<unsigned>()};
}
long long unit_launch_count(long long nf) { return g.perm_valid ? g.perm_units_bound : (nf + 1) / 2; }

// umbrella CSR (always) and faces-of-faces (resampling only); the face list does not change between OB02 steps
void build_adjacency(bool need_fof) {
    const int nv = (int)g.nv, nf = (int)g.nf;
    if (!g.adjacency_valid && g.topo_valid && g.want_topo && nf > 0) {
        // the mesh is K4's own: umbrellas (and faces-of-faces) straight from the case codes — two passes over the cut cubes and one scan
        g.d_deg.reserve((size_t)(nv + 1) * sizeof(unsigned));
        g.d_uoff.reserve((size_t)(nv + 2) * sizeof(unsigned));
        g.d_uitems.reserve((size_t)nf * 3 * sizeof(int) + 16);
        if (need_fof) g.d_fof.reserve((size_t)nf * 3 * sizeof(int) + 16);
        g_kernel_names[KT_ADJACENCY] = "adjacency(k_mc_adjacency<0>+k_uscan+k_mc_adjacency<1>)";
        const unsigned n_cubes = (unsigned)g.counters[2];
        // one thread per cut cube, no grid-stride cap: measured 0.343 against 0.394 ms at C4 1024^3 with the grid capped at 16 blocks per SM
        const unsigned blocks = (n_cubes + kAdjThreads - 1) / kAdjThreads;
        const unsigned* n_ptr = g.d_totals.as<unsigned>() + 2;
        ISD_TIMED(KT_ADJACENCY, k_mc_adjacency<false><<<blocks, kAdjThreads, 0, g.stream>>>(g.topo_m, g.topo_mb, g.topo_ea, g.d_signs.as<uint32_t>(), g.d_cubes.as<unsigned long long>(),
                                                                                   n_ptr, g.d_deg.as<unsigned>(), nullptr, nullptr, nullptr));
        exclusive_scan_u32(g.d_deg.as<unsigned>(), nv, g.d_uoff.as<unsigned>());
        ISD_TIMED(KT_ADJACENCY, k_mc_adjacency<true><<<blocks, kAdjThreads, 0, g.stream>>>(g.topo_m, g.topo_mb, g.topo_ea, g.d_signs.as<uint32_t>(), g.d_cubes.as<unsigned long long>(),
                                                                                  n_ptr, nullptr, g.d_uoff.as<unsigned>(), g.d_uitems.as<int>(),
                                                                                  need_fof ? g.d_fof.as<int>() : nullptr));
        g.adjacency_valid = true;
        g.fof_valid_ = need_fof;
    }
    if (!g.adjacency_valid) {
        g_kernel_names[KT_ADJACENCY] = "adjacency(k_vertex_degree+k_uscan+k_umbrella_fill+k_umbrella_sort)";
        g.d_deg.reserve((size_t)(nv + 1) * sizeof(unsigned));
        g.d_cursor.reserve((size_t)(nv + 1) * sizeof(unsigned));
        g.d_uoff.reserve((size_t)(nv + 2) * sizeof(unsigned));
        g.d_uitems.reserve((size_t)nf * 3 * sizeof(int) + 16);
        ISD_CUDA(cudaMemsetAsync(g.d_deg.ptr, 0, (size_t)(nv + 1) * sizeof(unsigned), g.stream));
        ISD_CUDA(cudaMemsetAsync(g.d_cursor.ptr, 0, (size_t)(nv + 1) * sizeof(unsigned), g.stream));
        ISD_TIMED(KT_ADJACENCY, k_vertex_degree<<<blocks_for((long long)nf * 3, 256), 256, 0, g.stream>>>(g.d_faces.as<int>(), nf, g.d_deg.as<unsigned>()));
        exclusive_scan_u32(g.d_deg.as<unsigned>(), nv, g.d_uoff.as<unsigned>());
        ISD_TIMED(KT_ADJACENCY, k_umbrella_fill<<<blocks_for((long long)nf * 3, 256), 256, 0, g.stream>>>(g.d_faces.as<int>(), nf, g.d_uoff.as<unsigned>(),
                                                                                 g.d_cursor.as<unsigned>(), g.d_uitems.as<int>()));
        ISD_TIMED(KT_ADJACENCY, k_umbrella_sort<<<blocks_for(nv, 128), 128, 0, g.stream>>>(nv, g.d_uoff.as<unsigned>(), g.d_uitems.as<int>()));
        g.adjacency_valid = true;
        g.fof_valid_ = false;
    }
    if (need_fof && !g.fof_valid_) {
        g.d_fof.reserve((size_t)nf * 3 * sizeof(int) + 16);
        ISD_TIMED(KT_FOF, k_faces_of_faces<<<blocks_for((long long)nf * 3, 256), 256, 0, g.stream>>>(g.d_faces.as<int>(), nf, g.d_uoff.as<unsigned>(),
                                                                                  g.d_uitems.as<int>(), g.d_fof.as<int>()));
        g.fof_valid_ = true;
    }
}

void compute_centroids() {
    const int nf = (int)g.nf;
    g.d_cen.reserve((size_t)nf * 3 * sizeof(float) + 16);
    k_centroids<<<blocks_for(nf, 128), 128, 0, g.stream>>>(g.d_verts.as<float>(), g.d_faces.as<int>(), nf, g.d_cen.as<float>());
    ISD_LAUNCH_CHECK();
}

// polygonize_step_1 (polygonizer_algorithm_ob02.hpp:94-106) -> process2_vertex_resampling_relaxation_v1
void vertex_resampling_step(float c) {
    const int nv = (int)g.nv, nf = (int)g.nf;
    if (nf == 0) return;
    build_adjacency(true);
    g.d_cen.reserve((size_t)nf * 3 * sizeof(float) + 16);
    g.d_nrm.reserve((size_t)nf * 3 * sizeof(float) + 16);
    g.d_w.reserve((size_t)nf * sizeof(float) + 16);
    g_kernel_names[KT_CENTROID_GRAD] = (g.point_spec_on && g.pair_kernels) ? "k_centroid_gradients_pair" : "k_centroid_gradients";
    if (g.point_spec_on && g.pair_kernels)
        ISD_TIMED(KT_CENTROID_GRAD, k_centroid_gradients_pair<<<blocks_for(unit_launch_count(nf), 128), 128, 0, g.stream>>>(g.sdf, g.d_verts.as<float>(), g.d_faces.as<int>(), nf,
                                                                                                                  g.d_cen.as<float>(), g.d_nrm.as<float>(), face_order()));
    else
        ISD_TIMED(KT_CENTROID_GRAD, (g.point_spec_on ? k_centroid_gradients<true> : k_centroid_gradients<false>)<<<blocks_for(nf, 128), 128, 0, g.stream>>>(g.sdf, g.d_verts.as<float>(), g.d_faces.as<int>(), nf, g.d_cen.as<float>(),
                                                                   g.d_nrm.as<float>()));
    ISD_TIMED(KT_FACE_WEIGHTS, k_face_weights<<<blocks_for(nf, 128), 128, 0, g.stream>>>(g.d_cen.as<float>(), g.d_nrm.as<float>(), g.d_fof.as<int>(), nf, c, g.d_w.as<float>()));
    ISD_TIMED(KT_RESAMPLE, k_resample<<<blocks_for(nv, 128), 128, 0, g.stream>>>(nv, g.d_uoff.as<unsigned>(), g.d_uitems.as<int>(), g.d_w.as<float>(), g.d_cen.as<float>(),
                                                          g.d_verts.as<float>()));
}

// make_alpha_list (centroids_projection.cpp:144-194)
// boost::random::mt11213b(seed 12) + uniform_01<float>, as make_random_pm1 uses them
// (vectorised_algorithms/make_random_pm1.hpp:15-28); published Matsumoto-Nishimura parameters.
struct Mt11213b {
    enum { N = 351, M = 175 };
    uint32_t x[N];
    int i;
    explicit Mt11213b(uint32_t seed) {
        x[0] = seed;
        for (int k = 1; k < N; ++k) x[k] = 1812433253u * (x[k - 1] ^ (x[k - 1] >> 30)) + (uint32_t)k;
        i = N;
    }
    uint32_t next() {
        if (i >= N) {
            const uint32_t upper = ~0u << 19, lower = ~upper;
            for (int k = 0; k < N; ++k) {
                const uint32_t y = (x[k] & upper) | (x[(k + 1) % N] & lower);
                x[k] = x[(k + M) % N] ^ (y >> 1) ^ ((y & 1u) ? 0xccab8ee7u : 0u);
            }
            i = 0;
        }
        uint32_t z = x[i++];
        z ^= (z >> 11);
        z ^= (z << 7) & 0x31b6ab00u;
        z ^= (z << 15) & 0xffe50000u;
        z ^= (z >> 17);
        return z;
    }
    float uniform01() {
        for (;;) {
            const float factor = 1.0f / ((float)0xffffffffu + 1.0f);
            const float r = (float)next() * factor;
            if (r < 1.0f) return r;
        }
    }
};

// Halo mode: global offsets of the slab (slab[2], slab[3]) and global sizes (slab[6], slab[7]) from the all-gather enqueued after
// the scan; waits for it if nobody has yet.
void resolve_halo_offsets() {
    if (!g.halo_mode) return;
    if (g.emulate) { g.slab[2] = g.emu_voff; g.slab[3] = g.emu_foff; g.slab[6] = -1; g.slab[7] = g.emu_gnf; return; }
    if (!g.halo_pending) return;
    if (!g.halo_gather_enqueued) enqueue_slab_gather(nullptr);      // no projection in this build: the counts travel alone
    ISD_CUDA(cudaStreamSynchronize(g.stream));
    const int64_t* hs = g.h_gather.as<int64_t>();
    long long voff = 0, foff = 0, gnv = 0, gnf = 0;
    for (int r = 0; r < g.world; ++r) {
        if (r == g.rank) { voff = gnv; foff = gnf; }
        gnv += hs[3 + kGatherWords * r]; gnf += hs[4 + kGatherWords * r];
    }
    g.halo_pending = false;
    if (gnv >= (1ll << 31) || gnf >= (1ll << 31)) {
        fprintf(stderr, "implisolid_b200: mesh too large for 32-bit indices (%lld vertices, %lld faces)\n", gnv, gnf);
        abort();
    }
    g.slab[2] = voff; g.slab[3] = foff; g.slab[6] = gnv; g.slab[7] = gnf;
}

// make_random_pm1(n, 3, 1e-6) seeds a fresh mt11213b(12) on every call (make_random_pm1.hpp:15-28), so draw j is the same
// number in every build: the stream is generated once (sequentially, on the host: a Mersenne twister with data-dependent
// rejections in uniform_01) and kept in HBM; a larger mesh only generates the tail it is missing.
const float* noise_stream(size_t count) {
    static Mt11213b rng(12);
    static size_t have = 0;
    if (count > have) {
        const size_t want = count + count / 4;
        DeviceBuffer bigger;
        bigger.reserve(want * sizeof(float));
        if (have) ISD_CUDA(cudaMemcpyAsync(bigger.ptr, g.d_noise.ptr, have * sizeof(float), cudaMemcpyDeviceToDevice, g.stream));
        std::vector<float> tail(want - have);
        const float amplitude = 0.000001f;
        for (size_t i = 0; i < tail.size(); ++i) tail[i] = (float)((rng.uniform01() * 2.0 - 1.0) * amplitude);
        ISD_CUDA(cudaMemcpyAsync(bigger.as<float>() + have, tail.data(), tail.size() * sizeof(float), cudaMemcpyHostToDevice, g.stream));
        ISD_CUDA(cudaStreamSynchronize(g.stream));      // `tail` is pageable and dies here
        g.d_noise.release();
        g.d_noise = bigger;
        have = want;
    }
    return g.d_noise.as<float>();
}

// polygonize_step_2 (polygonizer_algorithm_ob02.hpp:108-116) -> centroids_projection (centroids_projection.cpp:1219-1311)
// stage times and counters of projection steps that did not wait for the device (speculative builds): resolved by
// finish_deferred_projection() after the build's final synchronisation
struct DeferredSpan { int stage; cudaEvent_t a, b; };
std::vector<DeferredSpan> g_deferred_spans;
std::vector<int> g_deferred_slots;      // counter slots (= overall repeat) of the projection steps that ran speculatively

void projection_step(bool enable_qem, bool exact_reduce, int repeat) {
    const int nv = (int)g.nv, nf = (int)g.nf;
    if (nf == 0) {
        // A slab (plus halo) without faces still owes the other ranks its triple: every rank enqueues the collective once per
        // projection step, whatever it holds.
        if (g.halo_mode && !g.emulate && g.world > 1) {
            enqueue_slab_gather(nullptr);
        }
        return;
    }
    ISD_CUDA(cudaEventRecord(g.ev[6], g.stream));
    cudaEvent_t ev_start = nullptr;
    if (g_spec_used) { ev_start = g_kt.get(); ISD_CUDA(cudaEventRecord(ev_start, g.stream)); }
    const unsigned prep_blocks = blocks_for(nf, 256);
    {
        const void* before = g.d_scal.ptr;
        g.d_scal.reserve(sizeof(AlphaList) + (size_t)prep_blocks * sizeof(unsigned long long) + 64);
        if (g.d_scal.ptr != before) g.alpha_ticket_zeroed = false;      // a fresh allocation holds a garbage ticket
    }
    g.d_counters.reserve(sizeof(Ob02Counters));
    g.d_cen.reserve((size_t)nf * 3 * sizeof(float) + 16);
    AlphaList* d_al = g.d_scal.as<AlphaList>();
    unsigned long long* d_partials = reinterpret_cast<unsigned long long*>(g.d_scal.as<char>() + ((sizeof(AlphaList) + 15) & ~(size_t)15));
    if (!g.alpha_ticket_zeroed) {      // the last block of k_face_prepare leaves the ticket at zero again
        ISD_CUDA(cudaMemsetAsync(d_al, 0, sizeof(AlphaList), g.stream));
        g.alpha_ticket_zeroed = true;
    }
    float* d_len3 = nullptr;
    if (exact_reduce) {
        g.d_len3.reserve((size_t)nf * 3 * sizeof(float) + 16);
        d_len3 = g.d_len3.as<float>();
    }
    // centroids + mean edge length + alpha list, all on the device (centroids_projection.cpp:1230-1245)
    if (g.halo_mode) {
        // only the slab's own faces are summed; the all-reduce (8 bytes) gives every GPU the whole mesh's sum, in the stream
        ISD_TIMED(KT_FACE_PREPARE, k_face_prepare<<<prep_blocks, 256, 0, g.stream>>>(g.d_verts.as<float>(), g.d_faces.as<int>(), nf, g.d_cen.as<float>(), d_len3, d_partials, d_al,
                                                          (int)g.own_f0, (int)g.own_f1, 0));
        if (g.emulate) {
            ISD_CUDA(cudaMemcpyAsync(g.h_small.as<char>() + 3200, &d_al->edge_sum, sizeof(unsigned long long), cudaMemcpyDeviceToHost, g.stream));
            memcpy(g.h_small.as<char>() + 3208, &g.emu_edge_total, sizeof(unsigned long long));
            ISD_CUDA(cudaMemcpyAsync(&d_al->edge_total, g.h_small.as<char>() + 3208, sizeof(unsigned long long), cudaMemcpyHostToDevice, g.stream));
            k_alpha_from_sum<<<1, 1, 0, g.stream>>>(d_al, &d_al->edge_total, nullptr, g.world, g.emu_gnf);
        } else {
            enqueue_slab_gather(&d_al->edge_sum);
            k_alpha_from_gather<<<1, 1, 0, g.stream>>>(d_al, reinterpret_cast<const long long*>(g.d_gather.as<int64_t>()), g.world);
        }
        ISD_LAUNCH_CHECK();
    } else {
        ISD_TIMED(KT_FACE_PREPARE, k_face_prepare<<<prep_blocks, 256, 0, g.stream>>>(g.d_verts.as<float>(), g.d_faces.as<int>(), nf, g.d_cen.as<float>(), d_len3, d_partials, d_al,
                                                          0, nf, 1));
        if (g.world == 1 || g.emulate)      // the sum of a whole mesh: what emulated ranks are handed as the all-reduce result
            ISD_CUDA(cudaMemcpyAsync(g.h_small.as<char>() + 3200, &d_al->edge_sum, sizeof(unsigned long long), cudaMemcpyDeviceToHost, g.stream));
    }
    if (exact_reduce) {
        k_sum_serial_float<<<1, 32, 0, g.stream>>>(d_len3, (long long)nf * 3, nf, &d_al->avg_edge);
        ISD_LAUNCH_CHECK();
        k_alpha_from_mean<<<1, 1, 0, g.stream>>>(&d_al->avg_edge, d_al);
        ISD_LAUNCH_CHECK();
    }
    g.d_best.reserve((size_t)nf * sizeof(float4) + 16);
    g.d_fc.reserve((size_t)nf * sizeof(float) + 16);
    g.d_state.reserve((size_t)nf + 16);
    g.d_proj.reserve((size_t)nf * 3 * sizeof(float) + 16);
    g.d_pnrm.reserve((size_t)nf * 3 * sizeof(float) + 16);
    ISD_CUDA(cudaMemsetAsync(g.d_counters.ptr, 0, sizeof(Ob02Counters), g.stream));
    ProjectArgs a;
    a.verts = g.d_verts.as<float>(); a.faces = g.d_faces.as<int>(); a.cen = g.d_cen.as<float>(); a.noise = nullptr;
    a.al = d_al; a.best = g.d_best.as<float4>(); a.fc = g.d_fc.as<float>(); a.state = g.d_state.as<unsigned char>();
    a.out = g.d_proj.as<float>(); a.nrm_out = g.d_pnrm.as<float>(); a.counters = g.d_counters.as<Ob02Counters>();
    a.n = nf; a.speculative = 1; a.order = face_order();
    {   // units per ticket of the lane-refill kernels.  A ticket is a warp's serial work: where the projection is hard (C4's band of
        // failing first directions) one range of 128 units kept a warp busy for a millisecond while the rest of a thin z-slab's grid
        // had drained (k_search_trials_refill on an emulated 1/8 slab: 1.06 ms with 128, 0.30 ms with 32; whole mesh 0.222 / 0.205 /
        // 0.213 ms with 128 / 64 / 32).  64, or 32 when the mesh cannot give every resident warp four tickets of 64.
        const long long warps = (long long)g.sm_count * 6 * 4, per = unit_launch_count(nf) / (warps * 4);
        a.range = per >= kRefillRange ? kRefillRange : 32;
        if (g.refill_range > 0) a.range = (g.refill_range + 31) / 32 * 32;      // "gpu":{"refill_range":..}: A/B
    }
    // the list of left-over centroids lives in the output array: k_project_finish only writes it once there are none left to list
    a.leftover = reinterpret_cast<int*>(g.d_proj.as<float>());
    const bool pairs = g.point_spec_on && g.pair_kernels;
    g_kernel_names[KT_SEARCH] = pairs ? "k_project_search_pair" : "k_project_search";
    g_kernel_names[KT_FINISH] = !pairs ? "k_project_finish" : g.finish_split == 2 ? "k_finish_bisect_refill" : g.finish_split == 1 ? "k_finish_bisect_units" : "k_project_finish_pair";
    if (pairs && g.search_split) {
        ISD_TIMED(KT_SEARCH, k_project_search_pair<true><<<blocks_for(unit_launch_count(nf), 128), 128, 0, g.stream>>>(g.sdf, a));
        const unsigned want = blocks_for(unit_launch_count(nf), 4 * a.range), resident = (unsigned)g.sm_count * ISD_TRIALS_REFILL_MINBLOCKS;
        ISD_TIMED(KT_TRIALS, k_search_trials_refill<<<want < resident ? want : resident, 128, 0, g.stream>>>(g.sdf, a));
    } else if (pairs) ISD_TIMED(KT_SEARCH, k_project_search_pair<false><<<blocks_for(unit_launch_count(nf), 128), 128, 0, g.stream>>>(g.sdf, a));
    else ISD_TIMED(KT_SEARCH, (g.point_spec_on ? k_project_search<true> : k_project_search<false>)<<<blocks_for(nf, 128), 128, 0, g.stream>>>(g.sdf, a));
    // Phase B (direction types 2..6) is only needed when phase A left centroids over, which is rare and which only the device
    // knows: the rest of the step is launched right away and does nothing in that case (the kernels test the counter), so the
    // host does not have to wait here.
    // k_project_finish: one kernel, or (pair path) bisect from per-warp queues -> gradients of all
    auto finish_pass = [&](const ProjectArgs& pa) {
        if (pairs && g.finish_split) {
            if (g.finish_split == 2) {
                const unsigned want = blocks_for(unit_launch_count(nf), 4 * pa.range), resident = (unsigned)g.sm_count * ISD_FINISH_REFILL_MINBLOCKS;
                ISD_TIMED(KT_FINISH, k_finish_bisect_refill<<<want < resident ? want : resident, 128, 0, g.stream>>>(g.sdf, pa));
            }
            else ISD_TIMED(KT_FINISH, k_finish_bisect_units<<<blocks_for((nf + 1) / 2, 4 * kFinishRange), 128, 0, g.stream>>>(g.sdf, pa));
            ISD_TIMED(KT_UNIT_GRAD, k_unit_gradients_pair<<<blocks_for(unit_launch_count(nf), 128), 128, 0, g.stream>>>(g.sdf, pa.out, nf, pa.nrm_out,
                                                                                                            pa.speculative ? pa.counters : nullptr, pa.state, pa.order));
        } else if (pairs) {
            ISD_TIMED(KT_FINISH, k_project_finish_pair<<<blocks_for(unit_launch_count(nf), 128), 128, 0, g.stream>>>(g.sdf, pa));
        } else {
            ISD_TIMED(KT_FINISH, (g.point_spec_on ? k_project_finish<true> : k_project_finish<false>)<<<blocks_for(nf, 128), 128, 0, g.stream>>>(g.sdf, pa));
        }
    };
    Ob02Counters* hc = reinterpret_cast<Ob02Counters*>(g.h_small.as<char>() + 2048) + (repeat >= 0 && repeat < 8 ? repeat : 0);
    float* h_avg = reinterpret_cast<float*>(g.h_small.as<char>() + 1024);
    if (g_spec_used && repeat >= 0 && repeat < 8 && g_spec.leftovers[repeat] >= 0) {
        // Speculative build: the previous build of this very spec tells whether phase A leaves centroids over, and how many.  The
        // whole step is enqueued without waiting; the counter comes back with the build's last synchronisation and is checked.
        const long long n_left = g_spec.leftovers[repeat];
        if (enable_qem) build_adjacency(false);
        cudaEvent_t ev_m = g_kt.get(), ev_b = g_kt.get();
        if (n_left > 0) {
            a.noise = noise_stream((size_t)(g_spec.face_base[repeat] + nf) * 3) + g_spec.face_base[repeat] * 3;
            a.speculative = 0;
            const unsigned left_blocks = left_grid(n_left);
            ISD_TIMED(KT_LEFTOVERS, (g.point_spec_on ? k_project_leftovers<true> : k_project_leftovers<false>)<<<left_blocks, 128, 0, g.stream>>>(g.sdf, a, (int)n_left));
        }
        finish_pass(a);
        ISD_CUDA(cudaEventRecord(ev_m, g.stream));
        if (enable_qem)
            ISD_TIMED(KT_QEM, k_qem<<<blocks_for(nv, 128), 128, 0, g.stream>>>(nv, g.d_uoff.as<unsigned>(), g.d_uitems.as<int>(), g.d_proj.as<float>(),
                                                            g.d_pnrm.as<float>(), d_al, n_left > 0 ? nullptr : g.d_counters.as<Ob02Counters>(), g.d_verts.as<float>()));
        ISD_CUDA(cudaEventRecord(ev_b, g.stream));
        ISD_CUDA(cudaMemcpyAsync(hc, g.d_counters.ptr, sizeof(Ob02Counters), cudaMemcpyDeviceToHost, g.stream));
        ISD_CUDA(cudaMemcpyAsync(h_avg, &d_al->avg_edge, sizeof(float), cudaMemcpyDeviceToHost, g.stream));
        g_deferred_spans.push_back(DeferredSpan{T_PROJECT, ev_start, ev_m});
        g_deferred_spans.push_back(DeferredSpan{T_QEM, ev_m, ev_b});
        g_deferred_slots.push_back(repeat);
        g.proj_n = nf;
        return;
    }
    finish_pass(a);
    ISD_CUDA(cudaEventRecord(g.ev[7], g.stream));
    if (enable_qem) {   // centroids_projection.cpp:1272-1310: the mesh only changes when QEM is on
        build_adjacency(false);
        ISD_TIMED(KT_QEM, k_qem<<<blocks_for(nv, 128), 128, 0, g.stream>>>(nv, g.d_uoff.as<unsigned>(), g.d_uitems.as<int>(), g.d_proj.as<float>(),
                                                        g.d_pnrm.as<float>(), d_al, g.d_counters.as<Ob02Counters>(), g.d_verts.as<float>()));
    }
    ISD_CUDA(cudaEventRecord(g.ev[8], g.stream));
    ISD_CUDA(cudaMemcpyAsync(hc, g.d_counters.ptr, sizeof(Ob02Counters), cudaMemcpyDeviceToHost, g.stream));
    ISD_CUDA(cudaMemcpyAsync(h_avg, &d_al->avg_edge, sizeof(float), cudaMemcpyDeviceToHost, g.stream));
    ISD_CUDA(cudaStreamSynchronize(g.stream));
    float ms = 0;
    ISD_CUDA(cudaEventElapsedTime(&ms, g.ev[6], g.ev[7]));
    g.timings[T_PROJECT] += ms;
    ISD_CUDA(cudaEventElapsedTime(&ms, g.ev[7], g.ev[8]));
    g.timings[T_QEM] += ms;
    if (repeat >= 0 && repeat < 8) { g_spec.leftovers[repeat] = (long long)hc->c[10]; g_spec.face_base[repeat] = 0; }
    if (hc->c[10] > 0) {
        // rare: direction types 2..6 for the left-overs; type 2 needs make_random_pm1(n, 1e-6), centroid i <- draws 3i..3i+2
        ISD_CUDA(cudaEventRecord(g.ev[6], g.stream));
        // centroid i draws 3i..3i+2 of the stream, i being the face's index in the WHOLE mesh
        long long face_base = 0;
        if (g.halo_mode) { resolve_halo_offsets(); face_base = g.slab[3] - g.own_f0; }
        if (repeat >= 0 && repeat < 8) g_spec.face_base[repeat] = face_base;
        a.noise = noise_stream((size_t)(face_base + nf) * 3) + face_base * 3;
        a.speculative = 0;
        const int n_left = (int)hc->c[10];
        const unsigned left_blocks = left_grid(n_left);
        ISD_TIMED(KT_LEFTOVERS, (g.point_spec_on ? k_project_leftovers<true> : k_project_leftovers<false>)<<<left_blocks, 128, 0, g.stream>>>(g.sdf, a, n_left));
        finish_pass(a);
        ISD_CUDA(cudaEventRecord(g.ev[7], g.stream));
        if (enable_qem) {
            ISD_TIMED(KT_QEM, k_qem<<<blocks_for(nv, 128), 128, 0, g.stream>>>(nv, g.d_uoff.as<unsigned>(), g.d_uitems.as<int>(), g.d_proj.as<float>(),
                                                            g.d_pnrm.as<float>(), d_al, nullptr, g.d_verts.as<float>()));
        }
        ISD_CUDA(cudaEventRecord(g.ev[8], g.stream));
        ISD_CUDA(cudaMemcpyAsync(hc, g.d_counters.ptr, sizeof(Ob02Counters), cudaMemcpyDeviceToHost, g.stream));
        ISD_CUDA(cudaStreamSynchronize(g.stream));
        ISD_CUDA(cudaEventElapsedTime(&ms, g.ev[6], g.ev[7]));
        g.timings[T_PROJECT] += ms;
        ISD_CUDA(cudaEventElapsedTime(&ms, g.ev[7], g.ev[8]));
        g.timings[T_QEM] += ms;
    }
    g.average_edge = *h_avg;
    if (g.world == 1 || g.emulate) memcpy(&g.last_edge_sum, g.h_small.as<char>() + 3200, sizeof(unsigned long long));
    g.proj_n = nf;
    for (int i = 0; i < 7; ++i) g.counters[3 + i] += (int64_t)hc->c[i];
    g.counters[10] += (int64_t)hc->c[7];
    g.counters[11] += (int64_t)hc->c[8];
    g.counters[12] += (int64_t)hc->c[9];
}

// ---- polygonize_step_3 ---------------------------------------------------------------------------------------------------
// glibc's rand(): random_r.c TYPE_3 (x^31 + x^3 + 1 additive feedback), srandom_r's Lehmer fill (16807, Schrage) and its 310
// discarded outputs.  The reference calls rand() without ever seeding it, so its stream is glibc's seed-1 stream running on
// from build to build; a private copy keeps other users of the process-wide rand() out of it.
struct GlibcRand {
    uint32_t ring[31];
    int f = 3, b = 0;
    unsigned long long pending_skips = 0;      // draws made by builds with zero amplitude: their values are never needed
    GlibcRand() { seed(1); }
    void seed(unsigned s) {
        int32_t word = (int32_t)(s ? s : 1);
        ring[0] = (uint32_t)word;
        for (int i = 1; i < 31; ++i) {
            const long hi = word / 127773, lo = word % 127773;
            word = (int32_t)(16807 * lo - 2836 * hi);
            if (word < 0) word += 2147483647;
            ring[i] = (uint32_t)word;
        }
        f = 3; b = 0; pending_skips = 0;
        for (int i = 0; i < 310; ++i) next();
    }
    int next() {
        ring[f] += ring[b];
        const uint32_t r = ring[f] >> 1;
        if (++f >= 31) f = 0;
        if (++b >= 31) b = 0;
        return (int)r;
    }
    void catch_up() { for (; pending_skips; --pending_skips) next(); }
} g_rand;

// 1-to-4 subdivision of every face + randomize_verts (subdiv_kernels.cuh).  One device->host read (the number of new vertices).
void subdivision_step(float amplitude) {
    const int nv = (int)g.nv, nf = (int)g.nf;
    if (nf == 0) return;
    if ((long long)nf * 4 * 3 > 2147483647LL) { fprintf(stderr, "implisolid_b200: mesh too large to subdivide (32-bit face indices)\n"); abort(); }
    ISD_CUDA(cudaEventRecord(g.ev[11], g.stream));
    build_adjacency(false);
    const long long n3 = (long long)nf * 3;
    g.d_sub_owner.reserve((size_t)n3 * sizeof(unsigned));
    g.d_sub_flag.reserve((size_t)n3 * sizeof(unsigned));
    g.d_sub_rank.reserve((size_t)(n3 + 1) * sizeof(unsigned));
    k_subdiv_mark<<<blocks_for(n3, 256), 256, 0, g.stream>>>(g.d_faces.as<int>(), nf, g.d_uoff.as<unsigned>(), g.d_uitems.as<int>(),
                                                            g.d_sub_owner.as<unsigned>(), g.d_sub_flag.as<unsigned>());
    ISD_LAUNCH_CHECK();
    exclusive_scan_u32(g.d_sub_flag.as<unsigned>(), n3, g.d_sub_rank.as<unsigned>(), KT_OTHER);
    unsigned* h_new = reinterpret_cast<unsigned*>(g.h_small.as<char>() + 3072);
    ISD_CUDA(cudaMemcpyAsync(h_new, g.d_sub_rank.as<unsigned>() + n3, sizeof(unsigned), cudaMemcpyDeviceToHost, g.stream));
    ISD_CUDA(cudaStreamSynchronize(g.stream));
    const long long nv2 = (long long)nv + (long long)*h_new, nf2 = (long long)nf * 4;
    g.d_verts2.reserve((size_t)nv2 * 3 * sizeof(float) + 16);
    g.d_faces2.reserve((size_t)nf2 * 3 * sizeof(int) + 16);
    ISD_CUDA(cudaMemcpyAsync(g.d_verts2.ptr, g.d_verts.ptr, (size_t)nv * 3 * sizeof(float), cudaMemcpyDeviceToDevice, g.stream));
    k_subdiv_emit<<<blocks_for(nf, 128), 128, 0, g.stream>>>(g.d_verts.as<float>(), nv, g.d_faces.as<int>(), nf, g.d_sub_owner.as<unsigned>(),
                                                            g.d_sub_rank.as<unsigned>(), g.d_verts2.as<float>(), g.d_faces2.as<int>());
    ISD_LAUNCH_CHECK();
    std::swap(g.d_verts, g.d_verts2);
    std::swap(g.d_faces, g.d_faces2);
    g.nv = nv2; g.nf = nf2;
    g.adjacency_valid = false; g.fof_valid_ = false; g.perm_valid = false; g.topo_valid = false;
    g.host_valid = false;
    // randomize_verts draws 3*nv2 numbers whatever the amplitude
    const long long draws = nv2 * 3;
    if (amplitude == 0.0f) {
        g_rand.pending_skips += (unsigned long long)draws;      // v + (+-0) == v
    } else {
        g_rand.catch_up();
        std::vector<float> ratio((size_t)draws);
        for (long long i = 0; i < draws; ++i) ratio[(size_t)i] = static_cast<float>(g_rand.next()) / static_cast<float>(2147483647);
        g.d_ratio.reserve((size_t)draws * sizeof(float));
        ISD_CUDA(cudaMemcpyAsync(g.d_ratio.ptr, ratio.data(), (size_t)draws * sizeof(float), cudaMemcpyHostToDevice, g.stream));
        k_add_noise<<<blocks_for(draws, 256), 256, 0, g.stream>>>(g.d_verts.as<float>(), draws, g.d_ratio.as<float>(), amplitude);
        ISD_LAUNCH_CHECK();
        ISD_CUDA(cudaStreamSynchronize(g.stream));      // `ratio` is pageable and dies here
    }
    ISD_CUDA(cudaEventRecord(g.ev[12], g.stream));
}

// N2: -grad f/|grad f| at the final vertices (mode 2|4 of k_eval_points = calculate_implicit_gradients(true))
void compute_normals() {
    if (g.normals_valid) return;
    const long long nv = g.nv;
    g.d_normals.reserve((size_t)(nv > 0 ? nv : 1) * 3 * sizeof(float));
    if (nv) {
        upload_program(g.last_program);
        k_eval_points<<<blocks_for(nv, 128), 128, 0, g.stream>>>(g.sdf, g.d_verts.as<float>() + g.local_voff * 3, nv, nullptr,
                                                               g.d_normals.as<float>(), 2 | 4);
        ISD_LAUNCH_CHECK();
    }
    g.normals_valid = true;
    g.normals_host_valid = false;
}

void copy_mesh_to_host();

// N3: send_mesh_back_to_client (polygonizer_algorithm_ob02.hpp:180-220)
void report_progress() {
    if (!g.progress_fn) return;
    copy_mesh_to_host();
    g.progress_fn(g.h_verts.as<float>(), (int)g.nv, g.h_faces.as<int>(), (int)g.nf, g.call_specs[0], g.call_specs[1], g.call_specs[2], g.progress_user);
    g.host_valid = false;       // the next step changes the mesh
}

// Halo mode, end of the step: the collective enqueued after the scan has long finished; the slab's own vertices are a
// contiguous range of the local arrays and its own faces only need the constant shift to global indices.
void export_own_slab() {
    if (!g.halo_mode) return;
    resolve_halo_offsets();
    const long long voff = g.slab[2], foff = g.slab[3], gnv = g.slab[6], gnf = g.slab[7];
    const long long own_nv = g.own_v1 - g.own_v0, own_nf = g.own_f1 - g.own_f0;
    if (own_nf > 0) {
        k_rebase_faces<<<blocks_for(own_nf * 3, 256), 256, 0, g.stream>>>(g.d_faces.as<int>() + g.own_f0 * 3, own_nf * 3, (int)(voff - g.own_v0));
        ISD_LAUNCH_CHECK();
    }
    g.local_voff = g.own_v0; g.local_foff = g.own_f0;
    g.nv = own_nv; g.nf = own_nf;
    g.adjacency_valid = false; g.fof_valid_ = false; g.perm_valid = false; g.topo_valid = false;      // they describe the local mesh, which is gone
    g.host_valid = false;
    g.slab[2] = voff; g.slab[3] = foff; g.slab[4] = own_nv; g.slab[5] = own_nf; g.slab[6] = gnv; g.slab[7] = gnf;
}

// faces do not change after marching cubes: start their device->host copy now, on the side stream
void start_faces_copy() {
    if (g.nf == 0) return;
    g.h_faces.reserve((size_t)g.nf * 3 * sizeof(int));
    ISD_CUDA(cudaEventRecord(g.ev_faces, g.stream));
    ISD_CUDA(cudaStreamWaitEvent(g.copy_stream, g.ev_faces, 0));
    ISD_CUDA(cudaMemcpyAsync(g.h_faces.ptr, g.d_faces.as<int>() + g.local_foff * 3, (size_t)g.nf * 3 * sizeof(int), cudaMemcpyDeviceToHost, g.copy_stream));
    g.faces_in_flight = true;
}

void copy_mesh_to_host() {
    if (g.host_valid) return;
    g.h_verts.reserve((size_t)(g.nv > 0 ? g.nv : 1) * 3 * sizeof(float));
    g.h_faces.reserve((size_t)(g.nf > 0 ? g.nf : 1) * 3 * sizeof(int));
    ISD_CUDA(cudaEventRecord(g.ev[9], g.stream));
    if (g.nv) ISD_CUDA(cudaMemcpyAsync(g.h_verts.ptr, g.d_verts.as<float>() + g.local_voff * 3, (size_t)g.nv * 3 * sizeof(float), cudaMemcpyDeviceToHost, g.stream));
    if (g.nf && !g.faces_in_flight)
        ISD_CUDA(cudaMemcpyAsync(g.h_faces.ptr, g.d_faces.as<int>() + g.local_foff * 3, (size_t)g.nf * 3 * sizeof(int), cudaMemcpyDeviceToHost, g.stream));
    ISD_CUDA(cudaEventRecord(g.ev[10], g.stream));
    ISD_CUDA(cudaStreamSynchronize(g.stream));
    if (g.faces_in_flight) ISD_CUDA(cudaStreamSynchronize(g.copy_stream));
    g.faces_in_flight = false;
    float ms = 0;
    ISD_CUDA(cudaEventElapsedTime(&ms, g.ev[9], g.ev[10]));
    g.timings[T_COPY] = ms;
    g.host_valid = true;
}

// grand_algorithm (mcc2.cpp:309-444)
bool build_once(const char* shape_json, const McSettings& s, bool copy_out) {
    if (g.active) {
        fprintf(stderr, "Error: There are non-deallocated geometry resources. Call finit_geometry() first. Not doing anything.");
        fprintf(stderr, "build_geometry() called in a bad state.");
        return true;
    }
    const auto t0 = std::chrono::steady_clock::now();
    for (float& t : g.timings) t = 0;
    for (int64_t& c : g.counters) c = 0;
    g.launches = 0;
    g.pair_kernels = s.pair != 0;
    g.search_split = s.search_split != 0;
    g.refill_range = s.refill_range;
    g.want_topo = s.mc_adjacency != 0;
    g.want_perm = s.perm > 0;      // -1 (auto) is decided once the program is known, below
    g.finish_split = s.finish_split;
    g_kt.reset(s.kernel_times != 0);
    g_kt.start_counting(s.count_flops != 0);
    g.host_valid = false;
    g.faces_in_flight = false;
    g.adjacency_valid = false;
    g.fof_valid_ = false;
    g.perm_valid = false;
    g.proj_n = 0;
    g.normals_valid = false;
    g.normals_host_valid = false;
    g.last_edge_sum = 0;
    const ShapeCompiler::Program& prog = compile_shape(shape_json, s.ignore_root_matrix, s.cube_mode);
    upload_program(prog.words);
    g.last_program = prog.words;
    g.prog_stats[0] = prog.flops_per_point; g.prog_stats[1] = prog.num_nodes; g.prog_stats[2] = (double)prog.words.size();
    g.prog_stats[3] = prog.value_depth;
    if (s.perm < 0) g.want_perm = prog.num_nodes > 0 && prog.flops_per_point >= 40.0 * prog.num_nodes;
    g.keep_field = true;
    if (!g_ev_build_start) { ISD_CUDA(cudaEventCreate(&g_ev_build_start)); ISD_CUDA(cudaEventCreate(&g_ev_build_end)); }
    g_gather_spans.clear();
    ISD_CUDA(cudaEventRecord(g_ev_build_start, g.stream));
    run_marching_cubes(s, prog);                                   // polygonize_step_0
    bool any_ob02 = s.overall_repeats > 0 && (s.vresampl_iters > 0 || s.projection_enabled || s.subdiv_enabled);
    bool skip_ob02 = false;
    if (any_ob02 && g.world > 1 && g.emulate && !g.halo_mode) {
        // An emulated rank has nobody to gather from: its slab's faces carry global vertex numbers, some of them owned by the slab
        // below, so the replicated OB02 of a real run (gather_mesh() first) cannot be run on it.  The marching-cubes slab is returned.
        fprintf(stderr, "implisolid_b200: emulated rank %d of %d without halo mode: OB02 needs the gathered mesh, returning the marching-cubes slab\n", g.rank, g.world);
        any_ob02 = false;
        skip_ob02 = true;
    }
    if (any_ob02 && g.world > 1 && !g.emulate && !g.halo_mode) gather_mesh();      // OB02 runs on the stitched mesh, replicated
    if (any_ob02 && (g.nv > 2147483647LL / 3 || g.nf > 2147483647LL / 3)) {
        // the OB02, adjacency and subdivision kernels index corners as 3 * i in 32-bit integers
        fprintf(stderr, "implisolid_b200: mesh too large for the OB02 stages on one GPU (%lld vertices, %lld faces; limit %lld): use more z-slabs\n",
                g.nv, g.nf, 2147483647LL / 3);
        abort();
    }
    const bool faces_final = !(s.overall_repeats > 0 && s.subdiv_enabled) && !g.progress_fn;
    if (copy_out && any_ob02 && faces_final && !g.halo_mode) start_faces_copy();
    report_progress();                                             // mcc2.cpp:351
    bool resample_timed = false, subdiv_timed = false;
    for (int overall_iter = 0; !skip_ob02 && overall_iter < s.overall_repeats; ++overall_iter) {
        if (s.vresampl_iters > 0) {
            cudaEvent_t ra = g.ev[4], rb = g.ev[5];
            if (g_spec_used) { ra = g_kt.get(); rb = g_kt.get(); }      // nobody waits in mid-build: private events, read at the end
            ISD_CUDA(cudaEventRecord(ra, g.stream));
            for (int i = 0; i < s.vresampl_iters; ++i) vertex_resampling_step(s.vresampl_c);   // polygonize_step_1
            ISD_CUDA(cudaEventRecord(rb, g.stream));
            if (g_spec_used) g_deferred_spans.push_back(DeferredSpan{T_RESAMPLE, ra, rb});
            else resample_timed = true;
        }
        report_progress();                                                                 // mcc2.cpp:372
        if (s.projection_enabled) {
            projection_step(s.qem_enabled, s.exact_reduce != 0, overall_iter);             // polygonize_step_2
            report_progress();                                                             // mcc2.cpp:390
        }
        if (resample_timed && (s.projection_enabled || s.overall_repeats > 1)) {          // events are reused by the next repeat
            ISD_CUDA(cudaEventSynchronize(g.ev[5]));      // (projection_step has waited for the stream unless the mesh is empty)
            float ms = 0;
            ISD_CUDA(cudaEventElapsedTime(&ms, g.ev[4], g.ev[5]));
            g.timings[T_RESAMPLE] += ms;
            resample_timed = false;
        }
        const bool is_last = overall_iter == s.overall_repeats - 1;
        if (s.subdiv_enabled && (s.overall_repeats <= 1 || is_last)) {                     // polygonize_step_3, mcc2.cpp:396-410
            const float scale_noise_according_to_matrix = 1.0 * 10.0;
            subdivision_step(is_last ? s.post_subdiv_noise * scale_noise_according_to_matrix : 0.0f);
            subdiv_timed = true;
        }
    }
    export_own_slab();
    ISD_CUDA(cudaEventRecord(g_ev_build_end, g.stream));
    ISD_CUDA(cudaStreamSynchronize(g.stream));
    float ms = 0;
    if (g.world > 1 && !g.emulate && g.halo_mode && !g_gather_spans.empty()) {
        // feedback for the slab balance (balanced_layers): this build's busy time, and what the others reported in the gather.
        // Only builds that did not wait for the device in mid-step count (the first build of a slab waits and allocates).
        float total = 0, waited = 0;
        ISD_CUDA(cudaEventElapsedTime(&total, g_ev_build_start, g_ev_build_end));
        for (const auto& sp : g_gather_spans) { ISD_CUDA(cudaEventElapsedTime(&ms, sp.first, sp.second)); waited += ms; }
        if (g_spec_used) {
            g_bal.my_busy_us = (long long)((total - waited) * 1000.0f);
            g_bal.my_busy_gen = g_bal.b.gen;
        }
        const int64_t* hs = g.h_gather.as<int64_t>();
        if ((int)g_bal.table_busy.size() == g.world)
            for (int r = 0; r < g.world; ++r) {
                const long long w = hs[3 + kGatherWords * r + 3];
                g_bal.table_busy[(size_t)r] = w & ((1ll << 40) - 1);
                g_bal.table_gen[(size_t)r] = (int)(w >> 40) - 1;
            }
    }
    if (g_spec_used) {
        // What the device counted against what the build assumed.  A difference means arrays sized and grids launched for another
        // mesh: the result is dropped and the caller (grand_algorithm) builds again, waiting for the device at every step.
        bool held = memcmp(g.h_small.ptr, g_spec.h, sizeof(g_spec.h)) == 0;
        for (int slot : g_deferred_slots) {
            const Ob02Counters* hc = reinterpret_cast<const Ob02Counters*>(g.h_small.as<char>() + 2048) + slot;
            if ((long long)hc->c[10] != g_spec.leftovers[slot]) held = false;
        }
        if (!held) {
            g_deferred_spans.clear(); g_deferred_slots.clear();
            g_kt.collect(); g_kt.finish_counting(); g_kt.on = false;
            g_spec.valid = false; g_spec_used = false;
            return false;
        }
        for (const DeferredSpan& sp : g_deferred_spans) { ISD_CUDA(cudaEventElapsedTime(&ms, sp.a, sp.b)); g.timings[sp.stage] += ms; }
        for (int slot : g_deferred_slots) {
            const Ob02Counters* hc = reinterpret_cast<const Ob02Counters*>(g.h_small.as<char>() + 2048) + slot;
            for (int i = 0; i < 7; ++i) g.counters[3 + i] += (int64_t)hc->c[i];
            g.counters[10] += (int64_t)hc->c[7];
            g.counters[11] += (int64_t)hc->c[8];
            g.counters[12] += (int64_t)hc->c[9];
        }
        if (!g_deferred_slots.empty()) {
            g.average_edge = *reinterpret_cast<const float*>(g.h_small.as<char>() + 1024);
            if (g.world == 1 || g.emulate) memcpy(&g.last_edge_sum, g.h_small.as<char>() + 3200, sizeof(unsigned long long));
        }
        g_deferred_spans.clear(); g_deferred_slots.clear();
    }
    if (resample_timed) { ISD_CUDA(cudaEventElapsedTime(&ms, g.ev[4], g.ev[5])); g.timings[T_RESAMPLE] += ms; }
    ISD_CUDA(cudaEventElapsedTime(&ms, g.ev[0], g.ev[1])); g.timings[T_FIELD] = ms;
    ISD_CUDA(cudaEventElapsedTime(&ms, g.ev[1], g.ev[2])); g.timings[T_CLASSIFY] = ms;
    ISD_CUDA(cudaEventElapsedTime(&ms, g.ev[2], g.ev[3])); g.timings[T_EMIT] = ms;
    if (g.gathered && g.world > 1) { ISD_CUDA(cudaEventElapsedTime(&ms, g.ev[13], g.ev[14])); g.timings[T_GATHER] = ms; }
    if (subdiv_timed) { ISD_CUDA(cudaEventElapsedTime(&ms, g.ev[11], g.ev[12])); g.timings[T_SUBDIV] = ms; }
    g_kt.collect();
    g_kt.finish_counting();
    g_kt.on = false;      // lazily computed normals and the other entry points are not part of a build's table
    if (s.normals) compute_normals();
    if (copy_out) copy_mesh_to_host();
    g.active = true;                                               // polygonize_terminate
    g.counters[13] = g.launches;
    g.counters[16] = g_spec_used ? 1 : 0;
    g_spec_used = false;
    g.timings[T_TOTAL] = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return true;
}

void grand_algorithm(const char* shape_json, const McSettings& s, bool copy_out) {
    const GlibcRand rand_before = g_rand;      // a dropped build must not leave its draws behind
    if (build_once(shape_json, s, copy_out)) return;
    if (g.world > 1 && !g.emulate) {      // the other ranks are not building again: their collectives would not meet this rank's
        fprintf(stderr, "implisolid_b200: rank %d: the device's counts differ from the previous build of this spec\n", g.rank);
        abort();
    }
    fprintf(stderr, "implisolid_b200: the device's counts differ from the previous build of this spec; building again without speculation\n");
    g_rand = rand_before;
    if (!build_once(shape_json, s, copy_out)) { fprintf(stderr, "implisolid_b200: speculation used by a build that has none\n"); abort(); }
    g.counters[17] = 1;
}

}  // namespace

// =====================================================================================================================
extern "C" {

int isd_init(int device) { ensure_init(device); return 0; }
void isd_shutdown(void) {
    if (!g.ready) return;
    cudaStreamSynchronize(g.stream);
    if (g.comm) { g.nccl.CommDestroy(g.comm); g.comm = nullptr; }
}
uint64_t isd_stream(void) { ensure_init(); return (uint64_t)(uintptr_t)g.stream; }

void build_geometry(const char* shape_parameters_json, const char* mc_parameters_json) {
    const McSettings s = parse_mc_settings(mc_parameters_json);
    ensure_init();
    grand_algorithm(shape_parameters_json, s, /*copy_out=*/true);
}
void build_geometry_u(const char* shape_parameters_json, const char* mc_parameters_json, const char* call_specs) {
    // worker_call_sepcs_t (worker_call_specs.hpp:20-34): three optional ints, -1 when absent
    if (call_specs && *call_specs) {
        JsonDoc d(call_specs);
        g.call_specs[0] = d.get_int("progressCallback_id", -1);
        g.call_specs[2] = d.get_int("call_id", -1);
        g.call_specs[1] = d.get_int("shape_id", -1);
    }
    build_geometry(shape_parameters_json, mc_parameters_json);
    g.call_specs[0] = g.call_specs[1] = g.call_specs[2] = -1;
}
void isd_set_progress_callback(isd_progress_fn fn, void* user) { g.progress_fn = fn; g.progress_user = user; }
void isd_srand(unsigned seed) { g_rand.seed(seed); }

static void normals_to_host() {
    compute_normals();
    if (g.normals_host_valid) return;
    g.h_normals.reserve((size_t)(g.nv > 0 ? g.nv : 1) * 3 * sizeof(float));
    if (g.nv) ISD_CUDA(cudaMemcpyAsync(g.h_normals.ptr, g.d_normals.ptr, (size_t)g.nv * 3 * sizeof(float), cudaMemcpyDeviceToHost, g.stream));
    ISD_CUDA(cudaStreamSynchronize(g.stream));
    g.normals_host_valid = true;
}
void* isd_get_n_ptr(void) {
    if (!g.ready) return nullptr;
    normals_to_host();
    return g.h_normals.ptr;
}
void isd_get_n(REAL* n_out, int vcount) {
    if (!g.ready) return;
    normals_to_host();
    const size_t n = (size_t)g.nv * 3;
    if (n) memcpy(n_out, g.h_normals.ptr, n * sizeof(float));
    if ((long long)vcount * 3 != (long long)n) fprintf(stderr, "sizes dont match: %g %d\n", n / 3., vcount);
}
const REAL* isd_device_normals(void) {
    if (!g.ready) return nullptr;
    compute_normals();
    return g.d_normals.as<float>();
}
void isd_build_geometry_device(const char* shape_parameters_json, const char* mc_parameters_json) {
    const McSettings s = parse_mc_settings(mc_parameters_json);
    ensure_init();
    grand_algorithm(shape_parameters_json, s, /*copy_out=*/false);
}
const REAL* isd_device_verts(void) { return g.d_verts.as<float>() + g.local_voff * 3; }
const int* isd_device_faces(void) { return g.d_faces.as<int>() + g.local_foff * 3; }

int get_f_size(void) { return (int)g.nf; }
int get_v_size(void) { return (int)g.nv; }
void get_v(REAL* v_out, int vcount) {
    if (g.ready) copy_mesh_to_host();
    const size_t n = (size_t)g.nv * 3;
    if (n) memcpy(v_out, g.h_verts.ptr, n * sizeof(float));
    if ((long long)vcount * 3 != (long long)n) fprintf(stderr, "sizes dont match: %g %d\n", n / 3., vcount);
}
void get_f(int* f_out, int fcount) {
    if (g.ready) copy_mesh_to_host();
    const size_t n = (size_t)g.nf * 3;
    if (n) memcpy(f_out, g.h_faces.ptr, n * sizeof(int));
    if ((long long)fcount * 3 != (long long)n) fprintf(stderr, "sizes dont match: %g %d\n", n / 3., fcount);
}
void* get_v_ptr(void) {
    if (!g.ready) return nullptr;
    copy_mesh_to_host();
    return g.h_verts.ptr;
}
void* get_f_ptr(void) {
    if (!g.ready) return nullptr;
    copy_mesh_to_host();
    return g.h_faces.ptr;
}
void finish_geometry(void) { g.active = false; }

// ---- direct evaluation ------------------------------------------------------------------------------------------------
int set_object(const char* shape_parameters_json, bool ignore_root_matrix) {
    if (g.svc_object) {
        fprintf(stderr, "Error: You cannot unset() the object before a set_object(json).\n");
        return 0;
    }
    ensure_init();
    ShapeCompiler::Program p = compile_shape(shape_parameters_json, ignore_root_matrix, 0);
    g.svc_program = p.words;
    g.svc_object = true;
    return 1;
}
bool unset_object(int id) {
    if (!g.svc_object) {
        fprintf(stderr, "Error: You cannot unset() the object before a set_object(json).\n");
        return false;
    }
    if (id <= 0) { fprintf(stderr, "Incorrect ID: use the same id returned by set_object(json).\n"); return false; }
    if (id != 1) { fprintf(stderr, "Incorrect ID. For now, The only id is 1\n"); return false; }
    g.svc_program.clear();
    g.svc_object = false;
    return true;
}
bool set_x(void* verts, int n) {
    if (g.svc_x) {
        fprintf(stderr, "Error: You set() before unset()ing the previous set().\n");
        return false;
    }
    if (n < 0 || n >= 10000 * 5) {
        fprintf(stderr, "Error: n is outside [0, 50000].\n");
        return false;
    }
    ensure_init();
    const size_t n3 = (size_t)(n > 0 ? n : 1) * 3 * sizeof(float);
    g.d_svc_x.reserve(n3); g.d_svc_f.reserve(n3 / 3); g.d_svc_g.reserve(n3);
    g.h_svc_f.reserve(n3 / 3); g.h_svc_g.reserve(n3);
    memset(g.h_svc_f.ptr, 0, n3 / 3);
    memset(g.h_svc_g.ptr, 0, n3);
    if (n > 0) {
        ISD_CUDA(cudaMemcpyAsync(g.d_svc_x.ptr, verts, (size_t)n * 3 * sizeof(float), cudaMemcpyHostToDevice, g.stream));
        ISD_CUDA(cudaStreamSynchronize(g.stream));      // the caller's buffer is copied, like the reference does (:781-785)
    }
    g.svc_n = n;
    g.svc_x = true;
    return true;
}
void unset_x(void) {
    if (!g.svc_x) {
        fprintf(stderr, "Error: You cannot unset() before a set().\n");
        return;
    }
    g.svc_x = false;
    g.svc_n = 0;
}
static void svc_eval(int mode) {
    if (g.svc_n == 0) return;
    upload_program(g.svc_program);
    k_eval_points<<<blocks_for(g.svc_n, 128), 128, 0, g.stream>>>(g.sdf, g.d_svc_x.as<float>(), g.svc_n, g.d_svc_f.as<float>(),
                                                                 g.d_svc_g.as<float>(), mode);
    ISD_LAUNCH_CHECK();
    if (mode & 1) ISD_CUDA(cudaMemcpyAsync(g.h_svc_f.ptr, g.d_svc_f.ptr, (size_t)g.svc_n * sizeof(float), cudaMemcpyDeviceToHost, g.stream));
    if (mode & 2) ISD_CUDA(cudaMemcpyAsync(g.h_svc_g.ptr, g.d_svc_g.ptr, (size_t)g.svc_n * 3 * sizeof(float), cudaMemcpyDeviceToHost, g.stream));
    ISD_CUDA(cudaStreamSynchronize(g.stream));
}
void calculate_implicit_values(void) {
    if (!g.svc_x || !g.svc_object) {
        fprintf(stderr, "Error: You need to set_x() and set_object() first.\n");
        return;
    }
    svc_eval(1);
}
void* get_values_ptr(void) { return g.h_svc_f.ptr; }
int get_values_size(void) { return g.svc_n; }
void calculate_implicit_gradients(bool normalize_and_invert) {
    if (!g.svc_x || !g.svc_object) {
        fprintf(stderr, "Error: You need to set_x() and set_object() first.\n");
        return;
    }
    svc_eval(2 | (normalize_and_invert ? 4 : 0));
}
void* get_gradients_ptr(void) {
    if (!g.svc_x || !g.svc_object) {
        fprintf(stderr, "Error: You need to set_x() and set_object() first.\n");
        return nullptr;
    }
    return g.h_svc_g.ptr;
}
int get_gradients_size(void) {
    if (!g.svc_x || !g.svc_object) {
        fprintf(stderr, "Error: You need to set_x() and set_object() first.\n");
        return 0;
    }
    return g.svc_n * 3;
}

void* get_pointset_ptr(char*) { return nullptr; }
int get_pointset_size(char*) { return 0; }
void about(void) {
    fprintf(stderr, "Build Info: \n%s %s\n", __DATE__, __TIME__);
    fprintf(stderr, "implisolid_b200: sm_100a CUDA implementation of the ImpliSolid polygonization path (field eval, marching cubes, OB02)\n");
    fprintf(stderr, "CONFIG: ROOT_TOLERANCE=%g \n", (double)kRootTolerance);
}

// ---- extensions -----------------------------------------------------------------------------------------------------
int isd_last_timings(float* out_ms, int cap) {
    int n = cap < (int)T_COUNT ? cap : (int)T_COUNT;
    for (int i = 0; i < n; ++i) out_ms[i] = g.timings[i];
    return n;
}
int isd_last_counters(int64_t* out, int cap) {
    int n = cap < 18 ? cap : 18;
    for (int i = 0; i < n; ++i) out[i] = g.counters[i];
    return n;
}
int isd_last_kernel_times(float* out_ms, int* out_launches, int cap) {
    int n = cap < (int)KT_N ? cap : (int)KT_N;
    for (int i = 0; i < n; ++i) { out_ms[i] = g_kt.ms[i]; out_launches[i] = g_kt.launches[i]; }
    return n;
}
int isd_last_kernel_gaps(float* out_ms, int cap) {
    int n = cap < (int)KT_N ? cap : (int)KT_N;
    for (int i = 0; i < n; ++i) out_ms[i] = g_kt.gap_ms[i];
    return n;
}
int isd_last_kernel_flops(double* out, int cap) {
    int n = cap < (int)KT_N ? cap : (int)KT_N;
    for (int i = 0; i < n; ++i) out[i] = g_kt.flops[i];
    return n;
}
const char* isd_kernel_name(int id) { return (id >= 0 && id < (int)KT_N) ? g_kernel_names[id] : ""; }
int isd_last_program_stats(double* out4) {
    for (int i = 0; i < 4; ++i) out4[i] = g.prog_stats[i];
    return 4;
}
int isd_grid_size(void) { return g.S; }
int isd_get_field(float* out, int64_t count) {
    if (!g.ready || g.S == 0 || g.world != 1) return 1;
    const int S = g.S;
    if (count != (int64_t)S * S * S) return 2;
    ISD_CUDA(cudaMemcpy2DAsync(out, (size_t)S * sizeof(float), g.d_field.ptr, (size_t)g.grid.pitch * sizeof(float), (size_t)S * sizeof(float),
                               (size_t)S * S, cudaMemcpyDeviceToHost, g.stream));
    ISD_CUDA(cudaStreamSynchronize(g.stream));
    return 0;
}
int isd_get_case_codes(uint8_t* out, int64_t count) {
    if (!g.ready || !g.have_cases) return 1;
    const int64_t n = (int64_t)(g.mc.zc_end - g.mc.scan_z0) * g.mc.NX * g.mc.NX;
    if (count != n) return 2;
    ISD_CUDA(cudaMemcpyAsync(out, g.d_cases.ptr, (size_t)n, cudaMemcpyDeviceToHost, g.stream));
    ISD_CUDA(cudaStreamSynchronize(g.stream));
    return 0;
}
float isd_average_edge(void) { return g.average_edge; }
int isd_get_projected_centroids(float* out, int64_t count) {
    if (!g.ready || g.proj_n == 0 || count != g.proj_n * 3) return 1;
    ISD_CUDA(cudaMemcpyAsync(out, g.d_proj.ptr, (size_t)count * sizeof(float), cudaMemcpyDeviceToHost, g.stream));
    ISD_CUDA(cudaStreamSynchronize(g.stream));
    return 0;
}
int isd_debug_adjacency(int* offsets_nv1, int* items_3nf, int* fof_3nf) {
    if (!g.ready || !g.adjacency_valid || g.nf == 0) return 1;
    ISD_CUDA(cudaMemcpyAsync(offsets_nv1, g.d_uoff.ptr, (size_t)(g.nv + 1) * sizeof(int), cudaMemcpyDeviceToHost, g.stream));
    ISD_CUDA(cudaMemcpyAsync(items_3nf, g.d_uitems.ptr, (size_t)g.nf * 3 * sizeof(int), cudaMemcpyDeviceToHost, g.stream));
    if (fof_3nf) {
        if (!g.fof_valid_) return 2;
        ISD_CUDA(cudaMemcpyAsync(fof_3nf, g.d_fof.ptr, (size_t)g.nf * 3 * sizeof(int), cudaMemcpyDeviceToHost, g.stream));
    }
    ISD_CUDA(cudaStreamSynchronize(g.stream));
    return 0;
}
int isd_get_centroid_normals(float* out, int64_t count) {
    if (!g.ready || g.proj_n == 0 || count != g.proj_n * 3) return 1;
    ISD_CUDA(cudaMemcpyAsync(out, g.d_pnrm.ptr, (size_t)count * sizeof(float), cudaMemcpyDeviceToHost, g.stream));
    ISD_CUDA(cudaStreamSynchronize(g.stream));
    return 0;
}

void isd_set_sdf_grid(int nx, int ny, int nz, float h, float ox, float oy, float oz, const float* values) {
    ensure_init();
    const size_t n = (size_t)nx * ny * nz;
    g.d_sdf.reserve(n * sizeof(float));
    ISD_CUDA(cudaMemcpy(g.d_sdf.ptr, values, n * sizeof(float), cudaMemcpyHostToDevice));
    g.sdf = SdfGridView{g.d_sdf.as<float>(), nx, ny, nz, h, ox, oy, oz};
}

static void load_mesh(const float* verts, int nv, const int* faces, int nf) {
    g.local_voff = 0; g.local_foff = 0;
    g.d_verts.reserve((size_t)(nv > 0 ? nv : 1) * 3 * sizeof(float));
    g.d_faces.reserve((size_t)(nf > 0 ? nf : 1) * 3 * sizeof(int));
    if (verts && nv) ISD_CUDA(cudaMemcpyAsync(g.d_verts.ptr, verts, (size_t)nv * 3 * sizeof(float), cudaMemcpyHostToDevice, g.stream));
    if (nf) ISD_CUDA(cudaMemcpyAsync(g.d_faces.ptr, faces, (size_t)nf * 3 * sizeof(int), cudaMemcpyHostToDevice, g.stream));
    g.nv = nv; g.nf = nf;
    g.adjacency_valid = false; g.fof_valid_ = false; g.perm_valid = false; g.topo_valid = false; g.host_valid = false;
}

void isd_ob02_on_mesh(const char* shape_parameters_json, float* verts, int nv, const int* faces, int nf, int resample_iters, float c,
                      int projection, int qem) {
    ensure_init();
    ShapeCompiler::Program p = compile_shape(shape_parameters_json, false, 0);
    upload_program(p.words);
    for (int64_t& v : g.counters) v = 0;
    load_mesh(verts, nv, faces, nf);
    for (int i = 0; i < resample_iters; ++i) vertex_resampling_step(c);
    g_spec_used = false;
    if (projection) projection_step(qem != 0, true, -1);
    ISD_CUDA(cudaMemcpyAsync(verts, g.d_verts.ptr, (size_t)nv * 3 * sizeof(float), cudaMemcpyDeviceToHost, g.stream));
    ISD_CUDA(cudaStreamSynchronize(g.stream));
}
void isd_faces_of_faces(const int* faces, int nf, int nv, int* out_3nf) {
    ensure_init();
    load_mesh(nullptr, nv, faces, nf);
    build_adjacency(true);
    ISD_CUDA(cudaMemcpyAsync(out_3nf, g.d_fof.ptr, (size_t)nf * 3 * sizeof(int), cudaMemcpyDeviceToHost, g.stream));
    ISD_CUDA(cudaStreamSynchronize(g.stream));
}
void isd_umbrella(const int* faces, int nf, int nv, int* offsets_nv1, int* items_3nf) {
    ensure_init();
    load_mesh(nullptr, nv, faces, nf);
    build_adjacency(false);
    ISD_CUDA(cudaMemcpyAsync(offsets_nv1, g.d_uoff.ptr, (size_t)(nv + 1) * sizeof(int), cudaMemcpyDeviceToHost, g.stream));
    ISD_CUDA(cudaMemcpyAsync(items_3nf, g.d_uitems.ptr, (size_t)nf * 3 * sizeof(int), cudaMemcpyDeviceToHost, g.stream));
    ISD_CUDA(cudaStreamSynchronize(g.stream));
}

// ---- isd_umbrella without any allocation (no counterpart in the reference) -------
// For callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it): the umbrella CSR of
// a mesh whose faces are already in device memory, written to device arrays the caller owns.  The scratch the kernels need —
// per-vertex degree and fill cursor, the single-pass scan's block state — lives in ONE file-scope device array of fixed size
// (4.2 MB, held by every GPU that loads the library until the process exits) instead of the grow-only buffers of build_adjacency().
// In host code `s_noalloc_scratch` is only a stand-in for the symbol: its device address comes from cudaGetSymbolAddress, once per
// device, with that device current.  Every call shares the array, so calls that overlap on the GPU (different streams, or graph
// launches in flight together) race; what must start at zero is cleared by every call.
constexpr int kNoAllocMaxVertices = 524288;      // stated maximum: meshes of up to 2^19 vertices
constexpr size_t kNoAllocStateWords = 4 + 2 * (size_t)kUScanMaxBlocks;
constexpr size_t kNoAllocScratchBytes = (2 * ((size_t)kNoAllocMaxVertices + 1) + kNoAllocStateWords) * sizeof(unsigned);
}      // extern "C" (reopened below: the array has file scope and internal linkage)
static __device__ __align__(256) unsigned char s_noalloc_scratch[kNoAllocScratchBytes];
extern "C" {
int isd_umbrella_noalloc(const int* d_faces, int nf, int nv, int* d_offsets_nv1, int* d_items_3nf, void* cuda_stream) {
    if (!g.ready) {      // ensure_init() creates streams and pinned buffers: not something this entry point may do
        fprintf(stderr, "implisolid_b200: isd_umbrella_noalloc: call isd_init() first (this entry point allocates nothing, not even the library's streams)\n");
        return 2;
    }
    if (nv <= 0 || nf < 0 || d_offsets_nv1 == nullptr || (nf > 0 && (d_faces == nullptr || d_items_3nf == nullptr))) {
        fprintf(stderr, "implisolid_b200: isd_umbrella_noalloc: bad arguments (nv %d, nf %d)\n", nv, nf);
        return 1;
    }
    if (nv > kNoAllocMaxVertices || nf > 2147483647 / 3) {
        fprintf(stderr, "implisolid_b200: isd_umbrella_noalloc: mesh too large for the fixed scratch array (%d vertices, limit %d; %d faces, limit %d): use "
                        "isd_umbrella, which allocates\n", nv, kNoAllocMaxVertices, nf, 2147483647 / 3);
        return 1;
    }
#define ISD_NOALLOC_CUDA(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { \
        fprintf(stderr, "implisolid_b200: isd_umbrella_noalloc: %s at line %d: %s\n", cudaGetErrorName(e__), __LINE__, cudaGetErrorString(e__)); return 3; } } while (0)
    static unsigned char* scratch_of_device[64] = {nullptr};      // device address of s_noalloc_scratch, per device
    int dev = 0;
    ISD_NOALLOC_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) { fprintf(stderr, "implisolid_b200: isd_umbrella_noalloc: device %d out of range\n", dev); return 1; }
    if (scratch_of_device[dev] == nullptr) {
        unsigned char* ptr = nullptr;
        ISD_NOALLOC_CUDA(cudaGetSymbolAddress((void**)&ptr, s_noalloc_scratch));
        scratch_of_device[dev] = ptr;
    }
    cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : g.stream;
    unsigned* deg = reinterpret_cast<unsigned*>(scratch_of_device[dev]);
    unsigned* cursor = deg + (kNoAllocMaxVertices + 1);
    unsigned* state = cursor + (kNoAllocMaxVertices + 1);
    unsigned* offsets = reinterpret_cast<unsigned*>(d_offsets_nv1);
    if (nf == 0) {
        ISD_NOALLOC_CUDA(cudaMemsetAsync(offsets, 0, ((size_t)nv + 1) * sizeof(unsigned), st));
        return 0;
    }
    // the array keeps the previous call's contents: degree, cursor and scan state start at zero
    ISD_NOALLOC_CUDA(cudaMemsetAsync(deg, 0, ((size_t)nv + 1) * sizeof(unsigned), st));
    ISD_NOALLOC_CUDA(cudaMemsetAsync(cursor, 0, ((size_t)nv + 1) * sizeof(unsigned), st));
    ISD_NOALLOC_CUDA(cudaMemsetAsync(state, 0, kNoAllocStateWords * sizeof(unsigned), st));
    const unsigned corner_blocks = (unsigned)(((long long)nf * 3 + 255) / 256);
    k_vertex_degree<<<corner_blocks, 256, 0, st>>>(d_faces, nf, deg);
    ISD_NOALLOC_CUDA(cudaGetLastError());
    {   // exclusive_scan_u32's launch geometry
        int blocks = (int)(((long long)nv + kUScanThreads * 4 - 1) / (kUScanThreads * 4));
        if (blocks > kUScanMaxBlocks) blocks = kUScanMaxBlocks;
        const long long per_block = (((long long)nv + blocks - 1) / blocks + kUScanThreads - 1) / kUScanThreads * kUScanThreads;
        k_uscan<<<blocks, kUScanThreads, 0, st>>>(deg, (long long)nv, per_block, state, offsets);
        ISD_NOALLOC_CUDA(cudaGetLastError());
    }
    k_umbrella_fill<<<corner_blocks, 256, 0, st>>>(d_faces, nf, offsets, cursor, d_items_3nf);
    ISD_NOALLOC_CUDA(cudaGetLastError());
    k_umbrella_sort<<<(unsigned)((nv + 127) / 128), 128, 0, st>>>(nv, offsets, d_items_3nf);
    ISD_NOALLOC_CUDA(cudaGetLastError());
#undef ISD_NOALLOC_CUDA
    return 0;
}
int isd_umbrella_noalloc_max_vertices(void) { return kNoAllocMaxVertices; }

void isd_eval_points(const char* shape_parameters_json, int ignore_root_matrix, const float* xyz, int64_t n, float* f_out, float* grad_out) {
    ensure_init();
    if (n <= 0) return;
    ShapeCompiler::Program p = compile_shape(shape_parameters_json, ignore_root_matrix != 0, 0);
    upload_program(p.words);
    DeviceBuffer dx, df, dg;
    dx.reserve((size_t)n * 3 * sizeof(float)); df.reserve((size_t)n * sizeof(float)); dg.reserve((size_t)n * 3 * sizeof(float));
    ISD_CUDA(cudaMemcpyAsync(dx.ptr, xyz, (size_t)n * 3 * sizeof(float), cudaMemcpyHostToDevice, g.stream));
    k_eval_points<<<blocks_for(n, 128), 128, 0, g.stream>>>(g.sdf, dx.as<float>(), n, df.as<float>(), dg.as<float>(), grad_out ? 3 : 1);
    ISD_LAUNCH_CHECK();
    ISD_CUDA(cudaMemcpyAsync(f_out, df.ptr, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, g.stream));
    if (grad_out) ISD_CUDA(cudaMemcpyAsync(grad_out, dg.ptr, (size_t)n * 3 * sizeof(float), cudaMemcpyDeviceToHost, g.stream));
    ISD_CUDA(cudaStreamSynchronize(g.stream));
    dx.release(); df.release(); dg.release();
}

// ---- z-slab sharding ----------------------------------------------------------------------------------------------------
int isd_dist_unique_id(uint8_t* id128) {
    if (!g.nccl.load()) return 1;
    NcclApi::unique_id id;
    g.nccl.check(g.nccl.GetUniqueId(&id), "ncclGetUniqueId");
    memcpy(id128, &id, 128);
    return 0;
}
int isd_dist_init(int rank, int world, const uint8_t* id128) {
    ensure_init();
    if (world <= 1) { g.rank = 0; g.world = 1; return 0; }
    if (rank < 0 || rank >= world || world > 4096) {
        fprintf(stderr, "implisolid_b200: isd_dist_init: rank %d / world %d out of range\n", rank, world);
        return 1;
    }
    g.h_gather.reserve((size_t)(3 + kGatherWords * world) * sizeof(int64_t));
    if (!g.nccl.load()) abort();
    NcclApi::unique_id id;
    memcpy(&id, id128, 128);
    g.nccl.check(g.nccl.CommInitRank(&g.comm, world, id, rank), "ncclCommInitRank");
    g.rank = rank; g.world = world; g.emulate = false;
    g.comm_rank = rank; g.comm_world = world;
    return 0;
}
void isd_debug_spoil_speculation(int what) {
    // tests only: make the next speculative build of the cached spec assume a wrong left-over count for its first projection step
    // (the one count whose misprediction cannot overrun an array), so that the check-and-build-again path runs
    if (what == 0) { g_spec.valid = false; return; }
    if (g_spec.valid && g_spec.leftovers[0] >= 0) g_spec.leftovers[0] = g_spec.leftovers[0] > 0 ? 0 : 1;
}
void isd_dist_suspend(void) { g.rank = 0; g.world = 1; g.emulate = false; }
void isd_dist_resume(void) {
    if (g.comm) { g.rank = g.comm_rank; g.world = g.comm_world; g.emulate = false; }
}
void isd_dist_finalize(void) {
    if (g.comm) { g.nccl.CommDestroy(g.comm); g.comm = nullptr; }
    g.rank = 0; g.world = 1; g.emulate = false; g.comm_rank = 0; g.comm_world = 1;
}
int isd_dist_emulate(int rank, int world) {
    if (world > 1 && (rank < 0 || rank >= world)) {
        fprintf(stderr, "implisolid_b200: isd_dist_emulate: rank %d / world %d out of range\n", rank, world);
        return 1;
    }
    g.rank = rank; g.world = world; g.emulate = world > 1;
    g.emu_voff = 0; g.emu_foff = 0;
    return 0;
}
void isd_dist_set_offsets(int64_t voff, int64_t foff) { g.emu_voff = voff; g.emu_foff = foff; }
int isd_dist_last_slab(int64_t* out8) {
    for (int i = 0; i < 8; ++i) out8[i] = g.slab[i];
    return 0;
}
void isd_dist_gather_mesh(void) { gather_mesh(); }
void isd_dist_set_edge_total(uint64_t edge_total_fixed_point, int64_t global_faces) { g.emu_edge_total = edge_total_fixed_point; g.emu_gnf = global_faces; }
uint64_t isd_dist_last_edge_sum(void) { return g.last_edge_sum; }
int isd_dist_last_halo(void) { return g.halo_mode ? g.halo : 0; }

}  // extern "C"
