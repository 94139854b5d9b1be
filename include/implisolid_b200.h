/* implisolid_b200 — C ABI of libimplisolid_b200.so
 *
 * Group 1-3 are, symbol for symbol, the extern "C" block the reference exports from mcc2.cpp and
 * that its Emscripten/JS layer binds with Module.cwrap (implisolid_main.js:38-59,
 * js/worker_api.js:121-127).  A maintainer swaps the library, not the callers.  Group 4 ("isd_")
 * is ours: device-resident results, stage timings, z-slab sharding and parity introspection.
 *
 * All entry points are plain C: pointers and sizes, no C++/torch types.  Paths below are relative
 * to /root/reference/js_iteration_2/.
 *
 * Protocol (mcc2.cpp:165-193, :309-319, :540-552): singleton, not re-entrant, one caller thread.
 *   build_geometry -> get_*_size / get_* / get_*_ptr -> finish_geometry -> next build_geometry.
 * A second build_geometry without finish_geometry logs and returns without building.  Malformed
 * settings or an unknown shape type abort() the process after a message on stderr, as the
 * reference does; a CUDA failure does the same.  There is no CPU fallback: without a usable
 * sm_100 device every compute entry point aborts with a message.
 */
#ifndef IMPLISOLID_B200_H
#define IMPLISOLID_B200_H

#include <stdbool.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef float REAL; /* basic_data_structures.hpp:60 */

/* ---- group 1: polygonization ------------------------------------------------------------ */
/* mcc2.cpp:446-464 (decl :89).  shape_parameters_json: one node of an mp5 document
 * (object_factory.hpp:56-758); mc_parameters_json: mc_settings (polygoniser_settings.hpp:147-305). */
void build_geometry(const char* shape_parameters_json, const char* mc_parameters_json);
/* mcc2.cpp:90 (declared, body commented out at :296-302; JS cwraps it, js/worker_api.js:126).
 * call_specs = {"progressCallback_id","call_id","shape_id"} (worker_call_specs.hpp:20-34): the ids handed back with every
 * progress report (isd_set_progress_callback below), the native stand-in for send_mesh_back_to_client
 * (polygonizer_algorithm_ob02.hpp:180-220). */
void build_geometry_u(const char* shape_parameters_json, const char* mc_parameters_json, const char* call_specs);
int get_v_size(void);              /* mcc2.cpp:470-473  number of vertices  */
int get_f_size(void);              /* mcc2.cpp:466-469  number of triangles */
void get_v(REAL* v_out, int vcount); /* mcc2.cpp:474-492  copies 3*Nv floats; logs on size mismatch */
void get_f(int* f_out, int fcount);  /* mcc2.cpp:496-511  copies 3*Nf ints;   logs on size mismatch */
void* get_v_ptr(void);             /* mcc2.cpp:515-519  borrowed, valid until the next build_geometry */
void* get_f_ptr(void);             /* mcc2.cpp:521-525 */
void finish_geometry(void);        /* mcc2.cpp:540-552 */

/* ---- group 2: direct evaluation of the implicit function (normals) ------------------------ */
int set_object(const char* shape_parameters_json, bool ignore_root_matrix); /* mcc2.cpp:726-745, returns id 1 (0 on error) */
bool unset_object(int id);                                                  /* mcc2.cpp:747-763 */
bool set_x(void* verts_xyz_f32, int n);                                     /* mcc2.cpp:765-798, rejects n<0 or n>=50000 */
void unset_x(void);                                                         /* mcc2.cpp:800-812 */
void calculate_implicit_values(void);                                       /* mcc2.cpp:815-825 */
void* get_values_ptr(void);                                                 /* mcc2.cpp:827-829, n floats */
int get_values_size(void);                                                  /* mcc2.cpp:830-832 */
void calculate_implicit_gradients(bool normalize_and_invert);               /* mcc2.cpp:834-889 */
void* get_gradients_ptr(void);                                              /* mcc2.cpp:891-900, 3n floats */
int get_gradients_size(void);                                               /* mcc2.cpp:902-911 */

/* ---- group 3: misc ------------------------------------------------------------------------ */
void* get_pointset_ptr(char* id);  /* mcc2.cpp:203-208  debug point sets: always NULL here */
int get_pointset_size(char* id);   /* mcc2.cpp:210-215  always 0 */
void about(void);                  /* mcc2.cpp:574-617 */

/* ---- group 4: extensions (ours) ------------------------------------------------------------ */
/* Progressive results (SURVEY §8 f-N3).  The reference reports the mesh back to its JS client after marching cubes, after
 * the resampling iterations and after the projection of every repeat (mcc2.cpp:351,372,390 -> send_mesh_back_to_client,
 * polygonizer_algorithm_ob02.hpp:180-220, an EM_ASM block with no native meaning).  Natively the same three points call
 * `fn` with host pointers to the current mesh (what get_v_ptr/get_f_ptr return at that moment) and the ids of
 * build_geometry_u's call_specs (-1 for build_geometry).  NULL (the default) disables reporting and its copies. */
typedef void (*isd_progress_fn)(const REAL* verts, int nverts, const int* faces, int nfaces, int progress_callback_id,
                                int shape_id, int call_id, void* user);
void isd_set_progress_callback(isd_progress_fn fn, void* user);

/* Per-vertex normals of the last build (SURVEY §8 f-N2): -grad f/|grad f| at the final vertices with the -42 rule of
 * calculate_implicit_gradients(true) (mcc2.cpp:852-870) — what the JS flow asks for after every build through
 * set_object/set_x/calculate_implicit_gradients (implisolid_main.js:551-600), without the second tree upload, the
 * 50 000-point cap of set_x (mcc2.cpp:770-773) or the round trip.  Computed on the device right after the last step when
 * mc_parameters_json holds "gpu":{"normals":1}, otherwise on first use.  3*Nv floats; valid until the next build. */
void* isd_get_n_ptr(void);
void isd_get_n(REAL* n_out, int vcount);
const REAL* isd_device_normals(void);

/* polygonize_step_3's noise uses libc rand(), never seeded (basic_functions.hpp:551-557): the library keeps a private
 * glibc-compatible generator that starts like srand(1) and runs on across builds as the reference's does; this resets it. */
void isd_srand(unsigned seed);
/* Device selection / lifetime.  isd_init is implicit on first use (device = $ISD_DEVICE or
 * $LOCAL_RANK or 0).  Returns 0 on success; any failure aborts. */
int isd_init(int device);
void isd_shutdown(void);
/* cudaStream_t all kernels are launched on, as an integer handle (for CUDA-event timing). */
uint64_t isd_stream(void);

/* build_geometry without the device->host copy of the mesh: results stay in HBM
 * (isd_device_verts / isd_device_faces).  get_v/get_f/get_*_ptr after this call copy lazily. */
void isd_build_geometry_device(const char* shape_parameters_json, const char* mc_parameters_json);
const REAL* isd_device_verts(void); /* device pointer, 3*Nv floats */
const int* isd_device_faces(void);  /* device pointer, 3*Nf ints   */

/* Per-stage device times of the last build, milliseconds, CUDA events on isd_stream().
 * out[0]=field eval, [1]=mc classify+scan, [2]=mc emit, [3]=vertex resampling, [4]=projection,
 * [5]=qem, [6]=device->host mesh copy, [7]=whole build (host wall clock).  Returns entries written. */
int isd_last_timings(float* out_ms, int cap);
/* Counters of the last build: [0]=nodes evaluated, [1]=cubes visited, [2]=cut cubes,
 * [3..9]=centroids resolved by projection direction type 0..6, [10]=projection failed,
 * [11]=bisection capped, [12]=bisected, [13]=kernel launches, [14]=field bytes, [15]=nodes
 * evaluated with the full program (others were proven off-surface and skipped), [16]=1 when the build
 * ran without waiting for the device in mid-step (sizes taken from the previous build of the same spec
 * and checked at the end, "gpu":{"speculate":1}), [17]=1 when such a build was dropped and done again. */
int isd_last_counters(int64_t* out, int cap);
/* Tests: what=0 forgets the previous build's counts (the next build waits for the device); what=1 makes the next
 * speculative build of the same spec assume a wrong left-over count, so that its end-of-build check fails and
 * the build is done again. */
void isd_debug_spoil_speculation(int what);
/* Shape program of the last build: out = {flops per point (mul and add counted separately), tree nodes,
 * program words, value-stack depth}. */
int isd_last_program_stats(double* out4);
/* With "gpu":{"kernel_times":1} in mc_parameters_json: device time (ms, summed over its launches) and launch count of every
 * kernel of the last build, each launch between its own pair of events on isd_stream().  isd_kernel_name(i) names entry i
 * ("" past the end).  Returns entries written.  Without the setting every entry is zero. */
int isd_last_kernel_times(float* out_ms, int* out_launches, int cap);
/* With "gpu":{"count_flops":1}: FP32 operations executed by each kernel's CSG evaluations in the last build (mul and add counted
 * separately; the instruction lists each evaluation really ran, not the whole tree).  Same indices as isd_last_kernel_times. */
int isd_last_kernel_flops(double* out, int cap);
/* Stream time between the end of the previous timed kernel and the start of kernel i, summed over its launches (memsets,
 * copies, host waits and collectives that precede it).  Same indices and setting as isd_last_kernel_times. */
int isd_last_kernel_gaps(float* out_ms, int cap);
const char* isd_kernel_name(int id);

/* Parity introspection (needs "gpu":{"keep_field":1} in mc_parameters_json for the field). */
int isd_grid_size(void);                       /* S = resolution + 5 nodes per axis */
int isd_get_field(float* out, int64_t count);  /* S^3 floats, x fastest (marching_cubes.hpp:264-268); 0 on success */
int isd_get_case_codes(uint8_t* out, int64_t count); /* (S-3)^3 cube codes in scan order z,y,x (marching_cubes.hpp:1041-1057) */
float isd_average_edge(void);                         /* mean edge length used by the last projection */
int isd_get_projected_centroids(float* out, int64_t count); /* 3*Nf floats of the last projection */
/* Debug (parity tests): the adjacency tables the last build's OB02 stages used, for the mesh they were built on (before any
 * subdivision): umbrella CSR (nv+1 offsets, 3*nf items) and, if fof_3nf != NULL and it was built, faces-of-faces.  0 on success. */
int isd_debug_adjacency(int* offsets_nv1, int* items_3nf, int* fof_3nf);
int isd_get_centroid_normals(float* out, int64_t count);

/* HEAD-literal `cube` (cube.hpp:65-72,215-270): the SDF table it interpolates, supplied by the
 * caller ("gpu":{"cube_mode":1} selects it). */
void isd_set_sdf_grid(int nx, int ny, int nz, float h, float ox, float oy, float oz, const float* values);

/* OB02 stages on a caller-supplied mesh (fixtures: mcc2.cpp:254-290 make_a_cube / make_a_square).
 * verts is updated in place. */
void isd_ob02_on_mesh(const char* shape_parameters_json, float* verts, int nv, const int* faces, int nf,
                      int resample_iters, float c, int projection, int qem);
/* mesh adjacency as the OB02 stages see it (mesh_algorithms.hpp:50-209) */
void isd_faces_of_faces(const int* faces, int nf, int nv, int* out_3nf);
void isd_umbrella(const int* faces, int nf, int nv, int* offsets_nv1, int* items_3nf);
/* isd_umbrella for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  No
 * counterpart in the reference.  All pointers are DEVICE pointers the caller owns
 * (faces: 3*nf ints in; offsets: nv+1 ints out; items: 3*nf ints out); the work is enqueued on `cuda_stream` (a cudaStream_t, NULL =
 * the library's stream) and NOT waited for.  The call allocates nothing: its scratch is one fixed 4.2 MB device array inside the
 * library, so nv <= isd_umbrella_noalloc_max_vertices() (524288), every call shares that array and calls that overlap on the GPU
 * race.  isd_init() must have been called.  Returns 0, or 1 bad arguments / mesh beyond the limit, 2 library not initialised,
 * 3 CUDA error (each with a message on stderr). */
int isd_umbrella_noalloc(const int* d_faces, int nf, int nv, int* d_offsets_nv1, int* d_items_3nf, void* cuda_stream);
int isd_umbrella_noalloc_max_vertices(void);

/* Points without the 50 000 cap of set_x: f and grad f at n host points (device kernel). */
void isd_eval_points(const char* shape_parameters_json, int ignore_root_matrix, const float* xyz, int64_t n,
                     float* f_out, float* grad_out_or_null);

/* ---- z-slab sharding across the GPUs of one box (one process per GPU) ---------------------- */
/* The library calls NCCL itself (dlopen of libnccl.so.2): one ncclAllGather of {nv,nf} per build.
 * id128 is an ncclUniqueId made by isd_dist_unique_id on rank 0 and handed to every rank by the
 * launcher (bench.py broadcasts it with torch.distributed).  After isd_dist_init every
 * build_geometry evaluates only this rank's z-slab; vertex and face indices are global. */
int isd_dist_unique_id(uint8_t* id128);
int isd_dist_init(int rank, int world, const uint8_t* id128);
void isd_dist_finalize(void);
/* Build the next geometries whole on this GPU (world 1) while keeping the communicator, and go back to this rank's slab:
 * lets one rank produce the single-GPU mesh a sharded run is checked against (bench.py's parity_ok). */
void isd_dist_suspend(void);
void isd_dist_resume(void);
/* Test hook: emulate rank `rank` of `world` without NCCL; the caller supplies the gathered counts
 * global vertex/face offsets of this slab through isd_dist_set_offsets (see tests/test_gpu_slabs.py). */
int isd_dist_emulate(int rank, int world);
void isd_dist_set_offsets(int64_t global_vertex_offset, int64_t global_face_offset);
/* Slab owned by this rank in the last build: out = {first cube layer, end cube layer, global vertex
 * offset, global face offset, local Nv, local Nf, global Nv, global Nf}. */
int isd_dist_last_slab(int64_t* out8);
/* Gather the stitched mesh on every rank (one group of in-place ncclBroadcasts, rank r's slab from rank r, at
 * the offsets the {nv,nf} all-gather gave); afterwards get_v/get_f return the global mesh on every rank. */
void isd_dist_gather_mesh(void);
/* z-slab runs with OB02 ("halo mode"): every GPU optimises its own slab plus a halo of recomputed cube layers,
 * the only exchanges being the {nv,nf} all-gather and an 8-byte all-reduce of the fixed-point edge-length sum.  When ranks
 * are emulated on one GPU (isd_dist_emulate) the test driver plays the all-reduce: it hands every rank the whole mesh's
 * sum and face count, and reads a rank's share back. */
void isd_dist_set_edge_total(uint64_t edge_total_fixed_point, int64_t global_faces);
uint64_t isd_dist_last_edge_sum(void);
int isd_dist_last_halo(void); /* halo layers used by the last build (0: none / replicated OB02) */

#ifdef __cplusplus
}
#endif
#endif /* IMPLISOLID_B200_H */
