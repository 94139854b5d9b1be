/* Smallest native caller of the drop-in library: the calls a g++ build of the reference's mcc2.cpp main() makes
 * (mcc2.cpp:916 ff.: build_geometry, get_v_size/get_f_size, get_v/get_f, finish_geometry), here in plain C99, writing the
 * mesh as a Wavefront OBJ.  Needs a CUDA device at run time (the library has no CPU fallback).
 *
 *   gcc -std=c99 -O2 -Iinclude examples/polygonize_to_obj.c -Limplisolid_b200 -limplisolid_b200 \
 *       -Wl,-rpath,$PWD/implisolid_b200 -o polygonize_to_obj
 *   ./polygonize_to_obj 128 > cube_minus_cylinder.obj
 */
#include <stdio.h>
#include <stdlib.h>
#include "implisolid_b200.h"

int main(int argc, char** argv) {
    const int resolution = argc > 1 ? atoi(argv[1]) : 64;
    /* the JSON the reference's JS layer sends (implisolid_main.js:152-199, js/worker_api.js:121-127) */
    const char* shape =
        "{\"type\":\"Difference\",\"matrix\":[1,0,0,0, 0,1,0,0, 0,0,1,0, 0,0,0,1],\"children\":["
        "{\"type\":\"cube\",\"matrix\":[1,0,0,0, 0,1,0,0, 0,0,1,0, 0,0,0,1]},"
        "{\"type\":\"cylinder\",\"matrix\":[0.6,0,0,0, 0,0.6,0,0, 0,0,1.4,0, 0,0,0,1]}]}";
    char mc[512];
    snprintf(mc, sizeof mc,
             "{\"resolution\":%d,\"box\":{\"xmin\":-0.8,\"xmax\":0.8,\"ymin\":-0.8,\"ymax\":0.8,\"zmin\":-0.8,\"zmax\":0.8},"
             "\"ignore_root_matrix\":0,\"overall_repeats\":1,\"vresampl\":{\"iters\":1,\"c\":1.0},"
             "\"projection\":{\"enabled\":1},\"qem\":{\"enabled\":1},\"subdiv\":{\"enabled\":0}}",
             resolution);
    build_geometry(shape, mc);
    const int nv = get_v_size(), nf = get_f_size();
    float* v = (float*)malloc(sizeof(float) * 3 * (size_t)(nv > 0 ? nv : 1));
    int* f = (int*)malloc(sizeof(int) * 3 * (size_t)(nf > 0 ? nf : 1));
    if (!v || !f) return 1;
    get_v(v, nv);
    get_f(f, nf);
    finish_geometry();
    for (int i = 0; i < nv; ++i) printf("v %.9g %.9g %.9g\n", v[3 * i], v[3 * i + 1], v[3 * i + 2]);
    for (int i = 0; i < nf; ++i) printf("f %d %d %d\n", f[3 * i] + 1, f[3 * i + 1] + 1, f[3 * i + 2] + 1);
    fprintf(stderr, "%d vertices, %d triangles\n", nv, nf);
    free(v);
    free(f);
    return 0;
}
