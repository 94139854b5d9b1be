// polygonize_step_3: 1-to-4 subdivision of every face (paths relative to /root/reference/js_iteration_2/):
//   polygonizer_algorithm_ob02.hpp:119-156  ->  my_subdiv_ (centroids_projection.cpp:1314-1367)
//   -> subdivide_given_faces (subdivision/do_subdivision.hpp:26-51, returns after the 1-to-4 pass)
//   -> subdivide_multiple_facets_1to4 (subdivision/subdiv_1to4.hpp:33-330), then randomize_verts (basic_functions.hpp:551-557).
//
// The reference walks the faces in ascending order with a std::map from the sorted vertex pair to the midpoint's index: an
// edge's midpoint is created by the FIRST face that holds the edge, in slot order e01, e12, e20, and numbered in that order of
// creation after the old vertices.  Without a map: the umbrella of one end vertex (ascending face indices, ob02_kernels.cuh)
// gives the lowest face holding the edge, a flag marks the (face, slot) pairs that create their midpoint, and an exclusive
// scan of the flags is the creation order.  The midpoint is a row of the reference's 3x3 product with {H,O,H; H,H,O; O,H,H}
// (subdiv_1to4.hpp:198-214), written out term by term.  Faces (subdiv_1to4.hpp:262-318): face i becomes (m12, m20, m01) and
// (v0,m01,m20), (v1,m12,m01), (v2,m20,m12) are appended at nf + 3i.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace isd {

// owner[3*fi + e] = 3*fj + e' : the first (face, slot) in the reference's walk that holds edge e of face fi;
// flag = 1 when that is (fi, e) itself.
__global__ void __launch_bounds__(256) k_subdiv_mark(const int* __restrict__ faces, int nf, const unsigned* __restrict__ uoff,
                                                     const int* __restrict__ uitems, unsigned* __restrict__ owner, unsigned* __restrict__ flag) {
    const long long t = (long long)blockIdx.x * 256 + threadIdx.x;
    if (t >= (long long)nf * 3) return;
    const int fi = (int)(t / 3), e = (int)(t - (long long)fi * 3);
    const int a = faces[fi * 3 + e], b = faces[fi * 3 + (e + 1) % 3];
    unsigned own = (unsigned)t;
    const unsigned lo = uoff[a], hi = uoff[a + 1];
    for (unsigned k = lo; k < hi; ++k) {
        const int fj = uitems[k];
        if (fj > fi) break;
        const int w0 = faces[fj * 3], w1 = faces[fj * 3 + 1], w2 = faces[fj * 3 + 2];
        int slot = -1;
        if ((w0 == a && w1 == b) || (w0 == b && w1 == a)) slot = 0;
        else if ((w1 == a && w2 == b) || (w1 == b && w2 == a)) slot = 1;
        else if ((w2 == a && w0 == b) || (w2 == b && w0 == a)) slot = 2;
        if (slot >= 0) { own = (unsigned)fj * 3u + (unsigned)slot; break; }
    }
    owner[t] = own;
    flag[t] = (own == (unsigned)t) ? 1u : 0u;
}

__global__ void __launch_bounds__(128) k_subdiv_emit(const float* __restrict__ verts, int nv, const int* __restrict__ faces, int nf,
                                                     const unsigned* __restrict__ owner, const unsigned* __restrict__ rank,
                                                     float* __restrict__ verts_out, int* __restrict__ faces_out) {
    const int fi = blockIdx.x * 128 + threadIdx.x;
    if (fi >= nf) return;
    const int v[3] = {faces[fi * 3], faces[fi * 3 + 1], faces[fi * 3 + 2]};
    int m[3];
    const float H_ = 0.5f, O_ = 0.0f;
    const float maker[3][3] = {{H_, O_, H_}, {H_, H_, O_}, {O_, H_, H_}};
#pragma unroll
    for (int e = 0; e < 3; ++e) {
        const unsigned t = (unsigned)fi * 3u + (unsigned)e, own = owner[t];
        m[e] = nv + (int)rank[own];
        if (own == t) {
#pragma unroll
            for (int d = 0; d < 3; ++d)
                verts_out[(size_t)m[e] * 3 + d] = verts[(size_t)v[0] * 3 + d] * maker[0][e] + verts[(size_t)v[1] * 3 + d] * maker[1][e] +
                                                  verts[(size_t)v[2] * 3 + d] * maker[2][e];
        }
    }
    const int m01 = m[0], m12 = m[1], m20 = m[2];
    int* o = faces_out + (size_t)fi * 3;
    o[0] = m12; o[1] = m20; o[2] = m01;
    int* q = faces_out + ((size_t)nf + (size_t)fi * 3) * 3;
    q[0] = v[0]; q[1] = m01; q[2] = m20;
    q[3] = v[1]; q[4] = m12; q[5] = m01;
    q[6] = v[2]; q[7] = m20; q[8] = m12;
}

// randomize_verts: v += (rand()/RAND_MAX - 0.5) * amplitude, the bracket and the sum in double (0.5 is a double literal),
// rounded once to binary32 at the store.  `ratio` holds float(rand()) / float(RAND_MAX) in the reference's draw order.
__global__ void __launch_bounds__(256) k_add_noise(float* __restrict__ verts, long long n, const float* __restrict__ ratio, float amplitude) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    verts[i] = (float)((double)verts[i] + ((double)ratio[i] - 0.5) * (double)amplitude);
}

}  // namespace isd

namespace isd {
// z-slab halo mode: a slab's own faces leave the local numbering (scan order of the slab plus its halo) for the global one,
// a constant shift because both are scan order.
__global__ void __launch_bounds__(256) k_rebase_faces(int* __restrict__ faces, long long n, int delta) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i < n) faces[i] += delta;
}
}  // namespace isd
