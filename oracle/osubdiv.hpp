// ORACLE — TEST INFRASTRUCTURE ONLY (see ojson.hpp header).
//
// CPU restatement of polygonize_step_3 — the 1-to-4 subdivision of every face that grand_algorithm runs last when
// "subdiv.enabled" (the default) — paths relative to /root/reference/js_iteration_2/:
//   polygonize_step_3                polygonizer_algorithm_ob02.hpp:119-156   noise = post_subdiv_noise * 10 on the last repeat
//   my_subdiv_                       centroids_projection.cpp:1314-1367        every face requested, then randomize_verts
//   subdivide_given_faces            subdivision/do_subdivision.hpp:26-51      returns right after the 1-to-4 pass ("premature")
//   subdivide_multiple_facets_1to4   subdivision/subdiv_1to4.hpp:33-330        midpoint map keyed by the sorted vertex pair
//   easy_edge / encode_edge__sort    mesh_algorithms.hpp:35-48
//   randomize_verts                  basic_functions.hpp:551-557               libc rand(), never seeded => glibc's seed 1
// The midpoints come from an Eigen 3x3 product with the constant matrix {H,O,H; H,H,O; O,H,H} (subdiv_1to4.hpp:198-214): each
// is va*0.5 + vb*0.5 (+ vc*0), a single rounding whatever the summation order, so Eigen's absence does not matter here.
// glibc's rand() (TYPE_3 additive feedback, r[i] = r[i-3] + r[i-31]) is restated with a private state so that the oracle and
// the library under test do not disturb each other through the process-wide generator; tests check it against libc.
#pragma once
#include <cstdint>
#include <map>
#include <set>
#include <utility>
#include <vector>
#include "oshapes.hpp"

namespace oracle {

struct GlibcRand {            // stdlib/random_r.c: __srandom_r + __random_r for TYPE_3 (degree 31, separation 3)
    int32_t r[34];
    uint32_t ring[31];
    int f, b;
    GlibcRand() { seed(1); }
    void seed(unsigned s) {
        int32_t word = (int32_t)(s ? s : 1);
        int32_t st[31];
        st[0] = word;
        for (int i = 1; i < 31; i++) {
            long hi = word / 127773, lo = word % 127773;
            word = (int32_t)(16807 * lo - 2836 * hi);
            if (word < 0) word += 2147483647;
            st[i] = word;
        }
        for (int i = 0; i < 31; i++) ring[i] = (uint32_t)st[i];
        f = 3; b = 0;
        for (int i = 0; i < 310; i++) next();
    }
    int next() {
        ring[f] += ring[b];
        const uint32_t result = ring[f] >> 1;
        if (++f >= 31) f = 0;
        if (++b >= 31) b = 0;
        return (int)result;
    }
};
inline GlibcRand& subdiv_rand() { static GlibcRand g; return g; }

inline void randomize_verts(std::vector<REAL>& verts, REAL amplitude) {
    GlibcRand& g = subdiv_rand();
    for (size_t i = 0; i < verts.size(); ++i)
        verts[i] += (static_cast<REAL>(g.next()) / static_cast<REAL>(2147483647) - 0.5) * amplitude;
}

typedef long edge_pair_type;
inline edge_pair_type easy_edge(int v1, int v2) {       // encode_edge__sort with CONFIG_C::edgecode_base = 1000000
    const edge_pair_type base = 1000000;
    edge_pair_type e1 = v1, e2 = v2;
    if (e1 > e2) std::swap(e1, e2);
    return e1 + base * e2;
}

// subdivide_multiple_facets_1to4 with every face requested and an empty midpoint map
inline void subdivide_all_1to4(std::vector<REAL>& verts, std::vector<int>& faces) {
    const int nf = (int)(faces.size() / 3), nv0 = (int)(verts.size() / 3);
    std::set<int> requested;
    for (int i = 0; i < nf; ++i) requested.insert(i);
    // NB: the reference's key base is 10^6 (mesh_algorithms.hpp:35-48): beyond a million vertices its keys collide.  The
    // oracle keys on the exact pair instead, which is what the key means.
    std::map<std::pair<int, int>, int> midpoint_map;
    std::vector<int> actual(3 * (size_t)nf), original(3 * (size_t)nf);
    std::vector<char> insert_which(3 * (size_t)nf);
    int total = nv0, req = 0;
    for (int fi : requested) {
        const int v[3] = {faces[fi * 3], faces[fi * 3 + 1], faces[fi * 3 + 2]};
        for (int k = 0; k < 3; ++k) original[req * 3 + k] = v[k];
        const std::pair<int, int> e[3] = {std::minmax(v[0], v[1]), std::minmax(v[1], v[2]), std::minmax(v[2], v[0])};
        for (int vii = 0; vii < 3; ++vii) {
            auto it = midpoint_map.find(e[vii]);
            if (it != midpoint_map.end()) {
                insert_which[req * 3 + vii] = 0;
                actual[req * 3 + vii] = it->second;
            } else {
                insert_which[req * 3 + vii] = 1;
                actual[req * 3 + vii] = total;
                midpoint_map[e[vii]] = total;
                ++total;
            }
        }
        ++req;
    }
    std::vector<REAL> nverts((size_t)total * 3);
    for (size_t i = 0; i < verts.size(); ++i) nverts[i] = verts[i];
    const REAL H_ = 0.5, O_ = 0.0;
    const REAL maker[3][3] = {{H_, O_, H_}, {H_, H_, O_}, {O_, H_, H_}};
    int at = nv0;
    for (int r = 0; r < nf; ++r) {
        const int v0 = original[r * 3], v1 = original[r * 3 + 1], v2 = original[r * 3 + 2];
        for (int vii = 0; vii < 3; ++vii) {
            if (!insert_which[r * 3 + vii]) continue;
            for (int d = 0; d < 3; ++d)
                nverts[(size_t)at * 3 + d] = verts[(size_t)v0 * 3 + d] * maker[0][vii] + verts[(size_t)v1 * 3 + d] * maker[1][vii] + verts[(size_t)v2 * 3 + d] * maker[2][vii];
            ++at;
        }
    }
    std::vector<int> nfaces((size_t)nf * 4 * 3);
    for (size_t i = 0; i < faces.size(); ++i) nfaces[i] = faces[i];
    int out = nf;
    for (int r = 0; r < nf; ++r) {
        const int v0 = original[r * 3], v1 = original[r * 3 + 1], v2 = original[r * 3 + 2];
        const int m01 = actual[r * 3], m12 = actual[r * 3 + 1], m20 = actual[r * 3 + 2];
        nfaces[r * 3] = m12; nfaces[r * 3 + 1] = m20; nfaces[r * 3 + 2] = m01;
        const int tri[3][3] = {{v0, m01, m20}, {v1, m12, m01}, {v2, m20, m12}};
        for (int t = 0; t < 3; ++t, ++out)
            for (int k = 0; k < 3; ++k) nfaces[(size_t)out * 3 + k] = tri[t][k];
    }
    verts.swap(nverts);
    faces.swap(nfaces);
}

// polygonize_step_3
inline void polygonize_step_3(std::vector<REAL>& verts, std::vector<int>& faces, REAL post_subdiv_noise, bool is_last) {
    const REAL scale_noise_according_to_matrix = 1.0 * 10.0;
    const REAL actual_noise = is_last ? post_subdiv_noise * scale_noise_according_to_matrix : 0;
    subdivide_all_1to4(verts, faces);
    randomize_verts(verts, actual_noise);
}

}  // namespace oracle
