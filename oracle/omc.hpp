// ORACLE — TEST INFRASTRUCTURE ONLY (see ojson.hpp header).
//
// CPU restatement of MarchingCubes<>::produce_mesh (marching_cubes.hpp, paths relative to
// /root/reference/js_iteration_2/):
//   ctor/init         :174-365     grid (R+5)^3, width = (max-min)/R
//   prepare_grid      :1662-1698   node coordinate (REAL)i*factor + min - 2*width
//   eval_shape        :1702-1725   field += f(grid)      (field starts at 0)
//   seal_exterior     :895-963     index in {0,1,S-2,S-1} on any axis <- -1e7
//   render_geometry   :1019-1072   z,y,x in [1,S-2)
//   polygonize_single_cube :518-718, VIntX/Y/Z :401-495, posnormtriv :727-821
//   flush_geometry_queue :1520-1658  std::map<edge key,int> first-appearance numbering
// The 4096-entry queue is an implementation detail with no effect on output order and is
// not restated.  Evaluation of the grid is done in chunks of whole z-planes (pointwise
// identical to the reference's single whole-grid call) so large grids fit in memory.
#pragma once
#include <cstdint>
#include <map>
#include <vector>
#include "omc_tables.h"
#include "oshapes.hpp"

namespace oracle {

struct bounding_box { REAL xmin, xmax, ymin, ymax, zmin, zmax; };

struct MarchingCubes {
    static const int skip_count_l = 2, skip_count_h = 3;
    int S;                        // nodes per axis = R + 5
    REAL widthx, widthy, widthz;
    bounding_box box;
    std::vector<REAL> field;      // x fastest, then y, then z
    // results
    std::vector<REAL> result_verts;
    std::vector<int> result_faces;
    std::vector<int64_t> result_keys;          // edge key (3*node + axis) of each vertex, for canonical ordering
    std::vector<uint8_t> case_codes;           // per visited cube, scan order (z,y,x in [1,S-2))
    std::map<int64_t, int> e3map;

    // Window of cube layers [cz0, cz1) (ours, for bounded samples of a large grid and for slab parity checks): only node
    // planes [cz0, cz1] are stored and evaluated, cube layers outside the window are not visited; everything else — node
    // coordinates, seal, scan order inside the window, edge keys (global node numbers) — is what the whole grid gives.
    // The default window is the whole grid: the reference's behaviour.
    int cz0, cz1;                 // cube layers visited: [cz0, cz1) within [1, S-2)
    int zb, nzp;                  // node planes stored: [zb, zb + nzp)
    bool keep_cases = true;

    MarchingCubes(int R, const bounding_box& b, int layer0 = -1, int layer1 = -1) : S(R + skip_count_l + skip_count_h), box(b) {
        widthx = (b.xmax - b.xmin) / (REAL)R;
        widthy = (b.ymax - b.ymin) / (REAL)R;
        widthz = (b.zmax - b.zmin) / (REAL)R;
        if (layer0 < 0) { cz0 = 1; cz1 = S - 2; zb = 0; nzp = S; }
        else {
            cz0 = std::max(1, layer0); cz1 = std::min(S - 2, layer1);
            zb = cz0; nzp = cz1 - cz0 + 1;
        }
        field.assign((size_t)S * S * nzp, 0.0f);
    }
    int64_t zoff() const { return (int64_t)zb * (int64_t)zstride(); }
    size_t ystride() const { return (size_t)S; }
    size_t zstride() const { return (size_t)S * S; }

    void eval_shape(const implicit_function& object) {
        REAL wx = box.xmax - box.xmin, wy = box.ymax - box.ymin, wz = box.zmax - box.zmin;
        int e = S - skip_count_l - skip_count_h;
        REAL xfactor = wx / (REAL)e, yfactor = wy / (REAL)e, zfactor = wz / (REAL)e;
        const size_t plane = zstride();
        int planes_per_chunk = (int)std::max<size_t>(1, (size_t)2000000 / plane);
        const int zend = zb + nzp, nchunks = (nzp + planes_per_chunk - 1) / planes_per_chunk;
        // chunks are independent (pointwise evaluation, disjoint slices of the field): threads() > 1 runs them concurrently
        parallel_for(nchunks, [&](long ci) {
            int z0 = zb + (int)ci * planes_per_chunk;
            int z1 = std::min(zend, z0 + planes_per_chunk);
            vects grid((size_t)(z1 - z0) * plane);
            for (int z = z0; z < z1; z++)
                for (int y = 0; y < S; y++)
                    for (int x = 0; x < S; x++) {
                        size_t i = (size_t)x + (size_t)y * ystride() + (size_t)(z - z0) * plane;
                        grid[i][0] = (REAL)x * xfactor + box.xmin - skip_count_l * widthx;
                        grid[i][1] = (REAL)y * yfactor + box.ymin - skip_count_l * widthy;
                        grid[i][2] = (REAL)z * zfactor + box.zmin - skip_count_l * widthz;
                    }
            scalars f(grid.n());
            object.eval_implicit(grid, &f);
            for (size_t i = 0; i < f.size(); i++) field[(size_t)(z0 - zb) * plane + i] += f[i];
        });
    }

    void seal_exterior(const REAL exterior_value) {
        for (int x = 0; x < S; x++)
            for (int y = 0; y < S; y++)
                for (int z = zb; z < zb + nzp; z++) {
                    bool border0 = (x == 0) || (x == S - 1) || (y == 0) || (y == S - 1) || (z == 0) || (z == S - 1);
                    bool border = (x == 1) || (x == S - 2) || (y == 1) || (y == S - 2) || (z == 1) || (z == S - 2);
                    if (border || border0) field[zstride() * (size_t)(z - zb) + x + y * ystride()] = exterior_value;
                }
    }

    // posnormtriv + flush_geometry_queue, fused: one triangle corner
    void corner(const REAL* vlist, const int64_t* e3list, int o) {
        int64_t key = e3list[o / 3];
        auto ins = e3map.emplace(key, (int)(result_verts.size() / 3));
        if (ins.second) {
            result_verts.push_back(vlist[o]);
            result_verts.push_back(vlist[o + 1]);
            result_verts.push_back(vlist[o + 2]);
            result_keys.push_back(key);
        }
        result_faces.push_back(ins.first->second);
    }

    int polygonize_single_cube(REAL fx, REAL fy, REAL fz, int64_t q) {
        const REAL isol = 0.0;
        const int64_t ys = (int64_t)ystride(), zs = (int64_t)zstride();
        int64_t qx = q + 1, qy = q + ys, qz = q + zs, qxy = qx + ys, qxz = qx + zs, qyz = q + ys + zs, qxyz = qx + ys + zs;
        unsigned cubeindex = 0;
        const REAL* fw = field.data() - zoff();      // q is the global node number (edge keys); the window starts at plane zb
        REAL field0 = fw[q], field1 = fw[qx], field2 = fw[qy], field3 = fw[qxy], field4 = fw[qz],
             field5 = fw[qxz], field6 = fw[qyz], field7 = fw[qxyz];
        if (field0 < isol) cubeindex |= 1;
        if (field1 < isol) cubeindex |= 2;
        if (field2 < isol) cubeindex |= 8;
        if (field3 < isol) cubeindex |= 4;
        if (field4 < isol) cubeindex |= 16;
        if (field5 < isol) cubeindex |= 32;
        if (field6 < isol) cubeindex |= 128;
        if (field7 < isol) cubeindex |= 64;
        if (keep_cases) case_codes.push_back((uint8_t)cubeindex);
        int bits = OMC_EDGE_MASK[cubeindex];
        if (bits == 0) return 0;
        REAL fx2 = fx + widthx, fy2 = fy + widthy, fz2 = fz + widthz;
        REAL vlist[36];
        int64_t e3[12];
        auto VIntX = [&](int off, REAL x, REAL y, REAL z, REAL v1, REAL v2, int64_t ijk) {
            REAL mu = (isol - v1) / (v2 - v1);
            vlist[off] = x + mu * widthx; vlist[off + 1] = y; vlist[off + 2] = z;
            e3[off / 3] = ijk * 3;
        };
        auto VIntY = [&](int off, REAL x, REAL y, REAL z, REAL v1, REAL v2, int64_t ijk) {
            REAL mu = (isol - v1) / (v2 - v1);
            vlist[off] = x; vlist[off + 1] = y + mu * widthy; vlist[off + 2] = z;
            e3[off / 3] = ijk * 3 + 1;
        };
        auto VIntZ = [&](int off, REAL x, REAL y, REAL z, REAL v1, REAL v2, int64_t ijk) {
            REAL mu = (isol - v1) / (v2 - v1);
            vlist[off] = x; vlist[off + 1] = y; vlist[off + 2] = z + mu * widthz;
            e3[off / 3] = ijk * 3 + 2;
        };
        if (bits & 1) VIntX(0, fx, fy, fz, field0, field1, q);
        if (bits & 2) VIntY(3, fx2, fy, fz, field1, field3, qx);
        if (bits & 4) VIntX(6, fx, fy2, fz, field2, field3, qy);
        if (bits & 8) VIntY(9, fx, fy, fz, field0, field2, q);
        if (bits & 16) VIntX(12, fx, fy, fz2, field4, field5, qz);
        if (bits & 32) VIntY(15, fx2, fy, fz2, field5, field7, qxz);
        if (bits & 64) VIntX(18, fx, fy2, fz2, field6, field7, qyz);
        if (bits & 128) VIntY(21, fx, fy, fz2, field4, field6, qz);
        if (bits & 256) VIntZ(24, fx, fy, fz, field0, field4, q);
        if (bits & 512) VIntZ(27, fx2, fy, fz, field1, field5, qx);
        if (bits & 1024) VIntZ(30, fx2, fy2, fz, field3, field7, qxy);
        if (bits & 2048) VIntZ(33, fx, fy2, fz, field2, field6, qy);
        int base = (int)cubeindex << 4, i = 0, numtris = 0;
        while (OMC_TRI[base + i] != -1) {
            corner(vlist, e3, 3 * OMC_TRI[base + i]);
            corner(vlist, e3, 3 * OMC_TRI[base + i + 1]);
            corner(vlist, e3, 3 * OMC_TRI[base + i + 2]);
            i += 3;
            numtris++;
        }
        return numtris;
    }

    void render_geometry() {
        const REAL xi0 = (+box.xmin) / widthx - skip_count_l;
        const REAL yi0 = (+box.ymin) / widthy - skip_count_l;
        const REAL zi0 = (+box.zmin) / widthz - skip_count_l;
        const int end = S - 2;
        if (keep_cases) case_codes.reserve((size_t)(cz1 - cz0) * (end - 1) * (end - 1));
        for (int zi = cz0; zi < cz1; zi++) {
            int64_t z_offset = (int64_t)zstride() * zi;
            REAL fz = (zi + zi0) * widthz;
            for (int yi = 1; yi < end; yi++) {
                int64_t y_offset = z_offset + (int64_t)ystride() * yi;
                REAL fy = (yi + yi0) * widthy;
                for (int xi = 1; xi < end; xi++) {
                    REAL fx = (xi + xi0) * widthx;
                    polygonize_single_cube(fx, fy, fz, y_offset + xi);
                }
            }
        }
    }

    void produce_mesh(const implicit_function& object) {
        eval_shape(object);
        seal_exterior(-10000000.0);
        render_geometry();
    }
};

}  // namespace oracle
