"""Test infrastructure: loads oracle/liboracle.so (CPU restatement of the reference) and tests/emu/libemu.so."""
import ctypes
import os
import subprocess

import numpy as np

from implisolid_b200.api import GeometryLibrary

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
EMU_DIR = os.path.join(ROOT, "tests", "emu")


class Oracle(GeometryLibrary):
    progress_symbol = "oracle_set_progress_callback"
    srand_symbol = "oracle_srand"

    def __init__(self, path):
        super().__init__(path, extensions=False)
        L = self.lib
        L.oracle_grid_size.restype = ctypes.c_int
        L.oracle_field_ptr.restype = ctypes.POINTER(ctypes.c_float)
        L.oracle_num_cubes.restype = ctypes.c_long
        L.oracle_case_codes_ptr.restype = ctypes.POINTER(ctypes.c_uint8)
        L.oracle_vertex_keys_ptr.restype = ctypes.POINTER(ctypes.c_int64)
        L.oracle_projection_count.restype = ctypes.c_int
        L.oracle_projected_centroids_ptr.restype = ctypes.POINTER(ctypes.c_float)
        L.oracle_centroid_normals_ptr.restype = ctypes.POINTER(ctypes.c_float)
        L.oracle_average_edge.restype = ctypes.c_float
        L.oracle_alpha_list.restype = ctypes.c_int
        L.oracle_get_threads.restype = ctypes.c_int

    def set_threads(self, n):
        """Host threads for the pointwise whole-array passes (1 = the reference's single-threaded calls; same bits either way)."""
        self.lib.oracle_set_threads(int(n))

    def build_layers(self, shape_json, mc_json, layer0, layer1):
        """build_geometry restricted to cube layers [layer0, layer1) of the grid (MarchingCubes' window, oracle/omc.hpp)."""
        self.lib.oracle_build_layers(shape_json.encode(), mc_json.encode(), int(layer0), int(layer1))

    def polygonize_layers(self, shape_json, mc_json, layer0, layer1):
        self.build_layers(shape_json, mc_json, layer0, layer1)
        v, f = self.get_v(), self.get_f()
        self.finish_geometry()
        return v, f

    def stage_seconds(self):
        buf = (ctypes.c_double * 4)()
        self.lib.oracle_stage_seconds(buf)
        return {"eval": buf[0], "marching_cubes": buf[1], "resample": buf[2], "projection_qem": buf[3]}

    def grid_size(self):
        return self.lib.oracle_grid_size()

    def get_field(self):
        S = self.grid_size()
        return np.ctypeslib.as_array(self.lib.oracle_field_ptr(), shape=(S, S, S)).copy()

    def get_case_codes(self):
        n = self.grid_size() - 3
        return np.ctypeslib.as_array(self.lib.oracle_case_codes_ptr(), shape=(n, n, n)).copy()

    def vertex_keys(self):
        n = self.get_v_size()
        return np.ctypeslib.as_array(self.lib.oracle_vertex_keys_ptr(), shape=(n,)).copy()

    def stats(self):
        buf = (ctypes.c_long * 9)()
        self.lib.oracle_get_stats(buf)
        return {"projection_by_type": list(buf[0:7]), "projection_failed": buf[7], "bisection_capped": buf[8]}

    def invariants(self):
        """The reference's ASSERT_USED checks around the bisection, counted by the last build (oob02.hpp Stats)."""
        buf = (ctypes.c_long * 2)()
        self.lib.oracle_get_invariants(buf)
        return {"not_conjugated": buf[0], "unresolved_after_bisection": buf[1]}

    def tetrahedron_integrity(self, corners, m12=None):
        c = np.ascontiguousarray(corners, dtype=np.float32).reshape(12)
        m = np.ascontiguousarray(m12 if m12 is not None else [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0], dtype=np.float32)
        return int(self.lib.oracle_tetrahedron_integrity(c.ctypes.data_as(ctypes.c_void_p), m.ctypes.data_as(ctypes.c_void_p)))

    def tetrahedron_planes(self, corners, m12=None):
        c = np.ascontiguousarray(corners, dtype=np.float32).reshape(12)
        m = np.ascontiguousarray(m12 if m12 is not None else [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0], dtype=np.float32)
        out = np.zeros((4, 4), dtype=np.float32)
        self.lib.oracle_tetrahedron_planes(c.ctypes.data_as(ctypes.c_void_p), m.ctypes.data_as(ctypes.c_void_p), out.ctypes.data_as(ctypes.c_void_p))
        return out

    def sym_eig3(self, A):
        a = np.ascontiguousarray(A, dtype=np.float32).reshape(9)
        w = np.zeros(3, dtype=np.float32)
        V = np.zeros((3, 3), dtype=np.float32)
        self.lib.oracle_sym_eig3(a.ctypes.data_as(ctypes.c_void_p), w.ctypes.data_as(ctypes.c_void_p), V.ctypes.data_as(ctypes.c_void_p))
        return w, V

    def mt11213b_raw(self, seed, n):
        out = np.zeros(n, dtype=np.uint32)
        self.lib.oracle_mt11213b_raw(ctypes.c_uint(seed), int(n), out.ctypes.data_as(ctypes.c_void_p))
        return out

    def projected_centroids(self):
        n = self.lib.oracle_projection_count()
        c = np.ctypeslib.as_array(self.lib.oracle_projected_centroids_ptr(), shape=(n, 3)).copy()
        nr = np.ctypeslib.as_array(self.lib.oracle_centroid_normals_ptr(), shape=(n, 3)).copy()
        return c, nr

    def average_edge(self):
        return float(self.lib.oracle_average_edge())

    def faces_of_faces(self, faces, nv=None):
        f = np.ascontiguousarray(faces, dtype=np.int32).reshape(-1, 3)
        out = np.empty_like(f)
        self.lib.oracle_faces_of_faces(f.ctypes.data_as(ctypes.c_void_p), f.shape[0], out.ctypes.data_as(ctypes.c_void_p))
        return out

    def umbrella(self, faces, nv):
        f = np.ascontiguousarray(faces, dtype=np.int32).reshape(-1, 3)
        off = np.empty(nv + 1, dtype=np.int32)
        items = np.empty(f.size, dtype=np.int32)
        self.lib.oracle_umbrella(f.ctypes.data_as(ctypes.c_void_p), f.shape[0], int(nv), off.ctypes.data_as(ctypes.c_void_p),
                                 items.ctypes.data_as(ctypes.c_void_p))
        return off, items

    def alpha_list(self, average_edge):
        buf = np.zeros(1024, dtype=np.float32)
        n = self.lib.oracle_alpha_list(ctypes.c_float(average_edge), buf.ctypes.data_as(ctypes.c_void_p), 1024)
        return buf[:n].copy()

    def invert_matrix(self, m12):
        a = np.ascontiguousarray(m12, dtype=np.float32)
        out = np.zeros(12, dtype=np.float32)
        self.lib.oracle_invert_matrix(a.ctypes.data_as(ctypes.c_void_p), out.ctypes.data_as(ctypes.c_void_p))
        return out

    def random_pm1(self, n, amplitude):
        out = np.zeros((n, 3), dtype=np.float32)
        self.lib.oracle_random_pm1(int(n), ctypes.c_float(amplitude), out.ctypes.data_as(ctypes.c_void_p))
        return out

    def ob02_on_mesh(self, shape_json, verts, faces, resample_iters=0, c=1.0, projection=0, qem=0):
        v = np.ascontiguousarray(verts, dtype=np.float32).reshape(-1, 3).copy()
        f = np.ascontiguousarray(faces, dtype=np.int32).reshape(-1, 3)
        self.lib.oracle_ob02_on_mesh(shape_json.encode(), v.ctypes.data_as(ctypes.c_void_p), v.shape[0], f.ctypes.data_as(ctypes.c_void_p),
                                     f.shape[0], int(resample_iters), ctypes.c_float(c), int(projection), int(qem))
        return v

    def set_sdf_grid(self, values, h, origin):
        v = np.ascontiguousarray(values, dtype=np.float32)
        nz, ny, nx = v.shape
        self.lib.oracle_set_sdf_grid(nx, ny, nz, ctypes.c_float(h), ctypes.c_float(origin[0]), ctypes.c_float(origin[1]),
                                     ctypes.c_float(origin[2]), v.ctypes.data_as(ctypes.c_void_p))


def build_oracle():
    subprocess.run(["make", "-C", ORACLE_DIR], check=True, stdout=subprocess.DEVNULL)
    return os.path.join(ORACLE_DIR, "liboracle.so")


def _cpu_tag():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    import hashlib
                    return hashlib.sha1(line.encode()).hexdigest()
    except OSError:
        pass
    return "unknown"


def build_oracle_mt():
    """liboracle_mt.so is built -march=native: rebuild it when this machine's CPU is not the one it was built on (the .so
    travels to the GPU box with the snapshot)."""
    tag_path, so = os.path.join(ORACLE_DIR, "liboracle_mt.cpu"), os.path.join(ORACLE_DIR, "liboracle_mt.so")
    tag = _cpu_tag()
    have = open(tag_path).read().strip() if os.path.exists(tag_path) else ""
    if have != tag and os.path.exists(so):
        os.remove(so)
    subprocess.run(["make", "-C", ORACLE_DIR, "mt"], check=True, stdout=subprocess.DEVNULL)
    with open(tag_path, "w") as f:
        f.write(tag)
    return so


_oracle = None
_oracle_mt = None


def load_oracle():
    global _oracle
    if _oracle is None:
        _oracle = Oracle(build_oracle())
    return _oracle


def host_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def load_oracle_mt(threads=None):
    """The multi-threaded -O3 -march=native build of the same oracle sources (bit-identical results): the timed CPU baseline
    and parity cases too large for one thread."""
    global _oracle_mt
    if _oracle_mt is None:
        _oracle_mt = Oracle(build_oracle_mt())
    _oracle_mt.set_threads(threads if threads else host_threads())
    return _oracle_mt


class Emu:
    def __init__(self, path):
        self.lib = ctypes.CDLL(path)
        self.lib.emu_compile.restype = ctypes.c_int
        self.lib.emu_eval_masked.restype = ctypes.c_int

    def compile(self, shape_json, ignore_root_matrix=False):
        words = np.zeros(16384, dtype=np.uint32)
        stats = np.zeros(4, dtype=np.float64)
        n = self.lib.emu_compile(shape_json.encode(), int(ignore_root_matrix), words.ctypes.data_as(ctypes.c_void_p), 16384,
                                 stats.ctypes.data_as(ctypes.c_void_p))
        return words[:n].copy(), {"value_depth": int(stats[0]), "frame_depth": int(stats[1]), "nodes": int(stats[2]), "flops": stats[3]}

    def eval(self, shape_json, pts, ignore_root_matrix=False):
        p = np.ascontiguousarray(pts, dtype=np.float32).reshape(-1, 3)
        n = p.shape[0]
        f = np.zeros(n, np.float32); f4 = np.zeros(n, np.float32); g = np.zeros((n, 3), np.float32)
        self.lib.emu_eval(shape_json.encode(), int(ignore_root_matrix), p.ctypes.data_as(ctypes.c_void_p), n,
                          f.ctypes.data_as(ctypes.c_void_p), f4.ctypes.data_as(ctypes.c_void_p), g.ctypes.data_as(ctypes.c_void_p))
        return f, f4, g

    def slab_balance(self, hist, layers, world, cost_us, fixed_us, rounds, threshold_permille=1030):
        """csrc/slab_balance.hpp on a synthetic cost model: (cuts per round [rounds+1, world+1], busy times [rounds+1, world], re-cuts)."""
        hist = np.ascontiguousarray(hist, dtype=np.uint32)
        cost = np.ascontiguousarray(cost_us, dtype=np.int64)
        assert cost.shape[0] == layers and hist.shape[0] >= ((layers + 1) >> 3) + 1
        cuts = np.zeros((rounds + 1, world + 1), dtype=np.int32)
        busy = np.zeros((rounds + 1, world), dtype=np.int64)
        self.lib.emu_slab_balance.restype = ctypes.c_int
        n = self.lib.emu_slab_balance(hist.ctypes.data_as(ctypes.c_void_p), int(hist.shape[0]), int(layers), int(world), cost.ctypes.data_as(ctypes.c_void_p),
                                      ctypes.c_longlong(int(fixed_us)), int(rounds), int(threshold_permille), cuts.ctypes.data_as(ctypes.c_void_p),
                                      busy.ctypes.data_as(ctypes.c_void_p))
        return cuts, busy, n

    def eval_masked(self, shape_json, pts):
        """(tree value, interpreter value with only the selected children live, live node count per point, node count)."""
        p = np.ascontiguousarray(pts, dtype=np.float32).reshape(-1, 3)
        n = p.shape[0]
        ft = np.zeros(n, np.float32); fm = np.zeros(n, np.float32); live = np.zeros(n, np.int32)
        nodes = self.lib.emu_eval_masked(shape_json.encode(), p.ctypes.data_as(ctypes.c_void_p), n, ft.ctypes.data_as(ctypes.c_void_p),
                                         fm.ctypes.data_as(ctypes.c_void_p), live.ctypes.data_as(ctypes.c_void_p))
        return ft, fm, live, nodes

    def eval_packed_mismatches(self, shape_json, pts):
        """Packed-pair interpreters (two points per instruction stream) against the scalar list interpreters and the full tree;
        returns the number of mismatching results (-1: the tree cannot be specialised)."""
        p = np.ascontiguousarray(pts, dtype=np.float32).reshape(-1, 3)
        self.lib.emu_eval_packed.restype = ctypes.c_int
        return int(self.lib.emu_eval_packed(shape_json.encode(), p.ctypes.data_as(ctypes.c_void_p), p.shape[0]))

    def bounds_check(self, shape_json, boxes, samples=16, seed=1):
        """Interval bounds / liveness masks of csg_bounds.cuh against the exact interpreters on random points of every box
        (centre xyz, half extents xyz).  Returns (node count or -count when the tree cannot be specialised, dict of counters)."""
        b = np.ascontiguousarray(boxes, dtype=np.float32).reshape(-1, 6)
        out = np.zeros(6, dtype=np.int64)
        self.lib.emu_bounds_check.restype = ctypes.c_int
        nodes = self.lib.emu_bounds_check(shape_json.encode(), b.ctypes.data_as(ctypes.c_void_p), b.shape[0], int(samples), int(seed),
                                          out.ctypes.data_as(ctypes.c_void_p))
        keys = ["bounds_violations", "value_mismatches", "gradient_mismatches", "looser_second_level", "live_nodes", "boxes_decided"]
        return nodes, dict(zip(keys, (int(v) for v in out)))

    def affine_inverse(self, m12):
        a = np.ascontiguousarray(m12, dtype=np.float32)
        out = np.zeros(12, dtype=np.float32)
        self.lib.emu_affine_inverse(a.ctypes.data_as(ctypes.c_void_p), out.ctypes.data_as(ctypes.c_void_p))
        return out


_emu = None


def load_emu():
    global _emu
    if _emu is None:
        out = os.path.join(EMU_DIR, "libemu.so")
        src = os.path.join(EMU_DIR, "emu_eval.cpp")
        subprocess.run(["g++", "-std=c++17", "-O2", "-fPIC", "-ffp-contract=off", "-shared", "-x", "c++", "-o", out, src], check=True)
        _emu = Emu(out)
    return _emu
