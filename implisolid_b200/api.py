"""ctypes binding of libimplisolid_b200.so, named after the reference's extern "C" API (mcc2.cpp:88-134).

Mirrors what `implisolid_main.js:38-59` / `js/worker_api.js:121-127` do with `Module.cwrap`: every method is the C
entry point of the same name, taking the two JSON strings and returning numpy views/copies of the result buffers
(`HEAPF32.subarray` / `HEAPU32.subarray` in the reference, `implisolid_main.js:230-237`).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libimplisolid_b200.so")


def library_path():
    return _LIB


def build_library(force=False):
    """Compile csrc/ into libimplisolid_b200.so with nvcc for sm_100a (cross-compiles without a GPU)."""
    csrc = os.path.join(_HERE, "csrc")
    if force and os.path.exists(_LIB):
        os.remove(_LIB)
    subprocess.run(["make", "-C", csrc], check=True, stdout=subprocess.DEVNULL)
    return _LIB


_c_float_p = ctypes.POINTER(ctypes.c_float)
_c_int_p = ctypes.POINTER(ctypes.c_int)

# name -> (restype, argtypes); the same list tests/test_abi.py checks against include/implisolid_b200.h
SIGNATURES = {
    "build_geometry": (None, [ctypes.c_char_p, ctypes.c_char_p]),
    "build_geometry_u": (None, [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p]),
    "get_v_size": (ctypes.c_int, []),
    "get_f_size": (ctypes.c_int, []),
    "get_v": (None, [ctypes.c_void_p, ctypes.c_int]),
    "get_f": (None, [ctypes.c_void_p, ctypes.c_int]),
    "get_v_ptr": (ctypes.c_void_p, []),
    "get_f_ptr": (ctypes.c_void_p, []),
    "finish_geometry": (None, []),
    "set_object": (ctypes.c_int, [ctypes.c_char_p, ctypes.c_bool]),
    "unset_object": (ctypes.c_bool, [ctypes.c_int]),
    "set_x": (ctypes.c_bool, [ctypes.c_void_p, ctypes.c_int]),
    "unset_x": (None, []),
    "calculate_implicit_values": (None, []),
    "get_values_ptr": (ctypes.c_void_p, []),
    "get_values_size": (ctypes.c_int, []),
    "calculate_implicit_gradients": (None, [ctypes.c_bool]),
    "get_gradients_ptr": (ctypes.c_void_p, []),
    "get_gradients_size": (ctypes.c_int, []),
    "get_pointset_ptr": (ctypes.c_void_p, [ctypes.c_char_p]),
    "get_pointset_size": (ctypes.c_int, [ctypes.c_char_p]),
    "about": (None, []),
}
EXT_SIGNATURES = {
    "isd_init": (ctypes.c_int, [ctypes.c_int]),
    "isd_shutdown": (None, []),
    "isd_stream": (ctypes.c_uint64, []),
    "isd_build_geometry_device": (None, [ctypes.c_char_p, ctypes.c_char_p]),
    "isd_device_verts": (ctypes.c_void_p, []),
    "isd_device_faces": (ctypes.c_void_p, []),
    "isd_last_timings": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]),
    "isd_last_counters": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]),
    "isd_debug_spoil_speculation": (None, [ctypes.c_int]),
    "isd_last_program_stats": (ctypes.c_int, [ctypes.c_void_p]),
    "isd_last_kernel_times": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]),
    "isd_kernel_name": (ctypes.c_char_p, [ctypes.c_int]),
    "isd_last_kernel_flops": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]),
    "isd_last_kernel_gaps": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]),
    "isd_grid_size": (ctypes.c_int, []),
    "isd_get_field": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int64]),
    "isd_get_case_codes": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int64]),
    "isd_average_edge": (ctypes.c_float, []),
    "isd_get_projected_centroids": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int64]),
    "isd_debug_adjacency": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]),
    "isd_get_centroid_normals": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int64]),
    "isd_set_sdf_grid": (None, [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_float, ctypes.c_float, ctypes.c_float,
                                ctypes.c_float, ctypes.c_void_p]),
    "isd_ob02_on_mesh": (None, [ctypes.c_char_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                ctypes.c_float, ctypes.c_int, ctypes.c_int]),
    "isd_faces_of_faces": (None, [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]),
    "isd_umbrella": (None, [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]),
    "isd_umbrella_noalloc": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]),
    "isd_umbrella_noalloc_max_vertices": (ctypes.c_int, []),
    "isd_eval_points": (None, [ctypes.c_char_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p]),
    "isd_set_progress_callback": (None, [ctypes.c_void_p, ctypes.c_void_p]),
    "isd_get_n_ptr": (ctypes.c_void_p, []),
    "isd_get_n": (None, [ctypes.c_void_p, ctypes.c_int]),
    "isd_device_normals": (ctypes.c_void_p, []),
    "isd_srand": (None, [ctypes.c_uint]),
    "isd_dist_unique_id": (ctypes.c_int, [ctypes.c_void_p]),
    "isd_dist_init": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.c_void_p]),
    "isd_dist_finalize": (None, []),
    "isd_dist_suspend": (None, []),
    "isd_dist_resume": (None, []),
    "isd_dist_emulate": (ctypes.c_int, [ctypes.c_int, ctypes.c_int]),
    "isd_dist_set_offsets": (None, [ctypes.c_int64, ctypes.c_int64]),
    "isd_dist_last_slab": (ctypes.c_int, [ctypes.c_void_p]),
    "isd_dist_gather_mesh": (None, []),
    "isd_dist_set_edge_total": (None, [ctypes.c_uint64, ctypes.c_int64]),
    "isd_dist_last_edge_sum": (ctypes.c_uint64, []),
    "isd_dist_last_halo": (ctypes.c_int, []),
}

TIMING_NAMES = ["field", "classify_scan", "emit", "resample", "project", "qem", "d2h", "total_host", "subdiv", "gather"]


def _b(s):
    return s if isinstance(s, bytes) else s.encode("utf-8")


class GeometryLibrary:
    """One loaded copy of a library exporting the reference's geometry API (ours, or any drop-in with the same ABI)."""

    def __init__(self, path, extensions=True):
        if not os.path.exists(path):
            raise RuntimeError(
                "%s is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a). There is no CPU fallback." % path)
        self.path = path
        self.lib = ctypes.CDLL(path)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(self.lib, name)
            fn.restype, fn.argtypes = res, args
        self.has_extensions = extensions
        if extensions:
            for name, (res, args) in EXT_SIGNATURES.items():
                fn = getattr(self.lib, name)
                fn.restype, fn.argtypes = res, args

    # ---- group 1 ----
    def build_geometry(self, shape_json, mc_json):
        self.lib.build_geometry(_b(shape_json), _b(mc_json))

    def build_geometry_u(self, shape_json, mc_json, call_specs="{}"):
        self.lib.build_geometry_u(_b(shape_json), _b(mc_json), _b(call_specs))

    def get_v_size(self):
        return self.lib.get_v_size()

    def get_f_size(self):
        return self.lib.get_f_size()

    def get_v(self):
        n = self.get_v_size()
        out = np.empty((n, 3), dtype=np.float32)
        self.lib.get_v(out.ctypes.data, n)
        return out

    def get_f(self):
        n = self.get_f_size()
        out = np.empty((n, 3), dtype=np.int32)
        self.lib.get_f(out.ctypes.data, n)
        return out

    def get_v_view(self):
        """Borrowed view on the library's own storage (valid until the next build), like HEAPF32.subarray."""
        n = self.get_v_size()
        p = self.lib.get_v_ptr()
        return np.ctypeslib.as_array(ctypes.cast(p, _c_float_p), shape=(n, 3)) if n else np.empty((0, 3), np.float32)

    def get_f_view(self):
        n = self.get_f_size()
        p = self.lib.get_f_ptr()
        return np.ctypeslib.as_array(ctypes.cast(p, _c_int_p), shape=(n, 3)) if n else np.empty((0, 3), np.int32)

    def finish_geometry(self):
        self.lib.finish_geometry()

    # ---- progressive results (isd_set_progress_callback; the oracle exports the same under its own prefix) ----
    PROGRESS_FN = ctypes.CFUNCTYPE(None, ctypes.POINTER(ctypes.c_float), ctypes.c_int, ctypes.POINTER(ctypes.c_int), ctypes.c_int,
                                   ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p)
    progress_symbol = "isd_set_progress_callback"
    srand_symbol = "isd_srand"

    def polygonize_progressive(self, shape_json, mc_json, call_specs=None):
        """build_geometry_u with a progress callback installed: returns the list of reports
        [(verts, faces, (progress_callback_id, shape_id, call_id)), ...] and the final (verts, faces)."""
        reports = []

        def on_progress(v, nv, f, nf, cb_id, shape_id, call_id, _user):
            vv = np.ctypeslib.as_array(v, shape=(nv, 3)).copy() if nv else np.empty((0, 3), np.float32)
            ff = np.ctypeslib.as_array(f, shape=(nf, 3)).copy() if nf else np.empty((0, 3), np.int32)
            reports.append((vv, ff, (cb_id, shape_id, call_id)))

        cb = self.PROGRESS_FN(on_progress)
        setter = getattr(self.lib, self.progress_symbol)
        setter.restype, setter.argtypes = None, [ctypes.c_void_p, ctypes.c_void_p]
        setter(ctypes.cast(cb, ctypes.c_void_p), None)
        try:
            if call_specs is None:
                self.build_geometry(shape_json, mc_json)
            else:
                self.build_geometry_u(shape_json, mc_json, call_specs)
            v, f = self.get_v(), self.get_f()
            self.finish_geometry()
        finally:
            setter(None, None)
        return reports, (v, f)

    def srand(self, seed=1):
        fn = getattr(self.lib, self.srand_symbol)
        fn.restype, fn.argtypes = None, [ctypes.c_uint]
        fn(int(seed))

    def polygonize(self, shape_json, mc_json):
        """build_geometry -> copies of (verts, faces) -> finish_geometry: impli2.make_geometry (implisolid_main.js:201-249)."""
        self.build_geometry(shape_json, mc_json)
        v, f = self.get_v(), self.get_f()
        self.finish_geometry()
        return v, f

    # ---- group 2 ----
    def query_implicit_values(self, shape_json, points, ignore_root_matrix=False):
        """implisolid_main.js:291-331."""
        pts = np.ascontiguousarray(points, dtype=np.float32).reshape(-1, 3)
        oid = self.lib.set_object(_b(shape_json), bool(ignore_root_matrix))
        if not self.lib.set_x(pts.ctypes.data, pts.shape[0]):
            self.lib.unset_object(oid)
            raise ValueError("set_x rejected %d points" % pts.shape[0])
        self.lib.calculate_implicit_values()
        n = self.lib.get_values_size()
        out = np.ctypeslib.as_array(ctypes.cast(self.lib.get_values_ptr(), _c_float_p), shape=(n,)).copy() if n else np.empty(0, np.float32)
        self.lib.unset_x()
        self.lib.unset_object(oid)
        return out

    def query_normals(self, shape_json, points, ignore_root_matrix=False, normalize_and_invert=True):
        """implisolid_main.js:169-189."""
        pts = np.ascontiguousarray(points, dtype=np.float32).reshape(-1, 3)
        oid = self.lib.set_object(_b(shape_json), bool(ignore_root_matrix))
        if not self.lib.set_x(pts.ctypes.data, pts.shape[0]):
            self.lib.unset_object(oid)
            raise ValueError("set_x rejected %d points" % pts.shape[0])
        self.lib.calculate_implicit_gradients(bool(normalize_and_invert))
        n = self.lib.get_gradients_size()
        out = (np.ctypeslib.as_array(ctypes.cast(self.lib.get_gradients_ptr(), _c_float_p), shape=(n,)).copy().reshape(-1, 3)
               if n else np.empty((0, 3), np.float32))
        self.lib.unset_x()
        self.lib.unset_object(oid)
        return out

    # ---- extensions ----
    def build_geometry_device(self, shape_json, mc_json):
        self.lib.isd_build_geometry_device(_b(shape_json), _b(mc_json))

    def get_n(self):
        """Per-vertex normals of the last build (isd_get_n), shape (Nv, 3)."""
        n = self.get_v_size()
        out = np.empty((n, 3), dtype=np.float32)
        self.lib.isd_get_n(out.ctypes.data, n)
        return out

    def get_n_view(self):
        n = self.get_v_size()
        p = self.lib.isd_get_n_ptr()
        return np.ctypeslib.as_array(ctypes.cast(p, _c_float_p), shape=(n, 3)) if n else np.empty((0, 3), np.float32)

    def last_timings(self):
        buf = np.zeros(len(TIMING_NAMES), dtype=np.float32)
        self.lib.isd_last_timings(buf.ctypes.data, len(TIMING_NAMES))
        return dict(zip(TIMING_NAMES, [float(x) for x in buf]))

    def last_counters(self):
        buf = np.zeros(18, dtype=np.int64)
        self.lib.isd_last_counters(buf.ctypes.data, 18)
        return {"nodes": int(buf[0]), "cubes": int(buf[1]), "cut_cubes": int(buf[2]),
                "projection_by_type": [int(x) for x in buf[3:10]], "projection_failed": int(buf[10]),
                "bisection_capped": int(buf[11]), "bisected": int(buf[12]), "launches": int(buf[13]),
                "field_bytes": int(buf[14]), "nodes_full_eval": int(buf[15]),
                "speculative": int(buf[16]), "speculation_redone": int(buf[17])}

    def spoil_speculation(self, what=1):
        self.lib.isd_debug_spoil_speculation(what)

    def last_kernel_times(self):
        """{kernel name: (ms summed over its launches, launches)} of the last build; needs "gpu":{"kernel_times":1}."""
        ms = np.zeros(64, dtype=np.float32)
        cnt = np.zeros(64, dtype=np.int32)
        n = self.lib.isd_last_kernel_times(ms.ctypes.data, cnt.ctypes.data, 64)
        return {self.lib.isd_kernel_name(i).decode(): (float(ms[i]), int(cnt[i])) for i in range(n) if cnt[i]}

    def last_kernel_gaps(self):
        """{kernel name: ms the stream spent between the previous timed kernel and this one} (memsets, copies, host waits)."""
        ms = np.zeros(64, dtype=np.float32)
        n = self.lib.isd_last_kernel_gaps(ms.ctypes.data, 64)
        return {self.lib.isd_kernel_name(i).decode(): float(ms[i]) for i in range(n) if ms[i] > 0}

    def last_kernel_flops(self):
        """{kernel name: FP32 operations its CSG evaluations executed} for the last build; needs "gpu":{"count_flops":1}."""
        fl = np.zeros(64, dtype=np.float64)
        n = self.lib.isd_last_kernel_flops(fl.ctypes.data, 64)
        return {self.lib.isd_kernel_name(i).decode(): float(fl[i]) for i in range(n) if fl[i] > 0}

    def last_program_stats(self):
        buf = np.zeros(4, dtype=np.float64)
        self.lib.isd_last_program_stats(buf.ctypes.data)
        return {"flops_per_point": float(buf[0]), "nodes": int(buf[1]), "words": int(buf[2]), "value_depth": int(buf[3])}

    def grid_size(self):
        return self.lib.isd_grid_size()

    def get_field(self):
        S = self.grid_size()
        out = np.empty((S, S, S), dtype=np.float32)      # [z][y][x]
        rc = self.lib.isd_get_field(out.ctypes.data, out.size)
        if rc:
            raise RuntimeError("isd_get_field failed (%d)" % rc)
        return out

    def get_case_codes(self):
        n = self.grid_size() - 3
        out = np.empty((n, n, n), dtype=np.uint8)
        rc = self.lib.isd_get_case_codes(out.ctypes.data, out.size)
        if rc:
            raise RuntimeError("isd_get_case_codes failed (%d): pass \"gpu\":{\"debug_cases\":1}" % rc)
        return out

    def projected_centroids(self):
        n = self.get_f_size()
        c = np.empty((n, 3), dtype=np.float32)
        nr = np.empty((n, 3), dtype=np.float32)
        if self.lib.isd_get_projected_centroids(c.ctypes.data, c.size) or self.lib.isd_get_centroid_normals(nr.ctypes.data, nr.size):
            raise RuntimeError("no projection in the last build")
        return c, nr

    def debug_adjacency(self, nv, nf, fof=True):
        """(umbrella offsets, umbrella items, faces-of-faces or None) the last build's OB02 stages used."""
        off = np.empty(nv + 1, dtype=np.int32)
        items = np.empty(nf * 3, dtype=np.int32)
        ff = np.empty(nf * 3, dtype=np.int32) if fof else None
        rc = self.lib.isd_debug_adjacency(off.ctypes.data, items.ctypes.data, ff.ctypes.data if fof else None)
        if rc:
            raise RuntimeError("no adjacency tables from the last build (%d)" % rc)
        return off, items, ff

    def average_edge(self):
        return float(self.lib.isd_average_edge())

    def ob02_on_mesh(self, shape_json, verts, faces, resample_iters=0, c=1.0, projection=0, qem=0):
        v = np.ascontiguousarray(verts, dtype=np.float32).reshape(-1, 3).copy()
        f = np.ascontiguousarray(faces, dtype=np.int32).reshape(-1, 3)
        self.lib.isd_ob02_on_mesh(_b(shape_json), v.ctypes.data, v.shape[0], f.ctypes.data, f.shape[0], int(resample_iters), float(c),
                                  int(projection), int(qem))
        return v

    def faces_of_faces(self, faces, nv):
        f = np.ascontiguousarray(faces, dtype=np.int32).reshape(-1, 3)
        out = np.empty_like(f)
        self.lib.isd_faces_of_faces(f.ctypes.data, f.shape[0], int(nv), out.ctypes.data)
        return out

    def umbrella(self, faces, nv):
        f = np.ascontiguousarray(faces, dtype=np.int32).reshape(-1, 3)
        off = np.empty(nv + 1, dtype=np.int32)
        items = np.empty(f.size, dtype=np.int32)
        self.lib.isd_umbrella(f.ctypes.data, f.shape[0], int(nv), off.ctypes.data, items.ctypes.data)
        return off, items

    def eval_points(self, shape_json, points, gradients=False, ignore_root_matrix=False):
        pts = np.ascontiguousarray(points, dtype=np.float32).reshape(-1, 3)
        f = np.empty(pts.shape[0], dtype=np.float32)
        g = np.empty((pts.shape[0], 3), dtype=np.float32) if gradients else None
        self.lib.isd_eval_points(_b(shape_json), int(ignore_root_matrix), pts.ctypes.data, pts.shape[0], f.ctypes.data,
                                 g.ctypes.data if gradients else None)
        return (f, g) if gradients else f

    def set_sdf_grid(self, values, h, origin):
        v = np.ascontiguousarray(values, dtype=np.float32)        # [z][y][x]
        nz, ny, nx = v.shape
        self.lib.isd_set_sdf_grid(nx, ny, nz, float(h), float(origin[0]), float(origin[1]), float(origin[2]), v.ctypes.data)

    def last_slab(self):
        buf = np.zeros(8, dtype=np.int64)
        self.lib.isd_dist_last_slab(buf.ctypes.data)
        keys = ["z_begin", "z_end", "voff", "foff", "nv", "nf", "global_nv", "global_nf"]
        return dict(zip(keys, [int(x) for x in buf]))


_loaded = None


def load_library():
    """The process-wide library (the C API is a singleton, mcc2.cpp:193)."""
    global _loaded
    if _loaded is None:
        _loaded = GeometryLibrary(_LIB)
    return _loaded
