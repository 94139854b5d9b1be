"""GPU: isd_umbrella_noalloc (no cudaMalloc: scratch in a fixed device array inside the library) against isd_umbrella, through the
C ABI, on calls in a row that share the scratch array.  Collected last on purpose."""
import ctypes

import numpy as np
import pytest

from implisolid_b200 import shapes

pytestmark = pytest.mark.gpu


def _device_umbrella(L, torch, faces, nv, stream=None):
    f = torch.from_numpy(np.ascontiguousarray(faces, dtype=np.int32)).cuda()
    off = torch.full((nv + 1,), -7, dtype=torch.int32, device="cuda")
    items = torch.full((max(faces.size, 1),), -7, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    rc = L.isd_umbrella_noalloc(ctypes.c_void_p(f.data_ptr()), int(faces.shape[0]), int(nv), ctypes.c_void_p(off.data_ptr()),
                                ctypes.c_void_p(items.data_ptr()), ctypes.c_void_p(stream) if stream else None)
    torch.cuda.synchronize()      # the entry point only enqueues
    return rc, off.cpu().numpy(), items.cpu().numpy()[:faces.size]


def test_umbrella_noalloc_matches_umbrella_on_calls_in_a_row(gpu_lib):
    import torch
    L = gpu_lib.lib
    assert L.isd_init(0) == 0
    meshes = []
    for sj, R, half in ((shapes.c4_deep_chain(), 72, 1.0), (shapes.c2_cube_minus_cylinder(), 40, 0.8), (shapes.c1_sphere(), 24, 0.55)):
        v, f = gpu_lib.polygonize(sj, shapes.mc_settings(R, half))
        assert 0 < v.shape[0] <= L.isd_umbrella_noalloc_max_vertices()
        meshes.append((f.copy(), int(v.shape[0])))
    expected = [gpu_lib.umbrella(f, nv) for f, nv in meshes]
    # calls in a row share the scratch array, which keeps the previous call's contents: a larger mesh, smaller ones, the larger again
    side = torch.cuda.Stream()
    for k in (0, 1, 2, 0, 0):
        f, nv = meshes[k]
        rc, off, items = _device_umbrella(L, torch, f, nv, stream=side.cuda_stream if k == 1 else None)
        assert rc == 0
        assert np.array_equal(off, expected[k][0])          # integer work: bit-exact, as in tests/test_gpu_ob02.py
        assert np.array_equal(items, expected[k][1])
    # a mesh without faces: all offsets zero
    rc, off, _ = _device_umbrella(L, torch, np.zeros((0, 3), np.int32), 5)
    assert rc == 0 and (off == 0).all()


def test_umbrella_noalloc_rejects_what_does_not_fit(gpu_lib):
    import torch
    L = gpu_lib.lib
    assert L.isd_init(0) == 0
    limit = L.isd_umbrella_noalloc_max_vertices()
    assert limit == 524288
    f = torch.zeros(3, dtype=torch.int32, device="cuda")
    out = torch.zeros(8, dtype=torch.int32, device="cuda")
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    assert L.isd_umbrella_noalloc(p(f), 1, limit + 1, p(out), p(out), None) == 1      # beyond the fixed scratch: refused, nothing launched
    assert L.isd_umbrella_noalloc(p(f), 1, 0, p(out), p(out), None) == 1
    assert L.isd_umbrella_noalloc(None, 1, 4, p(out), p(out), None) == 1
    torch.cuda.synchronize()
    assert (out == 0).all()
