"""CPU: the C-ABI library loads without a GPU, exports every symbol include/implisolid_b200.h declares, fails loudly
(no CPU fallback) when asked to compute without a device, and keeps the reference's settings/abort behaviour."""
import ctypes
import os
import re
import subprocess
import sys

import pytest

import implisolid_b200
from implisolid_b200 import api, shapes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "implisolid_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"typedef[^;]*;", "", src)          # function-pointer typedefs are not exports
    return re.findall(r"\b([a-z_][a-z0-9_]*)\s*\([^;{]*\)\s*;", src)


def test_every_declared_symbol_is_exported_and_bound():
    if not os.path.exists(implisolid_b200.library_path()):
        implisolid_b200.build_library()
    lib = ctypes.CDLL(implisolid_b200.library_path())
    names = _declared_symbols()
    assert len(names) >= 45
    for n in names:
        assert hasattr(lib, n), "missing export: " + n
    bound = set(api.SIGNATURES) | set(api.EXT_SIGNATURES)
    assert set(names) == bound, (set(names) ^ bound)
    # the reference's extern "C" block (mcc2.cpp:88-134), symbol for symbol
    for n in ["build_geometry", "build_geometry_u", "get_v_size", "get_f_size", "get_f", "get_v", "finish_geometry", "get_f_ptr",
              "get_v_ptr", "set_object", "unset_object", "set_x", "unset_x", "calculate_implicit_values", "get_values_ptr",
              "get_values_size", "calculate_implicit_gradients", "get_gradients_ptr", "get_gradients_size", "get_pointset_ptr",
              "get_pointset_size", "about"]:
        assert n in api.SIGNATURES


def test_getters_before_any_build():
    lib = implisolid_b200.load_library()
    assert lib.get_v_size() == 0 and lib.get_f_size() == 0       # mcc2.cpp:466-473 before any build
    assert lib.lib.get_pointset_ptr(b"x") is None and lib.lib.get_pointset_size(b"x") == 0


def _run(code):
    return subprocess.run([sys.executable, "-c", code], cwd=ROOT, capture_output=True, text=True, timeout=300)


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = _run("import implisolid_b200 as m; from implisolid_b200 import shapes as s;"
             "m.load_library().build_geometry(s.c1_sphere(), s.mc_settings(8, 0.55))")
    assert r.returncode != 0 and "no CPU path" in r.stderr


@pytest.mark.parametrize("mc,needle", [
    ('{"resolution": 8}', "missing or incorrect values"),                                                   # no box
    ('{"resolution": 8.5, "box": {"xmin":-1,"xmax":1,"ymin":-1,"ymax":1,"zmin":-1,"zmax":1}}', "must be integer"),
    ('{"resolution": 2, "box": {"xmin":-1,"xmax":1,"ymin":-1,"ymax":1,"zmin":-1,"zmax":1}}', "missing or incorrect values"),
    ('{"resolution": 8, "qem": {"enabled": 1}, "box": {"xmin":-1,"xmax":1,"ymin":-1,"ymax":1,"zmin":-1,"zmax":1}}', "QEM will not be applied"),
    ('{"resolution": 8, "qem": {"enable": 1}, "box": {"xmin":-1,"xmax":1,"ymin":-1,"ymax":1,"zmin":-1,"zmax":1}}', "instead of using"),
])
def test_bad_settings_abort_like_the_reference(mc, needle):
    """polygoniser_settings.hpp:164-301: message on stderr, then abort()."""
    for loader in ("import implisolid_b200 as m; L = m.load_library()", "from tests.oracle_lib import load_oracle; L = load_oracle()"):
        r = _run("%s\nfrom implisolid_b200 import shapes as s\nL.build_geometry(s.c1_sphere(), %r)" % (loader, mc))
        assert r.returncode != 0
        assert needle in (r.stderr + r.stdout), (loader, r.stderr[-400:])


def test_unknown_shape_aborts():
    import torch
    if torch.cuda.is_available():
        pytest.skip("shape is compiled after device init")
    r = _run("from tests.oracle_lib import load_oracle; L = load_oracle()\nfrom implisolid_b200 import shapes as s\n"
             "L.build_geometry('{\"type\":\"teapot\",\"matrix\":[1,0,0,0,0,1,0,0,0,0,1,0]}', s.mc_settings(8, 0.5))")
    assert r.returncode != 0 and "Invalid object" in r.stderr


def test_header_is_plain_c_and_cpp():
    """include/implisolid_b200.h is the drop-in boundary: it must compile on its own as C99 and as C++11 (no torch or CUDA
    types in any signature)."""
    import subprocess
    hdr = os.path.join(ROOT, "include", "implisolid_b200.h")
    subprocess.run(["gcc", "-fsyntax-only", "-x", "c", "-std=c99", "-Wall", "-Werror", hdr], check=True)
    subprocess.run(["g++", "-fsyntax-only", "-x", "c++", "-std=c++11", "-Wall", "-Werror", hdr], check=True)
    import re
    code = re.sub(r"/\*.*?\*/", "", open(hdr).read(), flags=re.S)        # comments may mention the launcher's torch.distributed
    assert "torch" not in code and "#include <cuda" not in code


def test_example_caller_compiles_and_links(tmp_path):
    """examples/polygonize_to_obj.c — plain C99 against include/ and the built library (link only: running needs a GPU)."""
    import subprocess
    from implisolid_b200 import build_library
    lib_dir = os.path.dirname(build_library())
    out = str(tmp_path / "polygonize_to_obj")
    subprocess.run(["gcc", "-std=c99", "-O2", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "examples", "polygonize_to_obj.c"), "-L" + lib_dir, "-limplisolid_b200",
                    "-Wl,-rpath," + lib_dir, "-o", out], check=True)
    assert os.path.exists(out)
