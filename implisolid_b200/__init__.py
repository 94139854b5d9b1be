"""implisolid_b200 — B200-native polygonization path of ImpliSolid behind the reference's C geometry API.

The product is `libimplisolid_b200.so` (CUDA, sm_100a; sources in `csrc/`, C ABI in `include/implisolid_b200.h`).
This package is the thin host-side mirror of the reference's JavaScript service layer
(`implisolid_main.js:38-59` cwraps the same symbols): it loads the library with ctypes and exposes the calls under
their reference names.  There is no CPU fallback; importing works without a GPU, computing does not.
"""
from .api import GeometryLibrary, load_library, library_path, build_library  # noqa: F401
from . import shapes, slabs  # noqa: F401

__all__ = ["GeometryLibrary", "load_library", "library_path", "build_library", "shapes", "slabs"]
