"""raytracer2_b200 — B200 (sm_100a) CUDA implementation of RayTracer2's path-tracing hot path.

The product is `librt2cuda.so` (C ABI in include/rt2_abi.h, sources in raytracer2_b200/csrc).
This package is the thin host-side mirror of the reference's pass objects (api.py) plus ctypes
bindings (_ffi.py).  Importing the package does not load CUDA; creating a Renderer does, and it
fails loudly if the extension is missing — there is no CPU fallback.
"""
from .api import *  # noqa: F401,F403
from . import _ffi  # noqa: F401
