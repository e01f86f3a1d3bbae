"""ctypes binding of the CPU oracle (oracle/librt2oracle.so).  TEST INFRASTRUCTURE: import only
from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "librt2oracle.so")

ORC_BRUTE_FORCE = 1
ORC_TRIG_LIBM = 2

u32, u64, f32 = C.c_uint32, C.c_uint64, C.c_float
VP = C.c_void_p
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} missing: run `make -C oracle`")
        L = C.CDLL(LIB_PATH)
        L.orc_scene_create.restype = VP; L.orc_scene_create.argtypes = [VP]
        L.orc_scene_destroy.restype = None; L.orc_scene_destroy.argtypes = [VP]
        L.orc_scene_num_triangles.restype = u64; L.orc_scene_num_triangles.argtypes = [VP]
        L.orc_trace_primary.restype = C.c_int
        L.orc_trace_primary.argtypes = [VP, VP, u32, u32, u32, C.c_int, VP]
        L.orc_render_reference.restype = C.c_int
        L.orc_render_reference.argtypes = [VP, VP, u32, u32, u32, u32, u32, u32, u32, u32, C.c_int, VP, VP, VP, C.POINTER(u64)]
        L.orc_zcurve_to_linear.restype = u32; L.orc_zcurve_to_linear.argtypes = [u32, u32]
        L.orc_linear_to_zcurve.restype = None; L.orc_linear_to_zcurve.argtypes = [u32, C.POINTER(u32), C.POINTER(u32)]
        L.orc_jenkins_hash.restype = u32; L.orc_jenkins_hash.argtypes = [u32]
        L.orc_rng_stream.restype = None; L.orc_rng_stream.argtypes = [u32, u32, u32, u32, u32, VP, VP]
        L.orc_f32_to_f16.restype = C.c_uint16; L.orc_f32_to_f16.argtypes = [f32]
        L.orc_f16_to_f32.restype = f32; L.orc_f16_to_f32.argtypes = [C.c_uint16]
        L.orc_sincos.restype = None; L.orc_sincos.argtypes = [f32, u32, C.POINTER(f32), C.POINTER(f32)]
        L.orc_ray_triangle.restype = C.c_int
        L.orc_ray_triangle.argtypes = [VP, VP, f32, f32, VP, VP, VP, C.POINTER(f32), C.POINTER(f32), C.POINTER(f32)]
        L.orc_primary_ray.restype = None; L.orc_primary_ray.argtypes = [VP, u32, u32, VP, VP]
        L.orc_cos_hemisphere_dir.restype = None; L.orc_cos_hemisphere_dir.argtypes = [VP, f32, f32, u32, VP]
        L.orc_sample_texture.restype = None; L.orc_sample_texture.argtypes = [VP, u32, f32, f32, VP]
        _lib = L
    return _lib


class OracleScene:
    def __init__(self, scene_desc_ptr):
        self._L = lib()
        self._p = self._L.orc_scene_create(C.cast(scene_desc_ptr, VP))
        if not self._p:
            raise RuntimeError("orc_scene_create failed")

    def close(self):
        if self._p:
            self._L.orc_scene_destroy(self._p)
            self._p = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def trace_primary(self, gconst, width, height, flags=0, threads=0):
        out = np.zeros((height, width, 8), dtype=np.uint32)
        rc = self._L.orc_trace_primary(self._p, C.cast(C.byref(gconst), VP), width, height, flags, threads, out.ctypes.data_as(VP))
        assert rc == 0
        return out

    def render_reference(self, gconst, frame, width, height, spp=0, max_bounces=0, row0=0, row1=0, flags=0, threads=0,
                         diffuse_f16=None, specular_f16=None, diffuse_f32=None):
        d16 = np.zeros((height, width, 4), np.uint16) if diffuse_f16 is None else diffuse_f16
        s16 = np.zeros((height, width, 4), np.uint16) if specular_f16 is None else specular_f16
        d32 = np.zeros((height, width, 4), np.float32) if diffuse_f32 is None else diffuse_f32
        rays = u64(0)
        rc = self._L.orc_render_reference(self._p, C.cast(C.byref(gconst), VP), frame, width, height, spp, max_bounces,
                                          row0, row1, flags, threads, d16.ctypes.data_as(VP), s16.ctypes.data_as(VP),
                                          d32.ctypes.data_as(VP), C.byref(rays))
        assert rc == 0
        return d16, s16, d32, rays.value
