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
ORC_PRESAMPLING = 4   # sampler as compiled with RTXDI_ENABLE_PRESAMPLING 1 (POWER_RIS tiles of K11)

u32, u64, f32 = C.c_uint32, C.c_uint64, C.c_float
VP = C.c_void_p
_lib = None


class OracleError(AssertionError):
    """a negative return code of an orc_* call (args[0]); an AssertionError, as the wrappers' asserts raise"""


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} missing: run `make -C oracle`")
        L = C.CDLL(LIB_PATH)
        L.orc_scene_create.restype = VP; L.orc_scene_create.argtypes = [VP]
        L.orc_scene_destroy.restype = None; L.orc_scene_destroy.argtypes = [VP]
        L.orc_scene_num_triangles.restype = u64; L.orc_scene_num_triangles.argtypes = [VP]
        L.orc_trace_rays.restype = C.c_int
        L.orc_trace_rays.argtypes = [VP, VP, u32, u32, VP]
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
        L.orc_prepare_lights.restype = C.c_int
        L.orc_prepare_lights.argtypes = [VP, VP, C.POINTER(u32), VP, VP, u32, C.POINTER(u32), VP, u32, u32]
        L.orc_generate_mips.restype = C.c_int; L.orc_generate_mips.argtypes = [VP, C.c_int, VP, u32, u32, u32]
        L.orc_postprocess.restype = C.c_int; L.orc_postprocess.argtypes = [VP, VP, u32, u32] + [VP] * 8
        L.orc_pack_r11g11b10.restype = u32; L.orc_pack_r11g11b10.argtypes = [VP]
        L.orc_unpack_r11g11b10.restype = None; L.orc_unpack_r11g11b10.argtypes = [u32, VP]
        L.orc_pack_rgba8_gamma.restype = u32; L.orc_pack_rgba8_gamma.argtypes = [VP]
        L.orc_unpack_rgba8_gamma.restype = None; L.orc_unpack_rgba8_gamma.argtypes = [u32, VP]
        L.orc_ndir_to_oct_unorm32.restype = u32; L.orc_ndir_to_oct_unorm32.argtypes = [VP]
        L.orc_oct_unorm32_to_ndir.restype = None; L.orc_oct_unorm32_to_ndir.argtypes = [u32, VP]
        L.orc_store_triangle_light.restype = None; L.orc_store_triangle_light.argtypes = [VP, VP, VP, VP, VP, C.POINTER(f32)]
        L.orc_create_triangle_light.restype = None; L.orc_create_triangle_light.argtypes = [VP] * 6 + [C.POINTER(f32)]
        L.orc_tonemap.restype = u32; L.orc_tonemap.argtypes = [VP]
        L.orc_gbuffer.restype = C.c_int; L.orc_gbuffer.argtypes = [VP, VP, u32, u32, u32, C.c_int] + [VP] * 7 + [C.POINTER(u64)]
        L.orc_di_fused_resampling.restype = C.c_int
        L.orc_di_fused_resampling.argtypes = [VP, VP, u32, u32, u32, u32, C.c_int] + [VP] * 5 + [VP, u32, VP, VP, u32, u32, u32, VP, VP, VP, C.POINTER(u64), VP, VP, u64, VP, u32, u32, u32]
        L.orc_reservoir_position_to_pointer.restype = u32; L.orc_reservoir_position_to_pointer.argtypes = [u32] * 5
        L.orc_brdf_rays.restype = C.c_int; L.orc_brdf_rays.argtypes = [VP, VP, u32, u32, u32, u32, C.c_int, VP, VP, VP, VP, C.POINTER(u64)]
        L.orc_shade_secondary_surfaces.restype = C.c_int
        L.orc_shade_secondary_surfaces.argtypes = [VP, VP, u32, u32, u32, u32, C.c_int, VP, VP, VP, VP, VP, VP, C.POINTER(u64)]
        L.orc_gi_final_shading.restype = C.c_int; L.orc_gi_final_shading.argtypes = [VP, VP, u32, u32, u32, C.c_int, VP, VP, VP, VP, VP, C.POINTER(u64)]
        L.orc_gi_temporal_resampling.restype = C.c_int
        L.orc_gi_temporal_resampling.argtypes = [VP, VP, u32, u32, u32, u32, C.c_int, VP, VP, VP, VP, C.POINTER(u64)]
        L.orc_gi_spatial_resampling.restype = C.c_int
        L.orc_gi_spatial_resampling.argtypes = [VP, VP, u32, u32, u32, u32, C.c_int, VP, VP, u32, VP, C.POINTER(u64)]
        L.orc_fill_neighbor_offset_buffer.restype = None; L.orc_fill_neighbor_offset_buffer.argtypes = [u32, VP]
        L.orc_presample.restype = C.c_int; L.orc_presample.argtypes = [VP, u32, C.c_int, VP, u32, u32, u32, VP, u32, VP, VP]
        L.orc_encode_logluv.restype = u32; L.orc_encode_logluv.argtypes = [VP]
        L.orc_decode_logluv.restype = None; L.orc_decode_logluv.argtypes = [u32, VP]
        L.orc_encode_snorm_oct.restype = u32; L.orc_encode_snorm_oct.argtypes = [VP]
        L.orc_decode_snorm_oct.restype = None; L.orc_decode_snorm_oct.argtypes = [u32, VP]
        _lib = L
    return _lib


def _f(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def mip_dims(w, h, mips):
    return [(max(1, w >> k), max(1, h >> k)) for k in range(mips)]


def split_mips(packed, w, h, mips):
    out, off = [], 0
    for mw, mh in mip_dims(w, h, mips):
        out.append(packed[off:off + mw * mh].reshape(mh, mw)); off += mw * mh
    return out


class _SceneHead(C.Structure):   # leading fields of rt2_scene_desc (include/rt2_abi.h)
    _fields_ = [("vertices", VP), ("num_vertices", u64), ("indices", VP), ("num_indices", u64),
                ("geometries", VP), ("index_counts", VP), ("num_geometries", u32)]


class GBufferIn(C.Structure):   # orc_gbuffer_in
    _fields_ = [(k, VP) for k in ("depth", "normals", "geo_normals", "diffuse_albedo", "specular_rough")]


class LightIn(C.Structure):     # orc_light_in
    _fields_ = [("lights", VP), ("num_lights", u32), ("geom_to_light", VP), ("pdf_packed", VP), ("pdf_w", u32), ("pdf_h", u32), ("pdf_mips", u32),
                ("ris", VP), ("ris_light_data", VP), ("ris_entries", u64), ("env_pdf_packed", VP), ("env_w", u32), ("env_h", u32), ("env_mips", u32)]


def presample(gconst, frame, which, pdf_packed, w, h, mips, lights=None, ris=None, ris_light_data=None, entries=2 * 128 * 1024):
    """K11.  Returns (ris uint32 [entries, 2], ris_light_data uint32 [entries, 8]); pass the arrays back in to fill the other segment."""
    ris = np.zeros((entries, 2), np.uint32) if ris is None else ris
    data = np.zeros((entries, 8), np.uint32) if ris_light_data is None else ris_light_data
    pdf = np.ascontiguousarray(pdf_packed)
    li = np.ascontiguousarray(lights) if lights is not None else None
    rc = lib().orc_presample(C.cast(C.byref(gconst), VP), frame, which, pdf.ctypes.data_as(VP), w, h, mips,
                             li.ctypes.data_as(VP) if li is not None else None, len(li) if li is not None else 0,
                             ris.ctypes.data_as(VP), data.ctypes.data_as(VP))
    assert rc == 0, rc
    return ris, data


def fill_neighbor_offset_buffer(count=8192):
    out = np.zeros((count, 2), np.int8)
    lib().orc_fill_neighbor_offset_buffer(count, out.ctypes.data_as(VP))
    return out


class OracleScene:
    def __init__(self, scene_desc_ptr):
        self._L = lib()
        self._p = self._L.orc_scene_create(C.cast(scene_desc_ptr, VP))
        if not self._p:
            raise RuntimeError("orc_scene_create failed")
        self.num_geometries = int(C.cast(scene_desc_ptr, C.POINTER(_SceneHead)).contents.num_geometries)

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

    def trace_rays(self, rays, flags=0):
        """rays [n, 8] float32 = (o.xyz, tmin, d.xyz, tmax) -> hit records [n, 8] uint32 views"""
        r = np.ascontiguousarray(rays, np.float32)
        out = np.zeros((r.shape[0], 8), np.uint32)
        assert self._L.orc_trace_rays(self._p, r.ctypes.data_as(VP), r.shape[0], flags, out.ctypes.data_as(VP)) == 0
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

    # ---- K10 / K12 / K13 ---------------------------------------------------------------------------------
    def prepare_lights(self, pdf_w=0, pdf_h=0, env=False):
        """returns (tasks[n,4] u32, geom_to_light u32, lights[n,12] u32, pdf_mip0 f32 or None).  env: size the light buffer for
        lights + 2 records, so that the environment light lands at index lights + 1 (scenes with a sky; SURVEY Q7 pinned to intent)."""
        L = self._L
        nt, nl = u32(0), u32(0)
        assert L.orc_prepare_lights(self._p, None, C.byref(nt), None, None, 0, C.byref(nl), None, 0, 0) == 0
        ng = self.num_geometries
        tasks = np.zeros((max(ng, 1), 4), np.uint32); g2l = np.zeros(max(ng, 1), np.uint32)
        lights = np.zeros((max(nl.value, 1) + (2 if env else 0), 12), np.uint32)
        if env:
            lights = lights[:nl.value + 2]
        pdf = np.zeros((pdf_h, pdf_w), np.float32) if pdf_w and pdf_h else None
        rc = L.orc_prepare_lights(self._p, tasks.ctypes.data_as(VP), C.byref(nt), g2l.ctypes.data_as(VP), lights.ctypes.data_as(VP),
                                  lights.shape[0], C.byref(nl), pdf.ctypes.data_as(VP) if pdf is not None else None, pdf_w, pdf_h)
        assert rc == 0
        return tasks[:nt.value], g2l[:ng], lights[:nl.value + (2 if env else 0)], pdf

    def generate_mips(self, which, packed, w, h, mips):
        assert packed.flags["C_CONTIGUOUS"]
        rc = self._L.orc_generate_mips(self._p, which, packed.ctypes.data_as(VP), w, h, mips)
        assert rc == 0, rc
        return packed

    def postprocess(self, gconst, w, h, diffuse_f16, specular_f16=None, gb_depth=None, gb_albedo=None, gb_specrough=None, gb_emissive=None):
        out = np.zeros((h, w, 4), np.uint8); mv = np.zeros((h, w, 4), np.float32)
        p = lambda a: a.ctypes.data_as(VP) if a is not None else None
        rc = self._L.orc_postprocess(self._p, C.cast(C.byref(gconst), VP), w, h, p(diffuse_f16), p(specular_f16), p(gb_depth), p(gb_albedo),
                                     p(gb_specrough), p(gb_emissive), p(mv), p(out))
        assert rc == 0, rc
        return out, mv

    # ---- K3 / K9 ---------------------------------------------------------------------------------------------
    def gbuffer(self, gconst, w, h, flags=0, threads=0):
        gb = {"depth": np.zeros((h, w), np.float32), "normals": np.zeros((h, w), np.uint32), "geo_normals": np.zeros((h, w), np.uint32),
              "diffuse_albedo": np.zeros((h, w), np.uint32), "specular_rough": np.zeros((h, w), np.uint32),
              "emissive": np.zeros((h, w, 4), np.uint16), "motion": np.zeros((h, w, 4), np.float32)}
        rays = u64(0)
        rc = self._L.orc_gbuffer(self._p, C.cast(C.byref(gconst), VP), w, h, flags, threads, *[gb[k].ctypes.data_as(VP) for k in
                                 ("depth", "normals", "geo_normals", "diffuse_albedo", "specular_rough", "emissive", "motion")], C.byref(rays))
        assert rc == 0, rc
        return gb

    def di_fused_resampling(self, gconst, frame, w, h, gb, lights, geom_to_light, pdf_packed, pdf_w, pdf_h, pdf_mips, array_pitch,
                            diffuse_f16=None, specular_f16=None, flags=0, threads=0, ris=None, ris_light_data=None, env_pdf=None, env_dims=(0, 0, 0)):
        """K9.  ris / ris_light_data: K11's outputs (presample()), read with flags & ORC_PRESAMPLING and local_light_sampling_mode 1.
        env_pdf / env_dims = (w, h, mips): K12's r16f environment pyramid (uint16, packed), read with ORC_PRESAMPLING + an environment light."""
        res = np.zeros((array_pitch * 2, 6), np.uint32)
        d16 = np.zeros((h, w, 4), np.uint16) if diffuse_f16 is None else diffuse_f16
        s16 = np.zeros((h, w, 4), np.uint16) if specular_f16 is None else specular_f16
        rays = u64(0)
        lights = np.ascontiguousarray(lights); geom_to_light = np.ascontiguousarray(geom_to_light); pdf_packed = np.ascontiguousarray(pdf_packed, np.float32)
        rc = self._L.orc_di_fused_resampling(self._p, C.cast(C.byref(gconst), VP), frame, w, h, flags, threads,
                                             *[np.ascontiguousarray(gb[k]).ctypes.data_as(VP) for k in ("depth", "normals", "geo_normals", "diffuse_albedo", "specular_rough")],
                                             lights.ctypes.data_as(VP), len(lights), geom_to_light.ctypes.data_as(VP), pdf_packed.ctypes.data_as(VP), pdf_w, pdf_h, pdf_mips,
                                             res.ctypes.data_as(VP), d16.ctypes.data_as(VP), s16.ctypes.data_as(VP), C.byref(rays),
                                             ris.ctypes.data_as(VP) if ris is not None else None,
                                             ris_light_data.ctypes.data_as(VP) if ris_light_data is not None else None, u64(len(ris) if ris is not None else 0),
                                             env_pdf.ctypes.data_as(VP) if env_pdf is not None else None, *env_dims)
        if rc != 0:
            raise OracleError(rc)
        return res, d16, s16, rays.value

    # ---- the ReSTIR-GI chain K4 -> K5 -> K8 ----------------------------------------------------------------
    @staticmethod
    def _gb_in(gb):
        keep = [np.ascontiguousarray(gb[k]) for k in ("depth", "normals", "geo_normals", "diffuse_albedo", "specular_rough")]
        return GBufferIn(*[a.ctypes.data_as(VP) for a in keep]), keep

    @staticmethod
    def _light_in(lights, geom_to_light, pdf_packed, pdf_w, pdf_h, pdf_mips, ris=None, ris_light_data=None, env_pdf=None, env_dims=(0, 0, 0)):
        if lights is None:
            return None, []
        keep = [np.ascontiguousarray(lights), np.ascontiguousarray(geom_to_light), np.ascontiguousarray(pdf_packed, np.float32), ris, ris_light_data, env_pdf]
        return LightIn(keep[0].ctypes.data_as(VP), len(keep[0]), keep[1].ctypes.data_as(VP), keep[2].ctypes.data_as(VP), pdf_w, pdf_h, pdf_mips,
                       ris.ctypes.data_as(VP) if ris is not None else None, ris_light_data.ctypes.data_as(VP) if ris_light_data is not None else None,
                       len(ris) if ris is not None else 0, env_pdf.ctypes.data_as(VP) if env_pdf is not None else None, *env_dims), keep

    def brdf_rays(self, gconst, frame, w, h, gb, secondary, diffuse_f16, specular_f16, flags=0, threads=0):
        """K4.  secondary: uint32 [arrayPitch, 12] (in/out); lighting images rgba16f as uint16 [h, w, 4] (in/out).  Returns rays."""
        gin, keep = self._gb_in(gb)
        rays = u64(0)
        rc = self._L.orc_brdf_rays(self._p, C.cast(C.byref(gconst), VP), frame, w, h, flags, threads, C.byref(gin),
                                   secondary.ctypes.data_as(VP), diffuse_f16.ctypes.data_as(VP), specular_f16.ctypes.data_as(VP), C.byref(rays))
        assert rc == 0, rc
        return rays.value

    def shade_secondary_surfaces(self, gconst, frame, w, h, gb, secondary, gi_reservoirs, diffuse_f16, specular_f16,
                                 lights=None, geom_to_light=None, pdf_packed=None, pdf_dims=(0, 0, 0), flags=0, threads=0, ris=None, ris_light_data=None,
                                 env_pdf=None, env_dims=(0, 0, 0)):
        """K5.  gi_reservoirs: uint32 [arrayPitch * 2, 8] (in/out)."""
        gin, keep = self._gb_in(gb)
        lin, keep2 = self._light_in(lights, geom_to_light, pdf_packed, *pdf_dims, ris=ris, ris_light_data=ris_light_data, env_pdf=env_pdf, env_dims=env_dims)
        rays = u64(0)
        rc = self._L.orc_shade_secondary_surfaces(self._p, C.cast(C.byref(gconst), VP), frame, w, h, flags, threads, C.byref(gin),
                                                  C.byref(lin) if lin is not None else None, secondary.ctypes.data_as(VP), gi_reservoirs.ctypes.data_as(VP),
                                                  diffuse_f16.ctypes.data_as(VP), specular_f16.ctypes.data_as(VP), C.byref(rays))
        assert rc == 0, rc
        return rays.value

    def gi_final_shading(self, gconst, w, h, gb, secondary, gi_reservoirs, diffuse_f16, specular_f16, flags=0, threads=0):
        """K8."""
        gin, keep = self._gb_in(gb)
        rays = u64(0)
        rc = self._L.orc_gi_final_shading(self._p, C.cast(C.byref(gconst), VP), w, h, flags, threads, C.byref(gin),
                                          secondary.ctypes.data_as(VP), gi_reservoirs.ctypes.data_as(VP),
                                          diffuse_f16.ctypes.data_as(VP), specular_f16.ctypes.data_as(VP), C.byref(rays))
        assert rc == 0, rc
        return rays.value

    def gi_temporal_resampling(self, gconst, frame, w, h, gb, gb_prev, motion, gi_reservoirs, flags=0, threads=0):
        """K6.  motion: float32 [h, w, 4]; gi_reservoirs in/out."""
        gin, keep = self._gb_in(gb)
        gpv, keep2 = self._gb_in(gb_prev)
        mv = np.ascontiguousarray(motion, np.float32)
        rays = u64(0)
        rc = self._L.orc_gi_temporal_resampling(self._p, C.cast(C.byref(gconst), VP), frame, w, h, flags, threads, C.byref(gin), C.byref(gpv),
                                                mv.ctypes.data_as(VP), gi_reservoirs.ctypes.data_as(VP), C.byref(rays))
        assert rc == 0, rc
        return rays.value

    def gi_spatial_resampling(self, gconst, frame, w, h, gb, neighbor_offsets, gi_reservoirs, flags=0, threads=0):
        """K7.  neighbor_offsets: int8 [count, 2]."""
        gin, keep = self._gb_in(gb)
        no = np.ascontiguousarray(neighbor_offsets, np.int8)
        rays = u64(0)
        rc = self._L.orc_gi_spatial_resampling(self._p, C.cast(C.byref(gconst), VP), frame, w, h, flags, threads, C.byref(gin),
                                               no.ctypes.data_as(VP), no.size // 2, gi_reservoirs.ctypes.data_as(VP), C.byref(rays))
        assert rc == 0, rc
        return rays.value

    def gi_chain(self, gconst, frame, w, h, gb, array_pitch, light_inputs=None, state=None, flags=0, threads=0, gb_prev=None, motion=None, neighbor_offsets=None):
        """K4 -> K5 -> [K6] -> [K7] -> K8 on fresh (or carried-over `state`) buffers, as LightPasses::execute orders them
        (src/light_passes.rs:621-659).  Returns dict(secondary, gi_reservoirs, diffuse, specular, rays)."""
        st = state or {"secondary": np.zeros((array_pitch, 12), np.uint32), "gi_reservoirs": np.zeros((array_pitch * 2, 8), np.uint32),
                       "diffuse": np.zeros((h, w, 4), np.uint16), "specular": np.zeros((h, w, 4), np.uint16)}
        li = light_inputs or {}
        r4 = self.brdf_rays(gconst, frame, w, h, gb, st["secondary"], st["diffuse"], st["specular"], flags, threads)
        r5 = self.shade_secondary_surfaces(gconst, frame, w, h, gb, st["secondary"], st["gi_reservoirs"], st["diffuse"], st["specular"],
                                           li.get("lights"), li.get("geom_to_light"), li.get("pdf_packed"), li.get("pdf_dims", (0, 0, 0)), flags, threads,
                                           li.get("ris"), li.get("ris_light_data"), li.get("env_pdf"), li.get("env_dims", (0, 0, 0)))
        r6 = r7 = 0
        if gconst.enable_temporal_resampling == 1:
            r6 = self.gi_temporal_resampling(gconst, frame, w, h, gb, gb_prev, motion, st["gi_reservoirs"], flags, threads)
        if gconst.enable_spatial_resampling == 1:
            r7 = self.gi_spatial_resampling(gconst, frame, w, h, gb, neighbor_offsets, st["gi_reservoirs"], flags, threads)
        r8 = self.gi_final_shading(gconst, w, h, gb, st["secondary"], st["gi_reservoirs"], st["diffuse"], st["specular"], flags, threads)
        st["rays"] = (r4, r5, r6, r7, r8)
        return st
