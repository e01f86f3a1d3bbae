"""Host-side mirror of the reference's pass objects over the C ABI.

Names and call order follow the reference's host (src/main.rs:661-700):
    Renderer            <- Renderer::new                 (src/context.rs:120)
    Model.from_scene    <- Model::from_gltf + TLAS       (src/model.rs:185, src/main.rs:153-191)
    LightPasses         <- LightPasses::{new,update_uniform,execute}   (src/light_passes.rs:226,665,550)
    PrepareLightsTasks  <- PrepareLightsTasks::execute   (src/prepare_lights.rs:172)
    GenerateMipsPass    <- GenerateMipsPass::execute_mip_generation    (src/mip_pass.rs:122)
    PostProcessPass     <- PostProcessPass::execute      (src/postprocess.rs:225)
Every method is one call into librt2cuda.so; errors raise Rt2Error with rt2_last_error()
(the reference's constructors return anyhow::Result / panic on unwrap).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _ffi
from ._ffi import (Camera, GConst, HitRecord, LaunchParams, PlanarViewConstants, SceneDesc, Stats)

# rt2_buffer ids (include/rt2_abi.h)
BUF_DIFFUSE_F16, BUF_SPECULAR_F16, BUF_DIFFUSE_F32, BUF_HIT_IDS, BUF_OUTPUT_RGBA8 = 0, 1, 2, 3, 4
BUF_LIGHT_INFO, BUF_GEOM_TO_LIGHT, BUF_LIGHT_TASKS = 5, 6, 7
BUF_LOCAL_PDF_MIP0, BUF_ENV_PDF_MIP0 = 16, 48
BUF_GBUFFER_DEPTH, BUF_GBUFFER_NORMALS, BUF_GBUFFER_GEO_NORMALS = 80, 81, 82
BUF_GBUFFER_DIFFUSE_ALBEDO, BUF_GBUFFER_SPECULAR_ROUGH, BUF_GBUFFER_EMISSIVE = 83, 84, 85
BUF_MOTION_VECTORS, BUF_DI_RESERVOIRS, BUF_GI_RESERVOIRS, BUF_SECONDARY_GBUFFER = 86, 87, 88, 89
BUF_NEIGHBOR_OFFSETS = 90
BUF_RIS, BUF_RIS_LIGHT_DATA = 91, 92

BUF_BLAS_NODES, BUF_BLAS_TRIS, BUF_BLAS_SLOT_TO_PRIM, BUF_TLAS_NODES, BUF_TLAS_SLOT_TO_INST = 100, 101, 102, 103, 104

LAUNCH_TRIG_LIBDEVICE = 1
LAUNCH_PROFILE = 2
FILTER_NEAREST, FILTER_LINEAR = 0, 1   # rt2_sampler_desc (src/model.rs:165-183)
LAUNCH_PRESAMPLING = 4   # light sampler as compiled with RTXDI_ENABLE_PRESAMPLING 1: POWER_RIS reads K11's tiles

SCENE_CORNELL, SCENE_HALL, SCENE_MANYLIGHT, SCENE_INSTANCED, SCENE_FURNACE, SCENE_RANDOM = 1, 2, 3, 5, 6, 7


class Rt2Error(RuntimeError):
    pass


def load_exr(path: str):
    """RGBA32F numpy array [h, w, 4] (DynamicImage::to_rgba32f of image::open, src/context.rs:1476)."""
    h = _ffi.host_lib()
    w_, h_ = C.c_uint32(), C.c_uint32()
    err = C.create_string_buffer(512)
    p = h.rt2h_load_exr(path.encode(), C.byref(w_), C.byref(h_), err, len(err))
    if not p:
        raise Rt2Error(f"cannot load {path}: {err.value.decode(errors='replace')}")
    try:
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=(h_.value, w_.value, 4)).copy()
    finally:
        h.rt2h_free(p)


# ---- host library wrappers ------------------------------------------------------------------------------
class Scene:
    """A procedurally generated scene in the reference's upload layout (rt2h_scene_generate)."""

    def __init__(self, kind: int, scale: float = 1.0, texture_size: int = 0, seed: int = 0):
        self._h = _ffi.host_lib()
        self._p = self._h.rt2h_scene_generate(kind, scale, texture_size, seed)
        if not self._p:
            raise Rt2Error(f"unknown scene kind {kind}")
        self.kind = kind

    @classmethod
    def from_gltf(cls, path: str, sky_exr: str | None = None) -> "Scene":
        """gltf::load_file + image::open of the sky box (src/main.rs:62-63) through the C++ readers (rt2h_scene_load_gltf)."""
        self = cls.__new__(cls)
        self._h = _ffi.host_lib()
        err = C.create_string_buffer(512)
        self._p = self._h.rt2h_scene_load_gltf(path.encode(), sky_exr.encode() if sky_exr else None, err, len(err))
        if not self._p:
            raise Rt2Error(f"cannot load {path}: {err.value.decode(errors='replace')}")
        self.kind = 0
        return self

    @property
    def desc(self) -> "C.POINTER(SceneDesc)":
        return self._h.rt2h_scene_desc(self._p)

    def camera(self, aspect_ratio: float = 0.0) -> Camera:
        cam = Camera()
        self._h.rt2h_scene_camera(self._p, aspect_ratio, C.byref(cam))
        return cam

    def add_sky(self, width: int = 512) -> "Scene":
        """Attach the procedural sky box (width x width/2) to a scene generated without one; call before uploading."""
        self._h.rt2h_scene_add_procedural_sky(self._p, width)
        return self

    @property
    def num_lights(self) -> int:
        return self._h.rt2h_scene_num_lights(self._p)

    @property
    def num_triangles(self) -> int:
        return self._h.rt2h_scene_num_triangles(self._p)

    def close(self):
        if self._p:
            self._h.rt2h_scene_destroy(self._p)
            self._p = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def default_gconst(cam: Camera, width: int, height: int, num_lights: int = 0, env_size=(0, 0)) -> GConst:
    """src/main.rs:237-400 defaults."""
    g = GConst()
    _ffi.host_lib().rt2h_default_gconst(C.byref(g), C.byref(cam), width, height, num_lights, env_size[0], env_size[1])
    return g


def overhead_camera(scene: "Scene", aspect_ratio: float, height: float = 60.0) -> Camera:
    """A second camera for the instanced field (BASELINE config C5): above the centre looking down (the generator's own camera
    looks across the field and sees mostly sky).  The view direction is -direction (the projection keeps w = +z_view,
    src/camera.rs:99-123), and UP = (0,-1,0) forbids an exactly vertical axis, hence the slight tilt."""
    cam = Camera()
    pos, d = (C.c_float * 3)(0.0, height, 0.0), (C.c_float * 3)(0.15, 1.0, 0.1)
    _ffi.host_lib().rt2h_camera_new(C.byref(cam), pos, d, 65.0, aspect_ratio, 0.1, 1000.0)
    return cam


def advance_frame(g: GConst, cam: Camera, width: int, height: int, frame: int) -> None:
    """src/main.rs:641-651."""
    _ffi.host_lib().rt2h_advance_frame(C.byref(g), C.byref(cam), width, height, frame)


def planar_view_constants(cam: Camera, width: int, height: int) -> PlanarViewConstants:
    v = PlanarViewConstants()
    _ffi.host_lib().rt2h_planar_view_constants(C.byref(cam), width, height, C.byref(v))
    return v


# ---- the CUDA context ---------------------------------------------------------------------------------------
class Renderer:
    """Owns one rt2_ctx (one CUDA device, one stream)."""

    def __init__(self, device: int = 0):
        self._lib = _ffi.cuda_lib()
        self._ctx = C.c_void_p()
        rc = self._lib.rt2_create(device, C.byref(self._ctx))
        if rc != 0:
            msg = self._lib.rt2_last_error(None)
            raise Rt2Error(f"rt2_create({device}) failed ({rc}): {msg.decode() if msg else ''}")
        self.width = self.height = 0

    def _check(self, rc: int, what: str):
        if rc != 0:
            msg = self._lib.rt2_last_error(self._ctx)
            raise Rt2Error(f"{what} failed ({rc}): {msg.decode() if msg else ''}")

    def set_stream(self, cuda_stream_handle: int):
        self._check(self._lib.rt2_set_stream(self._ctx, C.c_void_p(cuda_stream_handle)), "rt2_set_stream")

    def resize(self, width: int, height: int):
        self._check(self._lib.rt2_resize(self._ctx, width, height), "rt2_resize")
        self.width, self.height = width, height

    def sync(self):
        self._check(self._lib.rt2_sync(self._ctx), "rt2_sync")

    def buffer_size(self, which: int) -> int:
        return int(self._lib.rt2_buffer_size(self._ctx, which))

    def read(self, which: int, dtype=np.uint8) -> np.ndarray:
        n = self.buffer_size(which)
        if n == 0:
            raise Rt2Error(f"buffer {which} is not allocated")
        out = np.empty(n, dtype=np.uint8)
        self._check(self._lib.rt2_read(self._ctx, which, out.ctypes.data_as(C.c_void_p), n), "rt2_read")
        return out.view(dtype)

    def read_into(self, which: int, dst_ptr: int, nbytes: int):
        self._check(self._lib.rt2_read(self._ctx, which, C.c_void_p(dst_ptr), nbytes), "rt2_read")

    def read_into_async(self, which: int, dst_ptr: int, nbytes: int):
        """Frame-in-flight read into (pinned) host memory; valid after sync().  See rt2_read_async."""
        self._check(self._lib.rt2_read_async(self._ctx, which, C.c_void_p(dst_ptr), nbytes), "rt2_read_async")

    def read_fence(self):
        self._check(self._lib.rt2_read_fence(self._ctx), "rt2_read_fence")

    def write(self, which: int, arr: np.ndarray):
        a = np.ascontiguousarray(arr)
        self._check(self._lib.rt2_write(self._ctx, which, a.ctypes.data_as(C.c_void_p), a.nbytes), "rt2_write")

    def copy_to_device(self, which: int, device_ptr: int, nbytes: int, row0: int = 0, row1: int = 0):
        self._check(self._lib.rt2_copy_to_device(self._ctx, which, C.c_void_p(device_ptr), nbytes, row0, row1), "rt2_copy_to_device")

    def pack_bands(self, which: int, device_ptr: int, nbytes: int, band_rows: int, band_rank: int, band_world: int):
        self._check(self._lib.rt2_pack_bands(self._ctx, which, C.c_void_p(device_ptr), nbytes, band_rows, band_rank, band_world), "rt2_pack_bands")

    def stats(self) -> Stats:
        s = Stats()
        self._check(self._lib.rt2_get_stats(self._ctx, C.byref(s)), "rt2_get_stats")
        return s

    def close(self):
        if getattr(self, "_ctx", None) and self._ctx.value:
            if not getattr(self, "_borrowed", False):      # contexts of a Group belong to the group
                self._lib.rt2_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Group:
    """rt2_group: one frame over the GPUs of one node behind the C ABI — band plan, NCCL all-gather and the un-permute into the
    assembled frame all happen inside librt2cuda.so.  Group(devices=[0, 1, ...]) drives n devices from this process;
    Group.from_rank(device, rank, world, unique_id) is the one-process-per-GPU form (unique_id from Group.unique_id() on rank 0,
    distributed by the host's own means, e.g. torch.distributed.broadcast_object_list)."""

    def __init__(self, devices=None, _handle=None):
        self._lib = _ffi.cuda_lib()
        self._g = C.c_void_p()
        if _handle is not None:
            self._g = _handle
        else:
            arr = (C.c_int * len(devices))(*devices)
            rc = self._lib.rt2_group_create(arr, len(devices), C.byref(self._g))
            if rc != 0:
                raise Rt2Error(f"rt2_group_create failed ({rc}): {(self._lib.rt2_group_last_error(None) or b'').decode()}")
        self.width = self.height = 0

    @staticmethod
    def unique_id() -> bytes:
        buf = C.create_string_buffer(128)
        lib = _ffi.cuda_lib()
        rc = lib.rt2_group_unique_id(buf)
        if rc != 0:
            raise Rt2Error(f"rt2_group_unique_id failed ({rc}): {(lib.rt2_group_last_error(None) or b'').decode()}")
        return buf.raw

    @classmethod
    def from_rank(cls, device: int, rank: int, world: int, unique_id: bytes | None) -> "Group":
        lib = _ffi.cuda_lib()
        h = C.c_void_p()
        idbuf = C.create_string_buffer(unique_id, 128) if unique_id is not None else None
        rc = lib.rt2_group_create_rank(device, rank, world, idbuf, C.byref(h))
        if rc != 0:
            raise Rt2Error(f"rt2_group_create_rank failed ({rc}): {(lib.rt2_group_last_error(None) or b'').decode()}")
        return cls(_handle=h)

    def _check(self, rc: int, what: str):
        if rc != 0:
            raise Rt2Error(f"{what} failed ({rc}): {(self._lib.rt2_group_last_error(self._g) or b'').decode()}")

    @property
    def world(self) -> int:
        return self._lib.rt2_group_world(self._g)

    @property
    def local_count(self) -> int:
        return self._lib.rt2_group_local_count(self._g)

    def ctx(self, i: int = 0) -> "Renderer":
        """A non-owning Renderer view of local context i (stats, per-context reads, set_stream)."""
        r = Renderer.__new__(Renderer)
        r._lib = self._lib
        r._ctx = C.c_void_p(self._lib.rt2_group_ctx(self._g, i))
        r._borrowed = True
        r.width, r.height = self.width, self.height
        return r

    def upload_scene(self, scene_desc):
        self._check(self._lib.rt2_group_upload_scene(self._g, scene_desc), "rt2_group_upload_scene")
        self._check(self._lib.rt2_group_build_accel(self._g), "rt2_group_build_accel")

    def resize(self, width: int, height: int):
        self._check(self._lib.rt2_group_resize(self._g, width, height), "rt2_group_resize")
        self.width, self.height = width, height

    def update_uniform(self, g_const: GConst):
        self._check(self._lib.rt2_group_set_constants(self._g, C.byref(g_const), C.sizeof(g_const)), "rt2_group_set_constants")

    def prepare_lights(self):
        self._check(self._lib.rt2_group_prepare_lights(self._g), "rt2_group_prepare_lights")

    def generate_mips(self, which: int):
        self._check(self._lib.rt2_group_generate_mips(self._g, which), "rt2_group_generate_mips")

    def execute(self, frame: int, params: LaunchParams | None = None, gather: int = BUF_DIFFUSE_F32):
        p = C.byref(params) if params is not None else None
        self._check(self._lib.rt2_group_render(self._g, frame & 0xFFFFFFFF, p, gather), "rt2_group_render")

    def postprocess(self, frame: int):
        self._check(self._lib.rt2_group_postprocess(self._g, frame & 0xFFFFFFFF), "rt2_group_postprocess")

    def gather(self, which: int):
        self._check(self._lib.rt2_group_gather(self._g, which), "rt2_group_gather")

    def read(self, dtype=np.uint8) -> np.ndarray:
        n = C.c_size_t(0)
        if not self._lib.rt2_group_assembled(self._g, 0, C.byref(n)):
            raise Rt2Error("no gathered frame")
        out = np.empty(n.value, dtype=np.uint8)
        self._check(self._lib.rt2_group_read(self._g, out.ctypes.data_as(C.c_void_p), n.value), "rt2_group_read")
        return out.view(dtype)

    def read_into_async(self, dst_ptr: int, nbytes: int):
        self._check(self._lib.rt2_group_read_async(self._g, C.c_void_p(dst_ptr), nbytes), "rt2_group_read_async")

    def assembled_ptr(self, i: int = 0):
        n = C.c_size_t(0)
        p = self._lib.rt2_group_assembled(self._g, i, C.byref(n))
        return p, n.value

    def wait_gather(self):
        self._check(self._lib.rt2_group_wait_gather(self._g), "rt2_group_wait_gather")

    def sync(self):
        self._check(self._lib.rt2_group_sync(self._g), "rt2_group_sync")

    def close(self):
        if getattr(self, "_g", None) and self._g.value:
            self._lib.rt2_group_destroy(self._g)
            self._g = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Model:
    """Model::from_gltf (uploads) + the BLAS/TLAS build, on the device."""

    def __init__(self, ctx: Renderer, scene_desc):
        self.ctx = ctx
        ctx._check(ctx._lib.rt2_upload_scene(ctx._ctx, scene_desc), "rt2_upload_scene")
        ctx._check(ctx._lib.rt2_build_accel(ctx._ctx), "rt2_build_accel")

    @classmethod
    def from_scene(cls, ctx: Renderer, scene: Scene) -> "Model":
        return cls(ctx, scene.desc)


class LightPasses:
    def __init__(self, ctx: Renderer, width: int, height: int):
        self.ctx = ctx
        ctx.resize(width, height)

    def update_uniform(self, g_const: GConst):
        self.ctx._check(self.ctx._lib.rt2_set_constants(self.ctx._ctx, C.byref(g_const), C.sizeof(g_const)), "rt2_set_constants")

    def execute(self, frame: int, params: LaunchParams | None = None):
        p = C.byref(params) if params is not None else None
        self.ctx._check(self.ctx._lib.rt2_render(self.ctx._ctx, frame & 0xFFFFFFFF, p), "rt2_render")

    def execute_presampeling(self, frame: int, skybox_changed: bool) -> bool:
        """src/light_passes.rs:510-548 (K11, spelling as in the reference): local-light presampling every frame, the
        environment map only when the sky box changed; returns the reference's `skybox_dirty` follow-up value."""
        self.ctx._check(self.ctx._lib.rt2_presample(self.ctx._ctx, frame & 0xFFFFFFFF, 0), "rt2_presample")
        if skybox_changed:
            self.ctx._check(self.ctx._lib.rt2_presample(self.ctx._ctx, frame & 0xFFFFFFFF, 1), "rt2_presample")
            return False
        return True

    def trace_primary(self, params: LaunchParams | None = None):
        p = C.byref(params) if params is not None else None
        self.ctx._check(self.ctx._lib.rt2_trace_primary(self.ctx._ctx, p), "rt2_trace_primary")


def trace_primary_noalloc(ctx: "Renderer", device_hits_ptr: int, nbytes: int, device_hit_count_ptr: int = 0, params: LaunchParams | None = None):
    """rt2_trace_primary_noalloc: primary visibility into caller-owned device memory, allocating nothing (static scratch in the
    library, at most 2^23 padded pixels).  No counterpart in the reference."""
    p = C.byref(params) if params is not None else None
    ctx._check(ctx._lib.rt2_trace_primary_noalloc(ctx._ctx, p, C.c_void_p(device_hits_ptr), nbytes, C.c_void_p(device_hit_count_ptr or None)),
               "rt2_trace_primary_noalloc")


class PrepareLightsTasks:
    def __init__(self, ctx: Renderer):
        self.ctx = ctx

    def execute(self):
        self.ctx._check(self.ctx._lib.rt2_prepare_lights(self.ctx._ctx), "rt2_prepare_lights")


class GenerateMipsPass:
    LOCAL_LIGHTS, ENVIRONMENT = 0, 1

    def __init__(self, ctx: Renderer, which: int):
        self.ctx, self.which = ctx, which

    def execute_mip_generation(self):
        self.ctx._check(self.ctx._lib.rt2_generate_mips(self.ctx._ctx, self.which), "rt2_generate_mips")


class PostProcessPass:
    def __init__(self, ctx: Renderer):
        self.ctx = ctx

    def execute(self, frame: int):
        self.ctx._check(self.ctx._lib.rt2_postprocess(self.ctx._ctx, frame & 0xFFFFFFFF), "rt2_postprocess")
