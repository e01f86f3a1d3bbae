"""ctypes bindings: the C-ABI product library (librt2cuda.so) and the C++ host library
(librt2host.so).  Python is glue only — tests and bench drive the same C ABI the Rust host would
bind.  The CUDA library is loaded lazily and loudly: there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# RT2_CUDA_LIB: an alternative build of the same library (kernel A/B experiments, tools/gpu_ab.sh); never a different backend
CUDA_LIB_PATH = os.environ.get("RT2_CUDA_LIB") or os.path.join(_HERE, "librt2cuda.so")
HOST_LIB_PATH = os.path.join(_HERE, "librt2host.so")

RT2_GCONST_SIZE = 1424

u32, u64, f32, i32 = C.c_uint32, C.c_uint64, C.c_float, C.c_int32


def _fields(names, ty):
    return [(n, ty) for n in names.split()]


class Vertex(C.Structure):  # include/rt2_abi.h rt2_vertex (src/gltf/mod.rs:41-48)
    _fields_ = [("pos", f32 * 4), ("normal", f32 * 4), ("color", f32 * 4), ("uv", f32 * 2), ("_pad", f32 * 2)]


class GeometryInfo(C.Structure):  # rt2_geometry_info (src/model.rs:12-23)
    _fields_ = [("transform", f32 * 16), ("base_color", f32 * 4), ("base_color_texture_index", i32),
                ("metallic_factor", f32), ("index_offset", u32), ("vertex_offset", u32),
                ("emission", f32 * 4), ("roughness", f32), ("_pad", f32 * 3)]


class ImageDesc(C.Structure):
    _fields_ = [("rgba8", C.POINTER(C.c_uint8)), ("width", u32), ("height", u32)]


class SamplerDesc(C.Structure):
    _fields_ = [("mag_filter", u32), ("min_filter", u32)]


class TextureDesc(C.Structure):
    _fields_ = [("image_index", u32), ("sampler_index", u32)]


class SceneDesc(C.Structure):  # rt2_scene_desc
    _fields_ = [("vertices", C.POINTER(Vertex)), ("num_vertices", u64),
                ("indices", C.POINTER(u32)), ("num_indices", u64),
                ("geometries", C.POINTER(GeometryInfo)), ("index_counts", C.POINTER(u32)), ("num_geometries", u32),
                ("images", C.POINTER(ImageDesc)), ("num_images", u32),
                ("samplers", C.POINTER(SamplerDesc)), ("num_samplers", u32),
                ("textures", C.POINTER(TextureDesc)), ("num_textures", u32),
                ("sky_rgba32f", C.POINTER(f32)), ("sky_width", u32), ("sky_height", u32),
                ("instance_transforms", C.POINTER(f32)), ("num_instances", u32)]


class HitRecord(C.Structure):
    _fields_ = [("instance", u32), ("geometry", u32), ("primitive", u32), ("t", f32), ("u", f32), ("v", f32), ("_pad", u32 * 2)]


class LaunchParams(C.Structure):
    _fields_ = [("spp", u32), ("max_bounces", u32), ("band_rows", u32), ("band_rank", u32),
                ("band_world", u32), ("flags", u32), ("_reserved", u32 * 2)]


class Stats(C.Structure):
    _fields_ = _fields("rays_traced rays_hit kernel_launches wavefront_iterations bvh_nodes bvh_triangles device_bytes trace_ns shade_ns trace_launches sample_schedule rays_walked", u64)


# ---- GConst (include/rt2_gconst.h; src/shader_params.rs) ----------------------------------------------
class PlanarViewConstants(C.Structure):
    _fields_ = [("mat_world_to_view", f32 * 16), ("mat_view_to_clip", f32 * 16), ("mat_world_to_clip", f32 * 16),
                ("mat_clip_to_view", f32 * 16), ("mat_view_to_world", f32 * 16), ("mat_clip_to_world", f32 * 16),
                ("viewport_origin", f32 * 2), ("viewport_size", f32 * 2), ("viewport_size_inv", f32 * 2),
                ("pixel_offset", f32 * 2), ("clip_to_window_scale", f32 * 2), ("clip_to_window_bias", f32 * 2),
                ("window_to_clip_scale", f32 * 2), ("window_to_clip_bias", f32 * 2),
                ("camera_direction_or_position", f32 * 4)]


class RuntimeParameters(C.Structure):
    _fields_ = _fields("neighbor_offset_mask active_checkerboard_field pad1 pad2", u32)


class GITemporal(C.Structure):
    _fields_ = [("depth_threshold", f32), ("normal_threshold", f32), ("enable_permutation_sampling", u32),
                ("max_history_length", u32), ("max_reservoir_age", u32), ("enable_boiling_filter", u32),
                ("boiling_filter_strength", f32), ("enable_fallback_sampling", u32),
                ("temporal_bias_correction_mode", u32), ("uniform_random_number", u32), ("pad2", u32), ("pad3", u32)]


class GISpatial(C.Structure):
    _fields_ = [("spatial_depth_threshold", f32), ("spatial_normal_threshold", f32), ("num_spatial_samples", u32),
                ("spatial_sampling_radius", f32), ("spatial_bias_correction_mode", u32), ("pad1", u32), ("pad2", u32), ("pad3", u32)]


class GIFinal(C.Structure):
    _fields_ = _fields("enable_final_visibility enable_final_mis pad1 pad2", u32)


class GIBufferIndices(C.Structure):
    _fields_ = _fields("secondary_surface_restirdi_output_buffer_index temporal_resampling_input_buffer_index "
                       "temporal_resampling_output_buffer_index spatial_resampling_input_buffer_index "
                       "spatial_resampling_output_buffer_index final_shading_input_buffer_index pad1 pad2", u32)


class ReservoirBufferParams(C.Structure):
    _fields_ = _fields("reservoir_block_row_pitch reservoir_array_pitch pad1 pad2", u32)


class RestirGI(C.Structure):
    _fields_ = [("reservoir_buffer_params", ReservoirBufferParams), ("buffer_indices", GIBufferIndices),
                ("temporal_resampling_params", GITemporal), ("spatial_resampling_params", GISpatial),
                ("final_shading_params", GIFinal)]


class LightBufferRegion(C.Structure):
    _fields_ = _fields("first_light_index num_lights pad1 pad2", u32)


class EnvLightParams(C.Structure):
    _fields_ = _fields("light_present light_index pad1 pad2", u32)


class LightBufferParams(C.Structure):
    _fields_ = [("local_light_buffer_region", LightBufferRegion), ("infinite_light_buffer_region", LightBufferRegion),
                ("environment_light_params", EnvLightParams)]


class DIInitial(C.Structure):
    _fields_ = [("num_primary_local_light_samples", u32), ("num_primary_infinite_light_samples", u32),
                ("num_primary_environment_samples", u32), ("num_primary_brdf_samples", u32),
                ("brdf_cutoff", f32), ("enable_initial_visibility", u32),
                ("environment_map_importance_sampling", u32), ("local_light_sampling_mode", u32)]


class DIBufferIndices(C.Structure):
    _fields_ = _fields("initial_sampling_output_buffer_index temporal_resampling_input_buffer_index "
                       "temporal_resampling_output_buffer_index spatial_resampling_input_buffer_index "
                       "spatial_resampling_output_buffer_index shading_input_buffer_index pad1 pad2", u32)


class DITemporal(C.Structure):
    _fields_ = [("temporal_depth_threshold", f32), ("temporal_normal_threshold", f32), ("max_history_length", u32),
                ("temporal_bias_correction", u32), ("enable_permutation_sampling", u32),
                ("permutation_sampling_threshold", f32), ("enable_boiling_filter", u32), ("boiling_filter_strength", f32),
                ("discard_invisible_samples", u32), ("uniform_random_number", u32), ("pad2", u32), ("pad3", u32)]


class DISpatial(C.Structure):
    _fields_ = [("spatial_depth_threshold", f32), ("spatial_normal_threshold", f32), ("spatial_bias_correction", u32),
                ("num_spatial_samples", u32), ("num_disocclusion_boost_samples", u32), ("spatial_sampling_radius", f32),
                ("neighbor_offset_mask", u32), ("discount_naive_samples", u32)]


class DIShading(C.Structure):
    _fields_ = [("enable_final_visibility", u32), ("reuse_final_visibility", u32), ("final_visibility_max_age", u32),
                ("final_visibility_max_distance", f32), ("enable_denoiser_input_packing", u32),
                ("pad1", u32), ("pad2", u32), ("pad3", u32)]


class RestirDI(C.Structure):
    _fields_ = [("reservoir_buffer_params", ReservoirBufferParams), ("buffer_indices", DIBufferIndices),
                ("initial_sampling_params", DIInitial), ("temporal_resampling_params", DITemporal),
                ("spatial_resampling_params", DISpatial), ("shading_params", DIShading)]


class RisSegment(C.Structure):
    _fields_ = _fields("buffer_offset tile_size tile_count pad1", u32)


class GConst(C.Structure):
    _fields_ = [("view", PlanarViewConstants), ("prev_view", PlanarViewConstants), ("runtime_params", RuntimeParameters),
                ("enable_brdf_indirect", u32), ("enable_brdf_additive_blend", u32), ("enable_accumulation", u32), ("frame", u32),
                ("restir_gi", RestirGI), ("restir_di", RestirDI),
                ("enable_restir_di", u32), ("enable_restir_gi", u32), ("refrence_mode", u32), ("textures", u32),
                ("blend_factor", f32), ("enable_spatial_resampling", u32), ("enable_temporal_resampling", u32), ("environment", u32),
                ("light_buffer_params", LightBufferParams),
                ("local_lights_risbuffer_segment_params", RisSegment), ("environment_light_risbuffer_segment_params", RisSegment),
                ("environment_pdf_texture_size", u32 * 2), ("local_light_pdf_texture_size", u32 * 2)]


assert C.sizeof(PlanarViewConstants) == 464
assert C.sizeof(GConst) == RT2_GCONST_SIZE
assert C.sizeof(Vertex) == 64 and C.sizeof(GeometryInfo) == 128 and C.sizeof(HitRecord) == 32


class Camera(C.Structure):  # rt2h_camera (src/camera.rs:16-24)
    _fields_ = [("position", f32 * 3), ("direction", f32 * 3), ("fov", f32), ("aspect_ratio", f32),
                ("z_near", f32), ("z_far", f32)]


# ---- symbol tables (kept in one place so the CPU test can check every export of include/rt2_abi.h) -------
VP = C.c_void_p
ABI_SYMBOLS = {
    "rt2_abi_version": (C.c_int, []),
    "rt2_create": (C.c_int, [C.c_int, C.POINTER(VP)]),
    "rt2_destroy": (C.c_int, [VP]),
    "rt2_set_stream": (C.c_int, [VP, VP]),
    "rt2_last_error": (C.c_char_p, [VP]),
    "rt2_upload_scene": (C.c_int, [VP, C.POINTER(SceneDesc)]),
    "rt2_build_accel": (C.c_int, [VP]),
    "rt2_resize": (C.c_int, [VP, u32, u32]),
    "rt2_set_constants": (C.c_int, [VP, VP, C.c_size_t]),
    "rt2_prepare_lights": (C.c_int, [VP]),
    "rt2_generate_mips": (C.c_int, [VP, C.c_int]),
    "rt2_presample": (C.c_int, [VP, u32, C.c_int]),
    "rt2_render": (C.c_int, [VP, u32, C.POINTER(LaunchParams)]),
    "rt2_trace_primary": (C.c_int, [VP, C.POINTER(LaunchParams)]),
    "rt2_trace_primary_noalloc": (C.c_int, [VP, C.POINTER(LaunchParams), VP, C.c_size_t, VP]),
    "rt2_postprocess": (C.c_int, [VP, u32]),
    "rt2_sync": (C.c_int, [VP]),
    "rt2_buffer_size": (C.c_size_t, [VP, C.c_int]),
    "rt2_read": (C.c_int, [VP, C.c_int, VP, C.c_size_t]),
    "rt2_read_async": (C.c_int, [VP, C.c_int, VP, C.c_size_t]),
    "rt2_read_fence": (C.c_int, [VP]),
    "rt2_copy_to_device": (C.c_int, [VP, C.c_int, VP, C.c_size_t, u32, u32]),
    "rt2_pack_bands": (C.c_int, [VP, C.c_int, VP, C.c_size_t, u32, u32, u32]),
    "rt2_write": (C.c_int, [VP, C.c_int, VP, C.c_size_t]),
    "rt2_get_stats": (C.c_int, [VP, C.POINTER(Stats)]),
    # multi-GPU group (rt2_group.cu)
    "rt2_group_create": (C.c_int, [C.POINTER(C.c_int), C.c_int, C.POINTER(VP)]),
    "rt2_group_unique_id": (C.c_int, [VP]),
    "rt2_group_create_rank": (C.c_int, [C.c_int, C.c_int, C.c_int, VP, C.POINTER(VP)]),
    "rt2_group_destroy": (C.c_int, [VP]),
    "rt2_group_last_error": (C.c_char_p, [VP]),
    "rt2_group_world": (C.c_int, [VP]),
    "rt2_group_local_count": (C.c_int, [VP]),
    "rt2_group_ctx": (VP, [VP, C.c_int]),
    "rt2_group_upload_scene": (C.c_int, [VP, C.POINTER(SceneDesc)]),
    "rt2_group_build_accel": (C.c_int, [VP]),
    "rt2_group_resize": (C.c_int, [VP, u32, u32]),
    "rt2_group_set_constants": (C.c_int, [VP, VP, C.c_size_t]),
    "rt2_group_prepare_lights": (C.c_int, [VP]),
    "rt2_group_generate_mips": (C.c_int, [VP, C.c_int]),
    "rt2_group_render": (C.c_int, [VP, u32, C.POINTER(LaunchParams), C.c_int]),
    "rt2_group_gather": (C.c_int, [VP, C.c_int]),
    "rt2_group_postprocess": (C.c_int, [VP, u32]),
    "rt2_group_read": (C.c_int, [VP, VP, C.c_size_t]),
    "rt2_group_read_async": (C.c_int, [VP, VP, C.c_size_t]),
    "rt2_group_assembled": (VP, [VP, C.c_int, C.POINTER(C.c_size_t)]),
    "rt2_group_wait_gather": (C.c_int, [VP]),
    "rt2_group_sync": (C.c_int, [VP]),
}

HOST_SYMBOLS = {
    "rt2h_camera_new": (None, [C.POINTER(Camera), C.POINTER(f32), C.POINTER(f32), f32, f32, f32, f32]),
    "rt2h_camera_view_matrix": (None, [C.POINTER(Camera), C.POINTER(f32)]),
    "rt2h_perspective": (None, [f32, f32, f32, f32, C.POINTER(f32)]),
    "rt2h_planar_view_constants": (None, [C.POINTER(Camera), u32, u32, C.POINTER(PlanarViewConstants)]),
    "rt2h_fill_neighbor_offset_buffer": (u32, [u32, C.POINTER(C.c_uint8)]),
    "rt2h_compute_pdf_texture_size": (None, [u32, C.POINTER(u32), C.POINTER(u32), C.POINTER(u32)]),
    "rt2h_calculate_reservoir_buffer_parameters": (None, [u32, u32, C.POINTER(ReservoirBufferParams)]),
    "rt2h_default_gconst": (None, [C.POINTER(GConst), C.POINTER(Camera), u32, u32, u32, u32, u32]),
    "rt2h_advance_frame": (None, [C.POINTER(GConst), C.POINTER(Camera), u32, u32, u32]),
    "rt2h_scene_generate": (VP, [C.c_int, f32, u32, u64]),
    "rt2h_scene_load_gltf": (VP, [C.c_char_p, C.c_char_p, C.c_char_p, C.c_size_t]),
    "rt2h_load_exr": (VP, [C.c_char_p, C.POINTER(u32), C.POINTER(u32), C.c_char_p, C.c_size_t]),
    "rt2h_decode_image": (VP, [C.c_char_p, C.c_size_t, C.POINTER(u32), C.POINTER(u32), C.POINTER(u32), C.c_char_p, C.c_size_t]),
    "rt2h_free": (None, [VP]),
    "rt2h_scene_destroy": (None, [VP]),
    "rt2h_scene_desc": (C.POINTER(SceneDesc), [VP]),
    "rt2h_scene_camera": (None, [VP, f32, C.POINTER(Camera)]),
    "rt2h_scene_num_lights": (u32, [VP]),
    "rt2h_scene_add_procedural_sky": (None, [VP, u32]),
    "rt2h_scene_num_triangles": (u64, [VP]),
}


def _bind(lib, table):
    for name, (res, args) in table.items():
        fn = getattr(lib, name)  # AttributeError if the export is missing
        fn.restype = res
        fn.argtypes = args
    return lib


_cuda = None
_host = None


def host_lib():
    global _host
    if _host is None:
        if not os.path.exists(HOST_LIB_PATH):
            raise RuntimeError(f"{HOST_LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
        _host = _bind(C.CDLL(HOST_LIB_PATH), HOST_SYMBOLS)
    return _host


def cuda_lib():
    """The product library.  Raises if it is not built — never falls back to anything else."""
    global _cuda
    if _cuda is None:
        if not os.path.exists(CUDA_LIB_PATH):
            raise RuntimeError(f"{CUDA_LIB_PATH} is missing: the CUDA extension must be built "
                               "(`python -c 'import __graft_entry__ as g; g.build()'`); there is no CPU fallback")
        _cuda = _bind(C.CDLL(CUDA_LIB_PATH), ABI_SYMBOLS)
    return _cuda
