//! Raw bindings to `librt2cuda.so` — one item per declaration of `include/rt2_abi.h`.
//!
//! The structs are the reference's own `#[repr(C)]` upload types, so the host can pass its vectors without conversion:
//! `Rt2Vertex` == `gltf::Vertex` (src/gltf/mod.rs:41-48), `Rt2GeometryInfo` == `model::GeometryInfo` (src/model.rs:12-23),
//! and `rt2_set_constants` takes `&GConst` (src/shader_params.rs:243-274) as bytes.
#![allow(non_camel_case_types, non_upper_case_globals)]
use core::ffi::{c_char, c_int, c_void};

pub const RT2_ABI_VERSION: c_int = 1;
pub const RT2_OK: c_int = 0;
pub const RT2_E_INVALID: c_int = -1;
pub const RT2_E_CUDA: c_int = -2;
pub const RT2_E_OOM: c_int = -3;
pub const RT2_E_NO_DEVICE: c_int = -4;
pub const RT2_E_STATE: c_int = -5;
pub const RT2_GCONST_SIZE: usize = 1424;
pub const RT2_NEIGHBOR_OFFSET_COUNT: u32 = 8192;

pub const RT2_FILTER_NEAREST: u32 = 0;
pub const RT2_FILTER_LINEAR: u32 = 1;

pub const RT2_LAUNCH_TRIG_LIBDEVICE: u32 = 1;
pub const RT2_LAUNCH_PROFILE: u32 = 2;
pub const RT2_LAUNCH_PRESAMPLING: u32 = 4;

pub const RT2_SECONDARY_IS_SPECULAR_RAY: u32 = 1;
pub const RT2_SECONDARY_IS_DELTA_SURFACE: u32 = 2;
pub const RT2_SECONDARY_IS_ENVIRONMENT: u32 = 4;

// enum rt2_buffer
pub const RT2_BUF_DIFFUSE_F16: c_int = 0;
pub const RT2_BUF_SPECULAR_F16: c_int = 1;
pub const RT2_BUF_DIFFUSE_F32: c_int = 2;
pub const RT2_BUF_HIT_IDS: c_int = 3;
pub const RT2_BUF_OUTPUT_RGBA8: c_int = 4;
pub const RT2_BUF_LIGHT_INFO: c_int = 5;
pub const RT2_BUF_GEOM_TO_LIGHT: c_int = 6;
pub const RT2_BUF_LIGHT_TASKS: c_int = 7;
pub const RT2_BUF_LOCAL_PDF_MIP0: c_int = 16;
pub const RT2_BUF_ENV_PDF_MIP0: c_int = 48;
pub const RT2_BUF_GBUFFER_DEPTH: c_int = 80;
pub const RT2_BUF_GBUFFER_NORMALS: c_int = 81;
pub const RT2_BUF_GBUFFER_GEO_NORMALS: c_int = 82;
pub const RT2_BUF_GBUFFER_DIFFUSE_ALBEDO: c_int = 83;
pub const RT2_BUF_GBUFFER_SPECULAR_ROUGH: c_int = 84;
pub const RT2_BUF_GBUFFER_EMISSIVE: c_int = 85;
pub const RT2_BUF_MOTION_VECTORS: c_int = 86;
pub const RT2_BUF_DI_RESERVOIRS: c_int = 87;
pub const RT2_BUF_GI_RESERVOIRS: c_int = 88;
pub const RT2_BUF_SECONDARY_GBUFFER: c_int = 89;
pub const RT2_BUF_NEIGHBOR_OFFSETS: c_int = 90;
pub const RT2_BUF_RIS: c_int = 91;
pub const RT2_BUF_RIS_LIGHT_DATA: c_int = 92;
pub const RT2_BUF_BLAS_NODES: c_int = 100;
pub const RT2_BUF_BLAS_TRIS: c_int = 101;
pub const RT2_BUF_BLAS_SLOT_TO_PRIM: c_int = 102;
pub const RT2_BUF_TLAS_NODES: c_int = 103;
pub const RT2_BUF_TLAS_SLOT_TO_INST: c_int = 104;

/// 64 bytes — src/gltf/mod.rs:41-48 `Vertex`
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt2_vertex {
    pub pos: [f32; 4],
    pub normal: [f32; 4],
    pub color: [f32; 4],
    pub uv: [f32; 2],
    pub _pad: [f32; 2],
}

/// 128 bytes — src/model.rs:12-23 `GeometryInfo`
#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct rt2_geometry_info {
    pub transform: [f32; 16],
    pub base_color: [f32; 4],
    pub base_color_texture_index: i32,
    pub metallic_factor: f32,
    pub index_offset: u32,
    pub vertex_offset: u32,
    pub emission: [f32; 4],
    pub roughness: f32,
    pub _pad: [f32; 3],
}

#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct rt2_image_desc {
    pub rgba8: *const u8,
    pub width: u32,
    pub height: u32,
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt2_sampler_desc {
    pub mag_filter: u32,
    pub min_filter: u32,
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt2_texture_desc {
    pub image_index: u32,
    pub sampler_index: u32,
}

#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct rt2_scene_desc {
    pub vertices: *const rt2_vertex,
    pub num_vertices: u64,
    pub indices: *const u32,
    pub num_indices: u64,
    pub geometries: *const rt2_geometry_info,
    pub index_counts: *const u32,
    pub num_geometries: u32,
    pub images: *const rt2_image_desc,
    pub num_images: u32,
    pub samplers: *const rt2_sampler_desc,
    pub num_samplers: u32,
    pub textures: *const rt2_texture_desc,
    pub num_textures: u32,
    pub sky_rgba32f: *const f32,
    pub sky_width: u32,
    pub sky_height: u32,
    pub instance_transforms: *const f32,
    pub num_instances: u32,
}

/// 16 bytes — src/prepare_lights.rs:13-20
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt2_prepare_lights_task {
    pub geometry_index: u32,
    pub light_buffer_offset: u32,
    pub triangle_count: u32,
    pub pad: u32,
}

/// 48 bytes — PolymorphicLight.glsl:19-36 `RAB_LightInfo`
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt2_light_info {
    pub center: [f32; 3],
    pub color_type_and_flags: u32,
    pub direction1: u32,
    pub direction2: u32,
    pub scalars: u32,
    pub log_radiance: u32,
    pub ies_profile_index: u32,
    pub primary_axis: u32,
    pub cos_cone_angle_and_softness: u32,
    pub padding: u32,
}

/// 24 bytes — rtxdi/RtxdiParameters.h:97-105 `RTXDI_PackedDIReservoir`
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt2_packed_di_reservoir {
    pub light_data: u32,
    pub uv_data: u32,
    pub m_visibility: u32,
    pub distance_age: u32,
    pub target_pdf: f32,
    pub weight: f32,
}

/// 32 bytes — rtxdi/RtxdiParameters.h `RTXDI_PackedGIReservoir`
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt2_packed_gi_reservoir {
    pub position: [f32; 3],
    pub packed_misc_age_m: u32,
    pub packed_radiance: u32,
    pub weight: f32,
    pub packed_normal: u32,
    pub unused: f32,
}

/// 48 bytes — ShaderParameters.glsl `SecondaryGBufferData`
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt2_secondary_gbuffer_data {
    pub world_pos: [f32; 3],
    pub normal: u32,
    pub throughput_and_flags: [u32; 2],
    pub diffuse_albedo: u32,
    pub specular_and_roughness: u32,
    pub emission: [f32; 3],
    pub pdf: f32,
}

/// 32 bytes
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt2_hit_record {
    pub instance: u32,
    pub geometry: u32,
    pub primitive: u32,
    pub t: f32,
    pub u: f32,
    pub v: f32,
    pub _pad: [u32; 2],
}

/// 32 bytes; zeroed == the reference's hard-coded 12 spp x 5 bounces, whole frame
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt2_launch_params {
    pub spp: u32,
    pub max_bounces: u32,
    pub band_rows: u32,
    pub band_rank: u32,
    pub band_world: u32,
    pub flags: u32,
    pub _reserved: [u32; 2],
}

/// 96 bytes
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rt2_stats {
    pub rays_traced: u64,
    pub rays_hit: u64,
    pub kernel_launches: u64,
    pub wavefront_iterations: u64,
    pub bvh_nodes: u64,
    pub bvh_triangles: u64,
    pub device_bytes: u64,
    pub trace_ns: u64,
    pub shade_ns: u64,
    pub trace_launches: u64,
    pub sample_schedule: u64,
    pub rays_walked: u64,
}

#[repr(C)]
pub struct rt2_ctx {
    _private: [u8; 0],
}
/// Opaque multi-GPU group (include/rt2_abi.h, "multi-GPU").
#[repr(C)]
pub struct rt2_group {
    _private: [u8; 0],
}
pub const RT2_GROUP_BAND_ROWS_ROWS: u64 = 8;
pub const RT2_UNIQUE_ID_BYTES_LEN: u64 = 128;

extern "C" {
    pub fn rt2_abi_version() -> c_int;
    pub fn rt2_create(cuda_device: c_int, out: *mut *mut rt2_ctx) -> c_int;
    pub fn rt2_destroy(ctx: *mut rt2_ctx) -> c_int;
    pub fn rt2_set_stream(ctx: *mut rt2_ctx, cuda_stream: *mut c_void) -> c_int;
    pub fn rt2_last_error(ctx: *mut rt2_ctx) -> *const c_char;
    pub fn rt2_upload_scene(ctx: *mut rt2_ctx, scene: *const rt2_scene_desc) -> c_int;
    pub fn rt2_build_accel(ctx: *mut rt2_ctx) -> c_int;
    pub fn rt2_resize(ctx: *mut rt2_ctx, width: u32, height: u32) -> c_int;
    pub fn rt2_set_constants(ctx: *mut rt2_ctx, gconst: *const c_void, nbytes: usize) -> c_int;
    pub fn rt2_prepare_lights(ctx: *mut rt2_ctx) -> c_int;
    pub fn rt2_generate_mips(ctx: *mut rt2_ctx, which: c_int) -> c_int;
    pub fn rt2_presample(ctx: *mut rt2_ctx, frame: u32, which: c_int) -> c_int;
    pub fn rt2_render(ctx: *mut rt2_ctx, frame: u32, params: *const rt2_launch_params) -> c_int;
    pub fn rt2_trace_primary(ctx: *mut rt2_ctx, params: *const rt2_launch_params) -> c_int;
    /// No counterpart in the reference: allocates nothing, static scratch, at most 2^23 padded pixels.
    pub fn rt2_trace_primary_noalloc(ctx: *mut rt2_ctx, params: *const rt2_launch_params, device_hits_out: *mut c_void, nbytes: usize, device_hit_count_out: *mut u64) -> c_int;
    pub fn rt2_postprocess(ctx: *mut rt2_ctx, frame: u32) -> c_int;
    pub fn rt2_sync(ctx: *mut rt2_ctx) -> c_int;
    pub fn rt2_buffer_size(ctx: *mut rt2_ctx, which: c_int) -> usize;
    pub fn rt2_read(ctx: *mut rt2_ctx, which: c_int, dst: *mut c_void, nbytes: usize) -> c_int;
    pub fn rt2_read_async(ctx: *mut rt2_ctx, which: c_int, dst: *mut c_void, nbytes: usize) -> c_int;
    pub fn rt2_read_fence(ctx: *mut rt2_ctx) -> c_int;
    pub fn rt2_copy_to_device(ctx: *mut rt2_ctx, which: c_int, device_dst: *mut c_void, nbytes: usize, row0: u32, row1: u32) -> c_int;
    pub fn rt2_pack_bands(ctx: *mut rt2_ctx, which: c_int, device_dst: *mut c_void, nbytes: usize, band_rows: u32, band_rank: u32, band_world: u32) -> c_int;
    pub fn rt2_write(ctx: *mut rt2_ctx, which: c_int, src: *const c_void, nbytes: usize) -> c_int;
    pub fn rt2_get_stats(ctx: *mut rt2_ctx, out: *mut rt2_stats) -> c_int;

    // multi-GPU: one frame over the GPUs of one node (band plan + NCCL gather + un-permute inside the library)
    pub fn rt2_group_create(cuda_devices: *const c_int, n: c_int, out: *mut *mut rt2_group) -> c_int;
    pub fn rt2_group_unique_id(id128: *mut c_void) -> c_int;
    pub fn rt2_group_create_rank(cuda_device: c_int, rank: c_int, world: c_int, id128: *const c_void, out: *mut *mut rt2_group) -> c_int;
    pub fn rt2_group_destroy(g: *mut rt2_group) -> c_int;
    pub fn rt2_group_last_error(g: *mut rt2_group) -> *const c_char;
    pub fn rt2_group_world(g: *mut rt2_group) -> c_int;
    pub fn rt2_group_local_count(g: *mut rt2_group) -> c_int;
    pub fn rt2_group_ctx(g: *mut rt2_group, local_index: c_int) -> *mut rt2_ctx;
    pub fn rt2_group_upload_scene(g: *mut rt2_group, scene: *const rt2_scene_desc) -> c_int;
    pub fn rt2_group_build_accel(g: *mut rt2_group) -> c_int;
    pub fn rt2_group_resize(g: *mut rt2_group, width: u32, height: u32) -> c_int;
    pub fn rt2_group_set_constants(g: *mut rt2_group, gconst: *const c_void, nbytes: usize) -> c_int;
    pub fn rt2_group_prepare_lights(g: *mut rt2_group) -> c_int;
    pub fn rt2_group_generate_mips(g: *mut rt2_group, which: c_int) -> c_int;
    pub fn rt2_group_render(g: *mut rt2_group, frame: u32, params: *const rt2_launch_params, gather_which: c_int) -> c_int;
    pub fn rt2_group_gather(g: *mut rt2_group, which: c_int) -> c_int;
    pub fn rt2_group_postprocess(g: *mut rt2_group, frame: u32) -> c_int;
    pub fn rt2_group_read(g: *mut rt2_group, dst: *mut c_void, nbytes: usize) -> c_int;
    pub fn rt2_group_read_async(g: *mut rt2_group, dst: *mut c_void, nbytes: usize) -> c_int;
    pub fn rt2_group_assembled(g: *mut rt2_group, local_index: c_int, nbytes: *mut usize) -> *mut c_void;
    pub fn rt2_group_wait_gather(g: *mut rt2_group) -> c_int;
    pub fn rt2_group_sync(g: *mut rt2_group) -> c_int;
}

// Layout checks that `cargo build` evaluates at compile time; the same numbers are asserted against the C header by the
// Python test named in Cargo.toml, because this repository cannot run cargo.
const _: () = {
    assert!(core::mem::size_of::<rt2_vertex>() == 64);
    assert!(core::mem::size_of::<rt2_geometry_info>() == 128);
    assert!(core::mem::size_of::<rt2_prepare_lights_task>() == 16);
    assert!(core::mem::size_of::<rt2_light_info>() == 48);
    assert!(core::mem::size_of::<rt2_packed_di_reservoir>() == 24);
    assert!(core::mem::size_of::<rt2_packed_gi_reservoir>() == 32);
    assert!(core::mem::size_of::<rt2_secondary_gbuffer_data>() == 48);
    assert!(core::mem::size_of::<rt2_hit_record>() == 32);
    assert!(core::mem::size_of::<rt2_launch_params>() == 32);
    assert!(core::mem::size_of::<rt2_stats>() == 96);
};
