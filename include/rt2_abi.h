/* rt2_abi.h — C ABI of librt2cuda.so, the B200 (sm_100a) CUDA replacement for RayTracer2's
 * per-pixel path-tracing hot path.
 *
 * The reference (a Rust/Vulkan application) has no FFI surface of its own; the seam this ABI
 * sits behind is the Rust-internal boundary between its pass objects and Vulkan + SPIR-V
 * (SURVEY.md §8b).  Every entry point below names the reference interface it replaces
 * (paths relative to the reference tree).  All structs are plain-old-data with the byte layouts
 * the reference uploads to the GPU, so the Rust host can pass its own buffers unchanged.
 *
 * Conventions: 0 = OK, negative = error; nothing throws or aborts across the boundary; the
 * caller owns every host pointer (copied before the call returns); the context owns every device
 * allocation; a context is NOT thread-safe (the reference records from one thread,
 * src/context.rs:639-694); work is enqueued on the context's CUDA stream in call order (this
 * replaces the reference's ALL_COMMANDS barriers, src/light_passes.rs:552-558) and is
 * asynchronous until rt2_sync / rt2_read.  There is no CPU fallback.
 */
#ifndef RT2_ABI_H
#define RT2_ABI_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT2_ABI_VERSION 1

/* ---- error codes ------------------------------------------------------------------------- */
#define RT2_OK            0
#define RT2_E_INVALID    -1   /* bad argument / call order                                    */
#define RT2_E_CUDA       -2   /* a CUDA runtime call failed (see rt2_last_error)              */
#define RT2_E_OOM        -3   /* device or host allocation failed                             */
#define RT2_E_NO_DEVICE  -4   /* no usable sm_100 device: there is no CPU fallback            */
#define RT2_E_STATE      -5   /* scene / accel / size not set up yet                          */

/* ---- scene layouts (reference byte layouts) ------------------------------------------------ */

/* src/gltf/mod.rs:41-48 `Vertex` (GLSL view: src/shaders/common.glsl:5-10). 64 bytes. */
typedef struct rt2_vertex {
    float pos[4];      /* xyz used, w unused                                  @0  */
    float normal[4];   /* xyz used                                            @16 */
    float color[4];    /* rgb used                                            @32 */
    float uv[2];       /*                                                     @48 */
    float _pad[2];     /* Rust repr(C) tail padding to the 16-byte alignment  @56 */
} rt2_vertex;

/* src/model.rs:12-23 `GeometryInfo` (GLSL: common.glsl:12-21). 128 bytes. */
typedef struct rt2_geometry_info {
    float    transform[16];              /* column-major (glam Mat4)           @0   */
    float    base_color[4];              /*                                    @64  */
    int32_t  base_color_texture_index;   /* -1 = none                          @80  */
    float    metallic_factor;            /*                                    @84  */
    uint32_t index_offset;               /*                                    @88  */
    uint32_t vertex_offset;              /*                                    @92  */
    float    emission[4];                /* w = 1                              @96  */
    float    roughness;                  /*                                    @112 */
    float    _pad[3];                    /*                                    @116 */
} rt2_geometry_info;

/* src/gltf/image.rs:57-109 — images arrive expanded to RGBA8; uploaded as R8G8B8A8_SRGB, one mip
 * (src/model.rs:224-287). */
typedef struct rt2_image_desc {
    const uint8_t* rgba8;
    uint32_t       width;
    uint32_t       height;
} rt2_image_desc;

#define RT2_FILTER_NEAREST 0
#define RT2_FILTER_LINEAR  1
/* src/model.rs:165-183 — only mag/min filter survive the glTF→Vulkan mapping; address mode is the
 * Vulkan default REPEAT. */
typedef struct rt2_sampler_desc {
    uint32_t mag_filter;
    uint32_t min_filter;
} rt2_sampler_desc;

/* src/model.rs:346-350 `textures: Vec<(image_index, sampler_index)>` */
typedef struct rt2_texture_desc {
    uint32_t image_index;
    uint32_t sampler_index;
} rt2_texture_desc;

/* What `Model::from_gltf(ctx, gltf::Model)` consumes (src/model.rs:185-476) plus the sky box
 * (src/main.rs:63,145) and the TLAS instance list (src/main.rs:153-191; the reference always has
 * exactly one identity instance — more than one is the documented extension of SURVEY.md §8a A2).
 * Single-BLAS by design: the reference builds ONE bottom-level structure holding every geometry of the model
 * (src/model.rs:379-459) and instances it once (src/model.rs:63-75); the extension keeps that shape — every
 * instance is a transform of that one BLAS, there is no per-instance BLAS index.  A scene made of several
 * different meshes is, as in the reference, several geometries of the one BLAS.
 */
typedef struct rt2_scene_desc {
    const rt2_vertex*        vertices;       uint64_t num_vertices;
    const uint32_t*          indices;        uint64_t num_indices;
    const rt2_geometry_info* geometries;     /* one per (node, primitive)                      */
    const uint32_t*          index_counts;   /* src/model.rs:35 `index_counts`, per geometry   */
    uint32_t                 num_geometries;
    const rt2_image_desc*    images;         uint32_t num_images;
    const rt2_sampler_desc*  samplers;       uint32_t num_samplers;
    const rt2_texture_desc*  textures;       uint32_t num_textures;
    const float*             sky_rgba32f;    /* equirect RGBA32F, may be NULL                  */
    uint32_t                 sky_width;
    uint32_t                 sky_height;
    const float*             instance_transforms; /* VkTransformMatrixKHR: 12 floats, 3x4 row-major */
    uint32_t                 num_instances;  /* 0 => one identity instance (the reference)     */
} rt2_scene_desc;

/* src/prepare_lights.rs:13-20 / src/shaders/prepare_lights.comp:22-27. 16 bytes. */
typedef struct rt2_prepare_lights_task {
    uint32_t geometry_index;
    uint32_t light_buffer_offset;
    uint32_t triangle_count;
    uint32_t pad;
} rt2_prepare_lights_task;

/* src/shaders/PolymorphicLight.glsl:19-36 `RAB_LightInfo`. 48 bytes. */
typedef struct rt2_light_info {
    float    center[3];
    uint32_t color_type_and_flags;
    uint32_t direction1;
    uint32_t direction2;
    uint32_t scalars;
    uint32_t log_radiance;
    uint32_t ies_profile_index;
    uint32_t primary_axis;
    uint32_t cos_cone_angle_and_softness;
    uint32_t padding;
} rt2_light_info;

/* src/shaders/rtxdi/RtxdiParameters.h:97-105 `RTXDI_PackedDIReservoir`. 24 bytes. */
typedef struct rt2_packed_di_reservoir {
    uint32_t light_data;
    uint32_t uv_data;
    uint32_t m_visibility;
    uint32_t distance_age;
    float    target_pdf;
    float    weight;
} rt2_packed_di_reservoir;

#define RT2_NEIGHBOR_OFFSET_COUNT 8192u   /* src/main.rs:56 NEIGHBOR_OFFSET_COUNT */

/* src/shaders/rtxdi/ReSTIRGIParameters.h:17-30 `RTXDI_PackedGIReservoir`. 32 bytes. */
typedef struct rt2_packed_gi_reservoir {
    float    position[3];
    uint32_t packed_misc_age_m;   /* M bits 0-7, age bits 8-15, misc bits 16-31 (GIReservoir.hlsli:52-61) */
    uint32_t packed_radiance;     /* 32-bit LogLuv (RtxdiMath.hlsli:233-264) */
    float    weight;
    uint32_t packed_normal;       /* 2 x snorm16 octahedral (RtxdiMath.hlsli:184-196) */
    float    unused;
} rt2_packed_gi_reservoir;

/* src/shaders/ShaderParameters.glsl:49-60 `SecondaryGBufferData`. 48 bytes; block-linear addressing like the reservoirs,
 * arrayPitch entries (src/render_resources.rs:309-317). */
typedef struct rt2_secondary_gbuffer_data {
    float    world_pos[3];
    uint32_t normal;                  /* octahedral unorm 2x16 */
    uint32_t throughput_and_flags[2]; /* .x = throughput.rg fp16, .y = throughput.b fp16 | flags << 16 */
    uint32_t diffuse_albedo;          /* R11G11B10_UFLOAT */
    uint32_t specular_and_roughness;  /* R8G8B8A8 gamma */
    float    emission[3];
    float    pdf;
} rt2_secondary_gbuffer_data;
#define RT2_SECONDARY_IS_SPECULAR_RAY   1u   /* ShaderParameters.glsl:21-23 */
#define RT2_SECONDARY_IS_DELTA_SURFACE  2u
#define RT2_SECONDARY_IS_ENVIRONMENT    4u

/* The 1424-byte `GConst` block (src/shader_params.rs:243-274 == GLSL `ResamplingConstants`,
 * src/shaders/ShaderParameters.glsl:62-92) is passed verbatim as bytes. */
#define RT2_GCONST_SIZE 1424 /* SURVEY.md says 1456; summing src/shader_params.rs field by field gives 464 (PlanarViewConstants) and 1424 */

/* One primary-visibility record per pixel (rt2_trace_primary / RT2_BUF_HIT_IDS). 32 bytes.
 * geometry/primitive follow gl_GeometryIndexEXT / gl_PrimitiveID (rayhit.rchit:60-66); a miss has
 * geometry = primitive = instance = 0xFFFFFFFF (raymiss.rmiss:12-17) and t = 0. */
typedef struct rt2_hit_record {
    uint32_t instance;
    uint32_t geometry;
    uint32_t primitive;
    float    t;
    float    u;   /* barycentric weight of vertex 1 */
    float    v;   /* barycentric weight of vertex 2 */
    uint32_t _pad[2];
} rt2_hit_record;

/* Launch parameters for values the reference hard-codes (refrence.rgen:16-17 maxBounces = 5,
 * maxSamples = 12; launch size 1920x1080, src/light_passes.rs:59-68).  A zeroed struct (or NULL)
 * reproduces the reference. */
typedef struct rt2_launch_params {
    uint32_t spp;          /* 0 => 12                                                         */
    uint32_t max_bounces;  /* 0 => 5                                                          */
    uint32_t band_rows;    /* interleaved row-band sharding: 0 => whole frame     */
    uint32_t band_rank;    /* this context renders bands b with b % band_world == band_rank    */
    uint32_t band_world;   /* 0 or 1 => whole frame                                            */
    uint32_t flags;        /* RT2_LAUNCH_*                                                     */
    uint32_t _reserved[2];
} rt2_launch_params;

#define RT2_LAUNCH_TRIG_LIBDEVICE 1u  /* use CUDA sinf/cosf/atan2f/asinf instead of the pinned polynomials */
#define RT2_LAUNCH_PRESAMPLING    4u  /* ReSTIR light sampler as compiled with RTXDI_ENABLE_PRESAMPLING 1 (the reference ships 0,
                                       * RtxdiApplicationBridge.glsl:12): local_light_sampling_mode 1 = POWER_RIS draws from the tiles
                                       * rt2_presample(…, 0) filled (InitialSamplingFunctions.hlsli:241-254); the environment light
                                       * rt2_prepare_lights writes at index lights + 1 for scenes with a sky (src/main.rs:381-386) exists:
                                       * num_primary_environment_samples candidates per pixel from the tiles rt2_presample(…, 1) filled
                                       * (:465-499, after rt2_generate_mips(1)), and BRDF rays that miss select it (:556-564) */
#define RT2_LAUNCH_PROFILE        2u  /* bracket every kernel launch with CUDA events (rt2_stats.trace_ns/shade_ns) */

/* Buffers readable with rt2_read (all row-major W x H unless noted). */
enum rt2_buffer {
    RT2_BUF_DIFFUSE_F16   = 0,  /* DiffuseLighting, rgba16f, 8 B/px (render_resources.rs:294-300)  */
    RT2_BUF_SPECULAR_F16  = 1,  /* SpecularLighting, rgba16f, 8 B/px (render_resources.rs:301-307) */
    RT2_BUF_DIFFUSE_F32   = 2,  /* fp32 shadow of DiffuseLighting (rgba32f, 16 B/px), SURVEY §7.3#7 */
    RT2_BUF_HIT_IDS       = 3,  /* rt2_hit_record per pixel                                        */
    RT2_BUF_OUTPUT_RGBA8  = 4,  /* post-process output, 4 B/px (replaces the swapchain image)      */
    RT2_BUF_LIGHT_INFO    = 5,  /* rt2_light_info x nLights (t_LightDataBuffer); with a sky: + a zero record and the environment light at nLights + 1 (src/main.rs:383) */
    RT2_BUF_GEOM_TO_LIGHT = 6,  /* uint32 x nGeometries (GeometryInstanceToLight)                  */
    RT2_BUF_LIGHT_TASKS   = 7,  /* rt2_prepare_lights_task x nTasks                                */
    RT2_BUF_LOCAL_PDF_MIP0 = 16, /* +k = mip k of the local-light PDF texture (r32f)               */
    RT2_BUF_ENV_PDF_MIP0   = 48, /* +k = mip k of the environment PDF texture (r16f)               */
    RT2_BUF_GBUFFER_DEPTH  = 80, /* r32f                                                           */
    RT2_BUF_GBUFFER_NORMALS = 81,
    RT2_BUF_GBUFFER_GEO_NORMALS = 82,
    RT2_BUF_GBUFFER_DIFFUSE_ALBEDO = 83,
    RT2_BUF_GBUFFER_SPECULAR_ROUGH = 84,
    RT2_BUF_GBUFFER_EMISSIVE = 85, /* rgba16f                                                      */
    RT2_BUF_MOTION_VECTORS  = 86,  /* rgba32f                                                      */
    RT2_BUF_DI_RESERVOIRS   = 87,  /* rt2_packed_di_reservoir x arrayPitch x 2, block-linear       */
    RT2_BUF_GI_RESERVOIRS   = 88,  /* 32 B x arrayPitch x 2                                        */
    RT2_BUF_SECONDARY_GBUFFER = 89, /* 48-byte rt2_secondary_gbuffer_data x arrayPitch, block-linear */
    RT2_BUF_NEIGHBOR_OFFSETS = 90,  /* K7 neighbour-offset table: RT2_NEIGHBOR_OFFSET_COUNT signed byte pairs */
    RT2_BUF_RIS = 91,               /* K11: uint2 per RIS entry (light index | COMPACT bit or packed uv, asuint(1/pdf)) */
    RT2_BUF_RIS_LIGHT_DATA = 92,    /* K11: two uint4 per RIS entry, the compact light record (PolymorphicLight.glsl:95-112) */
    /* acceleration-structure introspection (tests / tools) */
    RT2_BUF_BLAS_NODES = 100,      /* 80-byte CWBVH nodes                                          */
    RT2_BUF_BLAS_TRIS = 101,       /* 48-byte leaf triangles in slot order                         */
    RT2_BUF_BLAS_SLOT_TO_PRIM = 102, /* uint32 slot -> triangle index in geometry order            */
    RT2_BUF_TLAS_NODES = 103,
    RT2_BUF_TLAS_SLOT_TO_INST = 104
};

/* Counters of the last rt2_render / rt2_trace_primary on this context. */
typedef struct rt2_stats {
    uint64_t rays_traced;      /* every trace() issued (primary + bounce + shadow), SURVEY §8d */
    uint64_t rays_hit;
    uint64_t kernel_launches;  /* CUDA kernels launched by the last call                       */
    uint64_t wavefront_iterations;
    uint64_t bvh_nodes;        /* CWBVH nodes (BLAS + TLAS)                                    */
    uint64_t bvh_triangles;
    uint64_t device_bytes;     /* device memory owned by the context                          */
    uint64_t trace_ns;         /* with RT2_LAUNCH_PROFILE: summed CUDA-event time of the traversal kernel launches */
    uint64_t shade_ns;         /* with RT2_LAUNCH_PROFILE: summed CUDA-event time of the shading kernel launches   */
    uint64_t trace_launches;   /* with RT2_LAUNCH_PROFILE: traversal kernel launches that had rays to trace         */
    uint64_t sample_schedule;  /* K1: 0 = one launch pair per (sample, bounce); 1 = speculative sample parallelism (every sample its own path) */
    uint64_t rays_walked;      /* BVH walks behind rays_traced.  K1's primary ray has no jitter (refrence.rgen:26, GBufferHelpers.glsl:12-27), so the
                                  spp samples of a pixel issue the same trace() spp times: it is walked once and its hit record consumed spp
                                  times (RT2_PRIMARY_REUSE=0 walks every one).  K9 with both visibility tests on: a clear initial-visibility ray answers the
                                  final one (RT2_VISIBILITY_REUSE=0 walks both).  Equal to rays_traced everywhere else. */
} rt2_stats;

typedef struct rt2_ctx rt2_ctx;

/* ---- lifetime ---------------------------------------------------------------------------------
 * replaces Renderer::new (src/context.rs:120) + RenderResources::new (src/render_resources.rs:130)
 */
int  rt2_abi_version(void);
int  rt2_create(int cuda_device, rt2_ctx** out);
int  rt2_destroy(rt2_ctx* ctx);                       /* replaces Renderer::destroy (context.rs:762) */
/* Run on a caller-owned CUDA stream (cudaStream_t as void*; NULL => the context's own stream). */
int  rt2_set_stream(rt2_ctx* ctx, void* cuda_stream);
const char* rt2_last_error(rt2_ctx* ctx);             /* ctx may be NULL: last create error          */

/* ---- scene ------------------------------------------------------------------------------------
 * rt2_upload_scene replaces Model::from_gltf's uploads (src/model.rs:185-453) and
 * Image::new_from_data for the sky (src/main.rs:145).  rt2_build_accel replaces the driver BLAS +
 * TLAS builds (src/model.rs:454-459, src/main.rs:153-191, src/context.rs:824-911) with a device
 * CWBVH8 build. */
int  rt2_upload_scene(rt2_ctx* ctx, const rt2_scene_desc* scene);
int  rt2_build_accel(rt2_ctx* ctx);

/* ---- per-size resources: RenderResources::new (src/render_resources.rs:130-341) ---------------- */
int  rt2_resize(rt2_ctx* ctx, uint32_t width, uint32_t height);

/* ---- per-frame constants: LightPasses::update_uniform (src/light_passes.rs:665-668) ------------ */
int  rt2_set_constants(rt2_ctx* ctx, const void* gconst, size_t nbytes /* == RT2_GCONST_SIZE */);

/* ---- passes ----------------------------------------------------------------------------------- */
/* PrepareLightsTasks::execute (src/prepare_lights.rs:172-257) + prepare_lights.comp:40-126 (K10) */
int  rt2_prepare_lights(rt2_ctx* ctx);
/* GenerateMipsPass::execute_mip_generation (src/mip_pass.rs:122-186) + generate_mip_levels/mips.glsl
 * (K12).  which: 0 = local-light PDF texture, 1 = environment PDF texture. */
int  rt2_generate_mips(rt2_ctx* ctx, int which);
/* LightPasses::execute_presampeling (src/light_passes.rs:510-548) + lighting_passes/presample_{locallights,environment}.comp
 * (K11): 128 tiles x 1024 samples drawn through the PDF mip pyramid into the RIS buffer.  which: 0 = local lights (after
 * rt2_prepare_lights + rt2_generate_mips(0)), 1 = environment map (after rt2_generate_mips(1)).  `frame` seeds the sampler. */
int  rt2_presample(rt2_ctx* ctx, uint32_t frame, int which);
/* LightPasses::execute (src/light_passes.rs:550-663).  `frame` is the push constant
 * (light_passes.rs:581-585: low 32 bits of the u64 frame counter, first frame = 1).  Dispatches on
 * GConst.refrence_mode / enable_restir_di / enable_restir_gi exactly like the reference. */
int  rt2_render(rt2_ctx* ctx, uint32_t frame, const rt2_launch_params* params);
/* Primary visibility only (setupPrimaryRay + trace, GBufferHelpers.glsl:12-27, bridge :74-81,
 * with K1's TMin = 0.001): fills RT2_BUF_HIT_IDS.  Parity probe; no reference pass of its own. */
int  rt2_trace_primary(rt2_ctx* ctx, const rt2_launch_params* params);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).  Same result as rt2_trace_primary —
 * one rt2_hit_record per owned pixel, row-major W x H of the last rt2_resize — but written to CALLER-owned device memory
 * `device_hits_out` (`nbytes` must be W*H*32; rows of other band ranks are left untouched) and, if `device_hit_count_out` is
 * not NULL, the number of hits to that caller-owned device word.  The call allocates nothing (no cudaMalloc / cudaFree, no
 * host synchronisation): the kernel's scratch is a static device array inside the library, about 1 MiB per GPU, held by
 * every GPU that loads the library until the process exits.  Limit: at most 2^23 = 8 388 608 padded pixels
 * (ceil(W/8)*ceil(rows/4)*32 — 3840x2160 fits); larger frames return RT2_E_INVALID.  All calls on one device share the
 * scratch: calls on one stream are ordered and safe, calls whose kernels overlap on the GPU race.  rt2_get_stats is not
 * updated.  Runs on the context's stream like every other pass. */
int  rt2_trace_primary_noalloc(rt2_ctx* ctx, const rt2_launch_params* params, void* device_hits_out, size_t nbytes,
                               uint64_t* device_hit_count_out);
/* PostProcessPass::execute (src/postprocess.rs:225-307) + post_processing.comp:152-191 (K13) */
int  rt2_postprocess(rt2_ctx* ctx, uint32_t frame);

/* ---- results ---------------------------------------------------------------------------------- */
int  rt2_sync(rt2_ctx* ctx);
size_t rt2_buffer_size(rt2_ctx* ctx, int which);                       /* bytes, 0 if absent      */
int  rt2_read(rt2_ctx* ctx, int which, void* dst, size_t nbytes);      /* device -> host, syncs   */
/* Frame-in-flight read (the reference presents one swapchain image while it records the next, src/context.rs:30,
 * 639-694): snapshots the buffer on the context's stream and sends the snapshot to `dst` (pinned host memory for a
 * truly asynchronous transfer) on the context's copy stream, so the next rt2_render overlaps the PCIe transfer.
 * Up to two reads are in flight; `dst` is valid after rt2_sync.  rt2_read_fence makes later work on the context's
 * stream (e.g. a timing event) wait for the reads issued so far without blocking the host. */
int  rt2_read_async(rt2_ctx* ctx, int which, void* dst, size_t nbytes);
int  rt2_read_fence(rt2_ctx* ctx);
/* device -> device copy on the context's stream into caller-owned device memory (used to hand
 * row bands to NCCL without a host round trip). `row0,row1` select rows [row0,row1) of a W x H
 * buffer; pass 0,0 for the whole buffer. */
int  rt2_copy_to_device(rt2_ctx* ctx, int which, void* device_dst, size_t nbytes,
                        uint32_t row0, uint32_t row1);
/* all row bands of one rank (band b of `band_rows` rows belongs to rank b mod band_world, as in rt2_launch_params) packed
 * back to back into caller-owned device memory, in frame order — the send buffer of the NCCL band gather. */
int  rt2_pack_bands(rt2_ctx* ctx, int which, void* device_dst, size_t nbytes,
                    uint32_t band_rows, uint32_t band_rank, uint32_t band_world);
/* host -> device write into a context buffer (used to scatter gathered bands back, and by tests
 * to seed accumulation state). */
int  rt2_write(rt2_ctx* ctx, int which, const void* src, size_t nbytes);
int  rt2_get_stats(rt2_ctx* ctx, rt2_stats* out);

/* ---- multi-GPU: one frame over the GPUs of one node -----------------------------
 * The reference renders on ONE Vulkan device (src/context.rs:120, one queue, src/context.rs:639-694); the group generalises
 * its frame loop (src/main.rs:661-700) to N GPUs: the scene and acceleration structure are replicated, the frame is split into
 * interleaved bands of RT2_GROUP_BAND_ROWS rows (band b -> rank b mod N; the RNG is keyed on absolute pixel + frame,
 * src/shaders/Helpers.glsl:13-23, so the N-GPU image equals the 1-GPU image bit for bit), every rank renders its bands and
 * the library all-gathers them with NCCL over NVLink on its own stream and un-permutes them into an assembled W x H frame on
 * every rank.  Two ways to build a group:
 *   rt2_group_create       one process drives n devices (a Rust host needs no launcher);
 *   rt2_group_create_rank  one process per GPU (torchrun / MPI): rank 0 calls rt2_group_unique_id, the host distributes the
 *                          128 bytes by its own means, every rank passes them in.
 * NCCL (libnccl.so.2) is resolved with dlopen when a group is created, so the library loads — and single-GPU contexts work —
 * without it.  A group is not thread-safe; calls enqueue work and return. */
typedef struct rt2_group rt2_group;
#define RT2_GROUP_BAND_ROWS 8u
#define RT2_UNIQUE_ID_BYTES 128
int  rt2_group_create(const int* cuda_devices, int n, rt2_group** out);
int  rt2_group_unique_id(void* id128);
int  rt2_group_create_rank(int cuda_device, int rank, int world, const void* id128, rt2_group** out);
int  rt2_group_destroy(rt2_group* g);
const char* rt2_group_last_error(rt2_group* g);
int  rt2_group_world(rt2_group* g);                    /* ranks in the frame split                                  */
int  rt2_group_local_count(rt2_group* g);              /* contexts this process owns (n, or 1 per process)          */
rt2_ctx* rt2_group_ctx(rt2_group* g, int local_index); /* for per-context calls (rt2_get_stats, rt2_read of a buffer) */
/* Model::from_gltf + TLAS on every GPU (src/model.rs:185-476, src/main.rs:153-191) */
int  rt2_group_upload_scene(rt2_group* g, const rt2_scene_desc* scene);
int  rt2_group_build_accel(rt2_group* g);
int  rt2_group_resize(rt2_group* g, uint32_t width, uint32_t height);          /* RenderResources::new, src/render_resources.rs:130 */
int  rt2_group_set_constants(rt2_group* g, const void* gconst, size_t nbytes); /* LightPasses::update_uniform, src/light_passes.rs:665 */
int  rt2_group_prepare_lights(rt2_group* g);                                   /* replicated: K10 is tiny (src/prepare_lights.rs:172) */
int  rt2_group_generate_mips(rt2_group* g, int which);                         /* replicated: K12 (src/mip_pass.rs:122)              */
/* LightPasses::execute (src/light_passes.rs:550-663) on every rank's bands (params->band_* are filled in by the group), then,
 * if gather_which >= 0 (an rt2_buffer id of a W x H frame buffer, e.g. RT2_BUF_DIFFUSE_F32), pack + ncclAllGather +
 * un-permute of that buffer on the group's communication stream: the gather of frame n overlaps the render of frame n + 1.
 * With GI temporal / spatial resampling on (K6 / K7, src/light_passes.rs:642-647) the frame runs in phases and the reservoir page and
 * G-buffer those passes read at other ranks' pixels are exchanged in place between them (ncclAllGather on the render streams); the
 * frames stay bit-identical to one GPU's.  In a single-process group every local GPU's launches are issued by a host thread of its own. */
int  rt2_group_render(rt2_group* g, uint32_t frame, const rt2_launch_params* params, int gather_which);
/* the gather step alone (e.g. RT2_BUF_OUTPUT_RGBA8 after rt2_group_postprocess: sharded K13 + gather of RGBA8) */
int  rt2_group_gather(rt2_group* g, int which);
/* PostProcessPass::execute (src/postprocess.rs:225-307) on every rank (its own bands' pixels are the valid ones) */
int  rt2_group_postprocess(rt2_group* g, uint32_t frame);
/* the assembled frame of the last gathered rt2_group_render, from local context 0: blocking / frame-in-flight (see rt2_read_async) */
int  rt2_group_read(rt2_group* g, void* dst, size_t nbytes);
int  rt2_group_read_async(rt2_group* g, void* dst, size_t nbytes);
/* device pointer of the assembled frame on local context `local_index` (valid after rt2_group_sync / on the render stream after
 * rt2_group_wait_gather) */
void* rt2_group_assembled(rt2_group* g, int local_index, size_t* nbytes);
int  rt2_group_wait_gather(rt2_group* g);   /* later work on every context's stream waits for the gathers issued so far (no host block) */
int  rt2_group_sync(rt2_group* g);

#ifdef __cplusplus
}
#endif
#endif /* RT2_ABI_H */
