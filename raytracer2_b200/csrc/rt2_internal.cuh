// rt2_internal.cuh — context object and cross-file declarations of librt2cuda.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include "../../include/rt2_abi.h"
#include "../../include/rt2_gconst.h"
#include "rt2_bvh.cuh"

namespace rt2 {

struct LaunchCounter {
    uint64_t n = 0;
    void add(uint64_t k) { n += k; }
};

// One acceleration structure level.
struct Accel {
    Node8* nodes = nullptr;
    TriPos* tris = nullptr;             // BLAS only: leaf triangles in slot order
    uint32_t* slot_to_prim = nullptr;   // slot -> primitive id (BLAS: triangle in geometry order; TLAS: instance id)
    uint32_t num_nodes = 0, num_prims = 0, depth = 0;
    float box_lo[3] = {0, 0, 0}, box_hi[3] = {0, 0, 0};
    void release() {
        if (nodes) cudaFree(nodes);
        if (tris) cudaFree(tris);
        if (slot_to_prim) cudaFree(slot_to_prim);
        nodes = nullptr; tris = nullptr; slot_to_prim = nullptr; num_nodes = num_prims = depth = 0;
    }
};

// Per-triangle shading record in leaf (slot) order: what GetGeometryFromHit (Hit.glsl:2-42) gathers
// through Indices[] and Vertices[] — same fp32 values, one indirection instead of three. 96 bytes.
struct __align__(16) TriShade {
    float n0[3], n1[3], n2[3];
    float uv0[2], uv1[2], uv2[2];
    float c0[3], c1[3], c2[3];
};
static_assert(sizeof(TriShade) == 96, "TriShade must be 96 bytes");

// Per-geometry shading record (from GeometryInfo, src/model.rs:12-23). 112 bytes.
struct __align__(16) GeomShade {
    float transform[16];      // column-major; used for normalize(transform * vec4(n, 0)) (Hit.glsl:27)
    float base_color[3];
    int32_t texture;          // baseColorTextureIndex
    float emission12[3];      // emission.xyz * 12.0 (Hit.glsl:41) — one fp32 multiply, done at upload
    float metallic;
    float emission[3];        // raw emission (prepare_lights)
    float roughness;
};
static_assert(sizeof(GeomShade) == 112, "GeomShade must be 112 bytes");

struct TextureRec { uint32_t offset; uint32_t width, height, filter; };   // offset in texels into the texel pool

struct DeviceScene {
    rt2_vertex* vertices = nullptr; uint64_t num_vertices = 0;
    uint32_t* indices = nullptr; uint64_t num_indices = 0;
    rt2_geometry_info* geoms = nullptr; uint32_t num_geoms = 0;
    uint32_t* geom_tri_offset = nullptr;      // num_geoms + 1
    GeomShade* geom_shade = nullptr;
    uint32_t* texels = nullptr;               // RGBA8 pool
    TextureRec* textures = nullptr; uint32_t num_textures = 0;
    float* srgb_lut = nullptr;                // 256 floats
    float4* sky = nullptr; uint32_t sky_w = 0, sky_h = 0;
    InstanceRec* instances = nullptr; uint32_t num_instances = 0;   // 0 => single identity instance
    uint32_t num_tris = 0;
    uint32_t num_lights = 0;
    TriShade* tri_shade = nullptr;            // slot order (after build)
    Accel blas, tlas;
    Accel flat; FlatLeaf* flat_leaf = nullptr;   // flattened instancing: one CWBVH over every (instance, triangle) pair (nodes + leaf records); empty when not built
    bool accel_built = false;
    // host copies needed by host-side steps (prepare_lights scan, src/prepare_lights.rs:182-201)
    std::vector<rt2_geometry_info> h_geoms;
    std::vector<uint32_t> h_index_counts;
};

// Per-size resources (RenderResources::new, src/render_resources.rs:130-341)
struct FrameBuffers {
    uint32_t width = 0, height = 0;
    ushort4* diffuse_f16 = nullptr;    // DiffuseLighting rgba16f
    ushort4* specular_f16 = nullptr;   // SpecularLighting rgba16f
    float4* diffuse_f32 = nullptr;     // fp32 shadow accumulator
    rt2_hit_record* hit_ids = nullptr;
    uchar4* output_rgba8 = nullptr;
    // wavefront state (one slot per owned pixel)
    uint32_t max_slots = 0;
    float4 *st_ray0 = nullptr, *st_ray1 = nullptr, *st_acc = nullptr;
    uint4* st_misc = nullptr;
    float4* st_hit = nullptr;
    uint32_t* st_hit_inst = nullptr;
    float4* st_hit0 = nullptr;         // K1 primary-hit reuse: the hit record of the pixel's primary ray (k1_shade_step)
    uint32_t* st_hit_inst0 = nullptr;
    uint32_t *queue0 = nullptr, *queue1 = nullptr;
    // speculative sample parallelism (render_reference_spec): one path per (pixel, sample); allocated on first use, grow-only
    struct Spec {
        float4 *ray0 = nullptr, *ray1 = nullptr, *hit = nullptr, *contrib = nullptr, *px_rad = nullptr;
        uint4* misc = nullptr; uint32_t *hit_inst = nullptr, *queue0 = nullptr, *queue1 = nullptr, *pending0 = nullptr, *pending1 = nullptr, *counters = nullptr, *start_index = nullptr;
        uint16_t* nhits = nullptr; uint2* px_prog = nullptr;
        size_t cap_paths = 0, cap_contrib = 0, cap_pixels = 0, cap_counters = 0;
    } spec;
    uint32_t* warp_pop = nullptr; size_t warp_pop_cap = 0;   // k_path_warp: warp-private slot queues (num_warps x P)
    // ReSTIR-side per-size resources (RenderResources::new, src/render_resources.rs:48-101,191-206,287-317); allocated on first use
    float* gb_depth[2] = {nullptr, nullptr};             // r32f, ping-pong by frame % 2 (src/light_passes.rs:565-577)
    uint32_t* gb_normals[2] = {nullptr, nullptr};        // r32ui oct
    uint32_t* gb_geo_normals[2] = {nullptr, nullptr};
    uint32_t* gb_diffuse_albedo[2] = {nullptr, nullptr}; // R11G11B10
    uint32_t* gb_specular_rough[2] = {nullptr, nullptr}; // RGBA8 gamma
    ushort4* gb_emissive[2] = {nullptr, nullptr};        // rgba16f
    float4* motion_vectors = nullptr;                    // rgba32f
    rt2_packed_di_reservoir* di_reservoirs = nullptr;    // arrayPitch x 2
    uint4* gi_reservoirs = nullptr;                      // 32 B x arrayPitch x 2
    uint4* secondary_gbuffer = nullptr;                  // 48 B x arrayPitch
    uint4* di_state = nullptr;                           // 8 planes x W x H: sampler state parked between the wavefront stages of K9
    char2* neighbor_offsets = nullptr;                   // RT2_NEIGHBOR_OFFSET_COUNT signed byte pairs (K7)
    uint32_t reservoir_block_row_pitch = 0, reservoir_array_pitch = 0;
    uint32_t* counters = nullptr;      // [0..1] u64 rays, [2..3] u64 hits, [4..5] u64 walks (K1), then (count, head) per wavefront iteration
    size_t counter_capacity = 0;
};

}  // namespace rt2

#define RT2_FRAMES_IN_FLIGHT 2
struct rt2_ctx {
    int device = 0;
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr;
    int num_sms = 0;
    std::string last_error;
    rt2::DeviceScene scene;
    rt2::FrameBuffers fb;
    rt2_gconst gconst{};
    bool have_gconst = false;
    rt2_gconst* d_gconst = nullptr;
    // frames in flight (the reference keeps swapchain images + one command buffer per image in flight, src/context.rs:30,639-694):
    // a ring of pinned staging copies for rt2_set_constants, and rt2_read_async's device snapshots + copy stream
    rt2_gconst* h_gconst_ring = nullptr; cudaEvent_t gconst_done[RT2_FRAMES_IN_FLIGHT] = {}; uint32_t gconst_next = 0;
    cudaStream_t copy_stream = nullptr;
    void* snap[RT2_FRAMES_IN_FLIGHT] = {}; size_t snap_bytes[RT2_FRAMES_IN_FLIGHT] = {};
    cudaEvent_t snap_ready[RT2_FRAMES_IN_FLIGHT] = {}, read_done[RT2_FRAMES_IN_FLIGHT] = {}; uint32_t snap_next = 0; bool reads_pending = false;
    rt2_stats stats{};
    bool stats_from_counters = false;   // rays/hits of the last render still live in fb.counters
    bool stats_walks_counted = false;   // ... and fb.counters[4..5] holds the walks (K1); otherwise rays_walked = rays_traced
    bool stats_reused_counted = false;  // ... and fb.counters[16..17] holds trace() calls answered without a walk (ReSTIR frames): rays_walked = rays_traced - that
    bool sparse_contrib = true;         // speculative K1: contributions that are +0 are neither stored nor added (RT2_SPARSE_CONTRIB=0: all of them)
    bool any_hit = true;                // visibility-ray batches of the ReSTIR passes end a walk at its first hit (RT2_ANY_HIT=0: closest hit always)
    bool visibility_reuse = true;       // K9: a clear initial-visibility ray answers the final one (RT2_VISIBILITY_REUSE=0: both are walked)
    bool primary_reuse = true;          // K1: walk a pixel's primary ray once for all its samples (RT2_PRIMARY_REUSE=0: every sample walks its own)
    uint64_t pending_rays_static = 0;   // rays known on the host (primary probe)
    int trace_blocks_per_sm[4] = {1, 1, 1, 1};          // [0] single-level queue kernel, [1] unified two-level, [2] nested two-level, [3] flattened instancing
    int trace_queue_blocks_per_sm = 1;       // resident blocks of k_trace_queue (the ReSTIR passes' ray-queue kernel)
    bool use_flat = true;                                // trace instanced scenes through the flattened tree when it was built (RT2_USE_FLAT=0: two-level walk)
    int trace_variant = 0;
    // K1 sample schedule: speculate = -1 auto (from the previous frame's miss statistics, read back without a sync), 0 never, 1 whenever
    // the paths fit (RT2_SPECULATE); spec_rounds = prediction fix-up rounds before the candidate round (RT2_SPEC_ROUNDS)
    int speculate = -1; bool spec_auto = false, spec_auto_few_misses = false; int spec_rounds = -1;     // spec_rounds < 0: 1, or 0 after a frame with < 0.1 % missed samples
    unsigned long long* h_k1_stats = nullptr; cudaEvent_t k1_stats_ready = nullptr; bool k1_stats_pending = false; uint64_t k1_stats_paths = 0;
    int path_mode = 0;                                   // K1 schedule: 0 = one launch per bounce (wavefront chains), 1 = warp-local wavefronts in one persistent launch (RT2_PATH_MODE)
    int path_blocks_per_sm[3] = {1, 1, 1}; int warp_blocks_per_sm = 0; uint32_t warp_paths = 0;   // k_path_warp occupancy; RT2_WARP_BLOCKS / RT2_WARP_P overrides
    uint32_t rays_per_lane = 0; bool rays_per_lane_set = false;   // cap of the persistent trace grid for small wavefronts (RT2_RAYS_PER_LANE overrides the rule)
    uint32_t chains = 0;                                 // wavefront chains of the path tracer; 0 = by frame size (RT2_CHAINS overrides)
    cudaStream_t chain_streams[8] = {}; cudaEvent_t chain_join[8] = {}; cudaEvent_t chain_fork = nullptr;
    bool gi_megakernels = false;                         // RT2_GI_MEGAKERNEL=1: K4/K5 with inlined traversal instead of the ray-queue kernel (A/B)
    std::vector<cudaEvent_t> prof_events;   // RT2_LAUNCH_PROFILE: 3 per wavefront iteration (before trace, after trace, after shade)
    uint64_t device_bytes = 0;
    // lights / pdf textures (K10/K12)
    rt2_light_info* light_info = nullptr; uint32_t light_capacity = 0;   // = the scene's light count
    float* codec_luts = nullptr;                                           // [0,256): gamma22_decode(byte); [256, 256 + 65536): unpack_light_radiance(code); then 16384 x logluv_luminance(exponent) — built on this device at rt2_create
    uint32_t post_band[3] = {0, 0, 0};                                     // set by rt2_group_postprocess: {band_rows, rank, world}; K13 then visits this rank's rows only
    uint32_t restir_phase = 0;                                             // set by rt2_group_render around a sharded frame with K6 / K7: which part of render_restir runs (0 = all)
    uint32_t light_records = 0;                                            // records allocated: light_capacity, + 2 with a sky (a zero record and the environment light)
    uint32_t* geom_to_light = nullptr;
    rt2_prepare_lights_task* light_tasks = nullptr; uint32_t num_light_tasks = 0;
    bool lights_prepared = false;
    uint32_t last_gbuffer_index = 0;     // frame % 2 of the last ReSTIR-mode render: which G-buffer set rt2_read returns
    float* local_pdf = nullptr; uint32_t local_pdf_w = 0, local_pdf_h = 0, local_pdf_mips = 0;   // all mips, packed
    uint2* ris_buffer = nullptr; uint4* ris_light_data = nullptr; uint64_t ris_entries = 0; uint32_t* presample_error = nullptr;   // K11 outputs
    uint16_t* env_pdf = nullptr; uint32_t env_pdf_w = 0, env_pdf_h = 0, env_pdf_mips = 0;        // r16f, packed mips
};

namespace rt2 {

int fail(rt2_ctx* ctx, int code, const std::string& msg);
#define RT2_MAX_CHAINS 8u
#define RT2_CUDA(ctx, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return rt2::fail((ctx), RT2_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)

// rt2_build.cu
struct BuildInput;
int build_blas(cudaStream_t st, const BuildInput& in, Accel* out, LaunchCounter* lc, std::string& err);
int build_cwbvh_from_boxes(cudaStream_t st, const float4* lo, const float4* hi, uint32_t n, Node8** nodes_out, uint32_t* num_nodes_out,
                           uint32_t** out_prim_out, uint32_t* depth_out, LaunchCounter* lc, std::string& err, uint32_t ploc_radius = 16u /* RT2_PLOC_RADIUS: instance-level boxes */);
// rt2_scene.cu
int scene_upload(rt2_ctx* ctx, const rt2_scene_desc* d);
int scene_build_accel(rt2_ctx* ctx);
void scene_release(rt2_ctx* ctx);
// rt2_render.cu
int frame_resize(rt2_ctx* ctx, uint32_t w, uint32_t h);
void frame_release(rt2_ctx* ctx);
int render_reference(rt2_ctx* ctx, uint32_t frame, const rt2_launch_params& lp);   // K1
int render_primary(rt2_ctx* ctx, const rt2_launch_params& lp);
int render_init(rt2_ctx* ctx);
// rt2_noalloc.cu
int codec_luts_init(rt2_ctx* ctx);                 // rt2_passes_common.cu: tabulates the two byte/half-word decoders that cost a binary64 transcendental each
int noalloc_init(rt2_ctx* ctx);                    // resolves the static scratch's device address, once per device (cudaGetSymbolAddress)

}  // namespace rt2
