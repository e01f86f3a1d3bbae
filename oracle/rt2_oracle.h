/* rt2_oracle.h — C API of the CPU oracle (TEST INFRASTRUCTURE, never shipped, never on the product
 * path).  PARITY UNPINNED: the reference has no tests, golden vectors or CPU implementation of its
 * shader math and cannot be built or run here (no Rust, no glslang, no Vulkan).
 * This oracle is a hand restatement of the reference's GLSL, function by function, with every
 * function citing the reference file:line it follows.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load it.
 */
#ifndef RT2_ORACLE_H
#define RT2_ORACLE_H
#include "../include/rt2_abi.h"
#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_scene orc_scene;

/* flags for orc_render_* / orc_trace_primary */
#define ORC_BRUTE_FORCE   1u   /* intersect by testing every triangle (no BVH)                    */
#define ORC_PRESAMPLING   4u   /* light sampler as compiled with RTXDI_ENABLE_PRESAMPLING 1: POWER_RIS tiles of K11 */
#define ORC_TRIG_LIBM     2u   /* glibc sinf/cosf/atan2f/asinf instead of the pinned polynomials  */

orc_scene* orc_scene_create(const rt2_scene_desc* scene);
void       orc_scene_destroy(orc_scene* s);
uint64_t   orc_scene_num_triangles(const orc_scene* s);

/* setupPrimaryRay + trace with TMin 0.001 (refrence.rgen:26-27) for every pixel. */
/* arbitrary rays, rays[i] = {o.xyz, tmin, d.xyz, tmax} (test utility) */
int orc_trace_rays(const orc_scene* s, const float* rays, uint32_t n, uint32_t flags, rt2_hit_record* out);
int orc_trace_primary(const orc_scene* s, const void* gconst, uint32_t width, uint32_t height,
                      uint32_t flags, int threads, rt2_hit_record* out);

/* K1 refrence.rgen::main for rows [row0,row1) (row1 = 0 => height).  diffuse_f16/specular_f16 are
 * the rgba16f images (in/out: prior frame state when accumulation is on), diffuse_f32 the fp32
 * shadow accumulator (may be NULL).  Returns the number of rays traced through *rays. */
int orc_render_reference(const orc_scene* s, const void* gconst, uint32_t frame,
                         uint32_t width, uint32_t height, uint32_t spp, uint32_t max_bounces,
                         uint32_t row0, uint32_t row1, uint32_t flags, int threads,
                         uint16_t* diffuse_f16, uint16_t* specular_f16, float* diffuse_f32,
                         uint64_t* rays);

/* ---- known-answer probes for pure functions ------------------------------------------------- */
uint32_t orc_zcurve_to_linear(uint32_t x, uint32_t y);            /* RtxdiMath.hlsli:55-58 */
void     orc_linear_to_zcurve(uint32_t i, uint32_t* x, uint32_t* y); /* RtxdiMath.hlsli:61-66 */
uint32_t orc_jenkins_hash(uint32_t a);                            /* RtxdiMath.hlsli:69-79 */
/* initRandomSampler(pixel, frame + 13*pass) then n draws (Helpers.glsl:13-64) */
void     orc_rng_stream(uint32_t px, uint32_t py, uint32_t frame, uint32_t pass, uint32_t n,
                        uint32_t* raw, float* uniform);
uint16_t orc_f32_to_f16(float f);
float    orc_f16_to_f32(uint16_t h);
void     orc_sincos(float angle, uint32_t flags, float* s, float* c);
/* watertight ray/triangle test (pinned; Vulkan leaves it to the driver). returns 1 on hit */
int      orc_ray_triangle(const float o[3], const float d[3], float tmin, float tmax,
                          const float v0[3], const float v1[3], const float v2[3],
                          float* t, float* u, float* v);
/* setupPrimaryRay (GBufferHelpers.glsl:12-27) */
void     orc_primary_ray(const void* gconst, uint32_t px, uint32_t py, float origin[3], float dir[3]);
/* RAB_GetSurfaceBrdfSample's cosine lobe (bridge :437-461) for given normal + 3 uniforms */
void     orc_cos_hemisphere_dir(const float n[3], float r1, float r2, uint32_t flags, float dir[3]);
/* texture(textures[idx], uv).rgb with the pinned LOD-0 filter (Hit.glsl:35-37) */
void     orc_sample_texture(const orc_scene* s, uint32_t texture_index, float u, float v, float rgb[3]);

/* ---- K10 prepare_lights (src/prepare_lights.rs:172-257 + prepare_lights.comp:40-126) ----------
 * tasks / geom_to_light: capacity num_geometries; lights may be NULL to query counts only;
 * pdf_mip0 (r32f, pdf_w x pdf_h, caller zero-fills) receives getPower() at the Z-curve texel. */
int orc_prepare_lights(const orc_scene* s, rt2_prepare_lights_task* tasks, uint32_t* num_tasks, uint32_t* geom_to_light,
                       rt2_light_info* lights, uint32_t light_capacity, uint32_t* num_lights,
                       float* pdf_mip0, uint32_t pdf_w, uint32_t pdf_h);
/* The environment light record (PolymorphicLight.glsl:359-424) orc_prepare_lights writes at index num_lights + 1 when
 * light_capacity >= num_lights + 2 and the scene has a sky (SURVEY Q7 pinned to intent; read only with ORC_PRESAMPLING). */
void orc_environment_light_info(uint32_t tex_w, uint32_t tex_h, rt2_light_info* out);
/* ---- K12 mip_pass (src/mip_pass.rs:122-186 + generate_mip_levels/mips.glsl) ----------------------
 * packed = all mips back to back (mip k is max(1,w>>k) x max(1,h>>k)); which 0: r32f local-light
 * PDF (mip 0 is the input), which 1: r16f environment PDF (mip 0 is written from the sky). */
int orc_generate_mips(const orc_scene* s, int which, void* packed, uint32_t w, uint32_t h, uint32_t mips);
/* ---- K13 postprocess (post_processing.comp:152-191) -------------------------------------------- */
int orc_postprocess(const orc_scene* s, const void* gconst, uint32_t width, uint32_t height,
                    const uint16_t* diffuse_f16, const uint16_t* specular_f16, const float* gb_depth,
                    const uint32_t* gb_diffuse_albedo, const uint32_t* gb_specular_rough, const uint16_t* gb_emissive,
                    float* motion_vectors, uint8_t* out_rgba8);
/* codec / light KAT probes (packing.glsl, Helpers.glsl:251-332, PolymorphicLight.glsl:320-357) */
uint32_t orc_pack_r11g11b10(const float rgb[3]);
void     orc_unpack_r11g11b10(uint32_t p, float rgb[3]);
uint32_t orc_pack_rgba8_gamma(const float c[4]);
void     orc_unpack_rgba8_gamma(uint32_t p, float c[4]);
uint32_t orc_ndir_to_oct_unorm32(const float n[3]);
void     orc_oct_unorm32_to_ndir(uint32_t p, float n[3]);
void     orc_store_triangle_light(const float base[3], const float edge1[3], const float edge2[3], const float radiance[3],
                                  rt2_light_info* out, float* power);
void     orc_create_triangle_light(const rt2_light_info* li, float base[3], float edge1[3], float edge2[3], float radiance[3],
                                   float normal[3], float* area);
uint32_t orc_tonemap(const float rgb[3]);

/* ---- K3 g_buffer (lighting_passes/g_buffer.rgen:14-47); all images row-major W x H --------------- */
int orc_gbuffer(const orc_scene* s, const void* gconst, uint32_t width, uint32_t height, uint32_t flags, int threads,
                float* depth, uint32_t* normals, uint32_t* geo_normals, uint32_t* diffuse_albedo, uint32_t* specular_rough,
                uint16_t* emissive_f16, float* motion_vectors, uint64_t* rays);
/* ---- K9 di_fused_resampling (lighting_passes/di_fused_resampling.rgen:16-93): reads the G-buffer, the
 * light list (K10) and the local-light PDF pyramid (K12, all mips packed), writes the packed DI
 * reservoirs (block-linear, arrayPitch x 2) and the rgba16f lighting images (in/out). */
int orc_di_fused_resampling(const orc_scene* s, const void* gconst, uint32_t frame, uint32_t width, uint32_t height, uint32_t flags, int threads,
                            const float* gb_depth, const uint32_t* gb_normals, const uint32_t* gb_geo_normals, const uint32_t* gb_diffuse_albedo,
                            const uint32_t* gb_specular_rough, const rt2_light_info* lights, uint32_t num_lights, const uint32_t* geom_to_light,
                            const float* pdf_packed, uint32_t pdf_w, uint32_t pdf_h, uint32_t pdf_mips,
                            rt2_packed_di_reservoir* reservoirs, uint16_t* diffuse_f16, uint16_t* specular_f16, uint64_t* rays,
                            const uint32_t* ris, const uint32_t* ris_light_data, uint64_t ris_entries,
                            const void* env_pdf_packed, uint32_t env_w, uint32_t env_h, uint32_t env_mips);
/* RTXDI_ReservoirPositionToPointer (rtxdi/RtxdiHelpers.hlsli:75-88) */
uint32_t orc_reservoir_position_to_pointer(uint32_t block_row_pitch, uint32_t array_pitch, uint32_t x, uint32_t y, uint32_t array_index);

/* ---- the default ReSTIR-GI chain (src/light_passes.rs:621-659) -------------------------------
 * G-buffer inputs as for K9.  `secondary` (arrayPitch entries, block-linear) and `gi_reservoirs` (arrayPitch x 2)
 * persist between the passes and between frames, so they are in/out.  Light inputs may be null when
 * enableRestirDI == 0 (K5 then traces one more cosine bounce instead of sampling the light list). */
typedef struct orc_gbuffer_in {
    const float* depth; const uint32_t* normals; const uint32_t* geo_normals; const uint32_t* diffuse_albedo; const uint32_t* specular_rough;
} orc_gbuffer_in;
typedef struct orc_light_in {
    const rt2_light_info* lights; uint32_t num_lights; const uint32_t* geom_to_light;
    const float* pdf_packed; uint32_t pdf_w, pdf_h, pdf_mips;
    const uint32_t* ris; const uint32_t* ris_light_data; uint64_t ris_entries;   /* K11 outputs, read with ORC_PRESAMPLING + POWER_RIS */
    const void* env_pdf_packed; uint32_t env_w, env_h, env_mips;                  /* K12's r16f environment pyramid (ORC_PRESAMPLING + an environment light) */
} orc_light_in;
/* K4 lighting_passes/brdf_rays.rgen:19-195 */
int orc_brdf_rays(const orc_scene* s, const void* gconst, uint32_t frame, uint32_t width, uint32_t height, uint32_t flags, int threads,
                  const orc_gbuffer_in* gb, rt2_secondary_gbuffer_data* secondary, uint16_t* diffuse_f16, uint16_t* specular_f16, uint64_t* rays);
/* K5 lighting_passes/shade_secondary_surfaces.rgen:26-157 */
int orc_shade_secondary_surfaces(const orc_scene* s, const void* gconst, uint32_t frame, uint32_t width, uint32_t height, uint32_t flags, int threads,
                                 const orc_gbuffer_in* gb, const orc_light_in* li, rt2_secondary_gbuffer_data* secondary,
                                 rt2_packed_gi_reservoir* gi_reservoirs, uint16_t* diffuse_f16, uint16_t* specular_f16, uint64_t* rays);
/* K8 lighting_passes/gi_final_shading.rgen:43-101 */
int orc_gi_final_shading(const orc_scene* s, const void* gconst, uint32_t width, uint32_t height, uint32_t flags, int threads,
                         const orc_gbuffer_in* gb, const rt2_secondary_gbuffer_data* secondary, const rt2_packed_gi_reservoir* gi_reservoirs,
                         uint16_t* diffuse_f16, uint16_t* specular_f16, uint64_t* rays);
/* K6 lighting_passes/temporal_resampling.rgen + rtxdi/GIResamplingFunctions.hlsli:186-359.  gb_prev = last frame's G-buffer
 * (interpreted through prevView); motion_vectors = rgba32f of K3.  Reads page temporalResamplingInputBufferIndex at
 * reprojected neighbours, writes page temporalResamplingOutputBufferIndex. */
int orc_gi_temporal_resampling(const orc_scene* s, const void* gconst, uint32_t frame, uint32_t width, uint32_t height, uint32_t flags, int threads,
                               const orc_gbuffer_in* gb, const orc_gbuffer_in* gb_prev, const float* motion_vectors,
                               rt2_packed_gi_reservoir* gi_reservoirs, uint64_t* rays);
/* K7 lighting_passes/spatial_resampling.rgen + rtxdi/GIResamplingFunctions.hlsli:391-553.  neighbor_offsets: signed byte
 * pairs (SURVEY Q11 pin), neighbor_offset_count entries. */
int orc_gi_spatial_resampling(const orc_scene* s, const void* gconst, uint32_t frame, uint32_t width, uint32_t height, uint32_t flags, int threads,
                              const orc_gbuffer_in* gb, const int8_t* neighbor_offsets, uint32_t neighbor_offset_count,
                              rt2_packed_gi_reservoir* gi_reservoirs, uint64_t* rays);
/* fill_neighbor_offset_buffer (src/light_passes.rs:671-698) with signed bytes; xy holds 2 * count bytes */
void orc_fill_neighbor_offset_buffer(uint32_t neighbor_offset_count, int8_t* xy);
/* K11 lighting_passes/presample_{locallights,environment}.comp + rtxdi/PresamplingFunctions.hlsli:30-162.  which 0: local
 * lights (pdf_packed = r32f pyramid), 1: environment (r16f pyramid).  ris: uint2 per entry, both segments; ris_light_data:
 * 2 x uvec4 per entry (local lights only).  Writes the reference's fixed 128 x 1024 grid. */
int orc_presample(const void* gconst, uint32_t frame, int which, const void* pdf_packed, uint32_t w, uint32_t h, uint32_t mips,
                  const rt2_light_info* lights, uint32_t num_lights, uint32_t* ris, uint32_t* ris_light_data);
/* KAT probes: rtxdi/RtxdiMath.hlsli:129-296 */
uint32_t orc_encode_logluv(const float rgb[3]);
void orc_decode_logluv(uint32_t packed, float rgb[3]);
uint32_t orc_encode_snorm_oct(const float n[3]);
void orc_decode_snorm_oct(uint32_t packed, float n[3]);

#ifdef __cplusplus
}
#endif
#endif
