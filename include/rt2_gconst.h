/* rt2_gconst.h — typed C mirror of the reference's per-frame constant block.
 *
 * Field order, names and sizes follow src/shader_params.rs:3-274 (Rust `repr(C)`) which is the
 * same layout as the GLSL std140 block `ResamplingConstants` (src/shaders/ShaderParameters.glsl:
 * 25-92, rtxdi/ReSTIRGIParameters.h, rtxdi/ReSTIRDIParameters.h, rtxdi/RtxdiParameters.h).
 * rt2_set_constants takes these RT2_GCONST_SIZE bytes verbatim.
 */
#ifndef RT2_GCONST_H
#define RT2_GCONST_H
#include <stdint.h>
#include "rt2_abi.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rt2_planar_view_constants {   /* shader_params.rs:3-26 — 464 bytes */
    float mat_world_to_view[16];             /* column-major (glam) */
    float mat_view_to_clip[16];
    float mat_world_to_clip[16];
    float mat_clip_to_view[16];
    float mat_view_to_world[16];
    float mat_clip_to_world[16];
    float viewport_origin[2];
    float viewport_size[2];
    float viewport_size_inv[2];
    float pixel_offset[2];
    float clip_to_window_scale[2];
    float clip_to_window_bias[2];
    float window_to_clip_scale[2];
    float window_to_clip_bias[2];
    float camera_direction_or_position[4];
} rt2_planar_view_constants;

typedef struct rt2_rtxdi_runtime_parameters {          /* shader_params.rs:28-35 */
    uint32_t neighbor_offset_mask, active_checkerboard_field, pad1, pad2;
} rt2_rtxdi_runtime_parameters;

typedef struct rt2_restirgi_temporal_params {          /* shader_params.rs:37-55 */
    float depth_threshold, normal_threshold; uint32_t enable_permutation_sampling, max_history_length;
    uint32_t max_reservoir_age, enable_boiling_filter; float boiling_filter_strength; uint32_t enable_fallback_sampling;
    uint32_t temporal_bias_correction_mode, uniform_random_number, pad2, pad3;
} rt2_restirgi_temporal_params;

typedef struct rt2_restirgi_spatial_params {           /* shader_params.rs:57-70 */
    float spatial_depth_threshold, spatial_normal_threshold; uint32_t num_spatial_samples; float spatial_sampling_radius;
    uint32_t spatial_bias_correction_mode, pad1, pad2, pad3;
} rt2_restirgi_spatial_params;

typedef struct rt2_restirgi_final_shading_params {     /* shader_params.rs:72-79 */
    uint32_t enable_final_visibility, enable_final_mis, pad1, pad2;
} rt2_restirgi_final_shading_params;

typedef struct rt2_restirgi_buffer_indices {           /* shader_params.rs:81-93 */
    uint32_t secondary_surface_restirdi_output_buffer_index, temporal_resampling_input_buffer_index,
             temporal_resampling_output_buffer_index, spatial_resampling_input_buffer_index;
    uint32_t spatial_resampling_output_buffer_index, final_shading_input_buffer_index, pad1, pad2;
} rt2_restirgi_buffer_indices;

typedef struct rt2_reservoir_buffer_params {           /* shader_params.rs:95-102 */
    uint32_t reservoir_block_row_pitch, reservoir_array_pitch, pad1, pad2;
} rt2_reservoir_buffer_params;

typedef struct rt2_restirgi_params {                   /* shader_params.rs:104-112 — 144 bytes */
    rt2_reservoir_buffer_params reservoir_buffer_params;
    rt2_restirgi_buffer_indices buffer_indices;
    rt2_restirgi_temporal_params temporal_resampling_params;
    rt2_restirgi_spatial_params spatial_resampling_params;
    rt2_restirgi_final_shading_params final_shading_params;
} rt2_restirgi_params;

typedef struct rt2_light_buffer_region {               /* shader_params.rs:114-121 */
    uint32_t first_light_index, num_lights, pad1, pad2;
} rt2_light_buffer_region;

typedef struct rt2_environment_light_buffer_params {   /* shader_params.rs:123-130 */
    uint32_t light_present, light_index, pad1, pad2;
} rt2_environment_light_buffer_params;

typedef struct rt2_light_buffer_params {               /* shader_params.rs:132-138 — 48 bytes */
    rt2_light_buffer_region local_light_buffer_region;
    rt2_light_buffer_region infinite_light_buffer_region;
    rt2_environment_light_buffer_params environment_light_params;
} rt2_light_buffer_params;

typedef struct rt2_restirdi_initial_sampling_params {  /* shader_params.rs:140-152 */
    uint32_t num_primary_local_light_samples, num_primary_infinite_light_samples,
             num_primary_environment_samples, num_primary_brdf_samples;
    float brdf_cutoff; uint32_t enable_initial_visibility, environment_map_importance_sampling, local_light_sampling_mode;
} rt2_restirdi_initial_sampling_params;

typedef struct rt2_restirdi_buffer_indices {           /* shader_params.rs:154-166 */
    uint32_t initial_sampling_output_buffer_index, temporal_resampling_input_buffer_index,
             temporal_resampling_output_buffer_index, spatial_resampling_input_buffer_index;
    uint32_t spatial_resampling_output_buffer_index, shading_input_buffer_index, pad1, pad2;
} rt2_restirdi_buffer_indices;

typedef struct rt2_restirdi_temporal_params {          /* shader_params.rs:168-185 */
    float temporal_depth_threshold, temporal_normal_threshold; uint32_t max_history_length, temporal_bias_correction;
    uint32_t enable_permutation_sampling; float permutation_sampling_threshold; uint32_t enable_boiling_filter; float boiling_filter_strength;
    uint32_t discard_invisible_samples, uniform_random_number, pad2, pad3;
} rt2_restirdi_temporal_params;

typedef struct rt2_restirdi_spatial_params {           /* shader_params.rs:187-199 */
    float spatial_depth_threshold, spatial_normal_threshold; uint32_t spatial_bias_correction, num_spatial_samples;
    uint32_t num_disocclusion_boost_samples; float spatial_sampling_radius; uint32_t neighbor_offset_mask, discount_naive_samples;
} rt2_restirdi_spatial_params;

typedef struct rt2_restirdi_shading_params {           /* shader_params.rs:201-213 */
    uint32_t enable_final_visibility, reuse_final_visibility, final_visibility_max_age; float final_visibility_max_distance;
    uint32_t enable_denoiser_input_packing, pad1, pad2, pad3;
} rt2_restirdi_shading_params;

typedef struct rt2_restirdi_params {                   /* shader_params.rs:215-224 — 192 bytes */
    rt2_reservoir_buffer_params reservoir_buffer_params;
    rt2_restirdi_buffer_indices buffer_indices;
    rt2_restirdi_initial_sampling_params initial_sampling_params;
    rt2_restirdi_temporal_params temporal_resampling_params;
    rt2_restirdi_spatial_params spatial_resampling_params;
    rt2_restirdi_shading_params shading_params;
} rt2_restirdi_params;

typedef struct rt2_ris_buffer_segment_params {         /* shader_params.rs:226-233 */
    uint32_t buffer_offset, tile_size, tile_count, pad1;
} rt2_ris_buffer_segment_params;

typedef struct rt2_gconst {                            /* shader_params.rs:243-274 — 1424 bytes */
    rt2_planar_view_constants view;                    /* @0    */
    rt2_planar_view_constants prev_view;               /* @464  */
    rt2_rtxdi_runtime_parameters runtime_params;       /* @928  */
    uint32_t enable_brdf_indirect;                     /* @944  */
    uint32_t enable_brdf_additive_blend;
    uint32_t enable_accumulation;
    uint32_t frame;
    rt2_restirgi_params restir_gi;                     /* @960  */
    rt2_restirdi_params restir_di;                     /* @1104 */
    uint32_t enable_restir_di;                         /* @1296 */
    uint32_t enable_restir_gi;
    uint32_t refrence_mode;                            /* sic */
    uint32_t textures;
    float    blend_factor;                             /* @1312 */
    uint32_t enable_spatial_resampling;
    uint32_t enable_temporal_resampling;
    uint32_t environment;
    rt2_light_buffer_params light_buffer_params;       /* @1328 */
    rt2_ris_buffer_segment_params local_lights_risbuffer_segment_params;        /* @1376 */
    rt2_ris_buffer_segment_params environment_light_risbuffer_segment_params;   /* @1392 */
    uint32_t environment_pdf_texture_size[2];          /* @1408 */
    uint32_t local_light_pdf_texture_size[2];          /* @1416 */
} rt2_gconst;

#ifdef __cplusplus
}
static_assert(sizeof(rt2_planar_view_constants) == 464, "PlanarViewConstants");
static_assert(sizeof(rt2_restirgi_params) == 144, "ReSTIRGI_Parameters");
static_assert(sizeof(rt2_restirdi_params) == 192, "ReSTIRDI_Parameters");
static_assert(sizeof(rt2_gconst) == RT2_GCONST_SIZE, "GConst");
#endif
#endif /* RT2_GCONST_H */
