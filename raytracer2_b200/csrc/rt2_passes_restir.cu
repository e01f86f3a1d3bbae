// rt2_passes_restir.cu — the non-reference branch of LightPasses::execute (src/light_passes.rs:598-659):
//   K3 g_buffer            lighting_passes/g_buffer.rgen:14-47
//   K9 di_fused_resampling lighting_passes/di_fused_resampling.rgen:16-93 (reservoir light sampling)
//   K4 brdf_rays / K5 shade_secondary_surfaces / K8 gi_final_shading (the default ReSTIR-GI chain)
//
// Both are one-thread-per-pixel kernels with the traversal inlined (8x4 pixel tiles per warp so the
// primary / BRDF rays of a warp start coherent).  K3 is bound like the primary wavefront of K1
// (instruction issue in traversal, 44 B written per pixel); K9 evaluates numLocal light candidates per
// pixel out of a 48 B/light list that sits in L2 (10k lights = 480 KB) and traces <= numBrdf + 2 rays.
#include "rt2_restir.cuh"

namespace rt2 {

int restir_ensure_resources(rt2_ctx* ctx);
int lights_ensure_resources(rt2_ctx* ctx);
int trace_queue_generic(rt2_ctx* ctx, const uint32_t* queue, const uint32_t* count, uint32_t* head, unsigned long long* stat_rays, bool any_hit = false);

// interleaved row bands (the same rule as rt2_render.cu's Band): the passes below are per-pixel independent, every buffer
// stays full-frame and addressed by absolute pixel, a rank just visits its own rows
struct BandRef { uint32_t band_rows, rank, world, local_rows; };
struct GBufferArgs {
    SceneRefs sc; const rt2_gconst* g; uint32_t width, height; BandRef band;
    float* depth; uint32_t* normals; uint32_t* geo_normals; uint32_t* diffuse_albedo; uint32_t* specular_rough; ushort4* emissive; float4* motion;
    unsigned long long* stat_hits;
};

// getMotionVector, GBufferHelpers.glsl:29-52
__device__ __forceinline__ V3 get_motion_vector(const rt2_planar_view_constants& view, const rt2_planar_view_constants& prev, V3 pos, V3 prev_pos) {
    float c[4], p[4];
    mat4_mul_vec4(view.mat_world_to_clip, pos.x, pos.y, pos.z, 1.0f, c);
    c[0] /= c[3]; c[1] /= c[3]; c[2] /= c[3];
    mat4_mul_vec4(prev.mat_world_to_clip, prev_pos.x, prev_pos.y, prev_pos.z, 1.0f, p);
    p[0] /= p[3]; p[1] /= p[3]; p[2] /= p[3];
    if (c[3] <= 0 || p[3] <= 0) return v3(0.f, 0.f, 0.f);
    V3 m;
    m.x = (p[0] - c[0]) / 2.0f * view.viewport_size[0];
    m.y = (p[1] - c[1]) / 2.0f * view.viewport_size[1];
    m.x += (view.pixel_offset[0] - prev.pixel_offset[0]);
    m.y += (view.pixel_offset[1] - prev.pixel_offset[1]);
    m.z = p[3] - c[3];
    return m;
}

__device__ __forceinline__ bool tile_pixel(uint32_t i, uint32_t width, const BandRef& b, uint32_t* x, uint32_t* y) {
    const uint32_t tiles_x = (width + 7u) >> 3;
    const uint32_t tile = i >> 5, within = i & 31u;
    const uint32_t ly = (tile / tiles_x) * 4u + (within >> 3);
    *x = (tile % tiles_x) * 8u + (within & 7u);
    *y = b.world <= 1u ? ly : ((ly / b.band_rows) * b.world + b.rank) * b.band_rows + (ly % b.band_rows);
    return *x < width && ly < b.local_rows;
}

// Register caps measured on the ReSTIR frames (profiles/r01d_ab_gi_regs.txt): K9 stages a/b at 8 blocks/SM (64 registers, DI-only frame
// 3.14 -> 3.01 ms), K6/K7 at 6 blocks/SM (80 registers, GI + temporal + spatial 3.91 -> 3.79 ms).
#ifndef RT2_DI_BLOCKS
#define RT2_DI_BLOCKS 8
#endif
#ifndef RT2_GIRS_BLOCKS
#define RT2_GIRS_BLOCKS 6
#endif
// 6 blocks/SM (80 registers instead of 96): default ReSTIR-GI frame 2.91 -> 2.84 ms together with K8 (profiles/r01d_ab_gi_regs.txt)
#ifndef RT2_GBUF_BLOCKS
#define RT2_GBUF_BLOCKS 6
#endif
__global__ void __launch_bounds__(128, RT2_GBUF_BLOCKS) k_gbuffer(GBufferArgs a) {
    __shared__ float lut[256];
    lut[threadIdx.x] = a.sc.srgb_lut[threadIdx.x]; lut[threadIdx.x + 128] = a.sc.srgb_lut[threadIdx.x + 128];
    __syncthreads();
    uint32_t x, y;
    if (!tile_pixel(blockIdx.x * blockDim.x + threadIdx.x, a.width, a.band, &x, &y)) return;
    const size_t i = (size_t)y * a.width + x;
    const rt2_gconst& g = *a.g;
    V3 o, d;
    setup_primary_ray(g.view, x, y, &o, &d);
    uint32_t tri_slot;
    const Hit h = trace_closest(a.sc, o, d, 0.0f, RT2_BACKGROUND_DEPTH, &tri_slot);      // TMin = 0 (GBufferHelpers.glsl:24)
    if (h.geom != 0xFFFFFFFFu) {
        const SurfaceFetch sf = get_geometry_from_hit(a.sc, lut, tri_slot, h.inst, h.u, h.v);
        const V3 world_pos = o + d * h.t;
        const V3 mv = get_motion_vector(g.view, g.prev_view, world_pos, world_pos);
        const uint32_t n = ndir_to_oct_unorm32(sf.normal);
        a.depth[i] = h.t; a.normals[i] = n; a.geo_normals[i] = n;
        a.diffuse_albedo[i] = pack_r11g11b10(sf.color);
        a.specular_rough[i] = pack_rgba8_gamma(sf.specular_f0.x, sf.specular_f0.y, sf.specular_f0.z, sf.roughness);
        a.emissive[i] = make_ushort4(f32_to_f16_bits(sf.emission.x), f32_to_f16_bits(sf.emission.y), f32_to_f16_bits(sf.emission.z), 0);
        a.motion[i] = make_float4(mv.x, mv.y, mv.z, 0.0f);
        atomicAdd(a.stat_hits, 1ull);
    } else {
        a.depth[i] = RT2_BACKGROUND_DEPTH; a.normals[i] = 0u; a.geo_normals[i] = 0u; a.diffuse_albedo[i] = 0u; a.specular_rough[i] = 0u;
        a.emissive[i] = make_ushort4(0, 0, 0, 0); a.motion[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
}

// ------------------------------------------------------------------------------------------------
// RTXDI_SampleLightsForSurface (rtxdi/InitialSamplingFunctions.hlsli:594-664) — shared by K9 and K5
// ------------------------------------------------------------------------------------------------
struct LightRefs {
    const rt2_light_info* lights; const uint32_t* geom_to_light;
    const float* pdf; uint32_t pdf_w, pdf_h, pdf_last_offset;
    // RT2_LAUNCH_PRESAMPLING: the sampler as compiled with RTXDI_ENABLE_PRESAMPLING 1; K11's tiles for POWER_RIS
    uint32_t presampling; const uint2* ris; const uint4* ris_light_data;
    // the environment light (RT2_LAUNCH_PRESAMPLING only; SURVEY Q7 pinned to intent): present when environmentLightParams.lightPresent
    // is set AND the scene has a sky; env_pdf = K12's r16f pyramid, env_last_offset = the texel offset of the mip the pdf's average is read from
    uint32_t env_present, env_index; const uint16_t* env_pdf; uint32_t env_w, env_h, env_last_offset;
};

// RAB_EvaluateLocalLightSourcePdf, bridge :420-434 (Q14 pinned to texelFetch semantics)
__device__ __forceinline__ float local_light_source_pdf(const LightRefs& a, uint32_t light_index) {
    uint32_t x, y;
    linear_to_zcurve(light_index, x, y);
    const float texel = (x < a.pdf_w && y < a.pdf_h) ? __ldg(a.pdf + (size_t)y * a.pdf_w + x) : 0.f;
    const int last = max(0, (int)floorf(log2_pinned((float)max(a.pdf_w, a.pdf_h))));
    const float average = __ldg(a.pdf + a.pdf_last_offset);
    const float sum = average * (float)(1u << (uint32_t)last << (uint32_t)last);
    return texel / sum;
}

// RAB_EvaluateEnvironmentMapSamplingPdf, bridge :397-418 (the Q14 quirk again, same pin: texelFetch; uv == 1.0 reads 0)
__device__ __forceinline__ float environment_map_sampling_pdf(const LightRefs& a, const rt2_gconst& g, V3 L) {
    if (g.restir_di.initial_sampling_params.environment_map_importance_sampling == 0u) return 1.0f;
    float u, v; direction_to_equirect_uv(L, &u, &v);                   // RAB_GetEnvironmentMapRandXYFromDir, bridge :389-393
    const uint32_t sx = g.environment_pdf_texture_size[0], sy = g.environment_pdf_texture_size[1];
    const uint32_t tx = (uint32_t)((float)sx * u), ty = (uint32_t)((float)sy * v);
    const float texel = (tx < a.env_w && ty < a.env_h) ? f16_bits_to_f32(__ldg(a.env_pdf + (size_t)ty * a.env_w + tx)) : 0.f;
    const int last = max(0, (int)floorf(log2_pinned((float)max(sx, sy))));
    const float average = f16_bits_to_f32(__ldg(a.env_pdf + a.env_last_offset));
    const float sum = average * (float)(1u << (uint32_t)last << (uint32_t)last);
    return texel / sum;
}

__device__ __forceinline__ bool visibility_ray_miss(const SceneRefs& sc, const Surface& s, V3 sample_position, float offset, uint32_t* rays) {
    const V3 L = sample_position - s.world_pos;                        // setupVisibilityRay, bridge :191-218
    const float tmax = fmaxf(offset, length3(L) - offset * 2);
    uint32_t slot;
    const Hit h = trace_closest(sc, s.world_pos, normalize3(L), offset, tmax, &slot, true);       // "hit or not" is all the shader reads (bridge :218-231)
    (*rays)++;
    return h.geom == 0xFFFFFFFFu;
}

// The sampler in three pieces, so that the megakernel form (trace inlined between them) and the wavefront form of K9
// (ray-queue kernel between them) run the very same arithmetic.
// (1) RTXDI_SampleLocalLights{,Internal}, :264-343 (uniform selection)
// With lr.presampling and localLightSamplingMode == POWER_RIS the lights come from one of K11's tiles, chosen by the tile-coherent
// generator (:241-254, LocalLightSelection.hlsli:43-94, RISBuffer.hlsli:22-45); tile_x/tile_y = launch pixel / RTXDI_TILE_SIZE_IN_PIXELS.
struct TileRng { Rng rng; bool started; };      // the tile-coherent generator (di_fused_resampling.rgen:22): created on first use, shared by the local and environment tile draws
__device__ __forceinline__ float tile_rng_next(TileRng& t, uint32_t tile_x, uint32_t tile_y, uint32_t frame) {
    if (!t.started) { t.rng = rng_init(tile_x, tile_y, frame, 1u); t.started = true; }
    return rng_next(t.rng);
}
__device__ __forceinline__ void sl_local_lights(const SceneRefs& sc, const LightRefs& lr, const rt2_gconst& g, Rng& rng, TileRng& tile_rng, uint32_t tile_x, uint32_t tile_y, uint32_t frame,
                                                const Surface& surface, const SampleParams& sp,
                                                DIReservoir* local_reservoir, LightSample* local_sample) {
    const uint32_t first_local = g.light_buffer_params.local_light_buffer_region.first_light_index;
    const uint32_t num_local = g.light_buffer_params.local_light_buffer_region.num_lights;
    *local_sample = empty_light_sample();
    *local_reservoir = empty_di_reservoir();
    if (num_local != 0u && sp.num_local != 0u) {
        DIReservoir state = empty_di_reservoir();
        const bool ris_mode = lr.presampling != 0u && g.restir_di.initial_sampling_params.local_light_sampling_mode == 1u;
        uint32_t tile_offset = 0, tile_size = 0;
        if (ris_mode) {                                                // RTXDI_RandomlySelectRISTile
            const rt2_ris_buffer_segment_params& seg = g.local_lights_risbuffer_segment_params;
            const uint32_t tile_index = (uint32_t)(tile_rng_next(tile_rng, tile_x, tile_y, frame) * (float)seg.tile_count);
            tile_offset = tile_index * seg.tile_size + seg.buffer_offset;
            tile_size = seg.tile_size;
        }
        for (uint32_t i = 0; i < sp.num_local; i++) {
            const float rnd = rng_next(rng);
            float inv_source_pdf;
            uint32_t light_index;
            rt2_light_info li;
            if (ris_mode) {                                            // RTXDI_RandomlySelectLocalLightFromRISTile
                const uint32_t ptr = min((uint32_t)floorf(rnd * (float)tile_size), tile_size - 1u) + tile_offset;
                const uint2 tile_data = __ldg(lr.ris + ptr);
                light_index = tile_data.x & 0x7FFFFFFFu;               // RTXDI_LIGHT_INDEX_MASK
                inv_source_pdf = __uint_as_float(tile_data.y);
                // RTXDI_LIGHT_COMPACT_BIT: the 32-byte record K11 parked next to the tile entry (RAB_LoadCompactLightInfo, bridge :562-568)
                li = (tile_data.x & 0x80000000u) ? load_light_info(reinterpret_cast<const rt2_light_info*>(lr.ris_light_data), 0u, ptr)
                                                 : load_light_info(lr.lights, light_index);
            } else {                                                   // UniformSampling.hlsli:14-25
                inv_source_pdf = (float)num_local;
                light_index = first_local + min((uint32_t)floorf(rnd * (float)num_local), num_local - 1u);
                li = load_light_info(lr.lights, light_index);
            }
            const float uvx = rng_next(rng), uvy = rng_next(rng);
            const LightSample cand = sample_polymorphic_light(sc, li, surface, uvx, uvy);       // :131-156
            const float blended = light_brdf_mis_weight(surface, cand, 1.0f / inv_source_pdf, sp.local_mis_weight, false, sp);
            const float target_pdf = light_sample_target_pdf(cand, surface);
            const float ris_rnd = rng_next(rng);
            if (blended == 0) continue;
            if (di_stream_sample(state, light_index, uvx, uvy, ris_rnd, target_pdf, 1.0f / blended)) *local_sample = cand;
        }
        di_finalize(state, 1.0f, (float)sp.num_mis);
        state.M = 1;
        *local_reservoir = state;
    }
}
// (1b) RTXDI_SampleEnvironmentMap, :465-499 (compiled with RTXDI_ENABLE_PRESAMPLING only): candidates from one of K11's environment
// tiles; every candidate is streamed, a zero source pdf included (:441-463 has no early-out, unlike the local-light one)
__device__ __forceinline__ void sl_environment_map(const SceneRefs& sc, const LightRefs& lr, const rt2_gconst& g, Rng& rng, TileRng& tile_rng, uint32_t tile_x, uint32_t tile_y,
                                                   uint32_t frame, const Surface& surface, const SampleParams& sp,
                                                   DIReservoir* env_reservoir, LightSample* env_sample) {
    *env_sample = empty_light_sample();
    *env_reservoir = empty_di_reservoir();
    if (lr.presampling == 0u || lr.env_present == 0u || sp.num_env == 0u) return;
    const rt2_ris_buffer_segment_params& seg = g.environment_light_risbuffer_segment_params;
    const uint32_t tile_index = (uint32_t)(tile_rng_next(tile_rng, tile_x, tile_y, frame) * (float)seg.tile_count);   // RTXDI_RandomlySelectRISTile
    const uint32_t tile_offset = tile_index * seg.tile_size + seg.buffer_offset, tile_size = seg.tile_size;
    const rt2_light_info li = load_light_info(lr.lights, lr.env_index);
    DIReservoir state = empty_di_reservoir();
    for (uint32_t i = 0; i < sp.num_env; i++) {
        const float rnd = rng_next(rng);                                // RTXDI_RandomlySelectLightDataFromRISTile, RISBuffer.hlsli:20-31
        const uint32_t ptr = min((uint32_t)floorf(rnd * (float)tile_size), tile_size - 1u) + tile_offset;
        const uint2 tile_data = __ldg(lr.ris + ptr);
        const float inv_source_pdf = __uint_as_float(tile_data.y);     // RTXDI_UnpackEnvironmentLightDataFromRISData, :417-426
        const float uvx = (float)(tile_data.x & 0xffffu) / (float)0xffff, uvy = (float)(tile_data.x >> 16) / (float)0xffff;
        const LightSample cand = sample_polymorphic_light(sc, li, surface, uvx, uvy);
        const float blended = light_brdf_mis_weight(surface, cand, 1.0f / inv_source_pdf, sp.env_mis_weight, true, sp);
        const float target_pdf = light_sample_target_pdf(cand, surface);
        const float ris_rnd = rng_next(rng);
        if (di_stream_sample(state, lr.env_index, uvx, uvy, ris_rnd, target_pdf, 1.0f / blended)) *env_sample = cand;
    }
    di_finalize(state, 1.0f, (float)sp.num_mis);
    state.M = 1;
    *env_reservoir = state;
}
// (2) RTXDI_SampleBrdf, :507-591, one iteration: the ray ...
struct BrdfProbe { bool sampled; V3 dir; float max_distance; };
__device__ __forceinline__ BrdfProbe sl_brdf_ray(const Surface& surface, Rng& rng, const SampleParams& sp) {
    BrdfProbe p; p.dir = v3(0.f, 0.f, 0.f); p.max_distance = 0.f;
    p.sampled = surface_brdf_sample(surface, rng, &p.dir);
    if (p.sampled) p.max_distance = brdf_max_distance_from_pdf(sp.brdf_cutoff, surface_brdf_pdf(surface, p.dir));
    return p;
}
// ... and what follows its hit (RAB_TraceRayForLocalLight, bridge :629-669): geom == ~0u is a miss
__device__ __forceinline__ void sl_brdf_hit(const SceneRefs& sc, const LightRefs& lr, const rt2_gconst& g, Rng& rng, const Surface& surface, const SampleParams& sp, bool sampled,
                                            V3 sample_dir, uint32_t geom, uint32_t prim, float hu, float hv, DIReservoir* brdf_reservoir, LightSample* brdf_sample) {
    float light_source_pdf = 0;
    uint32_t light_index = RT2_INVALID_LIGHT;
    float rx = 0, ry = 0;
    LightSample cand = empty_light_sample();
    if (sampled) {
        if (geom != 0xFFFFFFFFu) {
            uint32_t li = __ldg(lr.geom_to_light + geom);              // getLightIndex, bridge :621-627
            if (li != RT2_INVALID_LIGHT) li += prim;
            light_index = li;
        }
        if (light_index != RT2_INVALID_LIGHT) {
            const float bx = 1 - hu - hv, bz = hv;                     // hitUVToBarycentric, randomFromBarycentric (Helpers.glsl:76-86)
            const float sqrtx = 1 - bx;
            rx = sqrtx * sqrtx; ry = bz / sqrtx;
            cand = sample_polymorphic_light(sc, load_light_info(lr.lights, light_index), surface, rx, ry);
            if (sp.brdf_cutoff > 0.f) {
                const V3 to_light = cand.position - surface.world_pos;
                const float light_distance = length3(to_light);
                const float pdf2 = surface_brdf_pdf(surface, to_light / light_distance);
                if (light_distance > brdf_max_distance_from_pdf(sp.brdf_cutoff, pdf2)) light_index = RT2_INVALID_LIGHT;
            }
            if (light_index != RT2_INVALID_LIGHT) light_source_pdf = local_light_source_pdf(lr, light_index);
        }
        // BRDF miss + environmentLightParams.lightPresent, :556-564: the environment light (index lights + 1) is never written by the
        // reference (Q7) -> absent as shipped; with RT2_LAUNCH_PRESAMPLING the one K10 writes there is sampled
        else if (geom == 0xFFFFFFFFu && lr.env_present != 0u) {
            light_index = lr.env_index;
            direction_to_equirect_uv(sample_dir, &rx, &ry);            // RAB_GetEnvironmentMapRandXYFromDir
            cand = sample_polymorphic_light(sc, load_light_info(lr.lights, light_index), surface, rx, ry);
            light_source_pdf = environment_map_sampling_pdf(lr, g, sample_dir);
        }
    }
    if (light_source_pdf == 0) return;
    const bool is_env = lr.env_present != 0u && light_index == lr.env_index;       // :573
    const float target_pdf = light_sample_target_pdf(cand, surface);
    const float blended = light_brdf_mis_weight(surface, cand, light_source_pdf, is_env ? sp.env_mis_weight : sp.local_mis_weight, is_env, sp);
    const float ris_rnd = rng_next(rng);
    if (di_stream_sample(*brdf_reservoir, light_index, rx, ry, ris_rnd, target_pdf, 1.0f / blended)) *brdf_sample = cand;
}
// (3) finalize the BRDF reservoir and combine, :586-590, 637-663.  The infinite-light region is always empty
// (src/main.rs:375-380; rejected on the host otherwise), so its reservoir and sample are the empty ones.
__device__ __forceinline__ DIReservoir sl_combine(const LightRefs& lr, Rng& rng, const SampleParams& sp, const DIReservoir& local_reservoir, const LightSample& local_sample,
                                                  const DIReservoir& env_reservoir, const LightSample& env_sample,
                                                  DIReservoir brdf_reservoir, const LightSample& brdf_sample, LightSample* o_light_sample) {
    di_finalize(brdf_reservoir, 1.0f, (float)sp.num_mis);
    brdf_reservoir.M = 1;
    const LightSample infinite_sample = empty_light_sample();
    const DIReservoir infinite_reservoir = empty_di_reservoir();
    DIReservoir reservoir = empty_di_reservoir();
    di_combine(reservoir, local_reservoir, 0.5f, local_reservoir.target_pdf);
    const bool select_infinite = di_combine(reservoir, infinite_reservoir, rng_next(rng), infinite_reservoir.target_pdf);
    // :641-643 — with presampling compiled in the environment reservoir is combined too (one more draw, empty reservoir or not)
    bool select_env = false;
    if (lr.presampling != 0u) select_env = di_combine(reservoir, env_reservoir, rng_next(rng), env_reservoir.target_pdf);
    const bool select_brdf = di_combine(reservoir, brdf_reservoir, rng_next(rng), brdf_reservoir.target_pdf);
    di_finalize(reservoir, 1.0f, 1.0f);
    reservoir.M = 1;
    *o_light_sample = select_brdf ? brdf_sample : (select_env ? env_sample : (select_infinite ? infinite_sample : local_sample));
    return reservoir;
}

__device__ __noinline__ DIReservoir sample_lights_for_surface(const SceneRefs& sc, const LightRefs& lr, const rt2_gconst& g, Rng& rng,
                                                              uint32_t tile_x, uint32_t tile_y, uint32_t frame, const Surface& surface,
                                                              const SampleParams& sp, LightSample* o_light_sample, uint32_t& rays, uint32_t& hits) {
    LightSample local_sample; DIReservoir local_reservoir;
    TileRng tile_rng; tile_rng.started = false;
    sl_local_lights(sc, lr, g, rng, tile_rng, tile_x, tile_y, frame, surface, sp, &local_reservoir, &local_sample);
    LightSample env_sample; DIReservoir env_reservoir;
    sl_environment_map(sc, lr, g, rng, tile_rng, tile_x, tile_y, frame, surface, sp, &env_reservoir, &env_sample);
    LightSample brdf_sample = empty_light_sample();
    DIReservoir brdf_reservoir = empty_di_reservoir();
    for (uint32_t i = 0; i < sp.num_brdf; ++i) {
        const BrdfProbe p = sl_brdf_ray(surface, rng, sp);
        Hit h; h.geom = h.prim = 0xFFFFFFFFu; h.u = h.v = 0.f;
        if (p.sampled) {
            uint32_t slot;
            h = trace_closest(sc, surface.world_pos, p.dir, sp.brdf_ray_min_t, p.max_distance, &slot);
            rays++;
            if (h.geom != 0xFFFFFFFFu) hits++;
        }
        sl_brdf_hit(sc, lr, g, rng, surface, sp, p.sampled, p.dir, h.geom, h.prim, h.u, h.v, &brdf_reservoir, &brdf_sample);
    }
    return sl_combine(lr, rng, sp, local_reservoir, local_sample, env_reservoir, env_sample, brdf_reservoir, brdf_sample, o_light_sample);
}

// ShadeSurfaceWithLightSample, ShadingHelpers.glsl:2-58.  Visibility reuse needs reservoir.age > 0 (DIReservoir.hlsli:186-189);
// both callers hand in a freshly sampled reservoir (age 0), so the final-visibility ray is always traced.
__device__ __forceinline__ void shade_surface_with_light_sample(const SceneRefs& sc, const rt2_gconst& g, DIReservoir& reservoir, const Surface& surface,
                                                                LightSample light_sample, V3* diffuse, V3* specular, uint32_t& rays, uint32_t& hits) {
    *diffuse = v3(0.f, 0.f, 0.f); *specular = v3(0.f, 0.f, 0.f);
    if (light_sample.solid_angle_pdf <= 0) return;
    const rt2_restirdi_params& di = g.restir_di;
    if (di.shading_params.enable_final_visibility == 1u) {
        const bool vis = visibility_ray_miss(sc, surface, light_sample.position, 0.01f, &rays);
        if (!vis) hits++;
        const V3 visibility = vis ? v3(1.f, 1.f, 1.f) : v3(0.f, 0.f, 0.f);
        di_store_visibility(reservoir, visibility, di.temporal_resampling_params.discard_invisible_samples == 1u);
        light_sample.radiance = light_sample.radiance * visibility;
    }
    light_sample.radiance = light_sample.radiance * (reservoir.weight_sum / light_sample.solid_angle_pdf);
    if (light_sample.radiance.x > 0.0f || light_sample.radiance.y > 0.0f || light_sample.radiance.z > 0.0f) {
        const SplitBrdf brdf = evaluate_brdf(surface, light_sample.position);
        *diffuse = light_sample.radiance * brdf.demodulated_diffuse;
        *specular = brdf.specular * light_sample.radiance;
    }
}

__device__ __forceinline__ void launch_to_pixel(const rt2_gconst& g, uint32_t lx, uint32_t ly, uint32_t* px, uint32_t* py) {   // RtxdiHelpers.hlsli:52-61
    const uint32_t field = g.runtime_params.active_checkerboard_field;
    *px = lx; *py = ly;
    if (field != 0u) { *px = lx << 1; *px += ((ly + field) & 1u); }
}
__device__ __forceinline__ void flush_ray_stats(unsigned long long* stat_rays, unsigned long long* stat_hits, uint32_t rays, uint32_t hits) {
    const uint32_t am = __activemask();
    const uint32_t wr = __reduce_add_sync(am, rays), wh = __reduce_add_sync(am, hits);
    if ((threadIdx.x & 31u) == (uint32_t)(__ffs(am) - 1)) { atomicAdd(stat_rays, (unsigned long long)wr); atomicAdd(stat_hits, (unsigned long long)wh); }
}

// ------------------------------------------------------------------------------------------------
// K9
// ------------------------------------------------------------------------------------------------
struct DIArgs {
    SceneRefs sc; const rt2_gconst* g; uint32_t width, height, frame; BandRef band;
    GBufferRefs gb;
    LightRefs li;
    rt2_packed_di_reservoir* reservoirs; ushort4* diffuse; ushort4* specular; float4* diffuse_f32;
    unsigned long long* stat_rays; unsigned long long* stat_hits;
    // K9 visibility reuse (k_di_stage_c): trace() calls answered without a walk are counted here (rt2_stats.rays_walked = rays_traced - these)
    unsigned long long* stat_reused; uint32_t visibility_reuse;
};

__global__ void __launch_bounds__(128) k_di_fused(DIArgs a) {
    uint32_t lx, ly;
    if (!tile_pixel(blockIdx.x * blockDim.x + threadIdx.x, a.width, a.band, &lx, &ly)) return;
    const rt2_gconst& g = *a.g;
    const rt2_restirdi_params& di = g.restir_di;
    uint32_t rays = 0, hits = 0;
    uint32_t px, py; launch_to_pixel(g, lx, ly, &px, &py);
    Rng rng = rng_init(px, py, a.frame, 1u);
    const Surface surface = gbuffer_surface(a.gb, a.width, (int)px, (int)py, g.view);
    const SampleParams sp = init_sample_params(di.initial_sampling_params.num_primary_local_light_samples, di.initial_sampling_params.num_primary_infinite_light_samples,
                                               di.initial_sampling_params.num_primary_environment_samples, di.initial_sampling_params.num_primary_brdf_samples,
                                               di.initial_sampling_params.brdf_cutoff, 0.001f);
    LightSample light_sample;
    DIReservoir reservoir = sample_lights_for_surface(a.sc, a.li, g, rng, px >> 4, py >> 4, a.frame, surface, sp, &light_sample, rays, hits);
    // ---- di_fused_resampling.rgen:42-48: initial visibility
    if (di.initial_sampling_params.enable_initial_visibility == 1u && di_valid(reservoir)) {
        if (!visibility_ray_miss(a.sc, surface, light_sample.position, 0.001f, &rays)) { hits++; di_store_visibility(reservoir, v3(0.f, 0.f, 0.f), true); }
    }
    V3 diffuse = v3(0.f, 0.f, 0.f), specular = v3(0.f, 0.f, 0.f);
    if (di_valid(reservoir)) {
        shade_surface_with_light_sample(a.sc, g, reservoir, surface, light_sample, &diffuse, &specular, rays, hits);
        specular = demodulate_specular(surface.specular_f0, specular);
    }
    const uint32_t ptr = reservoir_pointer(di.reservoir_buffer_params.reservoir_block_row_pitch, di.reservoir_buffer_params.reservoir_array_pitch, lx, ly,
                                           di.buffer_indices.shading_input_buffer_index);
    const rt2_packed_di_reservoir packed = pack_di_reservoir(reservoir);
    uint2* dst = reinterpret_cast<uint2*>(a.reservoirs + ptr);
    dst[0] = make_uint2(packed.light_data, packed.uv_data); dst[1] = make_uint2(packed.m_visibility, packed.distance_age);
    dst[2] = make_uint2(__float_as_uint(packed.target_pdf), __float_as_uint(packed.weight));
    if (px < a.width) store_shading_output(g, (size_t)py * a.width + px, diffuse, specular, g.enable_restir_di == 1u, a.diffuse, a.specular, a.diffuse_f32);   // imageStore outside the image is discarded
    flush_ray_stats(a.stat_rays, a.stat_hits, rays, hits);
}

// ------------------------------------------------------------------------------------------------
// Wavefront form of K9 (numPrimaryBrdfSamples <= 1, the reference's setting): the BRDF ray and the one or two
// visibility rays go through the shared ray-queue kernel; the sampler state (reservoir, chosen light sample, RNG
// position: 80 bytes) is parked in five uint4 planes between the stages.
//   A  local-light RIS + BRDF direction          -> ray queue
//   B  BRDF hit -> light, combine                -> initial-visibility ray (if enabled, else final-visibility ray)
//   C  apply initial visibility                  -> final-visibility ray          (only if initial visibility is on)
//   D  apply final visibility, shade, store
// ------------------------------------------------------------------------------------------------
struct DIWaveArgs {
    DIArgs di;
    float4 *st_ray0, *st_ray1; float4* st_hit; const uint32_t* st_hit_inst;
    uint4* state; uint32_t plane;          // 8 planes of `plane` uint4 each (5..7: the environment reservoir, di_env_store)
    const uint4* secondary;                // K5 (DI on): sample at the secondary surface instead of the G-buffer surface
    uint32_t* queue; uint32_t* count;
};
__device__ __forceinline__ void di_state_store(const DIWaveArgs& w, uint32_t slot, const DIReservoir& r, const LightSample& ls, uint32_t rng_index, uint32_t flags) {
    w.state[slot] = make_uint4(r.light_data, r.uv_data, __float_as_uint(r.weight_sum), __float_as_uint(r.target_pdf));
    w.state[(size_t)w.plane + slot] = make_uint4(__float_as_uint(r.M), r.packed_visibility, (uint32_t)r.sdx, (uint32_t)r.sdy);
    w.state[2 * (size_t)w.plane + slot] = make_uint4(r.age, rng_index, flags, ls.light_type);
    w.state[3 * (size_t)w.plane + slot] = make_uint4(__float_as_uint(ls.position.x), __float_as_uint(ls.position.y), __float_as_uint(ls.position.z), __float_as_uint(ls.solid_angle_pdf));
    w.state[4 * (size_t)w.plane + slot] = make_uint4(__float_as_uint(ls.radiance.x), __float_as_uint(ls.radiance.y), __float_as_uint(ls.radiance.z), 0u);
}
__device__ __forceinline__ void di_state_load(const DIWaveArgs& w, uint32_t slot, DIReservoir* r, LightSample* ls, uint32_t* rng_index, uint32_t* flags) {
    const uint4 a = w.state[slot], b = w.state[(size_t)w.plane + slot], c = w.state[2 * (size_t)w.plane + slot];
    const uint4 d = w.state[3 * (size_t)w.plane + slot], e = w.state[4 * (size_t)w.plane + slot];
    r->light_data = a.x; r->uv_data = a.y; r->weight_sum = __uint_as_float(a.z); r->target_pdf = __uint_as_float(a.w);
    r->M = __uint_as_float(b.x); r->packed_visibility = b.y; r->sdx = (int)b.z; r->sdy = (int)b.w;
    r->age = c.x; *rng_index = c.y; *flags = c.z;
    *ls = empty_light_sample();                                        // the normal is not needed past sample_polymorphic_light
    ls->light_type = c.w;
    ls->position = v3(__uint_as_float(d.x), __uint_as_float(d.y), __uint_as_float(d.z)); ls->solid_angle_pdf = __uint_as_float(d.w);
    ls->radiance = v3(__uint_as_float(e.x), __uint_as_float(e.y), __uint_as_float(e.z));
}
// the environment reservoir and its sample wait in planes 5..7 while the BRDF ray is traced (K9 only: K5 asks for no environment samples).
// M is 1 after RTXDI_SampleEnvironmentMap's finalize; the sample's type follows from the reservoir (a selected sample set lightData).
__device__ __forceinline__ bool di_env_active(const LightRefs& lr, const SampleParams& sp) { return lr.presampling != 0u && lr.env_present != 0u && sp.num_env != 0u; }
__device__ __forceinline__ void di_env_store(const DIWaveArgs& w, uint32_t slot, const DIReservoir& r, const LightSample& ls) {
    w.state[5 * (size_t)w.plane + slot] = make_uint4(r.light_data, r.uv_data, __float_as_uint(r.weight_sum), __float_as_uint(r.target_pdf));
    w.state[6 * (size_t)w.plane + slot] = make_uint4(__float_as_uint(ls.position.x), __float_as_uint(ls.position.y), __float_as_uint(ls.position.z), __float_as_uint(ls.solid_angle_pdf));
    w.state[7 * (size_t)w.plane + slot] = make_uint4(__float_as_uint(ls.radiance.x), __float_as_uint(ls.radiance.y), __float_as_uint(ls.radiance.z), 0u);
}
__device__ __forceinline__ void di_env_load(const DIWaveArgs& w, uint32_t slot, DIReservoir* r, LightSample* ls) {
    const uint4 a = w.state[5 * (size_t)w.plane + slot], d = w.state[6 * (size_t)w.plane + slot], e = w.state[7 * (size_t)w.plane + slot];
    *r = empty_di_reservoir();
    r->light_data = a.x; r->uv_data = a.y; r->weight_sum = __uint_as_float(a.z); r->target_pdf = __uint_as_float(a.w); r->M = 1;
    *ls = empty_light_sample();
    ls->light_type = a.x != 0u ? RT2_LIGHT_ENVIRONMENT : 0u;
    ls->position = v3(__uint_as_float(d.x), __uint_as_float(d.y), __uint_as_float(d.z)); ls->solid_angle_pdf = __uint_as_float(d.w);
    ls->radiance = v3(__uint_as_float(e.x), __uint_as_float(e.y), __uint_as_float(e.z));
}
__device__ __forceinline__ SampleParams di_sample_params(const rt2_restirdi_params& di) {
    return init_sample_params(di.initial_sampling_params.num_primary_local_light_samples, di.initial_sampling_params.num_primary_infinite_light_samples,
                              di.initial_sampling_params.num_primary_environment_samples, di.initial_sampling_params.num_primary_brdf_samples,
                              di.initial_sampling_params.brdf_cutoff, 0.001f);
}
// setupVisibilityRay, bridge :191-218
__device__ __forceinline__ void write_visibility_ray(const DIWaveArgs& w, uint32_t slot, const Surface& s, V3 sample_position, float offset) {
    const V3 L = sample_position - s.world_pos;
    const float tmax = fmaxf(offset, length3(L) - offset * 2);
    const V3 d = normalize3(L);
    w.st_ray0[slot] = make_float4(s.world_pos.x, s.world_pos.y, s.world_pos.z, offset);
    w.st_ray1[slot] = make_float4(d.x, d.y, d.z, tmax);
}
#define DIW_BRDF_SAMPLED 1u
#define DIW_RAY_PENDING 2u
#define DIW_ACTIVE 4u

// The surface a stage works on.  SEC = false: K9, the G-buffer surface of the pixel.  SEC = true: K5 with DI on
// (shade_secondary_surfaces.rgen:48-64), the secondary surface unpacked from SecondaryGBufferData; only valid,
// non-environment secondary surfaces sample lights.
template <bool SEC>
__device__ __forceinline__ Surface di_stage_surface(const DIWaveArgs& w, const rt2_gconst& g, uint32_t lx, uint32_t ly, uint32_t px, uint32_t py, bool* active) {
    const DIArgs& a = w.di;
    if (!SEC) { *active = true; return gbuffer_surface(a.gb, a.width, (int)px, (int)py, g.view); }
    const uint32_t gbuffer_index = reservoir_pointer(g.restir_di.reservoir_buffer_params.reservoir_block_row_pitch, g.restir_di.reservoir_buffer_params.reservoir_array_pitch, lx, ly, 0u);
    const uint4* sgp = w.secondary + 3 * (size_t)gbuffer_index;
    const uint4 s0 = sgp[0], s1 = sgp[1];
    const float tx = f16_bits_to_f32((uint16_t)(s1.x & 0xffffu)), ty = f16_bits_to_f32((uint16_t)(s1.x >> 16)), tz = f16_bits_to_f32((uint16_t)(s1.y & 0xffffu));
    const uint32_t flags = s1.y >> 16;
    *active = (tx != 0 || ty != 0 || tz != 0) && (flags & RT2_SECONDARY_IS_ENVIRONMENT) == 0u;
    Surface sec = empty_surface();
    if (!*active) return sec;
    const Surface primary = gbuffer_surface(a.gb, a.width, (int)px, (int)py, g.view);
    sec.world_pos = v3(__uint_as_float(s0.x), __uint_as_float(s0.y), __uint_as_float(s0.z));
    sec.view_depth = 1.0f;
    sec.normal = oct_to_ndir_unorm32(s0.w);
    sec.geo_normal = sec.normal;
    sec.diffuse_albedo = unpack_r11g11b10(s1.z);
    const float4 sr = unpack_rgba8_gamma(s1.w, a.sc.gamma_lut);
    sec.specular_f0 = v3(sr.x, sr.y, sr.z);
    sec.roughness = sr.w;
    sec.diffuse_probability = 1.0f;
    sec.view_dir = normalize3(primary.world_pos - sec.world_pos);
    return sec;
}
template <bool SEC>
__device__ __forceinline__ SampleParams di_stage_params(const rt2_restirdi_params& di) {
    return SEC ? init_sample_params(0u, 0u, 0u, 1u, 0.f, 0.001f) : di_sample_params(di);
}

template <bool SEC>
__global__ void __launch_bounds__(128, RT2_DI_BLOCKS) k_di_stage_a(DIWaveArgs w) {
    const DIArgs& a = w.di;
    uint32_t lx, ly;
    const bool inside = tile_pixel(blockIdx.x * blockDim.x + threadIdx.x, a.width, a.band, &lx, &ly);
    bool traced = false;
    uint32_t slot = 0;
    if (inside) {
        slot = ly * a.width + lx;
        const rt2_gconst& g = *a.g;
        uint32_t px, py; launch_to_pixel(g, lx, ly, &px, &py);
        Rng rng = SEC ? rng_init(lx, ly, a.frame, 6u) : rng_init(px, py, a.frame, 1u);
        bool active;
        const Surface surface = di_stage_surface<SEC>(w, g, lx, ly, px, py, &active);
        const SampleParams sp = di_stage_params<SEC>(g.restir_di);
        LightSample local_sample = empty_light_sample(); DIReservoir local_reservoir = empty_di_reservoir();
        TileRng tile_rng; tile_rng.started = false;
        if (active) sl_local_lights(a.sc, a.li, g, rng, tile_rng, (SEC ? lx : px) >> 4, (SEC ? ly : py) >> 4, a.frame, surface, sp, &local_reservoir, &local_sample);
        if (!SEC && di_env_active(a.li, sp)) {
            LightSample env_sample; DIReservoir env_reservoir;
            sl_environment_map(a.sc, a.li, g, rng, tile_rng, px >> 4, py >> 4, a.frame, surface, sp, &env_reservoir, &env_sample);
            di_env_store(w, slot, env_reservoir, env_sample);
        }
        uint32_t flags = active ? DIW_ACTIVE : 0u;
        if (active && sp.num_brdf != 0u) {
            const BrdfProbe p = sl_brdf_ray(surface, rng, sp);
            if (p.sampled) {
                flags |= DIW_BRDF_SAMPLED;
                w.st_ray0[slot] = make_float4(surface.world_pos.x, surface.world_pos.y, surface.world_pos.z, sp.brdf_ray_min_t);
                w.st_ray1[slot] = make_float4(p.dir.x, p.dir.y, p.dir.z, p.max_distance);
                traced = true;
            }
        }
        di_state_store(w, slot, local_reservoir, local_sample, rng.index, flags);
    }
    warp_append(traced, slot, w.queue, w.count);
}

// B: finish the sampler; queue the first visibility ray that is due
template <bool SEC>
__global__ void __launch_bounds__(128, RT2_DI_BLOCKS) k_di_stage_b(DIWaveArgs w, uint32_t* queue_out, uint32_t* count_out) {
    const DIArgs& a = w.di;
    uint32_t lx, ly;
    const bool inside = tile_pixel(blockIdx.x * blockDim.x + threadIdx.x, a.width, a.band, &lx, &ly);
    bool traced = false;
    uint32_t slot = 0, hits = 0;
    if (inside) {
        slot = ly * a.width + lx;
        const rt2_gconst& g = *a.g;
        const rt2_restirdi_params& di = g.restir_di;
        uint32_t px, py; launch_to_pixel(g, lx, ly, &px, &py);
        bool active;
        const Surface surface = di_stage_surface<SEC>(w, g, lx, ly, px, py, &active);
        const SampleParams sp = di_stage_params<SEC>(di);
        DIReservoir local_reservoir; LightSample local_sample; uint32_t flags;
        Rng rng = SEC ? rng_init(lx, ly, a.frame, 6u) : rng_init(px, py, a.frame, 1u);
        di_state_load(w, slot, &local_reservoir, &local_sample, &rng.index, &flags);
        if (active) {
        LightSample brdf_sample = empty_light_sample();
        DIReservoir brdf_reservoir = empty_di_reservoir();
        if (sp.num_brdf != 0u) {
            const bool sampled = (flags & DIW_BRDF_SAMPLED) != 0u;
            uint32_t geom = 0xFFFFFFFFu, prim = 0xFFFFFFFFu; float hu = 0.f, hv = 0.f;
            V3 sample_dir = v3(0.f, 0.f, 0.f);
            if (sampled) {
                const float4 rd = w.st_ray1[slot];                       // the BRDF direction stage A queued (the trace kernel only reads it)
                sample_dir = v3(rd.x, rd.y, rd.z);
                const float4 h = w.st_hit[slot];
                const uint32_t tri_slot = __float_as_uint(h.w);
                if (tri_slot != 0xFFFFFFFFu) {
                    hits++;
                    const TriPos* tp = a.sc.tris + tri_slot;
                    geom = f2u(ld16f(&tp->v0).w); prim = f2u(ld16f(&tp->v1).w); hu = h.y; hv = h.z;
                }
            }
            sl_brdf_hit(a.sc, a.li, g, rng, surface, sp, sampled, sample_dir, geom, prim, hu, hv, &brdf_reservoir, &brdf_sample);
        }
        LightSample env_sample = empty_light_sample(); DIReservoir env_reservoir = empty_di_reservoir();
        if (!SEC && di_env_active(a.li, sp)) di_env_load(w, slot, &env_reservoir, &env_sample);
        LightSample light_sample;
        DIReservoir reservoir = sl_combine(a.li, rng, sp, local_reservoir, local_sample, env_reservoir, env_sample, brdf_reservoir, brdf_sample, &light_sample);
        flags = DIW_ACTIVE;
        // K9 runs the initial-visibility test of di_fused_resampling.rgen:42-48 first; K5 shades with enableVisibilityReuse false and has none
        if (di_valid(reservoir) || SEC) {
            if (!SEC && di.initial_sampling_params.enable_initial_visibility == 1u) { write_visibility_ray(w, slot, surface, light_sample.position, 0.001f); traced = true; }
            else if (!(light_sample.solid_angle_pdf <= 0) && di.shading_params.enable_final_visibility == 1u) { write_visibility_ray(w, slot, surface, light_sample.position, 0.01f); traced = true; }
        }
        if (traced) flags |= DIW_RAY_PENDING;
        di_state_store(w, slot, reservoir, light_sample, rng.index, flags);
        }
    }
    warp_append(traced, slot, queue_out, count_out);
    flush_ray_stats(a.stat_rays, a.stat_hits, 0u, hits);
}

// C (only when initial visibility is on): di_fused_resampling.rgen:42-48, then queue the final-visibility ray.
// Visibility reuse: the final-visibility ray (ShadingHelpers.glsl:35, offset 0.01) has the origin and direction of the initial one
// (bridge :218-222, offset 0.001) — same surface, same light sample — and its interval [0.01, max(0.01, |L| - 0.02)] lies inside
// [0.001, max(0.001, |L| - 0.002)] unless |L| < 0.012.  A triangle's t does not depend on the interval, so an initial ray that came
// back without a hit proves the final one has none either: it is answered "visible" here, counted as the trace() call it is (and in
// stat_reused as a walk saved), and only the degenerate short rays are queued.  Same reservoirs and images, bit for bit.
__global__ void __launch_bounds__(128) k_di_stage_c(DIWaveArgs w, uint32_t* queue_out, uint32_t* count_out) {
    const DIArgs& a = w.di;
    uint32_t lx, ly;
    const bool inside = tile_pixel(blockIdx.x * blockDim.x + threadIdx.x, a.width, a.band, &lx, &ly);
    bool traced = false;
    uint32_t slot = 0, hits = 0, reused = 0;
    if (inside) {
        slot = ly * a.width + lx;
        const rt2_gconst& g = *a.g;
        const rt2_restirdi_params& di = g.restir_di;
        uint32_t px, py; launch_to_pixel(g, lx, ly, &px, &py);
        DIReservoir reservoir; LightSample light_sample; uint32_t flags, rng_index;
        di_state_load(w, slot, &reservoir, &light_sample, &rng_index, &flags);
        bool initial_clear = false;                       // the initial-visibility ray was traced and found nothing
        if (flags & DIW_RAY_PENDING) {
            if (__float_as_uint(w.st_hit[slot].w) != 0xFFFFFFFFu) { hits++; di_store_visibility(reservoir, v3(0.f, 0.f, 0.f), true); }
            else initial_clear = true;
        }
        flags = 0u;
        if (di_valid(reservoir) && !(light_sample.solid_angle_pdf <= 0) && di.shading_params.enable_final_visibility == 1u) {
            const Surface surface = gbuffer_surface(a.gb, a.width, (int)px, (int)py, g.view);
            const float len = length3(light_sample.position - surface.world_pos);
            const bool inside_initial = fmaxf(0.01f, len - 0.01f * 2) <= fmaxf(0.001f, len - 0.001f * 2);     // the two tmax of setupVisibilityRay
            if (a.visibility_reuse && initial_clear && inside_initial) reused = 1u;          // visible: stage D reads "no ray pending" as that
            else {
                write_visibility_ray(w, slot, surface, light_sample.position, 0.01f);
                traced = true; flags |= DIW_RAY_PENDING;
            }
        }
        di_state_store(w, slot, reservoir, light_sample, rng_index, flags);
    }
    warp_append(traced, slot, queue_out, count_out);
    flush_ray_stats(a.stat_rays, a.stat_hits, reused, hits);
    if (a.visibility_reuse) {
        const uint32_t am = __activemask(), wr = __reduce_add_sync(am, reused);
        if (wr && (threadIdx.x & 31u) == (uint32_t)(__ffs(am) - 1)) atomicAdd(a.stat_reused, (unsigned long long)wr);
    }
}

// D: ShadeSurfaceWithLightSample with the traced final visibility, DemodulateSpecular, stores (di_fused_resampling.rgen:50-92)
__global__ void __launch_bounds__(128) k_di_stage_d(DIWaveArgs w, uint32_t final_ray_was_queued) {
    const DIArgs& a = w.di;
    uint32_t lx, ly;
    if (!tile_pixel(blockIdx.x * blockDim.x + threadIdx.x, a.width, a.band, &lx, &ly)) return;
    const uint32_t slot = ly * a.width + lx;
    const rt2_gconst& g = *a.g;
    const rt2_restirdi_params& di = g.restir_di;
    uint32_t px, py; launch_to_pixel(g, lx, ly, &px, &py);
    uint32_t hits = 0;
    DIReservoir reservoir; LightSample light_sample; uint32_t flags, rng_index;
    di_state_load(w, slot, &reservoir, &light_sample, &rng_index, &flags);
    const Surface surface = gbuffer_surface(a.gb, a.width, (int)px, (int)py, g.view);
    V3 diffuse = v3(0.f, 0.f, 0.f), specular = v3(0.f, 0.f, 0.f);
    if (di_valid(reservoir)) {
        if (!(light_sample.solid_angle_pdf <= 0)) {
            if (di.shading_params.enable_final_visibility == 1u) {
                const bool vis = !(final_ray_was_queued && (flags & DIW_RAY_PENDING) && __float_as_uint(w.st_hit[slot].w) != 0xFFFFFFFFu);
                if (!vis) hits++;
                const V3 visibility = vis ? v3(1.f, 1.f, 1.f) : v3(0.f, 0.f, 0.f);
                di_store_visibility(reservoir, visibility, di.temporal_resampling_params.discard_invisible_samples == 1u);
                light_sample.radiance = light_sample.radiance * visibility;
            }
            light_sample.radiance = light_sample.radiance * (reservoir.weight_sum / light_sample.solid_angle_pdf);
            if (light_sample.radiance.x > 0.0f || light_sample.radiance.y > 0.0f || light_sample.radiance.z > 0.0f) {
                const SplitBrdf brdf = evaluate_brdf(surface, light_sample.position);
                diffuse = light_sample.radiance * brdf.demodulated_diffuse;
                specular = brdf.specular * light_sample.radiance;
            }
        }
        specular = demodulate_specular(surface.specular_f0, specular);
    }
    const uint32_t ptr = reservoir_pointer(di.reservoir_buffer_params.reservoir_block_row_pitch, di.reservoir_buffer_params.reservoir_array_pitch, lx, ly,
                                           di.buffer_indices.shading_input_buffer_index);
    const rt2_packed_di_reservoir packed = pack_di_reservoir(reservoir);
    uint2* dst = reinterpret_cast<uint2*>(a.reservoirs + ptr);
    dst[0] = make_uint2(packed.light_data, packed.uv_data); dst[1] = make_uint2(packed.m_visibility, packed.distance_age);
    dst[2] = make_uint2(__float_as_uint(packed.target_pdf), __float_as_uint(packed.weight));
    if (px < a.width) store_shading_output(g, (size_t)py * a.width + px, diffuse, specular, g.enable_restir_di == 1u, a.diffuse, a.specular, a.diffuse_f32);
    flush_ray_stats(a.stat_rays, a.stat_hits, 0u, hits);
}

// ------------------------------------------------------------------------------------------------
// The default ReSTIR-GI chain (src/light_passes.rs:621-659): K4 -> K5 -> K8.
// One thread per launch index, 8x4 pixel tiles per warp.  K4 and K5 trace one incoherent ray each and are
// bound like a bounce wavefront of K1; K8 only evaluates BRDFs (plus an optional visibility ray).  The
// 48-byte secondary G-buffer record and the 32-byte GI reservoir move as 16-byte words.
// ------------------------------------------------------------------------------------------------
struct GIArgs {
    SceneRefs sc; const rt2_gconst* g; uint32_t width, height, frame; BandRef band;
    GBufferRefs gb;
    LightRefs li;
    uint4* secondary;        // rt2_secondary_gbuffer_data: 3 x uint4 per entry
    uint4* reservoirs;       // rt2_packed_gi_reservoir: 2 x uint4 per entry
    ushort4* diffuse; ushort4* specular; float4* diffuse_f32;
    unsigned long long* stat_rays; unsigned long long* stat_hits;
};

__device__ __forceinline__ V3 unpack_half3(uint32_t tx, uint32_t ty) {
    return v3(f16_bits_to_f32((uint16_t)(tx & 0xffffu)), f16_bits_to_f32((uint16_t)(tx >> 16)), f16_bits_to_f32((uint16_t)(ty & 0xffffu)));
}

// K4 — lighting_passes/brdf_rays.rgen:19-195
__global__ void __launch_bounds__(128) k_brdf_rays(GIArgs a) {
    __shared__ float lut[256];
    lut[threadIdx.x] = a.sc.srgb_lut[threadIdx.x]; lut[threadIdx.x + 128] = a.sc.srgb_lut[threadIdx.x + 128];
    __syncthreads();
    uint32_t lx, ly;
    if (!tile_pixel(blockIdx.x * blockDim.x + threadIdx.x, a.width, a.band, &lx, &ly)) return;
    const rt2_gconst& g = *a.g;
    uint32_t px, py; launch_to_pixel(g, lx, ly, &px, &py);
    const Surface surface = gbuffer_surface(a.gb, a.width, (int)px, (int)py, g.view);
    uint32_t rays = 0, hits = 0;
    if (surface.view_depth != RT2_BACKGROUND_DEPTH) {                  // RAB_IsSurfaceValid, bridge :346-349
        Rng rng = rng_init(lx, ly, a.frame, 5u);
        V3 tangent, bitangent; construct_onb(surface.normal, &tangent, &bitangent);        // branchlessONB, Helpers.glsl:271-278
        const V3 cam = v3(g.view.camera_direction_or_position[0], g.view.camera_direction_or_position[1], g.view.camera_direction_or_position[2]);
        const float depth = fmaxf(1.0f, 0.1f * length3(surface.world_pos - cam));
        const float tmin = 0.001f * depth;
        float tmax = RT2_BACKGROUND_DEPTH;
        const float randx = rng_next(rng), randy = rng_next(rng);
        const V3 V = normalize3(cam - surface.world_pos);
        const bool is_delta = surface.roughness == 0;
        bool is_specular_ray;
        float specular_pdf, overall_pdf;
        V3 brdf_over_pdf, dir;
        {
            const V3 Ve = v3(dot3(V, tangent), dot3(V, bitangent), dot3(V, surface.normal));
            const V3 He = sample_ggx_vndf(Ve, surface.roughness, randx, randy);
            const V3 H = is_delta ? surface.normal : normalize3((tangent * He.x + bitangent * He.y) + surface.normal * He.z);
            const V3 specular_direction = reflect3(neg3(V), H);
            const float HoV = saturatef(dot3(H, V)), NoV = saturatef(dot3(surface.normal, V));
            const V3 F = schlick_fresnel(surface.specular_f0, HoV);
            const float G1 = is_delta ? 1.0f : (NoV > 0) ? g1_smith(surface.roughness, NoV) : 0.f;
            const V3 specular_brdf_over_pdf = F * G1;
            const V3 local_direction = sample_cos_hemisphere(randx, randy, false);
            const V3 diffuse_direction = (tangent * local_direction.x + bitangent * local_direction.y) + surface.normal * local_direction.z;
            const float diffuse_brdf_over_pdf = 1.0f;
            specular_pdf = saturatef(calc_luminance(specular_brdf_over_pdf) / calc_luminance(specular_brdf_over_pdf + surface.diffuse_albedo * diffuse_brdf_over_pdf));
            is_specular_ray = rng_next(rng) < specular_pdf;
            if (is_specular_ray) { dir = specular_direction; brdf_over_pdf = specular_brdf_over_pdf / specular_pdf; }
            else { dir = diffuse_direction; const float d = diffuse_brdf_over_pdf / (1.0f - specular_pdf); brdf_over_pdf = v3(d, d, d); }
            const float specular_lobe_pdf = ggx_vndf_pdf(surface.roughness, surface.normal, V, dir);
            const float diffuse_lobe_pdf = saturatef(dot3(dir, surface.normal)) / RT2_RTXDI_PI;
            overall_pdf = is_delta ? diffuse_lobe_pdf : mixf(diffuse_lobe_pdf, specular_lobe_pdf, specular_pdf);
        }
        if (dot3(surface.geo_normal, dir) <= 0.0f) { brdf_over_pdf = v3(0.f, 0.f, 0.f); tmax = 0.f; }
        const V3 origin = surface.world_pos;
        V3 radiance = v3(0.f, 0.f, 0.f);
        Hit h; h.t = 0.f; h.u = h.v = 0.f; h.inst = 0u; h.geom = 0xFFFFFFFFu; h.prim = 0xFFFFFFFFu;
        uint32_t tri_slot = 0;
        if (tmin <= tmax) { h = trace_closest(a.sc, origin, dir, tmin, tmax, &tri_slot); rays++; }   // TMin > TMax: miss (oracle P2)
        V3 sec_position, sec_normal, sec_albedo, sec_f0; float sec_roughness; bool sec_is_env;
        const bool include_emissive = g.enable_restir_di == 0u || (is_specular_ray && is_delta);
        if (h.geom != 0xFFFFFFFFu) {
            hits++;
            const SurfaceFetch sf = get_geometry_from_hit(a.sc, lut, tri_slot, h.inst, h.u, h.v);
            if (include_emissive) radiance = radiance + sf.emission;
            sec_position = origin + dir * h.t;
            sec_normal = (dot3(sf.normal, dir) < 0) ? sf.normal : neg3(sf.normal);
            sec_albedo = sf.color; sec_f0 = sf.specular_f0; sec_roughness = sf.roughness; sec_is_env = false;
        } else {
            if (include_emissive) radiance = radiance + environment_radiance(a.sc, g, dir);
            sec_position = origin + dir * 1000.0f;                     // DISTANT_LIGHT_DISTANCE, PolymorphicLight.glsl:43
            sec_normal = neg3(dir);
            sec_albedo = v3(0.f, 0.f, 0.f); sec_f0 = v3(0.f, 0.f, 0.f); sec_roughness = 0.0f; sec_is_env = true;
        }
        if (g.enable_brdf_indirect == 1u) {
            uint32_t fl = 0;
            if (is_specular_ray) fl |= RT2_SECONDARY_IS_SPECULAR_RAY;
            if (is_delta) fl |= RT2_SECONDARY_IS_DELTA_SURFACE;
            if (sec_is_env) fl |= RT2_SECONDARY_IS_ENVIRONMENT;
            const uint32_t ptr = reservoir_pointer(g.restir_gi.reservoir_buffer_params.reservoir_block_row_pitch, g.restir_gi.reservoir_buffer_params.reservoir_array_pitch, lx, ly, 0u);
            uint4* dst = a.secondary + 3 * (size_t)ptr;
            dst[0] = make_uint4(__float_as_uint(sec_position.x), __float_as_uint(sec_position.y), __float_as_uint(sec_position.z), ndir_to_oct_unorm32(sec_normal));
            dst[1] = make_uint4(pack_r16g16_float(brdf_over_pdf.x, brdf_over_pdf.y), pack_r16g16_float(brdf_over_pdf.z, 0.0f) | (fl << 16),
                                pack_r11g11b10(sec_albedo), pack_rgba8_gamma(sec_f0.x, sec_f0.y, sec_f0.z, sec_roughness));
            dst[2] = make_uint4(__float_as_uint(radiance.x), __float_as_uint(radiance.y), __float_as_uint(radiance.z), __float_as_uint(overall_pdf));
            if (g.enable_restir_gi == 1u) radiance = v3(0.f, 0.f, 0.f);
        }
        if (radiance.x > 0.0f || radiance.y > 0.0f || radiance.z > 0.0f || g.enable_brdf_additive_blend == 0u) {
            const V3 diffuse = is_specular_ray ? v3(0.f, 0.f, 0.f) : radiance * brdf_over_pdf;
            V3 specular = is_specular_ray ? radiance * brdf_over_pdf : v3(0.f, 0.f, 0.f);
            specular = demodulate_specular(surface.specular_f0, specular);
            if (px < a.width) store_shading_output(g, (size_t)py * a.width + px, diffuse, specular, false, a.diffuse, a.specular, a.diffuse_f32);
        }
    }
    flush_ray_stats(a.stat_rays, a.stat_hits, rays, hits);
}

// K5 — lighting_passes/shade_secondary_surfaces.rgen:26-157
__global__ void __launch_bounds__(128) k_shade_secondary(GIArgs a) {
    __shared__ float lut[256];
    lut[threadIdx.x] = a.sc.srgb_lut[threadIdx.x]; lut[threadIdx.x + 128] = a.sc.srgb_lut[threadIdx.x + 128];
    __syncthreads();
    uint32_t lx, ly;
    if (!tile_pixel(blockIdx.x * blockDim.x + threadIdx.x, a.width, a.band, &lx, &ly)) return;
    const rt2_gconst& g = *a.g;
    uint32_t px, py; launch_to_pixel(g, lx, ly, &px, &py);
    uint32_t rays = 0, hits = 0;
    Rng rng = rng_init(lx, ly, a.frame, 6u);
    // sic: the secondary G-buffer is addressed with the DI pitches here (:35) and with the GI pitches in K4/K8
    const uint32_t gbuffer_index = reservoir_pointer(g.restir_di.reservoir_buffer_params.reservoir_block_row_pitch, g.restir_di.reservoir_buffer_params.reservoir_array_pitch, lx, ly, 0u);
    const Surface primary = gbuffer_surface(a.gb, a.width, (int)px, (int)py, g.view);
    uint4* sgp = a.secondary + 3 * (size_t)gbuffer_index;
    const uint4 s0 = sgp[0], s1 = sgp[1], s2 = sgp[2];
    const V3 throughput = unpack_half3(s1.x, s1.y);
    const uint32_t flags = s1.y >> 16;
    const bool valid_secondary = throughput.x != 0 || throughput.y != 0 || throughput.z != 0;
    const bool is_specular_ray = (flags & RT2_SECONDARY_IS_SPECULAR_RAY) != 0u;
    const bool is_delta = (flags & RT2_SECONDARY_IS_DELTA_SURFACE) != 0u;
    const bool is_env = (flags & RT2_SECONDARY_IS_ENVIRONMENT) != 0u;
    V3 radiance = v3(__uint_as_float(s2.x), __uint_as_float(s2.y), __uint_as_float(s2.z));
    const float pdf = __uint_as_float(s2.w);
    Surface sec = empty_surface();
    sec.world_pos = v3(__uint_as_float(s0.x), __uint_as_float(s0.y), __uint_as_float(s0.z));
    sec.view_depth = 1.0f;
    sec.normal = oct_to_ndir_unorm32(s0.w);
    sec.geo_normal = sec.normal;
    sec.diffuse_albedo = unpack_r11g11b10(s1.z);
    const float4 sr = unpack_rgba8_gamma(s1.w, a.sc.gamma_lut);
    sec.specular_f0 = v3(sr.x, sr.y, sr.z);
    sec.roughness = sr.w;
    sec.diffuse_probability = 1.0f;
    sec.view_dir = normalize3(primary.world_pos - sec.world_pos);
    if (valid_secondary && !is_env && g.enable_restir_di == 1u) {
        const SampleParams sp = init_sample_params(0u, 0u, 0u, 1u, 0.f, 0.001f);
        LightSample light_sample;
        DIReservoir reservoir = sample_lights_for_surface(a.sc, a.li, g, rng, lx >> 4, ly >> 4, a.frame, sec, sp, &light_sample, rays, hits);
        V3 indirect_diffuse, indirect_specular;
        shade_surface_with_light_sample(a.sc, g, reservoir, sec, light_sample, &indirect_diffuse, &indirect_specular, rays, hits);
        radiance = radiance + (indirect_diffuse * sec.diffuse_albedo + indirect_specular);
        const float c_max_indirect_radiance = 100.0f;                  // firefly suppression, :113-116
        const float lum = calc_luminance(radiance);
        if (lum > c_max_indirect_radiance) radiance = radiance * (c_max_indirect_radiance / lum);
    }
    if (g.enable_restir_di == 0u) {                                    // :119-128: one more cosine bounce
        V3 dir;
        surface_brdf_sample(sec, rng, &dir);
        uint32_t tri_slot;
        const Hit h = trace_closest(a.sc, sec.world_pos, dir, 0.001f, 1000.0f, &tri_slot);
        rays++;
        V3 emission;
        if (h.geom == 0xFFFFFFFFu) emission = environment_radiance(a.sc, g, dir);       // Q2 / oracle P4
        else { hits++; emission = get_geometry_from_hit(a.sc, lut, tri_slot, h.inst, h.u, h.v).emission; }
        radiance = radiance + emission * sec.diffuse_albedo;
    }
    GIReservoir reservoir = empty_gi_reservoir();
    const bool output_shading_result = is_specular_ray && is_delta;
    if (valid_secondary && !output_shading_result) reservoir = make_gi_reservoir(sec.world_pos, sec.normal, radiance, pdf);
    const uint32_t rx = g.runtime_params.active_checkerboard_field != 0u ? (px >> 1) : px;     // RTXDI_PixelPosToReservoirPos
    store_gi_reservoir(a.reservoirs, reservoir_pointer(g.restir_gi.reservoir_buffer_params.reservoir_block_row_pitch, g.restir_gi.reservoir_buffer_params.reservoir_array_pitch,
                                                       rx, py, g.restir_gi.buffer_indices.secondary_surface_restirdi_output_buffer_index), reservoir, 0u);
    const V3 saved = output_shading_result ? v3(0.f, 0.f, 0.f) : radiance;                    // initial sample radiance for the MIS in K8
    sgp[2] = make_uint4(__float_as_uint(saved.x), __float_as_uint(saved.y), __float_as_uint(saved.z), s2.w);
    if (output_shading_result) {
        const V3 specular = demodulate_specular(primary.specular_f0, radiance);              // is_specular_ray holds: diffuse = 0
        if (px < a.width) store_shading_output(g, (size_t)py * a.width + px, v3(0.f, 0.f, 0.f), specular, false, a.diffuse, a.specular, a.diffuse_f32);
    }
    flush_ray_stats(a.stat_rays, a.stat_hits, rays, hits);
}

// K8 — lighting_passes/gi_final_shading.rgen:18-101
__device__ __forceinline__ float gi_mis_weight(const SplitBrdf& rough, const SplitBrdf& tru, V3 albedo) {
    const float k_max = 1e4f;
    V3 cr = albedo * rough.demodulated_diffuse + rough.specular;
    V3 ct = albedo * tru.demodulated_diffuse + tru.specular;
    cr = v3(clampf(cr.x, 1e-4f, k_max), clampf(cr.y, 1e-4f, k_max), clampf(cr.z, 1e-4f, k_max));
    ct = v3(clampf(ct.x, 0.f, k_max), clampf(ct.y, 0.f, k_max), clampf(ct.z, 0.f, k_max));
    const float w = saturatef(calc_luminance(ct) / calc_luminance(ct + cr));
    return w * w * w;
}

// 6 blocks/SM (80 registers instead of 128)
#ifndef RT2_K8_BLOCKS
#define RT2_K8_BLOCKS 6
#endif
__global__ void __launch_bounds__(128, RT2_K8_BLOCKS) k_gi_final_shading(GIArgs a) {
    uint32_t lx, ly;
    if (!tile_pixel(blockIdx.x * blockDim.x + threadIdx.x, a.width, a.band, &lx, &ly)) return;
    const rt2_gconst& g = *a.g;
    uint32_t px, py; launch_to_pixel(g, lx, ly, &px, &py);
    uint32_t rays = 0, hits = 0;
    const Surface primary = gbuffer_surface(a.gb, a.width, (int)px, (int)py, g.view);
    const uint32_t rx = g.runtime_params.active_checkerboard_field != 0u ? (px >> 1) : px;
    const uint32_t row_pitch = g.restir_gi.reservoir_buffer_params.reservoir_block_row_pitch, array_pitch = g.restir_gi.reservoir_buffer_params.reservoir_array_pitch;
    // sic: reads the K5 output page, not finalShadingInputBufferIndex (:50)
    const GIReservoir reservoir = load_gi_reservoir(a.reservoirs, reservoir_pointer(row_pitch, array_pitch, rx, py, g.restir_gi.buffer_indices.secondary_surface_restirdi_output_buffer_index), a.sc.logluv_lut);
    V3 diffuse = v3(0.f, 0.f, 0.f), specular = v3(0.f, 0.f, 0.f);
    if (reservoir.M != 0u) {
        V3 radiance = reservoir.radiance * reservoir.weight_sum;
        V3 visibility = v3(1.f, 1.f, 1.f);
        if (g.restir_gi.final_shading_params.enable_final_visibility == 1u) {
            const bool miss = visibility_ray_miss(a.sc, primary, reservoir.position, 0.001f, &rays);
            if (!miss) hits++;
            const float vis = miss ? 1.0f : 0.0f;
            visibility = v3(vis, vis, vis);
        }
        radiance = radiance * visibility;
        const SplitBrdf brdf = evaluate_brdf(primary, reservoir.position);
        if (g.restir_gi.final_shading_params.enable_final_mis == 1u) {
            const uint4* sgp = a.secondary + 3 * (size_t)reservoir_pointer(row_pitch, array_pitch, rx, py, 0u);       // LoadInitialSampleReservoir, :30-41
            const uint4 s0 = sgp[0], s1 = sgp[1], s2 = sgp[2];
            const V3 normal = oct_to_ndir_unorm32(s0.w);
            const V3 throughput = unpack_half3(s1.x, s1.y);
            const GIReservoir initial = make_gi_reservoir(v3(__uint_as_float(s0.x), __uint_as_float(s0.y), __uint_as_float(s0.z)), normal,
                                                          v3(__uint_as_float(s2.x), __uint_as_float(s2.y), __uint_as_float(s2.z)) * throughput, __uint_as_float(s2.w));
            const SplitBrdf brdf0 = evaluate_brdf(primary, initial.position);
            Surface roughened = primary;
            roughened.roughness = fmaxf(roughened.roughness, 0.3f);                          // kMISRoughness
            const SplitBrdf rough_brdf = evaluate_brdf(roughened, reservoir.position);
            const SplitBrdf rough_brdf0 = evaluate_brdf(roughened, initial.position);
            const float final_weight = 1.0f - gi_mis_weight(rough_brdf, brdf, primary.diffuse_albedo);
            const float initial_weight = gi_mis_weight(rough_brdf0, brdf0, primary.diffuse_albedo);
            const V3 initial_radiance = initial.radiance * initial.weight_sum;
            diffuse = radiance * brdf.demodulated_diffuse * final_weight + initial_radiance * brdf0.demodulated_diffuse * initial_weight;
            specular = brdf.specular * radiance * final_weight + brdf0.specular * initial_radiance * initial_weight;
        } else {
            diffuse = radiance * brdf.demodulated_diffuse;
            specular = brdf.specular * radiance;
        }
        specular = demodulate_specular(primary.specular_f0, specular);
    }
    if (px < a.width) store_shading_output(g, (size_t)py * a.width + px, diffuse, specular, g.enable_restir_di == 0u, a.diffuse, a.specular, a.diffuse_f32);
    flush_ray_stats(a.stat_rays, a.stat_hits, rays, hits);
}

// ------------------------------------------------------------------------------------------------
// Wavefront form of K4 and of K5's DI-off branch: generate the ray (all shader code up to trace()), hand the
// incoherent rays to the shared ray-queue kernel (rt2_render.cu: postponing + lane refill, 2-3x the rate of a
// traversal inlined in a megakernel), then finish the shader from the hit record.  Same arithmetic, same order:
// the split is invisible in the outputs.  Per-pixel hand-over state lives in the path-state arrays of K1
// (st_ray0/1 = ray, st_acc = BRDF/pdf, st_misc.x = flags), slot = launch index.
// ------------------------------------------------------------------------------------------------
struct GIWaveArgs {
    GIArgs gi;
    float4 *st_ray0, *st_ray1, *st_acc; uint4* st_misc; float4* st_hit; const uint32_t* st_hit_inst;
    uint32_t* queue; uint32_t* count;
};
#define GIW_VALID 1u
#define GIW_SPECULAR 2u
#define GIW_DELTA 4u

__global__ void __launch_bounds__(128) k_brdf_rays_gen(GIWaveArgs w) {
    const GIArgs& a = w.gi;
    uint32_t lx, ly;
    const bool inside = tile_pixel(blockIdx.x * blockDim.x + threadIdx.x, a.width, a.band, &lx, &ly);
    bool traced = false;
    uint32_t slot = 0;
    if (inside) {
        slot = ly * a.width + lx;
        const rt2_gconst& g = *a.g;
        uint32_t px, py; launch_to_pixel(g, lx, ly, &px, &py);
        const Surface surface = gbuffer_surface(a.gb, a.width, (int)px, (int)py, g.view);
        uint32_t flags = 0u;
        if (surface.view_depth != RT2_BACKGROUND_DEPTH) {              // RAB_IsSurfaceValid, bridge :346-349
            Rng rng = rng_init(lx, ly, a.frame, 5u);
            V3 tangent, bitangent; construct_onb(surface.normal, &tangent, &bitangent);
            const V3 cam = v3(g.view.camera_direction_or_position[0], g.view.camera_direction_or_position[1], g.view.camera_direction_or_position[2]);
            const float depth = fmaxf(1.0f, 0.1f * length3(surface.world_pos - cam));
            const float tmin = 0.001f * depth;
            float tmax = RT2_BACKGROUND_DEPTH;
            const float randx = rng_next(rng), randy = rng_next(rng);
            const V3 V = normalize3(cam - surface.world_pos);
            const bool is_delta = surface.roughness == 0;
            bool is_specular_ray;
            float specular_pdf, overall_pdf;
            V3 brdf_over_pdf, dir;
            {
                const V3 Ve = v3(dot3(V, tangent), dot3(V, bitangent), dot3(V, surface.normal));
                const V3 He = sample_ggx_vndf(Ve, surface.roughness, randx, randy);
                const V3 H = is_delta ? surface.normal : normalize3((tangent * He.x + bitangent * He.y) + surface.normal * He.z);
                const V3 specular_direction = reflect3(neg3(V), H);
                const float HoV = saturatef(dot3(H, V)), NoV = saturatef(dot3(surface.normal, V));
                const V3 F = schlick_fresnel(surface.specular_f0, HoV);
                const float G1 = is_delta ? 1.0f : (NoV > 0) ? g1_smith(surface.roughness, NoV) : 0.f;
                const V3 specular_brdf_over_pdf = F * G1;
                const V3 local_direction = sample_cos_hemisphere(randx, randy, false);
                const V3 diffuse_direction = (tangent * local_direction.x + bitangent * local_direction.y) + surface.normal * local_direction.z;
                const float diffuse_brdf_over_pdf = 1.0f;
                specular_pdf = saturatef(calc_luminance(specular_brdf_over_pdf) / calc_luminance(specular_brdf_over_pdf + surface.diffuse_albedo * diffuse_brdf_over_pdf));
                is_specular_ray = rng_next(rng) < specular_pdf;
                if (is_specular_ray) { dir = specular_direction; brdf_over_pdf = specular_brdf_over_pdf / specular_pdf; }
                else { dir = diffuse_direction; const float d = diffuse_brdf_over_pdf / (1.0f - specular_pdf); brdf_over_pdf = v3(d, d, d); }
                const float specular_lobe_pdf = ggx_vndf_pdf(surface.roughness, surface.normal, V, dir);
                const float diffuse_lobe_pdf = saturatef(dot3(dir, surface.normal)) / RT2_RTXDI_PI;
                overall_pdf = is_delta ? diffuse_lobe_pdf : mixf(diffuse_lobe_pdf, specular_lobe_pdf, specular_pdf);
            }
            if (dot3(surface.geo_normal, dir) <= 0.0f) { brdf_over_pdf = v3(0.f, 0.f, 0.f); tmax = 0.f; }
            flags = GIW_VALID | (is_specular_ray ? GIW_SPECULAR : 0u) | (is_delta ? GIW_DELTA : 0u);
            w.st_ray0[slot] = make_float4(surface.world_pos.x, surface.world_pos.y, surface.world_pos.z, tmin);
            w.st_ray1[slot] = make_float4(dir.x, dir.y, dir.z, tmax);
            w.st_acc[slot] = make_float4(brdf_over_pdf.x, brdf_over_pdf.y, brdf_over_pdf.z, overall_pdf);
            traced = tmin <= tmax;                                     // TMin > TMax: miss without a ray (oracle P2)
            if (!traced) w.st_hit[slot] = make_float4(INFINITY, 0.f, 0.f, __uint_as_float(0xFFFFFFFFu));
        }
        w.st_misc[slot] = make_uint4(flags, 0u, 0u, 0u);
    }
    warp_append(traced, slot, w.queue, w.count);
}

__global__ void __launch_bounds__(128) k_brdf_rays_shade(GIWaveArgs w) {
    const GIArgs& a = w.gi;
    __shared__ float lut[256];
    lut[threadIdx.x] = a.sc.srgb_lut[threadIdx.x]; lut[threadIdx.x + 128] = a.sc.srgb_lut[threadIdx.x + 128];
    __syncthreads();
    uint32_t lx, ly;
    if (!tile_pixel(blockIdx.x * blockDim.x + threadIdx.x, a.width, a.band, &lx, &ly)) return;
    const uint32_t slot = ly * a.width + lx;
    const uint32_t flags = w.st_misc[slot].x;
    uint32_t hits = 0;
    if (flags & GIW_VALID) {
        const rt2_gconst& g = *a.g;
        uint32_t px, py; launch_to_pixel(g, lx, ly, &px, &py);
        const bool is_specular_ray = (flags & GIW_SPECULAR) != 0u, is_delta = (flags & GIW_DELTA) != 0u;
        const float4 r0 = w.st_ray0[slot], r1 = w.st_ray1[slot], ac = w.st_acc[slot], h = w.st_hit[slot];
        const V3 origin = v3(r0.x, r0.y, r0.z), dir = v3(r1.x, r1.y, r1.z);
        const V3 brdf_over_pdf = v3(ac.x, ac.y, ac.z);
        const float overall_pdf = ac.w;
        const uint32_t tri_slot = __float_as_uint(h.w);
        V3 radiance = v3(0.f, 0.f, 0.f);
        V3 sec_position, sec_normal, sec_albedo, sec_f0; float sec_roughness; bool sec_is_env;
        const bool include_emissive = g.enable_restir_di == 0u || (is_specular_ray && is_delta);
        if (tri_slot != 0xFFFFFFFFu) {
            hits++;
            const SurfaceFetch sf = get_geometry_from_hit(a.sc, lut, tri_slot, a.sc.instances ? w.st_hit_inst[slot] : 0u, h.y, h.z);
            if (include_emissive) radiance = radiance + sf.emission;
            sec_position = origin + dir * h.x;
            sec_normal = (dot3(sf.normal, dir) < 0) ? sf.normal : neg3(sf.normal);
            sec_albedo = sf.color; sec_f0 = sf.specular_f0; sec_roughness = sf.roughness; sec_is_env = false;
        } else {
            if (include_emissive) radiance = radiance + environment_radiance(a.sc, g, dir);
            sec_position = origin + dir * 1000.0f;                     // DISTANT_LIGHT_DISTANCE, PolymorphicLight.glsl:43
            sec_normal = neg3(dir);
            sec_albedo = v3(0.f, 0.f, 0.f); sec_f0 = v3(0.f, 0.f, 0.f); sec_roughness = 0.0f; sec_is_env = true;
        }
        if (g.enable_brdf_indirect == 1u) {
            uint32_t fl = 0;
            if (is_specular_ray) fl |= RT2_SECONDARY_IS_SPECULAR_RAY;
            if (is_delta) fl |= RT2_SECONDARY_IS_DELTA_SURFACE;
            if (sec_is_env) fl |= RT2_SECONDARY_IS_ENVIRONMENT;
            const uint32_t ptr = reservoir_pointer(g.restir_gi.reservoir_buffer_params.reservoir_block_row_pitch, g.restir_gi.reservoir_buffer_params.reservoir_array_pitch, lx, ly, 0u);
            uint4* dst = a.secondary + 3 * (size_t)ptr;
            dst[0] = make_uint4(__float_as_uint(sec_position.x), __float_as_uint(sec_position.y), __float_as_uint(sec_position.z), ndir_to_oct_unorm32(sec_normal));
            dst[1] = make_uint4(pack_r16g16_float(brdf_over_pdf.x, brdf_over_pdf.y), pack_r16g16_float(brdf_over_pdf.z, 0.0f) | (fl << 16),
                                pack_r11g11b10(sec_albedo), pack_rgba8_gamma(sec_f0.x, sec_f0.y, sec_f0.z, sec_roughness));
            dst[2] = make_uint4(__float_as_uint(radiance.x), __float_as_uint(radiance.y), __float_as_uint(radiance.z), __float_as_uint(overall_pdf));
            if (g.enable_restir_gi == 1u) radiance = v3(0.f, 0.f, 0.f);
        }
        if (radiance.x > 0.0f || radiance.y > 0.0f || radiance.z > 0.0f || g.enable_brdf_additive_blend == 0u) {
            const V3 diffuse = is_specular_ray ? v3(0.f, 0.f, 0.f) : radiance * brdf_over_pdf;
            V3 specular = is_specular_ray ? radiance * brdf_over_pdf : v3(0.f, 0.f, 0.f);
            V3 f0 = v3(0.f, 0.f, 0.f);                                 // the primary surface's specularF0 is all the shading tail needs
            if ((float)px < g.view.viewport_size[0] && (float)py < g.view.viewport_size[1]) { const float4 sr = unpack_rgba8_gamma(a.gb.specular_rough[(size_t)py * a.width + px], a.gb.gamma_lut); f0 = v3(sr.x, sr.y, sr.z); }
            specular = demodulate_specular(f0, specular);
            if (px < a.width) store_shading_output(g, (size_t)py * a.width + px, diffuse, specular, false, a.diffuse, a.specular, a.diffuse_f32);
        }
    }
    flush_ray_stats(a.stat_rays, a.stat_hits, 0u, hits);
}

// K5, DI off (shade_secondary_surfaces.rgen:119-128 is then the only ray): every launch index traces one cosine bounce
// from its secondary surface.
__global__ void __launch_bounds__(128) k_shade_secondary_gen(GIWaveArgs w) {
    const GIArgs& a = w.gi;
    uint32_t lx, ly;
    const bool inside = tile_pixel(blockIdx.x * blockDim.x + threadIdx.x, a.width, a.band, &lx, &ly);
    uint32_t slot = 0;
    if (inside) {
        slot = ly * a.width + lx;
        const rt2_gconst& g = *a.g;
        Rng rng = rng_init(lx, ly, a.frame, 6u);
        const uint32_t gbuffer_index = reservoir_pointer(g.restir_di.reservoir_buffer_params.reservoir_block_row_pitch, g.restir_di.reservoir_buffer_params.reservoir_array_pitch, lx, ly, 0u);
        const uint4 s0 = a.secondary[3 * (size_t)gbuffer_index];
        Surface sec = empty_surface();                                 // surface_brdf_sample only reads the normal
        sec.normal = oct_to_ndir_unorm32(s0.w);
        V3 dir;
        surface_brdf_sample(sec, rng, &dir);
        w.st_ray0[slot] = make_float4(__uint_as_float(s0.x), __uint_as_float(s0.y), __uint_as_float(s0.z), 0.001f);
        w.st_ray1[slot] = make_float4(dir.x, dir.y, dir.z, 1000.0f);
    }
    warp_append(inside, slot, w.queue, w.count);
}

__global__ void __launch_bounds__(128) k_shade_secondary_shade(GIWaveArgs w) {
    const GIArgs& a = w.gi;
    __shared__ float lut[256];
    lut[threadIdx.x] = a.sc.srgb_lut[threadIdx.x]; lut[threadIdx.x + 128] = a.sc.srgb_lut[threadIdx.x + 128];
    __syncthreads();
    uint32_t lx, ly;
    if (!tile_pixel(blockIdx.x * blockDim.x + threadIdx.x, a.width, a.band, &lx, &ly)) return;
    const uint32_t slot = ly * a.width + lx;
    const rt2_gconst& g = *a.g;
    uint32_t px, py; launch_to_pixel(g, lx, ly, &px, &py);
    uint32_t hits = 0;
    const uint32_t gbuffer_index = reservoir_pointer(g.restir_di.reservoir_buffer_params.reservoir_block_row_pitch, g.restir_di.reservoir_buffer_params.reservoir_array_pitch, lx, ly, 0u);
    uint4* sgp = a.secondary + 3 * (size_t)gbuffer_index;
    const uint4 s0 = sgp[0], s1 = sgp[1], s2 = sgp[2];
    const V3 throughput = unpack_half3(s1.x, s1.y);
    const uint32_t flags = s1.y >> 16;
    const bool valid_secondary = throughput.x != 0 || throughput.y != 0 || throughput.z != 0;
    const bool is_specular_ray = (flags & RT2_SECONDARY_IS_SPECULAR_RAY) != 0u;
    const bool is_delta = (flags & RT2_SECONDARY_IS_DELTA_SURFACE) != 0u;
    V3 radiance = v3(__uint_as_float(s2.x), __uint_as_float(s2.y), __uint_as_float(s2.z));
    const float pdf = __uint_as_float(s2.w);
    const V3 sec_world_pos = v3(__uint_as_float(s0.x), __uint_as_float(s0.y), __uint_as_float(s0.z));
    const V3 sec_normal = oct_to_ndir_unorm32(s0.w);
    const V3 sec_albedo = unpack_r11g11b10(s1.z);
    {
        const float4 h = w.st_hit[slot];
        const uint32_t tri_slot = __float_as_uint(h.w);
        V3 emission;
        if (tri_slot == 0xFFFFFFFFu) { const float4 r1 = w.st_ray1[slot]; emission = environment_radiance(a.sc, g, v3(r1.x, r1.y, r1.z)); }   // Q2 / oracle P4
        else { hits++; emission = get_geometry_from_hit(a.sc, lut, tri_slot, a.sc.instances ? w.st_hit_inst[slot] : 0u, h.y, h.z).emission; }
        radiance = radiance + emission * sec_albedo;
    }
    GIReservoir reservoir = empty_gi_reservoir();
    const bool output_shading_result = is_specular_ray && is_delta;
    if (valid_secondary && !output_shading_result) reservoir = make_gi_reservoir(sec_world_pos, sec_normal, radiance, pdf);
    const uint32_t rx = g.runtime_params.active_checkerboard_field != 0u ? (px >> 1) : px;
    store_gi_reservoir(a.reservoirs, reservoir_pointer(g.restir_gi.reservoir_buffer_params.reservoir_block_row_pitch, g.restir_gi.reservoir_buffer_params.reservoir_array_pitch,
                                                       rx, py, g.restir_gi.buffer_indices.secondary_surface_restirdi_output_buffer_index), reservoir, 0u);
    const V3 saved = output_shading_result ? v3(0.f, 0.f, 0.f) : radiance;
    sgp[2] = make_uint4(__float_as_uint(saved.x), __float_as_uint(saved.y), __float_as_uint(saved.z), s2.w);
    if (output_shading_result) {
        const Surface primary = gbuffer_surface(a.gb, a.width, (int)px, (int)py, g.view);
        const V3 specular = demodulate_specular(primary.specular_f0, radiance);
        if (px < a.width) store_shading_output(g, (size_t)py * a.width + px, v3(0.f, 0.f, 0.f), specular, false, a.diffuse, a.specular, a.diffuse_f32);
    }
    flush_ray_stats(a.stat_rays, a.stat_hits, 0u, hits);
}

// K5 with DI on, last stage (shade_secondary_surfaces.rgen:105-156): shade the secondary surface with the sampled light
// (visibility from the ray-queue kernel), firefly clamp, initial GI reservoir, MIS radiance hand-over.
__global__ void __launch_bounds__(128) k_shade_secondary_stage_d(DIWaveArgs w, GIArgs a) {
    uint32_t lx, ly;
    if (!tile_pixel(blockIdx.x * blockDim.x + threadIdx.x, a.width, a.band, &lx, &ly)) return;
    const uint32_t slot = ly * a.width + lx;
    const rt2_gconst& g = *a.g;
    const rt2_restirdi_params& di = g.restir_di;
    uint32_t px, py; launch_to_pixel(g, lx, ly, &px, &py);
    uint32_t hits = 0;
    const uint32_t gbuffer_index = reservoir_pointer(di.reservoir_buffer_params.reservoir_block_row_pitch, di.reservoir_buffer_params.reservoir_array_pitch, lx, ly, 0u);
    uint4* sgp = a.secondary + 3 * (size_t)gbuffer_index;
    const uint4 s0 = sgp[0], s1 = sgp[1], s2 = sgp[2];
    const V3 throughput = unpack_half3(s1.x, s1.y);
    const uint32_t sflags = s1.y >> 16;
    const bool valid_secondary = throughput.x != 0 || throughput.y != 0 || throughput.z != 0;
    const bool is_specular_ray = (sflags & RT2_SECONDARY_IS_SPECULAR_RAY) != 0u;
    const bool is_delta = (sflags & RT2_SECONDARY_IS_DELTA_SURFACE) != 0u;
    V3 radiance = v3(__uint_as_float(s2.x), __uint_as_float(s2.y), __uint_as_float(s2.z));
    const float pdf = __uint_as_float(s2.w);
    const V3 sec_world_pos = v3(__uint_as_float(s0.x), __uint_as_float(s0.y), __uint_as_float(s0.z));
    const V3 sec_normal = oct_to_ndir_unorm32(s0.w);
    DIReservoir reservoir; LightSample light_sample; uint32_t flags, rng_index;
    di_state_load(w, slot, &reservoir, &light_sample, &rng_index, &flags);
    if (flags & DIW_ACTIVE) {
        bool active;
        const Surface sec = di_stage_surface<true>(w, g, lx, ly, px, py, &active);
        V3 indirect_diffuse = v3(0.f, 0.f, 0.f), indirect_specular = v3(0.f, 0.f, 0.f);
        if (!(light_sample.solid_angle_pdf <= 0)) {                    // ShadeSurfaceWithLightSample, ShadingHelpers.glsl:2-58
            if (di.shading_params.enable_final_visibility == 1u) {
                const bool vis = !((flags & DIW_RAY_PENDING) && __float_as_uint(w.st_hit[slot].w) != 0xFFFFFFFFu);
                if (!vis) hits++;
                const V3 visibility = vis ? v3(1.f, 1.f, 1.f) : v3(0.f, 0.f, 0.f);
                di_store_visibility(reservoir, visibility, di.temporal_resampling_params.discard_invisible_samples == 1u);
                light_sample.radiance = light_sample.radiance * visibility;
            }
            light_sample.radiance = light_sample.radiance * (reservoir.weight_sum / light_sample.solid_angle_pdf);
            if (light_sample.radiance.x > 0.0f || light_sample.radiance.y > 0.0f || light_sample.radiance.z > 0.0f) {
                const SplitBrdf brdf = evaluate_brdf(sec, light_sample.position);
                indirect_diffuse = light_sample.radiance * brdf.demodulated_diffuse;
                indirect_specular = brdf.specular * light_sample.radiance;
            }
        }
        radiance = radiance + (indirect_diffuse * sec.diffuse_albedo + indirect_specular);
        const float c_max_indirect_radiance = 100.0f;                  // firefly suppression, :113-116
        const float lum = calc_luminance(radiance);
        if (lum > c_max_indirect_radiance) radiance = radiance * (c_max_indirect_radiance / lum);
    }
    GIReservoir gi_reservoir = empty_gi_reservoir();
    const bool output_shading_result = is_specular_ray && is_delta;
    if (valid_secondary && !output_shading_result) gi_reservoir = make_gi_reservoir(sec_world_pos, sec_normal, radiance, pdf);
    const uint32_t rx = g.runtime_params.active_checkerboard_field != 0u ? (px >> 1) : px;
    store_gi_reservoir(a.reservoirs, reservoir_pointer(g.restir_gi.reservoir_buffer_params.reservoir_block_row_pitch, g.restir_gi.reservoir_buffer_params.reservoir_array_pitch,
                                                       rx, py, g.restir_gi.buffer_indices.secondary_surface_restirdi_output_buffer_index), gi_reservoir, 0u);
    const V3 saved = output_shading_result ? v3(0.f, 0.f, 0.f) : radiance;
    sgp[2] = make_uint4(__float_as_uint(saved.x), __float_as_uint(saved.y), __float_as_uint(saved.z), s2.w);
    if (output_shading_result) {
        const Surface primary = gbuffer_surface(a.gb, a.width, (int)px, (int)py, g.view);
        const V3 specular = demodulate_specular(primary.specular_f0, radiance);
        if (px < a.width) store_shading_output(g, (size_t)py * a.width + px, v3(0.f, 0.f, 0.f), specular, false, a.diffuse, a.specular, a.diffuse_f32);
    }
    flush_ray_stats(a.stat_rays, a.stat_hits, 0u, hits);
}

// ------------------------------------------------------------------------------------------------
// N2 — GI temporal (K6) and spatial (K7) resampling: lighting_passes/temporal_resampling.rgen,
// spatial_resampling.rgen, rtxdi/GIResamplingFunctions.hlsli:186-553.  The first kernels with cross-pixel reads:
// neighbours come from last frame's G-buffer / reservoir page (K6) or from this frame's (K7); the page a kernel
// writes is never the page it gathers from (checked on the host), so stream order is the only dependency.
// ------------------------------------------------------------------------------------------------
struct GIResampleArgs {
    SceneRefs sc; const rt2_gconst* g; uint32_t width, height, frame; BandRef band;
    GBufferRefs gb, gb_prev;
    const float4* motion;
    const char2* neighbor_offsets;     // signed byte pairs, SNORM8 (SURVEY Q11 pin)
    uint4* reservoirs;
    unsigned long long* stat_rays; unsigned long long* stat_hits;
};

__global__ void __launch_bounds__(128, RT2_GIRS_BLOCKS) k_gi_temporal(GIResampleArgs a) {
    uint32_t lx, ly;
    if (!tile_pixel(blockIdx.x * blockDim.x + threadIdx.x, a.width, a.band, &lx, &ly)) return;
    const rt2_gconst& g = *a.g;
    const rt2_restirgi_params& gi = g.restir_gi;
    uint32_t px, py; launch_to_pixel(g, lx, ly, &px, &py);
    uint32_t rays = 0, hits = 0;
    Rng rng = rng_init(lx, ly, a.frame, 7u);
    const Surface surface = gbuffer_surface(a.gb, a.width, (int)px, (int)py, g.view);
    const uint32_t field = g.runtime_params.active_checkerboard_field;
    const uint32_t rx = field != 0u ? (px >> 1) : px;
    const uint32_t row_pitch = gi.reservoir_buffer_params.reservoir_block_row_pitch, array_pitch = gi.reservoir_buffer_params.reservoir_array_pitch;
    GIReservoir reservoir = load_gi_reservoir(a.reservoirs, reservoir_pointer(row_pitch, array_pitch, rx, py, gi.buffer_indices.secondary_surface_restirdi_output_buffer_index), a.sc.logluv_lut);
    V3 motion = v3(0.f, 0.f, 0.f);
    if (px < a.width && py < a.height) { const float4 m = a.motion[(size_t)py * a.width + px]; motion = v3(m.x, m.y, m.z); }
    {   // convertMotionVectorToPixelSpace, GBufferHelpers.glsl:69-80
        const float cx = (float)(int)px + 0.5f, cy = (float)(int)py + 0.5f;
        float ppx = cx + motion.x, ppy = cy + motion.y;
        ppx *= g.prev_view.viewport_size[0] * g.view.viewport_size_inv[0];
        ppy *= g.prev_view.viewport_size[1] * g.view.viewport_size_inv[1];
        motion.x = ppx - cx; motion.y = ppy - cy;
    }
    if (surface.view_depth != RT2_BACKGROUND_DEPTH) {
        const rt2_restirgi_temporal_params& tp = gi.temporal_resampling_params;
        const float age_scale = (float)tp.max_reservoir_age * (0.5f + rng_next(rng) * 0.5f);
        const uint32_t max_reservoir_age = isnan(age_scale) ? 0u : (uint32_t)age_scale;
        const GIReservoir input = reservoir;
        int prev_x = (int)rintf((float)px + motion.x), prev_y = (int)rintf((float)py + motion.y);
        const float expected_prev_depth = surface.view_depth + motion.z;
        const int radius = (field == 0u) ? 1 : 2;
        GIReservoir temporal = empty_gi_reservoir();
        bool found = false;
        const int start_idx = (int)(rng_next(rng) * 8);
        Surface temporal_surface = empty_surface();
        const int temporal_sample_count = 5;
        const int sample_count = temporal_sample_count + (tp.enable_fallback_sampling == 1u ? 1 : 0);
        for (int i = 0; i < sample_count; i++) {
            const bool is_first = i == 0, is_fallback = i == temporal_sample_count;
            int ox = 0, oy = 0;
            if (is_fallback) { prev_x = (int)px; prev_y = (int)py; }
            else if (!is_first) temporal_resampling_offset(start_idx + i, radius, &ox, &oy);
            int ix = prev_x + ox, iy = prev_y + oy;
            if ((tp.enable_permutation_sampling == 1u && is_first) || is_fallback) apply_permutation_sampling(&ix, &iy, tp.uniform_random_number);
            activate_checkerboard_pixel(&ix, &iy, true, field);
            temporal_surface = gbuffer_surface(a.gb_prev, a.width, ix, iy, g.prev_view);
            if (temporal_surface.view_depth == RT2_BACKGROUND_DEPTH) continue;
            if (!is_fallback && !is_valid_neighbor(surface.normal, temporal_surface.normal, expected_prev_depth, temporal_surface.view_depth, tp.normal_threshold, tp.depth_threshold)) continue;
            if (!materials_similar(surface, temporal_surface)) continue;
            const uint32_t prx = field != 0u ? ((uint32_t)ix >> 1) : (uint32_t)ix;
            temporal = load_gi_reservoir(a.reservoirs, reservoir_pointer(row_pitch, array_pitch, prx, (uint32_t)iy, gi.buffer_indices.temporal_resampling_input_buffer_index), a.sc.logluv_lut);
            if (temporal.M == 0u) continue;
            found = true;
            break;
        }
        GIReservoir cur = empty_gi_reservoir();
        float selected_target_pdf = 0;
        if (input.M != 0u) {
            selected_target_pdf = gi_target_pdf(input.position, input.radiance, surface);
            gi_combine(cur, input, 0.5f, selected_target_pdf);
        }
        if (found) {
            float jacobian = gi_jacobian(surface.world_pos, temporal_surface.world_pos, temporal);
            if (!validate_gi_jacobian(jacobian)) found = false;
            temporal.weight_sum *= jacobian;
            temporal.M = min(temporal.M, tp.max_history_length);
            ++temporal.age;
            if (temporal.age > max_reservoir_age) found = false;
        }
        bool selected_previous = false;
        if (found) {
            const float target_pdf = gi_target_pdf(temporal.position, temporal.radiance, surface);
            selected_previous = gi_combine(cur, temporal, rng_next(rng), target_pdf);
            if (selected_previous) selected_target_pdf = target_pdf;
        }
        if (tp.temporal_bias_correction_mode >= 1u) {                  // RTXDI_BIAS_CORRECTION_BASIC
            float pi = selected_target_pdf;
            float pi_sum = selected_target_pdf * (float)input.M;
            if (cur.M != 0u && found) {
                float temporal_p = gi_target_pdf(cur.position, cur.radiance, temporal_surface);
                if (tp.temporal_bias_correction_mode == 3u && temporal_p > 0) {        // RAY_TRACED: bridge :708-711, current BVH
                    if (!visibility_ray_miss(a.sc, temporal_surface, cur.position, 0.001f, &rays)) { hits++; temporal_p = 0; }
                }
                pi = selected_previous ? temporal_p : pi;
                pi_sum += temporal_p * (float)temporal.M;
            }
            gi_finalize(cur, pi, pi_sum * selected_target_pdf);
        } else {
            gi_finalize(cur, 1.0f, selected_target_pdf * (float)cur.M);
        }
        reservoir = cur;
    }
    store_gi_reservoir(a.reservoirs, reservoir_pointer(row_pitch, array_pitch, rx, py, gi.buffer_indices.temporal_resampling_output_buffer_index), reservoir, 0u);
    flush_ray_stats(a.stat_rays, a.stat_hits, rays, hits);
}

__device__ __forceinline__ void spatial_offset(const char2* table, int sample_idx, float radius, uint32_t mask, int* ox, int* oy) {   // :132-136
    sample_idx &= (int)mask;
    const char2 b = table[sample_idx];
    const float fx = fmaxf((float)b.x / 127.0f, -1.0f), fy = fmaxf((float)b.y / 127.0f, -1.0f);
    *ox = (int)(fx * radius); *oy = (int)(fy * radius);
}

__global__ void __launch_bounds__(128, RT2_GIRS_BLOCKS) k_gi_spatial(GIResampleArgs a) {
    uint32_t lx, ly;
    if (!tile_pixel(blockIdx.x * blockDim.x + threadIdx.x, a.width, a.band, &lx, &ly)) return;
    const rt2_gconst& g = *a.g;
    const rt2_restirgi_params& gi = g.restir_gi;
    uint32_t px, py; launch_to_pixel(g, lx, ly, &px, &py);
    uint32_t rays = 0, hits = 0;
    Rng rng = rng_init(lx, ly, a.frame, 8u);
    const Surface surface = gbuffer_surface(a.gb, a.width, (int)px, (int)py, g.view);
    const uint32_t field = g.runtime_params.active_checkerboard_field, mask = g.runtime_params.neighbor_offset_mask;
    const uint32_t rx = field != 0u ? (px >> 1) : px;
    const uint32_t row_pitch = gi.reservoir_buffer_params.reservoir_block_row_pitch, array_pitch = gi.reservoir_buffer_params.reservoir_array_pitch;
    const uint32_t src = gi.buffer_indices.spatial_resampling_input_buffer_index;
    GIReservoir reservoir = load_gi_reservoir(a.reservoirs, reservoir_pointer(row_pitch, array_pitch, rx, py, src), a.sc.logluv_lut);
    if (surface.view_depth != RT2_BACKGROUND_DEPTH) {
        const rt2_restirgi_spatial_params& sp = gi.spatial_resampling_params;
        const int num_samples = (int)sp.num_spatial_samples;
        const GIReservoir input = reservoir;
        GIReservoir cur = empty_gi_reservoir();
        float selected_target_pdf = 0;
        if (input.M != 0u) {
            selected_target_pdf = gi_target_pdf(input.position, input.radiance, surface);
            gi_combine(cur, input, 0.5f, selected_target_pdf);
        }
        uint32_t cached = 0;
        int selected = -1;
        const int start_idx = (int)(rng_next(rng) * (float)mask);
        for (int i = 0; i < num_samples; ++i) {
            int ox, oy; spatial_offset(a.neighbor_offsets, start_idx + i, sp.spatial_sampling_radius, mask, &ox, &oy);
            int ix = (int)px + ox, iy = (int)py + oy;
            clamp_sample_position_into_view(g, &ix, &iy);
            activate_checkerboard_pixel(&ix, &iy, false, field);
            const Surface ns = gbuffer_surface(a.gb, a.width, ix, iy, g.view);
            if (!is_valid_neighbor(surface.normal, ns.normal, surface.view_depth, ns.view_depth, sp.spatial_normal_threshold, sp.spatial_depth_threshold)) continue;
            if (!materials_similar(surface, ns)) continue;
            const uint32_t nrx = field != 0u ? ((uint32_t)ix >> 1) : (uint32_t)ix;
            const GIReservoir nr = load_gi_reservoir(a.reservoirs, reservoir_pointer(row_pitch, array_pitch, nrx, (uint32_t)iy, src), a.sc.logluv_lut);
            if (nr.M == 0u) continue;
            float jacobian = gi_jacobian(surface.world_pos, ns.world_pos, nr);
            const float target_pdf = gi_target_pdf(nr.position, nr.radiance, surface);
            if (!validate_gi_jacobian(jacobian)) continue;
            cached |= (1u << (uint32_t)i);
            if (gi_combine(cur, nr, rng_next(rng), target_pdf * jacobian)) { selected = i; selected_target_pdf = target_pdf; }
        }
        if (sp.spatial_bias_correction_mode >= 1u) {
            float pi = selected_target_pdf;
            float pi_sum = selected_target_pdf * (float)input.M;
            for (int i = 0; i < num_samples; ++i) {
                if ((cached & (1u << (uint32_t)i)) == 0u) continue;
                int ox, oy; spatial_offset(a.neighbor_offsets, start_idx + i, sp.spatial_sampling_radius, mask, &ox, &oy);
                int ix = (int)px + ox, iy = (int)py + oy;
                clamp_sample_position_into_view(g, &ix, &iy);
                activate_checkerboard_pixel(&ix, &iy, false, field);
                const Surface ns = gbuffer_surface(a.gb, a.width, ix, iy, g.view);
                const uint32_t nrx = field != 0u ? ((uint32_t)ix >> 1) : (uint32_t)ix;
                const uint4 head = a.reservoirs[2 * (size_t)reservoir_pointer(row_pitch, array_pitch, nrx, (uint32_t)iy, src)];   // only M is needed again
                float ps = gi_target_pdf(cur.position, cur.radiance, ns);
                if (sp.spatial_bias_correction_mode == 3u && ps > 0) {
                    if (!visibility_ray_miss(a.sc, ns, cur.position, 0.001f, &rays)) { hits++; ps = 0; }
                }
                pi = selected == i ? ps : pi;
                pi_sum += ps * (float)(head.w & 0xffu);
            }
            gi_finalize(cur, pi, selected_target_pdf * pi_sum);
        } else {
            gi_finalize(cur, 1.0f, (float)cur.M * selected_target_pdf);
        }
        reservoir = cur;
    }
    store_gi_reservoir(a.reservoirs, reservoir_pointer(row_pitch, array_pitch, rx, py, gi.buffer_indices.spatial_resampling_output_buffer_index), reservoir, 0u);
    flush_ray_stats(a.stat_rays, a.stat_hits, rays, hits);
}

// ------------------------------------------------------------------------------------------------
// host: LightPasses::execute, non-reference branch (src/light_passes.rs:598-659)
// ------------------------------------------------------------------------------------------------
int render_restir(rt2_ctx* ctx, uint32_t frame, const rt2_launch_params& lp) {
    if (!ctx->scene.accel_built) return fail(ctx, RT2_E_STATE, "no acceleration structure: call rt2_upload_scene + rt2_build_accel first");
    const rt2_gconst& g = ctx->gconst;
    if (lp.band_world > 1u) {
        if (lp.band_rows == 0u || lp.band_rank >= lp.band_world) return fail(ctx, RT2_E_INVALID, "bad band parameters");
        // K3 / K9 / K4 / K5 / K8 are per-pixel independent and shard by rows with no exchange; K6 / K7 read their
        // neighbours' G-buffer and reservoirs across band edges and would need a halo exchange between the ranks
        // across band edges (K6 anywhere in the previous frame): rt2_group_render runs the frame in phases and exchanges those inputs between
        // them (ctx->restir_phase, rt2_group.cu); a plain rt2_render of one band set has nobody to exchange with
        if (ctx->restir_phase == 0u && g.enable_restir_gi == 1u && (g.enable_temporal_resampling == 1u || g.enable_spatial_resampling == 1u))
            return fail(ctx, RT2_E_INVALID, "GI temporal / spatial resampling reads neighbouring rows: with row-band sharding it needs the halo exchange rt2_group_render performs");
    }
    // phases of a sharded frame with resampling: 1 = K3, K9, K4, K5; 2 = K6; 4 = K7, K8.  0 (a plain rt2_render) = everything.
    const uint32_t ph = ctx->restir_phase ? ctx->restir_phase : 7u;
    int rc = restir_ensure_resources(ctx); if (rc) return rc;
    FrameBuffers& f = ctx->fb;
    // counters: [0..1] rays, [2..3] hits
    // ... [4..15] the ray queues' (count, head) pairs, [16..17] u64 trace() calls answered without a walk (K9 visibility reuse)
    if (f.counter_capacity < 20) {
        if (f.counters) { RT2_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); cudaFree(f.counters); f.counters = nullptr; }
        RT2_CUDA(ctx, cudaMalloc((void**)&f.counters, 20 * 4)); f.counter_capacity = 20;
    }
    if (ph & 1u) RT2_CUDA(ctx, cudaMemsetAsync(f.counters, 0, 80, ctx->stream));
    unsigned long long* stat_rays = reinterpret_cast<unsigned long long*>(f.counters);
    unsigned long long* stat_hits = reinterpret_cast<unsigned long long*>(f.counters + 2);
    const uint32_t cur = frame & 1u;                   // ping-pong, src/light_passes.rs:565-577
    ctx->last_gbuffer_index = cur;
    BandRef band{lp.band_rows, lp.band_rank, lp.band_world, f.height};
    if (band.world <= 1u || band.band_rows == 0u) { band.world = 1; band.rank = 0; band.band_rows = f.height; }
    else {
        uint32_t rows = 0;
        for (uint32_t y0 = band.rank * band.band_rows; y0 < f.height; y0 += band.world * band.band_rows) rows += std::min(band.band_rows, f.height - y0);
        band.local_rows = rows;
    }
    const uint32_t tiles_x = (f.width + 7) / 8, tiles_y = (band.local_rows + 3) / 4;
    const uint32_t padded = tiles_x * tiles_y * 32;
    const uint32_t blocks = std::max(1u, (padded + 127) / 128);
    uint64_t launches = 0;
    auto pitches_ok = [&](const rt2_reservoir_buffer_params& p) {
        return p.reservoir_array_pitch == f.reservoir_array_pitch && p.reservoir_block_row_pitch == f.reservoir_block_row_pitch;
    };
    // ---- validation of everything the enabled passes will index with -------------------------------------------
    LightRefs lrefs{};
    if (g.enable_restir_di == 1u) {
        const uint32_t num_local = g.light_buffer_params.local_light_buffer_region.num_lights;
        if (g.light_buffer_params.infinite_light_buffer_region.num_lights != 0u) return fail(ctx, RT2_E_INVALID, "infinite lights have no producer in the reference (src/main.rs:375-380)");
        const uint32_t num_env = g.restir_di.initial_sampling_params.num_primary_environment_samples;
        if (num_env != 0u && !(lp.flags & RT2_LAUNCH_PRESAMPLING))
            return fail(ctx, RT2_E_INVALID, "environment RIS sampling is compiled out of the shipped shaders (RTXDI_ENABLE_PRESAMPLING 0): pass RT2_LAUNCH_PRESAMPLING");
        if (!pitches_ok(g.restir_di.reservoir_buffer_params)) return fail(ctx, RT2_E_INVALID, "GConst reservoir buffer parameters do not match the frame size (calculate_reservoir_buffer_parameters)");
        if (g.restir_di.buffer_indices.shading_input_buffer_index > 1u) return fail(ctx, RT2_E_INVALID, "reservoir buffer index out of range");
        if (num_local) {
            if (!ctx->lights_prepared) return fail(ctx, RT2_E_STATE, "light list not built: call rt2_prepare_lights (+ rt2_generate_mips(0)) first");
            if (g.light_buffer_params.local_light_buffer_region.first_light_index + (uint64_t)num_local > ctx->light_capacity) return fail(ctx, RT2_E_INVALID, "local light region exceeds the light buffer");
        }
        if ((rc = lights_ensure_resources(ctx))) return rc;
        lrefs.lights = ctx->light_info; lrefs.geom_to_light = ctx->geom_to_light;
        lrefs.pdf = ctx->local_pdf; lrefs.pdf_w = ctx->local_pdf_w; lrefs.pdf_h = ctx->local_pdf_h;
        const uint32_t m = std::max(ctx->local_pdf_w, ctx->local_pdf_h);
        uint32_t last = 0; while ((2u << last) <= m) last++;           // floor(log2(m))
        last = std::min(last, ctx->local_pdf_mips - 1);
        uint32_t off = 0;
        for (uint32_t k = 0; k < last; k++) off += std::max(1u, ctx->local_pdf_w >> k) * std::max(1u, ctx->local_pdf_h >> k);
        lrefs.pdf_last_offset = off;
        if (lp.flags & RT2_LAUNCH_PRESAMPLING) {
            lrefs.presampling = 1u;
            const uint32_t mode = g.restir_di.initial_sampling_params.local_light_sampling_mode;
            if (mode > 1u) return fail(ctx, RT2_E_INVALID, "local_light_sampling_mode 2 (ReGIR) is compiled out of the reference (RTXDI_REGIR_MODE disabled)");
            if (mode == 1u) {
                const rt2_ris_buffer_segment_params& seg = g.local_lights_risbuffer_segment_params;
                if (!ctx->ris_buffer || !ctx->ris_light_data) return fail(ctx, RT2_E_STATE, "POWER_RIS light sampling needs the presampled tiles: call rt2_presample(ctx, frame, 0) first");
                if (seg.tile_size == 0u || seg.tile_count == 0u || (uint64_t)seg.buffer_offset + (uint64_t)seg.tile_size * seg.tile_count > ctx->ris_entries)
                    return fail(ctx, RT2_E_INVALID, "local-light RIS segment exceeds the RIS buffer");
                lrefs.ris = ctx->ris_buffer; lrefs.ris_light_data = ctx->ris_light_data;
            }
            // the environment light: environmentLightParams.lightPresent (always 1 in the reference, src/main.rs:381-386) AND a sky in the scene
            const rt2_environment_light_buffer_params& ep = g.light_buffer_params.environment_light_params;
            if (ep.light_present != 0u && ctx->scene.sky != nullptr) {
                if (!ctx->lights_prepared || ctx->light_records != ctx->light_capacity + 2u)
                    return fail(ctx, RT2_E_STATE, "the environment light is not written: call rt2_prepare_lights first");
                if (ep.light_index != ctx->light_capacity + 1u) return fail(ctx, RT2_E_INVALID, "environment light index must be lights + 1 (src/main.rs:383): rt2_prepare_lights writes the environment light there");
                if (!ctx->env_pdf) return fail(ctx, RT2_E_STATE, "environment PDF pyramid not built: call rt2_generate_mips(1) first");
                if (g.environment_pdf_texture_size[0] != ctx->env_pdf_w || g.environment_pdf_texture_size[1] != ctx->env_pdf_h)
                    return fail(ctx, RT2_E_INVALID, "GConst environment_pdf_texture_size does not match the environment PDF texture");
                lrefs.env_present = 1u; lrefs.env_index = ep.light_index;
                lrefs.env_pdf = ctx->env_pdf; lrefs.env_w = ctx->env_pdf_w; lrefs.env_h = ctx->env_pdf_h;
                const uint32_t em = std::max(ctx->env_pdf_w, ctx->env_pdf_h);
                uint32_t elast = 0; while ((2u << elast) <= em) elast++;
                elast = std::min(elast, ctx->env_pdf_mips - 1);
                uint32_t eoff = 0;
                for (uint32_t k = 0; k < elast; k++) eoff += std::max(1u, ctx->env_pdf_w >> k) * std::max(1u, ctx->env_pdf_h >> k);
                lrefs.env_last_offset = eoff;
                if (num_env != 0u) {
                    const rt2_ris_buffer_segment_params& seg = g.environment_light_risbuffer_segment_params;
                    if (!ctx->ris_buffer) return fail(ctx, RT2_E_STATE, "environment-map sampling needs the presampled tiles: call rt2_presample(ctx, frame, 1) first");
                    if (seg.tile_size == 0u || seg.tile_count == 0u || (uint64_t)seg.buffer_offset + (uint64_t)seg.tile_size * seg.tile_count > ctx->ris_entries)
                        return fail(ctx, RT2_E_INVALID, "environment RIS segment exceeds the RIS buffer");
                    lrefs.ris = ctx->ris_buffer;
                }
            }
        }
    }
    if (g.enable_restir_gi == 1u) {
        if (!pitches_ok(g.restir_gi.reservoir_buffer_params) || !pitches_ok(g.restir_di.reservoir_buffer_params))   // K5 addresses the secondary G-buffer with the DI pitches
            return fail(ctx, RT2_E_INVALID, "GConst reservoir buffer parameters do not match the frame size (calculate_reservoir_buffer_parameters)");
        if (g.restir_gi.buffer_indices.secondary_surface_restirdi_output_buffer_index > 1u) return fail(ctx, RT2_E_INVALID, "GI reservoir buffer index out of range");
        if (g.runtime_params.active_checkerboard_field > 2u) return fail(ctx, RT2_E_INVALID, "active_checkerboard_field must be 0, 1 or 2");
        const rt2_restirgi_buffer_indices& bi = g.restir_gi.buffer_indices;
        if (g.enable_temporal_resampling == 1u) {
            if (bi.temporal_resampling_input_buffer_index > 1u || bi.temporal_resampling_output_buffer_index > 1u) return fail(ctx, RT2_E_INVALID, "GI reservoir buffer index out of range");
            if (bi.temporal_resampling_input_buffer_index == bi.temporal_resampling_output_buffer_index)
                return fail(ctx, RT2_E_INVALID, "temporal resampling would gather from the reservoir page it writes (a data race in the reference too)");
        }
        if (g.enable_spatial_resampling == 1u) {
            if (bi.spatial_resampling_input_buffer_index > 1u || bi.spatial_resampling_output_buffer_index > 1u) return fail(ctx, RT2_E_INVALID, "GI reservoir buffer index out of range");
            if (bi.spatial_resampling_input_buffer_index == bi.spatial_resampling_output_buffer_index)
                return fail(ctx, RT2_E_INVALID, "spatial resampling would gather from the reservoir page it writes (a data race in the reference too)");
            if (g.runtime_params.neighbor_offset_mask >= RT2_NEIGHBOR_OFFSET_COUNT) return fail(ctx, RT2_E_INVALID, "neighbor_offset_mask exceeds the neighbour-offset table (8192 entries, src/main.rs:56)");
            if (g.restir_gi.spatial_resampling_params.num_spatial_samples > 32u) return fail(ctx, RT2_E_INVALID, "num_spatial_samples must be <= 32 (bit mask of accepted neighbours)");
        }
    }
    // ---- K3
    GBufferArgs ga{};
    ga.sc = make_scene_refs(ctx); ga.g = ctx->d_gconst; ga.width = f.width; ga.height = f.height; ga.band = band;
    ga.depth = f.gb_depth[cur]; ga.normals = f.gb_normals[cur]; ga.geo_normals = f.gb_geo_normals[cur]; ga.diffuse_albedo = f.gb_diffuse_albedo[cur];
    ga.specular_rough = f.gb_specular_rough[cur]; ga.emissive = f.gb_emissive[cur]; ga.motion = f.motion_vectors; ga.stat_hits = stat_hits;
    if (ph & 1u) { k_gbuffer<<<blocks, 128, 0, ctx->stream>>>(ga); launches++; }
    uint64_t static_rays = (uint64_t)f.width * band.local_rows;      // K3: one primary ray per owned pixel
    const GBufferRefs gbrefs{f.gb_depth[cur], f.gb_normals[cur], f.gb_geo_normals[cur], f.gb_diffuse_albedo[cur], f.gb_specular_rough[cur], ctx->codec_luts};
    // ---- K9
    if (g.enable_restir_di == 1u && (ph & 1u)) {
        DIArgs da{};
        da.sc = ga.sc; da.g = ctx->d_gconst; da.width = f.width; da.height = f.height; da.frame = frame; da.band = band;
        da.gb = gbrefs; da.li = lrefs;
        da.reservoirs = f.di_reservoirs; da.diffuse = f.diffuse_f16; da.specular = f.specular_f16; da.diffuse_f32 = f.diffuse_f32;
        da.stat_rays = stat_rays; da.stat_hits = stat_hits;
        da.stat_reused = reinterpret_cast<unsigned long long*>(f.counters + 16); da.visibility_reuse = ctx->visibility_reuse ? 1u : 0u;
        if (ctx->gi_megakernels || g.restir_di.initial_sampling_params.num_primary_brdf_samples > 1u) {
            k_di_fused<<<blocks, 128, 0, ctx->stream>>>(da); launches++;
        } else {
            // counters: [8],[9] BRDF rays, [10],[11] first visibility stage, [12],[13] second
            DIWaveArgs dw{};
            dw.di = da; dw.st_ray0 = f.st_ray0; dw.st_ray1 = f.st_ray1; dw.st_hit = f.st_hit; dw.st_hit_inst = f.st_hit_inst;
            dw.state = f.di_state; dw.plane = (uint32_t)((size_t)f.width * f.height);
            dw.queue = f.queue0; dw.count = f.counters + 8;
            const bool initial_vis = g.restir_di.initial_sampling_params.enable_initial_visibility == 1u;
            const bool final_vis = g.restir_di.shading_params.enable_final_visibility == 1u;
            k_di_stage_a<false><<<blocks, 128, 0, ctx->stream>>>(dw); launches++;
            if (g.restir_di.initial_sampling_params.num_primary_brdf_samples != 0u) {
                if ((rc = trace_queue_generic(ctx, f.queue0, f.counters + 8, f.counters + 9, stat_rays))) return rc;
                launches++;
            }
            k_di_stage_b<false><<<blocks, 128, 0, ctx->stream>>>(dw, f.queue1, f.counters + 10); launches++;
            if (initial_vis || final_vis) { if ((rc = trace_queue_generic(ctx, f.queue1, f.counters + 10, f.counters + 11, stat_rays, true))) return rc; launches++; }
            if (initial_vis) {
                k_di_stage_c<<<blocks, 128, 0, ctx->stream>>>(dw, f.queue0, f.counters + 12); launches++;
                if (final_vis) { if ((rc = trace_queue_generic(ctx, f.queue0, f.counters + 12, f.counters + 13, stat_rays, true))) return rc; launches++; }
            }
            k_di_stage_d<<<blocks, 128, 0, ctx->stream>>>(dw, final_vis ? 1u : 0u); launches++;
        }
    }
    // ---- K4 -> K5 -> K8 (stream order is the ALL_COMMANDS barrier of src/light_passes.rs:621-659)
    if (g.enable_restir_gi == 1u) {
        GIArgs gi{};
        gi.sc = ga.sc; gi.g = ctx->d_gconst; gi.width = f.width; gi.height = f.height; gi.frame = frame; gi.band = band;
        gi.gb = gbrefs; gi.li = lrefs;
        gi.secondary = f.secondary_gbuffer; gi.reservoirs = f.gi_reservoirs;
        gi.diffuse = f.diffuse_f16; gi.specular = f.specular_f16; gi.diffuse_f32 = f.diffuse_f32;
        gi.stat_rays = stat_rays; gi.stat_hits = stat_hits;
        if (!(ph & 1u)) {
            // K4 / K5 ran in an earlier phase of this frame
        } else if (ctx->gi_megakernels) {
            k_brdf_rays<<<blocks, 128, 0, ctx->stream>>>(gi); launches++;
            k_shade_secondary<<<blocks, 128, 0, ctx->stream>>>(gi); launches++;
        } else {
            // generate -> shared ray-queue kernel -> finish; queue counters: [4] count, [5] head for K4, [6], [7] for K5
            GIWaveArgs wa{};
            wa.gi = gi; wa.st_ray0 = f.st_ray0; wa.st_ray1 = f.st_ray1; wa.st_acc = f.st_acc; wa.st_misc = f.st_misc; wa.st_hit = f.st_hit; wa.st_hit_inst = f.st_hit_inst;
            wa.queue = f.queue0; wa.count = f.counters + 4;
            k_brdf_rays_gen<<<blocks, 128, 0, ctx->stream>>>(wa);
            if ((rc = trace_queue_generic(ctx, wa.queue, wa.count, f.counters + 5, stat_rays))) return rc;
            k_brdf_rays_shade<<<blocks, 128, 0, ctx->stream>>>(wa); launches += 3;
            if (g.enable_restir_di == 1u) {                            // K5 samples the light list at the secondary surface: the K9 stages, pointed at SecondaryGBufferData
                DIWaveArgs dw{};
                dw.di.sc = gi.sc; dw.di.g = gi.g; dw.di.width = gi.width; dw.di.height = gi.height; dw.di.frame = gi.frame; dw.di.band = band; dw.di.gb = gi.gb; dw.di.li = gi.li;
                dw.di.stat_rays = stat_rays; dw.di.stat_hits = stat_hits;
                dw.st_ray0 = f.st_ray0; dw.st_ray1 = f.st_ray1; dw.st_hit = f.st_hit; dw.st_hit_inst = f.st_hit_inst;
                dw.state = f.di_state; dw.plane = (uint32_t)((size_t)f.width * f.height); dw.secondary = f.secondary_gbuffer;
                dw.queue = f.queue1; dw.count = f.counters + 6;
                k_di_stage_a<true><<<blocks, 128, 0, ctx->stream>>>(dw);
                if ((rc = trace_queue_generic(ctx, f.queue1, f.counters + 6, f.counters + 7, stat_rays))) return rc;
                k_di_stage_b<true><<<blocks, 128, 0, ctx->stream>>>(dw, f.queue0, f.counters + 14); launches += 3;
                if (g.restir_di.shading_params.enable_final_visibility == 1u) {
                    if ((rc = trace_queue_generic(ctx, f.queue0, f.counters + 14, f.counters + 15, stat_rays, true))) return rc;
                    launches++;
                }
                k_shade_secondary_stage_d<<<blocks, 128, 0, ctx->stream>>>(dw, gi); launches++;
            } else {
                wa.queue = f.queue1; wa.count = f.counters + 6;
                k_shade_secondary_gen<<<blocks, 128, 0, ctx->stream>>>(wa);
                if ((rc = trace_queue_generic(ctx, wa.queue, wa.count, f.counters + 7, stat_rays))) return rc;
                k_shade_secondary_shade<<<blocks, 128, 0, ctx->stream>>>(wa); launches += 3;
            }
        }
        if (g.enable_temporal_resampling == 1u || g.enable_spatial_resampling == 1u) {
            const uint32_t prev = cur ^ 1u;
            GIResampleArgs ra{};
            ra.sc = ga.sc; ra.g = ctx->d_gconst; ra.width = f.width; ra.height = f.height; ra.frame = frame; ra.band = band;
            ra.gb = gbrefs;
            ra.gb_prev = GBufferRefs{f.gb_depth[prev], f.gb_normals[prev], f.gb_geo_normals[prev], f.gb_diffuse_albedo[prev], f.gb_specular_rough[prev], ctx->codec_luts};
            ra.motion = f.motion_vectors; ra.neighbor_offsets = f.neighbor_offsets; ra.reservoirs = f.gi_reservoirs;
            ra.stat_rays = stat_rays; ra.stat_hits = stat_hits;
            // the reference records K6 and K7 without a barrier between them (src/light_passes.rs:642-647); stream order = K6 then K7
            if (g.enable_temporal_resampling == 1u && (ph & 2u)) { k_gi_temporal<<<blocks, 128, 0, ctx->stream>>>(ra); launches++; }
            if (g.enable_spatial_resampling == 1u && (ph & 4u)) { k_gi_spatial<<<blocks, 128, 0, ctx->stream>>>(ra); launches++; }
        }
        if (ph & 4u) { k_gi_final_shading<<<blocks, 128, 0, ctx->stream>>>(gi); launches++; }
    }
    RT2_CUDA(ctx, cudaGetLastError());
    ctx->stats.kernel_launches = (ph & 1u) ? launches : ctx->stats.kernel_launches + launches;
    ctx->stats.wavefront_iterations = 1;
    ctx->stats_from_counters = true;
    ctx->stats_walks_counted = false;
    ctx->stats_reused_counted = true;
    if (ph & 1u) ctx->pending_rays_static = static_rays;
    return RT2_OK;
}

}  // namespace rt2
