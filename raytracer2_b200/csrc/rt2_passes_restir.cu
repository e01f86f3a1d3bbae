// rt2_passes_restir.cu — the non-reference branch of LightPasses::execute (src/light_passes.rs:598-659):
//   K3 g_buffer            lighting_passes/g_buffer.rgen:14-47
//   K9 di_fused_resampling lighting_passes/di_fused_resampling.rgen:16-93 (reservoir light sampling)
//
// Both are one-thread-per-pixel kernels with the traversal inlined (8x4 pixel tiles per warp so the
// primary / BRDF rays of a warp start coherent).  K3 is bound like the primary wavefront of K1
// (instruction issue in traversal, 44 B written per pixel); K9 evaluates numLocal light candidates per
// pixel out of a 48 B/light list that sits in L2 (10k lights = 480 KB) and traces <= numBrdf + 2 rays.
#include "rt2_restir.cuh"

namespace rt2 {

int restir_ensure_resources(rt2_ctx* ctx);
int lights_ensure_resources(rt2_ctx* ctx);

struct GBufferArgs {
    SceneRefs sc; const rt2_gconst* g; uint32_t width, height;
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

__device__ __forceinline__ bool tile_pixel(uint32_t i, uint32_t width, uint32_t height, uint32_t* x, uint32_t* y) {
    const uint32_t tiles_x = (width + 7u) >> 3;
    const uint32_t tile = i >> 5, within = i & 31u;
    *x = (tile % tiles_x) * 8u + (within & 7u); *y = (tile / tiles_x) * 4u + (within >> 3);
    return *x < width && *y < height;
}

__global__ void __launch_bounds__(128) k_gbuffer(GBufferArgs a) {
    __shared__ float lut[256];
    lut[threadIdx.x] = a.sc.srgb_lut[threadIdx.x]; lut[threadIdx.x + 128] = a.sc.srgb_lut[threadIdx.x + 128];
    __syncthreads();
    uint32_t x, y;
    if (!tile_pixel(blockIdx.x * blockDim.x + threadIdx.x, a.width, a.height, &x, &y)) return;
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
// K9
// ------------------------------------------------------------------------------------------------
struct DIArgs {
    SceneRefs sc; const rt2_gconst* g; uint32_t width, height, frame;
    GBufferRefs gb;
    const rt2_light_info* lights; const uint32_t* geom_to_light;
    const float* pdf; uint32_t pdf_w, pdf_h, pdf_last_offset;
    rt2_packed_di_reservoir* reservoirs; ushort4* diffuse; ushort4* specular; float4* diffuse_f32;
    unsigned long long* stat_rays; unsigned long long* stat_hits;
};

// RAB_EvaluateLocalLightSourcePdf, bridge :420-434 (Q14 pinned to texelFetch semantics)
__device__ __forceinline__ float local_light_source_pdf(const DIArgs& a, uint32_t light_index) {
    uint32_t x, y;
    linear_to_zcurve(light_index, x, y);
    const float texel = (x < a.pdf_w && y < a.pdf_h) ? __ldg(a.pdf + (size_t)y * a.pdf_w + x) : 0.f;
    const int last = max(0, (int)floorf(log2_pinned((float)max(a.pdf_w, a.pdf_h))));
    const float average = __ldg(a.pdf + a.pdf_last_offset);
    const float sum = average * (float)(1u << (uint32_t)last << (uint32_t)last);
    return texel / sum;
}

__device__ __forceinline__ bool visibility_ray_miss(const DIArgs& a, const Surface& s, V3 sample_position, float offset, uint32_t* rays) {
    const V3 L = sample_position - s.world_pos;                        // setupVisibilityRay, bridge :191-218
    const float tmax = fmaxf(offset, length3(L) - offset * 2);
    uint32_t slot;
    const Hit h = trace_closest(a.sc, s.world_pos, normalize3(L), offset, tmax, &slot);
    (*rays)++;
    return h.geom == 0xFFFFFFFFu;
}

__global__ void __launch_bounds__(128) k_di_fused(DIArgs a) {
    uint32_t lx, ly;
    if (!tile_pixel(blockIdx.x * blockDim.x + threadIdx.x, a.width, a.height, &lx, &ly)) return;
    const rt2_gconst& g = *a.g;
    const rt2_restirdi_params& di = g.restir_di;
    uint32_t rays = 0, hits = 0;
    uint32_t px = lx, py = ly;                                         // RTXDI_ReservoirPosToPixelPos, RtxdiHelpers.hlsli:52-61
    const uint32_t field = g.runtime_params.active_checkerboard_field;
    if (field != 0u) { px = lx << 1; px += ((py + field) & 1u); }
    Rng rng = rng_init(px, py, a.frame, 1u);
    const Surface surface = gbuffer_surface(a.gb, a.width, (int)px, (int)py, g.view);
    const SampleParams sp = init_sample_params(di.initial_sampling_params.num_primary_local_light_samples, di.initial_sampling_params.num_primary_infinite_light_samples,
                                               di.initial_sampling_params.num_primary_environment_samples, di.initial_sampling_params.num_primary_brdf_samples,
                                               di.initial_sampling_params.brdf_cutoff, 0.001f);
    // ---- RTXDI_SampleLightsForSurface, InitialSamplingFunctions.hlsli:594-664 -----------------------------------
    const uint32_t first_local = g.light_buffer_params.local_light_buffer_region.first_light_index;
    const uint32_t num_local = g.light_buffer_params.local_light_buffer_region.num_lights;
    LightSample local_sample = empty_light_sample();
    DIReservoir local_reservoir = empty_di_reservoir();
    if (num_local != 0u && sp.num_local != 0u) {                       // RTXDI_SampleLocalLights{,Internal}, :264-343 (uniform selection)
        DIReservoir state = empty_di_reservoir();
        for (uint32_t i = 0; i < sp.num_local; i++) {
            const float rnd = rng_next(rng);                           // UniformSampling.hlsli:14-25
            const float inv_source_pdf = (float)num_local;
            const uint32_t light_index = first_local + min((uint32_t)floorf(rnd * (float)num_local), num_local - 1u);
            const rt2_light_info li = load_light_info(a.lights, light_index);
            const float uvx = rng_next(rng), uvy = rng_next(rng);
            const LightSample cand = sample_polymorphic_light(li, surface, uvx, uvy);           // :131-156
            const float blended = light_brdf_mis_weight(surface, cand, 1.0f / inv_source_pdf, sp.local_mis_weight, false, sp);
            const float target_pdf = light_sample_target_pdf(cand, surface);
            const float ris_rnd = rng_next(rng);
            if (blended == 0) continue;
            if (di_stream_sample(state, light_index, uvx, uvy, ris_rnd, target_pdf, 1.0f / blended)) local_sample = cand;
        }
        di_finalize(state, 1.0f, (float)sp.num_mis);
        state.M = 1;
        local_reservoir = state;
    }
    // RTXDI_SampleInfiniteLights: the infinite region is always empty (src/main.rs:375-380; rejected on the host otherwise)
    const LightSample infinite_sample = empty_light_sample();
    const DIReservoir infinite_reservoir = empty_di_reservoir();
    // ---- RTXDI_SampleBrdf, :507-591
    LightSample brdf_sample = empty_light_sample();
    DIReservoir brdf_reservoir = empty_di_reservoir();
    for (uint32_t i = 0; i < sp.num_brdf; ++i) {
        float light_source_pdf = 0;
        V3 sample_dir = v3(0.f, 0.f, 0.f);
        uint32_t light_index = RT2_INVALID_LIGHT;
        float rx = 0, ry = 0;
        LightSample cand = empty_light_sample();
        if (surface_brdf_sample(surface, rng, &sample_dir)) {
            const float brdf_pdf = surface_brdf_pdf(surface, sample_dir);
            const float max_distance = brdf_max_distance_from_pdf(sp.brdf_cutoff, brdf_pdf);
            uint32_t slot;                                             // RAB_TraceRayForLocalLight, bridge :629-669
            const Hit h = trace_closest(a.sc, surface.world_pos, sample_dir, sp.brdf_ray_min_t, max_distance, &slot);
            rays++;
            if (h.geom != 0xFFFFFFFFu) {
                hits++;
                uint32_t li = __ldg(a.geom_to_light + h.geom);         // getLightIndex, bridge :621-627
                if (li != RT2_INVALID_LIGHT) li += h.prim;
                light_index = li;
            }
            if (light_index != RT2_INVALID_LIGHT) {
                const float bx = 1 - h.u - h.v, bz = h.v;              // hitUVToBarycentric, randomFromBarycentric (Helpers.glsl:76-86)
                const float sqrtx = 1 - bx;
                rx = sqrtx * sqrtx; ry = bz / sqrtx;
                cand = sample_polymorphic_light(load_light_info(a.lights, light_index), surface, rx, ry);
                if (sp.brdf_cutoff > 0.f) {
                    const V3 to_light = cand.position - surface.world_pos;
                    const float light_distance = length3(to_light);
                    const float pdf2 = surface_brdf_pdf(surface, to_light / light_distance);
                    if (light_distance > brdf_max_distance_from_pdf(sp.brdf_cutoff, pdf2)) light_index = RT2_INVALID_LIGHT;
                }
                if (light_index != RT2_INVALID_LIGHT) light_source_pdf = local_light_source_pdf(a, light_index);
            }
            // BRDF miss + environmentLightParams.lightPresent: the environment light (index lights + 1) is never written (Q7) -> absent
        }
        if (light_source_pdf == 0) continue;
        const float target_pdf = light_sample_target_pdf(cand, surface);
        const float blended = light_brdf_mis_weight(surface, cand, light_source_pdf, sp.local_mis_weight, false, sp);
        const float ris_rnd = rng_next(rng);
        if (di_stream_sample(brdf_reservoir, light_index, rx, ry, ris_rnd, target_pdf, 1.0f / blended)) brdf_sample = cand;
    }
    di_finalize(brdf_reservoir, 1.0f, (float)sp.num_mis);
    brdf_reservoir.M = 1;
    // ---- combine, :637-663
    DIReservoir reservoir = empty_di_reservoir();
    di_combine(reservoir, local_reservoir, 0.5f, local_reservoir.target_pdf);
    const bool select_infinite = di_combine(reservoir, infinite_reservoir, rng_next(rng), infinite_reservoir.target_pdf);
    const bool select_brdf = di_combine(reservoir, brdf_reservoir, rng_next(rng), brdf_reservoir.target_pdf);
    di_finalize(reservoir, 1.0f, 1.0f);
    reservoir.M = 1;
    LightSample light_sample = select_brdf ? brdf_sample : (select_infinite ? infinite_sample : local_sample);
    // ---- di_fused_resampling.rgen:42-48: initial visibility
    if (di.initial_sampling_params.enable_initial_visibility == 1u && di_valid(reservoir)) {
        if (!visibility_ray_miss(a, surface, light_sample.position, 0.001f, &rays)) { hits++; di_store_visibility(reservoir, v3(0.f, 0.f, 0.f), true); }
    }
    // ---- shading: ShadingHelpers.glsl:2-58, DemodulateSpecular Helpers.glsl:312-315
    V3 diffuse = v3(0.f, 0.f, 0.f), specular = v3(0.f, 0.f, 0.f);
    if (di_valid(reservoir)) {
        if (!(light_sample.solid_angle_pdf <= 0)) {
            if (di.shading_params.enable_final_visibility == 1u) {
                // visibility reuse needs reservoir.age > 0 (DIReservoir.hlsli:186-189); this reservoir is fresh, so the ray is always traced
                const bool vis = visibility_ray_miss(a, surface, light_sample.position, 0.01f, &rays);
                if (!vis) hits++;
                const V3 visibility = vis ? v3(1.f, 1.f, 1.f) : v3(0.f, 0.f, 0.f);
                di_store_visibility(reservoir, visibility, di.temporal_resampling_params.discard_invisible_samples == 1u);
                light_sample.radiance = light_sample.radiance * visibility;
            }
            light_sample.radiance = light_sample.radiance * (reservoir.weight_sum / light_sample.solid_angle_pdf);
            if (light_sample.radiance.x > 0.0f || light_sample.radiance.y > 0.0f || light_sample.radiance.z > 0.0f) {
                const V3 L = normalize3(light_sample.position - surface.world_pos);             // EvaluateBrdf, bridge :146-159
                const float dd = lambert(surface.normal, neg3(L));
                const V3 spec = surface.roughness == 0 ? v3(0.f, 0.f, 0.f)
                                                       : ggx_times_ndotl(surface.view_dir, L, surface.normal, fmaxf(surface.roughness, RT2_MIN_ROUGHNESS), surface.specular_f0);
                diffuse = light_sample.radiance * dd;
                specular = spec * light_sample.radiance;
            }
        }
        specular = v3(specular.x / fmaxf(0.01f, surface.specular_f0.x), specular.y / fmaxf(0.01f, surface.specular_f0.y), specular.z / fmaxf(0.01f, surface.specular_f0.z));
    }
    const uint32_t ptr = reservoir_pointer(di.reservoir_buffer_params.reservoir_block_row_pitch, di.reservoir_buffer_params.reservoir_array_pitch, lx, ly,
                                           di.buffer_indices.shading_input_buffer_index);
    const rt2_packed_di_reservoir packed = pack_di_reservoir(reservoir);
    uint2* dst = reinterpret_cast<uint2*>(a.reservoirs + ptr);
    dst[0] = make_uint2(packed.light_data, packed.uv_data); dst[1] = make_uint2(packed.m_visibility, packed.distance_age);
    dst[2] = make_uint2(__float_as_uint(packed.target_pdf), __float_as_uint(packed.weight));
    store_shading_output(g, (size_t)ly * a.width + lx, diffuse, specular, g.enable_restir_di == 1u, a.diffuse, a.specular, a.diffuse_f32);
    // per-warp ray statistics
    const uint32_t am = __activemask();
    const uint32_t wr = __reduce_add_sync(am, rays), wh = __reduce_add_sync(am, hits);
    if ((threadIdx.x & 31u) == (uint32_t)(__ffs(am) - 1)) { atomicAdd(a.stat_rays, (unsigned long long)wr); atomicAdd(a.stat_hits, (unsigned long long)wh); }
}

// ------------------------------------------------------------------------------------------------
// host: LightPasses::execute, non-reference branch (src/light_passes.rs:598-659)
// ------------------------------------------------------------------------------------------------
int render_restir(rt2_ctx* ctx, uint32_t frame, const rt2_launch_params& lp) {
    if (!ctx->scene.accel_built) return fail(ctx, RT2_E_STATE, "no acceleration structure: call rt2_upload_scene + rt2_build_accel first");
    if (lp.band_world > 1u) return fail(ctx, RT2_E_INVALID, "row-band sharding is implemented for the reference path tracer only");
    const rt2_gconst& g = ctx->gconst;
    if (g.enable_restir_gi == 1u) return fail(ctx, RT2_E_STATE, "the ReSTIR GI chain (brdf_rays / shade_secondary / gi_final_shading) is not built yet: set enable_restir_gi = 0");
    int rc = restir_ensure_resources(ctx); if (rc) return rc;
    FrameBuffers& f = ctx->fb;
    // counters: [0..1] rays, [2..3] hits
    if (f.counter_capacity < 8) {
        if (f.counters) { RT2_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); cudaFree(f.counters); f.counters = nullptr; }
        RT2_CUDA(ctx, cudaMalloc((void**)&f.counters, 8 * 4)); f.counter_capacity = 8;
    }
    RT2_CUDA(ctx, cudaMemsetAsync(f.counters, 0, 16, ctx->stream));
    unsigned long long* stat_rays = reinterpret_cast<unsigned long long*>(f.counters);
    unsigned long long* stat_hits = reinterpret_cast<unsigned long long*>(f.counters + 2);
    const uint32_t cur = frame & 1u;                   // ping-pong, src/light_passes.rs:565-577
    ctx->last_gbuffer_index = cur;
    const uint32_t tiles_x = (f.width + 7) / 8, tiles_y = (f.height + 3) / 4;
    const uint32_t padded = tiles_x * tiles_y * 32;
    uint64_t launches = 0;
    // K3
    GBufferArgs ga{};
    ga.sc = make_scene_refs(ctx); ga.g = ctx->d_gconst; ga.width = f.width; ga.height = f.height;
    ga.depth = f.gb_depth[cur]; ga.normals = f.gb_normals[cur]; ga.geo_normals = f.gb_geo_normals[cur]; ga.diffuse_albedo = f.gb_diffuse_albedo[cur];
    ga.specular_rough = f.gb_specular_rough[cur]; ga.emissive = f.gb_emissive[cur]; ga.motion = f.motion_vectors; ga.stat_hits = stat_hits;
    k_gbuffer<<<(padded + 127) / 128, 128, 0, ctx->stream>>>(ga); launches++;
    uint64_t static_rays = (uint64_t)f.width * f.height;
    // K9
    if (g.enable_restir_di == 1u) {
        const uint32_t num_local = g.light_buffer_params.local_light_buffer_region.num_lights;
        if (g.light_buffer_params.infinite_light_buffer_region.num_lights != 0u) return fail(ctx, RT2_E_INVALID, "infinite lights have no producer in the reference (src/main.rs:375-380)");
        if (g.restir_di.initial_sampling_params.num_primary_environment_samples != 0u) return fail(ctx, RT2_E_INVALID, "environment RIS sampling is compiled out (RTXDI_ENABLE_PRESAMPLING 0)");
        if (g.restir_di.reservoir_buffer_params.reservoir_array_pitch != f.reservoir_array_pitch || g.restir_di.reservoir_buffer_params.reservoir_block_row_pitch != f.reservoir_block_row_pitch)
            return fail(ctx, RT2_E_INVALID, "GConst reservoir buffer parameters do not match the frame size (calculate_reservoir_buffer_parameters)");
        if (g.restir_di.buffer_indices.shading_input_buffer_index > 1u) return fail(ctx, RT2_E_INVALID, "reservoir buffer index out of range");
        if (num_local) {
            if (!ctx->lights_prepared) return fail(ctx, RT2_E_STATE, "light list not built: call rt2_prepare_lights (+ rt2_generate_mips(0)) first");
            if (g.light_buffer_params.local_light_buffer_region.first_light_index + (uint64_t)num_local > ctx->light_capacity) return fail(ctx, RT2_E_INVALID, "local light region exceeds the light buffer");
        }
        if ((rc = lights_ensure_resources(ctx))) return rc;
        DIArgs da{};
        da.sc = ga.sc; da.g = ctx->d_gconst; da.width = f.width; da.height = f.height; da.frame = frame;
        da.gb = GBufferRefs{f.gb_depth[cur], f.gb_normals[cur], f.gb_geo_normals[cur], f.gb_diffuse_albedo[cur], f.gb_specular_rough[cur]};
        da.lights = ctx->light_info; da.geom_to_light = ctx->geom_to_light;
        da.pdf = ctx->local_pdf; da.pdf_w = ctx->local_pdf_w; da.pdf_h = ctx->local_pdf_h;
        {
            const uint32_t m = std::max(ctx->local_pdf_w, ctx->local_pdf_h);
            uint32_t last = 0; while ((2u << last) <= m) last++;           // floor(log2(m))
            last = std::min(last, ctx->local_pdf_mips - 1);
            uint32_t off = 0;
            for (uint32_t k = 0; k < last; k++) off += std::max(1u, ctx->local_pdf_w >> k) * std::max(1u, ctx->local_pdf_h >> k);
            da.pdf_last_offset = off;
        }
        da.reservoirs = f.di_reservoirs; da.diffuse = f.diffuse_f16; da.specular = f.specular_f16; da.diffuse_f32 = f.diffuse_f32;
        da.stat_rays = stat_rays; da.stat_hits = stat_hits;
        k_di_fused<<<(padded + 127) / 128, 128, 0, ctx->stream>>>(da); launches++;
    }
    RT2_CUDA(ctx, cudaGetLastError());
    ctx->stats.kernel_launches = launches;
    ctx->stats.wavefront_iterations = 1;
    ctx->stats_from_counters = true;
    ctx->pending_rays_static = static_rays;
    return RT2_OK;
}

}  // namespace rt2
