// rt2_restir.cuh — device restatement of the RTXDI application bridge and the reachable part of the
// RTXDI SDK (SURVEY §8a rows G2, D1–D10): G-buffer surfaces, BRDF evaluation, triangle-light
// sampling, DI reservoirs and the initial light sampling.  Citations are relative to src/shaders/ of
// the reference tree ("bridge" = lighting_passes/RtxdiApplicationBridge.glsl).
#pragma once
#include "rt2_surface.cuh"
#include "rt2_codec.cuh"

namespace rt2 {

#define RT2_MIN_ROUGHNESS 0.05f          /* common.glsl:3 */
#define RT2_INVALID_LIGHT 0xFFFFFFFFu    /* rtxdi/RtxdiParameters.h:52 */

// RAB_Surface, bridge :83-94
struct Surface { V3 world_pos, view_dir; float view_depth; V3 normal, geo_normal, diffuse_albedo, specular_f0; float roughness, diffuse_probability; };
// RAB_EmptySurface leaves everything but viewDepth uninitialised (bridge :161-166): pinned to zero
__device__ __forceinline__ Surface empty_surface() {
    Surface s;
    s.world_pos = s.view_dir = s.normal = s.geo_normal = s.diffuse_albedo = s.specular_f0 = v3(0.f, 0.f, 0.f);
    s.roughness = s.diffuse_probability = 0.f; s.view_depth = RT2_BACKGROUND_DEPTH;
    return s;
}
struct GBufferRefs { const float* depth; const uint32_t* normals; const uint32_t* geo_normals; const uint32_t* diffuse_albedo; const uint32_t* specular_rough;
                     const float* gamma_lut; /* CodecLuts::gamma22, may be null */ };
// GetGBufferSurface, bridge :296-323; viewDepthToWorldPos, GBufferHelpers.glsl:54-67
__device__ __forceinline__ Surface gbuffer_surface(const GBufferRefs& gb, uint32_t width, int px, int py, const rt2_planar_view_constants& view) {
    Surface s = empty_surface();
    if ((float)px >= view.viewport_size[0] || (float)py >= view.viewport_size[1]) return s;
    if (px < 0 || py < 0) return s;      // imageLoad outside the image: pinned to "invalid surface" (oracle pin, K6/K7 neighbours)
    const size_t i = (size_t)py * width + (size_t)px;
    s.view_depth = gb.depth[i];
    if (s.view_depth == RT2_BACKGROUND_DEPTH) return s;
    s.normal = oct_to_ndir_unorm32(gb.normals[i]);
    s.geo_normal = oct_to_ndir_unorm32(gb.geo_normals[i]);
    s.diffuse_albedo = unpack_r11g11b10(gb.diffuse_albedo[i]);
    const float4 sr = unpack_rgba8_gamma(gb.specular_rough[i], gb.gamma_lut);
    s.specular_f0 = v3(sr.x, sr.y, sr.z); s.roughness = sr.w;
    V3 o, d;
    setup_primary_ray(view, (uint32_t)px, (uint32_t)py, &o, &d);
    s.world_pos = o + d * s.view_depth;
    s.view_dir = normalize3(o - s.world_pos);
    s.diffuse_probability = 1.0f;                 // getSurfaceDiffuseProbability, bridge :131-138
    return s;
}

// ---- BRDF: Helpers.glsl:182-239 with `#define square(a) a*a` (common.glsl:2) expanded literally (Q1) -----------
__device__ __forceinline__ float ggx_d_literal(float alpha, float NoH) {
    return alpha * alpha / (((RT2_RTXDI_PI * NoH * NoH * alpha * alpha) + (1 - NoH * NoH) * NoH * NoH * alpha * alpha) + (1 - NoH * NoH));
}
__device__ __forceinline__ float ggx_vndf_pdf(float roughness, V3 N, V3 V, V3 L) {
    const V3 H = normalize3(L + V);
    const float NoH = saturatef(dot3(N, H)), VoH = saturatef(dot3(V, H));
    const float alpha = roughness * roughness;
    const float D = ggx_d_literal(alpha, NoH);
    return (VoH > 0.0f) ? D / (4.0f * VoH) : 0.0f;
}
__device__ __forceinline__ float g_smith_over_ndotv(float roughness, float NdotV, float NdotL) {
    const float alpha = roughness * roughness;
    const float g1 = NdotV * sqrtf(alpha * alpha + (1.0f - alpha * alpha) * NdotL * NdotL);
    const float g2 = NdotL * sqrtf(alpha * alpha + (1.0f - alpha * alpha) * NdotV * NdotV);
    return 2.0f * NdotL / (g1 + g2);
}
__device__ __forceinline__ V3 ggx_times_ndotl(V3 V, V3 L, V3 N, float roughness, V3 F0) {
    const V3 H = normalize3(L + V);
    const float NoL = saturatef(dot3(N, L)), VoH = saturatef(dot3(V, H)), NoV = saturatef(dot3(N, V)), NoH = saturatef(dot3(N, H));
    if (NoL > 0) {
        const float G = g_smith_over_ndotv(roughness, NoV, NoL);
        const float alpha = roughness * roughness;
        const float D = ggx_d_literal(alpha, NoH);
        const float p = pow5_pinned(fmaxf(1 - VoH, 0.f));                 // Schlick_Fresnel
        const V3 F = v3(F0.x + (1 - F0.x) * p, F0.y + (1 - F0.y) * p, F0.z + (1 - F0.z) * p);
        return F * (D * G / 4);
    }
    return v3(0.f, 0.f, 0.f);
}
__device__ __forceinline__ float lambert(V3 normal, V3 light_incident) { return fmaxf(0.f, -dot3(normal, light_incident)) / RT2_RTXDI_PI; }
// RAB_GetSurfaceBrdfPdf, bridge :463-470
__device__ __forceinline__ float surface_brdf_pdf(const Surface& s, V3 dir) {
    const float cos_theta = saturatef(dot3(s.normal, dir));
    const float diffuse_pdf = cos_theta / RT2_RTXDI_PI;
    const float specular_pdf = ggx_vndf_pdf(fmaxf(s.roughness, RT2_MIN_ROUGHNESS), s.normal, s.view_dir, dir);
    return cos_theta > 0.f ? mixf(specular_pdf, diffuse_pdf, s.diffuse_probability) : 0.f;
}
// RAB_GetSurfaceBrdfSample, bridge :437-461 (diffuseProbability == 1: only the cosine lobe is reachable)
__device__ __forceinline__ bool surface_brdf_sample(const Surface& s, Rng& rng, V3* dir) {
    const float rx = rng_next(rng), ry = rng_next(rng), rz = rng_next(rng);
    (void)rx;
    *dir = tangent_to_world(s.normal, sample_cos_hemisphere(ry, rz, false));
    return dot3(s.normal, *dir) > 0.f;
}

// ---- light samples: bridge :96-103, :478-525; PolymorphicLight.glsl:266-294 ------------------------------------
struct LightSample { V3 position, normal, radiance; float solid_angle_pdf; uint32_t light_type; };
__device__ __forceinline__ LightSample empty_light_sample() {
    LightSample s; s.position = s.normal = s.radiance = v3(0.f, 0.f, 0.f); s.solid_angle_pdf = 0.f; s.light_type = 0u; return s;
}
// a light-buffer entry (48 B apart), or — with compact_index — one of K11's compact records (32 B apart, `lights` = their base)
__device__ __forceinline__ rt2_light_info load_light_info(const rt2_light_info* lights, uint32_t index, uint32_t compact_index = 0xFFFFFFFFu) {
    const uint4* p = compact_index != 0xFFFFFFFFu ? reinterpret_cast<const uint4*>(lights) + 2 * (size_t)compact_index
                                                  : reinterpret_cast<const uint4*>(lights + index);
    const uint4 a = __ldg(p), b = __ldg(p + 1);      // uint4[2] holds only shaping data (unused for triangle lights)
    rt2_light_info li;
    li.center[0] = __uint_as_float(a.x); li.center[1] = __uint_as_float(a.y); li.center[2] = __uint_as_float(a.z); li.color_type_and_flags = a.w;
    li.direction1 = b.x; li.direction2 = b.y; li.scalars = b.z; li.log_radiance = b.w;
    li.ies_profile_index = li.primary_axis = li.cos_cone_angle_and_softness = li.padding = 0u;
    return li;
}
#define RT2_DISTANT_LIGHT_DISTANCE 1000.0f   /* PolymorphicLight.glsl:43 */
// equirectUVToDirection, Helpers.glsl:334-345
__device__ __forceinline__ V3 equirect_uv_to_direction(float u, float v, float* cos_elevation) {
    const float azimuth = (u + 0.25f) * (2 * RT2_RTXDI_PI);
    const float elevation = (0.5f - v) * RT2_RTXDI_PI;
    float sa, ca, se, ce;
    sincos_pinned(azimuth, &sa, &ca); sincos_pinned(elevation, &se, &ce);
    *cos_elevation = ce;
    return v3(ca * ce, se, sa * ce);
}
// calcEnvironmentLightSample, PolymorphicLight.glsl:368-410 (importance-sampled branch: the only kind K10 writes)
__device__ __forceinline__ LightSample environment_light_sample(const SceneRefs& sc, const rt2_light_info& li, V3 viewer_position, float rx, float ry) {
    LightSample out = empty_light_sample();
    out.light_type = RT2_LIGHT_ENVIRONMENT;
    const int texture_index = (int)li.direction1;                         // CreateEnvironmentLight, :414-426
    const float rotation = f16_bits_to_f32((uint16_t)(li.scalars & 0xffffu));
    const uint32_t tex_w = li.direction2 & 0xffffu, tex_h = li.direction2 >> 16;
    const V3 radiance_scale = unpack_r8g8b8(li.color_type_and_flags) * unpack_light_radiance(li.log_radiance & 0xffffu, sc.radiance_lut);
    float cos_elevation;
    const V3 dir = equirect_uv_to_direction(rx + rotation, ry, &cos_elevation);
    out.solid_angle_pdf = (float)(tex_w * tex_h) / (2 * RT2_RTXDI_PI * RT2_RTXDI_PI * cos_elevation);
    V3 radiance = radiance_scale;
    if (texture_index >= 0) radiance = radiance * sample_sky(sc, rx, ry);
    const float sum = radiance.x + radiance.y + radiance.z;
    if (isinf(sum) || isnan(sum)) radiance = v3(0.f, 0.f, 0.f);
    out.position = viewer_position + dir * RT2_DISTANT_LIGHT_DISTANCE;
    out.normal = neg3(dir);
    out.radiance = radiance;
    return out;
}
__device__ __forceinline__ LightSample sample_polymorphic_light(const SceneRefs& sc, const rt2_light_info& li, const Surface& surface, float uvx, float uvy) {
    LightSample out = empty_light_sample();
    out.light_type = (li.color_type_and_flags >> RT2_LIGHT_TYPE_SHIFT) & RT2_LIGHT_TYPE_MASK;
    if (out.light_type == RT2_LIGHT_ENVIRONMENT && sc.sky != nullptr && (li.scalars >> 16) != 0u) {
        out = environment_light_sample(sc, li, surface.world_pos, uvx, uvy);
        if (out.solid_angle_pdf > 0) out.radiance = out.radiance * v3(1.0f, 1.0f, 1.0f);   // evaluateLightShaping with the shaping bit clear
        return out;
    }
    if (out.light_type != RT2_LIGHT_TRIANGLE) return out;
    const TriangleLight self = create_triangle_light(li, sc.radiance_lut);
    const float sqrtx = sqrtf(uvx);                                       // sampleTriangle, Helpers.glsl:66-74
    const float by = sqrtx * (1 - uvy), bz = sqrtx * uvy;
    out.position = (self.base + self.edge1 * by) + self.edge2 * bz;
    out.normal = self.normal;
    V3 L = out.position - surface.world_pos;                             // calcTriangleSolidAnglePdf
    const float Ldist = length3(L);
    L = L / Ldist;
    const float area_pdf = 1.0f / self.surface_area;
    const float cos_theta = clampf(dot3(L, neg3(out.normal)), 0.0f, 1.0f);
    out.solid_angle_pdf = area_pdf * (Ldist * Ldist) / cos_theta;        // pdfAtoW
    out.radiance = self.radiance;
    if (out.solid_angle_pdf > 0) out.radiance = out.radiance * v3(1.0f, 1.0f, 1.0f);   // evaluateLightShaping with isSpot == 0
    return out;
}
__device__ __forceinline__ float light_sample_target_pdf(const LightSample& ls, const Surface& s) {
    if (ls.solid_angle_pdf <= 0) return 0;
    const V3 L = normalize3(ls.position - s.world_pos);
    if (dot3(L, s.geo_normal) <= 0) return 0;
    const float d = lambert(s.normal, neg3(L));
    const V3 sp = s.roughness == 0 ? v3(0.f, 0.f, 0.f) : ggx_times_ndotl(s.view_dir, L, s.normal, fmaxf(s.roughness, RT2_MIN_ROUGHNESS), s.specular_f0);
    const V3 reflected = ls.radiance * (s.diffuse_albedo * d + sp);
    return calc_luminance(reflected) / ls.solid_angle_pdf;
}

// ---- DI reservoirs: rtxdi/DIReservoir.hlsli:28-340 ----------------------------------------------------------------
struct DIReservoir { uint32_t light_data, uv_data; float weight_sum, target_pdf, M; uint32_t packed_visibility; int sdx, sdy; uint32_t age; };
__device__ __forceinline__ DIReservoir empty_di_reservoir() {
    DIReservoir r; r.light_data = r.uv_data = 0u; r.weight_sum = r.target_pdf = r.M = 0.f; r.packed_visibility = 0u; r.sdx = r.sdy = 0; r.age = 0u; return r;
}
__device__ __forceinline__ bool di_valid(const DIReservoir& r) { return r.light_data != 0u; }
__device__ __forceinline__ rt2_packed_di_reservoir pack_di_reservoir(const DIReservoir& r) {
    const int cx = min(max(r.sdx, -127), 127), cy = min(max(r.sdy, -127), 127);
    rt2_packed_di_reservoir d;
    d.light_data = r.light_data; d.uv_data = r.uv_data;
    d.m_visibility = r.packed_visibility | (min((uint32_t)r.M, 0x3fffu) << 18);
    d.distance_age = ((uint32_t)cx & 0xffu) | (((uint32_t)cy & 0xffu) << 8) | (min(r.age, 0xffu) << 16);
    d.target_pdf = r.target_pdf; d.weight = r.weight_sum;
    return d;
}
// RTXDI_ReservoirPositionToPointer, rtxdi/RtxdiHelpers.hlsli:75-88 (16x16 block-linear)
__device__ __forceinline__ uint32_t reservoir_pointer(uint32_t block_row_pitch, uint32_t array_pitch, uint32_t x, uint32_t y, uint32_t array_index) {
    return array_index * array_pitch + (y / 16u) * block_row_pitch + (x / 16u) * 256u + (y % 16u) * 16u + (x % 16u);
}
__device__ __forceinline__ void di_store_visibility(DIReservoir& r, V3 vis, bool discard_if_invisible) {
    r.packed_visibility = (uint32_t)(saturatef(vis.x) * 63.0f) | ((uint32_t)(saturatef(vis.y) * 63.0f) << 6) | ((uint32_t)(saturatef(vis.z) * 63.0f) << 12);
    r.sdx = r.sdy = 0; r.age = 0u;
    if (discard_if_invisible && vis.x == 0 && vis.y == 0 && vis.z == 0) { r.light_data = 0u; r.weight_sum = 0.f; }
}
__device__ __forceinline__ bool di_stream_sample(DIReservoir& r, uint32_t light_index, float uvx, float uvy, float random, float target_pdf, float inv_source_pdf) {
    const float ris_weight = target_pdf * inv_source_pdf;
    r.M += 1;
    r.weight_sum += ris_weight;
    const bool select = (random * r.weight_sum < ris_weight);
    if (select) {
        r.light_data = light_index | 0x80000000u;
        r.uv_data = (uint32_t)(saturatef(uvx) * 65535.0f) | ((uint32_t)(saturatef(uvy) * 65535.0f) << 16);
        r.target_pdf = target_pdf;
    }
    return select;
}
__device__ __forceinline__ bool di_combine(DIReservoir& r, const DIReservoir& n, float random, float target_pdf) {
    const float ris_weight = target_pdf * (n.weight_sum * n.M);
    r.M += n.M;
    r.weight_sum += ris_weight;
    const bool select = (random * r.weight_sum < ris_weight);
    if (select) {
        r.light_data = n.light_data; r.uv_data = n.uv_data; r.target_pdf = target_pdf;
        r.packed_visibility = n.packed_visibility; r.sdx = n.sdx; r.sdy = n.sdy; r.age = n.age;
    }
    return select;
}
__device__ __forceinline__ void di_finalize(DIReservoir& r, float num, float den) {
    const float denominator = r.target_pdf * den;
    r.weight_sum = (denominator == 0.0f) ? 0.0f : (r.weight_sum * num) / denominator;
}

// ---- initial sampling: rtxdi/InitialSamplingFunctions.hlsli:50-115 -------------------------------------------------
struct SampleParams { uint32_t num_local, num_infinite, num_env, num_brdf, num_mis; float local_mis_weight, env_mis_weight, brdf_mis_weight, brdf_cutoff, brdf_ray_min_t; };
__device__ __forceinline__ SampleParams init_sample_params(uint32_t nl, uint32_t ni, uint32_t ne, uint32_t nb, float brdf_cutoff, float brdf_ray_min_t) {
    SampleParams r;
    r.num_local = nl; r.num_infinite = ni; r.num_env = ne; r.num_brdf = nb; r.num_mis = nl + ne + nb;
    r.local_mis_weight = (float)nl / (float)r.num_mis; r.env_mis_weight = (float)ne / (float)r.num_mis; r.brdf_mis_weight = (float)nb / (float)r.num_mis;
    r.brdf_cutoff = brdf_cutoff; r.brdf_ray_min_t = brdf_ray_min_t;
    return r;
}
__device__ __forceinline__ float brdf_max_distance_from_pdf(float brdf_cutoff, float pdf) {
    return brdf_cutoff > 0.f ? sqrtf((1.f / brdf_cutoff - 1.f) * pdf) : 3.402823466e+38F;
}
__device__ __forceinline__ float light_brdf_mis_weight(const Surface& s, const LightSample& ls, float light_selection_pdf, float light_mis_weight,
                                                       bool is_environment_map, const SampleParams& sp) {
    const float pdf = ls.solid_angle_pdf;
    const bool analytic = ls.light_type != RT2_LIGHT_TRIANGLE && ls.light_type != RT2_LIGHT_ENVIRONMENT;   // RAB_IsAnalyticLightSample, bridge :544-548
    if (sp.brdf_mis_weight == 0 || analytic || pdf <= 0 || isinf(pdf) || isnan(pdf)) return light_mis_weight * light_selection_pdf;
    V3 light_dir; float light_distance;                                   // RAB_GetLightDirDistance, bridge :527-542
    if (ls.light_type == RT2_LIGHT_ENVIRONMENT) { light_dir = neg3(ls.normal); light_distance = RT2_DISTANT_LIGHT_DISTANCE; }
    else { const V3 to_light = ls.position - s.world_pos; light_distance = length3(to_light); light_dir = to_light / light_distance; }
    float brdf_pdf = surface_brdf_pdf(s, light_dir);
    const float max_distance = brdf_max_distance_from_pdf(sp.brdf_cutoff, brdf_pdf);
    if (!is_environment_map && light_distance > max_distance) brdf_pdf = 0.f;
    const float source_pdf_wrt_solid_angle = light_selection_pdf * pdf;
    const float blended = light_mis_weight * source_pdf_wrt_solid_angle + sp.brdf_mis_weight * brdf_pdf;
    return blended / pdf;
}

// ---- EvaluateBrdf (bridge :146-159), DemodulateSpecular (Helpers.glsl:312-315) ---------------------------------
struct SplitBrdf { float demodulated_diffuse; V3 specular; };
__device__ __forceinline__ SplitBrdf evaluate_brdf(const Surface& s, V3 sample_position) {
    const V3 L = normalize3(sample_position - s.world_pos);
    SplitBrdf b;
    b.demodulated_diffuse = lambert(s.normal, neg3(L));
    b.specular = s.roughness == 0 ? v3(0.f, 0.f, 0.f) : ggx_times_ndotl(s.view_dir, L, s.normal, fmaxf(s.roughness, RT2_MIN_ROUGHNESS), s.specular_f0);
    return b;
}
__device__ __forceinline__ V3 demodulate_specular(V3 f0, V3 specular) {
    return v3(specular.x / fmaxf(0.01f, f0.x), specular.y / fmaxf(0.01f, f0.y), specular.z / fmaxf(0.01f, f0.z));
}
__device__ __forceinline__ V3 schlick_fresnel(V3 F0, float VdotH) {                      // Helpers.glsl:199-202
    const float p = pow5_pinned(fmaxf(1 - VdotH, 0.f));
    return v3(F0.x + (1 - F0.x) * p, F0.y + (1 - F0.y) * p, F0.z + (1 - F0.z) * p);
}

// ---- VNDF sampling pieces of K4: Helpers.glsl:281-310 ------------------------------------------------------------
__device__ __forceinline__ V3 sample_ggx_vndf(V3 Ve, float roughness, float randx, float randy) {
    const float alpha = roughness * roughness;
    const V3 Vh = normalize3(v3(alpha * Ve.x, alpha * Ve.y, Ve.z));
    const float lensq = Vh.x * Vh.x + Vh.y * Vh.y;
    const V3 T1 = lensq > 0.0f ? v3(-Vh.y, Vh.x, 0.0f) / sqrtf(lensq) : v3(1.0f, 0.0f, 0.0f);
    const V3 T2 = cross3(Vh, T1);
    const float r = sqrtf(randx);
    const float phi = 2.0f * RT2_RTXDI_PI * randy;
    float sn, cs; sincos_pinned(phi, &sn, &cs);
    const float t1 = r * cs;
    float t2 = r * sn;
    const float sv = 0.5f * (1.0f + Vh.z);
    t2 = (1.0f - sv) * sqrtf(1.0f - t1 * t1) + sv * t2;
    const V3 Nh = (T1 * t1 + T2 * t2) + Vh * sqrtf(fmaxf(0.0f, (1.0f - t1 * t1) - t2 * t2));
    return v3(alpha * Nh.x, alpha * Nh.y, fmaxf(0.0f, Nh.z));
}
__device__ __forceinline__ float g1_smith(float roughness, float NdotL) {
    const float alpha = roughness * roughness;
    return 2.0f * NdotL / (NdotL + sqrtf(alpha * alpha + (1.0f - alpha * alpha) * NdotL * NdotL));
}
__device__ __forceinline__ V3 reflect3(V3 I, V3 N) { const float k = 2.0f * dot3(N, I); return I - N * k; }

// ---- GI reservoir codecs: rtxdi/RtxdiMath.hlsli:129-296 (round() pinned to half-to-even, oracle P1) -------------
__device__ __forceinline__ uint32_t pack_snorm2x16(float vx, float vy) {
    if (isnan(vx) || isnan(vy)) { vx = 0.f; vy = 0.f; } else { vx = clampf(vx, -1.0f, 1.0f); vy = clampf(vy, -1.0f, 1.0f); }
    const int ix = (int)rintf(vx * 32767.0f), iy = (int)rintf(vy * 32767.0f);
    return ((uint32_t)ix & 0x0000ffffu) | ((uint32_t)iy << 16);
}
__device__ __forceinline__ uint32_t encode_snorm_oct(V3 n) {
    const float inv = 1.0f / ((fabsf(n.x) + fabsf(n.y)) + fabsf(n.z));
    float px = n.x * inv, py = n.y * inv;
    if (n.z < 0.0f) {
        const float qx = (1.0f - fabsf(py)) * (px >= 0.0f ? 1.0f : -1.0f);
        const float qy = (1.0f - fabsf(px)) * (py >= 0.0f ? 1.0f : -1.0f);
        px = qx; py = qy;
    }
    return pack_snorm2x16(px, py);
}
__device__ __forceinline__ V3 decode_snorm_oct(uint32_t packed) {
    const int bx = (int)(packed << 16) >> 16, by = (int)packed >> 16;
    const float px = fmaxf((float)bx / 32767.0f, -1.0f), py = fmaxf((float)by / 32767.0f, -1.0f);
    V3 n = v3(px, py, (1.0f - fabsf(px)) - fabsf(py));
    if (n.z < 0.0f) {
        const float qx = (1.0f - fabsf(n.y)) * (n.x >= 0.0f ? 1.0f : -1.0f);
        const float qy = (1.0f - fabsf(n.x)) * (n.y >= 0.0f ? 1.0f : -1.0f);
        n.x = qx; n.y = qy;
    }
    return normalize3(n);
}
__device__ __forceinline__ uint32_t encode_logluv(V3 c) {
    const float X = dot3(c, v3(0.4123907992659595f, 0.3575843393838780f, 0.1804807884018343f));
    const float Y = dot3(c, v3(0.2126390058715104f, 0.7151686787677559f, 0.0721923153607337f));
    const float Z = dot3(c, v3(0.0193308187155918f, 0.1191947797946259f, 0.9505321522496608f));
    const float logY = 409.6f * (log2_pinned(Y) + 20.0f);
    const float cl = clampf(logY, 0.0f, 16383.0f);
    const uint32_t Le = isnan(cl) ? 0u : (uint32_t)cl;
    if (Le == 0u) return 0u;
    const float inv_denom = 1.0f / ((-2.0f * X + 12.0f * Y) + 3.0f * ((X + Y) + Z));
    const float u = 4.0f * X * inv_denom, v = 9.0f * Y * inv_denom;
    const float cu = clampf(820.0f * u, 0.0f, 511.0f), cv = clampf(820.0f * v, 0.0f, 511.0f);
    const uint32_t ue = isnan(cu) ? 0u : (uint32_t)cu, ve = isnan(cv) ? 0u : (uint32_t)cv;
    return (Le << 18) | (ue << 9) | ve;
}
// the luminance of a 14-bit LogLuv exponent field (tabulated as CodecLuts::logluv_y)
__device__ __forceinline__ float logluv_luminance(uint32_t Le) {
    const float logY = ((float)Le + 0.5f) / 409.6f - 20.0f;
    return exp2_pinned(logY);                                              // pow(2.0, logY), same pin (binary64, rounded once)
}
__device__ __forceinline__ V3 decode_logluv(uint32_t packed, const float* __restrict__ y_lut = nullptr) {
    const uint32_t Le = packed >> 18;
    if (Le == 0u) return v3(0.f, 0.f, 0.f);
    const float Y = y_lut ? __ldg(y_lut + Le) : logluv_luminance(Le);
    const uint32_t ue = (packed >> 9) & 0x1ffu, ve = packed & 0x1ffu;
    const float u = ((float)ue + 0.5f) / 820.0f, v = ((float)ve + 0.5f) / 820.0f;
    const float inv_denom = 1.0f / ((6.0f * u - 16.0f * v) + 12.0f);
    const float x = 9.0f * u * inv_denom, y = 4.0f * v * inv_denom;
    const float sc = Y / y;
    const V3 XYZ = v3(sc * x, Y, sc * ((1.f - x) - y));
    const float r = dot3(XYZ, v3(3.240969941904522f, -1.537383177570094f, -0.4986107602930032f));
    const float g = dot3(XYZ, v3(-0.9692436362808803f, 1.875967501507721f, 0.04155505740717569f));
    const float b = dot3(XYZ, v3(0.05563007969699373f, -0.2039769588889765f, 1.056971514242878f));
    return v3(fmaxf(r, 0.0f), fmaxf(g, 0.0f), fmaxf(b, 0.0f));
}

// ---- GI reservoirs: rtxdi/GIReservoir.hlsli:29-185, GIResamplingFunctions.hlsli:97-111 ---------------------------
struct GIReservoir { V3 position, normal, radiance; float weight_sum; uint32_t M, age; };
__device__ __forceinline__ GIReservoir empty_gi_reservoir() {
    GIReservoir r; r.position = r.normal = r.radiance = v3(0.f, 0.f, 0.f); r.weight_sum = 0.f; r.M = 0u; r.age = 0u; return r;
}
__device__ __forceinline__ GIReservoir make_gi_reservoir(V3 pos, V3 normal, V3 radiance, float pdf) {
    GIReservoir r; r.position = pos; r.normal = normal; r.radiance = radiance;
    r.weight_sum = pdf > 0.0f ? 1.0f / pdf : 0.0f; r.M = 1u; r.age = 0u;
    return r;
}
// 32-byte packed form as two 16-byte words: {pos.xyz, misc_age_M}, {radiance, weight, normal, unused}
__device__ __forceinline__ void store_gi_reservoir(uint4* buf, uint32_t ptr, const GIReservoir& r, uint32_t misc) {
    const uint32_t mam = (misc & 0xffff0000u) | (min(r.age, 0xffu) << 8) | min(r.M, 0xffu);
    buf[2 * (size_t)ptr] = make_uint4(__float_as_uint(r.position.x), __float_as_uint(r.position.y), __float_as_uint(r.position.z), mam);
    buf[2 * (size_t)ptr + 1] = make_uint4(encode_logluv(r.radiance), __float_as_uint(r.weight_sum), encode_snorm_oct(r.normal), 0u);
}
__device__ __forceinline__ GIReservoir load_gi_reservoir(const uint4* buf, uint32_t ptr, const float* __restrict__ y_lut = nullptr) {
    const uint4 a = buf[2 * (size_t)ptr], b = buf[2 * (size_t)ptr + 1];
    GIReservoir r;
    r.position = v3(__uint_as_float(a.x), __uint_as_float(a.y), __uint_as_float(a.z));
    r.normal = decode_snorm_oct(b.z);
    r.radiance = decode_logluv(b.x, y_lut);
    r.weight_sum = __uint_as_float(b.y);
    r.M = a.w & 0xffu; r.age = (a.w >> 8) & 0xffu;
    return r;
}

// ---- GI resampling: rtxdi/GIResamplingFunctions.hlsli:25-136, RtxdiMath.hlsli:18-29,120-123, bridge :252-265,600-616,673-694
__device__ __forceinline__ float rtxdi_luminance(V3 rgb) { return dot3(rgb, v3(0.2126f, 0.7152f, 0.0722f)); }
__device__ __forceinline__ bool compare_relative_difference(float reference, float candidate, float threshold) {
    return (threshold <= 0) || fabsf(reference - candidate) <= threshold * fmaxf(reference, candidate);
}
__device__ __forceinline__ bool is_valid_neighbor(V3 our_norm, V3 their_norm, float our_depth, float their_depth, float normal_threshold, float depth_threshold) {
    return (dot3(their_norm, our_norm) >= normal_threshold) && compare_relative_difference(our_depth, their_depth, depth_threshold);
}
__device__ __forceinline__ bool materials_similar(const Surface& a, const Surface& b) {
    if (!compare_relative_difference(a.roughness, b.roughness, 0.5f)) return false;
    if (fabsf(calc_luminance(a.specular_f0) - calc_luminance(b.specular_f0)) > 0.25f) return false;
    if (fabsf(calc_luminance(a.diffuse_albedo) - calc_luminance(b.diffuse_albedo)) > 0.25f) return false;
    return true;
}
__device__ __forceinline__ bool validate_gi_jacobian(float& jacobian) {
    if (jacobian > 10.0f || jacobian < 1 / 10.0f) return false;
    jacobian = clampf(jacobian, 1 / 3.0f, 3.0f);
    return true;
}
__device__ __forceinline__ float gi_target_pdf(V3 sample_position, V3 sample_radiance, const Surface& s) {
    const SplitBrdf brdf = evaluate_brdf(s, sample_position);
    return rtxdi_luminance(sample_radiance * (s.diffuse_albedo * brdf.demodulated_diffuse + brdf.specular));
}
__device__ __forceinline__ bool gi_combine(GIReservoir& r, const GIReservoir& n, float random, float target_pdf) {
    const float ris_weight = target_pdf * n.weight_sum * (float)n.M;
    r.M += n.M;
    r.weight_sum += ris_weight;
    const bool select = (random * r.weight_sum <= ris_weight);
    if (select) { r.position = n.position; r.normal = n.normal; r.radiance = n.radiance; r.age = n.age; }
    return select;
}
__device__ __forceinline__ void gi_finalize(GIReservoir& r, float num, float den) {
    r.weight_sum = (den == 0.0f) ? 0.0f : (r.weight_sum * num) / den;
}
__device__ __forceinline__ void gi_partial_jacobian(V3 receiver_pos, V3 sample_pos, V3 sample_normal, float* distance, float* cosine) {
    const V3 vec = receiver_pos - sample_pos;
    *distance = length3(vec);
    *cosine = saturatef(dot3(sample_normal, vec / *distance));
}
__device__ __forceinline__ float gi_jacobian(V3 receiver_pos, V3 neighbor_receiver_pos, const GIReservoir& n) {
    float od, oc, nd, nc;
    gi_partial_jacobian(receiver_pos, n.position, n.normal, &nd, &nc);
    gi_partial_jacobian(neighbor_receiver_pos, n.position, n.normal, &od, &oc);
    float j = (nc * od * od) / (oc * nd * nd);
    if (isinf(j) || isnan(j)) j = 0;
    return j;
}
__device__ __forceinline__ void temporal_resampling_offset(int sample_idx, int radius, int* ox, int* oy) {
    sample_idx &= 7;
    const int mask2 = sample_idx >> 1 & 0x01, mask4 = 1 - (sample_idx >> 2 & 0x01);
    const int tmp0 = -1 + 2 * (sample_idx & 0x01), tmp1 = 1 - 2 * mask2, tmp2 = mask4 | mask2, tmp3 = mask4 | (1 - mask2);
    *ox = tmp0 * tmp2 * radius; *oy = tmp0 * tmp1 * tmp3 * radius;
}
__device__ __forceinline__ void apply_permutation_sampling(int* x, int* y, uint32_t uniform_random_number) {       // RtxdiHelpers.hlsli:63-73
    const int ox = (int)(uniform_random_number & 3u), oy = (int)((uniform_random_number >> 2) & 3u);
    *x += ox; *y += oy; *x ^= 3; *y ^= 3; *x -= ox; *y -= oy;
}
__device__ __forceinline__ void activate_checkerboard_pixel(int* x, int* y, bool previous_frame, uint32_t field) {  // RtxdiHelpers.hlsli:16-43
    if (field == 0u) return;
    uint32_t ux = (uint32_t)*x;
    const uint32_t uy = (uint32_t)*y;
    if (((ux + uy + (previous_frame ? 1u : 0u)) & 1u) == (field & 1u)) return;
    if (previous_frame) ux += (uint32_t)((int)field * 2 - 3);
    else ux += (uy & 1u) != 0u ? 1u : 0xFFFFFFFFu;
    *x = (int)ux;
}
__device__ __forceinline__ void clamp_sample_position_into_view(const rt2_gconst& g, int* x, int* y) {
    const int width = (int)g.view.viewport_size[0], height = (int)g.view.viewport_size[1];
    if (*x < 0) *x = -*x;
    if (*y < 0) *y = -*y;
    if (*x >= width) *x = 2 * width - *x - 1;
    if (*y >= height) *y = 2 * height - *y - 1;
}

}  // namespace rt2
