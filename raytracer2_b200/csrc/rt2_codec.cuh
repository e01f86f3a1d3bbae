// rt2_codec.cuh — device codecs and light encode/decode used by the light, G-buffer, reservoir and
// post-process kernels.  Citations are relative to src/shaders/ of the reference tree.
//
// GLSL leaves log2 / exp2 / pow / atan / asin implementation-defined; here they are evaluated in
// binary64 and rounded once to binary32 (the same rule the CPU oracle uses), so the packed integers
// these feed (log-radiance, gamma-packed colours, RGBA8 output) are reproducible.  -fmad=false applies.
#pragma once
#include "rt2_math.cuh"
#include "../../include/rt2_abi.h"

namespace rt2 {

__device__ __forceinline__ float log2_pinned(float x) { return (float)log2((double)x); }
__device__ __forceinline__ float exp2_pinned(float x) { return (float)exp2((double)x); }
__device__ __forceinline__ float pow_pinned(float x, float y) { return (float)pow((double)x, (double)y); }
// pow(x, 5.0) under the same pin (binary64, rounded once to binary32): four binary64 multiplications carry <= 2 ulp(binary64) of error,
// as libm's pow does, so the binary32 result differs from it only when x^5 lies within ~2^-52 of a rounding boundary (once in ~10^8
// evaluations: inside the tolerance the reservoir parity tests state for the binary64 transcendentals)
__device__ __forceinline__ float pow5_pinned(float x) { const double d = (double)x, d2 = d * d; return (float)(d2 * d2 * d); }
__device__ __forceinline__ float cos_pinned(float a) { float s, c; sincos_pinned(a, &s, &c); return c; }
__device__ __forceinline__ float sin_pinned(float a) { float s, c; sincos_pinned(a, &s, &c); return s; }
__device__ __forceinline__ float length3(V3 v) { return sqrtf(dot3(v, v)); }
__device__ __forceinline__ V3 cross3(V3 x, V3 y) { return v3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y); }
__device__ __forceinline__ V3 neg3(V3 v) { return v3(-v.x, -v.y, -v.z); }
__device__ __forceinline__ float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }

// ---- packing.glsl:3-79, Helpers.glsl:317-332 ----------------------------------------------------------
__device__ __forceinline__ uint32_t pack_ufloat(float r, uint32_t size, float d) {
    const uint32_t mask = (1u << size) - 1u;
    return (uint32_t)(long long)floorf(r * (float)mask + d) & mask;
}
__device__ __forceinline__ float unpack_ufloat(uint32_t r, uint32_t size) {
    const uint32_t mask = (1u << size) - 1u;
    return (float)(r & mask) / (float)mask;
}
__device__ __forceinline__ V3 unpack_r11g11b10(uint32_t p) { return v3(unpack_ufloat(p, 11), unpack_ufloat(p >> 11, 11), unpack_ufloat(p >> 22, 10)); }
__device__ __forceinline__ uint32_t pack_r11g11b10(V3 c) { return pack_ufloat(c.x, 11, 0.5f) | (pack_ufloat(c.y, 11, 0.5f) << 11) | (pack_ufloat(c.z, 10, 0.5f) << 22); }
__device__ __forceinline__ uint32_t pack_rgba8_gamma(float r, float g, float b, float a) {
    const float ig = 1.0f / 2.2f;
    // pow(0, y) = 0 and pow(1, y) = 1 exactly: dielectrics have F0 = 0 and the reference's roughness is the constant 1 (Hit.glsl:40), so
    // most G-buffer pixels never need the binary64 pow
    auto enc = [&](float x) { x = clampf(x, 0.f, 1.f); return x == 0.f ? 0.f : (x == 1.f ? 1.f : pow_pinned(x, ig)); };
    r = enc(r); g = enc(g); b = enc(b); a = enc(a);
    return pack_ufloat(r, 8, 0.5f) | (pack_ufloat(g, 8, 0.5f) << 8) | (pack_ufloat(b, 8, 0.5f) << 16) | (pack_ufloat(a, 8, 0.5f) << 24);
}
__device__ __forceinline__ float gamma22_decode(uint32_t byte) { return pow_pinned(saturatef(unpack_ufloat(byte, 8)), 2.2f); }
// `lut` = CodecLuts::gamma22 (the 256 values of gamma22_decode, tabulated on this device by k_build_codec_luts): the G-buffer's
// specular/roughness word is decoded at every stage of every ReSTIR pass and for every neighbour K6 / K7 look at — four binary64
// pow() per decode otherwise.  Bit-identical to the inline form by construction.
__device__ __forceinline__ float4 unpack_rgba8_gamma(uint32_t p, const float* __restrict__ lut = nullptr) {
    if (lut) return make_float4(__ldg(lut + (p & 0xffu)), __ldg(lut + ((p >> 8) & 0xffu)), __ldg(lut + ((p >> 16) & 0xffu)), __ldg(lut + (p >> 24)));
    return make_float4(gamma22_decode(p), gamma22_decode(p >> 8), gamma22_decode(p >> 16), gamma22_decode(p >> 24));
}
__device__ __forceinline__ V3 unpack_r8g8b8(uint32_t p) { return v3(unpack_ufloat(p, 8), unpack_ufloat(p >> 8, 8), unpack_ufloat(p >> 16, 8)); }
__device__ __forceinline__ uint32_t pack_r8g8b8(V3 c) { return pack_ufloat(c.x, 8, 0.5f) | (pack_ufloat(c.y, 8, 0.5f) << 8) | (pack_ufloat(c.z, 8, 0.5f) << 16); }
__device__ __forceinline__ uint32_t pack_r16g16_float(float r, float g) { return (uint32_t)f32_to_f16_bits(r) | ((uint32_t)f32_to_f16_bits(g) << 16); }

// ---- octahedral normals: Helpers.glsl:251-268, packing.glsl:114-133 --------------------------------
__device__ __forceinline__ uint32_t ndir_to_oct_unorm32(V3 n) {
    const float inv = 1.f / ((fabsf(n.x) + fabsf(n.y)) + fabsf(n.z));
    float x = n.x * inv, y = n.y * inv;
    if (n.z < 0.f) {                                  // octWrap: one sign for both lanes, as written
        const float sgn = (x >= 0.f || y >= 0.f) ? 1.f : -1.f;
        const float wx = (1.f - fabsf(y)) * sgn, wy = (1.f - fabsf(x)) * sgn;
        x = wx; y = wy;
    }
    x = saturatef(x * 0.5f + 0.5f); y = saturatef(y * 0.5f + 0.5f);
    return (uint32_t)(x * 65534.0f) | ((uint32_t)(y * 65534.0f) << 16);
}
__device__ __forceinline__ V3 oct_to_ndir_signed(float px, float py) {
    V3 n = v3(px, py, (1.0f - fabsf(px)) - fabsf(py));
    const float t = fmaxf(0.f, -n.z);
    const float add = (n.x >= 0.0f || n.y >= 0.0f) ? -t : t;
    n.x += add; n.y += add;
    return normalize3(n);
}
__device__ __forceinline__ V3 oct_to_ndir_unorm32(uint32_t p) {
    const float x = clampf((float)(p & 0xffffu) / 65534.0f, 0.f, 1.f), y = clampf((float)(p >> 16) / 65534.0f, 0.f, 1.f);
    return oct_to_ndir_signed(x * 2.0f - 1.0f, y * 2.0f - 1.0f);
}

// ---- PolymorphicLight.glsl (triangle lights) ---------------------------------------------------------
#define RT2_LIGHT_TRIANGLE 4u
#define RT2_LIGHT_TYPE_SHIFT 24u
#define RT2_LIGHT_TYPE_MASK 0xfu
#define RT2_LIGHT_MIN_LOG2 (-8.f)
#define RT2_LIGHT_MAX_LOG2 (40.f)
__device__ __forceinline__ float calc_luminance(V3 c) { return dot3(c, v3(0.299f, 0.587f, 0.114f)); }
// `lut` = CodecLuts::light_radiance: the function below for all 65536 codes (a binary64 exp2 per light candidate otherwise)
__device__ __forceinline__ float unpack_light_radiance(uint32_t lr, const float* __restrict__ lut = nullptr) {   // :65-68
    if (lut) return __ldg(lut + (lr & 0xffffu));
    if (lr == 0u) return 0.f;
    return exp2_pinned(((float)(lr - 1u) / 65534.0f) * (RT2_LIGHT_MAX_LOG2 - RT2_LIGHT_MIN_LOG2) + RT2_LIGHT_MIN_LOG2);
}
struct TriangleLight { V3 base, edge1, edge2, radiance, normal; float surface_area; };
__device__ __forceinline__ TriangleLight create_triangle_light(const rt2_light_info& li, const float* __restrict__ radiance_lut = nullptr) {    // :320-343
    TriangleLight t;
    t.edge1 = oct_to_ndir_unorm32(li.direction1) * f16_bits_to_f32((uint16_t)(li.scalars & 0xffffu));
    t.edge2 = oct_to_ndir_unorm32(li.direction2) * f16_bits_to_f32((uint16_t)(li.scalars >> 16));
    t.base = v3(li.center[0], li.center[1], li.center[2]) - (t.edge1 + t.edge2) / 3.0f;
    t.radiance = unpack_r8g8b8(li.color_type_and_flags) * unpack_light_radiance(li.log_radiance & 0xffffu, radiance_lut);
    const V3 ln = cross3(t.edge1, t.edge2);
    const float len = length3(ln);
    if (len > 0.0f) { t.surface_area = 0.5f * len; t.normal = ln / len; }
    else { t.surface_area = 0.0f; t.normal = v3(0.f, 0.f, 0.f); }
    return t;
}
__device__ __forceinline__ rt2_light_info store_triangle_light(const TriangleLight& s) {      // :345-357, :77-93 (Q3: zero-initialised)
    rt2_light_info li;
    li.color_type_and_flags = 0u; li.log_radiance = 0u; li.ies_profile_index = 0u; li.primary_axis = 0u;
    li.cos_cone_angle_and_softness = 0u; li.padding = 0u;
    const float intensity = fmaxf(s.radiance.x, fmaxf(s.radiance.y, s.radiance.z));
    if (intensity > 0.0f) {
        const float lr = saturatef((log2_pinned(intensity) - RT2_LIGHT_MIN_LOG2) / (RT2_LIGHT_MAX_LOG2 - RT2_LIGHT_MIN_LOG2));
        const uint32_t packed = min((uint32_t)ceilf(lr * 65534.0f) + 1u, 0xffffu);
        const float unpacked = unpack_light_radiance(packed);
        li.log_radiance |= packed;
        li.color_type_and_flags |= pack_r8g8b8(v3(saturatef(s.radiance.x / unpacked), saturatef(s.radiance.y / unpacked), saturatef(s.radiance.z / unpacked)));
    }
    const V3 c = s.base + (s.edge1 + s.edge2) / 3.0f;
    li.center[0] = c.x; li.center[1] = c.y; li.center[2] = c.z;
    li.direction1 = ndir_to_oct_unorm32(normalize3(s.edge1));
    li.direction2 = ndir_to_oct_unorm32(normalize3(s.edge2));
    li.scalars = (uint32_t)f32_to_f16_bits(length3(s.edge1)) | ((uint32_t)f32_to_f16_bits(length3(s.edge2)) << 16);
    li.color_type_and_flags |= RT2_LIGHT_TRIANGLE << RT2_LIGHT_TYPE_SHIFT;
    return li;
}
// EnvironmentLight, :359-424.  The reference names one at index lights + 1 (src/main.rs:381-386) and never writes it;
// K10 writes this record there (pinned to intent, read only with RT2_LAUNCH_PRESAMPLING): radianceScale 1 through packLightColor,
// importance sampled, texture = the SkyBox binding, size = the environment PDF texture's.  rotation = 0.25 turns reconciles
// equirectUVToDirection's azimuth (u + 0.25) * 2pi (Helpers.glsl:336) with the u = 0.5 + atan2(z, x) / 2pi every sky lookup
// uses (Helpers.glsl:244) — see oracle/rt2_oracle_passes.inc StoreEnvironmentLight.
#define RT2_LIGHT_ENVIRONMENT 6u
__device__ __forceinline__ rt2_light_info store_environment_light(uint32_t tex_w, uint32_t tex_h) {
    rt2_light_info li;
    li.center[0] = li.center[1] = li.center[2] = 0.f;
    li.color_type_and_flags = 0u; li.log_radiance = 0u; li.ies_profile_index = 0u; li.primary_axis = 0u;
    li.cos_cone_angle_and_softness = 0u; li.padding = 0u;
    const float intensity = 1.0f;                                                              // packLightColor(vec3(1)), :77-93
    const float lr = saturatef((log2_pinned(intensity) - RT2_LIGHT_MIN_LOG2) / (RT2_LIGHT_MAX_LOG2 - RT2_LIGHT_MIN_LOG2));
    const uint32_t packed = min((uint32_t)ceilf(lr * 65534.0f) + 1u, 0xffffu);
    const float unpacked = unpack_light_radiance(packed);
    li.log_radiance |= packed;
    li.color_type_and_flags |= pack_r8g8b8(v3(saturatef(1.0f / unpacked), saturatef(1.0f / unpacked), saturatef(1.0f / unpacked)));
    li.color_type_and_flags |= RT2_LIGHT_ENVIRONMENT << RT2_LIGHT_TYPE_SHIFT;
    li.direction1 = 0u;                                                                        // textureIndex >= 0
    li.scalars = (uint32_t)f32_to_f16_bits(0.25f) | (1u << 16);                                // rotation | importanceSampled
    li.direction2 = (tex_w & 0xffffu) | (tex_h << 16);
    return li;
}
__device__ __forceinline__ float light_power(const rt2_light_info& li) {                      // getPower :454-471
    if (((li.color_type_and_flags >> RT2_LIGHT_TYPE_SHIFT) & RT2_LIGHT_TYPE_MASK) != RT2_LIGHT_TRIANGLE) return 0.f;
    const TriangleLight t = create_triangle_light(li);
    return t.surface_area * RT2_RTXDI_PI * calc_luminance(t.radiance);
}

}  // namespace rt2
