// rt2_math.cuh — device math for the path, with a pinned operation order.
//
// Everything here is compiled with -fmad=false: a*b+c is two roundings unless rt2_fma() is written,
// which is what makes results bit-identical to the CPU oracle (built with -ffp-contract=off).
// Functions are __host__ __device__ so tests/hostemu can exercise the same logic without a GPU;
// the shipped library only ever calls them from kernels.
//
// Reference citations are relative to src/shaders/ of the reference tree.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <math.h>
#include <stdint.h>

#define RT2_HD __host__ __device__ __forceinline__

namespace rt2 {

RT2_HD float rt2_fma(float a, float b, float c) {
#ifdef __CUDA_ARCH__
    return __fmaf_rn(a, b, c);
#else
    return fmaf(a, b, c);
#endif
}
RT2_HD uint32_t f2u(float f) {
#ifdef __CUDA_ARCH__
    return __float_as_uint(f);
#else
    uint32_t u; memcpy(&u, &f, 4); return u;
#endif
}
RT2_HD float u2f(uint32_t u) {
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}

struct V3 { float x, y, z; };
RT2_HD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
RT2_HD V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
RT2_HD V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
RT2_HD V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
RT2_HD V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
RT2_HD V3 operator/(V3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
RT2_HD float dot3(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
// GLSL normalize(vec3), pinned: v / sqrt((x*x + y*y) + z*z)   (IEEE sqrt and divide)
RT2_HD V3 normalize3(V3 v) { return v / sqrtf(dot3(v, v)); }
RT2_HD float saturatef(float x) { return fminf(fmaxf(x, 0.0f), 1.0f); }
RT2_HD float comp(const V3& v, int k) { return k == 0 ? v.x : (k == 1 ? v.y : v.z); }
// comp() as two SELs.  In the triangle loop of k_trace nvcc turned the nine comp() calls of ray_triangle into seven
// branch diamonds (64 of the loop's 158 SASS instructions, run at 6-7 active lanes); selp keeps them straight-line.
RT2_HD float comp_sel(const V3& v, int k) {
#ifdef __CUDA_ARCH__
    float r;
    asm("{\n\t.reg .pred p0, p1;\n\t.reg .f32 t;\n\tsetp.eq.s32 p0, %4, 0;\n\tsetp.eq.s32 p1, %4, 1;\n\t"
        "selp.f32 t, %2, %3, p1;\n\tselp.f32 %0, %1, t, p0;\n\t}" : "=f"(r) : "f"(v.x), "f"(v.y), "f"(v.z), "r"(k));
    return r;
#else
    return comp(v, k);
#endif
}

// GLSL `M * v` for column-major M (m[c*4+r]); pinned order ((c0*x + c1*y) + c2*z) + c3*w
RT2_HD void mat4_mul_vec4(const float* m, float x, float y, float z, float w, float out[4]) {
    out[0] = ((m[0] * x + m[4] * y) + m[8] * z) + m[12] * w;
    out[1] = ((m[1] * x + m[5] * y) + m[9] * z) + m[13] * w;
    out[2] = ((m[2] * x + m[6] * y) + m[10] * z) + m[14] * w;
    out[3] = ((m[3] * x + m[7] * y) + m[11] * z) + m[15] * w;
}
// VkTransformMatrixKHR (3x4 row-major) * point / vector
RT2_HD V3 xform_point(const float* m, V3 p) {
    return v3(((m[0] * p.x + m[1] * p.y) + m[2] * p.z) + m[3],
              ((m[4] * p.x + m[5] * p.y) + m[6] * p.z) + m[7],
              ((m[8] * p.x + m[9] * p.y) + m[10] * p.z) + m[11]);
}
RT2_HD V3 xform_vector(const float* m, V3 v) {
    return v3((m[0] * v.x + m[1] * v.y) + m[2] * v.z,
              (m[4] * v.x + m[5] * v.y) + m[6] * v.z,
              (m[8] * v.x + m[9] * v.y) + m[10] * v.z);
}

// ---- rtxdi/RtxdiMath.hlsli:33-79 ---------------------------------------------------------------------
RT2_HD uint32_t integer_explode(uint32_t x) {
    x = (x | (x << 8)) & 0x00FF00FFu;
    x = (x | (x << 4)) & 0x0F0F0F0Fu;
    x = (x | (x << 2)) & 0x33333333u;
    x = (x | (x << 1)) & 0x55555555u;
    return x;
}
RT2_HD uint32_t integer_compact(uint32_t x) {
    x = (x & 0x11111111u) | ((x & 0x44444444u) >> 1);
    x = (x & 0x03030303u) | ((x & 0x30303030u) >> 2);
    x = (x & 0x000F000Fu) | ((x & 0x0F000F00u) >> 4);
    x = (x & 0x000000FFu) | ((x & 0x00FF0000u) >> 8);
    return x;
}
RT2_HD uint32_t zcurve_to_linear(uint32_t x, uint32_t y) { return integer_explode(x) | (integer_explode(y) << 1); }
RT2_HD void linear_to_zcurve(uint32_t i, uint32_t& x, uint32_t& y) { x = integer_compact(i); y = integer_compact(i >> 1); }
RT2_HD uint32_t jenkins_hash(uint32_t a) {
    a = (a + 0x7ed55d16u) + (a << 12);
    a = (a ^ 0xc761c23cu) ^ (a >> 19);
    a = (a + 0x165667b1u) + (a << 5);
    a = (a + 0xd3a2646cu) ^ (a << 9);
    a = (a + 0xfd7046c5u) + (a << 3);
    a = (a ^ 0xb55a4f09u) ^ (a >> 16);
    return a;
}

// ---- RNG: Helpers.glsl:7-64; RAB_InitRandomSampler bridge :378-381 -----------------------------------
struct Rng { uint32_t seed, index; };
RT2_HD Rng rng_init(uint32_t px, uint32_t py, uint32_t frame, uint32_t pass) {
    Rng r; r.index = 1u; r.seed = jenkins_hash(zcurve_to_linear(px, py)) + (frame + pass * 13u); return r;
}
RT2_HD uint32_t rotl32(uint32_t x, uint32_t y) { return (x << y) | (x >> (32u - y)); }
RT2_HD uint32_t murmur3(Rng& r) {
    uint32_t hash = r.seed;
    uint32_t k = r.index++;
    k *= 0xcc9e2d51u; k = rotl32(k, 15u); k *= 0x1b873593u;
    hash ^= k;
    hash = rotl32(hash, 13u) * 5u + 0xe6546b64u;
    hash ^= 4u;
    hash ^= (hash >> 16);
    hash *= 0x85ebca6bu;
    hash ^= (hash >> 13);
    hash *= 0xc2b2ae35u;
    hash ^= (hash >> 16);
    return hash;
}
RT2_HD float rng_next(Rng& r) { return u2f((murmur3(r) & 0x007FFFFFu) | 0x3F800000u) - 1.0f; }

// ---- pinned sincos (see oracle/rt2_oracle.cpp sincos_pinned: same constants, same op order) ----------
RT2_HD void sincos_pinned(float a, float* s, float* c) {
    const float TWO_OVER_PI = 0.636619772f;
    const float PIO2_HI = 1.57079625f;
    const float PIO2_LO = 7.54978942e-08f;
    float kf = rintf(a * TWO_OVER_PI);
    float r = rt2_fma(-kf, PIO2_HI, a);
    r = rt2_fma(-kf, PIO2_LO, r);
    float z = r * r;
    float ps = rt2_fma(z, -1.9515295891e-4f, 8.3321608736e-3f);
    ps = rt2_fma(ps, z, -1.6666654611e-1f);
    float sr = rt2_fma(ps * z, r, r);
    float pc = rt2_fma(z, 2.443315711809948e-5f, -1.388731625493765e-3f);
    pc = rt2_fma(pc, z, 4.166664568298827e-2f);
    float cr = rt2_fma(pc * z, z, rt2_fma(-0.5f, z, 1.0f));
    int q = (int)kf & 3;
    float ss = (q & 1) ? cr : sr;
    float cc = (q & 1) ? sr : cr;
    if (q == 1 || q == 2) cc = -cc;
    if (q >= 2) ss = -ss;
    *s = ss; *c = cc;
}

// ---- fp16 store (rgba16f images): round-to-nearest-even --------------------------------------------
RT2_HD uint16_t f32_to_f16_bits(float f) {
#ifdef __CUDA_ARCH__
    return __half_as_ushort(__float2half_rn(f));
#else
    uint32_t x = f2u(f), sign = (x >> 16) & 0x8000u, ax = x & 0x7FFFFFFFu;
    if (ax >= 0x7F800000u) return (uint16_t)(sign | 0x7C00u | ((ax > 0x7F800000u) ? (0x0200u | ((ax >> 13) & 0x3FFu)) : 0u));
    if (ax >= 0x477FF000u) return (uint16_t)(sign | 0x7C00u);
    if (ax < 0x33000001u) return (uint16_t)sign;
    int32_t e = (int32_t)(ax >> 23) - 127;
    uint32_t man = (ax & 0x7FFFFFu) | 0x800000u, shift, hb;
    if (e < -14) { shift = (uint32_t)(13 + (-14 - e)); hb = man >> shift; }
    else { shift = 13; hb = ((uint32_t)(e + 15) << 10) + ((man >> 13) - 0x400u); }
    uint32_t rem = man & ((1u << shift) - 1u), half = 1u << (shift - 1);
    if (rem > half || (rem == half && (hb & 1u))) hb++;
    return (uint16_t)(sign | hb);
#endif
}
RT2_HD float f16_bits_to_f32(uint16_t h) {
#ifdef __CUDA_ARCH__
    return __half2float(__ushort_as_half(h));
#else
    uint32_t sign = (uint32_t)(h & 0x8000u) << 16, e = (h >> 10) & 0x1Fu, m = h & 0x3FFu;
    if (e == 0) { if (m == 0) return u2f(sign); float v = (float)m * 5.9604644775390625e-08f; return sign ? -v : v; }
    if (e == 31) return u2f(sign | 0x7F800000u | (m << 13));
    return u2f(sign | ((e + 112u) << 23) | (m << 13));
#endif
}

// ---- watertight ray/triangle (pinned; same op order as the oracle's ray_triangle) ---------------------
struct RayShear { int kx, ky, kz; float Sx, Sy, Sz; };
RT2_HD RayShear make_shear(V3 d) {
    RayShear s;
    float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    int kz = 0; float am = ax;
    if (ay > am) { kz = 1; am = ay; }
    if (az > am) { kz = 2; }
    int kx = kz + 1; if (kx == 3) kx = 0;
    int ky = kx + 1; if (ky == 3) ky = 0;
    float dz = comp(d, kz);
    if (dz < 0.0f) { int t = kx; kx = ky; ky = t; }
    s.kx = kx; s.ky = ky; s.kz = kz;
    s.Sx = comp(d, kx) / dz;
    s.Sy = comp(d, ky) / dz;
    s.Sz = 1.0f / dz;
    return s;
}
RT2_HD bool ray_triangle(V3 o, const RayShear& s, float tmin, float tmax, V3 p0, V3 p1, V3 p2,
                         float* t_out, float* u_out, float* v_out) {
    V3 A = p0 - o, B = p1 - o, C = p2 - o;
    float Akz = comp_sel(A, s.kz), Bkz = comp_sel(B, s.kz), Ckz = comp_sel(C, s.kz);
    float Ax = rt2_fma(-s.Sx, Akz, comp_sel(A, s.kx));
    float Ay = rt2_fma(-s.Sy, Akz, comp_sel(A, s.ky));
    float Bx = rt2_fma(-s.Sx, Bkz, comp_sel(B, s.kx));
    float By = rt2_fma(-s.Sy, Bkz, comp_sel(B, s.ky));
    float Cx = rt2_fma(-s.Sx, Ckz, comp_sel(C, s.kx));
    float Cy = rt2_fma(-s.Sy, Ckz, comp_sel(C, s.ky));
    float U = Cx * By - Cy * Bx;      // two roundings per product: keeps edge functions antisymmetric
    float V = Ax * Cy - Ay * Cx;
    float W = Bx * Ay - By * Ax;
    // an edge function that is exactly zero is recomputed in binary64 (Woop / Benthin / Wald 2013, section 3.3; same rule as the
    // oracle): fp32 products are exact in binary64 and the difference is rounded once, so the sign is the true one.  Rare on
    // the device (a BVH only offers triangles the ray passes close to), so the branch costs one compare on the common path.
    if (U == 0.0f || V == 0.0f || W == 0.0f) {
        U = (float)((double)Cx * (double)By - (double)Cy * (double)Bx);
        V = (float)((double)Ax * (double)Cy - (double)Ay * (double)Cx);
        W = (float)((double)Bx * (double)Ay - (double)By * (double)Ax);
    }
    if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f)) return false;
    float det = (U + V) + W;
    if (det == 0.0f) return false;
    float Az = s.Sz * Akz, Bz = s.Sz * Bkz, Cz = s.Sz * Ckz;
    float T = rt2_fma(U, Az, rt2_fma(V, Bz, W * Cz));
    float rcp = 1.0f / det;
    float t = T * rcp;
    if (!(t > tmin && t < tmax)) return false;
    *t_out = t; *u_out = V * rcp; *v_out = W * rcp;
    return true;
}

// Hit record carried between the trace and shade stages.  Total order (t, instance, geometry,
// primitive) makes the closest hit independent of traversal order.
struct Hit { float t, u, v; uint32_t inst, geom, prim; };
RT2_HD bool hit_closer(float t, uint32_t inst, uint32_t geom, uint32_t prim, const Hit& h) {
    if (t < h.t) return true;
    if (t > h.t) return false;
    if (inst != h.inst) return inst < h.inst;
    if (geom != h.geom) return geom < h.geom;
    return prim < h.prim;
}

// ---- Helpers.glsl:112-126,171-179; bridge :118-128 ----------------------------------------------------
#define RT2_RTXDI_PI 3.1415926535f        /* RtxdiMath.hlsli:14 */
RT2_HD void construct_onb(V3 n, V3* tangent, V3* bitangent) {
    float sign = (n.z >= 0.0f) ? 1.0f : -1.0f;
    float a = -1.0f / (sign + n.z);
    float b = n.x * n.y * a;
    *tangent = v3(1.0f + sign * n.x * n.x * a, sign * b, -sign * n.x);
    *bitangent = v3(b, sign + n.y * n.y * a, -n.y);
}
RT2_HD V3 tangent_to_world(V3 normal, V3 h) {
    V3 t, b; construct_onb(normal, &t, &b);
    return (b * h.x + t * h.y) + normal * h.z;
}
// SampleCosHemisphere(random = (rx, ry)) with the pinned or libdevice sincos
RT2_HD V3 sample_cos_hemisphere(float rx, float ry, bool libdevice_trig) {
    float angle = 2.0f * RT2_RTXDI_PI * rx;
    float sn, cs;
    if (libdevice_trig) { sn = sinf(angle); cs = cosf(angle); } else sincos_pinned(angle, &sn, &cs);
    float r = sqrtf(ry);
    float elevation = sqrtf(saturatef(1.0f - ry));
    return v3(cs * r, sn * r, elevation);
}

RT2_HD float mixf(float x, float y, float a) { return x * (1.0f - a) + y * a; }

}  // namespace rt2
