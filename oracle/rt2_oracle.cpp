// rt2_oracle.cpp — CPU oracle for RayTracer2's path-tracing hot path.
//
// TEST INFRASTRUCTURE ONLY.  PARITY UNPINNED (see rt2_oracle.h): a scalar C++17 restatement of the
// reference's GLSL, written from the shader sources, function by function, same names, same
// operation order.  Compile with -ffp-contract=off so every a*b+c below is two roundings unless
// std::fmaf is written explicitly.  Paths in the comments are relative to the reference tree
// (src/shaders/ unless they start with src/).
//
// Three things Vulkan leaves to the driver are DEFINED here and the CUDA kernels
// must reproduce them bit for bit:
//   (1) ray/triangle: Woop-Benthin-Wald watertight test in fp32, inclusive edge rule, hit iff
//       tmin < t < tmax, closest hit by (t, instance, geometry, primitive) total order;
//   (2) LOD-0 texture sampling: Vulkan §16 unnormalised coords u = s*W - 0.5, REPEAT wrap, fp32
//       weights, sRGB decode by the IEC 61966-2-1 formula evaluated in double then rounded;
//   (3) rgba16f stores: round-to-nearest-even.
// GLSL sin/cos are pinned to a Cody-Waite + Cephes-polynomial sincos (ORC_TRIG_LIBM switches to
// glibc for the statistical cross-check); normalize(v) is pinned to v / sqrt((x*x+y*y)+z*z).

#include "rt2_oracle.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <thread>
#include <vector>

namespace {

// ------------------------------------------------------------------------------------------------
// small vector types
// ------------------------------------------------------------------------------------------------
struct V3 { float x, y, z; };
struct V4 { float x, y, z, w; };
static inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
static inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline V3 operator*(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
static inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
static inline V3 operator/(V3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
static inline float dot3(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
static inline float comp(const V3& v, int k) { return k == 0 ? v.x : (k == 1 ? v.y : v.z); }
// GLSL normalize(vec3), pinned: v / sqrt((x*x + y*y) + z*z)
static inline V3 normalize3(V3 v) { return v / std::sqrt(dot3(v, v)); }
static inline float saturate(float x) { return std::fmin(std::fmax(x, 0.0f), 1.0f); }

static inline uint32_t f2u(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
static inline float u2f(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }

// column-major 4x4 (glam Mat4 / GLSL mat4): m[c*4 + r]
struct M4 { float m[16]; };
// GLSL `M * v`, pinned summation order ((c0*x + c1*y) + c2*z) + c3*w
static inline V4 mul(const M4& M, V4 v) {
    V4 r;
    r.x = ((M.m[0] * v.x + M.m[4] * v.y) + M.m[8] * v.z) + M.m[12] * v.w;
    r.y = ((M.m[1] * v.x + M.m[5] * v.y) + M.m[9] * v.z) + M.m[13] * v.w;
    r.z = ((M.m[2] * v.x + M.m[6] * v.y) + M.m[10] * v.z) + M.m[14] * v.w;
    r.w = ((M.m[3] * v.x + M.m[7] * v.y) + M.m[11] * v.z) + M.m[15] * v.w;
    return r;
}

// ------------------------------------------------------------------------------------------------
// GConst views (src/shader_params.rs:3-26,243-274; ShaderParameters.glsl:25-47,62-92)
// ------------------------------------------------------------------------------------------------
struct PlanarView {          // 464 bytes
    M4 matWorldToView, matViewToClip, matWorldToClip, matClipToView, matViewToWorld, matClipToWorld;
    float viewportOrigin[2], viewportSize[2];
    float viewportSizeInv[2], pixelOffset[2];
    float clipToWindowScale[2], clipToWindowBias[2];
    float windowToClipScale[2], windowToClipBias[2];
    float cameraDirectionOrPosition[4];
};
static_assert(sizeof(PlanarView) == 464, "PlanarViewConstants layout");
struct GConstView {
    PlanarView view, prevView;          // 0, 464
    uint32_t runtimeParams[4];          // 928
    uint32_t enableBrdfIndirect, enableBrdfAdditiveBlend, enableAccumulation, frame;  // 944
    uint32_t restirGI[36];              // 960  (144 bytes)
    uint32_t restirDI[48];              // 1104 (192 bytes)
    uint32_t enableRestirDI, enableRestirGI, refrenceMode, textures;   // 1296
    float blendFactor; uint32_t enableSpatialResampling, enableTemporalResampling, environment; // 1312
    uint32_t lightBufferParams[12];     // 1328
    uint32_t localLightsRIS[4], environmentLightRIS[4];  // 1376, 1392
    uint32_t environmentPdfTextureSize[2], localLightPdfTextureSize[2];  // 1408
};
static_assert(sizeof(GConstView) == RT2_GCONST_SIZE, "GConst layout");

// ------------------------------------------------------------------------------------------------
// integer helpers — rtxdi/RtxdiMath.hlsli:33-79
// ------------------------------------------------------------------------------------------------
static inline uint32_t RTXDI_IntegerExplode(uint32_t x) {         // RtxdiMath.hlsli:33-40
    x = (x | (x << 8)) & 0x00FF00FFu;
    x = (x | (x << 4)) & 0x0F0F0F0Fu;
    x = (x | (x << 2)) & 0x33333333u;
    x = (x | (x << 1)) & 0x55555555u;
    return x;
}
static inline uint32_t RTXDI_IntegerCompact(uint32_t x) {         // RtxdiMath.hlsli:45-52
    x = (x & 0x11111111u) | ((x & 0x44444444u) >> 1);
    x = (x & 0x03030303u) | ((x & 0x30303030u) >> 2);
    x = (x & 0x000F000Fu) | ((x & 0x0F000F00u) >> 4);
    x = (x & 0x000000FFu) | ((x & 0x00FF0000u) >> 8);
    return x;
}
static inline uint32_t RTXDI_ZCurveToLinearIndex(uint32_t x, uint32_t y) {   // :55-58
    return RTXDI_IntegerExplode(x) | (RTXDI_IntegerExplode(y) << 1);
}
static inline uint32_t RTXDI_JenkinsHash(uint32_t a) {            // RtxdiMath.hlsli:69-79
    a = (a + 0x7ed55d16u) + (a << 12);
    a = (a ^ 0xc761c23cu) ^ (a >> 19);
    a = (a + 0x165667b1u) + (a << 5);
    a = (a + 0xd3a2646cu) ^ (a << 9);
    a = (a + 0xfd7046c5u) + (a << 3);
    a = (a ^ 0xb55a4f09u) ^ (a >> 16);
    return a;
}

// ------------------------------------------------------------------------------------------------
// RNG — Helpers.glsl:7-64, bridge :378-387
// ------------------------------------------------------------------------------------------------
struct RAB_RandomSamplerState { uint32_t seed, index; };
static inline RAB_RandomSamplerState initRandomSampler(uint32_t px, uint32_t py, uint32_t frameIndex) {
    RAB_RandomSamplerState s;                                       // Helpers.glsl:13-23
    uint32_t linearPixelIndex = RTXDI_ZCurveToLinearIndex(px, py);
    s.index = 1;
    s.seed = RTXDI_JenkinsHash(linearPixelIndex) + frameIndex;
    return s;
}
static inline RAB_RandomSamplerState RAB_InitRandomSampler(uint32_t px, uint32_t py, uint32_t pass, uint32_t frame) {
    return initRandomSampler(px, py, frame + pass * 13u);           // bridge :378-381
}
static inline uint32_t rot32(uint32_t x, uint32_t y) { return (x << y) | (x >> (32 - y)); }
static inline uint32_t murmur3(RAB_RandomSamplerState& r) {         // Helpers.glsl:25-55
    const uint32_t c1 = 0xcc9e2d51u, c2 = 0x1b873593u, r1 = 15, r2 = 13, m = 5, n = 0xe6546b64u;
    uint32_t hash = r.seed;
    uint32_t k = r.index++;
    k *= c1; k = rot32(k, r1); k *= c2;
    hash ^= k;
    hash = rot32(hash, r2) * m + n;
    hash ^= 4;
    hash ^= (hash >> 16);
    hash *= 0x85ebca6bu;
    hash ^= (hash >> 13);
    hash *= 0xc2b2ae35u;
    hash ^= (hash >> 16);
    return hash;
}
static inline float sampleUniformRng(RAB_RandomSamplerState& r) {   // Helpers.glsl:57-63
    uint32_t v = murmur3(r);
    const uint32_t one = f2u(1.0f);
    const uint32_t mask = (1u << 23) - 1u;
    return u2f((mask & v) | one) - 1.0f;
}

// ------------------------------------------------------------------------------------------------
// pinned sincos (GLSL sin/cos precision is implementation-defined; Vulkan asks 2^-11 abs)
//   k = rint(a * 2/pi); r = a - k*pi/2 (two-step Cody-Waite with fma); Cephes sinf/cosf minimax
//   polynomials on [-pi/4, pi/4]; quadrant fix-up.  Valid for |a| < ~1e4 (we only see [0, 2pi)).
// ------------------------------------------------------------------------------------------------
static inline void sincos_pinned(float a, float* s, float* c) {
    const float TWO_OVER_PI = 0.636619772f;
    const float PIO2_HI = 1.57079625f;           // 0x3FC90FDA
    const float PIO2_LO = 7.54978942e-08f;       // pi/2 - PIO2_HI
    float kf = std::rint(a * TWO_OVER_PI);
    float r = std::fmaf(-kf, PIO2_HI, a);
    r = std::fmaf(-kf, PIO2_LO, r);
    float z = r * r;
    float ps = std::fmaf(z, -1.9515295891e-4f, 8.3321608736e-3f);
    ps = std::fmaf(ps, z, -1.6666654611e-1f);
    float sr = std::fmaf(ps * z, r, r);
    float pc = std::fmaf(z, 2.443315711809948e-5f, -1.388731625493765e-3f);
    pc = std::fmaf(pc, z, 4.166664568298827e-2f);
    float cr = std::fmaf(pc * z, z, std::fmaf(-0.5f, z, 1.0f));
    int q = (int)kf & 3;
    float ss = (q & 1) ? cr : sr;
    float cc = (q & 1) ? sr : cr;
    if (q == 1 || q == 2) cc = -cc;
    if (q >= 2) ss = -ss;
    *s = ss; *c = cc;
}
static inline void sincos_sel(float a, uint32_t flags, float* s, float* c) {
    if (flags & ORC_TRIG_LIBM) { *s = std::sin(a); *c = std::cos(a); }
    else sincos_pinned(a, s, c);
}

// ------------------------------------------------------------------------------------------------
// fp16 (rgba16f image stores: round-to-nearest-even)
// ------------------------------------------------------------------------------------------------
static inline uint16_t f32_to_f16(float f) {
    uint32_t x = f2u(f);
    uint32_t sign = (x >> 16) & 0x8000u;
    uint32_t ax = x & 0x7FFFFFFFu;
    if (ax >= 0x7F800000u) {                       // inf / nan
        return (uint16_t)(sign | 0x7C00u | ((ax > 0x7F800000u) ? (0x0200u | ((ax >> 13) & 0x3FFu)) : 0u));
    }
    if (ax >= 0x477FF000u) return (uint16_t)(sign | 0x7C00u);      // rounds to inf (>= 65520)
    if (ax < 0x33000001u) return (uint16_t)sign;                   // rounds to zero (<= 2^-25)
    int32_t e = (int32_t)(ax >> 23) - 127;
    uint32_t man = (ax & 0x7FFFFFu) | 0x800000u;
    uint32_t shift, half_bits;
    if (e < -14) {                                  // subnormal half
        shift = (uint32_t)(13 + (-14 - e));
        half_bits = man >> shift;
    } else {
        shift = 13;
        half_bits = ((uint32_t)(e + 15) << 10) + ((man >> 13) - 0x400u);
    }
    uint32_t rem = man & ((1u << shift) - 1u);
    uint32_t halfway = 1u << (shift - 1);
    if (rem > halfway || (rem == halfway && (half_bits & 1u))) half_bits++;
    return (uint16_t)(sign | half_bits);
}
static inline float f16_to_f32(uint16_t h) {
    uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
    uint32_t e = (h >> 10) & 0x1Fu, m = h & 0x3FFu;
    if (e == 0) {
        if (m == 0) return u2f(sign);
        float v = (float)m * 5.9604644775390625e-08f;  // 2^-24
        return (sign ? -v : v);
    }
    if (e == 31) return u2f(sign | 0x7F800000u | (m << 13));
    return u2f(sign | ((e + 112u) << 23) | (m << 13));
}

// ------------------------------------------------------------------------------------------------
// watertight ray/triangle test (pinned; replaces the driver's traceRayEXT intersection)
// ------------------------------------------------------------------------------------------------
struct RayShear { int kx, ky, kz; float Sx, Sy, Sz; };
static inline RayShear make_shear(V3 d) {
    RayShear s;
    float ax = std::fabs(d.x), ay = std::fabs(d.y), az = std::fabs(d.z);
    int kz = 0; float am = ax;
    if (ay > am) { kz = 1; am = ay; }
    if (az > am) { kz = 2; }
    int kx = kz + 1; if (kx == 3) kx = 0;
    int ky = kx + 1; if (ky == 3) ky = 0;
    if (comp(d, kz) < 0.0f) std::swap(kx, ky);
    s.kx = kx; s.ky = ky; s.kz = kz;
    s.Sx = comp(d, kx) / comp(d, kz);
    s.Sy = comp(d, ky) / comp(d, kz);
    s.Sz = 1.0f / comp(d, kz);
    return s;
}
static inline bool ray_triangle(V3 o, const RayShear& s, float tmin, float tmax,
                                V3 p0, V3 p1, V3 p2, float* t_out, float* u_out, float* v_out) {
    V3 A = p0 - o, B = p1 - o, C = p2 - o;
    float Akz = comp(A, s.kz), Bkz = comp(B, s.kz), Ckz = comp(C, s.kz);
    float Ax = std::fmaf(-s.Sx, Akz, comp(A, s.kx));
    float Ay = std::fmaf(-s.Sy, Akz, comp(A, s.ky));
    float Bx = std::fmaf(-s.Sx, Bkz, comp(B, s.kx));
    float By = std::fmaf(-s.Sy, Bkz, comp(B, s.ky));
    float Cx = std::fmaf(-s.Sx, Ckz, comp(C, s.kx));
    float Cy = std::fmaf(-s.Sy, Ckz, comp(C, s.ky));
    // edge functions: two roundings per product (an fma here would break edge antisymmetry)
    float U = Cx * By - Cy * Bx;
    float V = Ax * Cy - Ay * Cx;
    float W = Bx * Ay - By * Ax;
    // Woop / Benthin / Wald 2013, section 3.3: an edge function that comes out exactly zero is recomputed in binary64 (the
    // fp32 products are exact there, the difference is rounded once).  Without this a triangle seen edge-on from hundreds of
    // units away can cancel two edge functions to +-0 and be "hit" by a ray that passes nowhere near it (found by the
    // full-size brute-force comparison on the instanced field).
    if (U == 0.0f || V == 0.0f || W == 0.0f) {
        U = (float)((double)Cx * (double)By - (double)Cy * (double)Bx);
        V = (float)((double)Ax * (double)Cy - (double)Ay * (double)Cx);
        W = (float)((double)Bx * (double)Ay - (double)By * (double)Ax);
    }
    if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f)) return false;
    float det = (U + V) + W;
    if (det == 0.0f) return false;
    float Az = s.Sz * Akz, Bz = s.Sz * Bkz, Cz = s.Sz * Ckz;
    float T = std::fmaf(U, Az, std::fmaf(V, Bz, W * Cz));
    float rcp = 1.0f / det;
    float t = T * rcp;
    if (!(t > tmin && t < tmax)) return false;
    *t_out = t; *u_out = V * rcp; *v_out = W * rcp;
    return true;
}

// ------------------------------------------------------------------------------------------------
// scene
// ------------------------------------------------------------------------------------------------
struct Tri { V3 p0, p1, p2; uint32_t geom, prim; };
struct Hit { float t, u, v; uint32_t inst, geom, prim; };
static inline bool closer(float t, uint32_t inst, uint32_t geom, uint32_t prim, const Hit& h) {
    if (t < h.t) return true;
    if (t > h.t) return false;
    if (inst != h.inst) return inst < h.inst;
    if (geom != h.geom) return geom < h.geom;
    return prim < h.prim;
}

struct Box { V3 lo, hi; };
static inline Box empty_box() {
    const float inf = std::numeric_limits<float>::infinity();
    return {{inf, inf, inf}, {-inf, -inf, -inf}};
}
static inline void grow(Box& b, V3 p) {
    b.lo.x = std::fmin(b.lo.x, p.x); b.lo.y = std::fmin(b.lo.y, p.y); b.lo.z = std::fmin(b.lo.z, p.z);
    b.hi.x = std::fmax(b.hi.x, p.x); b.hi.y = std::fmax(b.hi.y, p.y); b.hi.z = std::fmax(b.hi.z, p.z);
}
static inline void grow(Box& b, const Box& o) { grow(b, o.lo); grow(b, o.hi); }
static inline float half_area(const Box& b) {
    float dx = b.hi.x - b.lo.x, dy = b.hi.y - b.lo.y, dz = b.hi.z - b.lo.z;
    return dx * dy + dy * dz + dz * dx;
}

struct BvhNode { Box box; uint32_t left, right, first, count; };   // count > 0 => leaf

// Binned-SAH BVH2 over arbitrary boxed items (triangles or instances).
struct Bvh2 {
    std::vector<BvhNode> nodes;
    std::vector<uint32_t> order;
    void build(const std::vector<Box>& boxes, uint32_t leaf_size) {
        uint32_t n = (uint32_t)boxes.size();
        order.resize(n);
        for (uint32_t i = 0; i < n; i++) order[i] = i;
        nodes.clear();
        nodes.reserve(2 * (size_t)n + 1);
        if (n == 0) return;
        std::vector<V3> cent(n);
        for (uint32_t i = 0; i < n; i++) cent[i] = (boxes[i].lo + boxes[i].hi) * 0.5f;
        struct Job { uint32_t node, first, count; };
        std::vector<Job> stack;
        nodes.push_back(BvhNode{});
        stack.push_back({0, 0, n});
        while (!stack.empty()) {
            Job j = stack.back(); stack.pop_back();
            Box bb = empty_box(), cb = empty_box();
            for (uint32_t i = j.first; i < j.first + j.count; i++) { grow(bb, boxes[order[i]]); grow(cb, cent[order[i]]); }
            BvhNode& nd = nodes[j.node];
            nd.box = bb; nd.first = j.first; nd.count = j.count; nd.left = nd.right = 0;
            if (j.count <= leaf_size) continue;
            V3 ext = cb.hi - cb.lo;
            int axis = 0; if (ext.y > ext.x) axis = 1; if (ext.z > comp(ext, axis)) axis = 2;
            float lo = comp(cb.lo, axis), e = comp(ext, axis);
            uint32_t mid;
            if (!(e > 0.0f)) {
                mid = j.first + j.count / 2;
            } else {
                const int NB = 16;
                Box bbx[NB]; uint32_t cnt[NB];
                for (int b = 0; b < NB; b++) { bbx[b] = empty_box(); cnt[b] = 0; }
                float scale = (float)NB / e;
                auto bin_of = [&](uint32_t id) { int b = (int)((comp(cent[id], axis) - lo) * scale); return b < 0 ? 0 : (b >= NB ? NB - 1 : b); };
                for (uint32_t i = j.first; i < j.first + j.count; i++) { int b = bin_of(order[i]); cnt[b]++; grow(bbx[b], boxes[order[i]]); }
                float la[NB], ra[NB]; uint32_t lc[NB], rc[NB];
                Box acc = empty_box(); uint32_t c = 0;
                for (int b = 0; b < NB; b++) { if (cnt[b]) grow(acc, bbx[b]); c += cnt[b]; la[b] = c ? half_area(acc) : 0.f; lc[b] = c; }
                acc = empty_box(); c = 0;
                for (int b = NB - 1; b >= 0; b--) { if (cnt[b]) grow(acc, bbx[b]); c += cnt[b]; ra[b] = c ? half_area(acc) : 0.f; rc[b] = c; }
                float best = std::numeric_limits<float>::infinity(); int bs = -1;
                for (int b = 0; b < NB - 1; b++) {
                    if (lc[b] == 0 || rc[b + 1] == 0) continue;
                    float cost = la[b] * (float)lc[b] + ra[b + 1] * (float)rc[b + 1];
                    if (cost < best) { best = cost; bs = b; }
                }
                if (bs < 0) mid = j.first + j.count / 2;
                else {
                    auto it = std::partition(order.begin() + j.first, order.begin() + j.first + j.count,
                                             [&](uint32_t id) { return bin_of(id) <= bs; });
                    mid = (uint32_t)(it - order.begin());
                    if (mid == j.first || mid == j.first + j.count) mid = j.first + j.count / 2;
                }
            }
            uint32_t l = (uint32_t)nodes.size();
            nodes.push_back(BvhNode{}); nodes.push_back(BvhNode{});
            nodes[j.node].left = l; nodes[j.node].right = l + 1; nodes[j.node].count = 0;
            stack.push_back({l, j.first, mid - j.first});
            stack.push_back({l + 1, mid, j.first + j.count - mid});
        }
    }
};

// conservative slab test (PBRT "robust" bounds): never rejects a box the exact ray touches
static inline bool ray_box(const Box& b, V3 o, V3 inv, float tmin, float tmax) {
    const float kGrow = 1.0f + 2.0f * (3.0f * 5.9604645e-08f) / (1.0f - 3.0f * 5.9604645e-08f);
    float t0 = tmin, t1 = tmax;
    for (int a = 0; a < 3; a++) {
        float io = comp(inv, a), oo = comp(o, a);
        float tn = (comp(b.lo, a) - oo) * io;
        float tf = (comp(b.hi, a) - oo) * io;
        if (tn > tf) std::swap(tn, tf);
        tf *= kGrow;
        tn = tn - std::fabs(tn) * 4e-7f;    // also loosen the near side
        t0 = tn > t0 ? tn : t0;
        t1 = tf < t1 ? tf : t1;
        if (t0 > t1) return false;
    }
    return true;
}

struct Texture { uint32_t w, h, filter; std::vector<uint8_t> rgba; };

}  // namespace

struct orc_scene {
    std::vector<rt2_vertex> vertices;
    std::vector<uint32_t> indices;
    std::vector<rt2_geometry_info> geoms;
    std::vector<uint32_t> index_counts;
    std::vector<Texture> textures;          // resolved (image, sampler) pairs
    std::vector<float> sky; uint32_t sky_w = 0, sky_h = 0;
    std::vector<Tri> tris;                  // BLAS-space triangles (geometry transform applied)
    Bvh2 blas;
    // instances
    bool instanced = false;
    std::vector<float> inst_fwd;            // 12 floats each, row-major 3x4
    std::vector<float> inst_inv;            // 12 floats each
    Bvh2 tlas;
    float srgb_lut[256];
};

namespace {

// VkTransformMatrixKHR (row-major 3x4) point / vector transforms; pinned order ((r0*x + r1*y) + r2*z) + r3
static inline V3 xform_point(const float* m, V3 p) {
    return { ((m[0] * p.x + m[1] * p.y) + m[2] * p.z) + m[3],
             ((m[4] * p.x + m[5] * p.y) + m[6] * p.z) + m[7],
             ((m[8] * p.x + m[9] * p.y) + m[10] * p.z) + m[11] };
}
static inline V3 xform_vector(const float* m, V3 v) {
    return { (m[0] * v.x + m[1] * v.y) + m[2] * v.z,
             (m[4] * v.x + m[5] * v.y) + m[6] * v.z,
             (m[8] * v.x + m[9] * v.y) + m[10] * v.z };
}
// inverse of an affine 3x4 in double (adjugate / determinant), rounded to float. Pinned order.
static void invert_affine(const float* m, float* out) {
    double a = m[0], b = m[1], c = m[2], d = m[4], e = m[5], f = m[6], g = m[8], h = m[9], i = m[10];
    double tx = m[3], ty = m[7], tz = m[11];
    double A = e * i - f * h, B = -(d * i - f * g), C = d * h - e * g;
    double det = a * A + b * B + c * C;
    double id = 1.0 / det;
    double r00 = A * id, r01 = -(b * i - c * h) * id, r02 = (b * f - c * e) * id;
    double r10 = B * id, r11 = (a * i - c * g) * id, r12 = -(a * f - c * d) * id;
    double r20 = C * id, r21 = -(a * h - b * g) * id, r22 = (a * e - b * d) * id;
    out[0] = (float)r00; out[1] = (float)r01; out[2] = (float)r02; out[3] = (float)(-(r00 * tx + r01 * ty + r02 * tz));
    out[4] = (float)r10; out[5] = (float)r11; out[6] = (float)r12; out[7] = (float)(-(r10 * tx + r11 * ty + r12 * tz));
    out[8] = (float)r20; out[9] = (float)r21; out[10] = (float)r22; out[11] = (float)(-(r20 * tx + r21 * ty + r22 * tz));
}

// closest hit inside one BLAS (ray already in BLAS space)
static void blas_intersect(const orc_scene& s, V3 o, V3 d, float tmin, float tmax, uint32_t inst,
                           bool brute, Hit& best) {
    RayShear sh = make_shear(d);
    if (brute || s.blas.nodes.empty()) {
        for (const Tri& tr : s.tris) {
            float t, u, v;
            if (ray_triangle(o, sh, tmin, tmax, tr.p0, tr.p1, tr.p2, &t, &u, &v) && closer(t, inst, tr.geom, tr.prim, best))
                best = Hit{t, u, v, inst, tr.geom, tr.prim};
        }
        return;
    }
    V3 inv = {1.0f / d.x, 1.0f / d.y, 1.0f / d.z};
    uint32_t stack[128]; int sp = 0;
    stack[sp++] = 0;
    while (sp) {
        const BvhNode& n = s.blas.nodes[stack[--sp]];
        float far = best.t < tmax ? best.t : tmax;
        // keep nodes whose entry equals the current best t (ties are resolved by id)
        if (!ray_box(n.box, o, inv, tmin, far)) continue;
        if (n.count) {
            for (uint32_t k = n.first; k < n.first + n.count; k++) {
                const Tri& tr = s.tris[s.blas.order[k]];
                float t, u, v;
                if (ray_triangle(o, sh, tmin, tmax, tr.p0, tr.p1, tr.p2, &t, &u, &v) && closer(t, inst, tr.geom, tr.prim, best))
                    best = Hit{t, u, v, inst, tr.geom, tr.prim};
            }
        } else {
            stack[sp++] = n.left; stack[sp++] = n.right;
        }
    }
}

// trace(ray): traceRayEXT(SceneBVH, opaque, 0xff, ...) — bridge :74-81; rayhit.rchit:60-66; raymiss.rmiss:12-17
static Hit trace(const orc_scene& s, V3 o, V3 d, float tmin, float tmax, bool brute) {
    Hit best; best.t = std::numeric_limits<float>::infinity(); best.u = best.v = 0.f;
    best.inst = best.geom = best.prim = 0xFFFFFFFFu;
    if (!s.instanced) { blas_intersect(s, o, d, tmin, tmax, 0, brute, best); return best; }
    uint32_t ninst = (uint32_t)(s.inst_fwd.size() / 12);
    auto visit = [&](uint32_t inst) {
        const float* iv = &s.inst_inv[(size_t)inst * 12];
        V3 oo = xform_point(iv, o), dd = xform_vector(iv, d);
        blas_intersect(s, oo, dd, tmin, tmax, inst, brute, best);
    };
    if (brute || s.tlas.nodes.empty()) { for (uint32_t i = 0; i < ninst; i++) visit(i); return best; }
    V3 inv = {1.0f / d.x, 1.0f / d.y, 1.0f / d.z};
    uint32_t stack[128]; int sp = 0; stack[sp++] = 0;
    while (sp) {
        const BvhNode& n = s.tlas.nodes[stack[--sp]];
        float far = best.t < tmax ? best.t : tmax;
        if (!ray_box(n.box, o, inv, tmin, far)) continue;
        if (n.count) { for (uint32_t k = n.first; k < n.first + n.count; k++) visit(s.tlas.order[k]); }
        else { stack[sp++] = n.left; stack[sp++] = n.right; }
    }
    return best;
}

// ------------------------------------------------------------------------------------------------
// texture(textures[idx], uv).rgb at implicit LOD 0 (ray-gen stage) — Hit.glsl:35-37;
// R8G8B8A8_SRGB, REPEAT (src/model.rs:224-287,165-183); Vulkan spec §16 texel selection.
// ------------------------------------------------------------------------------------------------
static inline int wrap_repeat(int i, int n) { int m = i % n; return m < 0 ? m + n : m; }
static V3 sample_texture(const orc_scene& s, uint32_t ti, float su, float sv) {
    const Texture& tx = s.textures[ti];
    const int W = (int)tx.w, H = (int)tx.h;
    auto texel = [&](int x, int y) -> V3 {
        const uint8_t* p = &tx.rgba[((size_t)y * tx.w + x) * 4];
        return { s.srgb_lut[p[0]], s.srgb_lut[p[1]], s.srgb_lut[p[2]] };
    };
    if (tx.filter == RT2_FILTER_NEAREST) {
        int x = wrap_repeat((int)std::floor(su * (float)W), W);
        int y = wrap_repeat((int)std::floor(sv * (float)H), H);
        return texel(x, y);
    }
    float u = su * (float)W - 0.5f, v = sv * (float)H - 0.5f;
    float fu = std::floor(u), fv = std::floor(v);
    float a = u - fu, b = v - fv;
    int x0 = wrap_repeat((int)fu, W), y0 = wrap_repeat((int)fv, H);
    int x1 = x0 + 1 == W ? 0 : x0 + 1, y1 = y0 + 1 == H ? 0 : y0 + 1;
    V3 t00 = texel(x0, y0), t10 = texel(x1, y0), t01 = texel(x0, y1), t11 = texel(x1, y1);
    float ia = 1.0f - a, ib = 1.0f - b;
    V3 top = t00 * ia + t10 * a;
    V3 bot = t01 * ia + t11 * a;
    return top * ib + bot * b;
}

// ------------------------------------------------------------------------------------------------
// GBufferHelpers.glsl:12-27 setupPrimaryRay
// ------------------------------------------------------------------------------------------------
struct RayDesc { V3 Origin, Direction; float TMin, TMax; };
static const float BACKGROUND_DEPTH = 100000.0f;                  // ShaderParameters.glsl:12
static RayDesc setupPrimaryRay(uint32_t px, uint32_t py, const PlanarView& view) {
    float pcx = (float)px + 0.5f, pcy = (float)py + 0.5f;
    float ux = pcx / view.viewportSize[0], uy = pcy / view.viewportSize[1];
    float dx = ux * 2.0f - 1.0f, dy = uy * 2.0f - 1.0f;
    V4 target = mul(view.matClipToView, V4{dx, dy, 1.0f, 1.0f});
    V3 nt = normalize3(v3(target.x, target.y, target.z));
    V4 direction = mul(view.matViewToWorld, V4{nt.x, nt.y, nt.z, 0.0f});
    RayDesc r;
    r.Origin = v3(view.cameraDirectionOrPosition[0], view.cameraDirectionOrPosition[1], view.cameraDirectionOrPosition[2]);
    r.Direction = v3(direction.x, direction.y, direction.z);
    r.TMin = 0.0f; r.TMax = BACKGROUND_DEPTH;
    return r;
}

// ------------------------------------------------------------------------------------------------
// Hit.glsl:2-42 GetGeometryFromHit
// ------------------------------------------------------------------------------------------------
struct Surface {        // RAB_Surface, bridge :83-94
    V3 worldPos, viewDir; float viewDepth; V3 normal, geoNormal, diffuseAlbedo, specularF0;
    float roughness, diffuseProbability;
};
static void GetGeometryFromHit(const orc_scene& s, const Hit& h, V3* normal, V3* specularF0, float* roughness,
                               V3* color, V3* emission) {
    const rt2_geometry_info& g = s.geoms[h.geom];
    uint32_t indexOffset = g.index_offset + 3u * h.prim;
    uint32_t i0 = g.vertex_offset + s.indices[indexOffset];
    uint32_t i1 = g.vertex_offset + s.indices[indexOffset + 1];
    uint32_t i2 = g.vertex_offset + s.indices[indexOffset + 2];
    const rt2_vertex &v0 = s.vertices[i0], &v1 = s.vertices[i1], &v2 = s.vertices[i2];
    float bx = 1.0f - h.u - h.v, by = h.u, bz = h.v;
    auto interp = [&](const float* a, const float* b, const float* c, int k) { return (a[k] * bx + b[k] * by) + c[k] * bz; };
    V3 n = normalize3(v3(interp(v0.normal, v1.normal, v2.normal, 0), interp(v0.normal, v1.normal, v2.normal, 1), interp(v0.normal, v1.normal, v2.normal, 2)));
    // normalize(geometryInfo.transform * vec4(normal, 0.0)).xyz — a vec4 normalize (Hit.glsl:27)
    M4 M; std::memcpy(M.m, g.transform, 64);
    V4 tn = mul(M, V4{n.x, n.y, n.z, 0.0f});
    float len = std::sqrt(((tn.x * tn.x + tn.y * tn.y) + tn.z * tn.z) + tn.w * tn.w);
    n = v3(tn.x / len, tn.y / len, tn.z / len);
    if (s.instanced) {  // A2 extension: shading normal additionally through the instance matrix
        const float* fw = &s.inst_fwd[(size_t)h.inst * 12];
        n = normalize3(xform_vector(fw, n));
    }
    float su = interp(v0.uv, v1.uv, v2.uv, 0), sv = interp(v0.uv, v1.uv, v2.uv, 1);
    V3 vc = v3(interp(v0.color, v1.color, v2.color, 0), interp(v0.color, v1.color, v2.color, 1), interp(v0.color, v1.color, v2.color, 2));
    V3 c = v3(g.base_color[0], g.base_color[1], g.base_color[2]) * vc;
    if (g.base_color_texture_index > -1) c = c * sample_texture(s, (uint32_t)g.base_color_texture_index, su, sv);
    *normal = n;
    float m = g.metallic_factor;    // mix(vec3(0), color, metallic) = 0*(1-a) + color*a
    *specularF0 = v3(0.0f * (1.0f - m) + c.x * m, 0.0f * (1.0f - m) + c.y * m, 0.0f * (1.0f - m) + c.z * m);
    *roughness = 1.0f;
    *color = c;
    *emission = v3(g.emission[0] * 12.0f, g.emission[1] * 12.0f, g.emission[2] * 12.0f);
}

// ------------------------------------------------------------------------------------------------
// environment — bridge :618-627, Helpers.glsl:242-248 (PI here is common.glsl:31's 3.1415926)
// sampler: LINEAR, CLAMP_TO_BORDER black, normalised coords (src/main.rs:123-138 builds the
// unnormalised one for the mip pass; the SkyBox binding of the lighting passes uses the default).
// ------------------------------------------------------------------------------------------------
// directionToEquirectUV, Helpers.glsl:242-248
static void directionToEquirectUV(V3 dir, float* u, float* v) {
    const float PI_ = 3.1415926f;
    // transcendental pin: binary64, rounded once (GLSL leaves atan / asin precision to the implementation)
    *u = 0.5f + (float)std::atan2((double)dir.z, (double)dir.x) / (2.0f * PI_);
    *v = 0.5f - (float)std::asin((double)dir.y) / PI_;
}
// texture(SkyBox, uv).rgb / textureLod(SkyBox, uv, 0).rgb (ray-generation shaders have no derivatives: LOD 0 either way)
static V3 sample_sky(const orc_scene& s, float u, float v) {
    const int W = (int)s.sky_w, H = (int)s.sky_h;
    float x = u * (float)W - 0.5f, y = v * (float)H - 0.5f;
    float fx = std::floor(x), fy = std::floor(y);
    float a = x - fx, b = y - fy;
    int x0 = (int)fx, y0 = (int)fy;
    auto texel = [&](int xx, int yy) -> V3 {
        if (xx < 0 || yy < 0 || xx >= W || yy >= H) return v3(0, 0, 0);
        const float* p = &s.sky[((size_t)yy * W + xx) * 4];
        return v3(p[0], p[1], p[2]);
    };
    V3 t00 = texel(x0, y0), t10 = texel(x0 + 1, y0), t01 = texel(x0, y0 + 1), t11 = texel(x0 + 1, y0 + 1);
    float ia = 1.0f - a, ib = 1.0f - b;
    V3 top = t00 * ia + t10 * a, bot = t01 * ia + t11 * a;
    return top * ib + bot * b;
}
static V3 GetEnvironmentRadiance(const orc_scene& s, const GConstView& g, V3 dir) {
    if (g.environment != 1u || s.sky.empty()) return v3(0, 0, 0);
    float u, v; directionToEquirectUV(dir, &u, &v);
    return sample_sky(s, u, v);
}

// ------------------------------------------------------------------------------------------------
// BRDF sampling — Helpers.glsl:112-126,171-179; bridge :118-128,437-461
// ------------------------------------------------------------------------------------------------
static const float RTXDI_PI = 3.1415926535f;                     // RtxdiMath.hlsli:14
static void ConstructONB(V3 n, V3* tangent, V3* bitangent) {       // Helpers.glsl:112-119
    float sign = (n.z >= 0.0f) ? 1.0f : -1.0f;
    float a = -1.0f / (sign + n.z);
    float b = n.x * n.y * a;
    *tangent = v3(1.0f + sign * n.x * n.x * a, sign * b, -sign * n.x);
    *bitangent = v3(b, sign + n.y * n.y * a, -n.y);
}
static V3 tangentToWorld(V3 normal, V3 h) {                        // bridge :118-128
    V3 t, b; ConstructONB(normal, &t, &b);
    return (b * h.x + t * h.y) + normal * h.z;
}
static V3 SampleCosHemisphere(float rx, float ry, uint32_t flags, float* pdf) {  // Helpers.glsl:171-179
    float angle = 2.0f * RTXDI_PI * rx;                            // SampleDisk, Helpers.glsl:122-126
    float sn, cs; sincos_sel(angle, flags, &sn, &cs);
    float r = std::sqrt(ry);
    float elevation = std::sqrt(saturate(1.0f - ry));
    *pdf = elevation / RTXDI_PI;
    return v3(cs * r, sn * r, elevation);
}
// RAB_GetSurfaceBrdfSample, bridge :437-461.  diffuseProbability == 1 (bridge :131-138) and
// rand.x < 1 always, so only the cosine lobe is reachable from K1.
static bool RAB_GetSurfaceBrdfSample(const Surface& surface, RAB_RandomSamplerState& rng, uint32_t flags, V3* dir) {
    float rx = sampleUniformRng(rng), ry = sampleUniformRng(rng), rz = sampleUniformRng(rng);
    (void)rx;
    float pdf;
    V3 h = SampleCosHemisphere(ry, rz, flags, &pdf);
    *dir = tangentToWorld(surface.normal, h);
    return dot3(surface.normal, *dir) > 0.0f;
}

// ------------------------------------------------------------------------------------------------
// ShadingHelpers.glsl:61-88 StoreShadingOutput (rgba16f images + an fp32 shadow of DiffuseLighting)
// ------------------------------------------------------------------------------------------------
static inline float mixf(float x, float y, float a) { return x * (1.0f - a) + y * a; }
static void StoreShadingOutput(const GConstView& g, size_t px, V3 diffuse, V3 specular, bool isFirstPass,
                               uint16_t* dimg, uint16_t* simg, float* d32) {
    V3 d_f32 = diffuse;
    if (g.enableAccumulation == 1u) {
        V3 pd = v3(f16_to_f32(dimg[px * 4 + 0]), f16_to_f32(dimg[px * 4 + 1]), f16_to_f32(dimg[px * 4 + 2]));
        float bf = g.blendFactor;
        diffuse = v3(mixf(pd.x, diffuse.x, bf), mixf(pd.y, diffuse.y, bf), mixf(pd.z, diffuse.z, bf));
        specular = v3(mixf(pd.x, diffuse.x, bf), mixf(pd.y, diffuse.y, bf), mixf(pd.z, diffuse.z, bf));  // sic (Q4)
        if (d32) d_f32 = v3(mixf(d32[px * 4 + 0], d_f32.x, bf), mixf(d32[px * 4 + 1], d_f32.y, bf), mixf(d32[px * 4 + 2], d_f32.z, bf));
    } else if (!isFirstPass) {
        V3 pd = v3(f16_to_f32(dimg[px * 4 + 0]), f16_to_f32(dimg[px * 4 + 1]), f16_to_f32(dimg[px * 4 + 2]));
        V3 ps = v3(f16_to_f32(simg[px * 4 + 0]), f16_to_f32(simg[px * 4 + 1]), f16_to_f32(simg[px * 4 + 2]));
        diffuse = diffuse + pd; specular = specular + ps;
        if (d32) d_f32 = d_f32 + v3(d32[px * 4 + 0], d32[px * 4 + 1], d32[px * 4 + 2]);
    }
    dimg[px * 4 + 0] = f32_to_f16(diffuse.x); dimg[px * 4 + 1] = f32_to_f16(diffuse.y);
    dimg[px * 4 + 2] = f32_to_f16(diffuse.z); dimg[px * 4 + 3] = 0;
    simg[px * 4 + 0] = f32_to_f16(specular.x); simg[px * 4 + 1] = f32_to_f16(specular.y);
    simg[px * 4 + 2] = f32_to_f16(specular.z); simg[px * 4 + 3] = 0;
    if (d32) { d32[px * 4 + 0] = d_f32.x; d32[px * 4 + 1] = d_f32.y; d32[px * 4 + 2] = d_f32.z; d32[px * 4 + 3] = 0.0f; }
}

// ------------------------------------------------------------------------------------------------
// refrence.rgen:18-47 main (K1)
// ------------------------------------------------------------------------------------------------
static uint64_t reference_pixel(const orc_scene& s, const GConstView& g, uint32_t frame, uint32_t px, uint32_t py,
                                uint32_t width, uint32_t spp, uint32_t maxBounces, uint32_t flags,
                                uint16_t* dimg, uint16_t* simg, float* d32) {
    const bool brute = (flags & ORC_BRUTE_FORCE) != 0;
    uint64_t rays = 0;
    RAB_RandomSamplerState rng = RAB_InitRandomSampler(px, py, 1, frame);
    V3 radiance = v3(0, 0, 0);
    for (uint32_t smp = 0; smp < spp; smp++) {
        V3 throughput = v3(1, 1, 1);
        RayDesc ray = setupPrimaryRay(px, py, g.view);
        ray.TMin = 0.001f;
        for (uint32_t i = 0; i < maxBounces; i++) {
            Hit h = trace(s, ray.Origin, ray.Direction, ray.TMin, ray.TMax, brute);   // GetSurface, Hit.glsl:44-71
            rays++;
            if (h.geom == 0xFFFFFFFFu) {
                V3 emission = GetEnvironmentRadiance(s, g, ray.Direction);
                radiance = radiance + throughput * emission;
                break;
            }
            Surface surface;
            V3 emission;
            GetGeometryFromHit(s, h, &surface.normal, &surface.specularF0, &surface.roughness, &surface.diffuseAlbedo, &emission);
            surface.geoNormal = surface.normal;
            surface.worldPos = ray.Origin + ray.Direction * h.t;
            surface.viewDir = ray.Direction;
            surface.viewDepth = h.t;
            surface.diffuseProbability = 1.0f;
            radiance = radiance + throughput * emission;
            throughput = throughput * surface.diffuseAlbedo;
            RAB_GetSurfaceBrdfSample(surface, rng, flags, &ray.Direction);
            ray.Origin = surface.worldPos;
        }
    }
    StoreShadingOutput(g, (size_t)py * width + px, radiance / (float)spp, v3(0, 0, 0), true, dimg, simg, d32);
    return rays;
}

template <class F>
static void parallel_rows(uint32_t row0, uint32_t row1, int threads, F f) {
    if (threads <= 0) threads = (int)std::thread::hardware_concurrency();
    if (threads <= 0) threads = 1;
    std::atomic<uint32_t> next(row0);
    auto work = [&]() { for (;;) { uint32_t y = next.fetch_add(1); if (y >= row1) break; f(y); } };
    if (threads == 1) { work(); return; }
    std::vector<std::thread> pool;
    for (int i = 0; i < threads; i++) pool.emplace_back(work);
    for (auto& t : pool) t.join();
}

}  // namespace

// ================================================================================================
// C API
// ================================================================================================
extern "C" {

orc_scene* orc_scene_create(const rt2_scene_desc* d) {
    if (!d) return nullptr;
    orc_scene* s = new orc_scene();
    s->vertices.assign(d->vertices, d->vertices + d->num_vertices);
    s->indices.assign(d->indices, d->indices + d->num_indices);
    s->geoms.assign(d->geometries, d->geometries + d->num_geometries);
    s->index_counts.assign(d->index_counts, d->index_counts + d->num_geometries);
    for (int i = 0; i < 256; i++) {
        double c = i / 255.0;
        double l = c <= 0.04045 ? c / 12.92 : std::pow((c + 0.055) / 1.055, 2.4);
        s->srgb_lut[i] = (float)l;
    }
    for (uint32_t t = 0; t < d->num_textures; t++) {
        const rt2_texture_desc& td = d->textures[t];
        Texture tx; tx.w = tx.h = 1; tx.filter = RT2_FILTER_LINEAR; tx.rgba.assign(4, 255);
        if (td.image_index < d->num_images) {
            const rt2_image_desc& im = d->images[td.image_index];
            tx.w = im.width; tx.h = im.height;
            tx.rgba.assign(im.rgba8, im.rgba8 + (size_t)im.width * im.height * 4);
        }
        if (td.sampler_index < d->num_samplers) tx.filter = d->samplers[td.sampler_index].mag_filter;  // LOD 0 => magnification
        s->textures.push_back(std::move(tx));
    }
    if (d->sky_rgba32f && d->sky_width && d->sky_height) {
        s->sky_w = d->sky_width; s->sky_h = d->sky_height;
        s->sky.assign(d->sky_rgba32f, d->sky_rgba32f + (size_t)d->sky_width * d->sky_height * 4);
    }
    // BLAS-space triangles: object->BLAS transform = node matrix, M*v (src/model.rs:193-212,439-445)
    for (uint32_t g = 0; g < d->num_geometries; g++) {
        const rt2_geometry_info& gi = s->geoms[g];
        M4 M; std::memcpy(M.m, gi.transform, 64);
        uint32_t ntri = s->index_counts[g] / 3;
        for (uint32_t p = 0; p < ntri; p++) {
            Tri tr; tr.geom = g; tr.prim = p;
            V3* dst[3] = {&tr.p0, &tr.p1, &tr.p2};
            for (int k = 0; k < 3; k++) {
                const rt2_vertex& v = s->vertices[gi.vertex_offset + s->indices[gi.index_offset + 3 * p + k]];
                // VkTransformMatrixKHR row r = (m[0][r], m[1][r], m[2][r], m[3][r])
                float rows[12] = { M.m[0], M.m[4], M.m[8], M.m[12], M.m[1], M.m[5], M.m[9], M.m[13], M.m[2], M.m[6], M.m[10], M.m[14] };
                *dst[k] = xform_point(rows, v3(v.pos[0], v.pos[1], v.pos[2]));
            }
            s->tris.push_back(tr);
        }
    }
    std::vector<Box> boxes(s->tris.size());
    for (size_t i = 0; i < s->tris.size(); i++) {
        Box b = empty_box(); grow(b, s->tris[i].p0); grow(b, s->tris[i].p1); grow(b, s->tris[i].p2);
        // The BVH has to be conservative with respect to ray_triangle, whose sheared coordinates carry an absolute error of
        // about |vertex| * 2^-23: a ray leaving a 1e4-unit triangle is reported to hit that triangle again at t ~ 1e-3 (just above
        // TMin), outside its exact box.  Pad every triangle box on all axes by its largest coordinate * 2^-16
        // (tests/test_adversarial_geometry.py found the miss; the brute-force path is the definition).
        const float m = std::fmax(std::fmax(std::fmax(std::fabs(b.lo.x), std::fabs(b.hi.x)), std::fmax(std::fabs(b.lo.y), std::fabs(b.hi.y))),
                                  std::fmax(std::fabs(b.lo.z), std::fabs(b.hi.z)));
        const float pad = m * 1.52587890625e-05f;
        b.lo = b.lo - v3(pad, pad, pad); b.hi = b.hi + v3(pad, pad, pad);
        boxes[i] = b;
    }
    s->blas.build(boxes, 4);
    if (d->num_instances > 0 && d->instance_transforms) {
        s->instanced = true;
        s->inst_fwd.assign(d->instance_transforms, d->instance_transforms + (size_t)d->num_instances * 12);
        s->inst_inv.resize(s->inst_fwd.size());
        Box bb = s->blas.nodes.empty() ? empty_box() : s->blas.nodes[0].box;
        std::vector<Box> ib(d->num_instances);
        for (uint32_t i = 0; i < d->num_instances; i++) {
            invert_affine(&s->inst_fwd[(size_t)i * 12], &s->inst_inv[(size_t)i * 12]);
            Box w = empty_box();
            for (int c = 0; c < 8; c++) {
                V3 p = v3((c & 1) ? bb.hi.x : bb.lo.x, (c & 2) ? bb.hi.y : bb.lo.y, (c & 4) ? bb.hi.z : bb.lo.z);
                grow(w, xform_point(&s->inst_fwd[(size_t)i * 12], p));
            }
            // pad: the TLAS box is only a culling aid, it must never be tighter than the BLAS content
            V3 e = (w.hi - w.lo) * 1e-4f + v3(1e-6f, 1e-6f, 1e-6f);
            w.lo = w.lo - e; w.hi = w.hi + e;
            ib[i] = w;
        }
        s->tlas.build(ib, 1);
    }
    return s;
}

void orc_scene_destroy(orc_scene* s) { delete s; }
uint64_t orc_scene_num_triangles(const orc_scene* s) { return s ? s->tris.size() : 0; }

int orc_trace_primary(const orc_scene* s, const void* gconst, uint32_t width, uint32_t height,
                      uint32_t flags, int threads, rt2_hit_record* out) {
    if (!s || !gconst || !out) return -1;
    GConstView g; std::memcpy(&g, gconst, sizeof g);
    const bool brute = (flags & ORC_BRUTE_FORCE) != 0;
    parallel_rows(0, height, threads, [&](uint32_t y) {
        for (uint32_t x = 0; x < width; x++) {
            RayDesc ray = setupPrimaryRay(x, y, g.view);
            ray.TMin = 0.001f;
            Hit h = trace(*s, ray.Origin, ray.Direction, ray.TMin, ray.TMax, brute);
            rt2_hit_record& r = out[(size_t)y * width + x];
            std::memset(&r, 0, sizeof r);
            r.instance = h.inst; r.geometry = h.geom; r.primitive = h.prim;
            if (h.geom != 0xFFFFFFFFu) { r.t = h.t; r.u = h.u; r.v = h.v; }
        }
    });
    return 0;
}

// arbitrary rays (test utility): rays[i] = {o.xyz, tmin, d.xyz, tmax}
int orc_trace_rays(const orc_scene* s, const float* rays, uint32_t n, uint32_t flags, rt2_hit_record* out) {
    if (!s || !rays || !out) return -1;
    const bool brute = (flags & ORC_BRUTE_FORCE) != 0;
    for (uint32_t i = 0; i < n; i++) {
        const float* r8 = rays + (size_t)i * 8;
        Hit h = trace(*s, v3(r8[0], r8[1], r8[2]), v3(r8[4], r8[5], r8[6]), r8[3], r8[7], brute);
        rt2_hit_record& r = out[i];
        std::memset(&r, 0, sizeof r);
        r.instance = h.inst; r.geometry = h.geom; r.primitive = h.prim;
        if (h.geom != 0xFFFFFFFFu) { r.t = h.t; r.u = h.u; r.v = h.v; }
    }
    return 0;
}

int orc_render_reference(const orc_scene* s, const void* gconst, uint32_t frame,
                         uint32_t width, uint32_t height, uint32_t spp, uint32_t max_bounces,
                         uint32_t row0, uint32_t row1, uint32_t flags, int threads,
                         uint16_t* diffuse_f16, uint16_t* specular_f16, float* diffuse_f32, uint64_t* rays) {
    if (!s || !gconst || !diffuse_f16 || !specular_f16) return -1;
    GConstView g; std::memcpy(&g, gconst, sizeof g);
    if (spp == 0) spp = 12;                 // refrence.rgen:17
    if (max_bounces == 0) max_bounces = 5;  // refrence.rgen:16
    if (row1 == 0 || row1 > height) row1 = height;
    std::atomic<uint64_t> total(0);
    parallel_rows(row0, row1, threads, [&](uint32_t y) {
        uint64_t r = 0;
        for (uint32_t x = 0; x < width; x++)
            r += reference_pixel(*s, g, frame, x, y, width, spp, max_bounces, flags, diffuse_f16, specular_f16, diffuse_f32);
        total += r;
    });
    if (rays) *rays = total.load();
    return 0;
}

uint32_t orc_zcurve_to_linear(uint32_t x, uint32_t y) { return RTXDI_ZCurveToLinearIndex(x, y); }
void orc_linear_to_zcurve(uint32_t i, uint32_t* x, uint32_t* y) { *x = RTXDI_IntegerCompact(i); *y = RTXDI_IntegerCompact(i >> 1); }
uint32_t orc_jenkins_hash(uint32_t a) { return RTXDI_JenkinsHash(a); }
void orc_rng_stream(uint32_t px, uint32_t py, uint32_t frame, uint32_t pass, uint32_t n, uint32_t* raw, float* uniform) {
    RAB_RandomSamplerState a = RAB_InitRandomSampler(px, py, pass, frame), b = a;
    for (uint32_t i = 0; i < n; i++) { if (raw) raw[i] = murmur3(a); if (uniform) uniform[i] = sampleUniformRng(b); }
}
uint16_t orc_f32_to_f16(float f) { return f32_to_f16(f); }
float orc_f16_to_f32(uint16_t h) { return f16_to_f32(h); }
void orc_sincos(float angle, uint32_t flags, float* s, float* c) { sincos_sel(angle, flags, s, c); }
int orc_ray_triangle(const float o[3], const float d[3], float tmin, float tmax,
                     const float v0[3], const float v1[3], const float v2[3], float* t, float* u, float* v) {
    RayShear sh = make_shear(v3(d[0], d[1], d[2]));
    return ray_triangle(v3(o[0], o[1], o[2]), sh, tmin, tmax, v3(v0[0], v0[1], v0[2]), v3(v1[0], v1[1], v1[2]), v3(v2[0], v2[1], v2[2]), t, u, v) ? 1 : 0;
}
void orc_primary_ray(const void* gconst, uint32_t px, uint32_t py, float origin[3], float dir[3]) {
    GConstView g; std::memcpy(&g, gconst, sizeof g);
    RayDesc r = setupPrimaryRay(px, py, g.view);
    origin[0] = r.Origin.x; origin[1] = r.Origin.y; origin[2] = r.Origin.z;
    dir[0] = r.Direction.x; dir[1] = r.Direction.y; dir[2] = r.Direction.z;
}
void orc_cos_hemisphere_dir(const float n[3], float r1, float r2, uint32_t flags, float dir[3]) {
    float pdf; V3 h = SampleCosHemisphere(r1, r2, flags, &pdf);
    V3 d = tangentToWorld(v3(n[0], n[1], n[2]), h);
    dir[0] = d.x; dir[1] = d.y; dir[2] = d.z;
}
void orc_sample_texture(const orc_scene* s, uint32_t ti, float u, float v, float rgb[3]) {
    V3 c = sample_texture(*s, ti, u, v); rgb[0] = c.x; rgb[1] = c.y; rgb[2] = c.z;
}

}  // extern "C"

#include "rt2_oracle_passes.inc"
#include "rt2_oracle_gi.inc"
