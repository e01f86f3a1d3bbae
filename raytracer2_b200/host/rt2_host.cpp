// rt2_host.cpp — host-side mirror of the reference's CPU arithmetic for the path + procedural
// scene generators.  See rt2_host.h.  No CUDA here; pure C++17.
#include "rt2_host.h"
#include "rt2_ingest.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <memory>
#include <string>
#include <vector>

namespace {

struct F3 { float x, y, z; };
static inline F3 f3(float x, float y, float z) { return {x, y, z}; }
static inline F3 operator+(F3 a, F3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline F3 operator-(F3 a, F3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline F3 operator*(F3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
static inline float dot(F3 a, F3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline F3 cross(F3 a, F3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
static inline F3 normalize(F3 a) { float l = std::sqrt(dot(a, a)); return l > 0 ? a * (1.0f / l) : f3(0, 1, 0); }
static inline float lerp(float a, float b, float t) { return (1.0f - t) * a + t * b; }

// ---- 4x4 column-major helpers (glam Mat4 semantics) ----------------------------------------------
static void mat_mul(const float* a, const float* b, float* out) {   // out = a * b
    float r[16];
    for (int c = 0; c < 4; c++)
        for (int rr = 0; rr < 4; rr++) {
            float s = 0.f;
            for (int k = 0; k < 4; k++) s += a[k * 4 + rr] * b[c * 4 + k];
            r[c * 4 + rr] = s;
        }
    std::memcpy(out, r, sizeof r);
}
// general 4x4 inverse by cofactors (what glam::Mat4::inverse computes; rounding may differ in the
// last ulp from glam's SIMD schedule — this is host-side input preparation, both the oracle and
// the CUDA path receive the same bytes).
static void mat_inverse(const float* m, float* out) {
    float inv[16];
    inv[0] = m[5] * m[10] * m[15] - m[5] * m[11] * m[14] - m[9] * m[6] * m[15] + m[9] * m[7] * m[14] + m[13] * m[6] * m[11] - m[13] * m[7] * m[10];
    inv[4] = -m[4] * m[10] * m[15] + m[4] * m[11] * m[14] + m[8] * m[6] * m[15] - m[8] * m[7] * m[14] - m[12] * m[6] * m[11] + m[12] * m[7] * m[10];
    inv[8] = m[4] * m[9] * m[15] - m[4] * m[11] * m[13] - m[8] * m[5] * m[15] + m[8] * m[7] * m[13] + m[12] * m[5] * m[11] - m[12] * m[7] * m[9];
    inv[12] = -m[4] * m[9] * m[14] + m[4] * m[10] * m[13] + m[8] * m[5] * m[14] - m[8] * m[6] * m[13] - m[12] * m[5] * m[10] + m[12] * m[6] * m[9];
    inv[1] = -m[1] * m[10] * m[15] + m[1] * m[11] * m[14] + m[9] * m[2] * m[15] - m[9] * m[3] * m[14] - m[13] * m[2] * m[11] + m[13] * m[3] * m[10];
    inv[5] = m[0] * m[10] * m[15] - m[0] * m[11] * m[14] - m[8] * m[2] * m[15] + m[8] * m[3] * m[14] + m[12] * m[2] * m[11] - m[12] * m[3] * m[10];
    inv[9] = -m[0] * m[9] * m[15] + m[0] * m[11] * m[13] + m[8] * m[1] * m[15] - m[8] * m[3] * m[13] - m[12] * m[1] * m[11] + m[12] * m[3] * m[9];
    inv[13] = m[0] * m[9] * m[14] - m[0] * m[10] * m[13] - m[8] * m[1] * m[14] + m[8] * m[2] * m[13] + m[12] * m[1] * m[10] - m[12] * m[2] * m[9];
    inv[2] = m[1] * m[6] * m[15] - m[1] * m[7] * m[14] - m[5] * m[2] * m[15] + m[5] * m[3] * m[14] + m[13] * m[2] * m[7] - m[13] * m[3] * m[6];
    inv[6] = -m[0] * m[6] * m[15] + m[0] * m[7] * m[14] + m[4] * m[2] * m[15] - m[4] * m[3] * m[14] - m[12] * m[2] * m[7] + m[12] * m[3] * m[6];
    inv[10] = m[0] * m[5] * m[15] - m[0] * m[7] * m[13] - m[4] * m[1] * m[15] + m[4] * m[3] * m[13] + m[12] * m[1] * m[7] - m[12] * m[3] * m[5];
    inv[14] = -m[0] * m[5] * m[14] + m[0] * m[6] * m[13] + m[4] * m[1] * m[14] - m[4] * m[2] * m[13] - m[12] * m[1] * m[6] + m[12] * m[2] * m[5];
    inv[3] = -m[1] * m[6] * m[11] + m[1] * m[7] * m[10] + m[5] * m[2] * m[11] - m[5] * m[3] * m[10] - m[9] * m[2] * m[7] + m[9] * m[3] * m[6];
    inv[7] = m[0] * m[6] * m[11] - m[0] * m[7] * m[10] - m[4] * m[2] * m[11] + m[4] * m[3] * m[10] + m[8] * m[2] * m[7] - m[8] * m[3] * m[6];
    inv[11] = -m[0] * m[5] * m[11] + m[0] * m[7] * m[9] + m[4] * m[1] * m[11] - m[4] * m[3] * m[9] - m[8] * m[1] * m[7] + m[8] * m[3] * m[5];
    inv[15] = m[0] * m[5] * m[10] - m[0] * m[6] * m[9] - m[4] * m[1] * m[10] + m[4] * m[2] * m[9] + m[8] * m[1] * m[6] - m[8] * m[2] * m[5];
    float det = m[0] * inv[0] + m[1] * inv[4] + m[2] * inv[8] + m[3] * inv[12];
    float id = 1.0f / det;
    for (int i = 0; i < 16; i++) out[i] = inv[i] * id;
}
static void mat_identity(float* m) { std::memset(m, 0, 64); m[0] = m[5] = m[10] = m[15] = 1.0f; }

// ---- hashing / noise --------------------------------------------------------------------------------
static inline uint32_t pcg_hash(uint32_t v) {
    uint32_t s = v * 747796405u + 2891336453u;
    uint32_t w = ((s >> ((s >> 28) + 4)) ^ s) * 277803737u;
    return (w >> 22) ^ w;
}
static inline uint32_t hash3(uint32_t a, uint32_t b, uint32_t c) { return pcg_hash(a ^ pcg_hash(b ^ pcg_hash(c ^ 0x9E3779B9u))); }
static inline float hash01(uint32_t a, uint32_t b, uint32_t c) { return (float)(hash3(a, b, c) >> 8) * (1.0f / 16777216.0f); }
static float value_noise(float x, float y, uint32_t seed) {
    float fx = std::floor(x), fy = std::floor(y);
    int ix = (int)fx, iy = (int)fy;
    float tx = x - fx, ty = y - fy;
    tx = tx * tx * (3 - 2 * tx); ty = ty * ty * (3 - 2 * ty);
    float a = hash01((uint32_t)ix, (uint32_t)iy, seed), b = hash01((uint32_t)(ix + 1), (uint32_t)iy, seed);
    float c = hash01((uint32_t)ix, (uint32_t)(iy + 1), seed), d = hash01((uint32_t)(ix + 1), (uint32_t)(iy + 1), seed);
    return lerp(lerp(a, b, tx), lerp(c, d, tx), ty);
}
struct Lcg64 {   // geometry RNG: 64-bit LCG, seed 0x5254325F + config id
    uint64_t s;
    explicit Lcg64(uint64_t seed) : s(seed * 2862933555777941757ULL + 3037000493ULL) {}
    uint32_t next() { s = s * 6364136223846793005ULL + 1442695040888963407ULL; return (uint32_t)(s >> 33); }
    float uni() { return (float)(next() >> 7) * (1.0f / 16777216.0f); }
    float range(float a, float b) { return a + (b - a) * uni(); }
};

struct Material {
    float base_color[3] = {1, 1, 1};
    float emission[3] = {0, 0, 0};
    int texture = -1;
    float metallic = 0.0f;
    float roughness = 1.0f;
};

struct PatchPoint { F3 pos; F3 normal; };

}  // namespace

struct rt2h_scene {
    int kind = 0;
    std::vector<rt2_vertex> verts;
    std::vector<uint32_t> idx;
    std::vector<rt2_geometry_info> geoms;
    std::vector<uint32_t> index_counts;
    std::vector<std::vector<uint8_t>> image_data;
    std::vector<rt2_image_desc> images;
    std::vector<rt2_sampler_desc> samplers;
    std::vector<rt2_texture_desc> textures;
    std::vector<float> sky; uint32_t sky_w = 0, sky_h = 0;
    std::vector<float> instances;
    rt2_scene_desc desc{};
    rt2h_camera cam{};

    void begin_geometry(const Material& m, const float* transform_cols /* 16 or null */) {
        rt2_geometry_info g{};
        if (transform_cols) std::memcpy(g.transform, transform_cols, 64); else mat_identity(g.transform);
        g.base_color[0] = m.base_color[0]; g.base_color[1] = m.base_color[1]; g.base_color[2] = m.base_color[2]; g.base_color[3] = 1.0f;
        g.base_color_texture_index = m.texture;
        g.metallic_factor = m.metallic;
        g.index_offset = (uint32_t)idx.size();
        g.vertex_offset = (uint32_t)verts.size();
        g.emission[0] = m.emission[0]; g.emission[1] = m.emission[1]; g.emission[2] = m.emission[2]; g.emission[3] = 1.0f;
        g.roughness = m.roughness;
        geoms.push_back(g);
        index_counts.push_back(0);
    }
    uint32_t local_vertex_count() const { return (uint32_t)verts.size() - geoms.back().vertex_offset; }
    uint32_t add_vertex(F3 p, F3 n, float u, float v, F3 color = f3(1, 1, 1)) {
        rt2_vertex vx{};
        vx.pos[0] = p.x; vx.pos[1] = p.y; vx.pos[2] = p.z; vx.pos[3] = 0.f;
        vx.normal[0] = n.x; vx.normal[1] = n.y; vx.normal[2] = n.z; vx.normal[3] = 0.f;
        vx.color[0] = color.x; vx.color[1] = color.y; vx.color[2] = color.z; vx.color[3] = 1.f;
        vx.uv[0] = u; vx.uv[1] = v;
        uint32_t li = local_vertex_count();
        verts.push_back(vx);
        return li;
    }
    void add_tri(uint32_t a, uint32_t b, uint32_t c) { idx.push_back(a); idx.push_back(b); idx.push_back(c); index_counts.back() += 3; }
    // quad a-b-c-d (a->b->c->d around), normal n
    void add_quad(F3 a, F3 b, F3 c, F3 d, F3 n) {
        uint32_t i0 = add_vertex(a, n, 0, 0), i1 = add_vertex(b, n, 1, 0), i2 = add_vertex(c, n, 1, 1), i3 = add_vertex(d, n, 0, 1);
        add_tri(i0, i1, i2); add_tri(i0, i2, i3);
    }
    // (nu x nv) quad grid over f(u,v); normals from f; uv = (u*uscale, v*vscale)
    void add_patch(int nu, int nv, const std::function<PatchPoint(float, float)>& f, float uscale, float vscale, bool recompute_normals) {
        uint32_t base = local_vertex_count();
        std::vector<PatchPoint> pts((size_t)(nu + 1) * (nv + 1));
        for (int j = 0; j <= nv; j++)
            for (int i = 0; i <= nu; i++) pts[(size_t)j * (nu + 1) + i] = f((float)i / (float)nu, (float)j / (float)nv);
        auto P = [&](int i, int j) -> const PatchPoint& { i = std::clamp(i, 0, nu); j = std::clamp(j, 0, nv); return pts[(size_t)j * (nu + 1) + i]; };
        for (int j = 0; j <= nv; j++)
            for (int i = 0; i <= nu; i++) {
                const PatchPoint& p = P(i, j);
                F3 n = p.normal;
                if (recompute_normals) {
                    F3 du = P(i + 1, j).pos - P(i - 1, j).pos, dv = P(i, j + 1).pos - P(i, j - 1).pos;
                    F3 c = cross(du, dv);
                    if (dot(c, c) > 0) { c = normalize(c); if (dot(c, p.normal) < 0) c = c * -1.0f; n = c; }
                }
                add_vertex(p.pos, n, (float)i / (float)nu * uscale, (float)j / (float)nv * vscale);
            }
        for (int j = 0; j < nv; j++)
            for (int i = 0; i < nu; i++) {
                uint32_t a = base + (uint32_t)(j * (nu + 1) + i), b = a + 1, c = a + (uint32_t)(nu + 1) + 1, d = a + (uint32_t)(nu + 1);
                add_tri(a, b, c); add_tri(a, c, d);
            }
    }
    int add_texture(uint32_t w, uint32_t h, std::vector<uint8_t>&& px, uint32_t filter) {
        image_data.push_back(std::move(px));
        rt2_image_desc im{image_data.back().data(), w, h};
        images.push_back(im);
        // sampler 0 is the loader's default sampler (src/gltf/mod.rs); one more per filter kind
        if (samplers.empty()) { samplers.push_back({RT2_FILTER_LINEAR, RT2_FILTER_LINEAR}); samplers.push_back({RT2_FILTER_NEAREST, RT2_FILTER_NEAREST}); }
        textures.push_back({(uint32_t)images.size() - 1, filter == RT2_FILTER_LINEAR ? 0u : 1u});
        return (int)textures.size() - 1;
    }
    void finish() {
        // image_data vectors may have been moved while growing: refresh pointers
        for (size_t i = 0; i < images.size(); i++) images[i].rgba8 = image_data[i].data();
        desc.vertices = verts.data(); desc.num_vertices = verts.size();
        desc.indices = idx.data(); desc.num_indices = idx.size();
        desc.geometries = geoms.data(); desc.index_counts = index_counts.data(); desc.num_geometries = (uint32_t)geoms.size();
        desc.images = images.empty() ? nullptr : images.data(); desc.num_images = (uint32_t)images.size();
        desc.samplers = samplers.empty() ? nullptr : samplers.data(); desc.num_samplers = (uint32_t)samplers.size();
        desc.textures = textures.empty() ? nullptr : textures.data(); desc.num_textures = (uint32_t)textures.size();
        desc.sky_rgba32f = sky.empty() ? nullptr : sky.data(); desc.sky_width = sky_w; desc.sky_height = sky_h;
        desc.instance_transforms = instances.empty() ? nullptr : instances.data(); desc.num_instances = (uint32_t)(instances.size() / 12);
    }
};

namespace {

static void trs_matrix(F3 t, float yaw, float pitch, float scale, float* m /* col-major */) {
    float cy = std::cos(yaw), sy = std::sin(yaw), cp = std::cos(pitch), sp = std::sin(pitch);
    // R = Ry(yaw) * Rx(pitch)
    float r[9] = { cy, sy * sp, sy * cp,   0, cp, -sp,   -sy, cy * sp, cy * cp };   // row-major
    mat_identity(m);
    for (int c = 0; c < 3; c++) for (int rr = 0; rr < 3; rr++) m[c * 4 + rr] = r[rr * 3 + c] * scale;
    m[12] = t.x; m[13] = t.y; m[14] = t.z;
}

static std::vector<uint8_t> make_texture(uint32_t size, uint32_t material, uint32_t seed) {
    std::vector<uint8_t> px((size_t)size * size * 4);
    float c0[3], c1[3];
    for (int k = 0; k < 3; k++) { c0[k] = 0.45f + 0.5f * hash01(material, 11 + k, seed); c1[k] = 0.25f + 0.5f * hash01(material, 17 + k, seed); }
    uint32_t pattern = hash3(material, 5, seed) % 4;
    uint32_t cells = 4u << (hash3(material, 6, seed) % 3);
    for (uint32_t y = 0; y < size; y++)
        for (uint32_t x = 0; x < size; x++) {
            float u = (float)x / (float)size, v = (float)y / (float)size;
            float sel;
            switch (pattern) {
                case 0: sel = (float)(((uint32_t)(u * cells) + (uint32_t)(v * cells)) & 1u); break;                       // checker
                case 1: { float row = std::floor(v * cells); float uu = u * cells * 0.5f + 0.5f * (float)((int)row & 1);     // bricks
                          float fu = uu - std::floor(uu), fv = v * cells - row; sel = (fu < 0.06f || fv < 0.12f) ? 1.f : 0.f; } break;
                case 2: sel = 0.5f + 0.5f * std::sin(6.2831853f * (u + 0.3f * v) * (float)cells); break;                     // stripes
                default: sel = value_noise(u * cells * 2, v * cells * 2, seed + material); break;                            // blotches
            }
            float n = 0.85f + 0.3f * (hash01(x, y, material * 977u + seed) - 0.5f);
            uint8_t* p = &px[((size_t)y * size + x) * 4];
            for (int k = 0; k < 3; k++) {
                float c = lerp(c0[k], c1[k], sel) * n;
                p[k] = (uint8_t)std::clamp((int)(c * 255.0f + 0.5f), 0, 255);
            }
            p[3] = 255;
        }
    return px;
}

// ---- C1: Cornell box -----------------------------------------------------------------------------------
static void gen_cornell(rt2h_scene& s) {
    Material white, red, green, light;
    white.base_color[0] = white.base_color[1] = white.base_color[2] = 0.73f;
    red.base_color[0] = 0.65f; red.base_color[1] = 0.05f; red.base_color[2] = 0.05f;
    green.base_color[0] = 0.12f; green.base_color[1] = 0.45f; green.base_color[2] = 0.15f;
    light.base_color[0] = light.base_color[1] = light.base_color[2] = 0.78f;
    light.emission[0] = light.emission[1] = light.emission[2] = 1.0f;   // x12 in the shader (Hit.glsl:41)
    // room [-1,1]^3, open towards +z (the camera side)
    s.begin_geometry(white, nullptr); s.add_quad(f3(-1, -1, 1), f3(1, -1, 1), f3(1, -1, -1), f3(-1, -1, -1), f3(0, 1, 0));     // floor
    s.begin_geometry(white, nullptr); s.add_quad(f3(-1, 1, -1), f3(1, 1, -1), f3(1, 1, 1), f3(-1, 1, 1), f3(0, -1, 0));        // ceiling
    s.begin_geometry(white, nullptr); s.add_quad(f3(-1, -1, -1), f3(1, -1, -1), f3(1, 1, -1), f3(-1, 1, -1), f3(0, 0, 1));     // back
    s.begin_geometry(red, nullptr);   s.add_quad(f3(-1, -1, 1), f3(-1, -1, -1), f3(-1, 1, -1), f3(-1, 1, 1), f3(1, 0, 0));      // left
    s.begin_geometry(green, nullptr); s.add_quad(f3(1, -1, -1), f3(1, -1, 1), f3(1, 1, 1), f3(1, 1, -1), f3(-1, 0, 0));         // right
    s.begin_geometry(light, nullptr); s.add_quad(f3(-0.3f, 0.995f, -0.3f), f3(0.3f, 0.995f, -0.3f), f3(0.3f, 0.995f, 0.3f), f3(-0.3f, 0.995f, 0.3f), f3(0, -1, 0));
    auto box = [&](F3 c, float hx, float hy, float hz, float yaw) {
        float m[16]; trs_matrix(c, yaw, 0.f, 1.0f, m);
        s.begin_geometry(white, m);
        F3 v[8]; for (int i = 0; i < 8; i++) v[i] = f3((i & 1) ? hx : -hx, (i & 2) ? hy : -hy, (i & 4) ? hz : -hz);
        s.add_quad(v[2], v[3], v[7], v[6], f3(0, 1, 0));     // top
        s.add_quad(v[0], v[2], v[6], v[4], f3(-1, 0, 0));    // -x
        s.add_quad(v[1], v[5], v[7], v[3], f3(1, 0, 0));     // +x
        s.add_quad(v[0], v[1], v[3], v[2], f3(0, 0, -1));    // -z
        s.add_quad(v[4], v[6], v[7], v[5], f3(0, 0, 1));     // +z
    };
    box(f3(0.35f, -0.7f, 0.3f), 0.3f, 0.3f, 0.3f, -0.3f);
    box(f3(-0.35f, -0.4f, -0.3f), 0.3f, 0.6f, 0.3f, 0.35f);
    float pos[3] = {0.f, 0.f, 3.4f}, dir[3] = {0.f, 0.f, 1.f};   // rays leave along -direction
    rt2h_camera_new(&s.cam, pos, dir, 65.0f, 1.0f, 0.1f, 1000.0f);
}

// ---- furnace ---------------------------------------------------------------------------------------------
static void gen_furnace(rt2h_scene& s) {
    Material m; m.base_color[0] = m.base_color[1] = m.base_color[2] = 0.5f;
    m.emission[0] = m.emission[1] = m.emission[2] = 0.25f;
    s.begin_geometry(m, nullptr);
    s.add_quad(f3(-1, -1, 1), f3(1, -1, 1), f3(1, -1, -1), f3(-1, -1, -1), f3(0, 1, 0));
    s.add_quad(f3(-1, 1, -1), f3(1, 1, -1), f3(1, 1, 1), f3(-1, 1, 1), f3(0, -1, 0));
    s.add_quad(f3(-1, -1, -1), f3(1, -1, -1), f3(1, 1, -1), f3(-1, 1, -1), f3(0, 0, 1));
    s.add_quad(f3(-1, -1, 1), f3(-1, -1, -1), f3(-1, 1, -1), f3(-1, 1, 1), f3(1, 0, 0));
    s.add_quad(f3(1, -1, -1), f3(1, -1, 1), f3(1, 1, 1), f3(1, 1, -1), f3(-1, 0, 0));
    s.add_quad(f3(1, -1, 1), f3(-1, -1, 1), f3(-1, 1, 1), f3(1, 1, 1), f3(0, 0, -1));
    float pos[3] = {0.1f, -0.2f, 0.3f}, dir[3] = {0.2f, 0.1f, 1.f};
    rt2h_camera_new(&s.cam, pos, dir, 65.0f, 1.0f, 0.1f, 1000.0f);
}

// ---- random soup --------------------------------------------------------------------------------------------
static void gen_random(rt2h_scene& s, float scale, uint64_t seed) {
    Lcg64 rng(0x5254325FULL + 7 + seed);
    int ngeom = 6, per = std::max(4, (int)(400 * scale));
    for (int g = 0; g < ngeom; g++) {
        Material m; for (int k = 0; k < 3; k++) m.base_color[k] = rng.range(0.2f, 0.9f);
        if (g == 2) { m.emission[0] = 0.5f; m.emission[1] = 0.4f; m.emission[2] = 0.3f; }
        float mat[16];
        if (g == 2 || g == 0) mat_identity(mat);
        else trs_matrix(f3(rng.range(-0.3f, 0.3f), rng.range(-0.3f, 0.3f), rng.range(-0.3f, 0.3f)), rng.range(0, 6.28f), rng.range(-0.5f, 0.5f), rng.range(0.7f, 1.4f), mat);
        s.begin_geometry(m, mat);
        for (int t = 0; t < per; t++) {
            F3 c = f3(rng.range(-1, 1), rng.range(-1, 1), rng.range(-1, 1));
            float sz = rng.range(0.02f, 0.35f);
            F3 a = c + f3(rng.range(-sz, sz), rng.range(-sz, sz), rng.range(-sz, sz));
            F3 b = c + f3(rng.range(-sz, sz), rng.range(-sz, sz), rng.range(-sz, sz));
            F3 d = c + f3(rng.range(-sz, sz), rng.range(-sz, sz), rng.range(-sz, sz));
            F3 n = normalize(cross(b - a, d - a));
            uint32_t i0 = s.add_vertex(a, n, rng.uni(), rng.uni(), f3(rng.range(0.5f, 1), rng.range(0.5f, 1), rng.range(0.5f, 1)));
            uint32_t i1 = s.add_vertex(b, n, rng.uni(), rng.uni());
            uint32_t i2 = s.add_vertex(d, n, rng.uni(), rng.uni());
            s.add_tri(i0, i1, i2);
        }
    }
    float pos[3] = {0.f, 0.f, 3.0f}, dir[3] = {0.05f, -0.02f, 1.f};
    rt2h_camera_new(&s.cam, pos, dir, 65.0f, 1.0f, 0.1f, 1000.0f);
}

// ---- C2/C4: Sponza-scale hall ------------------------------------------------------------------------------
static void gen_hall(rt2h_scene& s, float scale, uint32_t texture_size, uint64_t seed) {
    const uint32_t sd = (uint32_t)(0x5254325FULL + 2 + seed);
    const int NU = std::max(2, (int)std::lround(64 * scale)), NV = std::max(1, (int)std::lround(32 * scale));
    const uint32_t TS = texture_size ? texture_size : 512u;
    const float X0 = -20, X1 = 20, Y0 = 0, Y1 = 12, Z0 = -8, Z1 = 8;
    int mat_id = 0;
    auto next_material = [&](bool emissive) {
        Material m;
        for (int k = 0; k < 3; k++) m.base_color[k] = 0.75f + 0.2f * hash01((uint32_t)mat_id, (uint32_t)k, sd);
        m.texture = s.add_texture(TS, TS, make_texture(TS, (uint32_t)mat_id, sd), (mat_id % 8 == 7) ? RT2_FILTER_NEAREST : RT2_FILTER_LINEAR);
        if (emissive) { m.emission[0] = 0.60f; m.emission[1] = 0.55f; m.emission[2] = 0.50f; }
        mat_id++;
        return m;
    };
    // displaced rectangular patch: corner o, edges eu, ev, inward normal n
    auto rect_patch = [&](F3 o, F3 eu, F3 ev, F3 n, float amp, bool emissive) {
        Material m = next_material(emissive);
        s.begin_geometry(m, nullptr);
        uint32_t nseed = sd + 31u * (uint32_t)mat_id;
        s.add_patch(NU, NV, [&](float u, float v) {
            // lerp form: exact at u,v in {0,1} so neighbouring patches share border coordinates
            F3 p = f3(lerp(o.x, o.x + eu.x, u) + (lerp(o.x, o.x + ev.x, v) - o.x),
                      lerp(o.y, o.y + eu.y, u) + (lerp(o.y, o.y + ev.y, v) - o.y),
                      lerp(o.z, o.z + eu.z, u) + (lerp(o.z, o.z + ev.z, v) - o.z));
            float win = std::min(1.0f, 6.0f * std::min(std::min(u, 1 - u), std::min(v, 1 - v)));
            win = win * win * (3 - 2 * win);
            float d = amp * win * (value_noise(u * 9.f, v * 5.f, nseed) + 0.5f * value_noise(u * 23.f, v * 13.f, nseed + 1));
            return PatchPoint{p + n * d, n};
        }, 4.0f, 2.0f, true);
    };
    const float amp = 0.12f;
    // floor: 4 x 2 tiles, normal +y
    for (int i = 0; i < 4; i++) for (int j = 0; j < 2; j++)
        rect_patch(f3(X0 + 10.f * i, Y0, Z0 + 8.f * j), f3(10, 0, 0), f3(0, 0, 8), f3(0, 1, 0), amp, false);
    // ceiling: 4 x 2 tiles, normal -y; 4 emissive strips (flat)
    for (int i = 0; i < 4; i++) for (int j = 0; j < 2; j++) {
        bool em = ((i + j) & 1) == 0;
        rect_patch(f3(X0 + 10.f * i, Y1, Z0 + 8.f * j), f3(10, 0, 0), f3(0, 0, 8), f3(0, -1, 0), em ? 0.0f : amp, em);
    }
    // long walls z = Z0 (normal +z) and z = Z1 (normal -z): 3 x 2 patches each
    for (int i = 0; i < 3; i++) for (int j = 0; j < 2; j++) {
        float w = (X1 - X0) / 3.0f;
        float xa = (i == 0) ? X0 : X0 + w * i, xb = (i == 2) ? X1 : X0 + w * (i + 1);
        rect_patch(f3(xa, Y0 + 6.f * j, Z0), f3(xb - xa, 0, 0), f3(0, 6, 0), f3(0, 0, 1), amp, false);
        rect_patch(f3(xa, Y0 + 6.f * j, Z1), f3(xb - xa, 0, 0), f3(0, 6, 0), f3(0, 0, -1), amp, false);
    }
    // short walls x = X0 (normal +x), x = X1 (normal -x): 1 x 2 patches each
    for (int j = 0; j < 2; j++) {
        rect_patch(f3(X0, Y0 + 6.f * j, Z0), f3(0, 0, 16), f3(0, 6, 0), f3(1, 0, 0), amp, false);
        rect_patch(f3(X1, Y0 + 6.f * j, Z0), f3(0, 0, 16), f3(0, 6, 0), f3(-1, 0, 0), amp, false);
    }
    // 16 columns (2 rows x 8): fluted cylinders in local space, placed by the node transform
    for (int r = 0; r < 2; r++) for (int c = 0; c < 8; c++) {
        Material m = next_material(false);
        float mtx[16]; trs_matrix(f3(-17.5f + 5.f * c, 0.f, r ? 4.5f : -4.5f), 0.37f * (float)(c + 8 * r), 0.f, 1.0f, mtx);
        s.begin_geometry(m, mtx);
        s.add_patch(NU, NV, [&](float u, float v) {
            float a = u * 6.2831853f;
            float rad = 0.55f + 0.04f * std::cos(a * 12.f) + 0.1f * std::exp(-30.f * v * v) + 0.1f * std::exp(-30.f * (1 - v) * (1 - v));
            F3 n = f3(std::cos(a), 0, std::sin(a));
            return PatchPoint{f3(rad * n.x, 8.0f * v, rad * n.z), n};
        }, 4.0f, 2.0f, true);
    }
    // 14 arches (2 rows x 7): half tori spanning neighbouring columns
    for (int r = 0; r < 2; r++) for (int c = 0; c < 7; c++) {
        Material m = next_material(false);
        float mtx[16]; trs_matrix(f3(-15.0f + 5.f * c, 8.0f, r ? 4.5f : -4.5f), 0.f, 0.f, 1.0f, mtx);
        s.begin_geometry(m, mtx);
        s.add_patch(NU, NV, [&](float u, float v) {
            float a = u * 3.14159265f, b = v * 6.2831853f;       // a: along the arch, b: around the tube
            float R = 2.5f, rt = 0.45f;
            F3 c0 = f3(-R * std::cos(a), R * std::sin(a) * 0.8f, 0);
            F3 radial = normalize(f3(-std::cos(a), std::sin(a) * 0.8f, 0));
            F3 n = normalize(radial * std::cos(b) + f3(0, 0, 1) * std::sin(b));
            return PatchPoint{c0 + n * rt, n};
        }, 4.0f, 2.0f, true);
    }
    // 2 spheres
    for (int k = 0; k < 2; k++) {
        Material m = next_material(false);
        float mtx[16]; trs_matrix(f3(k ? 6.f : -4.f, 1.5f, k ? -1.0f : 1.5f), 0.7f * k, 0.2f * k, 1.5f, mtx);
        s.begin_geometry(m, mtx);
        s.add_patch(NU, NV, [&](float u, float v) {
            float a = u * 6.2831853f, b = (v - 0.5f) * 3.14159265f;
            F3 n = f3(std::cos(a) * std::cos(b), std::sin(b), std::sin(a) * std::cos(b));
            return PatchPoint{n, n};
        }, 4.0f, 2.0f, false);
    }
    float pos[3] = {-17.0f, 5.0f, 0.6f}, dir[3] = {-1.0f, 0.08f, 0.15f};
    rt2h_camera_new(&s.cam, pos, dir, 65.0f, 16.0f / 9.0f, 0.1f, 1000.0f);
}

// ---- C3: many lights -------------------------------------------------------------------------------------
static void gen_manylight(rt2h_scene& s, float scale, uint64_t seed) {
    const uint32_t sd = (uint32_t)(0x5254325FULL + 3 + seed);
    const int NU = std::max(2, (int)std::lround(99 * scale)), NV = std::max(1, (int)std::lround(50 * scale));
    const int LU = std::max(1, (int)std::lround(10 * scale)), LV = std::max(1, (int)std::lround(5 * scale));
    const float TILE = 20.0f;
    // 100 ground tiles ("city blocks": smooth terrain + box-like towers)
    for (int ty = 0; ty < 10; ty++) for (int tx = 0; tx < 10; tx++) {
        Material m; for (int k = 0; k < 3; k++) m.base_color[k] = 0.45f + 0.4f * hash01((uint32_t)tx, (uint32_t)ty, sd + (uint32_t)k);
        s.begin_geometry(m, nullptr);
        F3 o = f3(-100.f + TILE * tx, 0, -100.f + TILE * ty);
        uint32_t nseed = sd + 101u * (uint32_t)(ty * 10 + tx);
        s.add_patch(NU, NV, [&](float u, float v) {
            float x = lerp(o.x, o.x + TILE, u), z = lerp(o.z, o.z + TILE, v);
            float win = std::min(1.0f, 8.0f * std::min(std::min(u, 1 - u), std::min(v, 1 - v)));
            float cellx = std::floor(u * 5.f), cellz = std::floor(v * 5.f);
            float fu = u * 5.f - cellx, fv = v * 5.f - cellz;
            float inside = (fu > 0.25f && fu < 0.75f && fv > 0.25f && fv < 0.75f) ? 1.f : 0.f;
            float tower = inside * (1.0f + 5.0f * hash01((uint32_t)cellx, (uint32_t)cellz, nseed));
            float h = win * (0.6f * value_noise(x * 0.2f, z * 0.2f, sd) + tower);
            return PatchPoint{f3(x, h, z), f3(0, 1, 0)};
        }, 4.0f, 2.0f, true);
    }
    // 100 emissive panels above the tiles, facing down, identity transform (Q6)
    for (int ty = 0; ty < 10; ty++) for (int tx = 0; tx < 10; tx++) {
        Material m; m.base_color[0] = m.base_color[1] = m.base_color[2] = 0.8f;
        m.emission[0] = 0.5f + 1.5f * hash01((uint32_t)tx, (uint32_t)ty, sd + 7);
        m.emission[1] = 0.5f + 1.5f * hash01((uint32_t)tx, (uint32_t)ty, sd + 8);
        m.emission[2] = 0.5f + 1.5f * hash01((uint32_t)tx, (uint32_t)ty, sd + 9);
        s.begin_geometry(m, nullptr);
        F3 c = f3(-90.f + TILE * tx + 4.f * hash01((uint32_t)tx, (uint32_t)ty, sd + 3), 9.0f + 3.f * hash01((uint32_t)tx, (uint32_t)ty, sd + 4), -90.f + TILE * ty);
        s.add_patch(LU, LV, [&](float u, float v) {
            return PatchPoint{f3(c.x + 4.0f * (u - 0.5f), c.y, c.z + 2.0f * (v - 0.5f)), f3(0, -1, 0)};
        }, 1.0f, 1.0f, false);
    }
    float pos[3] = {-60.0f, 14.0f, -70.0f}, dir[3] = {-0.6f, 0.28f, -0.75f};
    rt2h_camera_new(&s.cam, pos, dir, 65.0f, 16.0f / 9.0f, 0.1f, 1000.0f);
}

// procedural equirect sky: gradient + sun (width x width / 2)
static void gen_sky(rt2h_scene& s, uint32_t width) {
    s.sky_w = width; s.sky_h = s.sky_w / 2;
    s.sky.resize((size_t)s.sky_w * s.sky_h * 4);
    F3 sun = normalize(f3(0.4f, 0.6f, -0.5f));
    for (uint32_t y = 0; y < s.sky_h; y++) for (uint32_t x = 0; x < s.sky_w; x++) {
        float u = (x + 0.5f) / s.sky_w, v = (y + 0.5f) / s.sky_h;
        float az = (u - 0.5f) * 6.2831853f, el = (0.5f - v) * 3.14159265f;
        F3 d = f3(std::cos(az) * std::cos(el), std::sin(el), std::sin(az) * std::cos(el));
        float t = 0.5f + 0.5f * d.y;
        float sd_ = std::max(0.0f, dot(d, sun));
        float disk = std::pow(sd_, 800.0f) * 40.0f + std::pow(sd_, 8.0f) * 0.5f;
        float* p = &s.sky[((size_t)y * s.sky_w + x) * 4];
        p[0] = lerp(0.6f, 0.25f, t) + disk; p[1] = lerp(0.6f, 0.45f, t) + disk * 0.9f; p[2] = lerp(0.55f, 0.9f, t) + disk * 0.7f; p[3] = 1.0f;
    }
}

// ---- C5: instanced ------------------------------------------------------------------------------------------
static void gen_instanced(rt2h_scene& s, float scale, uint64_t seed) {
    const uint32_t sd = (uint32_t)(0x5254325FULL + 5 + seed);
    Lcg64 rng(0x5254325FULL + 5 + seed);
    const int N = 16;            // 16x16 quads = 512 tris per geometry, 4 geometries = 2048 tris
    const float CELL = 4.0f;
    Material ground; ground.base_color[0] = 0.35f; ground.base_color[1] = 0.5f; ground.base_color[2] = 0.25f;
    s.begin_geometry(ground, nullptr);
    s.add_patch(N, N, [&](float u, float v) { return PatchPoint{f3((u - 0.5f) * CELL * 1.8f, 0.0f, (v - 0.5f) * CELL * 1.8f), f3(0, 1, 0)}; }, 1, 1, false);
    Material trunk; trunk.base_color[0] = 0.4f; trunk.base_color[1] = 0.28f; trunk.base_color[2] = 0.18f;
    s.begin_geometry(trunk, nullptr);
    s.add_patch(N, N, [&](float u, float v) { float a = u * 6.2831853f; float r = 0.18f * (1.0f - 0.5f * v); F3 n = f3(std::cos(a), 0, std::sin(a)); return PatchPoint{f3(r * n.x, 1.6f * v, r * n.z), n}; }, 1, 1, true);
    Material canopy; canopy.base_color[0] = 0.2f; canopy.base_color[1] = 0.55f; canopy.base_color[2] = 0.2f;
    float mc[16]; trs_matrix(f3(0, 2.0f, 0), 0.3f, 0.1f, 0.9f, mc);
    s.begin_geometry(canopy, mc);
    s.add_patch(N, N, [&](float u, float v) { float a = u * 6.2831853f, b = (v - 0.5f) * 3.14159265f; F3 n = f3(std::cos(a) * std::cos(b), std::sin(b), std::sin(a) * std::cos(b));
        float r = 1.0f + 0.15f * value_noise(u * 8.f, v * 8.f, sd); return PatchPoint{n * r, n}; }, 1, 1, true);
    Material rock; rock.base_color[0] = rock.base_color[1] = rock.base_color[2] = 0.55f;
    float mr[16]; trs_matrix(f3(1.1f, 0.2f, 0.7f), 1.0f, 0.0f, 0.35f, mr);
    s.begin_geometry(rock, mr);
    s.add_patch(N, N, [&](float u, float v) { float a = u * 6.2831853f, b = (v - 0.5f) * 3.14159265f; F3 n = f3(std::cos(a) * std::cos(b), std::sin(b), std::sin(a) * std::cos(b));
        float r = 1.0f + 0.3f * value_noise(u * 5.f, v * 5.f, sd + 9); return PatchPoint{f3(n.x * r, n.y * r * 0.7f, n.z * r), n}; }, 1, 1, true);
    const int G = std::max(2, (int)std::lround(100 * scale));
    for (int j = 0; j < G; j++) for (int i = 0; i < G; i++) {
        float m[16];
        trs_matrix(f3((i - G * 0.5f + 0.5f) * CELL + rng.range(-0.8f, 0.8f), 0.0f, (j - G * 0.5f + 0.5f) * CELL + rng.range(-0.8f, 0.8f)),
                   rng.range(0, 6.2831853f), 0.0f, rng.range(0.8f, 1.25f), m);
        float rows[12] = { m[0], m[4], m[8], m[12], m[1], m[5], m[9], m[13], m[2], m[6], m[10], m[14] };
        s.instances.insert(s.instances.end(), rows, rows + 12);
    }
    gen_sky(s, std::max(64u, (uint32_t)(2048 * std::min(1.0f, scale * 2))));
    float ext = G * CELL * 0.5f;
    float pos[3] = {-ext * 0.6f, 9.0f, -ext * 0.7f}, dir[3] = {-0.6f, 0.3f, -0.74f};
    rt2h_camera_new(&s.cam, pos, dir, 65.0f, 16.0f / 9.0f, 0.1f, 1000.0f);
}

}  // namespace

extern "C" {

void rt2h_camera_new(rt2h_camera* c, const float position[3], const float direction[3], float fov,
                     float aspect_ratio, float z_near, float z_far) {
    F3 d = normalize(f3(direction[0], direction[1], direction[2]));
    c->position[0] = position[0]; c->position[1] = position[1]; c->position[2] = position[2];
    c->direction[0] = d.x; c->direction[1] = d.y; c->direction[2] = d.z;
    c->fov = fov; c->aspect_ratio = aspect_ratio; c->z_near = z_near; c->z_far = z_far;
}

void rt2h_camera_view_matrix(const rt2h_camera* c, float m[16]) {
    // glam Mat4::look_at_rh(eye, eye + dir, UP) == look_to_rh(eye, dir, UP); UP = (0,-1,0) (camera.rs:14)
    F3 eye = f3(c->position[0], c->position[1], c->position[2]);
    F3 f = normalize(f3(c->direction[0], c->direction[1], c->direction[2]));
    F3 s = normalize(cross(f, f3(0, -1, 0)));
    F3 u = cross(s, f);
    m[0] = s.x; m[1] = u.x; m[2] = -f.x; m[3] = 0;
    m[4] = s.y; m[5] = u.y; m[6] = -f.y; m[7] = 0;
    m[8] = s.z; m[9] = u.z; m[10] = -f.z; m[11] = 0;
    m[12] = -dot(s, eye); m[13] = -dot(u, eye); m[14] = dot(f, eye); m[15] = 1;
}

void rt2h_perspective(float fovy, float aspect, float z_near, float z_far, float m[16]) {
    float y_scale = 1.0f / std::tan(0.5f * fovy);
    float x_scale = y_scale / aspect;
    float z_scale = 1.0f / (z_far - z_near);
    float v[16] = { x_scale, 0, 0, 0,   0, y_scale, 0, 0,   0, 0, -(z_near + z_far) * z_scale, 1.0f,   0, 0, -2.0f * z_near * z_far * z_scale, 0 };
    std::memcpy(m, v, sizeof v);
}

void rt2h_planar_view_constants(const rt2h_camera* c, uint32_t width, uint32_t height, rt2_planar_view_constants* o) {
    std::memset(o, 0, sizeof *o);
    float ws[2] = {(float)width, (float)height};
    float view[16], proj[16], iview[16], iproj[16];
    rt2h_camera_view_matrix(c, view);
    rt2h_perspective(c->fov * 0.017453292519943295f, c->aspect_ratio, c->z_near, c->z_far, proj);
    mat_inverse(view, iview); mat_inverse(proj, iproj);
    std::memcpy(o->mat_world_to_view, view, 64);
    std::memcpy(o->mat_view_to_clip, proj, 64);
    mat_mul(proj, view, o->mat_world_to_clip);
    std::memcpy(o->mat_clip_to_view, iproj, 64);
    std::memcpy(o->mat_view_to_world, iview, 64);
    mat_mul(iproj, iview, o->mat_clip_to_world);     // sic: camera.rs:123 multiplies in this order
    o->viewport_origin[0] = o->viewport_origin[1] = 0;
    o->viewport_size[0] = ws[0]; o->viewport_size[1] = ws[1];
    o->viewport_size_inv[0] = 1.0f / ws[0]; o->viewport_size_inv[1] = 1.0f / ws[1];
    o->pixel_offset[0] = o->pixel_offset[1] = 0;
    o->clip_to_window_scale[0] = 0.5f * ws[0]; o->clip_to_window_scale[1] = -0.5f * ws[1];
    o->clip_to_window_bias[0] = ws[0] * 0.5f; o->clip_to_window_bias[1] = ws[1] * 0.5f;
    o->window_to_clip_scale[0] = 1.0f / o->clip_to_window_scale[0]; o->window_to_clip_scale[1] = 1.0f / o->clip_to_window_scale[1];
    o->window_to_clip_bias[0] = -o->clip_to_window_bias[0] * o->window_to_clip_scale[0];
    o->window_to_clip_bias[1] = -o->clip_to_window_bias[1] * o->window_to_clip_scale[1];
    o->camera_direction_or_position[0] = c->position[0]; o->camera_direction_or_position[1] = c->position[1];
    o->camera_direction_or_position[2] = c->position[2]; o->camera_direction_or_position[3] = 1.0f;
}

uint32_t rt2h_fill_neighbor_offset_buffer(uint32_t count, uint8_t* out) {
    // light_passes.rs:671-698, f32 arithmetic, `as u8` saturating cast (negatives -> 0), Q11
    uint32_t n = 0;
    const int R = 250;
    const float phi2 = (float)(1.0 / 1.3247179572447);
    float u = 0.5f, v = 0.5f;
    while (n < count * 2) {
        u += phi2; v += phi2 * phi2;
        if (u >= 1.0f) u -= 1.0f;
        if (v >= 1.0f) v -= 1.0f;
        float r_sq = (u - 0.5f) * (u - 0.5f) + (v - 0.5f) * (v - 0.5f);
        if (r_sq > 0.25f) continue;
        auto as_u8 = [](float f) -> uint8_t { if (!(f > 0.0f)) return 0; if (f >= 255.0f) return 255; return (uint8_t)f; };
        out[n++] = as_u8((u - 0.5f) * (float)R);
        out[n++] = as_u8((v - 0.5f) * (float)R);
    }
    return n;
}

void rt2h_compute_pdf_texture_size(uint32_t max_items, uint32_t* w, uint32_t* h, uint32_t* mips) {
    double tw = std::max(1.0, std::ceil(std::sqrt((double)max_items)));      // light_passes.rs:700-716
    tw = std::exp2(std::ceil(std::log2(tw)));
    double th = std::max(1.0, std::ceil((double)max_items / tw));
    th = std::exp2(std::ceil(std::log2(th)));
    double tm = std::max(1.0, std::log2(std::max(tw, th)) + 1.0);
    *w = (uint32_t)tw; *h = (uint32_t)th; *mips = (uint32_t)tm;
}

void rt2h_calculate_reservoir_buffer_parameters(uint32_t w, uint32_t h, rt2_reservoir_buffer_params* o) {
    const uint32_t B = 16;                                                    // RTXDI_RESERVOIR_BLOCK_SIZE
    uint32_t wb = (w + B - 1) / B, hb = (h + B - 1) / B;                       // light_passes.rs:718-731
    o->reservoir_block_row_pitch = wb * (B * B);
    o->reservoir_array_pitch = o->reservoir_block_row_pitch * hb;
    o->pad1 = o->pad2 = 0;
}

void rt2h_default_gconst(rt2_gconst* g, const rt2h_camera* cam, uint32_t width, uint32_t height,
                         uint32_t num_lights, uint32_t env_w, uint32_t env_h) {
    std::memset(g, 0, sizeof *g);
    const uint32_t NEIGHBOR_OFFSET_COUNT = 8192;                               // main.rs:57
    rt2h_planar_view_constants(cam, width, height, &g->view);
    g->prev_view = g->view;
    rt2_reservoir_buffer_params rp; rt2h_calculate_reservoir_buffer_parameters(width, height, &rp);
    g->restir_gi.reservoir_buffer_params = rp;
    g->restir_gi.buffer_indices = {0, 1, 0, 0, 1, 0, 0, 0};
    g->restir_gi.final_shading_params = {0, 1, 0, 0};
    g->restir_gi.spatial_resampling_params = {0.1f, 0.3f, 1, 3.0f, 2, 0, 0, 0};
    g->restir_gi.temporal_resampling_params = {0.1f, 0.3f, 0, 20, 50, 0, 0.0f, 1, 2, 0x9E3779B9u, 0, 0};
    g->runtime_params = {NEIGHBOR_OFFSET_COUNT - 1, 0, 0, 0};
    g->environment_pdf_texture_size[0] = env_w; g->environment_pdf_texture_size[1] = env_h;
    uint32_t lw, lh, lm; rt2h_compute_pdf_texture_size(std::max(1u, num_lights), &lw, &lh, &lm);
    g->local_light_pdf_texture_size[0] = lw; g->local_light_pdf_texture_size[1] = lh;
    g->environment_light_risbuffer_segment_params = {1024 * 128, 1024, 128, 0};
    g->local_lights_risbuffer_segment_params = {0, 1024, 128, 0};
    g->restir_di.reservoir_buffer_params = rp;
    g->restir_di.buffer_indices = {0, 1, 0, 0, 1, 0, 0, 0};
    g->restir_di.initial_sampling_params = {0, 0, 0, 1, 0.0f, 0, 0, 0};
    g->restir_di.temporal_resampling_params = {0.1f, 0.3f, 5, 2, 0, 0.0f, 0, 0.0f, 1, 0x85EBCA6Bu, 0, 0};
    g->restir_di.spatial_resampling_params = {0.1f, 0.3f, 2, 3, 2, 32.0f, NEIGHBOR_OFFSET_COUNT - 1, 0};
    g->restir_di.shading_params = {0, 0, 10, 1000.0f, 0, 0, 0, 0};
    g->light_buffer_params.local_light_buffer_region = {0, num_lights, 0, 0};
    g->light_buffer_params.infinite_light_buffer_region = {num_lights, 0, 0, 0};
    g->light_buffer_params.environment_light_params = {1, num_lights + 1, 0, 0};
    g->enable_accumulation = 0; g->enable_brdf_additive_blend = 1; g->enable_brdf_indirect = 1;
    g->enable_restir_di = 0; g->enable_restir_gi = 1; g->frame = 0; g->refrence_mode = 0; g->textures = 1;
    g->blend_factor = 0.1f; g->enable_spatial_resampling = 0; g->enable_temporal_resampling = 0; g->environment = 0;
}

void rt2h_advance_frame(rt2_gconst* g, const rt2h_camera* cam, uint32_t width, uint32_t height, uint32_t frame) {
    g->prev_view = g->view;                                                    // main.rs:644
    rt2h_planar_view_constants(cam, width, height, &g->view);                  // main.rs:645
    g->frame = frame;                                                          // main.rs:646
    g->restir_di.temporal_resampling_params.uniform_random_number = pcg_hash(frame * 2u + 1u);   // main.rs:647-648 (rand::random)
    g->restir_gi.temporal_resampling_params.uniform_random_number = pcg_hash(frame * 2u + 2u);
    // main.rs:649-651 swaps the DI temporal in/out buffer indices each frame
    std::swap(g->restir_di.buffer_indices.temporal_resampling_input_buffer_index,
              g->restir_di.buffer_indices.temporal_resampling_output_buffer_index);
}

rt2h_scene* rt2h_scene_generate(int kind, float scale, uint32_t texture_size, uint64_t seed) {
    if (!(scale > 0.0f)) scale = 1.0f;
    rt2h_scene* s = new rt2h_scene();
    s->kind = kind;
    switch (kind) {
        case RT2H_SCENE_CORNELL: gen_cornell(*s); break;
        case RT2H_SCENE_HALL: gen_hall(*s, scale, texture_size, seed); break;
        case RT2H_SCENE_MANYLIGHT: gen_manylight(*s, scale, seed); break;
        case RT2H_SCENE_INSTANCED: gen_instanced(*s, scale, seed); break;
        case RT2H_SCENE_FURNACE: gen_furnace(*s); break;
        case RT2H_SCENE_RANDOM: gen_random(*s, scale, seed); break;
        default: delete s; return nullptr;
    }
    s->finish();
    return s;
}
static void set_err(char* err, size_t err_len, const std::string& m) {
    if (err && err_len) { std::snprintf(err, err_len, "%s", m.c_str()); }
}
rt2h_scene* rt2h_scene_load_gltf(const char* gltf_path, const char* sky_exr_path, char* err, size_t err_len) {
    if (!gltf_path) { set_err(err, err_len, "gltf_path is NULL"); return nullptr; }
    rt2ingest::GltfModel m;
    std::string e;
    if (!rt2ingest::load_gltf(gltf_path, m, e)) { set_err(err, err_len, e); return nullptr; }
    std::unique_ptr<rt2h_scene> s(new rt2h_scene());
    s->kind = 0;
    s->verts = std::move(m.vertices); s->idx = std::move(m.indices);
    s->geoms = std::move(m.geometries); s->index_counts = std::move(m.index_counts);
    s->image_data = std::move(m.image_rgba8);
    for (size_t i = 0; i < s->image_data.size(); i++) s->images.push_back({s->image_data[i].data(), m.image_w[i], m.image_h[i]});
    s->samplers = std::move(m.samplers); s->textures = std::move(m.textures);
    if (sky_exr_path && !rt2ingest::load_exr(sky_exr_path, s->sky, s->sky_w, s->sky_h, e)) { set_err(err, err_len, e); return nullptr; }
    const float pos[3] = {0.0f, 0.0f, 10.0f}, dir[3] = {0.0f, 0.0f, 1.0f};      // src/main.rs:114-121
    rt2h_camera_new(&s->cam, pos, dir, 65.0f, 1920.0f / 1080.0f, 0.1f, 1000.0f);
    s->finish();
    return s.release();
}
float* rt2h_load_exr(const char* path, uint32_t* width, uint32_t* height, char* err, size_t err_len) {
    std::vector<float> px; uint32_t w = 0, h = 0; std::string e;
    if (!path || !rt2ingest::load_exr(path, px, w, h, e)) { set_err(err, err_len, path ? e : "path is NULL"); return nullptr; }
    float* out = (float*)std::malloc(px.size() * sizeof(float));
    if (!out) { set_err(err, err_len, "out of memory"); return nullptr; }
    std::memcpy(out, px.data(), px.size() * sizeof(float));
    if (width) *width = w;
    if (height) *height = h;
    return out;
}
uint8_t* rt2h_decode_image(const uint8_t* data, size_t size, uint32_t* width, uint32_t* height, uint32_t* channels, char* err, size_t err_len) {
    std::vector<uint8_t> px; uint32_t w = 0, h = 0, ch = 0; std::string e;
    if (!data || size < 4) { set_err(err, err_len, "no image data"); return nullptr; }
    const bool jpeg = data[0] == 0xFF && data[1] == 0xD8;
    if (!(jpeg ? rt2ingest::decode_jpeg(data, size, px, w, h, ch, e) : rt2ingest::decode_png(data, size, px, w, h, ch, e))) { set_err(err, err_len, e); return nullptr; }
    uint8_t* out = (uint8_t*)std::malloc(px.size() ? px.size() : 1);
    if (!out) { set_err(err, err_len, "out of memory"); return nullptr; }
    std::memcpy(out, px.data(), px.size());
    if (width) *width = w;
    if (height) *height = h;
    if (channels) *channels = ch;
    return out;
}
void rt2h_free(void* p) { std::free(p); }
void rt2h_scene_destroy(rt2h_scene* s) { delete s; }
const rt2_scene_desc* rt2h_scene_desc(const rt2h_scene* s) { return s ? &s->desc : nullptr; }
void rt2h_scene_camera(const rt2h_scene* s, float aspect_ratio, rt2h_camera* out) { *out = s->cam; if (aspect_ratio > 0) out->aspect_ratio = aspect_ratio; }
void rt2h_scene_add_procedural_sky(rt2h_scene* s, uint32_t width) {
    if (!s) return;
    gen_sky(*s, std::max(8u, std::min(width, 16384u)) & ~1u);
    s->finish();
}

uint32_t rt2h_scene_num_lights(const rt2h_scene* s) {
    uint32_t lights = 0;                                                       // model.rs:411-413
    for (size_t g = 0; g < s->geoms.size(); g++) {
        const float* e = s->geoms[g].emission;
        if (e[0] != 0.0f || e[1] != 0.0f || e[2] != 0.0f) lights += s->index_counts[g] / 3;
    }
    return lights;
}
uint64_t rt2h_scene_num_triangles(const rt2h_scene* s) {
    uint64_t t = s->idx.size() / 3;
    uint64_t ni = s->instances.size() / 12;
    return ni ? t * ni : t;
}

}  // extern "C"
