// rt2_surface.cuh — device functions shared by every ray-generation-style kernel (K1, K3, K9, …):
// primary rays, LOD-0 texture sampling, environment lookup, the surface fetch GetGeometryFromHit and
// StoreShadingOutput.  Citations are relative to src/shaders/ of the reference tree.
#pragma once
#include "rt2_internal.cuh"

namespace rt2 {

// device pointers of the uploaded scene + acceleration structures
struct SceneRefs {
    const Node8* blas_nodes; const TriPos* tris;
    const Node8* tlas_nodes; const uint32_t* tlas_slot_to_inst;
    const InstanceRec* instances;            // non-null => instanced (two-level traversal)
    const Node8* flat_nodes; const FlatLeaf* flat_leaf;   // non-null => flattened instancing: one tree over (instance, triangle) pairs
    const TriShade* tri_shade; const GeomShade* geom_shade;
    const uint32_t* texels; const TextureRec* textures; const float* srgb_lut;
    const float4* sky; uint32_t sky_w, sky_h;
    const float* gamma_lut; const float* radiance_lut; const float* logluv_lut;   // CodecLuts (rt2_internal.cuh): tabulated decoders, may be null
    uint32_t any_hit;            // inlined visibility rays stop at their first hit (rt2_ctx::any_hit)
};
inline SceneRefs make_scene_refs(const rt2_ctx* ctx) {
    const DeviceScene& s = ctx->scene;
    SceneRefs r{};
    r.blas_nodes = s.blas.nodes; r.tris = s.blas.tris; r.tlas_nodes = s.tlas.nodes; r.tlas_slot_to_inst = s.tlas.slot_to_prim;
    r.instances = s.num_instances ? s.instances : nullptr;
    if (s.flat.nodes && ctx->use_flat) { r.flat_nodes = s.flat.nodes; r.flat_leaf = s.flat_leaf; }
    r.tri_shade = s.tri_shade; r.geom_shade = s.geom_shade;
    r.texels = s.texels; r.textures = s.textures; r.srgb_lut = s.srgb_lut; r.sky = s.sky; r.sky_w = s.sky_w; r.sky_h = s.sky_h;
    r.any_hit = ctx->any_hit ? 1u : 0u;
    r.gamma_lut = ctx->codec_luts; r.radiance_lut = ctx->codec_luts ? ctx->codec_luts + 256 : nullptr; r.logluv_lut = ctx->codec_luts ? ctx->codec_luts + 256 + 65536 : nullptr;
    return r;
}

// ------------------------------------------------------------------------------------------------
// shared device helpers
// ------------------------------------------------------------------------------------------------
// GBufferHelpers.glsl:12-27 setupPrimaryRay (pixel centre, no jitter — Q12)
__device__ __forceinline__ void setup_primary_ray(const rt2_planar_view_constants& view, uint32_t px, uint32_t py, V3* o, V3* d) {
    const float pcx = (float)px + 0.5f, pcy = (float)py + 0.5f;
    const float ux = pcx / view.viewport_size[0], uy = pcy / view.viewport_size[1];
    const float dx = ux * 2.0f - 1.0f, dy = uy * 2.0f - 1.0f;
    float target[4];
    mat4_mul_vec4(view.mat_clip_to_view, dx, dy, 1.0f, 1.0f, target);
    const V3 nt = normalize3(v3(target[0], target[1], target[2]));
    float dir[4];
    mat4_mul_vec4(view.mat_view_to_world, nt.x, nt.y, nt.z, 0.0f, dir);
    *o = v3(view.camera_direction_or_position[0], view.camera_direction_or_position[1], view.camera_direction_or_position[2]);
    *d = v3(dir[0], dir[1], dir[2]);
}

#define RT2_BACKGROUND_DEPTH 100000.0f   /* ShaderParameters.glsl:12 */
#define RT2_K1_TMIN 0.001f               /* refrence.rgen:27 */
#define RT2_DEFAULT_TRACE_VARIANT 16

// texture(textures[idx], uv).rgb at implicit LOD 0 — Hit.glsl:35-37; R8G8B8A8_SRGB, REPEAT; pinned
// filter identical to the oracle's sample_texture.
__device__ __forceinline__ int wrap_repeat(int i, int n) { int m = i % n; return m < 0 ? m + n : m; }
__device__ __forceinline__ V3 decode_srgb(uint32_t texel, const float* lut) {
    return v3(lut[texel & 0xFFu], lut[(texel >> 8) & 0xFFu], lut[(texel >> 16) & 0xFFu]);
}
__device__ __forceinline__ V3 sample_texture(const uint32_t* __restrict__ texels, const TextureRec& tx, const float* lut, float su, float sv) {
    const int W = (int)tx.width, H = (int)tx.height;
    const uint32_t* base = texels + tx.offset;
    if (tx.filter == RT2_FILTER_NEAREST) {
        const int x = wrap_repeat((int)floorf(su * (float)W), W), y = wrap_repeat((int)floorf(sv * (float)H), H);
        return decode_srgb(__ldg(base + (size_t)y * W + x), lut);
    }
    const float u = su * (float)W - 0.5f, v = sv * (float)H - 0.5f;
    const float fu = floorf(u), fv = floorf(v);
    const float a = u - fu, b = v - fv;
    const int x0 = wrap_repeat((int)fu, W), y0 = wrap_repeat((int)fv, H);
    const int x1 = x0 + 1 == W ? 0 : x0 + 1, y1 = y0 + 1 == H ? 0 : y0 + 1;
    const uint32_t r00 = __ldg(base + (size_t)y0 * W + x0), r10 = __ldg(base + (size_t)y0 * W + x1);
    const uint32_t r01 = __ldg(base + (size_t)y1 * W + x0), r11 = __ldg(base + (size_t)y1 * W + x1);
    const V3 t00 = decode_srgb(r00, lut), t10 = decode_srgb(r10, lut), t01 = decode_srgb(r01, lut), t11 = decode_srgb(r11, lut);
    const float ia = 1.0f - a, ib = 1.0f - b;
    const V3 top = t00 * ia + t10 * a;
    const V3 bot = t01 * ia + t11 * a;
    return top * ib + bot * b;
}

// GetEnvironmentRadiance — bridge :618-627; directionToEquirectUV Helpers.glsl:242-248 (PI = 3.1415926,
// common.glsl:31); sampler LINEAR / CLAMP_TO_BORDER black / normalised (src/main.rs:123-138).
__device__ __forceinline__ void direction_to_equirect_uv(V3 dir, float* u, float* v) {
    const float PI_ = 3.1415926f;
    // atan / asin follow the transcendental pin (binary64, rounded once), so the lookup is bit-identical to the oracle's
    *u = 0.5f + (float)atan2((double)dir.z, (double)dir.x) / (2.0f * PI_);
    *v = 0.5f - (float)asin((double)dir.y) / PI_;
}
// texture(SkyBox, uv).rgb == textureLod(SkyBox, uv, 0).rgb in a ray-generation shader
__device__ __forceinline__ V3 sample_sky(const SceneRefs& a, float u, float v) {
    const int W = (int)a.sky_w, H = (int)a.sky_h;
    const float x = u * (float)W - 0.5f, y = v * (float)H - 0.5f;
    const float fx = floorf(x), fy = floorf(y);
    const float wa = x - fx, wb = y - fy;
    const int x0 = (int)fx, y0 = (int)fy;
    auto texel = [&](int xx, int yy) -> V3 {
        if (xx < 0 || yy < 0 || xx >= W || yy >= H) return v3(0.f, 0.f, 0.f);
        const float4 p = __ldg(a.sky + (size_t)yy * W + xx);
        return v3(p.x, p.y, p.z);
    };
    const V3 t00 = texel(x0, y0), t10 = texel(x0 + 1, y0), t01 = texel(x0, y0 + 1), t11 = texel(x0 + 1, y0 + 1);
    const float ia = 1.0f - wa, ib = 1.0f - wb;
    const V3 top = t00 * ia + t10 * wa, bot = t01 * ia + t11 * wa;
    return top * ib + bot * wb;
}
__device__ __forceinline__ V3 environment_radiance(const SceneRefs& a, const rt2_gconst& g, V3 dir) {
    if (g.environment != 1u || a.sky == nullptr) return v3(0.f, 0.f, 0.f);
    float u, v; direction_to_equirect_uv(dir, &u, &v);
    return sample_sky(a, u, v);
}

// ShadingHelpers.glsl:61-88 StoreShadingOutput: the two rgba16f images follow the shader literally
// (prior read back from fp16, Q4 included); diffuse_f32 is an fp32 shadow of DiffuseLighting.
__device__ __forceinline__ void store_shading_output(const rt2_gconst& g, size_t px, V3 diffuse, V3 specular, bool is_first_pass,
                                                     ushort4* dimg, ushort4* simg, float4* d32) {
    V3 d_f32 = diffuse;
    if (g.enable_accumulation == 1u) {
        const ushort4 p = dimg[px];
        const V3 pd = v3(f16_bits_to_f32(p.x), f16_bits_to_f32(p.y), f16_bits_to_f32(p.z));
        const float bf = g.blend_factor;
        diffuse = v3(mixf(pd.x, diffuse.x, bf), mixf(pd.y, diffuse.y, bf), mixf(pd.z, diffuse.z, bf));
        specular = v3(mixf(pd.x, diffuse.x, bf), mixf(pd.y, diffuse.y, bf), mixf(pd.z, diffuse.z, bf));
        const float4 q = d32[px];
        d_f32 = v3(mixf(q.x, d_f32.x, bf), mixf(q.y, d_f32.y, bf), mixf(q.z, d_f32.z, bf));
    } else if (!is_first_pass) {
        const ushort4 p = dimg[px], s = simg[px];
        diffuse = diffuse + v3(f16_bits_to_f32(p.x), f16_bits_to_f32(p.y), f16_bits_to_f32(p.z));
        specular = specular + v3(f16_bits_to_f32(s.x), f16_bits_to_f32(s.y), f16_bits_to_f32(s.z));
        const float4 q = d32[px];
        d_f32 = d_f32 + v3(q.x, q.y, q.z);
    }
    dimg[px] = make_ushort4(f32_to_f16_bits(diffuse.x), f32_to_f16_bits(diffuse.y), f32_to_f16_bits(diffuse.z), 0);
    simg[px] = make_ushort4(f32_to_f16_bits(specular.x), f32_to_f16_bits(specular.y), f32_to_f16_bits(specular.z), 0);
    d32[px] = make_float4(d_f32.x, d_f32.y, d_f32.z, 0.0f);
}


// trace(ray): traceRayEXT(SceneBVH, opaque, 0xff, ...) — bridge :74-81, rayhit.rchit:60-66, raymiss.rmiss:12-17.
// A miss leaves geom = prim = inst = ~0u and *tri_slot = ~0u.
// any = true: the caller only asks "hit or not" (visibility tests) — the walk stops at the first hit (sc.any_hit; RT2_ANY_HIT=0 turns it off)
__device__ __forceinline__ Hit trace_closest(const SceneRefs& sc, V3 o, V3 d, float tmin, float tmax, uint32_t* tri_slot, bool any = false) {
    TraceRay ray; ray.o = o; ray.d = d; ray.tmin = tmin; ray.tmax = tmax;
    Hit best; best.t = INFINITY; best.u = best.v = 0.f; best.inst = best.geom = best.prim = 0xFFFFFFFFu;
    *tri_slot = 0xFFFFFFFFu;
    any = any && sc.any_hit != 0u;
    if (sc.flat_nodes) flat_closest_hit(sc.flat_nodes, sc.flat_leaf, sc.instances, sc.tris, ray, best, tri_slot, nullptr, any);
    else if (sc.instances) tlas_closest_hit(sc.tlas_nodes, sc.tlas_slot_to_inst, sc.instances, sc.blas_nodes, sc.tris, ray, best, tri_slot, nullptr, any);
    else cwbvh_closest_hit(sc.blas_nodes, sc.tris, ray, 0u, best, tri_slot, nullptr, any);
    return best;
}

// GetGeometryFromHit, Hit.glsl:2-42, over the slot-ordered shading records (same fp32 values the shader
// gathers through Indices[] / Vertices[]).  `lut` = the sRGB decode table (shared memory).
struct SurfaceFetch { V3 normal, specular_f0, color, emission; float roughness; };
__device__ __forceinline__ SurfaceFetch get_geometry_from_hit(const SceneRefs& sc, const float* lut, uint32_t tri_slot, uint32_t inst, float hu, float hv) {
    const float4* ts = reinterpret_cast<const float4*>(sc.tri_shade + tri_slot);
    const float4 q0 = __ldg(ts), q1 = __ldg(ts + 1), q2 = __ldg(ts + 2), q3 = __ldg(ts + 3), q4 = __ldg(ts + 4), q5 = __ldg(ts + 5);
    const uint32_t geom = __float_as_uint(__ldg(&sc.tris[tri_slot].v0.w));
    const GeomShade* gs = sc.geom_shade + geom;
    const float bx = 1.0f - hu - hv, by = hu, bz = hv;
    // n0 = q0.xyz, n1 = (q0.w, q1.x, q1.y), n2 = (q1.z, q1.w, q2.x)
    V3 n = normalize3(v3((q0.x * bx + q0.w * by) + q1.z * bz, (q0.y * bx + q1.x * by) + q1.w * bz, (q0.z * bx + q1.y * by) + q2.x * bz));
    float tn[4];
    mat4_mul_vec4(gs->transform, n.x, n.y, n.z, 0.0f, tn);              // normalize(transform * vec4(n, 0)) is a vec4 normalize (Hit.glsl:27)
    const float len = sqrtf(((tn[0] * tn[0] + tn[1] * tn[1]) + tn[2] * tn[2]) + tn[3] * tn[3]);
    n = v3(tn[0] / len, tn[1] / len, tn[2] / len);
    if (sc.instances) n = normalize3(xform_vector(sc.instances[inst].fwd, n));   // A2 extension
    // uv0 = (q2.y, q2.z), uv1 = (q2.w, q3.x), uv2 = (q3.y, q3.z)
    const float su = (q2.y * bx + q2.w * by) + q3.y * bz, sv = (q2.z * bx + q3.x * by) + q3.z * bz;
    // c0 = (q3.w, q4.x, q4.y), c1 = (q4.z, q4.w, q5.x), c2 = (q5.y, q5.z, q5.w)
    const V3 vc = v3((q3.w * bx + q4.z * by) + q5.y * bz, (q4.x * bx + q4.w * by) + q5.z * bz, (q4.y * bx + q5.x * by) + q5.w * bz);
    V3 color = v3(gs->base_color[0], gs->base_color[1], gs->base_color[2]) * vc;
    const int32_t tex = gs->texture;
    if (tex > -1) color = color * sample_texture(sc.texels, sc.textures[tex], lut, su, sv);
    SurfaceFetch f;
    f.normal = n; f.color = color;
    const float m = gs->metallic;                                        // mix(vec3(0), color, metallic) = 0*(1-a) + color*a
    f.specular_f0 = v3(0.0f * (1.0f - m) + color.x * m, 0.0f * (1.0f - m) + color.y * m, 0.0f * (1.0f - m) + color.z * m);
    f.roughness = 1.0f;                                                  // Hit.glsl:40 (Q10)
    f.emission = v3(gs->emission12[0], gs->emission12[1], gs->emission12[2]);
    return f;
}

// one atomicAdd per warp: ballot + popc compaction of the lanes that stay alive
__device__ __forceinline__ void warp_append(bool alive, uint32_t slot, uint32_t* queue_out, uint32_t* count_out) {
    __syncwarp();
    const uint32_t mask = __ballot_sync(0xFFFFFFFFu, alive);   // every lane of the warp calls this
    if (mask == 0u) return;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t leader = __ffs(mask) - 1u;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(count_out, __popc(mask));
    base = __shfl_sync(0xFFFFFFFFu, base, leader);
    if (alive) queue_out[base + __popc(mask & ((1u << lane) - 1u))] = slot;
}


}  // namespace rt2
