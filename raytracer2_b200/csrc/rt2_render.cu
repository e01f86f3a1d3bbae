// rt2_render.cu — wavefront implementation of the reference path tracer K1
// (src/shaders/lighting_passes/refrence.rgen:18-47) and the primary-visibility probe.
//
// One path slot per owned pixel.  A slot walks the reference's loops as a state machine
// (sample s, bounce i); the RNG stream is per pixel and sequential over samples exactly like the
// shader's, so results do not depend on scheduling.  Per wavefront iteration:
//   k_trace  : persistent warps pull 32 queued slots at a time, CWBVH closest hit -> st_hit
//   k_shade  : GetGeometryFromHit + emission/throughput update + cosine-lobe sample + path
//              regeneration; surviving slots are appended to the next queue with one atomicAdd per
//              warp (ballot + popc compaction)
// State is SoA float4 so a warp's 32 slots load as full 128-byte lines.
#include "rt2_internal.cuh"

namespace rt2 {

// ------------------------------------------------------------------------------------------------
// pixel ownership: interleaved row bands
// ------------------------------------------------------------------------------------------------
struct Band {
    uint32_t width, height;
    uint32_t band_rows, rank, world;   // world == 1 => whole frame
    uint32_t local_rows;
};
__host__ __device__ __forceinline__ uint32_t band_local_to_global(const Band& b, uint32_t ly) {
    if (b.world <= 1u) return ly;
    return ((ly / b.band_rows) * b.world + b.rank) * b.band_rows + (ly % b.band_rows);
}
static Band make_band(uint32_t w, uint32_t h, const rt2_launch_params& lp) {
    Band b{w, h, lp.band_rows, lp.band_rank, lp.band_world, h};
    if (b.world <= 1u || b.band_rows == 0u) { b.world = 1; b.rank = 0; b.band_rows = h ? h : 1; b.local_rows = h; return b; }
    uint32_t rows = 0;
    for (uint32_t y0 = b.rank * b.band_rows; y0 < h; y0 += b.world * b.band_rows) rows += std::min(b.band_rows, h - y0);
    b.local_rows = rows;
    return b;
}

struct TraceArgs {
    const Node8* blas_nodes; const TriPos* tris;
    const Node8* tlas_nodes; const uint32_t* tlas_slot_to_inst; const InstanceRec* instances;
    const float4* st_ray0; const float4* st_ray1;
    float4* st_hit; uint32_t* st_hit_inst;
    const uint32_t* queue; const uint32_t* count; uint32_t* head;
    unsigned long long* stat_rays;
};

struct ShadeArgs {
    Band band;
    const rt2_gconst* g;
    uint32_t frame, spp, max_bounces, flags;
    const TriPos* tris; const TriShade* tri_shade; const GeomShade* geom_shade;
    const InstanceRec* instances;   // non-null => instanced
    const uint32_t* texels; const TextureRec* textures; const float* srgb_lut;
    const float4* sky; uint32_t sky_w, sky_h;
    float4 *st_ray0, *st_ray1, *st_acc; uint4* st_misc;
    const float4* st_hit; const uint32_t* st_hit_inst;
    const uint32_t* queue_in; const uint32_t* count_in;
    uint32_t* queue_out; uint32_t* count_out;
    ushort4* diffuse_f16; ushort4* specular_f16; float4* diffuse_f32;
    unsigned long long* stat_hits;
};

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
#define RT2_DEFAULT_TRACE_VARIANT 5

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
__device__ __forceinline__ V3 environment_radiance(const ShadeArgs& a, V3 dir) {
    if (a.g->environment != 1u || a.sky == nullptr) return v3(0.f, 0.f, 0.f);
    const float PI_ = 3.1415926f;
    const float u = 0.5f + atan2f(dir.z, dir.x) / (2.0f * PI_);
    const float v = 0.5f - asinf(dir.y) / PI_;
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

// ------------------------------------------------------------------------------------------------
// kernels
// ------------------------------------------------------------------------------------------------
// refrence.rgen:21-27: rng = RAB_InitRandomSampler(pixel, 1); first primary ray.  Threads are laid out
// as 8x4 pixel tiles per warp so primary rays of a warp are coherent.
__global__ void __launch_bounds__(256) k_k1_init(ShadeArgs a, uint32_t padded) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= padded) return;
    const uint32_t tiles_x = (a.band.width + 7u) >> 3;
    const uint32_t tile = i >> 5, within = i & 31u;
    const uint32_t x = (tile % tiles_x) * 8u + (within & 7u), ly = (tile / tiles_x) * 4u + (within >> 3);
    const bool alive = x < a.band.width && ly < a.band.local_rows;
    uint32_t slot = 0;
    if (alive) {
        slot = ly * a.band.width + x;
        const uint32_t y = band_local_to_global(a.band, ly);
        V3 o, d;
        setup_primary_ray(a.g->view, x, y, &o, &d);
        a.st_ray0[slot] = make_float4(o.x, o.y, o.z, d.x);
        a.st_ray1[slot] = make_float4(d.y, d.z, 1.0f, 1.0f);
        a.st_acc[slot] = make_float4(1.0f, 0.0f, 0.0f, 0.0f);
        a.st_misc[slot] = make_uint4(1u /* rng.index */, 0u /* sample | bounce << 16 */, 0u, 0u);
    }
    warp_append(alive, slot, a.queue_out, a.count_out);
}

template <bool INSTANCED>
__global__ void __launch_bounds__(128) k_trace(TraceArgs a) {
    const uint32_t count = *a.count;
    const uint32_t lane = threadIdx.x & 31u;
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(a.stat_rays, (unsigned long long)count);
    for (;;) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(a.head, 32u);
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        if (base >= count) break;
        const uint32_t idx = base + lane;
        if (idx < count) {
            const uint32_t slot = a.queue[idx];
            const float4 r0 = a.st_ray0[slot], r1 = a.st_ray1[slot];
            TraceRay ray;
            ray.o = v3(r0.x, r0.y, r0.z); ray.d = v3(r0.w, r1.x, r1.y);
            ray.tmin = RT2_K1_TMIN; ray.tmax = RT2_BACKGROUND_DEPTH;
            Hit best; best.t = INFINITY; best.u = best.v = 0.f; best.inst = best.geom = best.prim = 0xFFFFFFFFu;
            uint32_t tri_slot = 0xFFFFFFFFu;
            if (INSTANCED) tlas_closest_hit(a.tlas_nodes, a.tlas_slot_to_inst, a.instances, a.blas_nodes, a.tris, ray, best, &tri_slot);
            else cwbvh_closest_hit(a.blas_nodes, a.tris, ray, 0u, best, &tri_slot);
            a.st_hit[slot] = make_float4(best.t, best.u, best.v, __uint_as_float(tri_slot));
            if (INSTANCED) a.st_hit_inst[slot] = best.inst;
        }
    }
}

// Single-level traversal with dynamic ray fetch and triangle postponing (after Aila & Laine 2009 and
// Ylitie et al. 2017): a warp keeps its lanes busy by refilling finished lanes from the queue once
// enough lane-iterations have been lost, and a lane whose triangle group would run with fewer than
// POSTPONE_PCT % of the warp's live lanes pushes it back and keeps walking nodes.  The closest hit is a
// total order (hit_closer), so the result does not depend on any of this scheduling.
template <int POSTPONE_PCT, int REFILL_MIN, int CHECK_EVERY>
__global__ void __launch_bounds__(128) k_trace_dyn(TraceArgs a) {
    const uint32_t count = *a.count;
    const uint32_t lane = threadIdx.x & 31u;
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(a.stat_rays, (unsigned long long)count);
    uint2 stack[RT2_STACK_SIZE];
    int sp = 0;
    bool active = false, exhausted = false;
    uint32_t slot = 0, tri_slot = 0xFFFFFFFFu;
    RayPrep rp = make_ray_prep(v3(0.f, 0.f, 0.f), v3(1.f, 1.f, 1.f));
    RayShear sh = make_shear(v3(1.f, 1.f, 1.f));
    Hit best; best.t = INFINITY; best.u = best.v = 0.f; best.inst = best.geom = best.prim = 0xFFFFFFFFu;
    uint2 ngroup = make_uint2(0u, 0u), tgroup = make_uint2(0u, 0u);
    for (;;) {
        const uint32_t need = __ballot_sync(0xFFFFFFFFu, !active);
        if (!exhausted && (need == 0xFFFFFFFFu || (REFILL_MIN > 0 && __popc(need) >= REFILL_MIN))) {
            const uint32_t n = __popc(need);
            uint32_t base = 0;
            if (lane == 0) base = atomicAdd(a.head, n);
            base = __shfl_sync(0xFFFFFFFFu, base, 0);
            if (!active) {
                const uint32_t idx = base + __popc(need & ((1u << lane) - 1u));
                if (idx < count) {
                    slot = a.queue[idx];
                    const float4 r0 = a.st_ray0[slot], r1 = a.st_ray1[slot];
                    const V3 o = v3(r0.x, r0.y, r0.z), d = v3(r0.w, r1.x, r1.y);
                    rp = make_ray_prep(o, d);
                    sh = make_shear(d);
                    best.t = INFINITY; best.u = best.v = 0.f; best.geom = best.prim = 0xFFFFFFFFu;
                    tri_slot = 0xFFFFFFFFu;
                    sp = 0; ngroup = make_uint2(0u, 0x80000000u); tgroup = make_uint2(0u, 0u);
                    active = true;
                }
            }
            if (base + n >= count) exhausted = true;
        }
        const uint32_t act = __ballot_sync(0xFFFFFFFFu, active);
        if (act == 0u) break;
        int budget = CHECK_EVERY;
        if (active) {
            for (;;) {
                if (ngroup.y > 0x00FFFFFFu) {
                    cwbvh_node_step(a.blas_nodes, rp, RT2_K1_TMIN, fminf(best.t, RT2_BACKGROUND_DEPTH), ngroup, tgroup, stack, sp);
                } else {
                    tgroup = ngroup;
                    ngroup = make_uint2(0u, 0u);
                }
                const int in_loop = __popc(__activemask());
                while (tgroup.y != 0u) {
                    if (POSTPONE_PCT > 0) {
                        if (__popc(__activemask()) * 100 < in_loop * POSTPONE_PCT && sp < RT2_STACK_SIZE) { stack[sp++] = tgroup; break; }
                    }
                    const uint32_t ti = rt2_fhs(tgroup.y);
                    tgroup.y &= ~(1u << ti);
                    const uint32_t s = tgroup.x + ti;
                    const TriPos* tp = a.tris + s;
                    const float4 va = ld16f(&tp->v0), vb = ld16f(&tp->v1), vc = ld16f(&tp->v2);
                    float t, u, v;
                    if (ray_triangle(v3(rp.ox, rp.oy, rp.oz), sh, RT2_K1_TMIN, RT2_BACKGROUND_DEPTH, v3(va.x, va.y, va.z), v3(vb.x, vb.y, vb.z),
                                     v3(vc.x, vc.y, vc.z), &t, &u, &v)) {
                        const uint32_t geom = f2u(va.w), prim = f2u(vb.w);
                        if (hit_closer(t, 0u, geom, prim, best)) {
                            best.t = t; best.u = u; best.v = v; best.inst = 0u; best.geom = geom; best.prim = prim;
                            tri_slot = s;
                        }
                    }
                }
                if (ngroup.y <= 0x00FFFFFFu) {
                    if (sp == 0) {
                        a.st_hit[slot] = make_float4(best.t, best.u, best.v, __uint_as_float(tri_slot));
                        active = false;
                        break;
                    }
                    ngroup = stack[--sp];
                }
                if (REFILL_MIN > 0 && !exhausted && --budget == 0) break;
            }
        }
        __syncwarp();
    }
}

// refrence.rgen:28-46 for one traced ray of each queued slot
__global__ void __launch_bounds__(256) k_k1_shade(ShadeArgs a) {
    __shared__ float lut[256];
    lut[threadIdx.x] = a.srgb_lut[threadIdx.x];
    __syncthreads();
    const uint32_t count = *a.count_in;
    const uint32_t stride = gridDim.x * blockDim.x;
    // all lanes of a warp iterate together (warp_append uses warp collectives)
    for (uint32_t i0 = (blockIdx.x * blockDim.x + threadIdx.x) & ~31u; i0 < count; i0 += stride) {
        const uint32_t i = i0 + (threadIdx.x & 31u);
        bool alive = false;
        uint32_t slot = 0;
        if (i < count) {
            slot = a.queue_in[i];
            const float4 r0 = a.st_ray0[slot], r1 = a.st_ray1[slot], ac = a.st_acc[slot];
            const uint4 misc = a.st_misc[slot];
            const float4 h = a.st_hit[slot];
            const uint32_t x = slot % a.band.width, ly = slot / a.band.width;
            const uint32_t y = band_local_to_global(a.band, ly);
            V3 o = v3(r0.x, r0.y, r0.z), d = v3(r0.w, r1.x, r1.y);
            V3 thr = v3(r1.z, r1.w, ac.x), rad = v3(ac.y, ac.z, ac.w);
            Rng rng; rng.index = misc.x; rng.seed = jenkins_hash(zcurve_to_linear(x, y)) + (a.frame + 13u);
            uint32_t sample = misc.y & 0xFFFFu, bounce = misc.y >> 16;
            const uint32_t tri_slot = __float_as_uint(h.w);
            bool next_sample;
            if (tri_slot == 0xFFFFFFFFu) {                         // refrence.rgen:32-36
                const V3 emission = environment_radiance(a, d);
                rad = rad + thr * emission;
                next_sample = true;
            } else {
                // GetGeometryFromHit, Hit.glsl:2-42
                const float4* ts = reinterpret_cast<const float4*>(a.tri_shade + tri_slot);
                const float4 q0 = __ldg(ts), q1 = __ldg(ts + 1), q2 = __ldg(ts + 2), q3 = __ldg(ts + 3), q4 = __ldg(ts + 4), q5 = __ldg(ts + 5);
                const uint32_t geom = __float_as_uint(__ldg(&a.tris[tri_slot].v0.w));
                const GeomShade* gs = a.geom_shade + geom;
                const float bx = 1.0f - h.y - h.z, by = h.y, bz = h.z;
                // n0 = q0.xyz, n1 = (q0.w, q1.x, q1.y), n2 = (q1.z, q1.w, q2.x)
                V3 n = normalize3(v3((q0.x * bx + q0.w * by) + q1.z * bz, (q0.y * bx + q1.x * by) + q1.w * bz, (q0.z * bx + q1.y * by) + q2.x * bz));
                float tn[4];
                mat4_mul_vec4(gs->transform, n.x, n.y, n.z, 0.0f, tn);
                const float len = sqrtf(((tn[0] * tn[0] + tn[1] * tn[1]) + tn[2] * tn[2]) + tn[3] * tn[3]);
                n = v3(tn[0] / len, tn[1] / len, tn[2] / len);
                if (a.instances) n = normalize3(xform_vector(a.instances[a.st_hit_inst[slot]].fwd, n));   // A2 extension
                // uv0 = (q2.y, q2.z), uv1 = (q2.w, q3.x), uv2 = (q3.y, q3.z)
                const float su = (q2.y * bx + q2.w * by) + q3.y * bz, sv = (q2.z * bx + q3.x * by) + q3.z * bz;
                // c0 = (q3.w, q4.x, q4.y), c1 = (q4.z, q4.w, q5.x), c2 = (q5.y, q5.z, q5.w)
                const V3 vc = v3((q3.w * bx + q4.z * by) + q5.y * bz, (q4.x * bx + q4.w * by) + q5.z * bz, (q4.y * bx + q5.x * by) + q5.w * bz);
                V3 color = v3(gs->base_color[0], gs->base_color[1], gs->base_color[2]) * vc;
                const int32_t tex = gs->texture;
                if (tex > -1) color = color * sample_texture(a.texels, a.textures[tex], lut, su, sv);
                const V3 emission = v3(gs->emission12[0], gs->emission12[1], gs->emission12[2]);
                rad = rad + thr * emission;                          // refrence.rgen:38
                thr = thr * color;                                   // refrence.rgen:39
                // RAB_GetSurfaceBrdfSample, bridge :437-461 (diffuseProbability == 1: cosine lobe only)
                const float rx = rng_next(rng), ry = rng_next(rng), rz = rng_next(rng);
                (void)rx;
                const V3 hvec = sample_cos_hemisphere(ry, rz, (a.flags & RT2_LAUNCH_TRIG_LIBDEVICE) != 0u);
                const V3 world_pos = o + d * h.x;                    // Hit.glsl:67
                d = tangent_to_world(n, hvec);
                o = world_pos;                                       // refrence.rgen:42
                bounce++;
                next_sample = bounce >= a.max_bounces;
            }
            alive = true;
            if (next_sample) {
                sample++; bounce = 0;
                if (sample >= a.spp) {
                    // refrence.rgen:46
                    const float inv = (float)a.spp;
                    store_shading_output(*a.g, (size_t)y * a.band.width + x, v3(rad.x / inv, rad.y / inv, rad.z / inv), v3(0.f, 0.f, 0.f), true,
                                         a.diffuse_f16, a.specular_f16, a.diffuse_f32);
                    alive = false;
                } else {
                    thr = v3(1.f, 1.f, 1.f);
                    setup_primary_ray(a.g->view, x, y, &o, &d);
                }
            }
            if (alive) {
                a.st_ray0[slot] = make_float4(o.x, o.y, o.z, d.x);
                a.st_ray1[slot] = make_float4(d.y, d.z, thr.x, thr.y);
                a.st_acc[slot] = make_float4(thr.z, rad.x, rad.y, rad.z);
                a.st_misc[slot] = make_uint4(rng.index, sample | (bounce << 16), 0u, 0u);
            }
        }
        warp_append(alive, slot, a.queue_out, a.count_out);
    }
}

// primary visibility probe: one thread per pixel, 8x4 tiles per warp
template <bool INSTANCED>
__global__ void __launch_bounds__(128) k_primary(TraceArgs a, Band band, const rt2_gconst* g, rt2_hit_record* out, uint32_t padded,
                                                 unsigned long long* stat_hits) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= padded) return;
    const uint32_t tiles_x = (band.width + 7u) >> 3;
    const uint32_t tile = i >> 5, within = i & 31u;
    const uint32_t x = (tile % tiles_x) * 8u + (within & 7u), ly = (tile / tiles_x) * 4u + (within >> 3);
    if (x >= band.width || ly >= band.local_rows) return;
    const uint32_t y = band_local_to_global(band, ly);
    TraceRay ray;
    setup_primary_ray(g->view, x, y, &ray.o, &ray.d);
    ray.tmin = RT2_K1_TMIN; ray.tmax = RT2_BACKGROUND_DEPTH;
    Hit best; best.t = INFINITY; best.u = best.v = 0.f; best.inst = best.geom = best.prim = 0xFFFFFFFFu;
    if (INSTANCED) tlas_closest_hit(a.tlas_nodes, a.tlas_slot_to_inst, a.instances, a.blas_nodes, a.tris, ray, best, nullptr);
    else cwbvh_closest_hit(a.blas_nodes, a.tris, ray, 0u, best, nullptr);
    rt2_hit_record r;
    r.instance = best.inst; r.geometry = best.geom; r.primitive = best.prim;
    const bool hit = best.geom != 0xFFFFFFFFu;
    r.t = hit ? best.t : 0.f; r.u = hit ? best.u : 0.f; r.v = hit ? best.v : 0.f; r._pad[0] = r._pad[1] = 0u;
    out[(size_t)y * band.width + x] = r;
    if (hit) atomicAdd(stat_hits, 1ull);
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
template <class T>
static cudaError_t dev_alloc(rt2_ctx* ctx, T** p, size_t n) {
    cudaError_t e = cudaMalloc((void**)p, std::max<size_t>(n, 1) * sizeof(T));
    if (e == cudaSuccess) { ctx->device_bytes += std::max<size_t>(n, 1) * sizeof(T); e = cudaMemsetAsync(*p, 0, std::max<size_t>(n, 1) * sizeof(T), ctx->stream); }
    return e;
}

void frame_release(rt2_ctx* ctx) {
    FrameBuffers& f = ctx->fb;
    void* ptrs[] = {f.diffuse_f16, f.specular_f16, f.diffuse_f32, f.hit_ids, f.output_rgba8, f.st_ray0, f.st_ray1, f.st_acc, f.st_misc,
                    f.st_hit, f.st_hit_inst, f.queue0, f.queue1, f.counters};
    for (void* p : ptrs) if (p) cudaFree(p);
    f = FrameBuffers();
}

int frame_resize(rt2_ctx* ctx, uint32_t w, uint32_t h) {
    if (w == 0 || h == 0 || (uint64_t)w * h > 0x10000000ull || w > 65535u || h > 65535u) return fail(ctx, RT2_E_INVALID, "bad frame size");
    RT2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    frame_release(ctx);
    FrameBuffers& f = ctx->fb;
    const size_t n = (size_t)w * h;
    f.width = w; f.height = h; f.max_slots = (uint32_t)n;
    RT2_CUDA(ctx, dev_alloc(ctx, &f.diffuse_f16, n));
    RT2_CUDA(ctx, dev_alloc(ctx, &f.specular_f16, n));
    RT2_CUDA(ctx, dev_alloc(ctx, &f.diffuse_f32, n));
    RT2_CUDA(ctx, dev_alloc(ctx, &f.hit_ids, n));
    RT2_CUDA(ctx, dev_alloc(ctx, &f.output_rgba8, n));
    RT2_CUDA(ctx, dev_alloc(ctx, &f.st_ray0, n));
    RT2_CUDA(ctx, dev_alloc(ctx, &f.st_ray1, n));
    RT2_CUDA(ctx, dev_alloc(ctx, &f.st_acc, n));
    RT2_CUDA(ctx, dev_alloc(ctx, &f.st_misc, n));
    RT2_CUDA(ctx, dev_alloc(ctx, &f.st_hit, n));
    RT2_CUDA(ctx, dev_alloc(ctx, &f.st_hit_inst, n));
    RT2_CUDA(ctx, dev_alloc(ctx, &f.queue0, n));
    RT2_CUDA(ctx, dev_alloc(ctx, &f.queue1, n));
    f.counter_capacity = 0;
    return RT2_OK;
}

static int check_ready(rt2_ctx* ctx) {
    if (!ctx->scene.accel_built) return fail(ctx, RT2_E_STATE, "no acceleration structure: call rt2_upload_scene + rt2_build_accel first");
    if (!ctx->fb.width) return fail(ctx, RT2_E_STATE, "no frame buffers: call rt2_resize first");
    if (!ctx->have_gconst) return fail(ctx, RT2_E_STATE, "no constants: call rt2_set_constants first");
    return RT2_OK;
}

static TraceArgs make_trace_args(rt2_ctx* ctx) {
    const DeviceScene& s = ctx->scene; const FrameBuffers& f = ctx->fb;
    TraceArgs t{};
    t.blas_nodes = s.blas.nodes; t.tris = s.blas.tris;
    t.tlas_nodes = s.tlas.nodes; t.tlas_slot_to_inst = s.tlas.slot_to_prim; t.instances = s.instances;
    t.st_ray0 = f.st_ray0; t.st_ray1 = f.st_ray1; t.st_hit = f.st_hit; t.st_hit_inst = f.st_hit_inst;
    return t;
}

// counters layout: [0..1] u64 rays, [2..3] u64 hits, then per iteration it: count at 4 + 2*it, head at 5 + 2*it
static int ensure_counters(rt2_ctx* ctx, uint32_t iterations) {
    FrameBuffers& f = ctx->fb;
    const size_t need = 4 + 2 * ((size_t)iterations + 2);
    if (f.counter_capacity < need) {
        if (f.counters) { RT2_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); cudaFree(f.counters); f.counters = nullptr; }
        RT2_CUDA(ctx, cudaMalloc((void**)&f.counters, need * 4));
        f.counter_capacity = need;
    }
    RT2_CUDA(ctx, cudaMemsetAsync(f.counters, 0, need * 4, ctx->stream));
    return RT2_OK;
}

// single-level trace variants (RT2_TRACE_VARIANT, read once per context: A/B runs on the GPU box)
typedef void (*trace_kernel_t)(TraceArgs);
static const trace_kernel_t g_trace_variants[] = {
    k_trace<false>,            // 0: static 32-ray batches
    k_trace_dyn<20, 0, 0>,     // 1: postponing 20 %, whole-warp refill only
    k_trace_dyn<10, 0, 0>,     // 2
    k_trace_dyn<30, 0, 0>,     // 3
    k_trace_dyn<20, 8, 16>,    // 4: + refill when >= 8 lanes idle, checked every 16 iterations
    k_trace_dyn<20, 4, 8>,     // 5
    k_trace_dyn<20, 2, 4>,     // 6
    k_trace_dyn<25, 4, 4>,     // 7
    k_trace_dyn<15, 6, 8>,     // 8
    k_trace_dyn<20, 4, 12>,    // 9
};
static const int g_num_trace_variants = (int)(sizeof(g_trace_variants) / sizeof(g_trace_variants[0]));
static void launch_trace_single(rt2_ctx* ctx, const TraceArgs& t) {
    g_trace_variants[ctx->trace_variant]<<<ctx->trace_blocks_per_sm[0] * ctx->num_sms, 128, 0, ctx->stream>>>(t);
}

int render_primary(rt2_ctx* ctx, const rt2_launch_params& lp) {
    int rc = check_ready(ctx); if (rc) return rc;
    FrameBuffers& f = ctx->fb;
    if ((rc = ensure_counters(ctx, 0))) return rc;
    const Band band = make_band(f.width, f.height, lp);
    const uint32_t tiles_x = (band.width + 7) / 8, tiles_y = (band.local_rows + 3) / 4;
    const uint32_t padded = tiles_x * tiles_y * 32;
    TraceArgs t = make_trace_args(ctx);
    unsigned long long* hits = reinterpret_cast<unsigned long long*>(f.counters + 2);
    if (padded) {
        if (ctx->scene.num_instances) k_primary<true><<<(padded + 127) / 128, 128, 0, ctx->stream>>>(t, band, ctx->d_gconst, f.hit_ids, padded, hits);
        else k_primary<false><<<(padded + 127) / 128, 128, 0, ctx->stream>>>(t, band, ctx->d_gconst, f.hit_ids, padded, hits);
    }
    RT2_CUDA(ctx, cudaGetLastError());
    ctx->stats.kernel_launches = padded ? 1 : 0;
    ctx->stats.wavefront_iterations = 1;
    ctx->pending_rays_static = (uint64_t)band.width * band.local_rows;
    ctx->stats_from_counters = true;
    return RT2_OK;
}

int render_reference(rt2_ctx* ctx, uint32_t frame, const rt2_launch_params& lp) {
    int rc = check_ready(ctx); if (rc) return rc;
    FrameBuffers& f = ctx->fb; const DeviceScene& s = ctx->scene;
    const uint32_t spp = lp.spp ? lp.spp : 12u;                    // refrence.rgen:17
    const uint32_t bounces = lp.max_bounces ? lp.max_bounces : 5u; // refrence.rgen:16
    if (spp > 65535u || bounces > 65535u || (uint64_t)spp * bounces > (1u << 22)) return fail(ctx, RT2_E_INVALID, "spp x bounces too large");
    const uint32_t iterations = spp * bounces;
    if ((rc = ensure_counters(ctx, iterations))) return rc;
    const Band band = make_band(f.width, f.height, lp);
    const uint32_t tiles_x = (band.width + 7) / 8, tiles_y = (band.local_rows + 3) / 4;
    const uint32_t padded = tiles_x * tiles_y * 32;
    uint64_t launches = 0;
    ShadeArgs a{};
    a.band = band; a.g = ctx->d_gconst; a.frame = frame; a.spp = spp; a.max_bounces = bounces; a.flags = lp.flags;
    a.tris = s.blas.tris; a.tri_shade = s.tri_shade; a.geom_shade = s.geom_shade;
    a.instances = s.num_instances ? s.instances : nullptr;
    a.texels = s.texels; a.textures = s.textures; a.srgb_lut = s.srgb_lut; a.sky = s.sky; a.sky_w = s.sky_w; a.sky_h = s.sky_h;
    a.st_ray0 = f.st_ray0; a.st_ray1 = f.st_ray1; a.st_acc = f.st_acc; a.st_misc = f.st_misc; a.st_hit = f.st_hit; a.st_hit_inst = f.st_hit_inst;
    a.diffuse_f16 = f.diffuse_f16; a.specular_f16 = f.specular_f16; a.diffuse_f32 = f.diffuse_f32;
    a.stat_hits = reinterpret_cast<unsigned long long*>(f.counters + 2);
    TraceArgs t = make_trace_args(ctx);
    t.stat_rays = reinterpret_cast<unsigned long long*>(f.counters);
    uint32_t* q[2] = {f.queue0, f.queue1};
    if (padded) {
        a.queue_out = q[0]; a.count_out = f.counters + 4;
        k_k1_init<<<(padded + 255) / 256, 256, 0, ctx->stream>>>(a, padded); launches++;
        const int trace_blocks = ctx->trace_blocks_per_sm[s.num_instances ? 1 : 0] * ctx->num_sms;
        const int shade_blocks = ctx->num_sms * 8;
        const bool profile = (lp.flags & RT2_LAUNCH_PROFILE) != 0u;
        for (cudaEvent_t e : ctx->prof_events) cudaEventDestroy(e);
        ctx->prof_events.clear();
        if (profile) { ctx->prof_events.resize((size_t)iterations * 3); for (auto& e : ctx->prof_events) RT2_CUDA(ctx, cudaEventCreate(&e)); }
        for (uint32_t it = 0; it < iterations; it++) {
            if (profile) RT2_CUDA(ctx, cudaEventRecord(ctx->prof_events[3 * it], ctx->stream));
            t.queue = q[it & 1]; t.count = f.counters + 4 + 2 * it; t.head = f.counters + 5 + 2 * it;
            if (s.num_instances) k_trace<true><<<trace_blocks, 128, 0, ctx->stream>>>(t);
            else launch_trace_single(ctx, t);
            if (profile) RT2_CUDA(ctx, cudaEventRecord(ctx->prof_events[3 * it + 1], ctx->stream));
            a.queue_in = q[it & 1]; a.count_in = t.count; a.queue_out = q[(it + 1) & 1]; a.count_out = f.counters + 4 + 2 * (it + 1);
            k_k1_shade<<<shade_blocks, 256, 0, ctx->stream>>>(a);
            if (profile) RT2_CUDA(ctx, cudaEventRecord(ctx->prof_events[3 * it + 2], ctx->stream));
            launches += 2;
        }
    }
    RT2_CUDA(ctx, cudaGetLastError());
    ctx->stats.kernel_launches = launches;
    ctx->stats.wavefront_iterations = iterations;
    ctx->stats_from_counters = true;
    ctx->pending_rays_static = 0;
    return RT2_OK;
}

int render_init(rt2_ctx* ctx) {
    int b0 = 0, b1 = 0;
    ctx->trace_variant = RT2_DEFAULT_TRACE_VARIANT;
    if (const char* e = getenv("RT2_TRACE_VARIANT")) { int v = atoi(e); if (v >= 0 && v < g_num_trace_variants) ctx->trace_variant = v; }
    RT2_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b0, g_trace_variants[ctx->trace_variant], 128, 0));
    RT2_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b1, k_trace<true>, 128, 0));
    ctx->trace_blocks_per_sm[0] = std::max(b0, 1); ctx->trace_blocks_per_sm[1] = std::max(b1, 1);
    return RT2_OK;
}

}  // namespace rt2
