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
#include "rt2_surface.cuh"

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
    const Node8* flat_nodes; const FlatLeaf* flat_leaf;      // flattened instancing (rt2_bvh.cuh), null when not built
    const float4* st_ray0; const float4* st_ray1;
    float4* st_hit; uint32_t* st_hit_inst;
    const uint32_t* queue; const uint32_t* count; uint32_t* head;
    uint32_t postpone_limit;     // triangle groups may be postponed onto the stack only below this depth: RT2_STACK_SIZE - tree depth
    unsigned long long* stat_rays;
};

struct ShadeArgs {
    Band band;
    const rt2_gconst* g;
    uint32_t frame, spp, max_bounces, flags;
    SceneRefs sc;
    float4 *st_ray0, *st_ray1, *st_acc; uint4* st_misc;
    const float4* st_hit; const uint32_t* st_hit_inst;
    const uint32_t* queue_in; const uint32_t* count_in;
    uint32_t* queue_out; uint32_t* count_out;
    ushort4* diffuse_f16; ushort4* specular_f16; float4* diffuse_f32;
    unsigned long long* stat_hits;
    // primary-hit reuse (see k1_shade_step): the pixel's primary hit record, kept for its later samples; stat_rays counts the
    // hit records the shader consumes (= its trace() calls), the trace kernels count the walks
    float4* st_hit0; uint32_t* st_hit_inst0; unsigned long long* stat_rays; uint32_t primary_reuse;
};

// ------------------------------------------------------------------------------------------------
// kernels
// ------------------------------------------------------------------------------------------------
// refrence.rgen:21-27: rng = RAB_InitRandomSampler(pixel, 1); first primary ray.  Threads are laid out
// as 8x4 pixel tiles per warp so primary rays of a warp are coherent.
__global__ void __launch_bounds__(256) k_k1_init(ShadeArgs a, uint32_t first, uint32_t padded) {
    const uint32_t i = first + blockIdx.x * blockDim.x + threadIdx.x;    // [first, padded): the tile range of one wavefront chain
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

template <bool INSTANCED, bool GENERIC = false>
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
            ray.o = v3(r0.x, r0.y, r0.z); ray.d = GENERIC ? v3(r1.x, r1.y, r1.z) : v3(r0.w, r1.x, r1.y);
            ray.tmin = GENERIC ? r0.w : RT2_K1_TMIN; ray.tmax = GENERIC ? r1.w : RT2_BACKGROUND_DEPTH;
            Hit best; best.t = INFINITY; best.u = best.v = 0.f; best.inst = best.geom = best.prim = 0xFFFFFFFFu;
            uint32_t tri_slot = 0xFFFFFFFFu;
            if (INSTANCED) tlas_closest_hit(a.tlas_nodes, a.tlas_slot_to_inst, a.instances, a.blas_nodes, a.tris, ray, best, &tri_slot);
            else cwbvh_closest_hit(a.blas_nodes, a.tris, ray, 0u, best, &tri_slot);
            a.st_hit[slot] = make_float4(best.t, best.u, best.v, __uint_as_float(tri_slot));
            if (INSTANCED) a.st_hit_inst[slot] = best.inst;
        }
    }
}

// Ray-queue traversal with dynamic ray fetch and triangle postponing (after Aila & Laine 2009 and Ylitie et al.
// 2017): a warp keeps its lanes busy by refilling finished lanes from the queue (checked every CHECK_EVERY
// iterations once REFILL_MIN lanes are idle), and a lane whose triangle group would run with fewer than
// POSTPONE_PCT % of the warp's live lanes pushes it back and keeps walking nodes.
// INSTANCED: the TLAS and the BLAS are walked by ONE loop — a lane is either at the top level (inst == ~0) or
// inside an instance; entering an instance pushes what is left of the TLAS groups, re-expresses the ray in object
// space (o' = inv*o, d' = inv*d, t unchanged) and starts at the BLAS root; the instance is left when the stack
// is back at the depth it was entered with.  Lanes at different levels share every node step.
// GENERIC: rays carry their own [tmin, tmax] (ray0 = o|tmin, ray1 = d|tmax) instead of K1's constants.
// The closest hit is a total order (hit_closer), so the result does not depend on any of this scheduling.
__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
// leaf triangles are touched once per ray: loaded without allocating an L1 line so that the L1 keeps BVH nodes (RT2_TRI_NOALLOC)
__device__ __forceinline__ float4 ld16f_stream(const void* p) {
#ifdef RT2_TRI_NOALLOC
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
#else
    return ld16f(p);
#endif
}
#ifdef RT2_RAY_STREAM
#define RT2_LDRAY(p) __ldcs(p)
#define RT2_STHIT(p, v) __stcs((p), (v))
#else
#define RT2_LDRAY(p) (*(p))
#define RT2_STHIT(p, v) (*(p) = (v))
#endif

// LOCALQ: the queue belongs to this warp alone (k_path_warp): `queue`/`count` are passed in and the head is a warp-uniform
// register instead of a global atomic.
// FLAT (with INSTANCED = false): one tree over every (instance, triangle) pair walked with the world-space ray; the lane re-expresses
// its ray in an instance's object space only when a leaf of another instance comes up (rt2_bvh.cuh, flat_closest_hit).
// ANYHIT (GENERIC visibility rays): the caller reads only "hit or not" (p.geometryIndex == ~0u after trace(), bridge :218-231,
// ShadingHelpers.glsl:35-37), so the first triangle hit inside [tmin, tmax] ends the walk.  The record then names some hit, not the closest.
template <int POSTPONE_PCT, int REFILL_MIN, int CHECK_EVERY, bool INSTANCED, bool GENERIC, bool LOCALQ = false, bool FLAT = false, bool ANYHIT = false>
__device__ __forceinline__ void trace_dyn_body(const TraceArgs& a, const uint32_t* lq = nullptr, uint32_t lcount = 0u) {
    const uint32_t count = LOCALQ ? lcount : *a.count;
    const uint32_t* queue = LOCALQ ? lq : a.queue;      // no __restrict__: a warp-local queue was written earlier in this kernel
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t lhead = 0u;
    if (!LOCALQ && blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(a.stat_rays, (unsigned long long)count);
    uint2 stack[RT2_STACK_SIZE];
    int sp = 0, inst_sp = 0;
    bool active = false, exhausted = false;
    uint32_t slot = 0, tri_slot = 0xFFFFFFFFu, inst = 0xFFFFFFFFu;
    float tmin = RT2_K1_TMIN, tmax = RT2_BACKGROUND_DEPTH;
    RayPrep rp = make_ray_prep(v3(0.f, 0.f, 0.f), v3(1.f, 1.f, 1.f));
    RayShear sh = make_shear(v3(1.f, 1.f, 1.f));
    Hit best; best.t = INFINITY; best.u = best.v = 0.f; best.inst = best.geom = best.prim = 0xFFFFFFFFu;
    uint2 ngroup = make_uint2(0u, 0u), tgroup = make_uint2(0u, 0u);
    const int postpone_limit = (int)a.postpone_limit;
    // INSTANCED: a ray enters ~14 instances on C5 and tests a triangle in fewer than half of them, so the shear constants
    // (two divisions) are computed at the first triangle test inside an instance (sh_pending, object-space direction kept in
    // dobj), and the world-space reciprocal direction is kept (world_id) so that leaving an instance costs one 16-byte load and
    // six multiplies instead of three divisions.  C5: 577 -> 604 Mrays/s with the whole RayPrep kept (96 registers).
    V3 world_id = v3(1.f, 1.f, 1.f);
    V3 dobj = v3(1.f, 1.f, 1.f);
    bool sh_pending = false;
    V3 oobj = v3(0.f, 0.f, 0.f);              // FLAT: origin in the object space of instance `inst`
#ifdef RT2_DEBUG_STEP_CAP
    int dbg_steps = 0;
#endif
    for (;;) {
        const uint32_t need = __ballot_sync(0xFFFFFFFFu, !active);
        if (!exhausted && (need == 0xFFFFFFFFu || (REFILL_MIN > 0 && __popc(need) >= REFILL_MIN))) {
            const uint32_t n = __popc(need);
            uint32_t base = 0;
            if (LOCALQ) { base = lhead; lhead += n; }
            else {
                if (lane == 0) base = atomicAdd(a.head, n);
                base = __shfl_sync(0xFFFFFFFFu, base, 0);
            }
            if (!active) {
                const uint32_t idx = base + __popc(need & ((1u << lane) - 1u));
                if (idx < count) {
                    slot = queue[idx];
                    const float4 r0 = RT2_LDRAY(a.st_ray0 + slot), r1 = RT2_LDRAY(a.st_ray1 + slot);
                    const V3 o = v3(r0.x, r0.y, r0.z), d = GENERIC ? v3(r1.x, r1.y, r1.z) : v3(r0.w, r1.x, r1.y);
                    if (GENERIC) { tmin = r0.w; tmax = r1.w; }
                    if (INSTANCED) { world_id = ray_inv_dir(d); rp = ray_prep_from_inv(o, world_id); }
                    else rp = make_ray_prep(o, d);
                    if (!INSTANCED && !FLAT) sh = make_shear(d);
                    best.t = INFINITY; best.u = best.v = 0.f; best.inst = best.geom = best.prim = 0xFFFFFFFFu;
                    tri_slot = 0xFFFFFFFFu; inst = 0xFFFFFFFFu;
                    sp = 0; inst_sp = 0; ngroup = make_uint2(0u, 0x80000000u); tgroup = make_uint2(0u, 0u);
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
                    const Node8* level = FLAT ? a.flat_nodes : ((INSTANCED && inst == 0xFFFFFFFFu) ? a.tlas_nodes : a.blas_nodes);
                    cwbvh_node_step(level, rp, tmin, fminf(best.t, tmax), ngroup, tgroup, stack, sp);
#ifdef RT2_TRACE_PREFETCH_NODE   // measured -3.4 % on C2 (the address arithmetic costs more than the latency it hides): off
                    // the node this lane visits next is known now; its 80 bytes are wanted only after the triangle phase
                    if (ngroup.y > 0x00FFFFFFu) {
                        const uint32_t nslot = (rt2_fhs(ngroup.y) - 24u) ^ rp.oct_inv;
                        const uint32_t nrel = rt2_popc((ngroup.y & 0xFFu) & ~(0xFFFFFFFFu << nslot));
                        prefetch_l1(level + (ngroup.x + nrel));
                    }
#endif
                } else {
                    tgroup = ngroup;
                    ngroup = make_uint2(0u, 0u);
                }
                if (INSTANCED && inst == 0xFFFFFFFFu) {
                    if (tgroup.y != 0u) {                              // TLAS leaves = instance slots: enter one, keep the rest for later
                        const uint32_t ti = rt2_fhs(tgroup.y);
                        tgroup.y &= ~(1u << ti);
                        // each TLAS level can leave two entries (instance remnant + inner remainder): the launch gate
                        // (unified_walk_fits) sizes for that, the bounds checks keep a malformed tree from corrupting local memory
                        if (tgroup.y != 0u && sp < RT2_STACK_SIZE) stack[sp++] = tgroup;
                        if (ngroup.y > 0x00FFFFFFu && sp < RT2_STACK_SIZE) stack[sp++] = ngroup;
                        inst = __ldg(a.tlas_slot_to_inst + tgroup.x + ti);
                        const InstanceRec* ir = a.instances + inst;
                        float m[12];
#pragma unroll
                        for (int k = 0; k < 3; k++) { const float4 r = ld16f(ir->inv + 4 * k); m[4 * k] = r.x; m[4 * k + 1] = r.y; m[4 * k + 2] = r.z; m[4 * k + 3] = r.w; }
                        const float4 r0 = a.st_ray0[slot], r1 = a.st_ray1[slot];
                        const V3 o = xform_point(m, v3(r0.x, r0.y, r0.z)), d = xform_vector(m, GENERIC ? v3(r1.x, r1.y, r1.z) : v3(r0.w, r1.x, r1.y));
                        rp = make_ray_prep(o, d);
                        dobj = d; sh_pending = true;
                        inst_sp = sp;
                        ngroup = make_uint2(0u, 0x80000000u); tgroup = make_uint2(0u, 0u);
                        if (REFILL_MIN > 0 && !exhausted && --budget == 0) break;
                        continue;
                    }
                } else {
                    const int in_loop = __popc(__activemask());
                    while (tgroup.y != 0u) {
                        if (POSTPONE_PCT > 0) {
                            if (__popc(__activemask()) * 100 < in_loop * POSTPONE_PCT && sp < postpone_limit) {
#ifndef RT2_TRI_NOALLOC
                                if (FLAT) prefetch_l1(a.flat_leaf + (tgroup.x + rt2_fhs(tgroup.y)));
                                else prefetch_l1(a.tris + (tgroup.x + rt2_fhs(tgroup.y)));
#endif      // postponed: by the time it is tested it sits in L1 (+0.5-1 %)
                                stack[sp++] = tgroup; break;
                            }
                        }
                        const uint32_t ti = rt2_fhs(tgroup.y);
                        tgroup.y &= ~(1u << ti);
                        uint32_t s = tgroup.x + ti;
                        if (FLAT) {
                            const uint2 lf = __ldg(reinterpret_cast<const uint2*>(a.flat_leaf) + s);
                            s = lf.y;
                            if (lf.x != inst) {            // another instance: object-space origin + shear constants for it
                                inst = lf.x;
                                const InstanceRec* ir = a.instances + inst;
                                float m[12];
#pragma unroll
                                for (int k = 0; k < 3; k++) { const float4 r = ld16f(ir->inv + 4 * k); m[4 * k] = r.x; m[4 * k + 1] = r.y; m[4 * k + 2] = r.z; m[4 * k + 3] = r.w; }
                                const float4 r0 = a.st_ray0[slot], r1 = a.st_ray1[slot];
                                oobj = xform_point(m, v3(r0.x, r0.y, r0.z));
                                sh = make_shear(xform_vector(m, GENERIC ? v3(r1.x, r1.y, r1.z) : v3(r0.w, r1.x, r1.y)));
                            }
                        }
                        const TriPos* tp = a.tris + s;
                        const float4 va = ld16f_stream(&tp->v0), vb = ld16f_stream(&tp->v1), vc = ld16f_stream(&tp->v2);
                        if (INSTANCED && sh_pending) { sh = make_shear(dobj); sh_pending = false; }
                        float t, u, v;
                        if (ray_triangle(FLAT ? oobj : v3(rp.ox, rp.oy, rp.oz), sh, tmin, tmax, v3(va.x, va.y, va.z), v3(vb.x, vb.y, vb.z),
                                         v3(vc.x, vc.y, vc.z), &t, &u, &v)) {
                            const uint32_t geom = f2u(va.w), prim = f2u(vb.w);
                            const uint32_t hi = (INSTANCED || FLAT) ? inst : 0u;
                            if (hit_closer(t, hi, geom, prim, best)) {
                                best.t = t; best.u = u; best.v = v; best.inst = hi; best.geom = geom; best.prim = prim;
                                tri_slot = s;
                                if (ANYHIT) { tgroup.y = 0u; ngroup = make_uint2(0u, 0u); sp = 0; inst_sp = 0; }      // nothing left to visit: the ray retires below
                            }
                        }
                    }
                }
#ifdef RT2_DEBUG_STEP_CAP   // measurement only (wrong images): what would a launch cost if no ray walked more than this many rounds?
                if (++dbg_steps > RT2_DEBUG_STEP_CAP) { sp = 0; inst_sp = 0; ngroup.y = 0u; tgroup.y = 0u; dbg_steps = 0; }
#endif
                if (ngroup.y <= 0x00FFFFFFu) {
                    if (INSTANCED && inst != 0xFFFFFFFFu && sp == inst_sp) {    // this instance is done: back to the top level, world-space ray
                        inst = 0xFFFFFFFFu;
                        const float4 r0 = a.st_ray0[slot];
                        rp = ray_prep_from_inv(v3(r0.x, r0.y, r0.z), world_id);
                    }
                    if (sp == 0) {
                        RT2_STHIT(a.st_hit + slot, make_float4(best.t, best.u, best.v, __uint_as_float(tri_slot)));
                        if (INSTANCED || FLAT) a.st_hit_inst[slot] = best.inst;
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

template <int POSTPONE_PCT, int REFILL_MIN, int CHECK_EVERY, bool INSTANCED = false, bool GENERIC = false>
__global__ void __launch_bounds__(128) k_trace_dyn(TraceArgs a) { trace_dyn_body<POSTPONE_PCT, REFILL_MIN, CHECK_EVERY, INSTANCED, GENERIC>(a); }
// The two-level walk at 6 blocks/SM (80 registers; ptxas picks 96 = 5 blocks on its own): C5 619 -> 659 Mrays/s.  A separate
// entry point because __launch_bounds__(128, n) on the single-level kernels changes their code (C2 -0.4 % at n = 7).
template <int POSTPONE_PCT, int REFILL_MIN, int CHECK_EVERY, bool GENERIC = false, bool ANYHIT = false>
__global__ void __launch_bounds__(128, 6) k_trace_dyn_tlas(TraceArgs a) { trace_dyn_body<POSTPONE_PCT, REFILL_MIN, CHECK_EVERY, true, GENERIC, false, false, ANYHIT>(a); }
// flattened instancing: single-level walk over the (instance, triangle) tree
template <int POSTPONE_PCT, int REFILL_MIN, int CHECK_EVERY, bool GENERIC = false, bool ANYHIT = false>
#ifndef RT2_FLAT_POSTPONE
#define RT2_FLAT_POSTPONE 30   /* 20 / 25 / 30: 1143 / 1152 (with refill 6, check 6) / 1154 Mrays/s on C5 */
#endif
#ifndef RT2_FLAT_REFILL
#define RT2_FLAT_REFILL 4
#endif
#ifndef RT2_FLAT_CHECK
#define RT2_FLAT_CHECK 8
#endif
#define RT2_FLAT_PARAMS RT2_FLAT_POSTPONE, RT2_FLAT_REFILL, RT2_FLAT_CHECK     /* POSTPONE_PCT, REFILL_MIN, CHECK_EVERY of the flat walk */
#ifndef RT2_FLAT_MIN_BLOCKS
#define RT2_FLAT_MIN_BLOCKS 7   /* 72 registers, 28 warps per SM: the flat walk is latency-bound (L2 hit 72-80 %, DRAM 2 %); 5 / 6 / 7 / 8 blocks: 1021 / 1086 / 1144 / 1096 Mrays/s on C5 */
#endif
__global__ void __launch_bounds__(128, RT2_FLAT_MIN_BLOCKS) k_trace_dyn_flat(TraceArgs a) { trace_dyn_body<POSTPONE_PCT, REFILL_MIN, CHECK_EVERY, false, GENERIC, false, true, ANYHIT>(a); }

// The ray-queue kernel of the ReSTIR passes (GENERIC rays: origin | tmin, direction | tmax per slot) on single-level scenes: an entry
// point of its own so that its register cap and refill cadence can differ from K1's (80 registers / 36 % occupancy without the cap)
#ifndef RT2_GENERIC_POSTPONE
#define RT2_GENERIC_POSTPONE 25
#endif
#ifndef RT2_GENERIC_REFILL
#define RT2_GENERIC_REFILL 6
#endif
#ifndef RT2_GENERIC_CHECK
#define RT2_GENERIC_CHECK 6
#endif
#ifndef RT2_GENERIC_MIN_BLOCKS
#define RT2_GENERIC_MIN_BLOCKS 7
#endif
template <bool ANYHIT>
__global__ void __launch_bounds__(128, RT2_GENERIC_MIN_BLOCKS) k_trace_queue(TraceArgs a) {
    trace_dyn_body<RT2_GENERIC_POSTPONE, RT2_GENERIC_REFILL, RT2_GENERIC_CHECK, false, true, false, false, ANYHIT>(a);
}

// refrence.rgen:28-46 for the path in `slot` whose ray just came back (hit record in st_hit): emission / throughput update,
// cosine-lobe sample, path regeneration.  Returns true while the pixel still has rays to trace (state written back), false when
// its last sample finished (radiance stored).
// Primary-hit reuse: setupPrimaryRay has no jitter (GBufferHelpers.glsl:12-27: the pixel centre) and refrence.rgen:26 calls it anew
// for every sample, so the spp samples of a pixel issue the SAME first trace().  Its hit record is kept (st_hit0) when it comes back
// and the later samples consume it right here instead of walking the tree again: a sample costs max_bounces - 1 walks after the
// first, a sky pixel one walk altogether.  Same hit record => same image, bit for bit; `rays` counts every record consumed (the
// shader's trace() calls), the trace kernels count the walks (rt2_stats.rays_walked).
__device__ __forceinline__ bool k1_shade_step(const ShadeArgs& a, const float* lut, uint32_t slot, uint32_t* hits, uint32_t* rays) {
    const float4 r0 = a.st_ray0[slot], r1 = a.st_ray1[slot], ac = a.st_acc[slot];
    const uint4 misc = a.st_misc[slot];
    float4 h = a.st_hit[slot];
    uint32_t h_inst = a.sc.instances ? a.st_hit_inst[slot] : 0u;
    const uint32_t x = slot % a.band.width, ly = slot / a.band.width;
    const uint32_t y = band_local_to_global(a.band, ly);
    V3 o = v3(r0.x, r0.y, r0.z), d = v3(r0.w, r1.x, r1.y);
    V3 thr = v3(r1.z, r1.w, ac.x), rad = v3(ac.y, ac.z, ac.w);
    Rng rng; rng.index = misc.x; rng.seed = jenkins_hash(zcurve_to_linear(x, y)) + (a.frame + 13u);
    uint32_t sample = misc.y & 0xFFFFu, bounce = misc.y >> 16;
    if (a.primary_reuse && bounce == 0u) {                    // the one primary ray of this pixel that is walked
        a.st_hit0[slot] = h;
        if (a.sc.instances) a.st_hit_inst0[slot] = h_inst;
    }
    for (;;) {
        (*rays)++;
        const uint32_t tri_slot = __float_as_uint(h.w);
        bool next_sample;
        if (tri_slot == 0xFFFFFFFFu) {                         // refrence.rgen:32-36
            const V3 emission = environment_radiance(a.sc, *a.g, d);
            rad = rad + thr * emission;
            next_sample = true;
        } else {
            (*hits)++;
            const SurfaceFetch sf = get_geometry_from_hit(a.sc, lut, tri_slot, h_inst, h.y, h.z);
            const V3 n = sf.normal, color = sf.color;
            rad = rad + thr * sf.emission;                          // refrence.rgen:38
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
        if (!next_sample) break;
        sample++; bounce = 0;
        if (sample >= a.spp) {
            // refrence.rgen:46
            const float inv = (float)a.spp;
            store_shading_output(*a.g, (size_t)y * a.band.width + x, v3(rad.x / inv, rad.y / inv, rad.z / inv), v3(0.f, 0.f, 0.f), true,
                                 a.diffuse_f16, a.specular_f16, a.diffuse_f32);
            return false;
        }
        thr = v3(1.f, 1.f, 1.f);
        setup_primary_ray(a.g->view, x, y, &o, &d);
        if (!a.primary_reuse) break;
        h = a.st_hit0[slot];                                      // the next sample's first trace(): same ray, same record
        if (a.sc.instances) h_inst = a.st_hit_inst0[slot];
    }
    a.st_ray0[slot] = make_float4(o.x, o.y, o.z, d.x);
    a.st_ray1[slot] = make_float4(d.y, d.z, thr.x, thr.y);
    a.st_acc[slot] = make_float4(thr.z, rad.x, rad.y, rad.z);
    a.st_misc[slot] = make_uint4(rng.index, sample | (bounce << 16), 0u, 0u);
    return true;
}

// warp-aggregated counter flush (every lane of the warp calls this)
__device__ __forceinline__ void warp_flush_counter(unsigned long long* dst, uint32_t v) {
    const uint32_t sum = __reduce_add_sync(0xFFFFFFFFu, v);
    if ((threadIdx.x & 31u) == 0u && sum) atomicAdd(dst, (unsigned long long)sum);
}

// Queue append with ONE global atomic per block instead of one per warp: a wavefront of 8 M paths makes 260 k same-address
// atomics per launch with warp_append, which serialise in the L2.  Every thread of the block calls this the same number of times;
// `s_off` is 2 x 9 words of shared memory (double-buffered by `parity`, so one barrier pair per call is enough).
__device__ __forceinline__ uint32_t block_append_index(bool alive, uint32_t* count_out, uint32_t* s_off, uint32_t parity) {
    uint32_t* o = s_off + parity * 9u;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const uint32_t m = __ballot_sync(0xFFFFFFFFu, alive);
    if (lane == 0u) o[warp] = (uint32_t)__popc(m);
    __syncthreads();
    if (threadIdx.x == 0u) {
        uint32_t total = 0;
        for (uint32_t w = 0; w < nwarps; w++) total += o[w];
        uint32_t base = total ? atomicAdd(count_out, total) : 0u;
        for (uint32_t w = 0; w < nwarps; w++) { const uint32_t c = o[w]; o[w] = base; base += c; }
    }
    __syncthreads();
    return o[warp] + (uint32_t)__popc(m & ((1u << lane) - 1u));
}

// 4 blocks of 256 threads per SM (64 registers): the kernel waits on dependent gathers (state -> shading record -> texels),
// so resident warps buy bandwidth; measured 1 / 3 / 4 / 5 blocks: 2488 / 2486 / 2532 / 2501 Mrays/s on C2 (5 spills).
#ifndef RT2_SHADE_MIN_BLOCKS
#define RT2_SHADE_MIN_BLOCKS 4
#endif
__global__ void __launch_bounds__(256, RT2_SHADE_MIN_BLOCKS) k_k1_shade(ShadeArgs a) {
    __shared__ float lut[256];
    lut[threadIdx.x] = a.sc.srgb_lut[threadIdx.x];
    __syncthreads();
    __shared__ uint32_t s_off[18];
    const uint32_t count = *a.count_in;
    const uint32_t stride = gridDim.x * blockDim.x;
    uint32_t hits = 0, seen = 0, parity = 0;
    // all threads of a block iterate together (block_append_index uses block barriers)
    // the queue entry of the next iteration is fetched one iteration ahead: one dependent memory round trip less per ray
    uint32_t b0 = blockIdx.x * blockDim.x;
    uint32_t slot_ahead = b0 + threadIdx.x < count ? a.queue_in[b0 + threadIdx.x] : 0u;
    for (; b0 < count; b0 += stride, parity ^= 1u) {
        const uint32_t i = b0 + threadIdx.x;
        bool alive = false;
        const uint32_t slot = slot_ahead;
        if (i + stride < count) slot_ahead = a.queue_in[i + stride];
        if (i < count) alive = k1_shade_step(a, lut, slot, &hits, &seen);
        const uint32_t at = block_append_index(alive, a.count_out, s_off, parity);
        if (alive) a.queue_out[at] = slot;
    }
    // one atomic pair per block (a launch over 262 k pixels has 8 k warps: same-address atomics serialise in the L2)
    __shared__ uint32_t s_cnt[2][8];
    const uint32_t wh = __reduce_add_sync(0xFFFFFFFFu, hits), ws = __reduce_add_sync(0xFFFFFFFFu, seen);
    if ((threadIdx.x & 31u) == 0u) { s_cnt[0][threadIdx.x >> 5] = wh; s_cnt[1][threadIdx.x >> 5] = ws; }
    __syncthreads();
    if (threadIdx.x < 2u) {
        uint32_t sum = 0;
        for (uint32_t w = 0; w < (blockDim.x >> 5); w++) sum += s_cnt[threadIdx.x][w];
        unsigned long long* dst = threadIdx.x == 0u ? a.stat_hits : a.stat_rays;
        if (sum && dst) atomicAdd(dst, (unsigned long long)sum);
    }
}

// ------------------------------------------------------------------------------------------------
// Warp-local wavefronts: the whole frame in ONE persistent launch.
// Every warp owns a population of up to P paths (slots in its private queue).  It alternates between a shade phase — the
// k_k1_shade body over its population, 32 paths at a time at full lane width, finished pixels replaced by new ones pulled from
// a global pixel counter — and a trace phase, the postponing / lane-refilling traversal loop over the same queue.  No kernel
// boundary and no grid-wide barrier per bounce: a warp's tail (its last long rays) overlaps other warps' phases instead of
// idling the GPU, and a warp that shades (latency bound) shares its SM with warps that traverse (issue bound).  Traversal
// state is scoped to the trace phase, so the register budget is max(trace, shade), not the sum.  The image is bit-identical
// to the launch-per-bounce form: same per-pixel state machine, closest hit = total order, RNG keyed on pixel + frame.
// ------------------------------------------------------------------------------------------------
struct WarpPathArgs {
    uint32_t* pop;            // num_warps x P slot queue, warp-private rows
    uint32_t P;               // population per warp (multiple of 32)
    uint32_t* pixel_head;     // next unclaimed padded lane of [first, end)
    uint32_t first, end;
    // list mode (the speculative renderer's fallback): pixels come from `list` (*list_count slots) and resume at the sample, RNG
    // position and radiance recorded in px_prog / px_rad instead of starting fresh
    const uint32_t* list; const uint32_t* list_count; const uint2* px_prog; const float4* px_rad;
};

template <bool INSTANCED, bool FLAT>
__device__ __forceinline__ void path_warp_body(const TraceArgs& t, const ShadeArgs& a, const WarpPathArgs& w, const float* lut) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    uint32_t* q = w.pop + (size_t)warp * w.P;
    const uint32_t tiles_x = (a.band.width + 7u) >> 3;
    uint32_t n = 0, rays = 0, hits = 0, seen = 0;
    bool exhausted = false;
    for (;;) {
        // ---- shade phase: every path of the population has a hit record waiting --------------------------------
        uint32_t wr = 0;
        for (uint32_t i0 = 0; i0 < n; i0 += 32u) {
            const uint32_t i = i0 + lane;
            bool alive = false;
            uint32_t slot = 0;
            if (i < n) { slot = q[i]; alive = k1_shade_step(a, lut, slot, &hits, &seen); }
            const uint32_t m = __ballot_sync(0xFFFFFFFFu, alive);      // also orders this batch's queue reads before its writes
            if (alive) q[wr + __popc(m & ((1u << lane) - 1u))] = slot;
            wr += __popc(m);
        }
        // finished pixels are replaced: pull padded lanes (8x4 tiles, so a fresh batch of 32 is one coherent tile)
        while (!exhausted && wr < w.P) {
            const uint32_t want = min(32u, w.P - wr);
            uint32_t base = 0;
            if (lane == 0) base = atomicAdd(w.pixel_head, want);
            base = __shfl_sync(0xFFFFFFFFu, base, 0);
            const uint32_t i = w.first + base + lane;
            bool alive = false;
            uint32_t slot = 0;
            if (w.list) {
                const uint32_t n_list = *w.list_count;
                if (lane < want && base + lane < n_list) {
                    alive = true;
                    slot = w.list[base + lane];
                    const uint32_t x = slot % a.band.width, ly = slot / a.band.width;
                    const uint2 pr = w.px_prog[slot];
                    const float4 rd = w.px_rad[slot];
                    V3 o, d;
                    setup_primary_ray(a.g->view, x, band_local_to_global(a.band, ly), &o, &d);
                    a.st_ray0[slot] = make_float4(o.x, o.y, o.z, d.x);
                    a.st_ray1[slot] = make_float4(d.y, d.z, 1.0f, 1.0f);
                    a.st_acc[slot] = make_float4(1.0f, rd.x, rd.y, rd.z);
                    a.st_misc[slot] = make_uint4(pr.y, pr.x, 0u, 0u);
                }
                const uint32_t m = __ballot_sync(0xFFFFFFFFu, alive);
                if (alive) q[wr + __popc(m & ((1u << lane) - 1u))] = slot;
                wr += __popc(m);
                if (base + want >= n_list) exhausted = true;
                continue;
            }
            if (lane < want && i < w.end) {
                const uint32_t tile = i >> 5, within = i & 31u;
                const uint32_t x = (tile % tiles_x) * 8u + (within & 7u), ly = (tile / tiles_x) * 4u + (within >> 3);
                if (x < a.band.width && ly < a.band.local_rows) {
                    alive = true;
                    slot = ly * a.band.width + x;
                    V3 o, d;
                    setup_primary_ray(a.g->view, x, band_local_to_global(a.band, ly), &o, &d);
                    a.st_ray0[slot] = make_float4(o.x, o.y, o.z, d.x);
                    a.st_ray1[slot] = make_float4(d.y, d.z, 1.0f, 1.0f);
                    a.st_acc[slot] = make_float4(1.0f, 0.0f, 0.0f, 0.0f);
                    a.st_misc[slot] = make_uint4(1u, 0u, 0u, 0u);
                }
            }
            const uint32_t m = __ballot_sync(0xFFFFFFFFu, alive);
            if (alive) q[wr + __popc(m & ((1u << lane) - 1u))] = slot;
            wr += __popc(m);
            if (w.first + base + want >= w.end) exhausted = true;
        }
        n = wr;
        if (n == 0u) break;
        __syncwarp();          // queue + path state written by other lanes of this warp are visible from here on
        // ---- trace phase ------------------------------------------------------------------------------------------
        trace_dyn_body<20, 4, 8, INSTANCED, false, true, FLAT>(t, q, n);
        rays += n;
        __syncwarp();
    }
    if (lane == 0) { if (rays) atomicAdd(t.stat_rays, (unsigned long long)rays); }     // walks
    warp_flush_counter(a.stat_hits, hits);
    if (a.stat_rays) warp_flush_counter(a.stat_rays, seen);                               // hit records consumed = the shader's trace() calls
}

template <bool INSTANCED, bool FLAT = false>
__global__ void __launch_bounds__(128) k_path_warp(TraceArgs t, ShadeArgs a, WarpPathArgs w) {
    __shared__ float lut[256];
    lut[threadIdx.x] = a.sc.srgb_lut[threadIdx.x]; lut[threadIdx.x + 128] = a.sc.srgb_lut[threadIdx.x + 128];
    __syncthreads();
    path_warp_body<INSTANCED, FLAT>(t, a, w, lut);
}

// ------------------------------------------------------------------------------------------------
// Speculative sample parallelism (B200: 180 GB of HBM holds every sample of every pixel as a path of its own).
// The shader runs a pixel's samples one after the other on ONE RNG stream (refrence.rgen:21-46): sample s starts at stream
// position 1 + 3 * (hits of samples 0..s-1), three draws per hit.  A sample that never misses takes exactly max_bounces hits, so
// in a closed scene the start position of every sample is known in advance: all spp samples of all pixels are traced as
// independent paths — spp x more rays per wavefront, max_bounces wavefront iterations per frame instead of spp x max_bounces —
// each path recording its per-bounce radiance contributions thr * emission.  k_spec_reduce then walks a pixel's samples in
// order, adds the contributions in the shader's order (fp32 addition is not associative: same order => same bits) and stops
// at the first sample that ended on a miss: the samples after it started from a wrong stream position and are traced again
// (fix-up rounds over the pending pixels, then k_path_warp's list mode, which finishes whatever is left sequentially).
// Only valid samples are counted as rays.  The image is bit-identical to the sequential form.
// ------------------------------------------------------------------------------------------------
struct SpecArgs {
    Band band;
    const rt2_gconst* g;
    uint32_t frame, spp, max_bounces, flags;
    SceneRefs sc;
    float4 *ray0, *ray1; uint4* misc;                 // per path: o|d.x, d.yz|thr.xy, {rng index, bounce, pixel_slot * spp + sample, thr.z}
    const float4* hit; const uint32_t* hit_inst;
    float4* contrib;                                  // per (pixel, sample, bounce): thr * emission of that event
    uint16_t* nhits;                                  // per (pixel, sample): hits | 0x80 if the sample ended on a miss | which bounces left a contribution << 8
    // sparse contributions (max_bounces <= 8): a contribution that is +0 in every component is neither stored nor added — x + (+0) == x bit for
    // bit for every x the sum can hold (it starts at +0 and only a (-0) + (-0) could make it -0) — and in a closed scene almost all of them are
    uint32_t sparse;
    uint32_t* start_index;                            // per (pixel, sample): the RNG position the sample's path was started from
    // candidate round: misc.z / contrib / nhits are indexed by (pending-list position, candidate) = k * cand_K + m instead of
    // (pixel, sample), and the pixel is cand_list[k]
    const uint32_t* cand_list; uint32_t cand_K;
    uint32_t path_base;                               // first path / candidate slot of this wavefront chain's region
    uint2* px_prog; float4* px_rad;                   // per pixel: {next sample, RNG position}, radiance summed so far
    ushort4* diffuse_f16; ushort4* specular_f16; float4* diffuse_f32;
    unsigned long long *stat_rays, *stat_hits;
};

// pixel slot of list entry k: round 0 walks the padded 8x4-tile lanes [first, first + lanes), later rounds a pending list
template <bool IMPLICIT>
__device__ __forceinline__ bool spec_pixel(const SpecArgs& a, const uint32_t* list, uint32_t first, uint32_t k, uint32_t* slot) {
    if (!IMPLICIT) { *slot = list[k]; return true; }
    const uint32_t i = first + k;
    const uint32_t tiles_x = (a.band.width + 7u) >> 3;
    const uint32_t tile = i >> 5, within = i & 31u;
    const uint32_t x = (tile % tiles_x) * 8u + (within & 7u), ly = (tile / tiles_x) * 4u + (within >> 3);
    *slot = ly * a.band.width + x;
    return x < a.band.width && ly < a.band.local_rows;
}

// One path per (pending pixel, remaining sample) from a speculated RNG position: the pixel's position + 3 x the predicted hits of
// the samples in between.  Round 0 predicts max_bounces hits per sample (a path that never misses); a fix-up round predicts what
// the previous round observed for that sample — its path came from a wrong position, but what makes a sample short is mostly
// geometry (a primary ray that leaves through a crack misses whatever the random numbers are).
// Primary-hit reuse (k1_shade_step): the paths of a pixel start with the same primary ray, so only the first of them — the
// representative — is queued for the first walk (queue_rep), and bounce 0 of every path reads the representative's hit record.  The
// paths of a pixel are consecutive threads and alive as a prefix (sample < spp), block_append_index hands out consecutive places in
// thread order, so a path's representative sits `back` = min(j, threadIdx.x) places before it (a pixel whose paths straddle a block
// has one representative per part); `back` travels in the upper half of misc.y until bounce 0 is shaded.  queue_rep == nullptr: off.
template <bool IMPLICIT>
__global__ void __launch_bounds__(256) k_spec_gen(SpecArgs a, const uint32_t* list, const uint32_t* list_count, uint32_t first, uint32_t lanes,
                                                  uint32_t* queue_out, uint32_t* count_out, uint32_t path_base, uint32_t* queue_rep, uint32_t* count_rep) {
    const uint32_t n_list = IMPLICIT ? lanes : *list_count;
    const uint64_t n_items = (uint64_t)n_list * a.spp;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    __shared__ uint32_t s_off[18], s_rep[18];
    uint32_t parity = 0;
    for (uint64_t b0 = (uint64_t)blockIdx.x * blockDim.x; b0 < n_items; b0 += stride, parity ^= 1u) {
        const uint64_t i = b0 + threadIdx.x;
        bool alive = false;
        uint32_t slot = 0, sample = 0, index = 0, x = 0, ly = 0, back = 0;
        if (i < n_items) {
            const uint32_t k = (uint32_t)(i / a.spp), j = (uint32_t)(i % a.spp);
            if (queue_rep) back = min(j, threadIdx.x);
            if (spec_pixel<IMPLICIT>(a, list, first, k, &slot)) {
                uint2 pr = make_uint2(0u, 1u);                 // refrence.rgen:21: a fresh pixel starts sample 0 at position 1
                if (IMPLICIT) { if (j == 0u) { a.px_prog[slot] = pr; a.px_rad[slot] = make_float4(0.f, 0.f, 0.f, 0.f); } }
                else pr = a.px_prog[slot];
                sample = pr.x + j;
                if (sample < a.spp) {
                    alive = true; x = slot % a.band.width; ly = slot / a.band.width;
                    uint32_t predicted = a.max_bounces * j;
                    if (!IMPLICIT) { predicted = 0; for (uint32_t q = 0; q < j; q++) predicted += a.nhits[slot * a.spp + pr.x + q] & 0x7Fu; }
                    index = pr.y + 3u * predicted;
                }
            }
        }
        const uint32_t p = block_append_index(alive, count_out, s_off, parity);
        uint32_t rep_at = 0;
        if (queue_rep) rep_at = block_append_index(alive && back == 0u, count_rep, s_rep, parity);
        if (alive) {
            V3 o, d;
            setup_primary_ray(a.g->view, x, band_local_to_global(a.band, ly), &o, &d);
            a.ray0[path_base + p] = make_float4(o.x, o.y, o.z, d.x);
            a.ray1[path_base + p] = make_float4(d.y, d.z, 1.0f, 1.0f);
            a.misc[path_base + p] = make_uint4(index, back << 16, slot * a.spp + sample, __float_as_uint(1.0f));
            a.start_index[slot * a.spp + sample] = index;
            queue_out[p] = path_base + p;
            if (queue_rep && back == 0u) queue_rep[rep_at] = path_base + p;
        }
    }
}

// refrence.rgen:28-42 for one traced ray of each queued path; a path is one sample and ends with it
// shared_primary: this launch shades bounce 0 and the hit record is the representative's (k_spec_gen)
__global__ void __launch_bounds__(256, RT2_SHADE_MIN_BLOCKS) k_spec_shade(SpecArgs a, const uint32_t* queue_in, const uint32_t* count_in,
                                                                          uint32_t* queue_out, uint32_t* count_out, uint32_t shared_primary) {
    __shared__ float lut[256];
    lut[threadIdx.x] = a.sc.srgb_lut[threadIdx.x];
    __syncthreads();
    __shared__ uint32_t s_off[18];
    const uint32_t count = *count_in;
    const uint32_t stride = gridDim.x * blockDim.x;
    uint32_t b0 = blockIdx.x * blockDim.x, parity = 0;
    uint32_t p_ahead = b0 + threadIdx.x < count ? queue_in[b0 + threadIdx.x] : 0u;
    for (; b0 < count; b0 += stride, parity ^= 1u) {
        const uint32_t i = b0 + threadIdx.x;
        bool alive = false;
        const uint32_t p = p_ahead;
        if (i + stride < count) p_ahead = queue_in[i + stride];
        if (i < count) {
            const float4 r0 = a.ray0[p], r1 = a.ray1[p];
            const uint4 misc = a.misc[p];
            const uint32_t hp = shared_primary ? p - ((misc.y >> 16) & 0xFFu) : p;
            const float4 h = a.hit[hp];
            const uint32_t ps = misc.z, slot = a.cand_list ? a.cand_list[(ps - a.path_base) / a.cand_K] : ps / a.spp;
            const uint32_t x = slot % a.band.width, y = band_local_to_global(a.band, slot / a.band.width);
            V3 o = v3(r0.x, r0.y, r0.z), d = v3(r0.w, r1.x, r1.y);
            V3 thr = v3(r1.z, r1.w, __uint_as_float(misc.w));
            Rng rng; rng.index = misc.x; rng.seed = jenkins_hash(zcurve_to_linear(x, y)) + (a.frame + 13u);
            uint32_t bounce = misc.y & 0xFFFFu;
            uint32_t cmask = misc.y >> 24;                         // bounces of this path that left a contribution (bits 16..23 hold `back` at bounce 0)
            const uint32_t tri_slot = __float_as_uint(h.w);
            float4* c = a.contrib + ((size_t)ps * a.max_bounces + bounce);
            if (tri_slot == 0xFFFFFFFFu) {                         // refrence.rgen:32-36
                const V3 e = thr * environment_radiance(a.sc, *a.g, d);
                if (!a.sparse || (__float_as_uint(e.x) | __float_as_uint(e.y) | __float_as_uint(e.z)) != 0u) { *c = make_float4(e.x, e.y, e.z, 0.f); if (a.sparse) cmask |= 1u << bounce; }
                a.nhits[ps] = (uint16_t)(bounce | 0x80u | (cmask << 8));
            } else {
                const SurfaceFetch sf = get_geometry_from_hit(a.sc, lut, tri_slot, a.sc.instances ? a.hit_inst[hp] : 0u, h.y, h.z);
                const V3 e = thr * sf.emission;                    // refrence.rgen:38
                if (!a.sparse || (__float_as_uint(e.x) | __float_as_uint(e.y) | __float_as_uint(e.z)) != 0u) { *c = make_float4(e.x, e.y, e.z, 0.f); if (a.sparse) cmask |= 1u << bounce; }
                thr = thr * sf.color;                              // refrence.rgen:39
                const float rx = rng_next(rng), ry = rng_next(rng), rz = rng_next(rng);
                (void)rx;
                const V3 hvec = sample_cos_hemisphere(ry, rz, (a.flags & RT2_LAUNCH_TRIG_LIBDEVICE) != 0u);
                const V3 world_pos = o + d * h.x;                  // Hit.glsl:67
                d = tangent_to_world(sf.normal, hvec);
                o = world_pos;                                     // refrence.rgen:42
                bounce++;
                if (bounce >= a.max_bounces) a.nhits[ps] = (uint16_t)(bounce | (cmask << 8));
                else {
                    alive = true;
                    a.ray0[p] = make_float4(o.x, o.y, o.z, d.x);
                    a.ray1[p] = make_float4(d.y, d.z, thr.x, thr.y);
                    a.misc[p] = make_uint4(rng.index, bounce | (cmask << 24), ps, __float_as_uint(thr.z));
                }
            }
        }
        const uint32_t at = block_append_index(alive, count_out, s_off, parity);
        if (alive) queue_out[at] = p;
    }
}

// per pending pixel: contributions of its valid samples added in the shader's order.  A sample is valid iff its path was started
// from the position the stream really has there (1 + 3 x hits of the samples before it); the first sample that was not sends
// the pixel to the next round's list.  The first pending sample is always valid, so every round makes progress.  refrence.rgen:46 when done.
template <bool IMPLICIT>
__global__ void __launch_bounds__(256) k_spec_reduce(SpecArgs a, const uint32_t* list, const uint32_t* list_count, uint32_t first, uint32_t lanes,
                                                     uint32_t* pending_out, uint32_t* pending_count) {
    const uint32_t n_list = IMPLICIT ? lanes : *list_count;
    const uint32_t stride = gridDim.x * blockDim.x;
    uint32_t rays = 0, hits = 0;
    for (uint32_t k0 = (blockIdx.x * blockDim.x + threadIdx.x) & ~31u; k0 < n_list; k0 += stride) {
        const uint32_t k = k0 + (threadIdx.x & 31u);
        bool pending = false;
        uint32_t slot = 0;
        if (k < n_list && spec_pixel<IMPLICIT>(a, list, first, k, &slot)) {
            const uint2 pr = a.px_prog[slot];
            const float4 r4 = a.px_rad[slot];
            V3 rad = v3(r4.x, r4.y, r4.z);
            uint32_t sample = pr.x, index = pr.y;
            while (sample < a.spp) {
                const uint32_t ps = slot * a.spp + sample;
                if (a.start_index[ps] != index) break;
                const uint32_t nh = a.nhits[ps], h = nh & 0x7Fu, missed = (nh >> 7) & 1u, cm = a.sparse ? nh >> 8 : 0xFFFFFFFFu;
                const float4* c = a.contrib + (size_t)ps * a.max_bounces;
                for (uint32_t b = 0; b < h + missed; b++) if ((cm >> (b & 31u)) & 1u) { const float4 e = c[b]; rad = rad + v3(e.x, e.y, e.z); }
                index += 3u * h; rays += h + missed; hits += h;
                sample++;
            }
            a.px_prog[slot] = make_uint2(sample, index);
            a.px_rad[slot] = make_float4(rad.x, rad.y, rad.z, 0.f);
            if (sample >= a.spp) {
                const uint32_t x = slot % a.band.width, y = band_local_to_global(a.band, slot / a.band.width);
                const float inv = (float)a.spp;
                store_shading_output(*a.g, (size_t)y * a.band.width + x, v3(rad.x / inv, rad.y / inv, rad.z / inv), v3(0.f, 0.f, 0.f), true,
                                     a.diffuse_f16, a.specular_f16, a.diffuse_f32);
            } else pending = true;
        }
        warp_append(pending, slot, pending_out, pending_count);
    }
    warp_flush_counter(a.stat_rays, rays);
    warp_flush_counter(a.stat_hits, hits);
}

// Candidate round, for the pixels no prediction settles (next to a crack a bounce ray leaks at random: with 32 samples such a
// pixel mispredicts again and again, and finishing it one dependent ray after the other is a ~4 ms latency chain).  The path of a
// sample is a function of (pixel, start position) only — the primary ray has no jitter (Q12) and the stream is keyed on pixel +
// frame — and the remaining samples can only start at idx0 + 3 m, m = 0 .. max_bounces x (remaining - 1).  So ALL those
// candidate paths are traced in one parallel round, and k_spec_reduce_cand walks the chain m -> m + hits(m) in order.
__global__ void __launch_bounds__(256) k_spec_gen_cand(SpecArgs a, const uint32_t* list, const uint32_t* list_count, uint32_t max_pixels,
                                                       uint32_t* queue_out, uint32_t* count_out, uint32_t* queue_rep, uint32_t* count_rep) {
    __shared__ uint32_t s_off[18], s_rep[18];
    const uint32_t n_list = min(*list_count, max_pixels);
    const uint64_t n_items = (uint64_t)n_list * a.cand_K;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    uint32_t parity = 0;
    for (uint64_t b0 = (uint64_t)blockIdx.x * blockDim.x; b0 < n_items; b0 += stride, parity ^= 1u) {
        const uint64_t i = b0 + threadIdx.x;
        bool alive = false;
        uint32_t slot = 0, index = 0, back = 0;
        if (i < n_items) {
            const uint32_t k = (uint32_t)(i / a.cand_K), m = (uint32_t)(i % a.cand_K);
            if (queue_rep) back = min(m, threadIdx.x);            // the candidates of a pixel are alive as a prefix in m
            slot = list[k];
            const uint2 pr = a.px_prog[slot];
            if (pr.x < a.spp && m <= a.max_bounces * (a.spp - pr.x - 1u)) { alive = true; index = pr.y + 3u * m; }
        }
        const uint32_t p = block_append_index(alive, count_out, s_off, parity);
        uint32_t rep_at = 0;
        if (queue_rep) rep_at = block_append_index(alive && back == 0u, count_rep, s_rep, parity);
        if (alive) {
            V3 o, d;
            setup_primary_ray(a.g->view, slot % a.band.width, band_local_to_global(a.band, slot / a.band.width), &o, &d);
            a.ray0[a.path_base + p] = make_float4(o.x, o.y, o.z, d.x);
            a.ray1[a.path_base + p] = make_float4(d.y, d.z, 1.0f, 1.0f);
            a.misc[a.path_base + p] = make_uint4(index, back << 16, a.path_base + (uint32_t)i, __float_as_uint(1.0f));
            queue_out[p] = a.path_base + p;
            if (queue_rep && back == 0u) queue_rep[rep_at] = a.path_base + p;
        }
    }
}

__global__ void __launch_bounds__(256) k_spec_reduce_cand(SpecArgs a, const uint32_t* list, const uint32_t* list_count, uint32_t max_pixels,
                                                          uint32_t* overflow_out, uint32_t* overflow_count) {
    const uint32_t n_all = *list_count;
    const uint32_t stride = gridDim.x * blockDim.x;
    uint32_t rays = 0, hits = 0;
    for (uint32_t k0 = (blockIdx.x * blockDim.x + threadIdx.x) & ~31u; k0 < n_all; k0 += stride) {
        const uint32_t k = k0 + (threadIdx.x & 31u);
        bool overflow = false;
        uint32_t slot = 0;
        if (k < n_all) {
            slot = list[k];
            if (k >= max_pixels) overflow = true;                 // more pending pixels than candidate storage: sequential fallback
            else {
                const uint2 pr = a.px_prog[slot];
                const float4 r4 = a.px_rad[slot];
                V3 rad = v3(r4.x, r4.y, r4.z);
                uint32_t m = 0;
                for (uint32_t sample = pr.x; sample < a.spp; sample++) {
                    const size_t cps = (size_t)a.path_base + (size_t)k * a.cand_K + m;
                    const uint32_t nh = a.nhits[cps], h = nh & 0x7Fu, missed = (nh >> 7) & 1u, cm = a.sparse ? nh >> 8 : 0xFFFFFFFFu;
                    const float4* c = a.contrib + cps * a.max_bounces;
                    for (uint32_t b = 0; b < h + missed; b++) if ((cm >> (b & 31u)) & 1u) { const float4 e = c[b]; rad = rad + v3(e.x, e.y, e.z); }
                    m += h; rays += h + missed; hits += h;
                }
                a.px_prog[slot] = make_uint2(a.spp, pr.y + 3u * m);
                const uint32_t x = slot % a.band.width, y = band_local_to_global(a.band, slot / a.band.width);
                const float inv = (float)a.spp;
                store_shading_output(*a.g, (size_t)y * a.band.width + x, v3(rad.x / inv, rad.y / inv, rad.z / inv), v3(0.f, 0.f, 0.f), true,
                                     a.diffuse_f16, a.specular_f16, a.diffuse_f32);
            }
        }
        warp_append(overflow, slot, overflow_out, overflow_count);
    }
    warp_flush_counter(a.stat_rays, rays);
    warp_flush_counter(a.stat_hits, hits);
}

// primary visibility probe: one thread per pixel, 8x4 tiles per warp
template <bool INSTANCED, bool FLAT = false>
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
    if (FLAT) flat_closest_hit(a.flat_nodes, a.flat_leaf, a.instances, a.tris, ray, best, nullptr);
    else if (INSTANCED) tlas_closest_hit(a.tlas_nodes, a.tlas_slot_to_inst, a.instances, a.blas_nodes, a.tris, ray, best, nullptr);
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
                    f.st_hit, f.st_hit_inst, f.st_hit0, f.st_hit_inst0, f.queue0, f.queue1, f.counters, f.warp_pop,
                    f.spec.ray0, f.spec.ray1, f.spec.hit, f.spec.contrib, f.spec.px_rad, f.spec.misc, f.spec.hit_inst, f.spec.queue0, f.spec.queue1,
                    f.spec.pending0, f.spec.pending1, f.spec.counters, f.spec.nhits, f.spec.px_prog, f.spec.start_index};
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
    RT2_CUDA(ctx, dev_alloc(ctx, &f.st_hit0, n));
    RT2_CUDA(ctx, dev_alloc(ctx, &f.st_hit_inst0, n));
    // queues are indexed in the padded 8x4-tile lane range: a wavefront chain's window starts at its first padded lane, so
    // dead lanes of earlier chains shift later windows past w*h when w is not a multiple of 8 (or rows of 4)
    const size_t padded = (size_t)((w + 7u) / 8u) * ((h + 3u) / 4u) * 32u;
    RT2_CUDA(ctx, dev_alloc(ctx, &f.queue0, padded));
    RT2_CUDA(ctx, dev_alloc(ctx, &f.queue1, padded));
    f.counter_capacity = 0;
    return RT2_OK;
}

static int check_band(rt2_ctx* ctx, const rt2_launch_params& lp) {
    if (lp.band_world > 1u) {
        if (lp.band_rows == 0u) return fail(ctx, RT2_E_INVALID, "band_world > 1 needs band_rows > 0");
        if (lp.band_rank >= lp.band_world) return fail(ctx, RT2_E_INVALID, "band_rank must be < band_world");
    }
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
    const bool flat = s.flat.nodes && ctx->use_flat;
    t.flat_nodes = flat ? s.flat.nodes : nullptr; t.flat_leaf = flat ? s.flat_leaf : nullptr;
    t.st_ray0 = f.st_ray0; t.st_ray1 = f.st_ray1; t.st_hit = f.st_hit; t.st_hit_inst = f.st_hit_inst;
    if (flat) { const uint32_t d = s.flat.depth + 2u; t.postpone_limit = d < RT2_STACK_SIZE ? RT2_STACK_SIZE - d : 0u; return t; }
    const uint32_t depth = s.blas.depth + (s.num_instances ? 2u * s.tlas.depth + 2u : 0u) + 2u;    // node groups a ray can have stacked at once
    t.postpone_limit = depth < RT2_STACK_SIZE ? RT2_STACK_SIZE - depth : 0u;
    return t;
}

// counters layout: [0..1] u64 rays (the shader's trace() calls), [2..3] u64 hits, [4..5] u64 walks (K1: primary-hit reuse makes them
// fewer than the rays), then per iteration it: count at RT2_COUNTER_BASE + 2*it, head one after
#define RT2_COUNTER_BASE 6
static int ensure_counters(rt2_ctx* ctx, uint32_t iterations) {
    FrameBuffers& f = ctx->fb;
    const size_t need = RT2_COUNTER_BASE + RT2_MAX_CHAINS * 2 * ((size_t)iterations + 2);     // per chain and iteration: count, head
    if (f.counter_capacity < need) {
        if (f.counters) { RT2_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); cudaFree(f.counters); f.counters = nullptr; }
        RT2_CUDA(ctx, cudaMalloc((void**)&f.counters, need * 4));
        f.counter_capacity = need;
    }
    RT2_CUDA(ctx, cudaMemsetAsync(f.counters, 0, need * 4, ctx->stream));
    ctx->stats_walks_counted = false; ctx->stats_reused_counted = false;
    return RT2_OK;
}

// single-level trace variants (RT2_TRACE_VARIANT, read once per context: A/B runs on the GPU box)
// one-loop TLAS/BLAS walk: a TLAS level can hold two stack entries (instance remnant + inner remainder), a BLAS level one
static bool unified_walk_fits(const DeviceScene& s) { return 2u * s.tlas.depth + s.blas.depth + 6u <= RT2_STACK_SIZE; }

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
    k_trace_dyn<30, 4, 8>,     // 10
    k_trace_dyn<40, 4, 8>,     // 11
    k_trace_dyn<12, 4, 8>,     // 12
    k_trace_dyn<20, 6, 6>,     // 13
    k_trace_dyn<20, 3, 6>,     // 14
    k_trace_dyn<30, 4, 4>,     // 15 (r02b re-sweep after the tree-quality work: shorter walks favour checking more often)
    k_trace_dyn<25, 6, 6>,     // 16
    k_trace_dyn<30, 6, 6>,     // 17
    k_trace_dyn<25, 4, 6>,     // 18
    k_trace_dyn<30, 6, 4>,     // 19
    k_trace_dyn<35, 4, 4>,     // 20
    k_trace_dyn<25, 8, 4>,     // 21
};
static const int g_num_trace_variants = (int)(sizeof(g_trace_variants) / sizeof(g_trace_variants[0]));

int render_primary(rt2_ctx* ctx, const rt2_launch_params& lp) {
    int rc = check_ready(ctx); if (rc) return rc;
    if ((rc = check_band(ctx, lp))) return rc;
    FrameBuffers& f = ctx->fb;
    if ((rc = ensure_counters(ctx, 0))) return rc;
    const Band band = make_band(f.width, f.height, lp);
    const uint32_t tiles_x = (band.width + 7) / 8, tiles_y = (band.local_rows + 3) / 4;
    const uint32_t padded = tiles_x * tiles_y * 32;
    TraceArgs t = make_trace_args(ctx);
    unsigned long long* hits = reinterpret_cast<unsigned long long*>(f.counters + 2);
    if (padded) {
        if (ctx->scene.flat.nodes && ctx->use_flat) k_primary<false, true><<<(padded + 127) / 128, 128, 0, ctx->stream>>>(t, band, ctx->d_gconst, f.hit_ids, padded, hits);
        else if (ctx->scene.num_instances) k_primary<true><<<(padded + 127) / 128, 128, 0, ctx->stream>>>(t, band, ctx->d_gconst, f.hit_ids, padded, hits);
        else k_primary<false><<<(padded + 127) / 128, 128, 0, ctx->stream>>>(t, band, ctx->d_gconst, f.hit_ids, padded, hits);
    }
    RT2_CUDA(ctx, cudaGetLastError());
    ctx->stats.kernel_launches = padded ? 1 : 0;
    ctx->stats.wavefront_iterations = 1;
    ctx->pending_rays_static = (uint64_t)band.width * band.local_rows;
    ctx->stats_from_counters = true;
    return RT2_OK;
}

static int render_reference_seq(rt2_ctx* ctx, uint32_t frame, const rt2_launch_params& lp) {
    int rc = check_ready(ctx); if (rc) return rc;
    if ((rc = check_band(ctx, lp))) return rc;
    FrameBuffers& f = ctx->fb; const DeviceScene& s = ctx->scene;
    const uint32_t spp = lp.spp ? lp.spp : 12u;                    // refrence.rgen:17
    const uint32_t bounces = lp.max_bounces ? lp.max_bounces : 5u; // refrence.rgen:16
    if (spp > 65535u || bounces > 65535u || (uint64_t)spp * bounces > (1u << 22)) return fail(ctx, RT2_E_INVALID, "spp x bounces too large");
    // a pixel is a chain of spp x bounces dependent rays; with primary-hit reuse only the first sample walks its primary ray
    const bool reuse = ctx->primary_reuse && spp > 1u;
    const uint32_t iterations = reuse ? 1u + spp * (bounces - 1u) : spp * bounces;
    if ((rc = ensure_counters(ctx, iterations))) return rc;
    const Band band = make_band(f.width, f.height, lp);
    const uint32_t tiles_x = (band.width + 7) / 8, tiles_y = (band.local_rows + 3) / 4;
    const uint32_t padded = tiles_x * tiles_y * 32;
    uint64_t launches = 0;
    ShadeArgs a{};
    a.band = band; a.g = ctx->d_gconst; a.frame = frame; a.spp = spp; a.max_bounces = bounces; a.flags = lp.flags;
    a.sc = make_scene_refs(ctx);
    a.st_ray0 = f.st_ray0; a.st_ray1 = f.st_ray1; a.st_acc = f.st_acc; a.st_misc = f.st_misc; a.st_hit = f.st_hit; a.st_hit_inst = f.st_hit_inst;
    a.diffuse_f16 = f.diffuse_f16; a.specular_f16 = f.specular_f16; a.diffuse_f32 = f.diffuse_f32;
    a.stat_hits = reinterpret_cast<unsigned long long*>(f.counters + 2);
    // with reuse the shade steps count the trace() calls (hit records consumed) and the trace kernels the walks; without it they are the same number
    a.stat_rays = reuse ? reinterpret_cast<unsigned long long*>(f.counters) : nullptr;
    a.st_hit0 = f.st_hit0; a.st_hit_inst0 = f.st_hit_inst0; a.primary_reuse = reuse ? 1u : 0u;
    TraceArgs t = make_trace_args(ctx);
    t.stat_rays = reinterpret_cast<unsigned long long*>(f.counters + (reuse ? 4 : 0));
    ctx->stats_walks_counted = reuse;
    if (padded) {
        const int shade_blocks_full = ctx->num_sms * 8;
        const bool profile = (lp.flags & RT2_LAUNCH_PROFILE) != 0u;
        for (cudaEvent_t e : ctx->prof_events) cudaEventDestroy(e);
        ctx->prof_events.clear();
        if (profile) { ctx->prof_events.resize(ctx->path_mode == 1 ? 3 : (size_t)iterations * 3); for (auto& e : ctx->prof_events) RT2_CUDA(ctx, cudaEventCreate(&e)); }
        // Wavefront chains: the pixels of the frame are independent, so the tile range is cut into `chains` contiguous parts,
        // each running its own init -> (trace -> shade) x iterations sequence on its own stream with its own queues and
        // counters.  While one chain's trace kernel drains its last long rays (a ~45 us tail per launch that does not
        // shrink with the wavefront), the other chains' kernels fill the SMs.  Per-kernel event timing (LAUNCH_PROFILE)
        // needs serial kernels, so it runs one chain.
        // Small wavefronts (one rank's share of a sharded frame) also get a smaller persistent grid per chain: lane refill only
        // works while the queue is longer than the grid, so each chain launches about rays / (128 * rays_per_lane) blocks and the
        // chains together, not each alone, fill the resident block slots.
        // Measured on C2 (profiles/r01c_chain_sweep.txt, Mrays/s per GPU, 1 chain / 2 chains full grid / 8 chains capped grid):
        // 2.07 M pixels 2231 / 2322 / 2184, 1.04 M 2083 / 2165 / 2121, 691 k - / 1991 / 2060, 518 k 1856 / 1899 / 2014,
        // 346 k - / 1773 / 1882, 259 k (one of 8 ranks) 1488 / 1662 / 1752.  A scene whose trace launches last only microseconds
        // (C1, 32 triangles) is launch-rate bound and loses 10-25 % to the extra launches: one chain.
        const uint64_t pixels = (uint64_t)band.width * band.local_rows;
        const bool flat = s.flat.nodes && ctx->use_flat;
        if (ctx->path_mode == 1 && (!s.num_instances || flat || unified_walk_fits(s))) {
            // warp-local wavefronts: one persistent launch for the whole frame (see k_path_warp)
            const int occ = ctx->path_blocks_per_sm[flat ? 2 : (s.num_instances ? 1 : 0)];
            const int bps = ctx->warp_blocks_per_sm ? std::min(ctx->warp_blocks_per_sm, occ) : occ;
            int blocks = bps * ctx->num_sms;
            uint32_t P = ctx->warp_paths;
            if (!P) {
                // population per warp: 8 rays per lane when the frame is large enough, never below 2 per lane — a small frame
                // (one rank's bands) runs on fewer blocks instead
                const uint64_t per_warp = pixels / ((uint64_t)blocks * 4u);
                P = (uint32_t)std::min<uint64_t>(256u, std::max<uint64_t>(64u, per_warp & ~31ull));
                if (per_warp < 64u) blocks = std::max(ctx->num_sms, (int)std::min<uint64_t>((uint64_t)blocks, pixels / (64u * 4u) + 1u));
            }
            P = std::max(32u, P & ~31u);
            const size_t need = (size_t)blocks * 4u * P;
            if (f.warp_pop_cap < need) {
                if (f.warp_pop) { RT2_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); cudaFree(f.warp_pop); f.warp_pop = nullptr; f.warp_pop_cap = 0; }
                RT2_CUDA(ctx, cudaMalloc((void**)&f.warp_pop, need * 4));
                f.warp_pop_cap = need; ctx->device_bytes += need * 4;
            }
            WarpPathArgs w{f.warp_pop, P, f.counters + RT2_COUNTER_BASE, 0u, padded};
            if (profile) RT2_CUDA(ctx, cudaEventRecord(ctx->prof_events[0], ctx->stream));
            if (flat) k_path_warp<false, true><<<blocks, 128, 0, ctx->stream>>>(t, a, w);
            else if (s.num_instances) k_path_warp<true><<<blocks, 128, 0, ctx->stream>>>(t, a, w);
            else k_path_warp<false><<<blocks, 128, 0, ctx->stream>>>(t, a, w);
            if (profile) { RT2_CUDA(ctx, cudaEventRecord(ctx->prof_events[1], ctx->stream)); RT2_CUDA(ctx, cudaEventRecord(ctx->prof_events[2], ctx->stream)); }
            RT2_CUDA(ctx, cudaGetLastError());
            ctx->stats.kernel_launches = 1;
            ctx->stats.wavefront_iterations = iterations;
            ctx->stats_from_counters = true;
            ctx->pending_rays_static = 0;
            return RT2_OK;
        }
        const int full_blocks = ctx->trace_blocks_per_sm[flat ? 3 : (s.num_instances ? 1 : 0)] * ctx->num_sms;
        const bool small_wave = pixels < 900000u;
        const bool tiny_scene = s.num_tris < 4096u && !s.num_instances;      // an instanced scene's BLAS may be small; its trace launches are not
        const uint32_t auto_chains = tiny_scene ? 1u : (small_wave ? 8u : 2u);
        const uint32_t rays_per_lane = ctx->rays_per_lane_set ? ctx->rays_per_lane
                                     : (small_wave && !tiny_scene ? std::max<uint32_t>(1u, (uint32_t)(pixels / ((uint64_t)full_blocks * 128u))) : 0u);
        uint32_t chains = profile ? 1u : std::max(1u, std::min(ctx->chains ? ctx->chains : auto_chains, RT2_MAX_CHAINS));
        uint32_t per = ((padded + chains - 1) / chains + 255u) & ~255u;
        while (chains > 1u && (uint64_t)per * (chains - 1) >= padded) { chains--; per = ((padded + chains - 1) / chains + 255u) & ~255u; }
        if (chains > 1u) {
            if (!ctx->chain_fork) { RT2_CUDA(ctx, cudaEventCreateWithFlags(&ctx->chain_fork, cudaEventDisableTiming)); }
            for (uint32_t c = 1; c < chains; c++) {
                if (!ctx->chain_streams[c]) {
                    RT2_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->chain_streams[c], cudaStreamNonBlocking));
                    RT2_CUDA(ctx, cudaEventCreateWithFlags(&ctx->chain_join[c], cudaEventDisableTiming));
                }
            }
            RT2_CUDA(ctx, cudaEventRecord(ctx->chain_fork, ctx->stream));
        }
        const size_t counter_stride = 2 * ((size_t)iterations + 2);
        // Launches are enqueued iteration by iteration across the chains (not chain by chain), so that all chains advance
        // together and finish together instead of the last chain running alone at the end of the frame.
        struct Chain { cudaStream_t st; uint32_t first, end; uint32_t* cnt; uint32_t* q[2]; int shade_blocks, trace_blocks; } ch[RT2_MAX_CHAINS];
        for (uint32_t c = 0; c < chains; c++) {
            Chain& k = ch[c];
            k.st = c == 0 ? ctx->stream : ctx->chain_streams[c];
            if (c != 0) RT2_CUDA(ctx, cudaStreamWaitEvent(k.st, ctx->chain_fork, 0));
            k.first = c * per; k.end = std::min(padded, k.first + per);
            k.cnt = f.counters + RT2_COUNTER_BASE + c * counter_stride;
            k.q[0] = f.queue0 + k.first; k.q[1] = f.queue1 + k.first;
            k.shade_blocks = std::max(1, std::min(shade_blocks_full, (int)((k.end - k.first + 255u) / 256u)));
            // persistent grid: every resident slot, unless the chain is so small that each lane would get fewer than
            // rays_per_lane rays — lane refill needs a queue longer than the grid (RT2_RAYS_PER_LANE, 0 = always the full grid)
            k.trace_blocks = full_blocks;
            if (rays_per_lane && !profile) k.trace_blocks = std::max(ctx->num_sms, std::min(full_blocks, (int)((k.end - k.first) / (128u * rays_per_lane)) + 1));
            a.queue_out = k.q[0]; a.count_out = k.cnt;
            k_k1_init<<<(k.end - k.first + 255) / 256, 256, 0, k.st>>>(a, k.first, k.end); launches++;
        }
        for (uint32_t it = 0; it < iterations; it++) {
            for (uint32_t c = 0; c < chains; c++) {
                const Chain& k = ch[c];
                cudaStream_t st = k.st;
                if (profile) RT2_CUDA(ctx, cudaEventRecord(ctx->prof_events[3 * it], st));
                t.queue = k.q[it & 1]; t.count = k.cnt + 2 * it; t.head = k.cnt + 2 * it + 1;
                if (flat) {
                    k_trace_dyn_flat<RT2_FLAT_PARAMS><<<k.trace_blocks, 128, 0, st>>>(t);
                } else if (s.num_instances) {
                    // one-loop TLAS/BLAS walk unless the two trees together could overflow the shared stack (then: nested loops, own stack each)
                    const bool unified = ctx->trace_variant != 0 && unified_walk_fits(s);
                    if (unified) k_trace_dyn_tlas<20, 4, 8><<<k.trace_blocks, 128, 0, st>>>(t);
                    else k_trace<true><<<ctx->trace_blocks_per_sm[2] * ctx->num_sms, 128, 0, st>>>(t);
                } else {
                    g_trace_variants[ctx->trace_variant]<<<k.trace_blocks, 128, 0, st>>>(t);
                }
                if (profile) RT2_CUDA(ctx, cudaEventRecord(ctx->prof_events[3 * it + 1], st));
                a.queue_in = k.q[it & 1]; a.count_in = t.count; a.queue_out = k.q[(it + 1) & 1]; a.count_out = k.cnt + 2 * (it + 1);
                k_k1_shade<<<k.shade_blocks, 256, 0, st>>>(a);
                if (profile) RT2_CUDA(ctx, cudaEventRecord(ctx->prof_events[3 * it + 2], st));
                launches += 2;
            }
        }
        for (uint32_t c = 1; c < chains; c++) { RT2_CUDA(ctx, cudaEventRecord(ctx->chain_join[c], ch[c].st)); RT2_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->chain_join[c], 0)); }
    }
    RT2_CUDA(ctx, cudaGetLastError());
    ctx->stats.kernel_launches = launches;
    ctx->stats.wavefront_iterations = iterations;
    ctx->stats_from_counters = true;
    ctx->pending_rays_static = 0;
    return RT2_OK;
}

// ------------------------------------------------------------------------------------------------
// speculative sample parallelism: host side
// ------------------------------------------------------------------------------------------------
static void launch_trace_kernel(rt2_ctx* ctx, const TraceArgs& t, bool flat, int blocks, cudaStream_t st) {
    const DeviceScene& s = ctx->scene;
    if (flat) k_trace_dyn_flat<RT2_FLAT_PARAMS><<<blocks, 128, 0, st>>>(t);
    else if (s.num_instances) {
        if (ctx->trace_variant != 0 && unified_walk_fits(s)) k_trace_dyn_tlas<20, 4, 8><<<blocks, 128, 0, st>>>(t);
        else k_trace<true><<<ctx->trace_blocks_per_sm[2] * ctx->num_sms, 128, 0, st>>>(t);
    } else g_trace_variants[ctx->trace_variant]<<<blocks, 128, 0, st>>>(t);
}

template <class T>
static cudaError_t spec_grow(rt2_ctx* ctx, T** p, size_t n) {
    if (*p) { cudaFree(*p); *p = nullptr; }
    cudaError_t e = cudaMalloc((void**)p, std::max<size_t>(n, 1) * sizeof(T));
    if (e == cudaSuccess) ctx->device_bytes += std::max<size_t>(n, 1) * sizeof(T);
    return e;
}

static bool spec_eligible(rt2_ctx* ctx, const Band& band, uint32_t spp, uint32_t bounces) {
    const DeviceScene& s = ctx->scene;
    const uint64_t paths = (uint64_t)band.width * band.local_rows * spp;
    if (spp < 2u || bounces > 127u || paths == 0 || paths > (64ull << 20)) return false;
    if (paths * (16ull * 4 + 4 + 8 + 1 + 16ull * bounces) > (24ull << 30)) return false;
    if (s.num_instances && !(s.flat.nodes && ctx->use_flat) && !unified_walk_fits(s)) return false;     // the fallback needs the one-loop walk
    return true;
}

static int render_reference_spec(rt2_ctx* ctx, uint32_t frame, const rt2_launch_params& lp, uint32_t spp, uint32_t bounces) {
    FrameBuffers& f = ctx->fb; const DeviceScene& s = ctx->scene;
    FrameBuffers::Spec& sp = f.spec;
    int rc;
    if ((rc = ensure_counters(ctx, 0))) return rc;
    const Band band = make_band(f.width, f.height, lp);
    const uint32_t tiles_x = (band.width + 7) / 8, tiles_y = (band.local_rows + 3) / 4;
    const uint32_t padded = tiles_x * tiles_y * 32;
    const size_t band_pixels = (size_t)band.width * band.local_rows;
    const size_t paths = band_pixels * spp, contribs = paths * bounces;
    // prediction fix-up rounds before the candidate round: RT2_SPEC_ROUNDS, else 1 — or 0 when the previous frame says that so few samples
    // end on a miss (< 0.1 %) that the candidate round settles them for less than a round of near-empty launches costs (C2: 3041 -> 3095 Mrays/s)
    const uint32_t rounds = 1u + (uint32_t)(ctx->spec_rounds >= 0 ? ctx->spec_rounds : (ctx->spec_auto_few_misses ? 0 : 1));
    const uint32_t RB = 2u * (bounces + 1u) + 4u;                 // per round: (count, head) x (bounces + 1), (count, head) of the representatives' queue, pending count, pad
    const bool reuse = ctx->primary_reuse;
    const bool profile = (lp.flags & RT2_LAUNCH_PROFILE) != 0u;
    // two wavefront chains once a chain still gets >= 1 M paths (measured like the sequential form's rule); RT2_CHAINS overrides
    uint32_t nch = profile ? 1u : (ctx->chains ? std::min(ctx->chains, RT2_MAX_CHAINS) : (paths >= (2u << 20) ? 2u : 1u));
    nch = std::max(1u, std::min(nch, tiles_y ? tiles_y : 1u));
    const size_t CB = 8 + (size_t)(rounds + 1u) * RB;             // per chain: traced (u64), fallback head, pad; rounds + the candidate round
    const size_t n_counters = CB * nch;
    const bool flat = s.flat.nodes && ctx->use_flat;
    const int fb_blocks = ctx->num_sms;                            // fallback kernel: one block of 4 warps per SM
    const uint32_t fb_P = 64u;
    // grow-only buffers (first speculative frame of a size allocates; steady state allocates nothing)
    if (sp.cap_paths < paths || sp.cap_contrib < contribs || sp.cap_pixels < band_pixels || sp.cap_counters < n_counters ||
        f.warp_pop_cap < (size_t)fb_blocks * 4u * fb_P * RT2_MAX_CHAINS) {
        RT2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (sp.cap_paths < paths) {
            RT2_CUDA(ctx, spec_grow(ctx, &sp.ray0, paths)); RT2_CUDA(ctx, spec_grow(ctx, &sp.ray1, paths)); RT2_CUDA(ctx, spec_grow(ctx, &sp.hit, paths));
            RT2_CUDA(ctx, spec_grow(ctx, &sp.misc, paths)); RT2_CUDA(ctx, spec_grow(ctx, &sp.hit_inst, paths));
            RT2_CUDA(ctx, spec_grow(ctx, &sp.queue0, paths)); RT2_CUDA(ctx, spec_grow(ctx, &sp.queue1, paths)); RT2_CUDA(ctx, spec_grow(ctx, &sp.nhits, paths));
            RT2_CUDA(ctx, spec_grow(ctx, &sp.start_index, paths));
            sp.cap_paths = paths;
        }
        if (sp.cap_contrib < contribs) { RT2_CUDA(ctx, spec_grow(ctx, &sp.contrib, contribs)); sp.cap_contrib = contribs; }
        if (sp.cap_pixels < band_pixels) {
            RT2_CUDA(ctx, spec_grow(ctx, &sp.px_rad, band_pixels)); RT2_CUDA(ctx, spec_grow(ctx, &sp.px_prog, band_pixels));
            RT2_CUDA(ctx, spec_grow(ctx, &sp.pending0, band_pixels)); RT2_CUDA(ctx, spec_grow(ctx, &sp.pending1, band_pixels));
            sp.cap_pixels = band_pixels;
        }
        if (sp.cap_counters < n_counters) { RT2_CUDA(ctx, spec_grow(ctx, &sp.counters, n_counters)); sp.cap_counters = n_counters; }
        if (f.warp_pop_cap < (size_t)fb_blocks * 4u * fb_P * RT2_MAX_CHAINS) {
            if (f.warp_pop) cudaFree(f.warp_pop);
            f.warp_pop = nullptr; f.warp_pop_cap = 0;
            RT2_CUDA(ctx, cudaMalloc((void**)&f.warp_pop, (size_t)fb_blocks * 4u * fb_P * RT2_MAX_CHAINS * 4u));
            f.warp_pop_cap = (size_t)fb_blocks * 4u * fb_P * RT2_MAX_CHAINS; ctx->device_bytes += f.warp_pop_cap * 4;
        }
    }
    RT2_CUDA(ctx, cudaMemsetAsync(sp.counters, 0, n_counters * 4, ctx->stream));
    uint64_t launches = 0;
    if (padded) {
        SpecArgs a{};
        a.band = band; a.g = ctx->d_gconst; a.frame = frame; a.spp = spp; a.max_bounces = bounces; a.flags = lp.flags;
        a.sc = make_scene_refs(ctx);
        a.ray0 = sp.ray0; a.ray1 = sp.ray1; a.misc = sp.misc; a.hit = sp.hit; a.hit_inst = sp.hit_inst;
        a.sparse = bounces <= 8u && ctx->sparse_contrib ? 1u : 0u;
        a.contrib = sp.contrib; a.nhits = sp.nhits; a.start_index = sp.start_index; a.px_prog = sp.px_prog; a.px_rad = sp.px_rad;
        a.diffuse_f16 = f.diffuse_f16; a.specular_f16 = f.specular_f16; a.diffuse_f32 = f.diffuse_f32;
        a.stat_rays = reinterpret_cast<unsigned long long*>(f.counters);           // valid samples only (k_spec_reduce)
        a.stat_hits = reinterpret_cast<unsigned long long*>(f.counters + 2);
        unsigned long long* stat_walks = reinterpret_cast<unsigned long long*>(f.counters + 4);   // every walk, speculation losses included
        ctx->stats_walks_counted = true;
        TraceArgs t = make_trace_args(ctx);
        t.st_ray0 = sp.ray0; t.st_ray1 = sp.ray1; t.st_hit = sp.hit; t.st_hit_inst = sp.hit_inst;
        for (cudaEvent_t e : ctx->prof_events) cudaEventDestroy(e);
        ctx->prof_events.clear();
        if (profile) { ctx->prof_events.resize((size_t)bounces * 3); for (auto& e : ctx->prof_events) RT2_CUDA(ctx, cudaEventCreate(&e)); }
        const int full_trace = ctx->trace_blocks_per_sm[flat ? 3 : (s.num_instances ? 1 : 0)] * ctx->num_sms;
        const int full_shade = ctx->num_sms * 8;
        if (nch > 1u) {
            if (!ctx->chain_fork) { RT2_CUDA(ctx, cudaEventCreateWithFlags(&ctx->chain_fork, cudaEventDisableTiming)); }
            for (uint32_t c = 1; c < nch; c++) {
                if (!ctx->chain_streams[c]) {
                    RT2_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->chain_streams[c], cudaStreamNonBlocking));
                    RT2_CUDA(ctx, cudaEventCreateWithFlags(&ctx->chain_join[c], cudaEventDisableTiming));
                }
            }
            RT2_CUDA(ctx, cudaEventRecord(ctx->chain_fork, ctx->stream));
        }
        // Wavefront chains: the frame is cut at tile-row boundaries, so a chain owns the pixel slots [slot_lo, slot_hi) and with
        // them the regions [slot_lo * spp, slot_hi * spp) of the path, contribution and candidate arrays, its own queues, pending
        // lists and counters; one chain's trace tail overlaps the other's work.
        struct Chain {
            cudaStream_t st; uint32_t first, lanes, base; size_t cap_paths; uint32_t* cnt; uint32_t* q[2]; uint32_t* pend[2];
        } ch[RT2_MAX_CHAINS];
        for (uint32_t c = 0; c < nch; c++) {
            Chain& k = ch[c];
            const uint32_t tr0 = (uint32_t)((uint64_t)tiles_y * c / nch), tr1 = (uint32_t)((uint64_t)tiles_y * (c + 1) / nch);
            const uint32_t row0 = std::min(band.local_rows, tr0 * 4u), row1 = std::min(band.local_rows, tr1 * 4u);
            k.st = c == 0 ? ctx->stream : ctx->chain_streams[c];
            if (c != 0) RT2_CUDA(ctx, cudaStreamWaitEvent(k.st, ctx->chain_fork, 0));
            k.first = tr0 * tiles_x * 32u; k.lanes = (tr1 - tr0) * tiles_x * 32u;
            k.base = row0 * band.width * spp; k.cap_paths = (size_t)(row1 - row0) * band.width * spp;
            k.cnt = sp.counters + (size_t)c * CB;
            k.q[0] = sp.queue0 + k.base; k.q[1] = sp.queue1 + k.base;
            k.pend[0] = sp.pending0 + (size_t)row0 * band.width; k.pend[1] = sp.pending1 + (size_t)row0 * band.width;
        }
        auto round_counters = [&](const Chain& k, uint32_t r) { return k.cnt + 8 + (size_t)r * RB; };
        for (uint32_t r = 0; r < rounds; r++) {
            // fix-up rounds are small (the pixels that saw a miss): a fraction of the grid, every kernel is grid-stride / persistent
            for (uint32_t c = 0; c < nch; c++) {
                const Chain& k = ch[c];
                uint32_t* cnt = round_counters(k, r);
                const int gen_blocks = r ? ctx->num_sms * 2 : (int)std::max<uint64_t>(1, std::min<uint64_t>((uint64_t)ctx->num_sms * 16, ((uint64_t)k.lanes * spp + 255) / 256));
                SpecArgs ca = a; ca.path_base = k.base;
                // with primary-hit reuse the first walk of the round runs over the pixels' representatives, queued in q[1] (free until shade 0 writes it)
                uint32_t* q_rep = reuse ? k.q[1] : nullptr; uint32_t* c_rep = reuse ? cnt + (RB - 4) : nullptr;
                if (r == 0) k_spec_gen<true><<<gen_blocks, 256, 0, k.st>>>(ca, nullptr, nullptr, k.first, k.lanes, k.q[0], cnt, k.base, q_rep, c_rep);
                else k_spec_gen<false><<<gen_blocks, 256, 0, k.st>>>(ca, k.pend[(r - 1) & 1], round_counters(k, r - 1) + (RB - 2), 0u, 0u, k.q[0], cnt, k.base, q_rep, c_rep);
                launches++;
            }
            for (uint32_t it = 0; it < bounces; it++) {
                for (uint32_t c = 0; c < nch; c++) {
                    const Chain& k = ch[c];
                    uint32_t* cnt = round_counters(k, r);
                    const int trace_blocks = r ? ctx->num_sms : full_trace;
                    const int shade_blocks = r ? ctx->num_sms : (int)std::max<uint64_t>(1, std::min<uint64_t>((uint64_t)full_shade, (k.cap_paths + 255) / 256));
                    if (profile && r == 0) RT2_CUDA(ctx, cudaEventRecord(ctx->prof_events[3 * it], k.st));
                    t.queue = k.q[it & 1]; t.count = cnt + 2 * it; t.head = cnt + 2 * it + 1;
                    if (reuse && it == 0) { t.queue = k.q[1]; t.count = cnt + (RB - 4); t.head = cnt + (RB - 3); }
                    t.stat_rays = stat_walks;
                    launch_trace_kernel(ctx, t, flat, trace_blocks, k.st);
                    if (profile && r == 0) RT2_CUDA(ctx, cudaEventRecord(ctx->prof_events[3 * it + 1], k.st));
                    k_spec_shade<<<shade_blocks, 256, 0, k.st>>>(a, k.q[it & 1], cnt + 2 * it, k.q[(it + 1) & 1], cnt + 2 * (it + 1), reuse && it == 0 ? 1u : 0u);
                    if (profile && r == 0) RT2_CUDA(ctx, cudaEventRecord(ctx->prof_events[3 * it + 2], k.st));
                    launches += 2;
                }
            }
            for (uint32_t c = 0; c < nch; c++) {
                const Chain& k = ch[c];
                uint32_t* cnt = round_counters(k, r);
                if (r == 0) k_spec_reduce<true><<<std::max(1, std::min(ctx->num_sms * 8, (int)((k.lanes + 255) / 256))), 256, 0, k.st>>>(a, nullptr, nullptr, k.first, k.lanes, k.pend[0], cnt + (RB - 2));
                else k_spec_reduce<false><<<ctx->num_sms, 256, 0, k.st>>>(a, k.pend[(r - 1) & 1], round_counters(k, r - 1) + (RB - 2), 0u, 0u, k.pend[r & 1], cnt + (RB - 2));
                launches++;
            }
        }
        // candidate round over what is still pending; the chain's path region is free again and holds cap_paths / K pixels' candidates.
        // Pixels beyond that (a frame where speculation was the wrong choice) finish sequentially in one persistent launch.
        ShadeArgs sa{};
        sa.band = band; sa.g = ctx->d_gconst; sa.frame = frame; sa.spp = spp; sa.max_bounces = bounces; sa.flags = lp.flags;
        sa.sc = a.sc;
        sa.st_ray0 = f.st_ray0; sa.st_ray1 = f.st_ray1; sa.st_acc = f.st_acc; sa.st_misc = f.st_misc; sa.st_hit = f.st_hit; sa.st_hit_inst = f.st_hit_inst;
        sa.diffuse_f16 = f.diffuse_f16; sa.specular_f16 = f.specular_f16; sa.diffuse_f32 = f.diffuse_f32;
        sa.stat_hits = reinterpret_cast<unsigned long long*>(f.counters + 2);
        sa.stat_rays = reinterpret_cast<unsigned long long*>(f.counters);
        sa.st_hit0 = f.st_hit0; sa.st_hit_inst0 = f.st_hit_inst0; sa.primary_reuse = reuse ? 1u : 0u;
        TraceArgs ts = make_trace_args(ctx);
        ts.stat_rays = stat_walks;
        const uint32_t K = bounces * (spp - 1u) + 1u;
        for (uint32_t c = 0; c < nch; c++) {
            const Chain& k = ch[c];
            const uint32_t* pend_list = k.pend[(rounds - 1) & 1];
            const uint32_t* pend_count = round_counters(k, rounds - 1) + (RB - 2);
            uint32_t* overflow_list = k.pend[rounds & 1];
            uint32_t* ccnt = round_counters(k, rounds);
            const uint32_t max_pixels = (uint32_t)std::min<size_t>(k.cap_paths / K, 0x7FFFFFFFu / K);
            SpecArgs ca = a;
            ca.cand_list = pend_list; ca.cand_K = K; ca.path_base = k.base;
            k_spec_gen_cand<<<ctx->num_sms * 2, 256, 0, k.st>>>(ca, pend_list, pend_count, max_pixels, k.q[0], ccnt, reuse ? k.q[1] : nullptr, reuse ? ccnt + (RB - 4) : nullptr);
            launches++;
            for (uint32_t it = 0; it < bounces; it++) {
                t.queue = k.q[it & 1]; t.count = ccnt + 2 * it; t.head = ccnt + 2 * it + 1;
                if (reuse && it == 0) { t.queue = k.q[1]; t.count = ccnt + (RB - 4); t.head = ccnt + (RB - 3); }
                t.stat_rays = stat_walks;
                launch_trace_kernel(ctx, t, flat, ctx->num_sms * 2, k.st);
                k_spec_shade<<<ctx->num_sms, 256, 0, k.st>>>(ca, k.q[it & 1], ccnt + 2 * it, k.q[(it + 1) & 1], ccnt + 2 * (it + 1), reuse && it == 0 ? 1u : 0u);
                launches += 2;
            }
            k_spec_reduce_cand<<<ctx->num_sms, 256, 0, k.st>>>(ca, pend_list, pend_count, max_pixels, overflow_list, ccnt + (RB - 2));
            launches++;
            WarpPathArgs w{f.warp_pop + (size_t)c * fb_blocks * 4u * fb_P, fb_P, k.cnt + 2, 0u, 0u, overflow_list, ccnt + (RB - 2), sp.px_prog, sp.px_rad};
            if (flat) k_path_warp<false, true><<<fb_blocks, 128, 0, k.st>>>(ts, sa, w);
            else if (s.num_instances) k_path_warp<true><<<fb_blocks, 128, 0, k.st>>>(ts, sa, w);
            else k_path_warp<false><<<fb_blocks, 128, 0, k.st>>>(ts, sa, w);
            launches++;
        }
        for (uint32_t c = 1; c < nch; c++) { RT2_CUDA(ctx, cudaEventRecord(ctx->chain_join[c], ch[c].st)); RT2_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->chain_join[c], 0)); }
    }
    RT2_CUDA(ctx, cudaGetLastError());
    ctx->stats.kernel_launches = launches;
    ctx->stats.wavefront_iterations = (uint64_t)(rounds + 1u) * bounces;
    ctx->stats_from_counters = true;
    ctx->pending_rays_static = 0;
    return RT2_OK;
}

// K1 entry: picks the sample schedule, then leaves the frame's (rays, hits) where the next call can read them without a sync
int render_reference(rt2_ctx* ctx, uint32_t frame, const rt2_launch_params& lp) {
    int rc = check_ready(ctx); if (rc) return rc;
    if ((rc = check_band(ctx, lp))) return rc;
    const uint32_t spp = lp.spp ? lp.spp : 12u;                    // refrence.rgen:17
    const uint32_t bounces = lp.max_bounces ? lp.max_bounces : 5u; // refrence.rgen:16
    if (spp > 65535u || bounces > 65535u || (uint64_t)spp * bounces > (1u << 22)) return fail(ctx, RT2_E_INVALID, "spp x bounces too large");
    const Band band = make_band(ctx->fb.width, ctx->fb.height, lp);
    // auto: speculate when fewer than 2 % of the previous frame's samples ended on a miss (each costs its pixel a fix-up round)
    if (ctx->k1_stats_pending && cudaEventQuery(ctx->k1_stats_ready) == cudaSuccess) {
        ctx->k1_stats_pending = false;
        const unsigned long long rays = ctx->h_k1_stats[0], hits = ctx->h_k1_stats[1];
        if (ctx->k1_stats_paths) {
            ctx->spec_auto = (double)(rays - hits) < 0.02 * (double)ctx->k1_stats_paths;
            ctx->spec_auto_few_misses = (double)(rays - hits) < 0.001 * (double)ctx->k1_stats_paths;
        }
    }
    cudaGetLastError();
    const bool want = ctx->speculate > 0 || (ctx->speculate < 0 && ctx->spec_auto);
    const bool spec = want && ctx->path_mode == 0 && spec_eligible(ctx, band, spp, bounces);
    rc = spec ? render_reference_spec(ctx, frame, lp, spp, bounces) : render_reference_seq(ctx, frame, lp);
    if (rc != RT2_OK) return rc;
    ctx->stats.sample_schedule = spec ? 1u : 0u;
    if (ctx->speculate < 0) {
        if (!ctx->h_k1_stats) {
            RT2_CUDA(ctx, cudaHostAlloc((void**)&ctx->h_k1_stats, 32, cudaHostAllocDefault));
            RT2_CUDA(ctx, cudaEventCreateWithFlags(&ctx->k1_stats_ready, cudaEventDisableTiming));
        }
        if (!ctx->k1_stats_pending) {          // one read-back in flight: the pinned words are rewritten only after the host consumed them
            RT2_CUDA(ctx, cudaMemcpyAsync(ctx->h_k1_stats, ctx->fb.counters, 16, cudaMemcpyDeviceToHost, ctx->stream));
            RT2_CUDA(ctx, cudaEventRecord(ctx->k1_stats_ready, ctx->stream));
            ctx->k1_stats_pending = true;
            ctx->k1_stats_paths = (uint64_t)band.width * band.local_rows * spp;
        }
    }
    return RT2_OK;
}

// Traces a queue of GENERIC rays (st_ray0 = o|tmin, st_ray1 = d|tmax, slot = pixel-linear index) into st_hit / st_hit_inst:
// the ReSTIR passes hand their incoherent rays to the same postponing/refilling kernel the path tracer uses.
// any_hit: the rays are visibility tests — the consumer reads "hit or not" only — so a walk ends at its first hit (RT2_ANY_HIT=0: closest hit always)
int trace_queue_generic(rt2_ctx* ctx, const uint32_t* queue, const uint32_t* count, uint32_t* head, unsigned long long* stat_rays, bool any_hit) {
    const DeviceScene& s = ctx->scene;
    TraceArgs t = make_trace_args(ctx);
    t.queue = queue; t.count = count; t.head = head; t.stat_rays = stat_rays;
    any_hit = any_hit && ctx->any_hit;
    if (t.flat_nodes) {
        const int blocks = ctx->trace_blocks_per_sm[3] * ctx->num_sms;
        if (any_hit) k_trace_dyn_flat<RT2_FLAT_PARAMS, true, true><<<blocks, 128, 0, ctx->stream>>>(t);
        else k_trace_dyn_flat<RT2_FLAT_PARAMS, true><<<blocks, 128, 0, ctx->stream>>>(t);
    } else if (s.num_instances) {
        const int blocks = ctx->trace_blocks_per_sm[1] * ctx->num_sms;
        if (!unified_walk_fits(s)) k_trace<true, true><<<ctx->trace_blocks_per_sm[2] * ctx->num_sms, 128, 0, ctx->stream>>>(t);
        else if (any_hit) k_trace_dyn_tlas<20, 4, 8, true, true><<<blocks, 128, 0, ctx->stream>>>(t);
        else k_trace_dyn_tlas<20, 4, 8, true><<<blocks, 128, 0, ctx->stream>>>(t);
    } else {
        const int blocks = ctx->trace_queue_blocks_per_sm * ctx->num_sms;
        if (any_hit) k_trace_queue<true><<<blocks, 128, 0, ctx->stream>>>(t);
        else k_trace_queue<false><<<blocks, 128, 0, ctx->stream>>>(t);
    }
    RT2_CUDA(ctx, cudaGetLastError());
    return RT2_OK;
}

int render_init(rt2_ctx* ctx) {
    int b0 = 0, b1 = 0;
    ctx->trace_variant = RT2_DEFAULT_TRACE_VARIANT;
    if (const char* e = getenv("RT2_GI_MEGAKERNEL")) ctx->gi_megakernels = atoi(e) != 0;
    if (const char* e = getenv("RT2_RAYS_PER_LANE")) { int v = atoi(e); if (v >= 0 && v <= 64) { ctx->rays_per_lane = (uint32_t)v; ctx->rays_per_lane_set = true; } }
    if (const char* e = getenv("RT2_CHAINS")) { int v = atoi(e); if (v >= 1) ctx->chains = (uint32_t)v; }
    if (const char* e = getenv("RT2_TRACE_VARIANT")) { int v = atoi(e); if (v >= 0 && v < g_num_trace_variants) ctx->trace_variant = v; }
    RT2_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b0, g_trace_variants[ctx->trace_variant], 128, 0));
    int b2 = 0;
    RT2_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b1, k_trace_dyn_tlas<20, 4, 8>, 128, 0));
    RT2_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b2, k_trace<true>, 128, 0));
    if (const char* e = getenv("RT2_SPECULATE")) ctx->speculate = atoi(e);
    if (const char* e = getenv("RT2_SPEC_ROUNDS")) { int v = atoi(e); if (v >= 0 && v <= 16) ctx->spec_rounds = v; }
    if (const char* e = getenv("RT2_PATH_MODE")) ctx->path_mode = atoi(e);
    if (const char* e = getenv("RT2_PRIMARY_REUSE")) ctx->primary_reuse = atoi(e) != 0;
    if (const char* e = getenv("RT2_VISIBILITY_REUSE")) ctx->visibility_reuse = atoi(e) != 0;
    if (const char* e = getenv("RT2_ANY_HIT")) ctx->any_hit = atoi(e) != 0;
    if (const char* e = getenv("RT2_SPARSE_CONTRIB")) ctx->sparse_contrib = atoi(e) != 0;
    if (const char* e = getenv("RT2_WARP_P")) { int v = atoi(e); if (v >= 32 && v <= 4096) ctx->warp_paths = (uint32_t)v; }
    if (const char* e = getenv("RT2_WARP_BLOCKS")) { int v = atoi(e); if (v >= 1 && v <= 16) ctx->warp_blocks_per_sm = v; }
    int p0 = 0, p1 = 0, p2 = 0, b3 = 0;
    RT2_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&p0, k_path_warp<false>, 128, 0));
    RT2_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&p1, k_path_warp<true>, 128, 0));
    RT2_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&p2, k_path_warp<false, true>, 128, 0));
    RT2_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b3, k_trace_dyn_flat<RT2_FLAT_PARAMS>, 128, 0));
    ctx->path_blocks_per_sm[0] = std::max(p0, 1); ctx->path_blocks_per_sm[1] = std::max(p1, 1); ctx->path_blocks_per_sm[2] = std::max(p2, 1);
    ctx->trace_blocks_per_sm[3] = std::max(b3, 1);
    { int bq = 0; RT2_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bq, k_trace_queue<false>, 128, 0)); ctx->trace_queue_blocks_per_sm = std::max(bq, 1); }
    if (const char* e = getenv("RT2_USE_FLAT")) ctx->use_flat = atoi(e) != 0;
    ctx->trace_blocks_per_sm[0] = std::max(b0, 1); ctx->trace_blocks_per_sm[1] = std::max(b1, 1); ctx->trace_blocks_per_sm[2] = std::max(b2, 1);
    return RT2_OK;
}

}  // namespace rt2
