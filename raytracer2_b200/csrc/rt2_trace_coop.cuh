// rt2_trace_coop.cuh — ray-queue traversal with EIGHT LANES PER RAY (one lane per child slot of the 8-wide node).
//
// The lane-per-ray walk (trace_dyn_body, rt2_render.cu) tests a node's eight boxes one after the other in one lane: ~277 instructions
// per node step, 12-20 dependent steps per ray.  That is the right shape when a launch has many rays per lane (the postponing / refill
// logic keeps the warps full), and the wrong one when it has one or two — a rank's share of a strong-scaled frame, or any launch
// whose rays are long and few: the launch then lasts as long as its slowest lane, i.e. one ray's dependent chain.  Here a group of
// eight lanes walks ONE ray: every lane dequantises and slab-tests its own child box, the hit masks are combined with one
// __reduce_or_sync over the group, the leaf triangles of a node are tested eight at a time (one per lane) and the closest is found
// with __reduce_min_sync.  A node step is ~80 instructions per lane and runs at full SIMT efficiency inside the group, so a ray's
// latency drops by ~4x at ~1.3x the lane-instructions per ray.  Four rays per warp; the per-group traversal stack lives in shared
// memory.  Same node encoding, same conservative box arithmetic (un-folded form), same watertight triangle test and the same total
// order on hits as the lane-per-ray walk, so the results are the same bits.
#pragma once
// (included by rt2_render.cu inside namespace rt2, after TraceArgs and the lane-per-ray kernels)

#define RT2_COOP_WARPS 4      /* warps per block */

template <bool FLAT, bool GENERIC>
__device__ __forceinline__ void trace_coop_body(const TraceArgs& a) {
    __shared__ uint2 s_stack[RT2_COOP_WARPS][4][RT2_STACK_SIZE];
    const uint32_t lane = threadIdx.x & 31u, sub = lane & 7u, grp = lane >> 3;
    const uint32_t gmask = 0xFFu << (grp * 8u);
    uint2* stack = s_stack[threadIdx.x >> 5][grp];
    const Node8* __restrict__ nodes = FLAT ? a.flat_nodes : a.blas_nodes;
    const uint32_t count = *a.count;
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(a.stat_rays, (unsigned long long)count);

    bool active = false, exhausted = false;
    uint32_t slot = 0, tri_slot = 0xFFFFFFFFu;
    int sp = 0;
    float tmin = RT2_K1_TMIN, tmax = RT2_BACKGROUND_DEPTH;
    V3 ro = v3(0.f, 0.f, 0.f), rd = v3(1.f, 1.f, 1.f);
    RayPrep rp = make_ray_prep(ro, rd);
    RayShear sh = make_shear(rd);
    Hit best; best.t = INFINITY; best.u = best.v = 0.f; best.inst = best.geom = best.prim = 0xFFFFFFFFu;
    uint2 ngroup = make_uint2(0u, 0u), tgroup = make_uint2(0u, 0u);
    // which word / byte of a node's 8-byte planes this lane's child slot lives in
    const bool hi_word = sub >= 4u;
    const uint32_t bsel = sub & 3u;

    for (;;) {
        // ---- refill: every idle group takes the next ray of the queue (one atomic per warp)
        const uint32_t idle = __ballot_sync(0xFFFFFFFFu, !active) & 0x01010101u;          // bit 8g = group g is idle
        if (idle != 0u && !exhausted) {
            const uint32_t n = __popc(idle);
            uint32_t base = 0;
            if (lane == 0) base = atomicAdd(a.head, n);
            base = __shfl_sync(0xFFFFFFFFu, base, 0);
            if (base + n >= count) exhausted = true;
            if (!active) {
                const uint32_t idx = base + __popc(idle & ((1u << (grp * 8u)) - 1u));
                if (idx < count) {
                    slot = a.queue[idx];
                    const float4 r0 = a.st_ray0[slot], r1 = a.st_ray1[slot];
                    ro = v3(r0.x, r0.y, r0.z); rd = GENERIC ? v3(r1.x, r1.y, r1.z) : v3(r0.w, r1.x, r1.y);
                    tmin = GENERIC ? r0.w : RT2_K1_TMIN; tmax = GENERIC ? r1.w : RT2_BACKGROUND_DEPTH;
                    rp = make_ray_prep(ro, rd);
                    if (!FLAT) sh = make_shear(rd);
                    best.t = INFINITY; best.u = best.v = 0.f; best.inst = best.geom = best.prim = 0xFFFFFFFFu;
                    tri_slot = 0xFFFFFFFFu;
                    ngroup = make_uint2(0u, 0x80000000u); tgroup = make_uint2(0u, 0u);
                    sp = 0;
                    active = true;
                }
            }
        }
        if (__ballot_sync(0xFFFFFFFFu, active) == 0u) break;                               // nothing in flight and the queue is empty

        // ---- triangle step: up to eight leaf primitives of the current node, one per lane
        const bool do_tri = active && tgroup.y != 0u;
        if (__any_sync(0xFFFFFFFFu, do_tri)) {
            if (do_tri) {
                const uint32_t T = tgroup.y;
                const uint32_t ti = __fns(T, 0u, (int)sub + 1);                            // position of this lane's set bit, ~0u if there is none
                const uint32_t p8 = __fns(T, 0u, 8);
                tgroup.y = p8 == 0xFFFFFFFFu ? 0u : (T & ~((2u << p8) - 1u));              // the eight lowest set bits are taken
                float t = INFINITY, u = 0.f, v = 0.f;
                uint32_t hinst = 0xFFFFFFFFu, geom = 0xFFFFFFFFu, prim = 0xFFFFFFFFu, s = 0xFFFFFFFFu;
                if (ti != 0xFFFFFFFFu) {
                    s = tgroup.x + ti;
                    V3 oo = ro;
                    RayShear lsh = sh;
                    uint32_t inst = 0u;
                    if (FLAT) {
                        const uint2 lf = __ldg(reinterpret_cast<const uint2*>(a.flat_leaf) + s);
                        s = lf.y; inst = lf.x;
                        const InstanceRec* ir = a.instances + inst;
                        float m[12];
#pragma unroll
                        for (int k = 0; k < 3; k++) { const float4 r = ld16f(ir->inv + 4 * k); m[4 * k] = r.x; m[4 * k + 1] = r.y; m[4 * k + 2] = r.z; m[4 * k + 3] = r.w; }
                        oo = xform_point(m, ro);
                        lsh = make_shear(xform_vector(m, rd));
                    }
                    const TriPos* tp = a.tris + s;
                    const float4 va = ld16f(&tp->v0), vb = ld16f(&tp->v1), vc = ld16f(&tp->v2);
                    float tt, uu, vv;
                    if (ray_triangle(oo, lsh, tmin, tmax, v3(va.x, va.y, va.z), v3(vb.x, vb.y, vb.z), v3(vc.x, vc.y, vc.z), &tt, &uu, &vv)) {
                        t = tt; u = uu; v = vv; hinst = inst; geom = f2u(va.w); prim = f2u(vb.w);
                    }
                }
                // the closest of the eight under the same total order hit_closer applies: (t, inst, geom, prim)
                const uint32_t tb = __reduce_min_sync(gmask, __float_as_uint(t));          // t >= 0 or +inf: the bit pattern orders like the value
                if (tb != 0x7F800000u) {
                    bool cand = __float_as_uint(t) == tb;
                    const uint32_t mi = __reduce_min_sync(gmask, cand ? hinst : 0xFFFFFFFFu); cand = cand && hinst == mi;
                    const uint32_t mg = __reduce_min_sync(gmask, cand ? geom : 0xFFFFFFFFu); cand = cand && geom == mg;
                    const uint32_t mp = __reduce_min_sync(gmask, cand ? prim : 0xFFFFFFFFu); cand = cand && prim == mp;
                    const uint32_t winner = __ffs(__ballot_sync(gmask, cand) & gmask) - 1u;
                    const float wt = __uint_as_float(tb);
                    const float wu = __shfl_sync(gmask, u, winner), wv = __shfl_sync(gmask, v, winner);
                    const uint32_t ws = __shfl_sync(gmask, s, winner);
                    if (hit_closer(wt, mi, mg, mp, best)) {
                        best.t = wt; best.u = wu; best.v = wv; best.inst = mi; best.geom = mg; best.prim = mp;
                        tri_slot = ws;
                    }
                }
            }
        }

        // ---- node step: pop the nearest inner child of the current group, every lane tests one of its eight boxes
        const bool do_node = active && tgroup.y == 0u && ngroup.y > 0x00FFFFFFu;
        if (__any_sync(0xFFFFFFFFu, do_node)) {
            if (do_node) {
                const uint32_t hits = ngroup.y;
                const uint32_t imask = hits & 0xFFu;
                const uint32_t bit = rt2_fhs(hits);
                ngroup.y &= ~(1u << bit);
                if (ngroup.y > 0x00FFFFFFu) { if (sub == 0u && sp < RT2_STACK_SIZE) stack[sp] = ngroup; if (sp < RT2_STACK_SIZE) sp++; }
                const uint32_t cslot = (bit - 24u) ^ rp.oct_inv;
                const uint32_t rel = rt2_popc(imask & ~(0xFFFFFFFFu << cslot));
                const Node8* np = nodes + (ngroup.x + rel);
                const uint4 n0 = ld16(np), n1 = ld16(reinterpret_cast<const uint4*>(np) + 1);
                const uint4 n2 = ld16(reinterpret_cast<const uint4*>(np) + 2), n3 = ld16(reinterpret_cast<const uint4*>(np) + 3);
                const uint4 n4 = ld16(reinterpret_cast<const uint4*>(np) + 4);
                const float sx = u2f(rt2_byte(n0.w, 0) << 23), sy = u2f(rt2_byte(n0.w, 1) << 23), sz = u2f(rt2_byte(n0.w, 2) << 23);
                const float ax = u2f(n0.x) - rp.ox, ay = u2f(n0.y) - rp.oy, az = u2f(n0.z) - rp.oz;
                const uint32_t meta = rt2_byte(hi_word ? n1.w : n1.z, bsel);
                const float qlx = (float)rt2_byte(hi_word ? n2.y : n2.x, bsel), qly = (float)rt2_byte(hi_word ? n2.w : n2.z, bsel), qlz = (float)rt2_byte(hi_word ? n3.y : n3.x, bsel);
                const float qhx = (float)rt2_byte(hi_word ? n3.w : n3.z, bsel), qhy = (float)rt2_byte(hi_word ? n4.y : n4.x, bsel), qhz = (float)rt2_byte(hi_word ? n4.w : n4.z, bsel);
                // t = q * (s * idir) + (p - o) * idir, near planes with idir (1 - eps), far planes with idir (1 + eps): cwbvh_node_step's un-folded form
                const float tnx = rt2_fma(rp.negx ? qhx : qlx, sx * rp.inx, ax * rp.inx), tfx = rt2_fma(rp.negx ? qlx : qhx, sx * rp.ifx, ax * rp.ifx);
                const float tny = rt2_fma(rp.negy ? qhy : qly, sy * rp.iny, ay * rp.iny), tfy = rt2_fma(rp.negy ? qly : qhy, sy * rp.ify, ay * rp.ify);
                const float tnz = rt2_fma(rp.negz ? qhz : qlz, sz * rp.inz, az * rp.inz), tfz = rt2_fma(rp.negz ? qlz : qhz, sz * rp.ifz, az * rp.ifz);
                const float tn = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, tmin));
                const float tf = fminf(fminf(tfx, tfy), fminf(tfz, fminf(best.t, tmax)));
                const bool is_inner = (meta & (meta << 1) & 0x10u) != 0u;                  // bits 4 and 3 set: an inner child (index 24 + slot)
                const uint32_t bit_index = (meta ^ (is_inner ? rp.oct_inv : 0u)) & 0x1Fu;
                const uint32_t contrib = (tn <= tf) ? (((meta >> 5) & 7u) << bit_index) : 0u;
                const uint32_t hitmask = __reduce_or_sync(gmask, contrib);
                ngroup.x = n1.x;
                ngroup.y = (hitmask & 0xFF000000u) | (n0.w >> 24);
                tgroup.x = n1.y;
                tgroup.y = hitmask & 0x00FFFFFFu;
            }
        }
        __syncwarp();                                                                      // stack entries written by lane 0 of a group are read by all of it

        // ---- nothing left in the current groups: pop, or this ray is done
        if (active && tgroup.y == 0u && ngroup.y <= 0x00FFFFFFu) {
            if (sp == 0) {
                if (sub == 0u) {
                    a.st_hit[slot] = make_float4(best.t, best.u, best.v, __uint_as_float(tri_slot));
                    if (FLAT) a.st_hit_inst[slot] = best.inst;
                }
                active = false;
            } else {
                ngroup = stack[--sp];
            }
        }
    }
}

#ifndef RT2_COOP_MIN_BLOCKS
#define RT2_COOP_MIN_BLOCKS 6
#endif
template <bool FLAT, bool GENERIC>
__global__ void __launch_bounds__(32 * RT2_COOP_WARPS, RT2_COOP_MIN_BLOCKS) k_trace_coop(TraceArgs a) { trace_coop_body<FLAT, GENERIC>(a); }

