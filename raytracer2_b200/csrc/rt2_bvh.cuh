// rt2_bvh.cuh — compressed 8-wide BVH (CWBVH, after Ylitie/Karras/Laine 2017) node format and
// stack traversal.  Replaces the driver BLAS/TLAS + traceRayEXT of the reference
// (src/model.rs:379-459, src/main.rs:153-191, RtxdiApplicationBridge.glsl:74-81).
//
// 80-byte node, five 16-byte words:
//   w0: p.x p.y p.z (fp32 grid origin) | ex ey ez imask (4 x u8; grid step = 2^e, imask bit s = slot s is an inner child)
//   w1: child_base, tri_base, meta[8]   (meta: 0 empty | inner 001sssss with sssss = 24+slot | leaf ccc ooooo:
//                                         ccc = unary triangle count, ooooo = first triangle relative to tri_base)
//   w2: qlo.x[8] qlo.y[8]   w3: qlo.z[8] qhi.x[8]   w4: qhi.y[8] qhi.z[8]     (u8 grid coordinates per child)
// Child box = [p + qlo*2^e, p + qhi*2^e].  The builder keeps >= 2^-10 grid steps of slack around
// every child box and the slab test widens t by (1 +- 2^-18), so a ray the watertight triangle
// test accepts is never culled by a quantised box (bit-exact hit ids vs brute force).
//
// Leaf triangles are stored pre-transformed (node matrix M*v, src/model.rs:193-212) as three
// float4: (v0, geometryIndex) (v1, primitiveId) (v2, 0); ids ride in the .w lanes.
#pragma once
#include "rt2_math.cuh"

namespace rt2 {

struct __align__(16) Node8 {
    float px, py, pz;
    uint8_t ex, ey, ez, imask;
    uint32_t child_base, tri_base;
    uint8_t meta[8];
    uint8_t qlox[8], qloy[8];
    uint8_t qloz[8], qhix[8];
    uint8_t qhiy[8], qhiz[8];
};
static_assert(sizeof(Node8) == 80, "CWBVH node must be 80 bytes");

struct __align__(16) TriPos { float4 v0, v1, v2; };      // 48 bytes
static_assert(sizeof(TriPos) == 48, "leaf triangle must be 48 bytes");

#define RT2_SLAB_EPS 3.814697265625e-06f   /* 2^-18 */
#define RT2_STACK_SIZE 48

RT2_HD uint32_t rt2_popc(uint32_t x) {
#ifdef __CUDA_ARCH__
    return (uint32_t)__popc(x);
#else
    return (uint32_t)__builtin_popcount(x);
#endif
}
RT2_HD uint32_t rt2_fhs(uint32_t x) {   // index of highest set bit (x != 0)
#ifdef __CUDA_ARCH__
    return 31u - (uint32_t)__clz((int)x);
#else
    return 31u - (uint32_t)__builtin_clz(x);
#endif
}
RT2_HD uint32_t rt2_byte(uint32_t w, uint32_t i) { return (w >> (8u * i)) & 0xFFu; }
RT2_HD uint4 ld16(const void* p) {
#ifdef __CUDA_ARCH__
    return __ldg(reinterpret_cast<const uint4*>(p));
#else
    return *reinterpret_cast<const uint4*>(p);
#endif
}
RT2_HD float4 ld16f(const void* p) {
#ifdef __CUDA_ARCH__
    return __ldg(reinterpret_cast<const float4*>(p));
#else
    return *reinterpret_cast<const float4*>(p);
#endif
}
// sign-extend each byte's msb over the byte: 0x80 -> 0xFF, else 0x00
RT2_HD uint32_t sign_extend_s8x4(uint32_t x) {
#ifdef __CUDA_ARCH__
    // prmt selector nibbles 8..B = replicate the sign bit of byte 0..3 (__byte_perm masks that mode away)
    uint32_t r;
    asm("prmt.b32 %0, %1, 0, 0xBA98;" : "=r"(r) : "r"(x));
    return r;
#else
    uint32_t r = 0;
    for (int i = 0; i < 4; i++) if (x & (0x80u << (8 * i))) r |= 0xFFu << (8 * i);
    return r;
#endif
}
// u8 -> f32.  I2F.U8 issues to the XU pipe (16 lanes/clk/SM, 63 % busy in profiles/r01b).  The prmt variant drops the
// byte into the mantissa of 2^23 and subtracts 2^23 (exact, bit-identical) at the price of one more issue slot per
// value; measured on B200 it LOSES 3 % on C2 (2142 vs 2212 Mrays/s): the kernel is issue/latency bound before it is
// XU bound.  Kept behind RT2_BYTE_MAGIC for the record.
RT2_HD float byte_to_float(uint32_t w, uint32_t i) {
#if defined(__CUDA_ARCH__) && defined(RT2_BYTE_MAGIC)
    return u2f(__byte_perm(w, 0x4B000000u, 0x7540u | i)) - 8388608.0f;
#else
    return (float)rt2_byte(w, i);
#endif
}
// 2^23 + q through the byte permute unit (ALU pipe) instead of I2F.U8 (XU pipe); cwbvh_node_step folds the 2^23 into
// the plane offset of the planes it dequantises this way (RT2_NODE_FOLD_MASK).
RT2_HD float byte_to_float_biased(uint32_t w, uint32_t i) {
#ifdef __CUDA_ARCH__
    return u2f(__byte_perm(w, 0x4B000000u, 0x7540u | i));
#else
    return u2f(0x4B000000u | rt2_byte(w, i));
#endif
}
// Which of the six slab planes (bit 0..2 = near x,y,z; bit 3..5 = far x,y,z) are dequantised with the folded 2^23 form
// on the device.  0 = all through I2F.U8 (the oracle's arithmetic), 63 = none.  Measured on B200, C2 (gpurun A/B, Mrays/s):
// 0: 2318 | 9 (x near+far): 2346 | 5, 6: 2340 | 36, 7, 13, 3: 2322-2334 | 1: 2312 | 27: 2190 | 56: 2182 | 63: 2147.
#ifndef RT2_NODE_FOLD_MASK
#define RT2_NODE_FOLD_MASK 9
#endif

struct TraceRay {
    V3 o, d;
    float tmin, tmax;
};

// Per-ray slab constants: near planes use idir*(1-eps), far planes idir*(1+eps).
struct RayPrep {
    float ox, oy, oz;
    float inx, iny, inz, ifx, ify, ifz;
    uint32_t oct_inv;     // 7 - octant
    bool negx, negy, negz;
};
// 1/d with |1/d| capped at 2^60: a zero (or denormal-small) direction component would otherwise give inf and
// then NaN slab distances (0 * inf), which fminf/fmaxf drop — still conservative, but such a ray walks
// every node the other two axes admit (one ray in ~2^23 after a cosine bounce off an axis-aligned wall,
// costing milliseconds).  With the cap the slab test degenerates to "is the origin inside the slab".
RT2_HD V3 ray_inv_dir(const V3& d) {
    const float cap = 1.152921504606846976e18f;
    float idx = 1.0f / d.x, idy = 1.0f / d.y, idz = 1.0f / d.z;
    return v3(copysignf(fminf(fabsf(idx), cap), idx), copysignf(fminf(fabsf(idy), cap), idy), copysignf(fminf(fabsf(idz), cap), idz));
}
// The sign of 1/d is the sign of d (1/-0 = -inf), so the octant comes from the reciprocal alone: a kernel that keeps
// ray_inv_dir(d) rebuilds RayPrep with six multiplies instead of three divisions (the instanced walk does, per instance exit).
RT2_HD RayPrep ray_prep_from_inv(const V3& o, const V3& id) {
    RayPrep r;
    r.ox = o.x; r.oy = o.y; r.oz = o.z;
    r.inx = id.x * (1.0f - RT2_SLAB_EPS); r.iny = id.y * (1.0f - RT2_SLAB_EPS); r.inz = id.z * (1.0f - RT2_SLAB_EPS);
    r.ifx = id.x * (1.0f + RT2_SLAB_EPS); r.ify = id.y * (1.0f + RT2_SLAB_EPS); r.ifz = id.z * (1.0f + RT2_SLAB_EPS);
    r.negx = (f2u(id.x) >> 31) != 0; r.negy = (f2u(id.y) >> 31) != 0; r.negz = (f2u(id.z) >> 31) != 0;
    r.oct_inv = 7u - ((r.negx ? 1u : 0u) | (r.negy ? 2u : 0u) | (r.negz ? 4u : 0u));
    return r;
}
RT2_HD RayPrep make_ray_prep(const V3& o, const V3& d) { return ray_prep_from_inv(o, ray_inv_dir(d)); }

// One node step: pops the nearest-octant-first inner child of `ngroup` (which must have one),
// pushes the remainder on the stack, tests the child's 8 boxes and returns the new inner-hit group in
// `ngroup` and the triangle-hit group in `tgroup`.
RT2_HD void cwbvh_node_step(const Node8* __restrict__ nodes, const RayPrep& rp, float tmin, float tfar_limit,
                            uint2& ngroup, uint2& tgroup, uint2* stack, int& sp) {
    const uint32_t hits = ngroup.y;
    const uint32_t imask = hits & 0xFFu;
    const uint32_t bit = rt2_fhs(hits);
    ngroup.y &= ~(1u << bit);
    if (ngroup.y > 0x00FFFFFFu) { if (sp < RT2_STACK_SIZE) stack[sp++] = ngroup; }
    const uint32_t slot = (bit - 24u) ^ rp.oct_inv;
    const uint32_t rel = rt2_popc(imask & ~(0xFFFFFFFFu << slot));
    const Node8* np = nodes + (ngroup.x + rel);
    const uint4 n0 = ld16(np), n1 = ld16(reinterpret_cast<const uint4*>(np) + 1);
    const uint4 n2 = ld16(reinterpret_cast<const uint4*>(np) + 2), n3 = ld16(reinterpret_cast<const uint4*>(np) + 3);
    const uint4 n4 = ld16(reinterpret_cast<const uint4*>(np) + 4);
    const float sx = u2f(rt2_byte(n0.w, 0) << 23), sy = u2f(rt2_byte(n0.w, 1) << 23), sz = u2f(rt2_byte(n0.w, 2) << 23);
    const float ax = u2f(n0.x) - rp.ox, ay = u2f(n0.y) - rp.oy, az = u2f(n0.z) - rp.oz;
    // t = q * (s*idir) + (p-o)*idir ; s is a power of two so s*idir is exact
    const float snx = sx * rp.inx, sny = sy * rp.iny, snz = sz * rp.inz, sfx = sx * rp.ifx, sfy = sy * rp.ify, sfz = sz * rp.ifz;
    float anx = ax * rp.inx, any_ = ay * rp.iny, anz = az * rp.inz, afx = ax * rp.ifx, afy = ay * rp.ify, afz = az * rp.ifz;
#if defined(__CUDA_ARCH__)
    // Dequantisation split between the XU pipe (I2F.U8, 16 lanes/clk/SM: the busiest pipe of k_trace, 60 % in profiles/r01c)
    // and the ALU pipe: for the planes in RT2_NODE_FOLD_MASK, prmt drops the byte q into the mantissa of 2^23, giving the
    // float 2^23 + q, and the 2^23 is taken out of the per-node offset instead of out of every value:
    //     q*s + a == (2^23 + q)*s + (a - 2^23*s).
    // a - 2^23*s is rounded once (fma), to <= half a grid cell |s| — so the offset also moves a near plane one cell earlier
    // and a far plane one cell later (constants 2^23 +- 1, exact in binary32).  The box test stays conservative; for those
    // planes it is no longer the oracle's arithmetic, which is fine: the closest hit is a total order, independent of which
    // superfluous nodes are visited.  The CPU build of this header keeps the exact form.
    // All six planes folded measured -7 % on C2 (the 48 prmt land on the ALU pipe that FMNMX/LOP3/SEL already keep 46 % busy).
    if (RT2_NODE_FOLD_MASK & 1) anx = rt2_fma(snx, rp.negx ? -8388607.0f : -8388609.0f, anx);
    if (RT2_NODE_FOLD_MASK & 2) any_ = rt2_fma(sny, rp.negy ? -8388607.0f : -8388609.0f, any_);
    if (RT2_NODE_FOLD_MASK & 4) anz = rt2_fma(snz, rp.negz ? -8388607.0f : -8388609.0f, anz);
    if (RT2_NODE_FOLD_MASK & 8) afx = rt2_fma(sfx, rp.negx ? -8388609.0f : -8388607.0f, afx);
    if (RT2_NODE_FOLD_MASK & 16) afy = rt2_fma(sfy, rp.negy ? -8388609.0f : -8388607.0f, afy);
    if (RT2_NODE_FOLD_MASK & 32) afz = rt2_fma(sfz, rp.negz ? -8388609.0f : -8388607.0f, afz);
#define RT2_DQ(bit, w, j) ((RT2_NODE_FOLD_MASK & (bit)) ? byte_to_float_biased(w, j) : byte_to_float(w, j))
#else
#define RT2_DQ(bit, w, j) byte_to_float(w, j)
#endif
    const uint32_t oct_inv4 = rp.oct_inv * 0x01010101u;
    uint32_t hitmask = 0;
#pragma unroll
    for (int half = 0; half < 2; half++) {
        const uint32_t meta4 = half ? n1.w : n1.z;
        const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
        const uint32_t inner_mask4 = sign_extend_s8x4(is_inner4 << 3);
        const uint32_t bit_index4 = (meta4 ^ (oct_inv4 & inner_mask4)) & 0x1F1F1F1Fu;
        const uint32_t child_bits4 = (meta4 >> 5) & 0x07070707u;
        const uint32_t qlox = half ? n2.y : n2.x, qloy = half ? n2.w : n2.z, qloz = half ? n3.y : n3.x;
        const uint32_t qhix = half ? n3.w : n3.z, qhiy = half ? n4.y : n4.x, qhiz = half ? n4.w : n4.z;
        const uint32_t qnx = rp.negx ? qhix : qlox, qfx = rp.negx ? qlox : qhix;
        const uint32_t qny = rp.negy ? qhiy : qloy, qfy = rp.negy ? qloy : qhiy;
        const uint32_t qnz = rp.negz ? qhiz : qloz, qfz = rp.negz ? qloz : qhiz;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const float tnx = rt2_fma(RT2_DQ(1, qnx, j), snx, anx);
            const float tny = rt2_fma(RT2_DQ(2, qny, j), sny, any_);
            const float tnz = rt2_fma(RT2_DQ(4, qnz, j), snz, anz);
            const float tfx = rt2_fma(RT2_DQ(8, qfx, j), sfx, afx);
            const float tfy = rt2_fma(RT2_DQ(16, qfy, j), sfy, afy);
            const float tfz = rt2_fma(RT2_DQ(32, qfz, j), sfz, afz);
            const float tn = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, tmin));
            const float tf = fminf(fminf(tfx, tfy), fminf(tfz, tfar_limit));
            if (tn <= tf) hitmask |= rt2_byte(child_bits4, j) << rt2_byte(bit_index4, j);
        }
    }
#undef RT2_DQ
    ngroup.x = n1.x;
    ngroup.y = (hitmask & 0xFF000000u) | (n0.w >> 24);
    tgroup.x = n1.y;
    tgroup.y = hitmask & 0x00FFFFFFu;
}

// Generic CWBVH stack traversal.  `leaf(slot)` is called for every leaf primitive whose (padded)
// box the ray touches within [tmin, min(tmax, *t_limit)]; it may lower *t_limit (closest hit).
// `stop` (optional): the walk ends as soon as *stop is set by `leaf` — visibility tests that only ask "hit or not" (any-hit).
template <class LeafFn>
RT2_HD void cwbvh_traverse(const Node8* __restrict__ nodes, const TraceRay& ray, const float* t_limit, LeafFn&& leaf,
                           uint32_t* node_visits = nullptr, const bool* stop = nullptr) {
    const RayPrep rp = make_ray_prep(ray.o, ray.d);
    uint2 stack[RT2_STACK_SIZE];
    int sp = 0;
    uint2 ngroup = make_uint2(0u, 0x80000000u);

    for (;;) {
        uint2 tgroup;
        if (ngroup.y > 0x00FFFFFFu) {
            if (node_visits) (*node_visits)++;
            cwbvh_node_step(nodes, rp, ray.tmin, fminf(*t_limit, ray.tmax), ngroup, tgroup, stack, sp);
        } else {
            tgroup = ngroup;
            ngroup = make_uint2(0u, 0u);
        }
        while (tgroup.y != 0u) {
            const uint32_t ti = rt2_fhs(tgroup.y);
            tgroup.y &= ~(1u << ti);
            leaf(tgroup.x + ti);
            if (stop && *stop) return;
        }
        if (ngroup.y <= 0x00FFFFFFu) {
            if (sp == 0) break;
            ngroup = stack[--sp];
        }
    }
}

// Closest hit over one triangle CWBVH.  `inst` is written into the hit record (0 for the
// single-level case).  Caller initialises best.t = +inf and ids = ~0u.
// `any` (here and in the two functions below): stop at the first accepted hit — `best` then names some hit, not the closest.
RT2_HD void cwbvh_closest_hit(const Node8* __restrict__ nodes, const TriPos* __restrict__ tris,
                              const TraceRay& ray, uint32_t inst, Hit& best, uint32_t* slot_out, uint32_t* node_visits = nullptr, bool any = false) {
    const RayShear sh = make_shear(ray.d);
    bool found = false;
    cwbvh_traverse(nodes, ray, &best.t, [&](uint32_t slot) {
        const TriPos* tp = tris + slot;
        const float4 a = ld16f(&tp->v0), b = ld16f(&tp->v1), c = ld16f(&tp->v2);
        float t, u, v;
        if (ray_triangle(ray.o, sh, ray.tmin, ray.tmax, v3(a.x, a.y, a.z), v3(b.x, b.y, b.z), v3(c.x, c.y, c.z), &t, &u, &v)) {
            const uint32_t geom = f2u(a.w), prim = f2u(b.w);
            if (hit_closer(t, inst, geom, prim, best)) {
                best.t = t; best.u = u; best.v = v; best.inst = inst; best.geom = geom; best.prim = prim;
                if (slot_out) *slot_out = slot;
                found = true;
            }
        }
    }, node_visits, any ? &found : nullptr);
}

// Instance record of the TLAS (A2 extension): world->object 3x4, object->world 3x4. 96 bytes.
struct __align__(16) InstanceRec { float inv[12]; float fwd[12]; };

// Two-level closest hit: TLAS leaves are instance slots; the ray is re-expressed in object space
// (o' = inv*o, d' = inv*d, t unchanged) and the BLAS is traversed.
RT2_HD void tlas_closest_hit(const Node8* __restrict__ tlas_nodes, const uint32_t* __restrict__ tlas_slot_to_inst,
                             const InstanceRec* __restrict__ instances, const Node8* __restrict__ blas_nodes,
                             const TriPos* __restrict__ tris, const TraceRay& ray, Hit& best, uint32_t* slot_out,
                             uint32_t* node_visits = nullptr, bool any = false) {
    bool found = false;
    cwbvh_traverse(tlas_nodes, ray, &best.t, [&](uint32_t slot) {
        const uint32_t inst = tlas_slot_to_inst[slot];
        const InstanceRec* ir = instances + inst;
        float m[12];
#pragma unroll
        for (int k = 0; k < 3; k++) { float4 r = ld16f(ir->inv + 4 * k); m[4 * k] = r.x; m[4 * k + 1] = r.y; m[4 * k + 2] = r.z; m[4 * k + 3] = r.w; }
        TraceRay lr;
        lr.o = xform_point(m, ray.o); lr.d = xform_vector(m, ray.d); lr.tmin = ray.tmin; lr.tmax = ray.tmax;
        cwbvh_closest_hit(blas_nodes, tris, lr, inst, best, slot_out, node_visits, any);
        if (any && best.geom != 0xFFFFFFFFu) found = true;
    }, node_visits, any ? &found : nullptr);
}

// Flattened instancing (B200: 180 GB of HBM makes the memory a non-issue): ONE CWBVH over the world-space boxes of every
// (instance, triangle) pair, so a ray walks a single tree with the world-space ray — no instance entry / exit, no per-instance
// BLAS descent, lanes of a warp stay at comparable depths.  A leaf slot names its pair (FlatLeaf) and the triangle is tested in
// the instance's object space with the ray re-expressed exactly as the two-level walk does (o' = inv*o, d' = inv*d, t
// unchanged), so the accepted hits — and (t, u, v) — are the same; the closest hit is a total order, hence the same image.
// The object-space ray is cached per lane while consecutive leaves belong to the same instance.
struct FlatLeaf { uint32_t inst, tri; };      // tri = BLAS leaf slot (index into the object-space TriPos array)

RT2_HD void flat_closest_hit(const Node8* __restrict__ nodes, const FlatLeaf* __restrict__ leaves, const InstanceRec* __restrict__ instances,
                             const TriPos* __restrict__ tris, const TraceRay& ray, Hit& best, uint32_t* slot_out, uint32_t* node_visits = nullptr,
                             bool any = false) {
    uint32_t cur = 0xFFFFFFFFu;
    V3 oo = ray.o;
    RayShear sh = make_shear(ray.d);
    bool found = false;
    cwbvh_traverse(nodes, ray, &best.t, [&](uint32_t slot) {
        const FlatLeaf lf = leaves[slot];
        if (lf.inst != cur) {
            cur = lf.inst;
            const InstanceRec* ir = instances + cur;
            float m[12];
#pragma unroll
            for (int k = 0; k < 3; k++) { float4 r = ld16f(ir->inv + 4 * k); m[4 * k] = r.x; m[4 * k + 1] = r.y; m[4 * k + 2] = r.z; m[4 * k + 3] = r.w; }
            oo = xform_point(m, ray.o);
            sh = make_shear(xform_vector(m, ray.d));
        }
        const TriPos* tp = tris + lf.tri;
        const float4 a = ld16f(&tp->v0), b = ld16f(&tp->v1), c = ld16f(&tp->v2);
        float t, u, v;
        if (ray_triangle(oo, sh, ray.tmin, ray.tmax, v3(a.x, a.y, a.z), v3(b.x, b.y, b.z), v3(c.x, c.y, c.z), &t, &u, &v)) {
            const uint32_t geom = f2u(a.w), prim = f2u(b.w);
            if (hit_closer(t, cur, geom, prim, best)) {
                best.t = t; best.u = u; best.v = v; best.inst = cur; best.geom = geom; best.prim = prim;
                if (slot_out) *slot_out = lf.tri;
                found = true;
            }
        }
    }, node_visits, any ? &found : nullptr);
}

}  // namespace rt2
