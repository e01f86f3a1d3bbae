// rt2_build.cuh — per-element logic of the device BVH builder.
//
// Pipeline (all on the device; replaces vkCmdBuildAccelerationStructuresKHR, src/context.rs:824-911):
//   1. build_triangle      : indices + vertices + node matrix -> BLAS-space triangle, box, centroid
//   2. morton_key          : 63-bit Morton code of the centroid in the scene box
//   3. radix sort          : (rt2_build.cu)
//   4. PLOC                : parallel locally-ordered clustering (Meister & Bittner 2018) -> BVH2
//   5. collapse_node       : BVH2 -> 8-wide nodes (SAH-optimal slot distribution, 5a; <= 3 triangles per leaf child),
//                            octant-ordered slot assignment, conservative u8 quantisation -> Node8
// Each step is a __host__ __device__ function of one element so that tests/hostemu can run the same
// code serially on the CPU; the kernels in rt2_build.cu are thin grid-stride wrappers.
#pragma once
#include "rt2_bvh.cuh"
#include "../../include/rt2_abi.h"

namespace rt2 {

#ifdef __CUDA_ARCH__
#define RT2_ATOMIC_ADD(ptr, v) atomicAdd((ptr), (v))
#else
template <class T> inline T rt2_host_atomic_add(T* p, T v) { T o = *p; *p = o + v; return o; }
#define RT2_ATOMIC_ADD(ptr, v) rt2::rt2_host_atomic_add((ptr), (v))
#endif

struct BuildInput {
    const rt2_vertex* vertices;
    const uint32_t* indices;
    const rt2_geometry_info* geoms;
    const uint32_t* geom_tri_offset;   // exclusive scan of index_counts/3, num_geoms+1 entries
    uint32_t num_geoms;
    uint32_t num_tris;
};

// BVH2 node (PLOC output): leaves are ids < num_leaves
struct Bvh2 {
    float4* lo;        // xyz = box min, w = left child (uint bits)
    float4* hi;        // xyz = box max, w = right child (uint bits)
    uint32_t* count;   // triangles (or instances) below
};

RT2_HD uint32_t find_geometry(const uint32_t* offs, uint32_t n, uint32_t tri) {
    uint32_t lo = 0, hi = n;   // offs[lo] <= tri < offs[hi]
    while (hi - lo > 1) { uint32_t mid = (lo + hi) >> 1; if (offs[mid] <= tri) lo = mid; else hi = mid; }
    return lo;
}

// 1. triangle `i` in geometry order -> BLAS space (M*v with VkTransformMatrixKHR rows, pinned order)
RT2_HD void build_triangle(const BuildInput& in, uint32_t i, TriPos* out_tri, float4* out_lo, float4* out_hi) {
    const uint32_t g = find_geometry(in.geom_tri_offset, in.num_geoms, i);
    const uint32_t p = i - in.geom_tri_offset[g];
    const rt2_geometry_info& gi = in.geoms[g];
    const float* M = gi.transform;
    const float rows[12] = { M[0], M[4], M[8], M[12], M[1], M[5], M[9], M[13], M[2], M[6], M[10], M[14] };
    V3 P[3];
    for (int k = 0; k < 3; k++) {
        const rt2_vertex& v = in.vertices[gi.vertex_offset + in.indices[gi.index_offset + 3u * p + (uint32_t)k]];
        P[k] = xform_point(rows, v3(v.pos[0], v.pos[1], v.pos[2]));
    }
    TriPos t;
    t.v0 = make_float4(P[0].x, P[0].y, P[0].z, u2f(g));
    t.v1 = make_float4(P[1].x, P[1].y, P[1].z, u2f(p));
    t.v2 = make_float4(P[2].x, P[2].y, P[2].z, 0.0f);
    out_tri[i] = t;
    out_lo[i] = make_float4(fminf(P[0].x, fminf(P[1].x, P[2].x)), fminf(P[0].y, fminf(P[1].y, P[2].y)), fminf(P[0].z, fminf(P[1].z, P[2].z)), u2f(0xFFFFFFFFu));
    out_hi[i] = make_float4(fmaxf(P[0].x, fmaxf(P[1].x, P[2].x)), fmaxf(P[0].y, fmaxf(P[1].y, P[2].y)), fmaxf(P[0].z, fmaxf(P[1].z, P[2].z)), u2f(0xFFFFFFFFu));
}

RT2_HD uint64_t spread21(uint32_t v) {
    uint64_t x = v & 0x1FFFFFull;
    x = (x | (x << 32)) & 0x1F00000000FFFFull;
    x = (x | (x << 16)) & 0x1F0000FF0000FFull;
    x = (x | (x << 8)) & 0x100F00F00F00F00Full;
    x = (x | (x << 4)) & 0x10C30C30C30C30C3ull;
    x = (x | (x << 2)) & 0x1249249249249249ull;
    return x;
}
// 2. Morton key of a box centroid inside the scene box [slo, shi]
RT2_HD uint64_t morton_key(float4 lo, float4 hi, V3 slo, V3 shi) {
    float cx = 0.5f * (lo.x + hi.x), cy = 0.5f * (lo.y + hi.y), cz = 0.5f * (lo.z + hi.z);
    float ex = shi.x - slo.x, ey = shi.y - slo.y, ez = shi.z - slo.z;
    float nx = ex > 0.f ? (cx - slo.x) / ex : 0.f, ny = ey > 0.f ? (cy - slo.y) / ey : 0.f, nz = ez > 0.f ? (cz - slo.z) / ez : 0.f;
    const float S = 2097151.0f;
    uint32_t qx = (uint32_t)fminf(fmaxf(nx * S, 0.f), S), qy = (uint32_t)fminf(fmaxf(ny * S, 0.f), S), qz = (uint32_t)fminf(fmaxf(nz * S, 0.f), S);
    return (spread21(qx) << 2) | (spread21(qy) << 1) | spread21(qz);
}

RT2_HD float box_half_area(float lx, float ly, float lz, float hx, float hy, float hz) {
    float dx = hx - lx, dy = hy - ly, dz = hz - lz;
    return dx * dy + dy * dz + dz * dx;
}

// 4a. PLOC nearest-neighbour search for cluster position i (radius R). Distance = half area of the
// union box; ties are broken by (min index, max index) so the globally best pair is always mutual.
// Search radius: 16 for trees over instance-level boxes (TLAS, flattened instancing: heavily overlapping leaves), 8 for triangle
// BLASes — with the reinsertion rounds behind it a narrower search gives the better triangle tree (C2: radius 2 / 4 / 6 / 8 / 16 / 32 ->
// 2774 / 2796 / 2770 / 2805 / 2761 / 2738 Mrays/s; C5's flat tree: 955 / 998 / 1081 / 1016 / 1145 / -, profiles/r02b_trace_variant_sweep.txt)
#ifndef RT2_PLOC_RADIUS
#define RT2_PLOC_RADIUS 16
#endif
#ifndef RT2_PLOC_RADIUS_TRIANGLES
#define RT2_PLOC_RADIUS_TRIANGLES 8
#endif
RT2_HD uint32_t ploc_nearest(const Bvh2& b, const uint32_t* clusters, uint32_t n, uint32_t i, uint32_t radius = RT2_PLOC_RADIUS) {
    const uint32_t ci = clusters[i];
    const float4 li = b.lo[ci], hi_ = b.hi[ci];
    float best = INFINITY; uint32_t best_j = 0xFFFFFFFFu;
    uint32_t j0 = i > radius ? i - radius : 0u;
    uint32_t j1 = i + radius + 1u < n ? i + radius + 1u : n;
    for (uint32_t j = j0; j < j1; j++) {
        if (j == i) continue;
        const uint32_t cj = clusters[j];
        const float4 lj = b.lo[cj], hj = b.hi[cj];
        float a = box_half_area(fminf(li.x, lj.x), fminf(li.y, lj.y), fminf(li.z, lj.z), fmaxf(hi_.x, hj.x), fmaxf(hi_.y, hj.y), fmaxf(hi_.z, hj.z));
        bool better;
        if (best_j == 0xFFFFFFFFu || a < best) better = true;
        else if (a > best) better = false;
        else {
            // equal area: compare (min(i,j), max(i,j)) lexicographically
            uint32_t m0 = i < j ? i : j, M0 = i < j ? j : i;
            uint32_t m1 = i < best_j ? i : best_j, M1 = i < best_j ? best_j : i;
            better = (m0 < m1) || (m0 == m1 && M0 < M1);
        }
        if (better) { best = a; best_j = j; }
    }
    return best_j;
}

// 4b. PLOC merge for position i: returns 1 if position i stays valid (and may now hold a new node)
RT2_HD uint32_t ploc_merge(const Bvh2& b, uint32_t* clusters, const uint32_t* nearest, uint32_t n, uint32_t i, uint32_t* node_counter) {
    (void)n;
    const uint32_t j = nearest[i];
    if (j == 0xFFFFFFFFu || nearest[j] != i) return 1u;
    if (i > j) return 0u;
    const uint32_t ci = clusters[i], cj = clusters[j];
    const uint32_t id = RT2_ATOMIC_ADD(node_counter, 1u);
    const float4 li = b.lo[ci], hi_ = b.hi[ci], lj = b.lo[cj], hj = b.hi[cj];
    b.lo[id] = make_float4(fminf(li.x, lj.x), fminf(li.y, lj.y), fminf(li.z, lj.z), u2f(ci));
    b.hi[id] = make_float4(fmaxf(hi_.x, hj.x), fmaxf(hi_.y, hj.y), fmaxf(hi_.z, hj.z), u2f(cj));
    b.count[id] = b.count[ci] + b.count[cj];
    clusters[i] = id;
    return 1u;
}

// ---- 5. collapse -------------------------------------------------------------------------------------------
#ifndef RT2_LEAF_MAX
#define RT2_LEAF_MAX 3u
#endif

struct CollapseCtx {
    Bvh2 bvh2;
    uint32_t num_leaves;        // BVH2 leaf ids are [0, num_leaves)
    const uint32_t* leaf_prim;  // sorted position -> original primitive id
    Node8* nodes;               // output wide nodes
    uint32_t* node_counter;     // next free wide node
    uint32_t* prim_counter;     // next free output primitive slot
    uint32_t* out_prim;         // output slot -> original primitive id
    uint2* queue_out;           // (bvh2 id, wide index) of the next level
    uint32_t* queue_out_count;
    uint32_t max_nodes, max_prims;
    const uint32_t* dp_dec;     // decision words of the SAH-optimal collapse (5a); nullptr = open the largest child first
};

RT2_HD bool bvh2_is_leaf(const CollapseCtx& c, uint32_t id) { return id < c.num_leaves; }

// ---- 5a. SAH-optimal collapse (Ylitie, Karras, Laine 2017, "Efficient incoherent ray traversal on GPUs through compressed wide
// BVHs", §4.1), adapted to this node layout, where a subtree of <= RT2_LEAF_MAX primitives is a leaf SLOT of its parent:
//   C(n,1) = A(n) * count(n) * ct                 if count(n) <= RT2_LEAF_MAX      (a leaf slot)
//          = A(n) * cn + D(n,8)                   otherwise                         (a wide node of its own)
//   C(n,i) = min(C(n,i-1), D(n,i))   i = 2..7     n spread over up to i slots of an ancestor's wide node
//   D(n,j) = min_k C(left,k) + C(right,j-k)
// One bottom-up pass fills cost[node][7] and a decision word per node: bits 3(j-2)..+2 = argmin k - 1 of D(n,j), j = 2..8;
// bit 21+(i-2) = "C(n,i) is D(n,i)" for i = 2..7.  Measured on the C2 hall (tools/treelab.py): 40 % fewer wide nodes, 2-3 % fewer
// node steps per ray than opening the largest child first.
#ifndef RT2_DP_CN
#define RT2_DP_CN 1.0f
#endif
#ifndef RT2_DP_CT
#define RT2_DP_CT 0.3f
#endif
#ifdef __CUDA_ARCH__
#define RT2_LDCG_F(p) __ldcg(p)      /* written by another thread of the same launch: read past L1 */
#else
#define RT2_LDCG_F(p) (*(p))
#endif
struct CollapseDp { float* cost; uint32_t* dec; };

RT2_HD float bvh2_half_area(const Bvh2& b, uint32_t i) { const float4 l = b.lo[i], h = b.hi[i]; return box_half_area(l.x, l.y, l.z, h.x, h.y, h.z); }
RT2_HD void dp_solve_leaf(const CollapseDp& dp, const Bvh2& b, uint32_t n) {
    const float c = bvh2_half_area(b, n) * RT2_DP_CT;
    for (int i = 0; i < 7; i++) dp.cost[(size_t)n * 7 + i] = c;
    dp.dec[n] = 0u;
}
RT2_HD void dp_solve_inner(const CollapseDp& dp, const Bvh2& b, uint32_t n) {
    const uint32_t l = f2u(b.lo[n].w), r = f2u(b.hi[n].w);
    float cl[7], cr[7];
    for (int i = 0; i < 7; i++) { cl[i] = RT2_LDCG_F(dp.cost + (size_t)l * 7 + i); cr[i] = RT2_LDCG_F(dp.cost + (size_t)r * 7 + i); }
    float D[9]; uint32_t dec = 0u;
    for (int j = 2; j <= 8; j++) {
        float best = INFINITY; int bk = 1;
        for (int k = (j - 7 > 1 ? j - 7 : 1); k <= (j - 1 < 7 ? j - 1 : 7); k++) {
            const float v = cl[k - 1] + cr[j - k - 1];
            if (v < best) { best = v; bk = k; }
        }
        D[j] = best; dec |= (uint32_t)(bk - 1) << (3 * (j - 2));
    }
    const float A = bvh2_half_area(b, n);
    const uint32_t cnt = b.count[n];
    float prev = cnt <= RT2_LEAF_MAX ? A * (float)cnt * RT2_DP_CT : A * RT2_DP_CN + D[8];
    dp.cost[(size_t)n * 7] = prev;
    for (int i = 2; i <= 7; i++) {
        if (D[i] < prev) { prev = D[i]; dec |= 1u << (21 + (i - 2)); }
        dp.cost[(size_t)n * 7 + (i - 1)] = prev;
    }
    dp.dec[n] = dec;
}

// gather the (<= RT2_LEAF_MAX) primitives below `id`
RT2_HD uint32_t gather_leaves(const CollapseCtx& c, uint32_t id, uint32_t* out) {
    uint32_t st[8]; int sp = 0; uint32_t n = 0;
    st[sp++] = id;
    while (sp) {
        uint32_t x = st[--sp];
        if (bvh2_is_leaf(c, x)) { if (n < RT2_LEAF_MAX) out[n++] = x; }
        else { if (sp <= 6) { st[sp++] = f2u(c.bvh2.hi[x].w); st[sp++] = f2u(c.bvh2.lo[x].w); } }
    }
    return n;
}

// Conservative quantisation of one axis. Finds e and p such that every child interval [lo,hi] maps to
// u8 [qlo,qhi] with p + qlo*2^e <= lo - 2^-10*2^e and p + qhi*2^e >= hi + 2^-10*2^e.
RT2_HD void choose_grid(double lo, double hi, float* p_out, int* e_out) {
    double ext = hi - lo;
    double amax = fmax(fabs(lo), fabs(hi));
    double min_ext = fmax(amax * 9.5367431640625e-07 /* 2^-20 */, 1e-30);
    if (!(ext > min_ext)) ext = min_ext;
    int e;
    frexp(ext / 253.0, &e);              // 2^(e-1) <= ext/253 < 2^e  =>  253 * 2^e > ext
    if (e < -120) e = -120;
    if (e > 120) e = 120;
    for (;;) {
        double s = ldexp(1.0, e);
        float p = (float)(lo - s * (1.0 / 128.0));
        // float rounding may have moved p up: step it down until it is below lo - s/256
        while ((double)p > lo - s * (1.0 / 256.0)) p = u2f(p > 0.0f ? f2u(p) - 1u : (p < 0.0f ? f2u(p) + 1u : 0x80000001u));
        if (((hi - (double)p) / s) + (1.0 / 1024.0) <= 255.0) { *p_out = p; *e_out = e; return; }
        e++;
    }
}
RT2_HD uint32_t quant_lo(double v, float p, int e) {
    double q = floor((v - (double)p) / ldexp(1.0, e) - (1.0 / 1024.0));
    if (q < 0.0) q = 0.0;
    if (q > 255.0) q = 255.0;
    return (uint32_t)q;
}
RT2_HD uint32_t quant_hi(double v, float p, int e) {
    double q = ceil((v - (double)p) / ldexp(1.0, e) + (1.0 / 1024.0));
    if (q < 0.0) q = 0.0;
    if (q > 255.0) q = 255.0;
    return (uint32_t)q;
}

// Children of the wide node made from BVH2 node `bid`: open the largest inner child until 8 slots are used.
RT2_HD int collapse_pick_children(const CollapseCtx& c, uint32_t bid, uint32_t* ch) {
    int nch = 0;
    if (bvh2_is_leaf(c, bid) || c.bvh2.count[bid] <= RT2_LEAF_MAX) ch[nch++] = bid;
    else { ch[nch++] = f2u(c.bvh2.lo[bid].w); ch[nch++] = f2u(c.bvh2.hi[bid].w); }
    // open the largest inner child until 8 slots are used
    while (nch < 8) {
        int pick = -1; float pa = -1.0f;
        for (int k = 0; k < nch; k++) {
            uint32_t x = ch[k];
            if (bvh2_is_leaf(c, x) || c.bvh2.count[x] <= RT2_LEAF_MAX) continue;
            float4 l = c.bvh2.lo[x], h = c.bvh2.hi[x];
            float a = box_half_area(l.x, l.y, l.z, h.x, h.y, h.z);
            if (a > pa) { pa = a; pick = k; }
        }
        if (pick < 0) break;
        uint32_t x = ch[pick];
        ch[pick] = f2u(c.bvh2.lo[x].w);
        ch[nch++] = f2u(c.bvh2.hi[x].w);
    }
    return nch;
}

// Children of the wide node made from `bid` as the SAH-optimal collapse decided them: the 8 slots are split between the two
// subtrees as D(bid,8) says, and every subtree spreads over its share as its own decisions say.
RT2_HD int collapse_pick_children_dp(const CollapseCtx& c, uint32_t bid, uint32_t* ch) {
    int nch = 0;
    if (bvh2_is_leaf(c, bid) || c.bvh2.count[bid] <= RT2_LEAF_MAX) { ch[nch++] = bid; return nch; }
    uint32_t sn[16]; int si[16]; int sp = 0;
    {
        const int k = (int)((c.dp_dec[bid] >> (3 * 6)) & 7u) + 1;           // D(bid,8)
        sn[sp] = f2u(c.bvh2.hi[bid].w); si[sp++] = 8 - k;
        sn[sp] = f2u(c.bvh2.lo[bid].w); si[sp++] = k;
    }
    while (sp) {
        const uint32_t n = sn[--sp]; int i = si[sp];
        const uint32_t dec = bvh2_is_leaf(c, n) ? 0u : c.dp_dec[n];
        while (i > 1 && !((dec >> (21 + (i - 2))) & 1u)) i--;
        if (i == 1) { ch[nch++] = n; continue; }
        const int k = (int)((dec >> (3 * (i - 2))) & 7u) + 1;
        sn[sp] = f2u(c.bvh2.hi[n].w); si[sp++] = i - k;
        sn[sp] = f2u(c.bvh2.lo[n].w); si[sp++] = k;
    }
    return nch;
}

// Write wide node `wid` with the BVH2 subtrees ch[0..nch) as its children (subtrees of <= RT2_LEAF_MAX primitives become leaf slots).
RT2_HD void collapse_emit(const CollapseCtx& c, uint32_t wid, const uint32_t* ch, int nch) {
    // node box + centroid
    double nlo[3] = {1e300, 1e300, 1e300}, nhi[3] = {-1e300, -1e300, -1e300};
    float clo[8][3], chi[8][3];
    for (int k = 0; k < nch; k++) {
        float4 l = c.bvh2.lo[ch[k]], h = c.bvh2.hi[ch[k]];
        clo[k][0] = l.x; clo[k][1] = l.y; clo[k][2] = l.z; chi[k][0] = h.x; chi[k][1] = h.y; chi[k][2] = h.z;
        for (int a = 0; a < 3; a++) { nlo[a] = fmin(nlo[a], (double)clo[k][a]); nhi[a] = fmax(nhi[a], (double)chi[k][a]); }
    }
    // slot assignment: greedy maximum of (child centroid - node centroid) . octant direction
    int slot_of[8]; bool slot_used[8]; bool child_done[8];
    for (int k = 0; k < 8; k++) { slot_used[k] = false; child_done[k] = false; slot_of[k] = -1; }
    float nc[3] = {(float)(0.5 * (nlo[0] + nhi[0])), (float)(0.5 * (nlo[1] + nhi[1])), (float)(0.5 * (nlo[2] + nhi[2]))};
    for (int it = 0; it < nch; it++) {
        float bestc = -INFINITY; int bk = -1, bs = -1;
        for (int k = 0; k < nch; k++) {
            if (child_done[k]) continue;
            float dx = 0.5f * (clo[k][0] + chi[k][0]) - nc[0], dy = 0.5f * (clo[k][1] + chi[k][1]) - nc[1], dz = 0.5f * (clo[k][2] + chi[k][2]) - nc[2];
            for (int s = 0; s < 8; s++) {
                if (slot_used[s]) continue;
                float cost = ((s & 1) ? dx : -dx) + ((s & 2) ? dy : -dy) + ((s & 4) ? dz : -dz);
                if (cost > bestc) { bestc = cost; bk = k; bs = s; }
            }
        }
        child_done[bk] = true; slot_used[bs] = true; slot_of[bk] = bs;
    }
    // classify + allocate
    uint32_t prim_ids[8][RT2_LEAF_MAX]; uint32_t prim_n[8]; bool inner[8];
    uint32_t n_inner = 0, n_prims = 0;
    for (int k = 0; k < nch; k++) {
        uint32_t x = ch[k];
        inner[k] = !(bvh2_is_leaf(c, x) || c.bvh2.count[x] <= RT2_LEAF_MAX);
        prim_n[k] = 0;
        if (inner[k]) n_inner++;
        else { prim_n[k] = gather_leaves(c, x, prim_ids[k]); n_prims += prim_n[k]; }
    }
    uint32_t child_base = n_inner ? RT2_ATOMIC_ADD(c.node_counter, n_inner) : 0u;
    uint32_t prim_base = n_prims ? RT2_ATOMIC_ADD(c.prim_counter, n_prims) : 0u;
    if (child_base + n_inner > c.max_nodes || prim_base + n_prims > c.max_prims) return;   // cannot happen: sized 2N

    Node8 nd;
    float p[3]; int e[3];
    for (int a = 0; a < 3; a++) choose_grid(nlo[a], nhi[a], &p[a], &e[a]);
    nd.px = p[0]; nd.py = p[1]; nd.pz = p[2];
    nd.ex = (uint8_t)(e[0] + 127); nd.ey = (uint8_t)(e[1] + 127); nd.ez = (uint8_t)(e[2] + 127);
    nd.imask = 0; nd.child_base = child_base; nd.tri_base = prim_base;
    for (int s = 0; s < 8; s++) { nd.meta[s] = 0; nd.qlox[s] = nd.qloy[s] = nd.qloz[s] = 255; nd.qhix[s] = nd.qhiy[s] = nd.qhiz[s] = 0; }
    // children by slot order: inner children get consecutive node indices, leaf children consecutive prims
    uint32_t inner_rank = 0, prim_off = 0;
    for (int s = 0; s < 8; s++) {
        int k = -1;
        for (int kk = 0; kk < nch; kk++) if (slot_of[kk] == s) k = kk;
        if (k < 0) continue;
        nd.qlox[s] = (uint8_t)quant_lo(clo[k][0], p[0], e[0]); nd.qhix[s] = (uint8_t)quant_hi(chi[k][0], p[0], e[0]);
        nd.qloy[s] = (uint8_t)quant_lo(clo[k][1], p[1], e[1]); nd.qhiy[s] = (uint8_t)quant_hi(chi[k][1], p[1], e[1]);
        nd.qloz[s] = (uint8_t)quant_lo(clo[k][2], p[2], e[2]); nd.qhiz[s] = (uint8_t)quant_hi(chi[k][2], p[2], e[2]);
        if (inner[k]) {
            nd.imask |= (uint8_t)(1u << s);
            nd.meta[s] = (uint8_t)((1u << 5) | (24u + (uint32_t)s));
            uint32_t qi = RT2_ATOMIC_ADD(c.queue_out_count, 1u);
            c.queue_out[qi] = make_uint2(ch[k], child_base + inner_rank);
            inner_rank++;
        } else {
            uint32_t cnt = prim_n[k];
            uint32_t unary = cnt == 1 ? 1u : (cnt == 2 ? 3u : 7u);
            nd.meta[s] = (uint8_t)((unary << 5) | prim_off);
            for (uint32_t t = 0; t < cnt; t++) c.out_prim[prim_base + prim_off + t] = c.leaf_prim[prim_ids[k][t]];
            prim_off += cnt;
        }
    }
    c.nodes[wid] = nd;
}

// Turn BVH2 node `bid` into wide node `wid`.
RT2_HD void collapse_node(const CollapseCtx& c, uint32_t bid, uint32_t wid) {
    uint32_t ch[8];
    const int nch = c.dp_dec ? collapse_pick_children_dp(c, bid, ch) : collapse_pick_children(c, bid, ch);
    collapse_emit(c, wid, ch, nch);
}

}  // namespace rt2
