// rt2_reinsert.cuh — insertion-based optimisation of the PLOC BVH2, in parallel (Meister & Bittner 2018, "Parallel reinsertion for
// bounding volume hierarchy optimization"; the sequential idea is Bittner, Hapala, Havran 2013).  Runs between PLOC and the collapse
// (rt2_build.cuh steps 4 and 5).  One iteration:
//   find   every node looks, read-only, for the position where re-inserting its subtree shrinks the total node area most
//   lock   every node with a positive gain stamps (gain, id) with atomicMax on the six nodes its move rewires and, in a second array, on
//          the ancestors of its target up to the common ancestor
//   check  which moves keep all their stamps (and do not aim into a subtree a stronger move carries away)
//   apply  those moves are carried out (pointer surgery only: the removed parent node is re-used above the target)
//   refit  boxes and counts bottom-up
// Measured on the CPU with the product's traversal (tools/treelab.py, profiles/r02_tree_quality.txt): node steps per bounce ray
// -9 % on the C2 hall, -20 % on the flattened instanced scene.  Per-element __host__ __device__ functions, as in rt2_build.cuh: the
// kernels in rt2_build.cu are thin wrappers and tests/hostemu + tools/treelab run the same code serially.
#pragma once
#include "rt2_build.cuh"

namespace rt2 {

#ifdef __CUDA_ARCH__
#define RT2_ATOMIC_MAX64(ptr, v) atomicMax((ptr), (v))
#else
inline unsigned long long rt2_host_atomic_max64(unsigned long long* p, unsigned long long v) { unsigned long long o = *p; if (v > o) *p = v; return o; }
#define RT2_ATOMIC_MAX64(ptr, v) rt2::rt2_host_atomic_max64((ptr), (v))
#endif

struct Reinsert {
    Bvh2 b; uint32_t* parent; uint32_t root, num_leaves, num_nodes;
    uint32_t* target;                 // per node: where to go (0xFFFFFFFF = stay)
    uint32_t* lca;                    // per node: lowest common ancestor of the node and its target
    float* gain;                      // per node: area saved by the move
    unsigned long long* lock;         // per node: highest (gain bits << 32 | id) that wants to rewire it
    unsigned long long* pass;         // per node: highest key of a move whose target lies below it
};

#define RT2_REINSERT_STACK 48

RT2_HD float reinsert_union_area(const Bvh2& b, uint32_t y, const float4& l, const float4& h) {
    const float4 yl = b.lo[y], yh = b.hi[y];
    return box_half_area(fminf(yl.x, l.x), fminf(yl.y, l.y), fminf(yl.z, l.z), fmaxf(yh.x, h.x), fmaxf(yh.y, h.y), fmaxf(yh.z, h.z));
}
RT2_HD uint32_t reinsert_sibling(const Reinsert& r, uint32_t x) {
    const uint32_t p = r.parent[x];
    const uint32_t l = f2u(r.b.lo[p].w);
    return l == x ? f2u(r.b.hi[p].w) : l;
}

// branch and bound over the subtree of `out`: d = area already saved when `in` leaves its place and lands somewhere below `out`
RT2_HD void reinsert_search(const Reinsert& r, uint32_t out, float d, const float4& il, const float4& ih, float a_in, uint32_t lca_here,
                            uint32_t* best, float* d_best, uint32_t* best_lca) {
    uint32_t sn[RT2_REINSERT_STACK]; float sd[RT2_REINSERT_STACK]; int sp = 0;
    sn[0] = out; sd[0] = d; sp = 1;
    while (sp) {
        const uint32_t y = sn[--sp]; const float dy = sd[sp];
        if (dy - a_in <= *d_best) continue;                              // the direct cost is at least area(in)
        const float a_union = reinsert_union_area(r.b, y, il, ih);
        const float d_direct = dy - a_union;
        if (d_direct > *d_best) { *d_best = d_direct; *best = y; *best_lca = lca_here; }
        if (y >= r.num_leaves) {
            const float d_child = dy - (a_union - bvh2_half_area(r.b, y));
            if (d_child - a_in > *d_best && sp + 2 <= RT2_REINSERT_STACK) {
                sn[sp] = f2u(r.b.lo[y].w); sd[sp++] = d_child;
                sn[sp] = f2u(r.b.hi[y].w); sd[sp++] = d_child;
            }
        }
    }
}

// find: the best new position of subtree `in`
RT2_HD void reinsert_find(const Reinsert& r, uint32_t in) {
    r.target[in] = 0xFFFFFFFFu; r.gain[in] = 0.f; r.lca[in] = 0xFFFFFFFFu;
    if (in == r.root) return;
    const uint32_t p = r.parent[in];
    if (p == r.root) return;                                              // the root stays the root
    const float4 il = r.b.lo[in], ih = r.b.hi[in];
    const float a_in = box_half_area(il.x, il.y, il.z, ih.x, ih.y, ih.z);
    const float a_p = bvh2_half_area(r.b, p);
    uint32_t best = 0xFFFFFFFFu, best_lca = 0xFFFFFFFFu; float d_best = 0.f;
    const uint32_t s = reinsert_sibling(r, in);
    // below the sibling: the parent p goes away and comes back above the target, its old place holds the sibling grown by `in`
    reinsert_search(r, s, a_p, il, ih, a_in, p, &best, &d_best, &best_lca);
    if (best == s) { best = 0xFFFFFFFFu; d_best = 0.f; }                   // that is where it is now
    // climb: pivot_box = what the pivot's box shrinks to once `in` is gone
    uint32_t pivot = p;
    float4 pl = r.b.lo[s], ph = r.b.hi[s];
    float d = a_p;
    while (pivot != r.root) {
        const uint32_t up = r.parent[pivot];
        const uint32_t out = reinsert_sibling(r, pivot);
        reinsert_search(r, out, d, il, ih, a_in, up, &best, &d_best, &best_lca);
        const float4 ol = r.b.lo[out], oh = r.b.hi[out];
        pl = make_float4(fminf(pl.x, ol.x), fminf(pl.y, ol.y), fminf(pl.z, ol.z), 0.f);
        ph = make_float4(fmaxf(ph.x, oh.x), fmaxf(ph.y, oh.y), fmaxf(ph.z, oh.z), 0.f);
        d += bvh2_half_area(r.b, up) - box_half_area(pl.x, pl.y, pl.z, ph.x, ph.y, ph.z);
        pivot = up;
    }
    if (best != 0xFFFFFFFFu && d_best > 0.f) { r.target[in] = best; r.gain[in] = d_best; r.lca[in] = best_lca; }
}

// Which simultaneous moves may both be carried out?  A move rewires six nodes — `in`, its parent p, sibling s and grandparent gp, the
// target and the target's parent — and those must be its alone (`lock`: atomicMax of (gain, id)).  Beyond that the tree stays a tree
// unless a subtree ends up below itself, which takes a move whose target lies inside another moving subtree: so a move also stamps the
// ancestors of its target up to the common ancestor (`pass`), and of two moves where one's `in` / p sits on the other's target path
// the one with the smaller gain gives way.  Ancestors of `in` need no protection: they only shrink.  (The gains of the surviving
// moves are estimates once several of them reshape the same ancestors; the refit makes the boxes exact again.)
template <class Fn>
RT2_HD void reinsert_for_owned(const Reinsert& r, uint32_t in, Fn&& fn) {
    const uint32_t best = r.target[in], p = r.parent[in];
    fn(in); fn(p); fn(reinsert_sibling(r, in)); fn(r.parent[p]); fn(best); fn(r.parent[best]);
}
template <class Fn>
RT2_HD void reinsert_for_target_path(const Reinsert& r, uint32_t in, Fn&& fn) {
    const uint32_t lca = r.lca[in];
    for (uint32_t x = r.parent[r.target[in]]; x != lca; x = r.parent[x]) fn(x);
}
RT2_HD unsigned long long reinsert_key(const Reinsert& r, uint32_t in) { return ((unsigned long long)f2u(r.gain[in]) << 32) | (unsigned long long)in; }

RT2_HD void reinsert_lock(const Reinsert& r, uint32_t in) {
    if (r.target[in] == 0xFFFFFFFFu) return;
    const unsigned long long key = reinsert_key(r, in);
    reinsert_for_owned(r, in, [&](uint32_t x) { RT2_ATOMIC_MAX64(r.lock + x, key); });
    reinsert_for_target_path(r, in, [&](uint32_t x) { RT2_ATOMIC_MAX64(r.pass + x, key); });
}
// check (read-only, its own pass: a loser's paths may be rewired by the winners once they start): may this move go ahead?
RT2_HD bool reinsert_check(const Reinsert& r, uint32_t in) {
    if (r.target[in] == 0xFFFFFFFFu) return false;
    const unsigned long long key = reinsert_key(r, in);
    bool mine = true;
    reinsert_for_owned(r, in, [&](uint32_t x) { if (r.lock[x] != key) mine = false; });
    if (r.pass[in] > key || r.pass[r.parent[in]] > key) mine = false;             // a stronger move aims below this subtree
    reinsert_for_target_path(r, in, [&](uint32_t x) { if (r.lock[x] > key) mine = false; });   // a stronger move rewires an ancestor of the target
    return mine;
}
// apply a winning move: pointer surgery on nodes this move owns exclusively (boxes are refitted afterwards)
RT2_HD void reinsert_apply(const Reinsert& r, uint32_t in, uint32_t best) {
    const uint32_t p = r.parent[in], s = reinsert_sibling(r, in), gp = r.parent[p], bp = r.parent[best];
    // the sibling takes the parent's place
    if (f2u(r.b.lo[gp].w) == p) r.b.lo[gp].w = u2f(s); else r.b.hi[gp].w = u2f(s);
    r.parent[s] = gp;
    // the parent node comes back above the target (best == s was discarded in find, so bp != p)
    if (f2u(r.b.lo[bp].w) == best) r.b.lo[bp].w = u2f(p); else r.b.hi[bp].w = u2f(p);
    r.parent[p] = bp;
    r.b.lo[p].w = u2f(best); r.b.hi[p].w = u2f(in);
    r.parent[best] = p; r.parent[in] = p;
}
// refit of one inner node from its children.  On the device the children were written by other threads of the same launch (bottom-up
// climb): read them past L1.
#ifdef __CUDA_ARCH__
#define RT2_LDCG_F4(p) __ldcg(p)
#define RT2_LDCG_U(p) __ldcg(p)
#else
#define RT2_LDCG_F4(p) (*(p))
#define RT2_LDCG_U(p) (*(p))
#endif
RT2_HD void reinsert_refit_node(const Bvh2& b, uint32_t n) {
    const uint32_t l = f2u(b.lo[n].w), rr = f2u(b.hi[n].w);
    const float4 ll = RT2_LDCG_F4(b.lo + l), lh = RT2_LDCG_F4(b.hi + l), rl = RT2_LDCG_F4(b.lo + rr), rh = RT2_LDCG_F4(b.hi + rr);
    b.lo[n] = make_float4(fminf(ll.x, rl.x), fminf(ll.y, rl.y), fminf(ll.z, rl.z), u2f(l));
    b.hi[n] = make_float4(fmaxf(lh.x, rh.x), fmaxf(lh.y, rh.y), fmaxf(lh.z, rh.z), u2f(rr));
    b.count[n] = RT2_LDCG_U(b.count + l) + RT2_LDCG_U(b.count + rr);
}

}  // namespace rt2
