// treelab.cu — MEASUREMENT HARNESS, not product code.  Builds the product's CWBVH on the CPU from the product's own
// __host__ __device__ builder functions (rt2_build.cuh) with optional tree-quality passes between PLOC and the collapse, and
// counts node steps and triangle tests per ray with the product's traversal (rt2_bvh.cuh).  Used to decide whether a SAH
// optimisation pass is worth building on the device (VERDICT r1 task 6); results in profiles/r02_tree_quality.txt.
#include "../../raytracer2_b200/csrc/rt2_reinsert.cuh"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <queue>
#include <vector>

using namespace rt2;

struct Tree2 {
    std::vector<float4> lo, hi; std::vector<uint32_t> count, parent;
    uint32_t root = 0, nleaves = 0;
    uint32_t left(uint32_t i) const { return f2u(lo[i].w); }
    uint32_t right(uint32_t i) const { return f2u(hi[i].w); }
    bool leaf(uint32_t i) const { return i < nleaves; }
    float area(uint32_t i) const { return box_half_area(lo[i].x, lo[i].y, lo[i].z, hi[i].x, hi[i].y, hi[i].z); }
    float union_area(uint32_t i, const float4& l, const float4& h) const {
        return box_half_area(fminf(lo[i].x, l.x), fminf(lo[i].y, l.y), fminf(lo[i].z, l.z), fmaxf(hi[i].x, h.x), fmaxf(hi[i].y, h.y), fmaxf(hi[i].z, h.z));
    }
    void refit_one(uint32_t i) {
        const uint32_t a = left(i), b = right(i);
        lo[i] = make_float4(fminf(lo[a].x, lo[b].x), fminf(lo[a].y, lo[b].y), fminf(lo[a].z, lo[b].z), u2f(a));
        hi[i] = make_float4(fmaxf(hi[a].x, hi[b].x), fmaxf(hi[a].y, hi[b].y), fmaxf(hi[a].z, hi[b].z), u2f(b));
        count[i] = count[a] + count[b];
    }
    void refit_up(uint32_t i) { while (i != 0xFFFFFFFFu) { refit_one(i); i = parent[i]; } }
    void replace_child(uint32_t p, uint32_t oldc, uint32_t newc) {
        if (left(p) == oldc) lo[p].w = u2f(newc); else hi[p].w = u2f(newc);
        parent[newc] = p;
    }
    double sah() const {      // sum of inner-node areas / root area (node cost) — the quantity reinsertion minimises
        double s = 0; for (uint32_t i = nleaves; i < lo.size(); i++) if (i == root || parent[i] != 0xFFFFFFFFu) s += area(i);
        return s / area(root);
    }
};

// Bittner, Hapala, Havran 2013, "Fast insertion-based optimization of bounding volume hierarchies": remove a subtree, find the
// position that minimises the total area increase by branch and bound, re-insert.  Nodes are visited in order of decreasing area.
static void reinsertion_pass(Tree2& t, float fraction) {
    std::vector<uint32_t> order;
    for (uint32_t i = 0; i < t.lo.size(); i++) if (i != t.root && t.parent[i] != 0xFFFFFFFFu && t.parent[i] != t.root) order.push_back(i);
    std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return t.area(a) > t.area(b); });
    order.resize((size_t)(order.size() * fraction));
    struct QE { float induced; uint32_t node; bool operator<(const QE& o) const { return induced > o.induced; } };
    for (uint32_t x : order) {
        const uint32_t p = t.parent[x];
        if (p == 0xFFFFFFFFu || p == t.root) continue;
        const uint32_t s = t.left(p) == x ? t.right(p) : t.left(p), gp = t.parent[p];
        t.replace_child(gp, p, s);
        t.refit_up(gp);
        const float4 xl = t.lo[x], xh = t.hi[x];
        const float ax = t.area(x);
        uint32_t best = s; float best_cost = INFINITY;
        std::priority_queue<QE> q; q.push({0.f, t.root});
        while (!q.empty()) {
            const QE e = q.top(); q.pop();
            if (e.induced + ax >= best_cost) break;
            const float direct = t.union_area(e.node, xl, xh), total = e.induced + direct;
            if (total < best_cost) { best_cost = total; best = e.node; }
            if (!t.leaf(e.node)) {
                const float ci = total - t.area(e.node);
                if (ci + ax < best_cost) { q.push({ci, t.left(e.node)}); q.push({ci, t.right(e.node)}); }
            }
        }
        const uint32_t gb = t.parent[best];
        t.lo[p].w = u2f(best); t.hi[p].w = u2f(x);
        t.parent[best] = p; t.parent[x] = p;
        if (gb == 0xFFFFFFFFu) { t.root = p; t.parent[p] = 0xFFFFFFFFu; t.refit_one(p); }
        else { t.replace_child(gb, best, p); t.refit_up(p); }
    }
}

// Reference-quality BVH2: top-down, 32-bin SAH on the longest... on all three axes (Wald 2007), one primitive per leaf.  The ceiling
// PLOC + reinsertion is compared against.
static uint32_t build_topdown(Tree2& t, std::vector<uint32_t>& prims, uint32_t begin, uint32_t end, uint32_t& next_inner) {
    if (end - begin == 1) return prims[begin];
    float cl[3] = {INFINITY, INFINITY, INFINITY}, ch[3] = {-INFINITY, -INFINITY, -INFINITY};
    auto cen = [&](uint32_t p, int a) { const float* l = &t.lo[p].x; const float* h = &t.hi[p].x; return 0.5f * (l[a] + h[a]); };
    for (uint32_t i = begin; i < end; i++) for (int a = 0; a < 3; a++) { float c = cen(prims[i], a); cl[a] = fminf(cl[a], c); ch[a] = fmaxf(ch[a], c); }
    const int NB = 32;
    float best_cost = INFINITY; int best_axis = -1, best_split = -1;
    for (int a = 0; a < 3; a++) {
        const float ext = ch[a] - cl[a];
        if (!(ext > 0)) continue;
        struct Bin { float lo[3], hi[3]; uint32_t n; } bins[NB];
        for (auto& b : bins) { b.n = 0; for (int k = 0; k < 3; k++) { b.lo[k] = INFINITY; b.hi[k] = -INFINITY; } }
        for (uint32_t i = begin; i < end; i++) {
            const uint32_t p = prims[i];
            int b = (int)((cen(p, a) - cl[a]) / ext * NB); if (b >= NB) b = NB - 1; if (b < 0) b = 0;
            bins[b].n++;
            const float* l = &t.lo[p].x; const float* h = &t.hi[p].x;
            for (int k = 0; k < 3; k++) { bins[b].lo[k] = fminf(bins[b].lo[k], l[k]); bins[b].hi[k] = fmaxf(bins[b].hi[k], h[k]); }
        }
        float ra[NB]; uint32_t rn[NB];
        { float l[3] = {INFINITY, INFINITY, INFINITY}, h[3] = {-INFINITY, -INFINITY, -INFINITY}; uint32_t n = 0;
          for (int b = NB - 1; b > 0; b--) { for (int k = 0; k < 3; k++) { l[k] = fminf(l[k], bins[b].lo[k]); h[k] = fmaxf(h[k], bins[b].hi[k]); } n += bins[b].n;
              ra[b] = n ? box_half_area(l[0], l[1], l[2], h[0], h[1], h[2]) : 0.f; rn[b] = n; } }
        { float l[3] = {INFINITY, INFINITY, INFINITY}, h[3] = {-INFINITY, -INFINITY, -INFINITY}; uint32_t n = 0;
          for (int b = 0; b < NB - 1; b++) { for (int k = 0; k < 3; k++) { l[k] = fminf(l[k], bins[b].lo[k]); h[k] = fmaxf(h[k], bins[b].hi[k]); } n += bins[b].n;
              if (n == 0 || rn[b + 1] == 0) continue;
              const float cost = box_half_area(l[0], l[1], l[2], h[0], h[1], h[2]) * n + ra[b + 1] * rn[b + 1];
              if (cost < best_cost) { best_cost = cost; best_axis = a; best_split = b; } } }
    }
    uint32_t mid;
    if (best_axis < 0) mid = (begin + end) / 2;
    else {
        const float ext = ch[best_axis] - cl[best_axis];
        auto it = std::partition(prims.begin() + begin, prims.begin() + end, [&](uint32_t p) {
            int b = (int)((cen(p, best_axis) - cl[best_axis]) / ext * NB); if (b >= NB) b = NB - 1; if (b < 0) b = 0; return b <= best_split; });
        mid = (uint32_t)(it - prims.begin());
        if (mid == begin || mid == end) mid = (begin + end) / 2;
    }
    const uint32_t id = next_inner++;
    const uint32_t a = build_topdown(t, prims, begin, mid, next_inner), b = build_topdown(t, prims, mid, end, next_inner);
    t.lo[id].w = u2f(a); t.hi[id].w = u2f(b);
    t.refit_one(id);
    return id;
}

// SAH-optimal collapse (Ylitie, Karras, Laine 2017, "Efficient incoherent ray traversal on GPUs through compressed wide BVHs", §4.1),
// adapted to this node layout (a subtree of <= RT2_LEAF_MAX triangles is a leaf SLOT of its parent, not a node):
//   C(n,1)   = A(n) * count * ct                      if count(n) <= RT2_LEAF_MAX   (leaf slot)
//            = A(n) * cn + D(n,8)                     otherwise                      (a wide node of its own)
//   C(n,i)   = min(C(n,i-1), D(n,i))                  i = 2..7: n spread over up to i slots of an ancestor's wide node
//   D(n,j)   = min_k C(left,k) + C(right,j-k)
struct DpCollapse {
    std::vector<float> C;            // [node][7]: C(n,1..7)
    std::vector<uint8_t> splitk;     // [node][9]: argmin k of D(n,j), j = 2..8
    std::vector<uint8_t> use;        // [node][8]: for i = 2..7: 1 = C(n,i) is D(n,i), 0 = falls back to C(n,i-1)
    const Tree2* t = nullptr; float cn = 1.f, ct = 0.3f;
    float c(uint32_t n, int i) const { return C[(size_t)n * 7 + (i - 1)]; }
    void solve() {
        const Tree2& T = *t; const size_t N = T.lo.size();
        C.assign(N * 7, 0.f); splitk.assign(N * 9, 0); use.assign(N * 8, 0);
        // children have smaller ids than parents?  Not after reinsertion: do an explicit post-order
        std::vector<uint32_t> order; order.reserve(N);
        std::vector<uint32_t> st{T.root};
        while (!st.empty()) { uint32_t x = st.back(); st.pop_back(); order.push_back(x); if (!T.leaf(x)) { st.push_back(T.left(x)); st.push_back(T.right(x)); } }
        for (size_t oi = order.size(); oi-- > 0;) {
            const uint32_t n = order[oi];
            const float A = T.area(n);
            if (T.leaf(n)) { for (int i = 1; i <= 7; i++) C[(size_t)n * 7 + i - 1] = A * ct; continue; }
            const uint32_t l = T.left(n), r = T.right(n);
            float D[9];
            for (int j = 2; j <= 8; j++) {
                float best = INFINITY; int bk = 1;
                for (int k = 1; k < j; k++) {
                    if (k > 7 || j - k > 7) continue;
                    const float v = c(l, k) + c(r, j - k);
                    if (v < best) { best = v; bk = k; }
                }
                D[j] = best; splitk[(size_t)n * 9 + j] = (uint8_t)bk;
            }
            C[(size_t)n * 7] = T.count[n] <= RT2_LEAF_MAX ? A * (float)T.count[n] * ct : A * cn + D[8];
            for (int i = 2; i <= 7; i++) {
                const float prev = C[(size_t)n * 7 + i - 2];
                if (D[i] < prev) { C[(size_t)n * 7 + i - 1] = D[i]; use[(size_t)n * 8 + i] = 1; }
                else C[(size_t)n * 7 + i - 1] = prev;
            }
        }
    }
    // the slots subtree n occupies when given up to i of them
    void expand(uint32_t n, int i, uint32_t* ch, int& nch) const {
        const Tree2& T = *t;
        while (i > 1 && !use[(size_t)n * 8 + i]) i--;
        if (i == 1 || T.leaf(n)) { ch[nch++] = n; return; }
        const int k = splitk[(size_t)n * 9 + i];
        expand(T.left(n), k, ch, nch); expand(T.right(n), i - k, ch, nch);
    }
    int children(uint32_t n, uint32_t* ch) const {          // n becomes a wide node
        const Tree2& T = *t; int nch = 0;
        if (T.leaf(n) || T.count[n] <= RT2_LEAF_MAX) { ch[nch++] = n; return nch; }
        const int k = splitk[(size_t)n * 9 + 8];
        expand(T.left(n), k, ch, nch); expand(T.right(n), 8 - k, ch, nch);
        return nch;
    }
};

// The product's parallel reinsertion (rt2_reinsert.cuh), phases run one after the other over all nodes as the kernels do
static void parallel_reinsertion(Tree2& t, int iterations, int mod, bool verbose) {
    const uint32_t N = (uint32_t)t.lo.size();
    std::vector<uint32_t> target(N), lca(N), flags(N); std::vector<float> gain(N); std::vector<unsigned long long> lock(N), pass(N);
    std::vector<uint8_t> winner(N);
    for (int it = 0; it < iterations; it++) {
        Reinsert r{Bvh2{t.lo.data(), t.hi.data(), t.count.data()}, t.parent.data(), t.root, t.nleaves, N, target.data(), lca.data(), gain.data(), lock.data(), pass.data()};
        std::fill(lock.begin(), lock.end(), 0ull); std::fill(pass.begin(), pass.end(), 0ull);
        uint32_t wanted = 0, applied = 0;
        for (uint32_t i = 0; i < N; i++) { if (mod > 1 && (int)((i + it) % mod) != 0) { target[i] = 0xFFFFFFFFu; continue; } reinsert_find(r, i); if (target[i] != 0xFFFFFFFFu) wanted++; }
        for (uint32_t i = 0; i < N; i++) reinsert_lock(r, i);
        for (uint32_t i = 0; i < N; i++) winner[i] = reinsert_check(r, i) ? 1 : 0;
        for (uint32_t i = 0; i < N; i++) if (winner[i]) { reinsert_apply(r, i, target[i]); applied++; }
        // refit bottom-up: post-order from the root
        std::vector<uint32_t> order; order.reserve(N); std::vector<uint32_t> st{t.root};
        while (!st.empty()) { uint32_t x = st.back(); st.pop_back(); order.push_back(x); if (!t.leaf(x)) { st.push_back(t.left(x)); st.push_back(t.right(x)); } }
        if (order.size() != N) { fprintf(stderr, "reinsertion broke the tree: %zu of %u nodes reachable\n", order.size(), N); abort(); }
        for (size_t k = order.size(); k-- > 0;) if (!t.leaf(order[k])) reinsert_refit_node(r.b, order[k]);
        if (verbose) fprintf(stderr, "  iteration %d: %u moves wanted, %u applied, SAH %.3f\n", it, wanted, applied, t.sah());
    }
}

struct Lab {
    std::vector<Node8> nodes; std::vector<TriPos> tris; std::vector<uint32_t> slot_to_prim;
    double sah_before = 0, sah_after = 0; uint32_t depth = 0;
};

static bool build(Lab& L, const std::vector<float4>& lo, const std::vector<float4>& hi, int passes, float fraction, int collapse_mode) {
    (void)collapse_mode;
    const uint32_t n = (uint32_t)lo.size();
    V3 slo = v3(INFINITY, INFINITY, INFINITY), shi = v3(-INFINITY, -INFINITY, -INFINITY);
    for (uint32_t i = 0; i < n; i++) {
        slo = v3(fminf(slo.x, lo[i].x), fminf(slo.y, lo[i].y), fminf(slo.z, lo[i].z));
        shi = v3(fmaxf(shi.x, hi[i].x), fmaxf(shi.y, hi[i].y), fmaxf(shi.z, hi[i].z));
    }
    std::vector<std::pair<uint64_t, uint32_t>> kv(n);
    for (uint32_t i = 0; i < n; i++) kv[i] = {morton_key(lo[i], hi[i], slo, shi), i};
    std::stable_sort(kv.begin(), kv.end(), [](auto& a, auto& b) { return a.first < b.first; });
    std::vector<uint32_t> sorted(n);
    for (uint32_t i = 0; i < n; i++) sorted[i] = kv[i].second;
    Tree2 t; t.nleaves = n;
    t.lo.resize(2 * (size_t)n); t.hi.resize(2 * (size_t)n); t.count.resize(2 * (size_t)n);
    std::vector<uint32_t> cl0(n), cl1(n), nearest(n), valid(n);
    Bvh2 b2{t.lo.data(), t.hi.data(), t.count.data()};
    for (uint32_t i = 0; i < n; i++) {
        float4 l = lo[sorted[i]], h = hi[sorted[i]];
        l.w = u2f(0xFFFFFFFFu); h.w = u2f(0xFFFFFFFFu);
        t.lo[i] = l; t.hi[i] = h; t.count[i] = 1; cl0[i] = i;
    }
    uint32_t node_counter = n, cur = n;
    if (collapse_mode & 16) {          // build mode: top-down binned SAH instead of PLOC
        std::vector<uint32_t> prims(n); for (uint32_t i = 0; i < n; i++) prims[i] = i;
        cl0[0] = build_topdown(t, prims, 0, n, node_counter);
        cur = 1;
    }
    while (cur > 1) {
        for (uint32_t i = 0; i < cur; i++) nearest[i] = ploc_nearest(b2, cl0.data(), cur, i);
        for (uint32_t i = 0; i < cur; i++) valid[i] = ploc_merge(b2, cl0.data(), nearest.data(), cur, i, &node_counter);
        uint32_t m = 0;
        for (uint32_t i = 0; i < cur; i++) if (valid[i]) cl1[m++] = cl0[i];
        if (m >= cur || m == 0) return false;
        cur = m; std::swap(cl0, cl1);
    }
    t.root = cl0[0];
    t.lo.resize(node_counter); t.hi.resize(node_counter); t.count.resize(node_counter);
    t.parent.assign(node_counter, 0xFFFFFFFFu);
    for (uint32_t i = n; i < node_counter; i++) { t.parent[t.left(i)] = i; t.parent[t.right(i)] = i; }
    L.sah_before = t.sah();
    if (collapse_mode & 32) parallel_reinsertion(t, passes, (collapse_mode >> 16) & 0xff, true);
    else for (int k = 0; k < passes; k++) reinsertion_pass(t, fraction);
    L.sah_after = t.sah();
    b2 = Bvh2{t.lo.data(), t.hi.data(), t.count.data()};
    DpCollapse dp; const bool dp_on = (collapse_mode & 1) != 0;
    if (dp_on) { dp.t = &t; dp.ct = ((collapse_mode >> 8) & 0xff) ? (float)((collapse_mode >> 8) & 0xff) / 100.f : 0.3f; dp.solve(); }
    L.nodes.assign((size_t)n + 1, Node8{});
    L.slot_to_prim.assign(n, 0);
    std::vector<uint2> q0(n), q1(n);
    uint32_t wide_counter = 1, prim_counter = 0, qcount = 0;
    CollapseCtx cc{};
    cc.bvh2 = b2; cc.num_leaves = n; cc.leaf_prim = sorted.data(); cc.nodes = L.nodes.data(); cc.node_counter = &wide_counter;
    cc.prim_counter = &prim_counter; cc.out_prim = L.slot_to_prim.data(); cc.queue_out_count = &qcount; cc.max_nodes = n + 1; cc.max_prims = n;
    q0[0] = make_uint2(t.root, 0u);
    uint32_t qn = 1, depth = 0;
    while (qn) {
        depth++; qcount = 0; cc.queue_out = q1.data();
        for (uint32_t i = 0; i < qn; i++) {
            if (dp_on) { uint32_t ch[8]; const int nch = dp.children(q0[i].x, ch); collapse_emit(cc, q0[i].y, ch, nch); }
            else collapse_node(cc, q0[i].x, q0[i].y);
        }
        qn = qcount; std::swap(q0, q1);
        if (depth > 400) return false;
    }
    if (prim_counter != n) return false;
    L.nodes.resize(wide_counter); L.depth = depth;
    return true;
}

extern "C" {

void* treelab_build(const rt2_scene_desc* d, int reinsertion_passes, float fraction, int collapse_mode, double* sah_before, double* sah_after, uint32_t* wide_nodes, uint32_t* depth) {
    Lab* L = new Lab();
    std::vector<uint32_t> offs(d->num_geometries + 1, 0);
    for (uint32_t g = 0; g < d->num_geometries; g++) offs[g + 1] = offs[g] + d->index_counts[g] / 3;
    const uint32_t n = offs[d->num_geometries];
    BuildInput in{d->vertices, d->indices, d->geometries, offs.data(), d->num_geometries, n};
    std::vector<TriPos> tris_in(n); std::vector<float4> lo(n), hi(n);
    for (uint32_t i = 0; i < n; i++) build_triangle(in, i, tris_in.data(), lo.data(), hi.data());
    if (!build(*L, lo, hi, reinsertion_passes, fraction, collapse_mode)) { delete L; return nullptr; }
    L->tris.resize(n);
    for (uint32_t s = 0; s < n; s++) L->tris[s] = tris_in[L->slot_to_prim[s]];
    *sah_before = L->sah_before; *sah_after = L->sah_after; *wide_nodes = (uint32_t)L->nodes.size(); *depth = L->depth;
    return L;
}
// a world-space triangle soup (9 floats per triangle): the flattened-instancing tree of an instanced scene has exactly these leaves
void* treelab_build_soup(const float* tris9, uint32_t n, int reinsertion_passes, float fraction, int collapse_mode, double* sah_before, double* sah_after, uint32_t* wide_nodes, uint32_t* depth) {
    Lab* L = new Lab();
    std::vector<TriPos> tris_in(n); std::vector<float4> lo(n), hi(n);
    for (uint32_t i = 0; i < n; i++) {
        const float* t = tris9 + 9 * (size_t)i;
        tris_in[i].v0 = make_float4(t[0], t[1], t[2], u2f(0u)); tris_in[i].v1 = make_float4(t[3], t[4], t[5], u2f(i)); tris_in[i].v2 = make_float4(t[6], t[7], t[8], 0.f);
        lo[i] = make_float4(fminf(t[0], fminf(t[3], t[6])), fminf(t[1], fminf(t[4], t[7])), fminf(t[2], fminf(t[5], t[8])), u2f(0xFFFFFFFFu));
        hi[i] = make_float4(fmaxf(t[0], fmaxf(t[3], t[6])), fmaxf(t[1], fmaxf(t[4], t[7])), fmaxf(t[2], fmaxf(t[5], t[8])), u2f(0xFFFFFFFFu));
    }
    if (!build(*L, lo, hi, reinsertion_passes, fraction, collapse_mode)) { delete L; return nullptr; }
    L->tris.resize(n);
    for (uint32_t s = 0; s < n; s++) L->tris[s] = tris_in[L->slot_to_prim[s]];
    *sah_before = L->sah_before; *sah_after = L->sah_after; *wide_nodes = (uint32_t)L->nodes.size(); *depth = L->depth;
    return L;
}
void treelab_destroy(void* p) { delete (Lab*)p; }

// rays: o[3] d[3]; returns total node steps and triangle tests; hit t written per ray (inf = miss)
void treelab_trace(void* p, const float* rays, uint32_t n, float tmin, float tmax, float* t_out, uint64_t* node_steps, uint64_t* tri_tests) {
    Lab* L = (Lab*)p;
    uint64_t ns = 0, tt = 0;
    for (uint32_t i = 0; i < n; i++) {
        TraceRay ray; ray.o = v3(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]); ray.d = v3(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]);
        ray.tmin = tmin; ray.tmax = tmax;
        float best = INFINITY;
        uint32_t visits = 0;
        const RayShear sh = make_shear(ray.d);
        cwbvh_traverse(L->nodes.data(), ray, &best, [&](uint32_t slot) {
            tt++;
            const TriPos* tp = L->tris.data() + slot;
            float t, u, v;
            if (ray_triangle(ray.o, sh, ray.tmin, ray.tmax, v3(tp->v0.x, tp->v0.y, tp->v0.z), v3(tp->v1.x, tp->v1.y, tp->v1.z), v3(tp->v2.x, tp->v2.y, tp->v2.z), &t, &u, &v) && t < best) best = t;
        }, &visits);
        ns += visits;
        if (t_out) t_out[i] = best;
    }
    *node_steps = ns; *tri_tests = tt;
}

}  // extern "C"
