// hostemu.cu — TEST ONLY.  Runs the product's __host__ __device__ builder / traversal logic
// (raytracer2_b200/csrc/rt2_build.cuh, rt2_bvh.cuh) serially on the CPU so that the BVH build,
// the Node8 encoding and the CWBVH traversal can be checked against the oracle without a GPU.
// Not linked into librt2cuda.so; the product has no CPU path.
#include "../../raytracer2_b200/csrc/rt2_reinsert.cuh"
#include "../../include/rt2_gconst.h"

#include <algorithm>
#include <cstring>
#include <vector>

using namespace rt2;

struct HostAccel {
    std::vector<Node8> nodes;
    std::vector<TriPos> tris;
    std::vector<uint32_t> slot_to_prim;
    uint32_t depth = 0;
};

static bool build_from_boxes(const std::vector<float4>& lo, const std::vector<float4>& hi, std::vector<Node8>& nodes,
                             std::vector<uint32_t>& out_prim, uint32_t* depth_out, uint32_t ploc_radius = RT2_PLOC_RADIUS) {
    const uint32_t n = (uint32_t)lo.size();
    if (n == 0) { nodes.assign(1, Node8{}); memset(nodes.data(), 0, sizeof(Node8)); return true; }
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
    std::vector<float4> blo(2 * (size_t)n), bhi(2 * (size_t)n);
    std::vector<uint32_t> bcount(2 * (size_t)n), cl0(n), cl1(n), nearest(n), valid(n);
    Bvh2 b2{blo.data(), bhi.data(), bcount.data()};
    for (uint32_t i = 0; i < n; i++) {
        float4 l = lo[sorted[i]], h = hi[sorted[i]];
        l.w = u2f(0xFFFFFFFFu); h.w = u2f(0xFFFFFFFFu);
        blo[i] = l; bhi[i] = h; bcount[i] = 1; cl0[i] = i;
    }
    uint32_t node_counter = n, cur = n;
    while (cur > 1) {
        for (uint32_t i = 0; i < cur; i++) nearest[i] = ploc_nearest(b2, cl0.data(), cur, i, ploc_radius);
        // emulate the parallel kernel: every "thread" reads the pre-merge cluster array
        std::vector<uint32_t> snapshot(cl0.begin(), cl0.begin() + cur);
        for (uint32_t i = 0; i < cur; i++) {
            // ploc_merge reads clusters[i], clusters[j] and writes clusters[i]; j > i has not been written yet
            valid[i] = ploc_merge(b2, cl0.data(), nearest.data(), cur, i, &node_counter);
        }
        (void)snapshot;
        uint32_t m = 0;
        for (uint32_t i = 0; i < cur; i++) if (valid[i]) cl1[m++] = cl0[i];
        if (m >= cur || m == 0) return false;
        cur = m; std::swap(cl0, cl1);
    }
    const uint32_t root = cl0[0];
    nodes.assign((size_t)n + 1, Node8{});
    out_prim.assign(n, 0);
    std::vector<uint2> q0(n), q1(n);
    uint32_t wide_counter = 1, prim_counter = 0, qcount = 0;
    // insertion-based optimisation (rt2_reinsert.cuh), phase by phase as the kernels of rt2_build.cu run it
    std::vector<uint32_t> parent(node_counter, 0xFFFFFFFFu);
    for (uint32_t i = n; i < node_counter; i++) { parent[f2u(blo[i].w)] = i; parent[f2u(bhi[i].w)] = i; }
    auto post_order = [&]() {
        std::vector<uint32_t> order; order.reserve(node_counter); std::vector<uint32_t> st{root};
        while (!st.empty()) { uint32_t x = st.back(); st.pop_back(); order.push_back(x); if (x >= n) { st.push_back(f2u(blo[x].w)); st.push_back(f2u(bhi[x].w)); } }
        std::reverse(order.begin(), order.end());
        return order;
    };
    if (n >= 16) {
        std::vector<uint32_t> target(node_counter), lca(node_counter); std::vector<float> gain(node_counter);
        std::vector<unsigned long long> lock(node_counter), pass(node_counter); std::vector<uint8_t> win(node_counter);
        Reinsert r{b2, parent.data(), root, n, node_counter, target.data(), lca.data(), gain.data(), lock.data(), pass.data()};
        const int iters = n > (4u << 20) ? 6 : 12;
        for (int it = 0; it < iters; it++) {
            std::fill(lock.begin(), lock.end(), 0ull); std::fill(pass.begin(), pass.end(), 0ull);
            for (uint32_t i = 0; i < node_counter; i++) reinsert_find(r, i);
            for (uint32_t i = 0; i < node_counter; i++) reinsert_lock(r, i);
            uint32_t applied = 0;
            for (uint32_t i = 0; i < node_counter; i++) { win[i] = reinsert_check(r, i) ? 1 : 0; applied += win[i]; }
            for (uint32_t i = 0; i < node_counter; i++) if (win[i]) reinsert_apply(r, i, target[i]);
            const std::vector<uint32_t> order = post_order();
            if (order.size() != node_counter) return false;
            for (uint32_t x : order) if (x >= n) reinsert_refit_node(b2, x);
            if (applied * 2000ull < n) break;
        }
    }
    std::vector<float> dp_cost((size_t)node_counter * 7); std::vector<uint32_t> dp_dec(node_counter);
    CollapseDp dp{dp_cost.data(), dp_dec.data()};
    for (uint32_t x : post_order()) { if (x < n) dp_solve_leaf(dp, b2, x); else dp_solve_inner(dp, b2, x); }
    CollapseCtx cc{};
    cc.dp_dec = dp_dec.data();
    cc.bvh2 = b2; cc.num_leaves = n; cc.leaf_prim = sorted.data(); cc.nodes = nodes.data(); cc.node_counter = &wide_counter;
    cc.prim_counter = &prim_counter; cc.out_prim = out_prim.data(); cc.queue_out_count = &qcount; cc.max_nodes = n + 1; cc.max_prims = n;
    q0[0] = make_uint2(root, 0u);
    uint32_t qn = 1, depth = 0;
    while (qn) {
        depth++; qcount = 0; cc.queue_out = q1.data();
        for (uint32_t i = 0; i < qn; i++) collapse_node(cc, q0[i].x, q0[i].y);
        qn = qcount; std::swap(q0, q1);
        if (depth > 200) return false;
    }
    if (prim_counter != n) return false;
    nodes.resize(wide_counter);
    if (depth_out) *depth_out = depth;
    return true;
}

struct HostScene {
    HostAccel blas;
    std::vector<Node8> tlas_nodes; std::vector<uint32_t> tlas_slot_to_inst; std::vector<InstanceRec> instances;
    uint32_t tlas_depth = 0;
};

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

extern "C" {

void* hostemu_build(const rt2_scene_desc* d) {
    HostScene* hs = new HostScene();
    std::vector<uint32_t> offs(d->num_geometries + 1, 0);
    for (uint32_t g = 0; g < d->num_geometries; g++) offs[g + 1] = offs[g] + d->index_counts[g] / 3;
    const uint32_t n = offs[d->num_geometries];
    BuildInput in{d->vertices, d->indices, d->geometries, offs.data(), d->num_geometries, n};
    std::vector<TriPos> tris_in(n); std::vector<float4> lo(n), hi(n);
    for (uint32_t i = 0; i < n; i++) build_triangle(in, i, tris_in.data(), lo.data(), hi.data());
    if (!build_from_boxes(lo, hi, hs->blas.nodes, hs->blas.slot_to_prim, &hs->blas.depth, RT2_PLOC_RADIUS_TRIANGLES)) { delete hs; return nullptr; }
    hs->blas.tris.resize(n);
    for (uint32_t s = 0; s < n; s++) hs->blas.tris[s] = tris_in[hs->blas.slot_to_prim[s]];
    if (d->num_instances && d->instance_transforms) {
        V3 blo = v3(INFINITY, INFINITY, INFINITY), bhi = v3(-INFINITY, -INFINITY, -INFINITY);
        for (uint32_t i = 0; i < n; i++) { blo = v3(fminf(blo.x, lo[i].x), fminf(blo.y, lo[i].y), fminf(blo.z, lo[i].z)); bhi = v3(fmaxf(bhi.x, hi[i].x), fmaxf(bhi.y, hi[i].y), fmaxf(bhi.z, hi[i].z)); }
        hs->instances.resize(d->num_instances);
        std::vector<float4> ilo(d->num_instances), ihi(d->num_instances);
        for (uint32_t i = 0; i < d->num_instances; i++) {
            memcpy(hs->instances[i].fwd, d->instance_transforms + (size_t)i * 12, 48);
            invert_affine(hs->instances[i].fwd, hs->instances[i].inv);
            const float* m = hs->instances[i].fwd;
            V3 l = v3(INFINITY, INFINITY, INFINITY), h = v3(-INFINITY, -INFINITY, -INFINITY);
            for (int c = 0; c < 8; c++) {
                V3 p = xform_point(m, v3((c & 1) ? bhi.x : blo.x, (c & 2) ? bhi.y : blo.y, (c & 4) ? bhi.z : blo.z));
                l = v3(fminf(l.x, p.x), fminf(l.y, p.y), fminf(l.z, p.z)); h = v3(fmaxf(h.x, p.x), fmaxf(h.y, p.y), fmaxf(h.z, p.z));
            }
            V3 e = (h - l) * 1e-4f + v3(1e-6f, 1e-6f, 1e-6f);
            e = e + v3(fmaxf(fabsf(l.x), fabsf(h.x)), fmaxf(fabsf(l.y), fabsf(h.y)), fmaxf(fabsf(l.z), fabsf(h.z))) * 9.5367431640625e-07f;
            ilo[i] = make_float4(l.x - e.x, l.y - e.y, l.z - e.z, 0.f); ihi[i] = make_float4(h.x + e.x, h.y + e.y, h.z + e.z, 0.f);
        }
        if (!build_from_boxes(ilo, ihi, hs->tlas_nodes, hs->tlas_slot_to_inst, &hs->tlas_depth)) { delete hs; return nullptr; }
    }
    return hs;
}
void hostemu_destroy(void* p) { delete (HostScene*)p; }

// PLOC BVH2 of the scene's triangles + `iters` reinsertion rounds (rt2_reinsert.cuh), checking after every round that the result is
// still a binary tree over the same leaves with consistent parent pointers, enclosing boxes and additive counts.
// out[0] = SAH (sum of inner-node areas / root area) before, out[1] = after, out[2] = moves applied, out[3] = 1 if every check held.
void hostemu_reinsertion_probe(const rt2_scene_desc* d, int iters, double* out) {
    std::vector<uint32_t> offs(d->num_geometries + 1, 0);
    for (uint32_t g = 0; g < d->num_geometries; g++) offs[g + 1] = offs[g] + d->index_counts[g] / 3;
    const uint32_t n = offs[d->num_geometries];
    BuildInput in{d->vertices, d->indices, d->geometries, offs.data(), d->num_geometries, n};
    std::vector<TriPos> tris_in(n); std::vector<float4> lo(n), hi(n);
    for (uint32_t i = 0; i < n; i++) build_triangle(in, i, tris_in.data(), lo.data(), hi.data());
    V3 slo = v3(INFINITY, INFINITY, INFINITY), shi = v3(-INFINITY, -INFINITY, -INFINITY);
    for (uint32_t i = 0; i < n; i++) { slo = v3(fminf(slo.x, lo[i].x), fminf(slo.y, lo[i].y), fminf(slo.z, lo[i].z)); shi = v3(fmaxf(shi.x, hi[i].x), fmaxf(shi.y, hi[i].y), fmaxf(shi.z, hi[i].z)); }
    std::vector<std::pair<uint64_t, uint32_t>> kv(n);
    for (uint32_t i = 0; i < n; i++) kv[i] = {morton_key(lo[i], hi[i], slo, shi), i};
    std::stable_sort(kv.begin(), kv.end(), [](auto& a, auto& b) { return a.first < b.first; });
    std::vector<float4> blo(2 * (size_t)n), bhi(2 * (size_t)n);
    std::vector<uint32_t> bcount(2 * (size_t)n), cl0(n), cl1(n), nearest(n), valid(n);
    Bvh2 b2{blo.data(), bhi.data(), bcount.data()};
    for (uint32_t i = 0; i < n; i++) { float4 l = lo[kv[i].second], h = hi[kv[i].second]; l.w = u2f(0xFFFFFFFFu); h.w = u2f(0xFFFFFFFFu); blo[i] = l; bhi[i] = h; bcount[i] = 1; cl0[i] = i; }
    uint32_t node_counter = n, cur = n;
    while (cur > 1) {
        for (uint32_t i = 0; i < cur; i++) nearest[i] = ploc_nearest(b2, cl0.data(), cur, i, RT2_PLOC_RADIUS_TRIANGLES);
        for (uint32_t i = 0; i < cur; i++) valid[i] = ploc_merge(b2, cl0.data(), nearest.data(), cur, i, &node_counter);
        uint32_t m = 0;
        for (uint32_t i = 0; i < cur; i++) if (valid[i]) cl1[m++] = cl0[i];
        cur = m; std::swap(cl0, cl1);
    }
    const uint32_t root = cl0[0], N = node_counter;
    std::vector<uint32_t> parent(N, 0xFFFFFFFFu);
    for (uint32_t i = n; i < N; i++) { parent[f2u(blo[i].w)] = i; parent[f2u(bhi[i].w)] = i; }
    auto sah = [&]() { double s = 0; for (uint32_t i = n; i < N; i++) s += bvh2_half_area(b2, i); return s / bvh2_half_area(b2, root); };
    auto check = [&]() -> bool {
        std::vector<uint8_t> seen(N, 0); std::vector<uint32_t> st{root}; uint32_t reached = 0, leaves = 0;
        if (parent[root] != 0xFFFFFFFFu) return false;
        while (!st.empty()) {
            const uint32_t x = st.back(); st.pop_back();
            if (x >= N || seen[x]) return false;
            seen[x] = 1; reached++;
            if (x < n) { leaves++; if (bcount[x] != 1) return false; continue; }
            const uint32_t l = f2u(blo[x].w), r = f2u(bhi[x].w);
            if (l >= N || r >= N || l == r || parent[l] != x || parent[r] != x) return false;
            if (bcount[x] != bcount[l] + bcount[r]) return false;
            const float4 ll = blo[l], lh = bhi[l], rl = blo[r], rh = bhi[r];
            if (blo[x].x != fminf(ll.x, rl.x) || blo[x].y != fminf(ll.y, rl.y) || blo[x].z != fminf(ll.z, rl.z)) return false;
            if (bhi[x].x != fmaxf(lh.x, rh.x) || bhi[x].y != fmaxf(lh.y, rh.y) || bhi[x].z != fmaxf(lh.z, rh.z)) return false;
            st.push_back(l); st.push_back(r);
        }
        return reached == N && leaves == n && bcount[root] == n;
    };
    out[0] = sah(); out[2] = 0; out[3] = check() ? 1.0 : 0.0;
    std::vector<uint32_t> target(N), lca(N); std::vector<float> gain(N); std::vector<unsigned long long> lock(N), pass(N); std::vector<uint8_t> win(N);
    Reinsert r{b2, parent.data(), root, n, N, target.data(), lca.data(), gain.data(), lock.data(), pass.data()};
    for (int it = 0; it < iters && out[3] == 1.0; it++) {
        std::fill(lock.begin(), lock.end(), 0ull); std::fill(pass.begin(), pass.end(), 0ull);
        for (uint32_t i = 0; i < N; i++) reinsert_find(r, i);
        for (uint32_t i = 0; i < N; i++) reinsert_lock(r, i);
        for (uint32_t i = 0; i < N; i++) win[i] = reinsert_check(r, i) ? 1 : 0;
        for (uint32_t i = 0; i < N; i++) if (win[i]) { reinsert_apply(r, i, target[i]); out[2] += 1; }
        std::vector<uint32_t> order; std::vector<uint32_t> st{root};
        while (!st.empty() && order.size() <= N) { uint32_t x = st.back(); st.pop_back(); order.push_back(x); if (x >= n && x < N) { st.push_back(f2u(blo[x].w)); st.push_back(f2u(bhi[x].w)); } }
        if (order.size() != N) { out[3] = 0.0; break; }
        for (size_t k = order.size(); k-- > 0;) if (order[k] >= n) reinsert_refit_node(b2, order[k]);
        if (!check()) out[3] = 0.0;
    }
    out[1] = sah();
}
void hostemu_info(void* p, uint32_t* nodes, uint32_t* depth, uint32_t* tlas_nodes) {
    HostScene* hs = (HostScene*)p; *nodes = (uint32_t)hs->blas.nodes.size(); *depth = hs->blas.depth; *tlas_nodes = (uint32_t)hs->tlas_nodes.size();
}

// same math as setup_primary_ray in rt2_render.cu
static void primary_ray(const rt2_planar_view_constants& view, uint32_t px, uint32_t py, V3* o, V3* d) {
    const float pcx = (float)px + 0.5f, pcy = (float)py + 0.5f;
    const float ux = pcx / view.viewport_size[0], uy = pcy / view.viewport_size[1];
    const float dx = ux * 2.0f - 1.0f, dy = uy * 2.0f - 1.0f;
    float target[4]; mat4_mul_vec4(view.mat_clip_to_view, dx, dy, 1.0f, 1.0f, target);
    const V3 nt = normalize3(v3(target[0], target[1], target[2]));
    float dir[4]; mat4_mul_vec4(view.mat_view_to_world, nt.x, nt.y, nt.z, 0.0f, dir);
    *o = v3(view.camera_direction_or_position[0], view.camera_direction_or_position[1], view.camera_direction_or_position[2]);
    *d = v3(dir[0], dir[1], dir[2]);
}

// arbitrary rays: o[3], d[3] per ray
void hostemu_trace_rays(void* p, const float* rays, uint32_t n, float tmin, float tmax, rt2_hit_record* out, uint64_t* node_visits) {
    HostScene* hs = (HostScene*)p;
    uint32_t visits = 0; uint64_t total = 0;
    for (uint32_t i = 0; i < n; i++) {
        TraceRay ray; ray.o = v3(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]); ray.d = v3(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]);
        ray.tmin = tmin; ray.tmax = tmax;
        Hit best; best.t = INFINITY; best.u = best.v = 0.f; best.inst = best.geom = best.prim = 0xFFFFFFFFu;
        visits = 0;
        if (!hs->instances.empty()) tlas_closest_hit(hs->tlas_nodes.data(), hs->tlas_slot_to_inst.data(), hs->instances.data(), hs->blas.nodes.data(), hs->blas.tris.data(), ray, best, nullptr, &visits);
        else cwbvh_closest_hit(hs->blas.nodes.data(), hs->blas.tris.data(), ray, 0u, best, nullptr, &visits);
        total += visits;
        rt2_hit_record r; memset(&r, 0, sizeof r);
        r.instance = best.inst; r.geometry = best.geom; r.primitive = best.prim;
        if (best.geom != 0xFFFFFFFFu) { r.t = best.t; r.u = best.u; r.v = best.v; }
        out[i] = r;
    }
    if (node_visits) *node_visits = total;
}

void hostemu_trace_primary(void* p, const void* gconst, uint32_t w, uint32_t h, rt2_hit_record* out, uint64_t* node_visits) {
    rt2_gconst g; memcpy(&g, gconst, sizeof g);
    std::vector<float> rays((size_t)w * h * 6);
    for (uint32_t y = 0; y < h; y++) for (uint32_t x = 0; x < w; x++) {
        V3 o, d; primary_ray(g.view, x, y, &o, &d);
        float* r = &rays[((size_t)y * w + x) * 6];
        r[0] = o.x; r[1] = o.y; r[2] = o.z; r[3] = d.x; r[4] = d.y; r[5] = d.z;
    }
    hostemu_trace_rays(p, rays.data(), w * h, 0.001f, 100000.0f, out, node_visits);
}

}  // extern "C"

// traverse an externally supplied (device-built) single-level CWBVH on the CPU
extern "C" void hostemu_trace_external(const void* nodes, const void* tris, const void* gconst, uint32_t w, uint32_t h, rt2_hit_record* out) {
    rt2_gconst g; memcpy(&g, gconst, sizeof g);
    for (uint32_t y = 0; y < h; y++) for (uint32_t x = 0; x < w; x++) {
        TraceRay ray; primary_ray(g.view, x, y, &ray.o, &ray.d); ray.tmin = 0.001f; ray.tmax = 100000.0f;
        Hit best; best.t = INFINITY; best.u = best.v = 0.f; best.inst = best.geom = best.prim = 0xFFFFFFFFu;
        cwbvh_closest_hit((const Node8*)nodes, (const TriPos*)tris, ray, 0u, best, nullptr, nullptr);
        rt2_hit_record r; memset(&r, 0, sizeof r);
        r.instance = best.inst; r.geometry = best.geom; r.primitive = best.prim;
        if (best.geom != 0xFFFFFFFFu) { r.t = best.t; r.u = best.u; r.v = best.v; }
        out[(size_t)y * w + x] = r;
    }
}
