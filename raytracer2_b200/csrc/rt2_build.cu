// rt2_build.cu — device BVH builder: kernels + host orchestration.  See rt2_build.cuh.
#include "rt2_internal.cuh"
#include "rt2_build.cuh"
#include "rt2_reinsert.cuh"

#include <cstdio>
#include <vector>

namespace rt2 {

// ------------------------------------------------------------------------------------------------
// small device utilities: order-preserving float<->uint map, exclusive scan, radix sort
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t float_to_ordered(float f) { uint32_t u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
static inline float ordered_to_float_host(uint32_t u) { uint32_t v = (u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u; float f; memcpy(&f, &v, 4); return f; }

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

// exclusive scan of one tile per block; block totals go to block_sums (if non-null)
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_tiles(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, uint32_t n, uint32_t* __restrict__ block_sums) {
    __shared__ uint32_t warp_sums[SCAN_THREADS / 32];
    const uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    uint32_t v[SCAN_ITEMS]; uint32_t sum = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) { uint32_t idx = base + k; v[k] = idx < n ? in[idx] : 0u; sum += v[k]; }
    // inclusive warp scan of per-thread sums
    uint32_t inc = sum;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, d); if (lane >= (uint32_t)d) inc += t; }
    if (lane == 31u) warp_sums[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = lane < SCAN_THREADS / 32 ? warp_sums[lane] : 0u;
        uint32_t winc = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { uint32_t t = __shfl_up_sync(0xFFFFFFFFu, winc, d); if (lane >= (uint32_t)d) winc += t; }
        if (lane < SCAN_THREADS / 32) warp_sums[lane] = winc - w;          // exclusive warp offsets
        if (lane == SCAN_THREADS / 32 - 1 && block_sums) block_sums[blockIdx.x] = winc;
    }
    __syncthreads();
    uint32_t run = warp_sums[warp] + (inc - sum);
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) { uint32_t idx = base + k; if (idx < n) out[idx] = run; run += v[k]; }
}
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_add(uint32_t* __restrict__ out, uint32_t n, const uint32_t* __restrict__ block_offsets) {
    const uint32_t off = block_offsets[blockIdx.x];
    const uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) { uint32_t idx = base + k; if (idx < n) out[idx] += off; }
}

// scratch must hold scan_scratch_elems(n) uint32
static size_t scan_scratch_elems(size_t n) {
    size_t total = 0;
    while (n > 1) { size_t b = (n + SCAN_TILE - 1) / SCAN_TILE; total += 2 * b; n = b; if (b == 1) break; }
    return total + 8;
}
static void exclusive_scan(cudaStream_t st, const uint32_t* in, uint32_t* out, uint32_t n, uint32_t* scratch, LaunchCounter* lc) {
    if (n == 0) return;
    const uint32_t blocks = (n + SCAN_TILE - 1) / SCAN_TILE;
    if (blocks == 1) { k_scan_tiles<<<1, SCAN_THREADS, 0, st>>>(in, out, n, nullptr); lc->add(1); return; }
    uint32_t* sums = scratch; uint32_t* sums_scanned = scratch + blocks;
    k_scan_tiles<<<blocks, SCAN_THREADS, 0, st>>>(in, out, n, sums); lc->add(1);
    exclusive_scan(st, sums, sums_scanned, blocks, scratch + 2 * blocks, lc);
    k_scan_add<<<blocks, SCAN_THREADS, 0, st>>>(out, n, sums_scanned); lc->add(1);
}

// ---- LSD radix sort, 8-bit digits, 64-bit keys + 32-bit values ----------------------------------------------
constexpr int SORT_THREADS = 256;
constexpr int SORT_ITEMS = 8;
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS;

__global__ void __launch_bounds__(SORT_THREADS) k_sort_hist(const uint64_t* __restrict__ keys, uint32_t n, uint32_t shift, uint32_t* __restrict__ hist, uint32_t num_tiles) {
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t base = blockIdx.x * SORT_TILE;
#pragma unroll
    for (int k = 0; k < SORT_ITEMS; k++) {
        uint32_t idx = base + k * SORT_THREADS + threadIdx.x;
        if (idx < n) atomicAdd(&h[(uint32_t)(keys[idx] >> shift) & 0xFFu], 1u);
    }
    __syncthreads();
    hist[threadIdx.x * num_tiles + blockIdx.x] = h[threadIdx.x];
}
__global__ void __launch_bounds__(SORT_THREADS) k_sort_scatter(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                                                              uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                                                              uint32_t n, uint32_t shift, const uint32_t* __restrict__ hist_scanned, uint32_t num_tiles) {
    __shared__ uint32_t base[256];
    __shared__ uint32_t warp_cnt[SORT_THREADS / 32][256];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    base[threadIdx.x] = hist_scanned[threadIdx.x * num_tiles + blockIdx.x];
    const uint32_t tile_base = blockIdx.x * SORT_TILE;
    for (int k = 0; k < SORT_ITEMS; k++) {
#pragma unroll
        for (int w = 0; w < SORT_THREADS / 32; w++) warp_cnt[w][threadIdx.x] = 0;
        __syncthreads();
        const uint32_t idx = tile_base + k * SORT_THREADS + threadIdx.x;
        const bool valid = idx < n;
        uint64_t key = 0; uint32_t val = 0;
        if (valid) { key = keys_in[idx]; val = vals_in[idx]; }
        const uint32_t d = valid ? ((uint32_t)(key >> shift) & 0xFFu) : 256u;
        const uint32_t peers = __match_any_sync(0xFFFFFFFFu, d);
        const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
        if (valid && rank == 0) warp_cnt[warp][d] = __popc(peers);
        __syncthreads();
        {   // thread t owns digit t: exclusive prefix over warps, advance the running base
            uint32_t run = base[threadIdx.x];
#pragma unroll
            for (int w = 0; w < SORT_THREADS / 32; w++) { uint32_t c = warp_cnt[w][threadIdx.x]; warp_cnt[w][threadIdx.x] = run; run += c; }
            base[threadIdx.x] = run;
        }
        __syncthreads();
        if (valid) { uint32_t pos = warp_cnt[warp][d] + rank; keys_out[pos] = key; vals_out[pos] = val; }
        __syncthreads();
    }
}

// ---- builder kernels --------------------------------------------------------------------------------------
__global__ void k_build_triangles(BuildInput in, TriPos* tris, float4* lo, float4* hi, uint32_t* scene_bounds /* 6 ordered uints */) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    float l[3] = {INFINITY, INFINITY, INFINITY}, h[3] = {-INFINITY, -INFINITY, -INFINITY};
    if (i < in.num_tris) {
        build_triangle(in, i, tris, lo, hi);
        float4 a = lo[i], b = hi[i];
        l[0] = a.x; l[1] = a.y; l[2] = a.z; h[0] = b.x; h[1] = b.y; h[2] = b.z;
    }
#pragma unroll
    for (int k = 0; k < 3; k++) {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) { l[k] = fminf(l[k], __shfl_xor_sync(0xFFFFFFFFu, l[k], d)); h[k] = fmaxf(h[k], __shfl_xor_sync(0xFFFFFFFFu, h[k], d)); }
    }
    if ((threadIdx.x & 31u) == 0) {
#pragma unroll
        for (int k = 0; k < 3; k++) {
            if (l[k] < INFINITY) atomicMin(&scene_bounds[k], float_to_ordered(l[k]));
            if (h[k] > -INFINITY) atomicMax(&scene_bounds[3 + k], float_to_ordered(h[k]));
        }
    }
}
// generic: boxes already given (TLAS): only reduce the bounds
__global__ void k_reduce_bounds(const float4* lo, const float4* hi, uint32_t n, uint32_t* scene_bounds) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    float l[3] = {INFINITY, INFINITY, INFINITY}, h[3] = {-INFINITY, -INFINITY, -INFINITY};
    if (i < n) { float4 a = lo[i], b = hi[i]; l[0] = a.x; l[1] = a.y; l[2] = a.z; h[0] = b.x; h[1] = b.y; h[2] = b.z; }
#pragma unroll
    for (int k = 0; k < 3; k++) {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) { l[k] = fminf(l[k], __shfl_xor_sync(0xFFFFFFFFu, l[k], d)); h[k] = fmaxf(h[k], __shfl_xor_sync(0xFFFFFFFFu, h[k], d)); }
    }
    if ((threadIdx.x & 31u) == 0) {
#pragma unroll
        for (int k = 0; k < 3; k++) {
            if (l[k] < INFINITY) atomicMin(&scene_bounds[k], float_to_ordered(l[k]));
            if (h[k] > -INFINITY) atomicMax(&scene_bounds[3 + k], float_to_ordered(h[k]));
        }
    }
}
__global__ void k_morton(const float4* lo, const float4* hi, uint32_t n, V3 slo, V3 shi, uint64_t* keys, uint32_t* vals) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { keys[i] = morton_key(lo[i], hi[i], slo, shi); vals[i] = i; }
}
// leaves of the BVH2 in sorted order
__global__ void k_init_leaves(const float4* lo, const float4* hi, const uint32_t* sorted_prim, uint32_t n, Bvh2 b, uint32_t* clusters) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        uint32_t p = sorted_prim[i];
        float4 l = lo[p], h = hi[p];
        l.w = __uint_as_float(0xFFFFFFFFu); h.w = __uint_as_float(0xFFFFFFFFu);
        b.lo[i] = l; b.hi[i] = h; b.count[i] = 1u; clusters[i] = i;
    }
}
__global__ void k_ploc_nearest(Bvh2 b, const uint32_t* clusters, uint32_t n, uint32_t* nearest, uint32_t radius) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) nearest[i] = ploc_nearest(b, clusters, n, i, radius);
}
__global__ void k_ploc_merge(Bvh2 b, uint32_t* clusters, const uint32_t* nearest, uint32_t n, uint32_t* node_counter, uint32_t* valid) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) valid[i] = ploc_merge(b, clusters, nearest, n, i, node_counter);
}
__global__ void k_compact(const uint32_t* clusters, const uint32_t* valid, const uint32_t* pos, uint32_t n, uint32_t* out, uint32_t* total) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        if (valid[i]) out[pos[i]] = clusters[i];
        if (i == n - 1) *total = pos[i] + valid[i];
    }
}
// SAH-optimal collapse, bottom-up: a thread per leaf climbs; the second thread to arrive at an inner node solves it
__global__ void k_bvh2_parents(Bvh2 b, uint32_t num_leaves, uint32_t num_nodes, uint32_t* parent) {
    uint32_t i = num_leaves + blockIdx.x * blockDim.x + threadIdx.x;
    if (i < num_nodes) { parent[__float_as_uint(b.lo[i].w)] = i; parent[__float_as_uint(b.hi[i].w)] = i; }
}
__global__ void k_dp_bottom_up(Bvh2 b, uint32_t num_leaves, const uint32_t* __restrict__ parent, uint32_t* flags, CollapseDp dp, uint32_t root) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= num_leaves) return;
    dp_solve_leaf(dp, b, i);
    uint32_t cur = i;
    while (cur != root) {
        const uint32_t p = parent[cur];
        __threadfence();
        if (atomicAdd(&flags[p], 1u) == 0u) return;
        dp_solve_inner(dp, b, p);
        cur = p;
    }
}
// insertion-based optimisation of the BVH2 (rt2_reinsert.cuh): one iteration = find, lock, check, apply, refit
__global__ void __launch_bounds__(128) k_reinsert_find(Reinsert r) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < r.num_nodes) { r.lock[i] = 0ull; r.pass[i] = 0ull; reinsert_find(r, i); }
}
__global__ void k_reinsert_lock(Reinsert r) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < r.num_nodes) reinsert_lock(r, i);
}
__global__ void k_reinsert_check(Reinsert r, uint32_t* win, uint32_t* applied) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < r.num_nodes) { const bool w = reinsert_check(r, i); win[i] = w ? 1u : 0u; if (w) atomicAdd(applied, 1u); }
}
__global__ void k_reinsert_apply(Reinsert r, const uint32_t* win) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < r.num_nodes && win[i]) reinsert_apply(r, i, r.target[i]);
}
__global__ void k_refit_bottom_up(Bvh2 b, uint32_t num_leaves, const uint32_t* __restrict__ parent, uint32_t* flags, uint32_t root) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= num_leaves) return;
    uint32_t cur = i;
    while (cur != root) {
        const uint32_t p = parent[cur];
        __threadfence();
        if (atomicAdd(&flags[p], 1u) == 0u) return;
        reinsert_refit_node(b, p);
        cur = p;
    }
}
__global__ void k_collapse(CollapseCtx c, const uint2* queue_in, uint32_t count) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) { uint2 q = queue_in[i]; collapse_node(c, q.x, q.y); }
}
__global__ void k_permute_tris(const TriPos* in, const uint32_t* out_prim, uint32_t n, TriPos* out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = in[out_prim[i]];
}

// ------------------------------------------------------------------------------------------------
// host orchestration
// ------------------------------------------------------------------------------------------------
#define BCK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { err = std::string(#x) + ": " + cudaGetErrorString(e_); goto fail; } } while (0)

struct Scratch {
    std::vector<void*> ptrs;
    template <class T> cudaError_t alloc(T** p, size_t n) { cudaError_t e = cudaMalloc((void**)p, std::max<size_t>(n, 1) * sizeof(T)); if (e == cudaSuccess) ptrs.push_back(*p); return e; }
    ~Scratch() { for (void* p : ptrs) cudaFree(p); }
};

// Build a CWBVH over `n` boxed primitives (lo/hi on device). Produces nodes (cudaMalloc'd, owned by
// caller) and out_prim (slot -> primitive id, cudaMalloc'd, owned by caller).
int build_cwbvh_from_boxes(cudaStream_t st, const float4* lo, const float4* hi, uint32_t n, Node8** nodes_out, uint32_t* num_nodes_out,
                           uint32_t** out_prim_out, uint32_t* depth_out, LaunchCounter* lc, std::string& err, uint32_t ploc_radius) {
    Scratch sc;
    Node8* nodes = nullptr; uint32_t* out_prim = nullptr;
    const uint32_t T = 256;
    auto grid = [&](uint32_t m) { return (m + T - 1) / T; };
    uint32_t* bounds = nullptr; uint64_t *keys0 = nullptr, *keys1 = nullptr; uint32_t *vals0 = nullptr, *vals1 = nullptr;
    uint32_t *hist = nullptr, *hist_scanned = nullptr, *scan_scratch = nullptr;
    Bvh2 b2{}; uint32_t *cl0 = nullptr, *cl1 = nullptr, *nearest = nullptr, *valid = nullptr, *pos = nullptr, *counters = nullptr;
    uint2 *q0 = nullptr, *q1 = nullptr;
    CollapseDp dp{}; uint32_t *dp_parent = nullptr, *dp_flags = nullptr;
    static const bool use_dp = [] { const char* e = getenv("RT2_COLLAPSE_DP"); return !e || atoi(e) != 0; }();
    static const bool build_verbose = getenv("RT2_BUILD_VERBOSE") != nullptr;
    static const int reinsert_iters = [] { const char* e = getenv("RT2_REINSERT_ITERS"); return e ? atoi(e) : 12; }();
    uint32_t depth = 0;
    *nodes_out = nullptr; *out_prim_out = nullptr; *num_nodes_out = 0;

    if (n == 0) {   // empty scene: one node with no children -> every ray misses
        BCK(cudaMalloc((void**)&nodes, sizeof(Node8)));
        BCK(cudaMemsetAsync(nodes, 0, sizeof(Node8), st));
        BCK(cudaMalloc((void**)&out_prim, sizeof(uint32_t)));
        *nodes_out = nodes; *out_prim_out = out_prim; *num_nodes_out = 1; if (depth_out) *depth_out = 1;
        return RT2_OK;
    }
    {
        const uint32_t tiles = (n + SORT_TILE - 1) / SORT_TILE;
        const size_t hist_n = (size_t)tiles * 256;
        BCK(sc.alloc(&bounds, 8));
        BCK(sc.alloc(&keys0, n)); BCK(sc.alloc(&keys1, n)); BCK(sc.alloc(&vals0, n)); BCK(sc.alloc(&vals1, n));
        BCK(sc.alloc(&hist, hist_n)); BCK(sc.alloc(&hist_scanned, hist_n));
        BCK(sc.alloc(&scan_scratch, scan_scratch_elems(std::max<size_t>(hist_n, n))));
        BCK(sc.alloc(&b2.lo, 2 * (size_t)n)); BCK(sc.alloc(&b2.hi, 2 * (size_t)n)); BCK(sc.alloc(&b2.count, 2 * (size_t)n));
        BCK(sc.alloc(&cl0, n)); BCK(sc.alloc(&cl1, n)); BCK(sc.alloc(&nearest, n)); BCK(sc.alloc(&valid, n)); BCK(sc.alloc(&pos, n));
        BCK(sc.alloc(&counters, 8));
        BCK(sc.alloc(&q0, n)); BCK(sc.alloc(&q1, n));

        // scene bounds
        uint32_t init_bounds[8] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u, 0u, 0u, 0u};
        BCK(cudaMemcpyAsync(bounds, init_bounds, sizeof init_bounds, cudaMemcpyHostToDevice, st));
        k_reduce_bounds<<<grid(n), T, 0, st>>>(lo, hi, n, bounds); lc->add(1);
        uint32_t hb[8];
        BCK(cudaMemcpyAsync(hb, bounds, sizeof hb, cudaMemcpyDeviceToHost, st));
        BCK(cudaStreamSynchronize(st));
        V3 slo = v3(ordered_to_float_host(hb[0]), ordered_to_float_host(hb[1]), ordered_to_float_host(hb[2]));
        V3 shi = v3(ordered_to_float_host(hb[3]), ordered_to_float_host(hb[4]), ordered_to_float_host(hb[5]));
        // Morton keys + sort
        k_morton<<<grid(n), T, 0, st>>>(lo, hi, n, slo, shi, keys0, vals0); lc->add(1);
        for (uint32_t shift = 0; shift < 64; shift += 8) {
            k_sort_hist<<<tiles, SORT_THREADS, 0, st>>>(keys0, n, shift, hist, tiles); lc->add(1);
            exclusive_scan(st, hist, hist_scanned, (uint32_t)hist_n, scan_scratch, lc);
            k_sort_scatter<<<tiles, SORT_THREADS, 0, st>>>(keys0, vals0, keys1, vals1, n, shift, hist_scanned, tiles); lc->add(1);
            std::swap(keys0, keys1); std::swap(vals0, vals1);
        }
        // PLOC
        k_init_leaves<<<grid(n), T, 0, st>>>(lo, hi, vals0, n, b2, cl0); lc->add(1);
        uint32_t hc[8] = {n /* next bvh2 node id */, 0, 0, 0, 0, 0, 0, 0};
        BCK(cudaMemcpyAsync(counters, hc, sizeof hc, cudaMemcpyHostToDevice, st));
        uint32_t cur = n;
        int guard = 0;
        while (cur > 1) {
            k_ploc_nearest<<<grid(cur), T, 0, st>>>(b2, cl0, cur, nearest, ploc_radius); lc->add(1);
            k_ploc_merge<<<grid(cur), T, 0, st>>>(b2, cl0, nearest, cur, counters + 0, valid); lc->add(1);
            exclusive_scan(st, valid, pos, cur, scan_scratch, lc);
            k_compact<<<grid(cur), T, 0, st>>>(cl0, valid, pos, cur, cl1, counters + 1); lc->add(1);
            uint32_t next = 0;
            BCK(cudaMemcpyAsync(&next, counters + 1, 4, cudaMemcpyDeviceToHost, st));
            BCK(cudaStreamSynchronize(st));
            if (next >= cur || next == 0) { err = "PLOC made no progress"; goto fail; }
            cur = next; std::swap(cl0, cl1);
            if (++guard > 4096) { err = "PLOC iteration guard"; goto fail; }
        }
        uint32_t root = 0;
        BCK(cudaMemcpyAsync(&root, cl0, 4, cudaMemcpyDeviceToHost, st));
        BCK(cudaStreamSynchronize(st));
        const uint32_t num_b2 = 2 * n - 1;                              // PLOC allocates inner nodes [n, 2n - 1)
        if (use_dp || reinsert_iters > 0) {
            BCK(sc.alloc(&dp_parent, num_b2)); BCK(sc.alloc(&dp_flags, num_b2));
            BCK(cudaMemsetAsync(dp_parent, 0xFF, (size_t)num_b2 * 4, st));
            if (n > 1) { k_bvh2_parents<<<grid(n - 1), T, 0, st>>>(b2, n, num_b2, dp_parent); lc->add(1); }
        }
        // insertion-based optimisation of the BVH2 (rt2_reinsert.cuh)
        if (reinsert_iters > 0 && n >= 16) {
            Reinsert r{}; uint32_t* win = nullptr;
            r.b = b2; r.parent = dp_parent; r.root = root; r.num_leaves = n; r.num_nodes = num_b2;
            BCK(sc.alloc(&r.target, num_b2)); BCK(sc.alloc(&r.lca, num_b2)); BCK(sc.alloc(&r.gain, num_b2));
            BCK(sc.alloc(&r.lock, num_b2)); BCK(sc.alloc(&r.pass, num_b2)); BCK(sc.alloc(&win, num_b2));
            // big trees: fewer rounds (a round costs a branch-and-bound search per node)
            const int iters = n > (4u << 20) ? std::max(1, reinsert_iters / 2) : reinsert_iters;
            for (int it = 0; it < iters; it++) {
                BCK(cudaMemsetAsync(counters + 5, 0, 4, st));
                k_reinsert_find<<<(num_b2 + 127) / 128, 128, 0, st>>>(r); lc->add(1);
                k_reinsert_lock<<<grid(num_b2), T, 0, st>>>(r); lc->add(1);
                k_reinsert_check<<<grid(num_b2), T, 0, st>>>(r, win, counters + 5); lc->add(1);
                k_reinsert_apply<<<grid(num_b2), T, 0, st>>>(r, win); lc->add(1);
                BCK(cudaMemsetAsync(dp_flags, 0, (size_t)num_b2 * 4, st));
                k_refit_bottom_up<<<grid(n), T, 0, st>>>(b2, n, dp_parent, dp_flags, root); lc->add(1);
                uint32_t applied = 0;
                BCK(cudaMemcpyAsync(&applied, counters + 5, 4, cudaMemcpyDeviceToHost, st));
                BCK(cudaStreamSynchronize(st));
                if (build_verbose) fprintf(stderr, "[rt2 build] %u leaves, reinsertion round %d: %u moves applied\n", n, it, applied);
                if (applied * 2000ull < n) break;                       // fewer than 0.05 % of the leaves moved: converged
            }
        }
        // SAH-optimal slot distribution (rt2_build.cuh 5a): one bottom-up pass over the BVH2
        if (use_dp) {
            BCK(sc.alloc(&dp.cost, (size_t)num_b2 * 7)); BCK(sc.alloc(&dp.dec, num_b2));
            BCK(cudaMemsetAsync(dp_flags, 0, (size_t)num_b2 * 4, st));
            k_dp_bottom_up<<<grid(n), T, 0, st>>>(b2, n, dp_parent, dp_flags, dp, root); lc->add(1);
        }
        // collapse to 8-wide
        const uint32_t max_nodes = n + 1;   // every wide node holds >= 1 primitive-bearing or >= 2 children; n+1 is a safe bound
        BCK(cudaMalloc((void**)&nodes, (size_t)max_nodes * sizeof(Node8)));
        BCK(cudaMalloc((void**)&out_prim, (size_t)n * sizeof(uint32_t)));
        uint32_t hc2[8] = {0, 0, 1 /* wide node counter */, 0 /* prim counter */, 0 /* queue count */, 0, 0, 0};
        BCK(cudaMemcpyAsync(counters, hc2, sizeof hc2, cudaMemcpyHostToDevice, st));
        uint2 first = make_uint2(root, 0u);
        BCK(cudaMemcpyAsync(q0, &first, sizeof first, cudaMemcpyHostToDevice, st));
        uint32_t qn = 1;
        CollapseCtx cc{};
        cc.bvh2 = b2; cc.num_leaves = n; cc.leaf_prim = vals0; cc.nodes = nodes; cc.node_counter = counters + 2; cc.prim_counter = counters + 3;
        cc.out_prim = out_prim; cc.queue_out_count = counters + 4; cc.max_nodes = max_nodes; cc.max_prims = n;
        cc.dp_dec = use_dp ? dp.dec : nullptr;
        while (qn) {
            depth++;
            BCK(cudaMemsetAsync(counters + 4, 0, 4, st));
            cc.queue_out = q1;
            k_collapse<<<grid(qn), T, 0, st>>>(cc, q0, qn); lc->add(1);
            BCK(cudaMemcpyAsync(&qn, counters + 4, 4, cudaMemcpyDeviceToHost, st));
            BCK(cudaStreamSynchronize(st));
            std::swap(q0, q1);
            if (depth > RT2_STACK_SIZE - 2) { err = "BVH too deep for the traversal stack"; goto fail; }
        }
        uint32_t fin[2];
        BCK(cudaMemcpyAsync(fin, counters + 2, 8, cudaMemcpyDeviceToHost, st));
        BCK(cudaStreamSynchronize(st));
        if (fin[1] != n) { err = "collapse lost primitives"; goto fail; }
        *num_nodes_out = fin[0];
    }
    BCK(cudaGetLastError());
    *nodes_out = nodes; *out_prim_out = out_prim; if (depth_out) *depth_out = depth;
    return RT2_OK;
fail:
    if (nodes) cudaFree(nodes);
    if (out_prim) cudaFree(out_prim);
    return RT2_E_CUDA;
}

// BLAS: triangles from the uploaded scene
int build_blas(cudaStream_t st, const BuildInput& in, Accel* out, LaunchCounter* lc, std::string& err) {
    Scratch sc;
    TriPos* tris_in = nullptr; float4 *lo = nullptr, *hi = nullptr; uint32_t* bounds = nullptr;
    TriPos* tris = nullptr; Node8* nodes = nullptr; uint32_t* out_prim = nullptr;
    const uint32_t n = in.num_tris, T = 256;
    uint32_t num_nodes = 0, depth = 0;
    int rc;
    BCK(sc.alloc(&tris_in, n)); BCK(sc.alloc(&lo, n)); BCK(sc.alloc(&hi, n)); BCK(sc.alloc(&bounds, 8));
    if (n) {
        uint32_t init_bounds[8] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u, 0u, 0u, 0u};
        BCK(cudaMemcpyAsync(bounds, init_bounds, sizeof init_bounds, cudaMemcpyHostToDevice, st));
        k_build_triangles<<<(n + T - 1) / T, T, 0, st>>>(in, tris_in, lo, hi, bounds); lc->add(1);
    }
    rc = build_cwbvh_from_boxes(st, lo, hi, n, &nodes, &num_nodes, &out_prim, &depth, lc, err, RT2_PLOC_RADIUS_TRIANGLES);
    if (rc != RT2_OK) return rc;
    BCK(cudaMalloc((void**)&tris, std::max<size_t>(n, 1) * sizeof(TriPos)));
    if (n) { k_permute_tris<<<(n + T - 1) / T, T, 0, st>>>(tris_in, out_prim, n, tris); lc->add(1); }
    BCK(cudaStreamSynchronize(st));
    BCK(cudaGetLastError());
    {
        // BLAS root box (needed for instance boxes): read the unsorted bounds
        uint32_t hb[8] = {0};
        if (n) BCK(cudaMemcpy(hb, bounds, sizeof hb, cudaMemcpyDeviceToHost));
        for (int k = 0; k < 3; k++) { out->box_lo[k] = n ? ordered_to_float_host(hb[k]) : 0.f; out->box_hi[k] = n ? ordered_to_float_host(hb[3 + k]) : 0.f; }
    }
    out->nodes = nodes; out->tris = tris; out->slot_to_prim = out_prim; out->num_nodes = num_nodes; out->num_prims = n; out->depth = depth;
    return RT2_OK;
fail:
    if (nodes) cudaFree(nodes);
    if (out_prim) cudaFree(out_prim);
    if (tris) cudaFree(tris);
    return RT2_E_CUDA;
}

}  // namespace rt2
