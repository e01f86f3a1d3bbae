// rt2_noalloc.cu — rt2_trace_primary_noalloc: the primary-visibility probe for callers that cannot call cudaMalloc (code
// under CUDA graph capture, tight serving loops, harnesses that forbid allocation).  No counterpart in the
// reference.
//
// rt2_trace_primary (rt2_render.cu) writes into a context-owned buffer and counts hits with atomics into scratch that
// ensure_counters may cudaMalloc on the first call.  This variant allocates nothing: the caller passes the output buffer (and,
// optionally, a device word for the hit count), and the kernel's scratch — one hit count per warp, reduced by a second
// launch — lives in a file-scope `static __device__` array of fixed size.  Every GPU that loads the library holds that array
// (RT2_NOALLOC_SCRATCH_BYTES, about 1 MiB) until the process exits.
//
// The array's name is a host-side stand-in in host code, not a device address: its address is taken once per device with
// cudaGetSymbolAddress at set-up (noalloc_init, called by rt2_create with that device current) and passed to the kernels.
//
// Sharing: every call on a device uses the same array.  Nothing in it has to start at zero — each launch overwrites every
// slot it later reads (one per warp of the launch, then the total) — but the array keeps the last call's contents, and two
// calls whose kernels overlap on the GPU (different streams / contexts of one device) race on it.  Calls on one stream are
// ordered and safe.
#include "rt2_internal.cuh"
#include "rt2_surface.cuh"

#include <mutex>

namespace rt2 {

// Largest launch: 2^23 padded lanes (8x4-pixel tiles x 32) — covers 3840x2160 (8 294 400 lanes).  One u32 per warp + the total.
#define RT2_NOALLOC_MAX_LANES (1u << 23)
#define RT2_NOALLOC_WARP_SLOTS (RT2_NOALLOC_MAX_LANES / 32u)
#define RT2_NOALLOC_SCRATCH_WORDS (RT2_NOALLOC_WARP_SLOTS + 64u)      /* [0, WARP_SLOTS): per-warp hit counts; [WARP_SLOTS]: u64 total */
#define RT2_NOALLOC_SCRATCH_BYTES (RT2_NOALLOC_SCRATCH_WORDS * 4u)

static __device__ __align__(256) uint32_t g_noalloc_scratch[RT2_NOALLOC_SCRATCH_WORDS];

// device address of g_noalloc_scratch per CUDA device, resolved once per device
static std::mutex g_noalloc_mutex;
static uint32_t* g_noalloc_ptr[64] = {};

int noalloc_init(rt2_ctx* ctx) {
    if (ctx->device < 0 || ctx->device >= 64) return RT2_OK;      // no slot: the entry point reports it
    std::lock_guard<std::mutex> l(g_noalloc_mutex);
    if (g_noalloc_ptr[ctx->device]) return RT2_OK;
    void* p = nullptr;
    RT2_CUDA(ctx, cudaGetSymbolAddress(&p, g_noalloc_scratch));   // the device of ctx is current (rt2_create)
    if (!p) return fail(ctx, RT2_E_CUDA, "cudaGetSymbolAddress returned a null scratch address");
    g_noalloc_ptr[ctx->device] = static_cast<uint32_t*>(p);
    return RT2_OK;
}

// interleaved row bands, as rt2_render.cu's Band
struct NaBand { uint32_t width, band_rows, rank, world, local_rows; };
__device__ __forceinline__ uint32_t na_local_to_global(const NaBand& b, uint32_t ly) {
    if (b.world <= 1u) return ly;
    return ((ly / b.band_rows) * b.world + b.rank) * b.band_rows + (ly % b.band_rows);
}

// k_primary's body (rt2_render.cu) with the output in caller memory and the hit count kept per warp in `scratch`
__global__ void __launch_bounds__(128) k_primary_noalloc(SceneRefs sc, NaBand band, const rt2_gconst* g, rt2_hit_record* out, uint32_t padded,
                                                         uint32_t* scratch) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= padded) return;                                 // padded is a multiple of 32: whole warps leave together
    const uint32_t tiles_x = (band.width + 7u) >> 3;
    const uint32_t tile = i >> 5, within = i & 31u;
    const uint32_t x = (tile % tiles_x) * 8u + (within & 7u), ly = (tile / tiles_x) * 4u + (within >> 3);
    bool hit = false;
    if (x < band.width && ly < band.local_rows) {
        const uint32_t y = na_local_to_global(band, ly);
        V3 o, d;
        setup_primary_ray(g->view, x, y, &o, &d);
        uint32_t tri_slot;
        const Hit best = trace_closest(sc, o, d, RT2_K1_TMIN, RT2_BACKGROUND_DEPTH, &tri_slot);
        rt2_hit_record r;
        r.instance = best.inst; r.geometry = best.geom; r.primitive = best.prim;
        hit = best.geom != 0xFFFFFFFFu;
        r.t = hit ? best.t : 0.f; r.u = hit ? best.u : 0.f; r.v = hit ? best.v : 0.f; r._pad[0] = r._pad[1] = 0u;
        out[(size_t)y * band.width + x] = r;
    }
    const uint32_t m = __ballot_sync(0xFFFFFFFFu, hit);
    if ((threadIdx.x & 31u) == 0u) scratch[i >> 5] = (uint32_t)__popc(m);
}

// one block: sums the per-warp counts of the launch into the total slot (and the caller's word, if given)
__global__ void __launch_bounds__(1024) k_noalloc_reduce(uint32_t* scratch, uint32_t warps, unsigned long long* total_out) {
    __shared__ unsigned long long part[32];
    unsigned long long s = 0;
    for (uint32_t k = threadIdx.x; k < warps; k += blockDim.x) s += scratch[k];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, d);
    if ((threadIdx.x & 31u) == 0u) part[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32u) {
        s = part[threadIdx.x];
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, d);
        if (threadIdx.x == 0u) {
            *reinterpret_cast<unsigned long long*>(scratch + RT2_NOALLOC_WARP_SLOTS) = s;
            if (total_out) *total_out = s;
        }
    }
}

}  // namespace rt2

using namespace rt2;

extern "C" int rt2_trace_primary_noalloc(rt2_ctx* ctx, const rt2_launch_params* params, void* device_hits_out, size_t nbytes,
                                         uint64_t* device_hit_count_out) {
    if (!ctx) return RT2_E_INVALID;
    {
        cudaError_t e = cudaSetDevice(ctx->device);
        if (e != cudaSuccess) return fail(ctx, RT2_E_CUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(e));
    }
    rt2_launch_params lp{};
    if (params) lp = *params;
    if (!ctx->scene.accel_built) return fail(ctx, RT2_E_STATE, "no acceleration structure: call rt2_upload_scene + rt2_build_accel first");
    if (!ctx->fb.width) return fail(ctx, RT2_E_STATE, "no frame size: call rt2_resize first");
    if (!ctx->have_gconst) return fail(ctx, RT2_E_STATE, "no constants: call rt2_set_constants first");
    if (!device_hits_out) return fail(ctx, RT2_E_INVALID, "device_hits_out is NULL");
    const uint32_t w = ctx->fb.width, h = ctx->fb.height;
    if (nbytes != (size_t)w * h * sizeof(rt2_hit_record))
        return fail(ctx, RT2_E_INVALID, "size mismatch: the hit buffer is " + std::to_string((size_t)w * h * sizeof(rt2_hit_record)) + " bytes");
    if (lp.band_world > 1u && (lp.band_rows == 0u || lp.band_rank >= lp.band_world)) return fail(ctx, RT2_E_INVALID, "bad band parameters");
    NaBand band{w, lp.band_rows, lp.band_rank, lp.band_world, h};
    if (band.world <= 1u || band.band_rows == 0u) { band.world = 1; band.rank = 0; band.band_rows = h; }
    else {
        uint32_t rows = 0;
        for (uint32_t y0 = band.rank * band.band_rows; y0 < h; y0 += band.world * band.band_rows) rows += std::min(band.band_rows, h - y0);
        band.local_rows = rows;
    }
    const uint64_t padded64 = (uint64_t)((band.width + 7u) / 8u) * ((band.local_rows + 3u) / 4u) * 32u;
    if (padded64 > RT2_NOALLOC_MAX_LANES)
        return fail(ctx, RT2_E_INVALID, "rt2_trace_primary_noalloc: " + std::to_string(padded64) + " padded pixels exceed the static scratch (maximum " +
                                            std::to_string(RT2_NOALLOC_MAX_LANES) + ", e.g. 3840x2160); use rt2_trace_primary");
    uint32_t* scratch = (ctx->device >= 0 && ctx->device < 64) ? g_noalloc_ptr[ctx->device] : nullptr;
    if (!scratch) return fail(ctx, RT2_E_STATE, "static scratch address not resolved for this device");
    const uint32_t padded = (uint32_t)padded64;
    if (padded) k_primary_noalloc<<<(padded + 127u) / 128u, 128, 0, ctx->stream>>>(make_scene_refs(ctx), band, ctx->d_gconst,
                                                                                  static_cast<rt2_hit_record*>(device_hits_out), padded, scratch);
    k_noalloc_reduce<<<1, 1024, 0, ctx->stream>>>(scratch, padded / 32u, reinterpret_cast<unsigned long long*>(device_hit_count_out));
    RT2_CUDA(ctx, cudaGetLastError());
    return RT2_OK;
}
