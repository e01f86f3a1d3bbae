// rt2_passes_mips.cu — K12: the sampling-PDF mip pyramids (local lights r32f, environment r16f).
// Host loop: src/mip_pass.rs:122-186 (5 levels per dispatch); device: generate_mip_levels/mips.glsl.
//
// One 256-thread block reduces a 32x32 source tile to five levels: a 2x2 quad per thread, two shuffle
// stages inside the warp (threads sit on a Z curve, so four consecutive lanes are a 2x2 block and 16
// lanes a 4x4 block), then 16 partials through shared memory for the last two levels.  The summation
// order of every level is fixed by that tree — it is what makes the chain reproducible bit for bit.
// HBM-bound: each source texel is read once, 4/3 x texels written.  All mips live back to back in one
// allocation (mip k is max(1,w>>k) x max(1,h>>k)).
#include "rt2_internal.cuh"
#include "rt2_codec.cuh"

namespace rt2 {

size_t mip_chain_texels(uint32_t w, uint32_t h, uint32_t mips);
int lights_ensure_resources(rt2_ctx* ctx);

struct MipArgs {
    void* base;
    uint32_t w, h, mips;
    uint32_t offset[32];
    uint32_t source_mip;
    const float4* sky; uint32_t sky_w, sky_h;
};

template <bool F16>
__device__ __forceinline__ float mip_load(const MipArgs& a, uint32_t k, uint32_t x, uint32_t y) {
    const uint32_t mw = max(1u, a.w >> k), mh = max(1u, a.h >> k);
    if (k >= a.mips || x >= mw || y >= mh) return 0.f;      // out-of-range image loads pinned to 0
    const size_t i = (size_t)a.offset[k] + (size_t)y * mw + x;
    return F16 ? f16_bits_to_f32(reinterpret_cast<const uint16_t*>(a.base)[i]) : reinterpret_cast<const float*>(a.base)[i];
}
template <bool F16>
__device__ __forceinline__ void mip_store(const MipArgs& a, uint32_t k, uint32_t x, uint32_t y, float v) {
    const uint32_t mw = max(1u, a.w >> k), mh = max(1u, a.h >> k);
    if (k >= a.mips || x >= mw || y >= mh) return;          // out-of-range image stores are dropped
    const size_t i = (size_t)a.offset[k] + (size_t)y * mw + x;
    if (F16) reinterpret_cast<uint16_t*>(a.base)[i] = f32_to_f16_bits(v); else reinterpret_cast<float*>(a.base)[i] = v;
}

// mips.glsl:45-62 getPixelWeight.  The sky is read through an UNNORMALISED LINEAR sampler at integer
// coordinates (src/main.rs:123-136, Q15): a 2x2 box around the texel corner, CLAMP_TO_BORDER black.
__device__ __forceinline__ float env_pixel_weight(const MipArgs& a, uint32_t px, uint32_t py) {
    const int W = (int)a.sky_w, H = (int)a.sky_h;
    const float x = (float)px - 0.5f, y = (float)py - 0.5f;
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
    const V3 color = top * ib + bot * wb;
    const float luma = fmaxf(calc_luminance(color), 0.f);
    if (isinf(luma) || isnan(luma)) return 0.f;
    const float elevation = (((float)py + 0.5f) / (float)a.h - 0.5f) * 3.1415926535f;
    return clampf(luma * cos_pinned(elevation), 0.f, 65504.0f);
}

template <bool F16, bool ENV>
__global__ void __launch_bounds__(256) k_mip_pass(MipArgs a) {
    __shared__ float s_weights[16];
    const uint32_t t = threadIdx.x, lane = t & 31u;
    uint32_t lx, ly;
    linear_to_zcurve(t, lx, ly);
    const uint32_t gx = blockIdx.x * 16u + lx, gy = blockIdx.y * 16u + ly;
    const uint32_t sx = gx * 2u, sy = gy * 2u, sm = a.source_mip;
    float w0, w1, w2, w3;
    if (ENV && sm == 0u) {
        w0 = env_pixel_weight(a, sx, sy); w1 = env_pixel_weight(a, sx, sy + 1u);
        w2 = env_pixel_weight(a, sx + 1u, sy); w3 = env_pixel_weight(a, sx + 1u, sy + 1u);
        mip_store<F16>(a, 0u, sx, sy, w0); mip_store<F16>(a, 0u, sx, sy + 1u, w1);
        mip_store<F16>(a, 0u, sx + 1u, sy, w2); mip_store<F16>(a, 0u, sx + 1u, sy + 1u, w3);
    } else {
        w0 = mip_load<F16>(a, sm, sx, sy); w1 = mip_load<F16>(a, sm, sx, sy + 1u);
        w2 = mip_load<F16>(a, sm, sx + 1u, sy); w3 = mip_load<F16>(a, sm, sx + 1u, sy + 1u);
    }
    const uint32_t to_write = a.mips - sm - 1u;
    if (to_write < 1u) return;
    float weight = (((w0 + w1) + w2) + w3) * 0.25f;
    mip_store<F16>(a, sm + 1u, gx, gy, weight);
    if (to_write < 2u) return;
    weight = (((weight + __shfl_sync(0xFFFFFFFFu, weight, lane + 1u)) + __shfl_sync(0xFFFFFFFFu, weight, lane + 2u)) + __shfl_sync(0xFFFFFFFFu, weight, lane + 3u)) * 0.25f;
    if ((lane & 3u) == 0u) mip_store<F16>(a, sm + 2u, gx >> 1, gy >> 1, weight);
    if (to_write < 3u) return;
    weight = (((weight + __shfl_sync(0xFFFFFFFFu, weight, lane + 4u)) + __shfl_sync(0xFFFFFFFFu, weight, lane + 8u)) + __shfl_sync(0xFFFFFFFFu, weight, lane + 12u)) * 0.25f;
    if ((lane & 15u) == 0u) { mip_store<F16>(a, sm + 3u, gx >> 2, gy >> 2, weight); s_weights[t >> 4] = weight; }
    if (to_write < 4u) return;
    __syncthreads();
    if (t >= 32u) return;                 // whole first warp stays for the shuffles; lanes >= 16 carry zeros
    weight = t < 16u ? s_weights[t] : 0.f;
    const uint32_t ox = blockIdx.x * 2u + (lx >> 1), oy = blockIdx.y * 2u + (ly >> 1);
    weight = (((weight + __shfl_sync(0xFFFFFFFFu, weight, lane + 1u)) + __shfl_sync(0xFFFFFFFFu, weight, lane + 2u)) + __shfl_sync(0xFFFFFFFFu, weight, lane + 3u)) * 0.25f;
    if (t < 16u && (lane & 3u) == 0u) mip_store<F16>(a, sm + 4u, ox, oy, weight);
    if (to_write < 5u) return;
    weight = (((weight + __shfl_sync(0xFFFFFFFFu, weight, lane + 4u)) + __shfl_sync(0xFFFFFFFFu, weight, lane + 8u)) + __shfl_sync(0xFFFFFFFFu, weight, lane + 12u)) * 0.25f;
    if (lane == 0u) mip_store<F16>(a, sm + 5u, ox >> 1, oy >> 1, weight);
}

int mips_generate(rt2_ctx* ctx, int which) {
    if (which != 0 && which != 1) return fail(ctx, RT2_E_INVALID, "which must be 0 (local lights) or 1 (environment)");
    int rc = lights_ensure_resources(ctx); if (rc) return rc;
    MipArgs a{};
    if (which == 0) {
        a.base = ctx->local_pdf; a.w = ctx->local_pdf_w; a.h = ctx->local_pdf_h; a.mips = ctx->local_pdf_mips;
    } else {
        const DeviceScene& s = ctx->scene;
        if (!s.sky) return fail(ctx, RT2_E_STATE, "no sky box uploaded");
        if (!ctx->env_pdf) {   // src/render_resources.rs:208-223: R16_SFLOAT, skybox size, ceil(log2(max)) + 1 mips (f32 log, as written)
            ctx->env_pdf_w = s.sky_w; ctx->env_pdf_h = s.sky_h;
            ctx->env_pdf_mips = (uint32_t)std::ceil(std::log((float)std::max(s.sky_w, s.sky_h)) / std::log(2.0f)) + 1u;
            if (ctx->env_pdf_mips > 32u) return fail(ctx, RT2_E_INVALID, "sky box too large");
            const size_t texels = mip_chain_texels(ctx->env_pdf_w, ctx->env_pdf_h, ctx->env_pdf_mips);
            RT2_CUDA(ctx, cudaMalloc((void**)&ctx->env_pdf, texels * 2));
            RT2_CUDA(ctx, cudaMemsetAsync(ctx->env_pdf, 0, texels * 2, ctx->stream));
            ctx->device_bytes += texels * 2;
        }
        a.base = ctx->env_pdf; a.w = ctx->env_pdf_w; a.h = ctx->env_pdf_h; a.mips = ctx->env_pdf_mips;
        a.sky = s.sky; a.sky_w = s.sky_w; a.sky_h = s.sky_h;
    }
    if (a.mips > 32u) return fail(ctx, RT2_E_INVALID, "too many mip levels");
    uint32_t off = 0;
    for (uint32_t k = 0; k < a.mips; k++) { a.offset[k] = off; off += std::max(1u, a.w >> k) * std::max(1u, a.h >> k); }
    // src/mip_pass.rs:149-184; stream order replaces the barrier the reference forgot between dispatches
    uint32_t width = a.w, height = a.h, launches = 0;
    for (uint32_t mip = 0; mip < a.mips; mip += 5) {
        a.source_mip = mip;
        const dim3 grid((uint32_t)std::ceil((float)width / 32.0f), (uint32_t)std::ceil((float)height / 32.0f));
        if (which == 0) k_mip_pass<false, false><<<grid, 256, 0, ctx->stream>>>(a);
        else k_mip_pass<true, true><<<grid, 256, 0, ctx->stream>>>(a);
        launches++;
        width = std::max(1u, width >> 5); height = std::max(1u, height >> 5);
    }
    RT2_CUDA(ctx, cudaGetLastError());
    ctx->stats.kernel_launches = launches;
    ctx->stats_from_counters = false;
    return RT2_OK;
}

}  // namespace rt2
