// rt2_passes_presample.cu — K11: LightPasses::execute_presampeling (src/light_passes.rs:510-548):
//   lighting_passes/presample_locallights.comp, presample_environment.comp,
//   rtxdi/PresamplingFunctions.hlsli:30-162 (RTXDI_SamplePdfMipmap, RTXDI_PresampleLocalLights, RTXDI_PresampleEnvironmentMap)
// One thread per RIS entry (the reference's fixed grid: 128 tiles x 1024 samples): a descent through the PDF mip pyramid
// that K12 built (<= 13 levels, 4 texel loads each, all L2-resident), one 8-byte RIS entry and — for local lights — a
// 32-byte compact light record out.  131 072 threads: latency bound, microseconds.
//
// The reference dispatches these passes every frame although their consumer is compiled out (RTXDI_ENABLE_PRESAMPLING 0,
// bridge :12).  Pins (see oracle/rt2_oracle_gi.inc): the frame is the caller's (the compute passes never push it); texel
// fetches outside a mip return 0; the compact-light buffer is sized for the RIS segment (the reference under-allocates it).
#include "rt2_restir.cuh"

namespace rt2 {

struct PresampleArgs {
    const rt2_gconst* g; uint32_t frame;
    const float* pdf_f32; const uint16_t* pdf_f16; uint32_t w, h, mips;
    const rt2_light_info* lights; uint32_t light_capacity;
    uint2* ris; uint4* ris_light_data;
    uint32_t* error_flag;
};

template <bool ENV>
__device__ __forceinline__ float pdf_texel(const PresampleArgs& a, int x, int y, int mip, uint32_t mip_offset) {
    const uint32_t mw = max(1u, a.w >> mip), mh = max(1u, a.h >> mip);
    if (x < 0 || y < 0 || (uint32_t)x >= mw || (uint32_t)y >= mh) return 0.f;
    const size_t off = (size_t)mip_offset + (size_t)y * mw + x;
    return ENV ? f16_bits_to_f32(a.pdf_f16[off]) : a.pdf_f32[off];
}

// RTXDI_SamplePdfMipmap, :30-94
template <bool ENV>
__device__ __forceinline__ void sample_pdf_mipmap(const PresampleArgs& a, Rng& rng, uint32_t* posx, uint32_t* posy, float* pdf) {
    const int last = max(0, (int)floorf(log2_pinned((float)max(a.w, a.h))) - 1);
    uint32_t px = 0, py = 0;
    *pdf = 1.0f;
    for (int mip = last; mip >= 0; mip--) {
        px *= 2; py *= 2;
        uint32_t mip_offset = 0;
        for (int k = 0; k < mip; k++) mip_offset += max(1u, a.w >> k) * max(1u, a.h >> k);
        float sx = 0.f, sy = 0.f, sz = 0.f, sw = 0.f;
        if ((uint32_t)mip < a.mips) {
            sx = fmaxf(0.f, pdf_texel<ENV>(a, (int)px + 0, (int)py + 0, mip, mip_offset));
            sy = fmaxf(0.f, pdf_texel<ENV>(a, (int)px + 0, (int)py + 1, mip, mip_offset));
            sz = fmaxf(0.f, pdf_texel<ENV>(a, (int)px + 1, (int)py + 0, mip, mip_offset));
            sw = fmaxf(0.f, pdf_texel<ENV>(a, (int)px + 1, (int)py + 1, mip, mip_offset));
        }
        const float weight_sum = ((sx + sy) + sz) + sw;
        if (weight_sum <= 0) { *pdf = 0; *posx = px; *posy = py; return; }
        sx /= weight_sum; sy /= weight_sum; sz /= weight_sum; sw /= weight_sum;
        float rnd = rng_next(rng);
        if (rnd < sx) { *pdf *= sx; }
        else {
            rnd -= sx;
            if (rnd < sy) { py += 1; *pdf *= sy; }
            else {
                rnd -= sy;
                if (rnd < sz) { px += 1; *pdf *= sz; }
                else { px += 1; py += 1; *pdf *= sw; }
            }
        }
    }
    *posx = px; *posy = py;
}

template <bool ENV>
__global__ void __launch_bounds__(256) k_presample(PresampleArgs a) {
    const uint32_t sample_in_tile = blockIdx.x * blockDim.x + threadIdx.x;     // gl_GlobalInvocationID.x, 0..1023
    const uint32_t tile_index = blockIdx.y;                                     // gl_GlobalInvocationID.y, 0..127
    const rt2_gconst& g = *a.g;
    Rng rng = rng_init(sample_in_tile, tile_index, a.frame, 0u);
    uint32_t tx, ty; float pdf;
    sample_pdf_mipmap<ENV>(a, rng, &tx, &ty, &pdf);
    if (!ENV) {                                                                 // RTXDI_PresampleLocalLights, :96-133
        const rt2_ris_buffer_segment_params& seg = g.local_lights_risbuffer_segment_params;
        const uint32_t first = g.light_buffer_params.local_light_buffer_region.first_light_index;
        uint32_t light_index = zcurve_to_linear(tx, ty);
        const uint32_t ptr = sample_in_tile + tile_index * seg.tile_size;      // sic: bufferOffset is not added for local lights
        bool compact = false;
        float inv_source_pdf = 0;
        if (pdf > 0) {
            inv_source_pdf = 1.0f / pdf;
            if (light_index + first >= a.light_capacity) { atomicExch(a.error_flag, 1u); return; }   // PDF texture and light list disagree
            const rt2_light_info li = load_light_info(a.lights, light_index + first);
            compact = (li.color_type_and_flags & (1u << 28)) == 0u;            // packCompactLightInfo: only spot-shaped lights do not fit
            if (compact) {
                a.ris_light_data[2 * (size_t)ptr] = make_uint4(__float_as_uint(li.center[0]), __float_as_uint(li.center[1]), __float_as_uint(li.center[2]), li.color_type_and_flags);
                a.ris_light_data[2 * (size_t)ptr + 1] = make_uint4(li.direction1, li.direction2, li.scalars, li.log_radiance);
            }
        }
        light_index += first;
        if (compact) light_index |= 0x80000000u;                               // RTXDI_LIGHT_COMPACT_BIT
        a.ris[ptr] = make_uint2(light_index, __float_as_uint(inv_source_pdf));
    } else {                                                                    // RTXDI_PresampleEnvironmentMap, :135-162
        const rt2_ris_buffer_segment_params& seg = g.environment_light_risbuffer_segment_params;
        float fx = (float)tx, fy = (float)ty;
        fx += rng_next(rng);
        fy += rng_next(rng);
        const float u = fx / (float)a.w, v = fy / (float)a.h;
        const uint32_t packed_uv = (uint32_t)(saturatef(u) * 65535.0f) | ((uint32_t)(saturatef(v) * 65535.0f) << 16);
        const float inv_source_pdf = (pdf > 0) ? (1.0f / pdf) : 0.f;
        const uint32_t ptr = seg.buffer_offset + sample_in_tile + tile_index * seg.tile_size;
        a.ris[ptr] = make_uint2(packed_uv, __float_as_uint(inv_source_pdf));
    }
}

void presample_release(rt2_ctx* ctx) {
    if (ctx->ris_buffer) cudaFree(ctx->ris_buffer);
    if (ctx->ris_light_data) cudaFree(ctx->ris_light_data);
    if (ctx->presample_error) cudaFree(ctx->presample_error);
    ctx->ris_buffer = nullptr; ctx->ris_light_data = nullptr; ctx->presample_error = nullptr; ctx->ris_entries = 0;
}

// which: 0 = local lights (needs K10 + K12(0)), 1 = environment map (needs K12(1))
int presample(rt2_ctx* ctx, uint32_t frame, int which) {
    if (which != 0 && which != 1) return fail(ctx, RT2_E_INVALID, "rt2_presample: which must be 0 (local lights) or 1 (environment)");
    if (!ctx->have_gconst) return fail(ctx, RT2_E_STATE, "no constants: call rt2_set_constants first");
    const rt2_gconst& g = ctx->gconst;
    const rt2_ris_buffer_segment_params& seg = which == 0 ? g.local_lights_risbuffer_segment_params : g.environment_light_risbuffer_segment_params;
    const uint32_t TILES = 128, SAMPLES = 1024;                                 // the reference's fixed dispatch, src/light_passes.rs:538-542
    if (seg.tile_size < SAMPLES) return fail(ctx, RT2_E_INVALID, "RIS segment tile_size is smaller than the 1024 samples the pass writes per tile");
    const uint64_t end = (uint64_t)(which == 0 ? 0u : seg.buffer_offset) + (uint64_t)(TILES - 1) * seg.tile_size + SAMPLES;
    if (end > (1ull << 26)) return fail(ctx, RT2_E_INVALID, "RIS segment parameters address more than 64 M entries");
    if (ctx->ris_entries < end) {
        RT2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        uint2* nb = nullptr; uint4* nd = nullptr;
        RT2_CUDA(ctx, cudaMalloc((void**)&nb, end * sizeof(uint2)));
        RT2_CUDA(ctx, cudaMalloc((void**)&nd, end * 2 * sizeof(uint4)));
        RT2_CUDA(ctx, cudaMemsetAsync(nb, 0, end * sizeof(uint2), ctx->stream));
        RT2_CUDA(ctx, cudaMemsetAsync(nd, 0, end * 2 * sizeof(uint4), ctx->stream));
        if (ctx->ris_buffer) {      // keep the other segment's samples
            RT2_CUDA(ctx, cudaMemcpyAsync(nb, ctx->ris_buffer, ctx->ris_entries * sizeof(uint2), cudaMemcpyDeviceToDevice, ctx->stream));
            RT2_CUDA(ctx, cudaMemcpyAsync(nd, ctx->ris_light_data, ctx->ris_entries * 2 * sizeof(uint4), cudaMemcpyDeviceToDevice, ctx->stream));
            RT2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            cudaFree(ctx->ris_buffer); cudaFree(ctx->ris_light_data);
        }
        ctx->device_bytes += (end - ctx->ris_entries) * (sizeof(uint2) + 2 * sizeof(uint4));
        ctx->ris_buffer = nb; ctx->ris_light_data = nd; ctx->ris_entries = end;
    }
    if (!ctx->presample_error) { RT2_CUDA(ctx, cudaMalloc((void**)&ctx->presample_error, 4)); RT2_CUDA(ctx, cudaMemsetAsync(ctx->presample_error, 0, 4, ctx->stream)); }
    PresampleArgs a{};
    a.g = ctx->d_gconst; a.frame = frame; a.ris = ctx->ris_buffer; a.ris_light_data = ctx->ris_light_data; a.error_flag = ctx->presample_error;
    const dim3 grid(SAMPLES / 256, TILES);
    if (which == 0) {
        if (!ctx->lights_prepared || !ctx->local_pdf) return fail(ctx, RT2_E_STATE, "local-light PDF pyramid not built: call rt2_prepare_lights and rt2_generate_mips(0) first");
        if (g.local_light_pdf_texture_size[0] != ctx->local_pdf_w || g.local_light_pdf_texture_size[1] != ctx->local_pdf_h)
            return fail(ctx, RT2_E_INVALID, "GConst.localLightPdfTextureSize does not match the PDF texture (compute_pdf_texture_size)");
        a.pdf_f32 = ctx->local_pdf; a.w = ctx->local_pdf_w; a.h = ctx->local_pdf_h; a.mips = ctx->local_pdf_mips;
        a.lights = ctx->light_info; a.light_capacity = ctx->light_capacity;
        k_presample<false><<<grid, 256, 0, ctx->stream>>>(a);
    } else {
        if (!ctx->env_pdf) return fail(ctx, RT2_E_STATE, "environment PDF pyramid not built: call rt2_generate_mips(1) first");
        if (g.environment_pdf_texture_size[0] != ctx->env_pdf_w || g.environment_pdf_texture_size[1] != ctx->env_pdf_h)
            return fail(ctx, RT2_E_INVALID, "GConst.environmentPdfTextureSize does not match the PDF texture");
        a.pdf_f16 = ctx->env_pdf; a.w = ctx->env_pdf_w; a.h = ctx->env_pdf_h; a.mips = ctx->env_pdf_mips;
        k_presample<true><<<grid, 256, 0, ctx->stream>>>(a);
    }
    RT2_CUDA(ctx, cudaGetLastError());
    ctx->stats.kernel_launches = 1;
    return RT2_OK;
}

}  // namespace rt2
