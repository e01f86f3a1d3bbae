// rt2_passes_common.cu — resources shared by the light / G-buffer / reservoir / post passes and their
// read-out through rt2_read (RenderResources, src/render_resources.rs:130-341).
#include <vector>
#include "rt2_internal.cuh"
#include "rt2_restir.cuh"

namespace rt2 {

// The two decoders that cost a binary64 transcendental per call, tabulated with the very functions they replace (so a lookup is the
// same bits as the inline form): the gamma-2.2 byte of the G-buffer's specular/roughness word (256 values) and the 16-bit log radiance
// of a light record (65536 values); plus the luminance of a LogLuv exponent field (16384 values), decoded for every GI reservoir K6 / K7 / K8 load.
__global__ void k_build_codec_luts(float* luts) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 256u) luts[i] = gamma22_decode(i);
    if (i < 65536u) luts[256u + i] = unpack_light_radiance(i);
    if (i < 16384u) luts[256u + 65536u + i] = logluv_luminance(i);
}
int codec_luts_init(rt2_ctx* ctx) {
    if (ctx->codec_luts) return RT2_OK;
    RT2_CUDA(ctx, cudaMalloc((void**)&ctx->codec_luts, (256 + 65536 + 16384) * sizeof(float)));
    k_build_codec_luts<<<65536 / 256, 256, 0, ctx->stream>>>(ctx->codec_luts);
    RT2_CUDA(ctx, cudaGetLastError());
    return RT2_OK;
}

void lights_release(rt2_ctx* ctx);
size_t mip_chain_texels(uint32_t w, uint32_t h, uint32_t mips);

void passes_release_frame(rt2_ctx* ctx) {
    FrameBuffers& f = ctx->fb;
    for (int i = 0; i < 2; i++) {
        void* ptrs[] = {f.gb_depth[i], f.gb_normals[i], f.gb_geo_normals[i], f.gb_diffuse_albedo[i], f.gb_specular_rough[i], f.gb_emissive[i]};
        for (void* p : ptrs) if (p) cudaFree(p);
        f.gb_depth[i] = nullptr; f.gb_normals[i] = f.gb_geo_normals[i] = f.gb_diffuse_albedo[i] = f.gb_specular_rough[i] = nullptr; f.gb_emissive[i] = nullptr;
    }
    void* ptrs[] = {f.motion_vectors, f.di_reservoirs, f.gi_reservoirs, f.secondary_gbuffer, f.neighbor_offsets, f.di_state};
    for (void* p : ptrs) if (p) cudaFree(p);
    f.motion_vectors = nullptr; f.di_reservoirs = nullptr; f.gi_reservoirs = nullptr; f.secondary_gbuffer = nullptr; f.neighbor_offsets = nullptr; f.di_state = nullptr;
    f.reservoir_block_row_pitch = f.reservoir_array_pitch = 0;
}

void passes_release(rt2_ctx* ctx) { passes_release_frame(ctx); lights_release(ctx); }

template <class T>
static cudaError_t zalloc(rt2_ctx* ctx, T** p, size_t n) {
    cudaError_t e = cudaMalloc((void**)p, std::max<size_t>(n, 1) * sizeof(T));
    if (e == cudaSuccess) { ctx->device_bytes += std::max<size_t>(n, 1) * sizeof(T); e = cudaMemsetAsync(*p, 0, std::max<size_t>(n, 1) * sizeof(T), ctx->stream); }
    return e;
}

// G-buffers x2, motion vectors, DI/GI reservoirs, secondary G-buffer — allocated on the first ReSTIR-mode render
int restir_ensure_resources(rt2_ctx* ctx) {
    FrameBuffers& f = ctx->fb;
    if (!f.width) return fail(ctx, RT2_E_STATE, "no frame buffers: call rt2_resize first");
    if (f.gb_depth[0]) return RT2_OK;
    const size_t n = (size_t)f.width * f.height;
    for (int i = 0; i < 2; i++) {
        RT2_CUDA(ctx, zalloc(ctx, &f.gb_depth[i], n)); RT2_CUDA(ctx, zalloc(ctx, &f.gb_normals[i], n)); RT2_CUDA(ctx, zalloc(ctx, &f.gb_geo_normals[i], n));
        RT2_CUDA(ctx, zalloc(ctx, &f.gb_diffuse_albedo[i], n)); RT2_CUDA(ctx, zalloc(ctx, &f.gb_specular_rough[i], n)); RT2_CUDA(ctx, zalloc(ctx, &f.gb_emissive[i], n));
    }
    RT2_CUDA(ctx, zalloc(ctx, &f.motion_vectors, n));
    // src/light_passes.rs:718-731 calculate_reservoir_buffer_parameters (RTXDI_RESERVOIR_BLOCK_SIZE = 16)
    const uint32_t wb = (f.width + 15u) / 16u, hb = (f.height + 15u) / 16u;
    f.reservoir_block_row_pitch = wb * 256u; f.reservoir_array_pitch = f.reservoir_block_row_pitch * hb;
    RT2_CUDA(ctx, zalloc(ctx, &f.di_reservoirs, (size_t)f.reservoir_array_pitch * 2));    // src/render_resources.rs:191-198
    RT2_CUDA(ctx, zalloc(ctx, &f.gi_reservoirs, (size_t)f.reservoir_array_pitch * 2 * 2)); // 32 B each, :199-206
    {   // fill_neighbor_offset_buffer, src/light_passes.rs:671-698 (f32 arithmetic).  The reference stores `as u8` bytes and reads them as
        // vec2; pinned to the RTXDI SDK's intent: signed bytes, decoded as SNORM8 by K7.
        std::vector<signed char> tab; tab.reserve(RT2_NEIGHBOR_OFFSET_COUNT * 2);
        const float phi2 = 1.0f / 1.3247179572447f;
        float u = 0.5f, v = 0.5f;
        while (tab.size() < (size_t)RT2_NEIGHBOR_OFFSET_COUNT * 2) {
            u += phi2; v += phi2 * phi2;
            if (u >= 1.0f) u -= 1.0f;
            if (v >= 1.0f) v -= 1.0f;
            const float r_sq = (u - 0.5f) * (u - 0.5f) + (v - 0.5f) * (v - 0.5f);
            if (r_sq > 0.25f) continue;
            tab.push_back((signed char)((u - 0.5f) * 250.0f));
            tab.push_back((signed char)((v - 0.5f) * 250.0f));
        }
        RT2_CUDA(ctx, cudaMalloc((void**)&f.neighbor_offsets, tab.size())); ctx->device_bytes += tab.size();
        RT2_CUDA(ctx, cudaMemcpyAsync(f.neighbor_offsets, tab.data(), tab.size(), cudaMemcpyHostToDevice, ctx->stream));
        RT2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));     // `tab` is pageable host memory going out of scope
    }
    RT2_CUDA(ctx, zalloc(ctx, &f.di_state, n * 8));                                        // K9 wavefront hand-over state, 80 B per pixel + 48 B for the environment reservoir
    RT2_CUDA(ctx, zalloc(ctx, &f.secondary_gbuffer, (size_t)f.reservoir_array_pitch * 3)); // 48 B x arrayPitch (block-linear addressing), :309-317
    return RT2_OK;
}

bool pass_buffer(rt2_ctx* ctx, int which, void** ptr, size_t* bytes) {
    const FrameBuffers& f = ctx->fb;
    const size_t n = (size_t)f.width * f.height;
    const uint32_t cur = ctx->last_gbuffer_index & 1u;
    auto mip = [&](void* base, uint32_t w, uint32_t h, uint32_t mips, uint32_t k, size_t texel) -> bool {
        if (!base || k >= mips) return false;
        size_t off = 0;
        for (uint32_t j = 0; j < k; j++) off += (size_t)std::max(1u, w >> j) * std::max(1u, h >> j);
        *ptr = (char*)base + off * texel; *bytes = (size_t)std::max(1u, w >> k) * std::max(1u, h >> k) * texel;
        return true;
    };
    if (which >= RT2_BUF_LOCAL_PDF_MIP0 && which < RT2_BUF_LOCAL_PDF_MIP0 + 32)
        return mip(ctx->local_pdf, ctx->local_pdf_w, ctx->local_pdf_h, ctx->local_pdf_mips, (uint32_t)(which - RT2_BUF_LOCAL_PDF_MIP0), 4);
    if (which >= RT2_BUF_ENV_PDF_MIP0 && which < RT2_BUF_ENV_PDF_MIP0 + 32)
        return mip(ctx->env_pdf, ctx->env_pdf_w, ctx->env_pdf_h, ctx->env_pdf_mips, (uint32_t)(which - RT2_BUF_ENV_PDF_MIP0), 2);
    switch (which) {
        case RT2_BUF_LIGHT_INFO: *ptr = ctx->light_info; *bytes = (size_t)ctx->light_records * sizeof(rt2_light_info); return ctx->light_info && ctx->light_records;
        case RT2_BUF_GEOM_TO_LIGHT: *ptr = ctx->geom_to_light; *bytes = (size_t)ctx->scene.num_geoms * 4; return ctx->geom_to_light && ctx->scene.num_geoms;
        case RT2_BUF_LIGHT_TASKS: *ptr = ctx->light_tasks; *bytes = (size_t)ctx->num_light_tasks * sizeof(rt2_prepare_lights_task); return ctx->light_tasks && ctx->num_light_tasks;
        case RT2_BUF_GBUFFER_DEPTH: *ptr = f.gb_depth[cur]; *bytes = n * 4; return f.gb_depth[cur] != nullptr;
        case RT2_BUF_GBUFFER_NORMALS: *ptr = f.gb_normals[cur]; *bytes = n * 4; return f.gb_normals[cur] != nullptr;
        case RT2_BUF_GBUFFER_GEO_NORMALS: *ptr = f.gb_geo_normals[cur]; *bytes = n * 4; return f.gb_geo_normals[cur] != nullptr;
        case RT2_BUF_GBUFFER_DIFFUSE_ALBEDO: *ptr = f.gb_diffuse_albedo[cur]; *bytes = n * 4; return f.gb_diffuse_albedo[cur] != nullptr;
        case RT2_BUF_GBUFFER_SPECULAR_ROUGH: *ptr = f.gb_specular_rough[cur]; *bytes = n * 4; return f.gb_specular_rough[cur] != nullptr;
        case RT2_BUF_GBUFFER_EMISSIVE: *ptr = f.gb_emissive[cur]; *bytes = n * 8; return f.gb_emissive[cur] != nullptr;
        case RT2_BUF_MOTION_VECTORS: *ptr = f.motion_vectors; *bytes = n * 16; return f.motion_vectors != nullptr;
        case RT2_BUF_DI_RESERVOIRS: *ptr = f.di_reservoirs; *bytes = (size_t)f.reservoir_array_pitch * 2 * sizeof(rt2_packed_di_reservoir); return f.di_reservoirs != nullptr;
        case RT2_BUF_GI_RESERVOIRS: *ptr = f.gi_reservoirs; *bytes = (size_t)f.reservoir_array_pitch * 2 * 32; return f.gi_reservoirs != nullptr;
        case RT2_BUF_RIS: *ptr = ctx->ris_buffer; *bytes = (size_t)ctx->ris_entries * 8; return ctx->ris_buffer != nullptr;
        case RT2_BUF_RIS_LIGHT_DATA: *ptr = ctx->ris_light_data; *bytes = (size_t)ctx->ris_entries * 32; return ctx->ris_light_data != nullptr;
        case RT2_BUF_NEIGHBOR_OFFSETS: *ptr = f.neighbor_offsets; *bytes = (size_t)RT2_NEIGHBOR_OFFSET_COUNT * 2; return f.neighbor_offsets != nullptr;
        case RT2_BUF_SECONDARY_GBUFFER: *ptr = f.secondary_gbuffer; *bytes = (size_t)f.reservoir_array_pitch * 48; return f.secondary_gbuffer != nullptr;
        default: return false;
    }
}

}  // namespace rt2
