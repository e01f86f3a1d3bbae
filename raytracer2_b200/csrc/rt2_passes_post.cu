// rt2_passes_post.cu — K13: compose + AgX tonemap into the RGBA8 output
// (src/postprocess.rs:225-307, src/shaders/post_processing.comp:47-191).  The swapchain image is
// replaced by a device RGBA8 buffer (RT2_BUF_OUTPUT_RGBA8).
//
// HBM-bound, one thread per pixel, 8 rows x 32 columns per block so every warp reads and writes whole
// 128/256-byte lines: reference mode 8 B read + 4 B written per pixel, composed mode 28 + 4.
#include "rt2_internal.cuh"
#include "rt2_codec.cuh"

namespace rt2 {

struct PostArgs {
    const rt2_gconst* g;
    uint32_t width, height;
    const ushort4* diffuse; const ushort4* specular;
    const float* gb_depth; const uint32_t* gb_diffuse_albedo; const uint32_t* gb_specular_rough; const ushort4* gb_emissive;
    float4* motion_vectors;
    const float4* sky; uint32_t sky_w, sky_h;
    uchar4* out;
    const float* gamma_lut;      // CodecLuts::gamma22
    uint32_t band_rows, band_rank, band_world;   // band_world > 1: only the rows of this rank's bands (rt2_group_postprocess)
};

__device__ __forceinline__ float agx_contrast(float x) {          // agxDefaultContrastApprox, :47-58
    const float x2 = x * x, x4 = x2 * x2;
    return (((((15.5f * x4 * x2 - 40.14f * x4 * x) + 31.96f * x4) - 6.868f * x2 * x) + 0.4298f * x2) + 0.1191f * x) - 0.00232f;
}
__device__ __forceinline__ V3 mat3_mul(const float* m, V3 v) {    // GLSL mat3(c0,c1,c2) * v
    return v3((m[0] * v.x + m[3] * v.y) + m[6] * v.z, (m[1] * v.x + m[4] * v.y) + m[7] * v.z, (m[2] * v.x + m[5] * v.y) + m[8] * v.z);
}
// GLSL pow is undefined for x < 0: pinned to NaN (what exp2(y*log2(x)) yields on GPUs)
__device__ __forceinline__ float glsl_pow(float x, float y) { return x < 0.f ? __int_as_float(0x7FC00000) : pow_pinned(x, y); }
__device__ __forceinline__ uint32_t unorm8(float c) { return (uint32_t)floorf(clampf(c, 0.f, 1.f) * 255.0f + 0.5f); }

__device__ __forceinline__ uchar4 tonemap_store(V3 col) {
    const float agx_mat[9] = {0.842479062253094f, 0.0423282422610123f, 0.0423756549057051f, 0.0784335999999992f, 0.878468636469772f,
                              0.0784336f, 0.0792237451477643f, 0.0791661274605434f, 0.879142973793104f};
    const float agx_inv[9] = {1.19687900512017f, -0.0528968517574562f, -0.0529716355144438f, -0.0980208811401368f, 1.15190312990417f,
                              -0.0980434501171241f, -0.0990297440797205f, -0.0989611768448433f, 1.15107367264116f};
    const float min_ev = -12.47393f, max_ev = 4.026069f;
    // agx, :60-83
    V3 v = mat3_mul(agx_mat, col);
    v = v3(clampf(log2_pinned(v.x), min_ev, max_ev), clampf(log2_pinned(v.y), min_ev, max_ev), clampf(log2_pinned(v.z), min_ev, max_ev));
    v = v3((v.x - min_ev) / (max_ev - min_ev), (v.y - min_ev) / (max_ev - min_ev), (v.z - min_ev) / (max_ev - min_ev));
    v = v3(agx_contrast(v.x), agx_contrast(v.y), agx_contrast(v.z));
    // agxLook (AGX_LOOK 0), :100-124
    const float luma = dot3(v, v3(0.2126f, 0.7152f, 0.0722f));
    v = v3(glsl_pow(v.x * 1.0f + 0.0f, 1.0f), glsl_pow(v.y * 1.0f + 0.0f, 1.0f), glsl_pow(v.z * 1.0f + 0.0f, 1.0f));
    v = v3(luma + 1.0f * (v.x - luma), luma + 1.0f * (v.y - luma), luma + 1.0f * (v.z - luma));
    // agxEotf, :85-98
    v = mat3_mul(agx_inv, v);
    v = v3(glsl_pow(v.x, 2.2f), glsl_pow(v.y, 2.2f), glsl_pow(v.z, 2.2f));
    // max(col, 1e-6): fmaxf returns the non-NaN operand (FMNMX) — the pinned reading of GLSL max with NaN
    v = v3(fmaxf(v.x, 0.000001f), fmaxf(v.y, 0.000001f), fmaxf(v.z, 0.000001f));
    if (isnan(v.x) || isnan(v.y) || isnan(v.z)) v = v3(1.f, 0.f, 0.f);
    return make_uchar4((unsigned char)unorm8(v.x), (unsigned char)unorm8(v.y), (unsigned char)unorm8(v.z), 255);
}

__device__ __forceinline__ V3 h3(ushort4 p) { return v3(f16_bits_to_f32(p.x), f16_bits_to_f32(p.y), f16_bits_to_f32(p.z)); }

// post-process' own copy of GetEnvironmentRadiance (:38-47; same math as the bridge's)
__device__ __forceinline__ V3 post_environment(const PostArgs& a, V3 dir) {
    if (a.g->environment != 1u || a.sky == nullptr) return v3(0.f, 0.f, 0.f);
    const float PI_ = 3.1415926f;
    const float u = 0.5f + atan2f(dir.z, dir.x) / (2.0f * PI_);
    const float v = 0.5f - asinf(dir.y) / PI_;
    const int W = (int)a.sky_w, H = (int)a.sky_h;
    const float x = u * (float)W - 0.5f, y = v * (float)H - 0.5f;
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
    return top * ib + bot * wb;
}

__global__ void __launch_bounds__(256) k_postprocess(PostArgs a) {
    const uint32_t x = blockIdx.x * 32u + (threadIdx.x & 31u), y = blockIdx.y * 8u + (threadIdx.x >> 5);
    if (x >= a.width || y >= a.height) return;
    if (a.band_world > 1u && (y / a.band_rows) % a.band_world != a.band_rank) return;
    const size_t i = (size_t)y * a.width + x;
    const rt2_gconst& g = *a.g;
    V3 col = v3(0.f, 0.f, 0.f);
    if (g.refrence_mode == 1u) {
        col = h3(a.diffuse[i]);                                                         // :156-157
    } else if (a.gb_depth[i] != 100000.0f) {                                            // :158-174
        const V3 albedo = unpack_r11g11b10(a.gb_diffuse_albedo[i]);
        const float4 sf = unpack_rgba8_gamma(a.gb_specular_rough[i], a.gamma_lut);
        const V3 emissive = h3(a.gb_emissive[i]);
        V3 di = h3(a.diffuse[i]), si = h3(a.specular[i]);
        if (g.textures == 1u) { di = di * albedo; si = si * v3(fmaxf(0.01f, sf.x), fmaxf(0.01f, sf.y), fmaxf(0.01f, sf.z)); }
        col = (di + si) + emissive;
    } else {                                                                            // :175-180
        const rt2_planar_view_constants& view = g.view; const rt2_planar_view_constants& prev = g.prev_view;
        const float pcx = (float)x + 0.5f, pcy = (float)y + 0.5f;
        // setupPrimaryRay direction (GBufferHelpers.glsl:12-27)
        float target[4], dir[4];
        mat4_mul_vec4(view.mat_clip_to_view, (pcx / view.viewport_size[0]) * 2.0f - 1.0f, (pcy / view.viewport_size[1]) * 2.0f - 1.0f, 1.0f, 1.0f, target);
        const V3 nt = normalize3(v3(target[0], target[1], target[2]));
        mat4_mul_vec4(view.mat_view_to_world, nt.x, nt.y, nt.z, 0.0f, dir);
        // getEnvironmentMotionVector, :126-146
        float clip[4] = {view.window_to_clip_scale[0] * pcx + view.window_to_clip_bias[0], view.window_to_clip_scale[1] * pcy + view.window_to_clip_bias[1], 0.0f, 1.0f};
        float world[4], pc[4];
        mat4_mul_vec4(view.mat_clip_to_world, clip[0], clip[1], clip[2], clip[3], world);
        mat4_mul_vec4(prev.mat_world_to_clip, world[0], world[1], world[2], world[3], pc);
        pc[0] /= pc[3]; pc[1] /= pc[3];
        const float mx = view.clip_to_window_scale[0] * (pc[0] - clip[0]) + (view.pixel_offset[0] - prev.pixel_offset[0]);
        const float my = view.clip_to_window_scale[1] * (pc[1] - clip[1]) + (view.pixel_offset[1] - prev.pixel_offset[1]);
        if (a.motion_vectors) a.motion_vectors[i] = make_float4(mx, my, 0.0f, 0.0f);
        col = post_environment(a, v3(dir[0], dir[1], dir[2]));
    }
    a.out[i] = tonemap_store(col);
}

int post_process(rt2_ctx* ctx, uint32_t frame) {
    FrameBuffers& f = ctx->fb;
    if (!f.width) return fail(ctx, RT2_E_STATE, "no frame buffers: call rt2_resize first");
    if (!ctx->have_gconst) return fail(ctx, RT2_E_STATE, "no constants: call rt2_set_constants first");
    PostArgs a{};
    a.g = ctx->d_gconst; a.width = f.width; a.height = f.height;
    a.diffuse = f.diffuse_f16; a.specular = f.specular_f16; a.out = f.output_rgba8;
    a.sky = ctx->scene.sky; a.sky_w = ctx->scene.sky_w; a.sky_h = ctx->scene.sky_h; a.gamma_lut = ctx->codec_luts;
    a.band_rows = ctx->post_band[0]; a.band_rank = ctx->post_band[1]; a.band_world = ctx->post_band[0] ? ctx->post_band[2] : 0u;
    if (ctx->gconst.refrence_mode != 1u) {
        if (!f.gb_depth[0]) return fail(ctx, RT2_E_STATE, "no G-buffer: run rt2_render in ReSTIR mode first");
        const uint32_t cur = frame & 1u;           // source_descriptors[frame % 2], src/postprocess.rs:284
        a.gb_depth = f.gb_depth[cur]; a.gb_diffuse_albedo = f.gb_diffuse_albedo[cur]; a.gb_specular_rough = f.gb_specular_rough[cur];
        a.gb_emissive = f.gb_emissive[cur]; a.motion_vectors = f.motion_vectors;
    }
    const dim3 grid((f.width + 31u) / 32u, (f.height + 7u) / 8u);
    k_postprocess<<<grid, 256, 0, ctx->stream>>>(a);
    RT2_CUDA(ctx, cudaGetLastError());
    ctx->stats.kernel_launches = 1;
    ctx->stats_from_counters = false;
    return RT2_OK;
}

}  // namespace rt2
