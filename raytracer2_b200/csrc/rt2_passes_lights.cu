// rt2_passes_lights.cu — K10 prepare_lights: the emissive-triangle light list and mip 0 of the
// local-light PDF texture.  Host half: src/prepare_lights.rs:182-201 (exclusive scan over emissive
// geometries); device half: src/shaders/prepare_lights.comp:40-126.
//
// HBM-bound and tiny: 332 B read + 52 B written per light.  One thread per light; the
// task search runs over a task table that sits in L1 after the first warp touches it.
#include "rt2_internal.cuh"
#include "rt2_codec.cuh"

namespace rt2 {

// src/light_passes.rs:700-716 compute_pdf_texture_size (f64 arithmetic, as written)
void compute_pdf_texture_size(uint32_t max_items, uint32_t* w, uint32_t* h, uint32_t* mips) {
    double tw = std::max(1.0, std::ceil(std::sqrt((double)max_items)));
    tw = std::exp2(std::ceil(std::log2(tw)));
    double th = std::max(1.0, std::ceil((double)max_items / tw));
    th = std::exp2(std::ceil(std::log2(th)));
    double tm = std::max(1.0, std::log2(std::max(tw, th)) + 1.0);
    *w = (uint32_t)tw; *h = (uint32_t)th; *mips = (uint32_t)tm;
}

size_t mip_chain_texels(uint32_t w, uint32_t h, uint32_t mips) {
    size_t n = 0;
    for (uint32_t k = 0; k < mips; k++) n += (size_t)std::max(1u, w >> k) * std::max(1u, h >> k);
    return n;
}

__global__ void __launch_bounds__(256) k_prepare_lights(const rt2_prepare_lights_task* __restrict__ tasks, uint32_t num_tasks,
                                                        const rt2_geometry_info* __restrict__ geoms, const rt2_vertex* __restrict__ vertices,
                                                        const uint32_t* __restrict__ indices, rt2_light_info* __restrict__ lights,
                                                        float* __restrict__ pdf_mip0, uint32_t pdf_w, uint32_t pdf_h) {
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    // FindTask, prepare_lights.comp:40-72
    int left = 0, right = (int)num_tasks - 1;
    rt2_prepare_lights_task task;
    bool found = false;
    while (right >= left) {
        const int middle = (left + right) / 2;
        task = tasks[middle];
        const int tri = (int)tid - (int)task.light_buffer_offset;
        if (tri < 0) right = middle - 1;
        else if ((uint32_t)tri < task.triangle_count) { found = true; break; }
        else left = middle + 1;
    }
    if (!found) return;
    const uint32_t triangle = tid - task.light_buffer_offset;
    const rt2_geometry_info* g = geoms + task.geometry_index;
    const uint32_t index_offset = g->index_offset + 3u * triangle;
    float m[12];
#pragma unroll
    for (int c = 0; c < 3; c++) { m[c * 4] = g->transform[c * 4]; m[c * 4 + 1] = g->transform[c * 4 + 1]; m[c * 4 + 2] = g->transform[c * 4 + 2]; m[c * 4 + 3] = g->transform[c * 4 + 3]; }
    V3 p[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const float4 pos = *reinterpret_cast<const float4*>(vertices[g->vertex_offset + indices[index_offset + k]].pos);
        // (vec4(v.pos, 1) * transform).xyz: row vector x matrix, component c = dot(v4, column c)  (Q6, prepare_lights.comp:101-103)
        p[k] = v3(((pos.x * m[0] + pos.y * m[1]) + pos.z * m[2]) + 1.0f * m[3],
                  ((pos.x * m[4] + pos.y * m[5]) + pos.z * m[6]) + 1.0f * m[7],
                  ((pos.x * m[8] + pos.y * m[9]) + pos.z * m[10]) + 1.0f * m[11]);
    }
    TriangleLight tl;
    tl.base = p[0]; tl.edge1 = p[1] - p[0]; tl.edge2 = p[2] - p[0];
    tl.radiance = v3(g->emission[0] * 12.0f, g->emission[1] * 12.0f, g->emission[2] * 12.0f);
    const rt2_light_info li = store_triangle_light(tl);
    const uint32_t ptr = task.light_buffer_offset + triangle;
    uint4* dst = reinterpret_cast<uint4*>(lights + ptr);
    const uint4* src = reinterpret_cast<const uint4*>(&li);
    dst[0] = src[0]; dst[1] = src[1]; dst[2] = src[2];
    uint32_t x, y;
    linear_to_zcurve(ptr, x, y);
    if (x < pdf_w && y < pdf_h) pdf_mip0[(size_t)y * pdf_w + x] = light_power(li);
}

void presample_release(rt2_ctx* ctx);
// the environment light at index lights + 1 (src/main.rs:383; SURVEY Q7 pinned to intent, see store_environment_light)
__global__ void k_store_environment_light(rt2_light_info* lights, uint32_t num_lights, uint32_t tex_w, uint32_t tex_h) {
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    rt2_light_info zero; memset(&zero, 0, sizeof zero);
    lights[num_lights] = zero;
    lights[num_lights + 1] = store_environment_light(tex_w, tex_h);
}

void lights_release(rt2_ctx* ctx) {
    presample_release(ctx);
    void* ptrs[] = {ctx->light_info, ctx->geom_to_light, ctx->light_tasks, ctx->local_pdf, ctx->env_pdf};
    for (void* p : ptrs) if (p) cudaFree(p);
    ctx->light_info = nullptr; ctx->geom_to_light = nullptr; ctx->light_tasks = nullptr; ctx->local_pdf = nullptr; ctx->env_pdf = nullptr;
    ctx->light_capacity = ctx->light_records = ctx->num_light_tasks = 0;
    ctx->local_pdf_w = ctx->local_pdf_h = ctx->local_pdf_mips = ctx->env_pdf_w = ctx->env_pdf_h = ctx->env_pdf_mips = 0;
    ctx->lights_prepared = false;
}

// RenderResources::new for the scene-sized light resources (src/render_resources.rs:135-150,225-239)
int lights_ensure_resources(rt2_ctx* ctx) {
    const DeviceScene& s = ctx->scene;
    if (!s.geom_tri_offset) return fail(ctx, RT2_E_STATE, "no scene: call rt2_upload_scene first");
    if (ctx->light_info) return RT2_OK;
    const uint32_t n = s.num_lights;
    // lights + 2 records when the scene has a sky: the environment light lives at lights + 1 (src/main.rs:383)
    const bool env = s.sky != nullptr && s.sky_w <= 0xffffu && s.sky_h <= 0xffffu;
    const size_t records = (size_t)n + (env ? 2 : 0);
    RT2_CUDA(ctx, cudaMalloc((void**)&ctx->light_info, std::max<size_t>(records, 1) * sizeof(rt2_light_info)));
    RT2_CUDA(ctx, cudaMalloc((void**)&ctx->geom_to_light, std::max<size_t>(s.num_geoms, 1) * 4));
    RT2_CUDA(ctx, cudaMalloc((void**)&ctx->light_tasks, std::max<size_t>(s.num_geoms, 1) * sizeof(rt2_prepare_lights_task)));
    ctx->light_capacity = n; ctx->light_records = (uint32_t)records;
    compute_pdf_texture_size(n, &ctx->local_pdf_w, &ctx->local_pdf_h, &ctx->local_pdf_mips);
    const size_t texels = mip_chain_texels(ctx->local_pdf_w, ctx->local_pdf_h, ctx->local_pdf_mips);
    RT2_CUDA(ctx, cudaMalloc((void**)&ctx->local_pdf, texels * 4));
    // the reference never clears this image; texels no light writes are pinned to 0
    RT2_CUDA(ctx, cudaMemsetAsync(ctx->local_pdf, 0, texels * 4, ctx->stream));
    RT2_CUDA(ctx, cudaMemsetAsync(ctx->light_info, 0, std::max<size_t>(records, 1) * sizeof(rt2_light_info), ctx->stream));
    RT2_CUDA(ctx, cudaMemsetAsync(ctx->geom_to_light, 0xFF, std::max<size_t>(s.num_geoms, 1) * 4, ctx->stream));
    ctx->device_bytes += std::max<size_t>(records, 1) * sizeof(rt2_light_info) + std::max<size_t>(s.num_geoms, 1) * 20 + texels * 4;
    return RT2_OK;
}

int lights_prepare(rt2_ctx* ctx) {
    int rc = lights_ensure_resources(ctx); if (rc) return rc;
    const DeviceScene& s = ctx->scene;
    // src/prepare_lights.rs:182-201
    std::vector<uint32_t> geometry_to_light(s.num_geoms, 0xffffffffu);
    std::vector<rt2_prepare_lights_task> tasks;
    uint32_t light_buffer_offset = 0;
    for (uint32_t g = 0; g < s.num_geoms; g++) {
        const rt2_geometry_info& gi = s.h_geoms[g];
        if (gi.emission[0] == 0.0f && gi.emission[1] == 0.0f && gi.emission[2] == 0.0f) continue;
        geometry_to_light[g] = light_buffer_offset;
        const uint32_t triangle_count = s.h_index_counts[g] / 3;
        tasks.push_back(rt2_prepare_lights_task{g, light_buffer_offset, triangle_count, 0});
        light_buffer_offset += triangle_count;
    }
    ctx->num_light_tasks = (uint32_t)tasks.size();
    // staging copies are synchronous-with-respect-to-host: the vectors die at return
    if (!tasks.empty()) RT2_CUDA(ctx, cudaMemcpyAsync(ctx->light_tasks, tasks.data(), tasks.size() * sizeof(rt2_prepare_lights_task), cudaMemcpyHostToDevice, ctx->stream));
    if (s.num_geoms) RT2_CUDA(ctx, cudaMemcpyAsync(ctx->geom_to_light, geometry_to_light.data(), (size_t)s.num_geoms * 4, cudaMemcpyHostToDevice, ctx->stream));
    RT2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const uint32_t groups = (uint32_t)std::ceil((float)light_buffer_offset / 256.0f);   // src/prepare_lights.rs:250-255
    if (groups) {
        k_prepare_lights<<<groups, 256, 0, ctx->stream>>>(ctx->light_tasks, ctx->num_light_tasks, s.geoms, s.vertices, s.indices, ctx->light_info,
                                                          ctx->local_pdf, ctx->local_pdf_w, ctx->local_pdf_h);
        RT2_CUDA(ctx, cudaGetLastError());
    }
    if (ctx->light_records == ctx->light_capacity + 2) {
        k_store_environment_light<<<1, 32, 0, ctx->stream>>>(ctx->light_info, ctx->light_capacity, s.sky_w, s.sky_h);
        RT2_CUDA(ctx, cudaGetLastError());
    }
    ctx->stats.kernel_launches = (groups ? 1 : 0) + (ctx->light_records == ctx->light_capacity + 2 ? 1 : 0);
    ctx->stats_from_counters = false;
    ctx->lights_prepared = true;
    return RT2_OK;
}

}  // namespace rt2
