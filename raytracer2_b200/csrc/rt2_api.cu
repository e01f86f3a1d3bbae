// rt2_api.cu — the C ABI of librt2cuda.so (include/rt2_abi.h).  Every entry point catches its own
// failures and returns a code; nothing throws across the boundary.  There is no CPU path: without a
// CUDA device rt2_create fails with RT2_E_NO_DEVICE.
#include "rt2_internal.cuh"

#include <mutex>
#include <new>

namespace rt2 {
static std::mutex g_err_mutex;
static std::string g_create_error;

int fail(rt2_ctx* ctx, int code, const std::string& msg) {
    if (ctx) ctx->last_error = msg;
    else { std::lock_guard<std::mutex> l(g_err_mutex); g_create_error = msg; }
    return code;
}

// lights.cu / mips.cu / post.cu
int lights_prepare(rt2_ctx* ctx);
int mips_generate(rt2_ctx* ctx, int which);
int post_process(rt2_ctx* ctx, uint32_t frame);
void passes_release(rt2_ctx* ctx);          // everything the passes own
void passes_release_frame(rt2_ctx* ctx);    // per-size resources only
void lights_release(rt2_ctx* ctx);          // per-scene light resources
int presample(rt2_ctx* ctx, uint32_t frame, int which);
int render_restir(rt2_ctx* ctx, uint32_t frame, const rt2_launch_params& lp);
bool pass_buffer(rt2_ctx* ctx, int which, void** ptr, size_t* bytes);
}  // namespace rt2

using namespace rt2;

#define RT2_GUARD(ctx) do { if (!(ctx)) return RT2_E_INVALID; cudaError_t e_ = cudaSetDevice((ctx)->device); if (e_ != cudaSuccess) return rt2::fail((ctx), RT2_E_CUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(e_)); } while (0)

extern "C" {

int rt2_abi_version(void) { return RT2_ABI_VERSION; }

int rt2_create(int cuda_device, rt2_ctx** out) {
    if (!out) return RT2_E_INVALID;
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) return fail(nullptr, RT2_E_NO_DEVICE, std::string("no CUDA device (") + cudaGetErrorString(e) + "); librt2cuda has no CPU fallback");
    if (cuda_device < 0 || cuda_device >= n) return fail(nullptr, RT2_E_INVALID, "cuda_device out of range");
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, cuda_device)) != cudaSuccess) return fail(nullptr, RT2_E_CUDA, cudaGetErrorString(e));
    if (prop.major != 10) return fail(nullptr, RT2_E_NO_DEVICE, std::string("device is sm_") + std::to_string(prop.major * 10 + prop.minor) + "; this library is built for sm_100a (B200) only");
    if ((e = cudaSetDevice(cuda_device)) != cudaSuccess) return fail(nullptr, RT2_E_CUDA, cudaGetErrorString(e));
    rt2_ctx* ctx = new (std::nothrow) rt2_ctx();
    if (!ctx) return fail(nullptr, RT2_E_OOM, "out of host memory");
    ctx->device = cuda_device;
    ctx->num_sms = prop.multiProcessorCount;
    if ((e = cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking)) != cudaSuccess) { delete ctx; return fail(nullptr, RT2_E_CUDA, cudaGetErrorString(e)); }
    ctx->stream = ctx->own_stream;
    if ((e = cudaMalloc((void**)&ctx->d_gconst, sizeof(rt2_gconst))) != cudaSuccess) { cudaStreamDestroy(ctx->own_stream); delete ctx; return fail(nullptr, RT2_E_CUDA, cudaGetErrorString(e)); }
    for (int i = 0; i < RT2_FRAMES_IN_FLIGHT && e == cudaSuccess; i++) {
        if ((e = cudaEventCreateWithFlags(&ctx->gconst_done[i], cudaEventDisableTiming)) != cudaSuccess) break;
        if ((e = cudaEventCreateWithFlags(&ctx->snap_ready[i], cudaEventDisableTiming)) != cudaSuccess) break;
        e = cudaEventCreateWithFlags(&ctx->read_done[i], cudaEventDisableTiming);
    }
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaHostAlloc((void**)&ctx->h_gconst_ring, RT2_FRAMES_IN_FLIGHT * sizeof(rt2_gconst), cudaHostAllocDefault);
    if (e != cudaSuccess) { std::string m = cudaGetErrorString(e); rt2_destroy(ctx); return fail(nullptr, RT2_E_CUDA, m); }
    int rc = render_init(ctx);
    if (rc == RT2_OK) rc = noalloc_init(ctx);
    if (rc == RT2_OK) rc = codec_luts_init(ctx);
    if (rc != RT2_OK) { std::string m = ctx->last_error; rt2_destroy(ctx); return fail(nullptr, rc, m); }
    *out = ctx;
    return RT2_OK;
}

int rt2_destroy(rt2_ctx* ctx) {
    if (!ctx) return RT2_E_INVALID;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->copy_stream) { cudaStreamSynchronize(ctx->copy_stream); cudaStreamDestroy(ctx->copy_stream); }
    for (int i = 0; i < RT2_FRAMES_IN_FLIGHT; i++) {
        if (ctx->gconst_done[i]) cudaEventDestroy(ctx->gconst_done[i]);
        if (ctx->snap_ready[i]) cudaEventDestroy(ctx->snap_ready[i]);
        if (ctx->read_done[i]) cudaEventDestroy(ctx->read_done[i]);
        if (ctx->snap[i]) cudaFree(ctx->snap[i]);
    }
    if (ctx->h_gconst_ring) cudaFreeHost(ctx->h_gconst_ring);
    if (ctx->h_k1_stats) cudaFreeHost(ctx->h_k1_stats);
    if (ctx->codec_luts) cudaFree(ctx->codec_luts);
    if (ctx->k1_stats_ready) cudaEventDestroy(ctx->k1_stats_ready);
    scene_release(ctx); frame_release(ctx); passes_release(ctx);
    for (cudaEvent_t e : ctx->prof_events) cudaEventDestroy(e);
    if (ctx->d_gconst) cudaFree(ctx->d_gconst);
    for (uint32_t c = 0; c < RT2_MAX_CHAINS; c++) { if (ctx->chain_streams[c]) cudaStreamDestroy(ctx->chain_streams[c]); if (ctx->chain_join[c]) cudaEventDestroy(ctx->chain_join[c]); }
    if (ctx->chain_fork) cudaEventDestroy(ctx->chain_fork);
    if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
    delete ctx;
    return RT2_OK;
}

int rt2_set_stream(rt2_ctx* ctx, void* cuda_stream) {
    RT2_GUARD(ctx);
    RT2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->stream = cuda_stream ? (cudaStream_t)cuda_stream : ctx->own_stream;
    return RT2_OK;
}

const char* rt2_last_error(rt2_ctx* ctx) {
    if (ctx) return ctx->last_error.c_str();
    std::lock_guard<std::mutex> l(g_err_mutex);
    static thread_local std::string copy;
    copy = g_create_error;
    return copy.c_str();
}

int rt2_upload_scene(rt2_ctx* ctx, const rt2_scene_desc* scene) { RT2_GUARD(ctx); RT2_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); lights_release(ctx); return scene_upload(ctx, scene); }
int rt2_build_accel(rt2_ctx* ctx) { RT2_GUARD(ctx); return scene_build_accel(ctx); }
int rt2_resize(rt2_ctx* ctx, uint32_t w, uint32_t h) { RT2_GUARD(ctx); RT2_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); passes_release_frame(ctx); return frame_resize(ctx, w, h); }

int rt2_set_constants(rt2_ctx* ctx, const void* gconst, size_t nbytes) {
    RT2_GUARD(ctx);
    if (!gconst || nbytes != RT2_GCONST_SIZE) return fail(ctx, RT2_E_INVALID, "GConst must be exactly 1424 bytes");
    memcpy(&ctx->gconst, gconst, RT2_GCONST_SIZE);
    // stream-ordered copy from a context-owned pinned staging slot (replaces the UBO upload, light_passes.rs:665-668).
    // The slots form a ring, one per frame in flight: the call waits only for the copy that last used its slot, so the
    // host can record frame n+1 while frame n still renders (the reference's per-swapchain-image command buffers).
    const uint32_t slot = ctx->gconst_next++ % RT2_FRAMES_IN_FLIGHT;
    RT2_CUDA(ctx, cudaEventSynchronize(ctx->gconst_done[slot]));
    memcpy(&ctx->h_gconst_ring[slot], gconst, RT2_GCONST_SIZE);
    RT2_CUDA(ctx, cudaMemcpyAsync(ctx->d_gconst, &ctx->h_gconst_ring[slot], RT2_GCONST_SIZE, cudaMemcpyHostToDevice, ctx->stream));
    RT2_CUDA(ctx, cudaEventRecord(ctx->gconst_done[slot], ctx->stream));
    ctx->have_gconst = true;
    return RT2_OK;
}

int rt2_prepare_lights(rt2_ctx* ctx) { RT2_GUARD(ctx); return lights_prepare(ctx); }
int rt2_generate_mips(rt2_ctx* ctx, int which) { RT2_GUARD(ctx); return mips_generate(ctx, which); }
int rt2_postprocess(rt2_ctx* ctx, uint32_t frame) { RT2_GUARD(ctx); return post_process(ctx, frame); }
int rt2_presample(rt2_ctx* ctx, uint32_t frame, int which) { RT2_GUARD(ctx); return presample(ctx, frame, which); }

int rt2_render(rt2_ctx* ctx, uint32_t frame, const rt2_launch_params* params) {
    RT2_GUARD(ctx);
    rt2_launch_params lp{};
    if (params) lp = *params;
    if (!ctx->have_gconst) return fail(ctx, RT2_E_STATE, "no constants: call rt2_set_constants first");
    // LightPasses::execute dispatch, src/light_passes.rs:587-659
    if (ctx->gconst.refrence_mode == 1u) return render_reference(ctx, frame, lp);
    return render_restir(ctx, frame, lp);
}

int rt2_trace_primary(rt2_ctx* ctx, const rt2_launch_params* params) {
    RT2_GUARD(ctx);
    rt2_launch_params lp{};
    if (params) lp = *params;
    return render_primary(ctx, lp);
}

int rt2_sync(rt2_ctx* ctx) {
    RT2_GUARD(ctx);
    RT2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->reads_pending) { RT2_CUDA(ctx, cudaStreamSynchronize(ctx->copy_stream)); ctx->reads_pending = false; }
    RT2_CUDA(ctx, cudaGetLastError());
    return RT2_OK;
}

static bool buffer_lookup(rt2_ctx* ctx, int which, void** ptr, size_t* bytes) {
    const FrameBuffers& f = ctx->fb;
    const size_t n = (size_t)f.width * f.height;
    switch (which) {
        case RT2_BUF_DIFFUSE_F16: *ptr = f.diffuse_f16; *bytes = n * 8; return f.diffuse_f16 != nullptr;
        case RT2_BUF_SPECULAR_F16: *ptr = f.specular_f16; *bytes = n * 8; return f.specular_f16 != nullptr;
        case RT2_BUF_DIFFUSE_F32: *ptr = f.diffuse_f32; *bytes = n * 16; return f.diffuse_f32 != nullptr;
        case RT2_BUF_HIT_IDS: *ptr = f.hit_ids; *bytes = n * sizeof(rt2_hit_record); return f.hit_ids != nullptr;
        case RT2_BUF_OUTPUT_RGBA8: *ptr = f.output_rgba8; *bytes = n * 4; return f.output_rgba8 != nullptr;
        case RT2_BUF_BLAS_NODES: *ptr = ctx->scene.blas.nodes; *bytes = (size_t)ctx->scene.blas.num_nodes * sizeof(Node8); return ctx->scene.accel_built;
        case RT2_BUF_BLAS_TRIS: *ptr = ctx->scene.blas.tris; *bytes = (size_t)ctx->scene.blas.num_prims * sizeof(TriPos); return ctx->scene.accel_built && ctx->scene.blas.num_prims;
        case RT2_BUF_BLAS_SLOT_TO_PRIM: *ptr = ctx->scene.blas.slot_to_prim; *bytes = (size_t)ctx->scene.blas.num_prims * 4; return ctx->scene.accel_built && ctx->scene.blas.num_prims;
        case RT2_BUF_TLAS_NODES: *ptr = ctx->scene.tlas.nodes; *bytes = (size_t)ctx->scene.tlas.num_nodes * sizeof(Node8); return ctx->scene.accel_built && ctx->scene.tlas.nodes;
        case RT2_BUF_TLAS_SLOT_TO_INST: *ptr = ctx->scene.tlas.slot_to_prim; *bytes = (size_t)ctx->scene.tlas.num_prims * 4; return ctx->scene.accel_built && ctx->scene.tlas.nodes;
        default: return pass_buffer(ctx, which, ptr, bytes);
    }
}

size_t rt2_buffer_size(rt2_ctx* ctx, int which) {
    if (!ctx) return 0;
    void* p; size_t b = 0;
    return buffer_lookup(ctx, which, &p, &b) ? b : 0;
}

int rt2_read(rt2_ctx* ctx, int which, void* dst, size_t nbytes) {
    RT2_GUARD(ctx);
    void* p; size_t b = 0;
    if (!dst) return fail(ctx, RT2_E_INVALID, "dst is NULL");
    if (!buffer_lookup(ctx, which, &p, &b)) return fail(ctx, RT2_E_STATE, "buffer not allocated");
    if (nbytes != b) return fail(ctx, RT2_E_INVALID, "size mismatch: buffer is " + std::to_string(b) + " bytes");
    RT2_CUDA(ctx, cudaMemcpyAsync(dst, p, b, cudaMemcpyDeviceToHost, ctx->stream));
    RT2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return RT2_OK;
}

// Frame-in-flight read: the buffer is snapshotted on the context's stream (device -> device, microseconds) and the
// snapshot travels to the host on the context's copy stream, so the next rt2_render overlaps the PCIe transfer.
int rt2_read_async(rt2_ctx* ctx, int which, void* dst, size_t nbytes) {
    RT2_GUARD(ctx);
    void* p; size_t b = 0;
    if (!dst) return fail(ctx, RT2_E_INVALID, "dst is NULL");
    if (!buffer_lookup(ctx, which, &p, &b)) return fail(ctx, RT2_E_STATE, "buffer not allocated");
    if (nbytes != b) return fail(ctx, RT2_E_INVALID, "size mismatch: buffer is " + std::to_string(b) + " bytes");
    const uint32_t i = ctx->snap_next++ % RT2_FRAMES_IN_FLIGHT;
    RT2_CUDA(ctx, cudaEventSynchronize(ctx->read_done[i]));          // the transfer that last used this snapshot (two reads ago)
    if (ctx->snap_bytes[i] < b) {
        if (ctx->snap[i]) { RT2_CUDA(ctx, cudaFree(ctx->snap[i])); ctx->device_bytes -= ctx->snap_bytes[i]; ctx->snap[i] = nullptr; ctx->snap_bytes[i] = 0; }
        if (cudaMalloc(&ctx->snap[i], b) != cudaSuccess) { cudaGetLastError(); return fail(ctx, RT2_E_OOM, "out of device memory for the read snapshot"); }
        ctx->snap_bytes[i] = b; ctx->device_bytes += b;
    }
    RT2_CUDA(ctx, cudaMemcpyAsync(ctx->snap[i], p, b, cudaMemcpyDeviceToDevice, ctx->stream));
    RT2_CUDA(ctx, cudaEventRecord(ctx->snap_ready[i], ctx->stream));
    RT2_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, ctx->snap_ready[i], 0));
    RT2_CUDA(ctx, cudaMemcpyAsync(dst, ctx->snap[i], b, cudaMemcpyDeviceToHost, ctx->copy_stream));
    RT2_CUDA(ctx, cudaEventRecord(ctx->read_done[i], ctx->copy_stream));
    ctx->reads_pending = true;
    return RT2_OK;
}

// Orders everything enqueued on the context's stream after this call behind the asynchronous reads issued so far
// (an event recorded on the stream afterwards then also covers their transfers).  Does not block the host.
int rt2_read_fence(rt2_ctx* ctx) {
    RT2_GUARD(ctx);
    if (!ctx->reads_pending) return RT2_OK;
    for (int i = 0; i < RT2_FRAMES_IN_FLIGHT; i++) RT2_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->read_done[i], 0));
    return RT2_OK;
}

int rt2_copy_to_device(rt2_ctx* ctx, int which, void* device_dst, size_t nbytes, uint32_t row0, uint32_t row1) {
    RT2_GUARD(ctx);
    void* p; size_t b = 0;
    if (!device_dst) return fail(ctx, RT2_E_INVALID, "device_dst is NULL");
    if (!buffer_lookup(ctx, which, &p, &b)) return fail(ctx, RT2_E_STATE, "buffer not allocated");
    size_t off = 0, len = b;
    if (row0 != 0 || row1 != 0) {
        const uint32_t H = ctx->fb.height;
        if (which > RT2_BUF_OUTPUT_RGBA8 || row1 > H || row0 >= row1) return fail(ctx, RT2_E_INVALID, "bad row range");
        const size_t row_bytes = b / H;
        off = row_bytes * row0; len = row_bytes * (row1 - row0);
    }
    if (nbytes != len) return fail(ctx, RT2_E_INVALID, "size mismatch: range is " + std::to_string(len) + " bytes");
    RT2_CUDA(ctx, cudaMemcpyAsync(device_dst, (const char*)p + off, len, cudaMemcpyDeviceToDevice, ctx->stream));
    return RT2_OK;
}

// All row bands of one rank (band b -> rank b mod world) packed back to back into caller-owned device memory:
// one strided copy for the full-height bands plus one for a ragged last band, instead of a copy per band.
int rt2_pack_bands(rt2_ctx* ctx, int which, void* device_dst, size_t nbytes, uint32_t band_rows, uint32_t band_rank, uint32_t band_world) {
    RT2_GUARD(ctx);
    void* p; size_t b = 0;
    if (!device_dst) return fail(ctx, RT2_E_INVALID, "device_dst is NULL");
    if (which > RT2_BUF_OUTPUT_RGBA8) return fail(ctx, RT2_E_INVALID, "rt2_pack_bands: not a W x H frame buffer");
    if (!buffer_lookup(ctx, which, &p, &b)) return fail(ctx, RT2_E_STATE, "buffer not allocated");
    if (band_rows == 0 || band_world == 0 || band_rank >= band_world) return fail(ctx, RT2_E_INVALID, "bad band parameters");
    const uint32_t H = ctx->fb.height;
    const size_t row_bytes = b / H;
    const uint32_t num_bands = (H + band_rows - 1) / band_rows;
    uint32_t mine = 0, full = 0, last_rows = 0;
    for (uint32_t band = band_rank; band < num_bands; band += band_world) {
        const uint32_t rows = std::min(band_rows, H - band * band_rows);
        if (rows == band_rows) full++; else last_rows = rows;      // only the last band of the frame can be ragged
        mine++;
    }
    const size_t need = ((size_t)full * band_rows + last_rows) * row_bytes;
    if (nbytes != need) return fail(ctx, RT2_E_INVALID, "size mismatch: this rank's bands are " + std::to_string(need) + " bytes");
    const char* src = (const char*)p + (size_t)band_rank * band_rows * row_bytes;
    if (full) RT2_CUDA(ctx, cudaMemcpy2DAsync(device_dst, (size_t)band_rows * row_bytes, src, (size_t)band_rows * band_world * row_bytes,
                                             (size_t)band_rows * row_bytes, full, cudaMemcpyDeviceToDevice, ctx->stream));
    if (last_rows) RT2_CUDA(ctx, cudaMemcpyAsync((char*)device_dst + (size_t)full * band_rows * row_bytes, src + (size_t)full * band_rows * band_world * row_bytes,
                                                 (size_t)last_rows * row_bytes, cudaMemcpyDeviceToDevice, ctx->stream));
    (void)mine;
    return RT2_OK;
}

int rt2_write(rt2_ctx* ctx, int which, const void* src, size_t nbytes) {
    RT2_GUARD(ctx);
    void* p; size_t b = 0;
    if (!src) return fail(ctx, RT2_E_INVALID, "src is NULL");
    if (!buffer_lookup(ctx, which, &p, &b)) return fail(ctx, RT2_E_STATE, "buffer not allocated");
    if (nbytes != b) return fail(ctx, RT2_E_INVALID, "size mismatch: buffer is " + std::to_string(b) + " bytes");
    RT2_CUDA(ctx, cudaMemcpyAsync(p, src, b, cudaMemcpyHostToDevice, ctx->stream));
    RT2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return RT2_OK;
}

int rt2_get_stats(rt2_ctx* ctx, rt2_stats* out) {
    RT2_GUARD(ctx);
    if (!out) return fail(ctx, RT2_E_INVALID, "out is NULL");
    if (ctx->stats_from_counters && ctx->fb.counters) {
        unsigned long long c[3] = {0, 0, 0};
        RT2_CUDA(ctx, cudaMemcpyAsync(c, ctx->fb.counters, 24, cudaMemcpyDeviceToHost, ctx->stream));
        RT2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        ctx->stats.rays_traced = c[0] + ctx->pending_rays_static;
        ctx->stats.rays_hit = c[1];
        ctx->stats.rays_walked = ctx->stats_walks_counted ? c[2] : ctx->stats.rays_traced;
        if (ctx->stats_reused_counted && !ctx->stats_walks_counted) {
            unsigned long long reused = 0;
            RT2_CUDA(ctx, cudaMemcpyAsync(&reused, ctx->fb.counters + 16, 8, cudaMemcpyDeviceToHost, ctx->stream));
            RT2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            ctx->stats.rays_walked = ctx->stats.rays_traced - reused;
        }
    }
    ctx->stats.trace_ns = ctx->stats.shade_ns = ctx->stats.trace_launches = 0;
    if (!ctx->prof_events.empty()) {
        RT2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        double tr = 0, sh = 0; uint64_t n = 0;
        for (size_t it = 0; it + 2 < ctx->prof_events.size() + 0 && 3 * it + 2 < ctx->prof_events.size(); it++) {
            float a = 0, b = 0;
            RT2_CUDA(ctx, cudaEventElapsedTime(&a, ctx->prof_events[3 * it], ctx->prof_events[3 * it + 1]));
            RT2_CUDA(ctx, cudaEventElapsedTime(&b, ctx->prof_events[3 * it + 1], ctx->prof_events[3 * it + 2]));
            tr += a; sh += b; n++;
        }
        ctx->stats.trace_ns = (uint64_t)(tr * 1e6); ctx->stats.shade_ns = (uint64_t)(sh * 1e6); ctx->stats.trace_launches = n;
    }
    ctx->stats.device_bytes = ctx->device_bytes;
    ctx->stats.bvh_nodes = (uint64_t)ctx->scene.blas.num_nodes + ctx->scene.tlas.num_nodes;
    ctx->stats.bvh_triangles = ctx->scene.num_tris;
    *out = ctx->stats;
    return RT2_OK;
}

}  // extern "C"
