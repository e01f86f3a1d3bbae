// rt2_group.cu — one frame over the GPUs of one node behind the C ABI (include/rt2_abi.h, "multi-GPU").
// Band plan (interleaved row bands, band b -> rank b mod N) in C++, one rt2_ctx per GPU, NCCL all-gather of the finished
// bands issued by the library on a communication stream of its own, double-buffered so that the gather of frame n overlaps
// the render of frame n + 1, and a device kernel that un-permutes the rank-major slots into the assembled frame.
// NCCL is resolved with dlopen("libnccl.so.2") when a group is created: librt2cuda.so has no link-time dependency on it.
#include "rt2_internal.cuh"

#include <dlfcn.h>

#include <algorithm>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <new>
#include <thread>

namespace rt2 {

// ---- NCCL, resolved at run time (prototypes as in nccl.h 2.x; only what the band gather needs) ---------------------
typedef struct ncclComm* nccl_comm_t;
typedef struct { char internal[128]; } nccl_unique_id;
struct NcclApi {
    void* lib = nullptr;
    int (*GetUniqueId)(nccl_unique_id*) = nullptr;
    int (*CommInitRank)(nccl_comm_t*, int, nccl_unique_id, int) = nullptr;
    int (*CommInitAll)(nccl_comm_t*, int, const int*) = nullptr;
    int (*CommDestroy)(nccl_comm_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int /* ncclDataType_t */, nccl_comm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
static std::mutex g_nccl_mutex;
static NcclApi g_nccl;
static std::string g_group_error;
static const int NCCL_INT8 = 0;      // ncclInt8 / ncclChar

static bool nccl_load(std::string& err) {
    std::lock_guard<std::mutex> l(g_nccl_mutex);
    if (g_nccl.lib) return true;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);       // a host that already loaded NCCL (torch) gets that copy
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) { err = std::string("cannot load NCCL (libnccl.so.2): ") + dlerror(); return false; }
    NcclApi a; a.lib = h;
#define RT2_NCCL_SYM(field, name) *(void**)(&a.field) = dlsym(h, name); if (!a.field) { err = std::string("NCCL symbol missing: ") + name; dlclose(h); return false; }
    RT2_NCCL_SYM(GetUniqueId, "ncclGetUniqueId") RT2_NCCL_SYM(CommInitRank, "ncclCommInitRank") RT2_NCCL_SYM(CommInitAll, "ncclCommInitAll")
    RT2_NCCL_SYM(CommDestroy, "ncclCommDestroy") RT2_NCCL_SYM(AllGather, "ncclAllGather") RT2_NCCL_SYM(GroupStart, "ncclGroupStart")
    RT2_NCCL_SYM(GroupEnd, "ncclGroupEnd") RT2_NCCL_SYM(GetErrorString, "ncclGetErrorString")
#undef RT2_NCCL_SYM
    g_nccl = a;
    return true;
}

// ---- band plan (the C++ side of raytracer2_b200/sharding.py; same rule the kernels apply, rt2_launch_params.band_*) ----
struct BandPlanC {
    uint32_t height = 0, band_rows = RT2_GROUP_BAND_ROWS, world = 1;
    uint32_t rows(uint32_t rank) const {
        uint32_t n = 0;
        for (uint32_t y0 = rank * band_rows; y0 < height; y0 += world * band_rows) n += std::min(band_rows, height - y0);
        return n;
    }
    uint32_t max_rows() const { uint32_t m = 0; for (uint32_t r = 0; r < world; r++) m = std::max(m, rows(r)); return m; }
};

// gathered [world][max_rows][row_bytes] (rank-major slots, a rank's bands packed top to bottom) -> frame [height][row_bytes]
__global__ void __launch_bounds__(256) k_unpermute_bands(const uint4* __restrict__ gathered, uint4* __restrict__ frame, uint32_t height, uint32_t row_vec,
                                                         uint32_t band_rows, uint32_t world, uint32_t max_rows) {
    const uint64_t n = (uint64_t)height * row_vec;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t y = (uint32_t)(i / row_vec), x = (uint32_t)(i % row_vec);
        const uint32_t band = y / band_rows, rank = band % world;
        const uint32_t packed_row = (band / world) * band_rows + (y % band_rows);
        frame[i] = gathered[((uint64_t)rank * max_rows + packed_row) * row_vec + x];
    }
}

// ---- in-frame exchange of per-pixel buffers (K6 / K7 under sharding) ---------------------------------------------------------------
// K6 reads the previous frame's G-buffer and GI reservoirs wherever the motion vectors point, K7 the current ones within its
// sampling radius (rtxdi/GIResamplingFunctions.hlsli:186-359, 391-553) — rows other ranks own.  Each rank packs the entries of ITS
// pixels, one ncclAllGather moves them, each rank writes the other ranks' entries where they belong.  An entry is `words` uint32;
// layout 0 = row-major W x H, 1 = the reservoirs' 16 x 16 block-linear order (RTXDI_ReservoirPositionToPointer, RtxdiHelpers.hlsli:75-88).
struct ExchangeSeg { uint32_t* base; uint32_t words, layout; size_t slot_offset_words; };

__device__ __forceinline__ size_t exchange_entry(uint32_t x, uint32_t y, uint32_t width, uint32_t layout, uint32_t block_row_pitch) {
    return layout == 0u ? (size_t)y * width + x : (size_t)(y >> 4) * block_row_pitch + (size_t)(x >> 4) * 256u + (y & 15u) * 16u + (x & 15u);
}
#define RT2_EXCHANGE_MAX_SEGS 6
struct ExchangeSegs { ExchangeSeg s[RT2_EXCHANGE_MAX_SEGS]; uint32_t n; };

// one thread per pixel, every segment of the exchange in one launch; 8-word entries (reservoirs) move as two uint4
__device__ __forceinline__ void exchange_copy(uint32_t* __restrict__ dst, const uint32_t* __restrict__ src, uint32_t words) {
    if (words == 8u) {
        const uint4 a = reinterpret_cast<const uint4*>(src)[0], b = reinterpret_cast<const uint4*>(src)[1];
        reinterpret_cast<uint4*>(dst)[0] = a; reinterpret_cast<uint4*>(dst)[1] = b;
    } else {
        for (uint32_t w = 0; w < words; w++) dst[w] = src[w];
    }
}
__global__ void __launch_bounds__(256) k_exchange_pack(ExchangeSegs segs, uint32_t* __restrict__ send, uint32_t width, uint32_t local_rows, uint32_t band_rows,
                                                       uint32_t rank, uint32_t world, uint32_t block_row_pitch) {
    const uint32_t n = local_rows * width;
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x) {
        const uint32_t packed_row = e / width, x = e - packed_row * width;
        const uint32_t y = ((packed_row / band_rows) * world + rank) * band_rows + packed_row % band_rows;
        for (uint32_t k = 0; k < segs.n; k++) {
            const ExchangeSeg& sg = segs.s[k];
            exchange_copy(send + sg.slot_offset_words + (size_t)e * sg.words, sg.base + exchange_entry(x, y, width, sg.layout, block_row_pitch) * sg.words, sg.words);
        }
    }
}
__global__ void __launch_bounds__(256) k_exchange_unpack(ExchangeSegs segs, const uint32_t* __restrict__ recv, size_t slot_words, uint32_t width, uint32_t height,
                                                         uint32_t band_rows, uint32_t rank, uint32_t world, uint32_t block_row_pitch) {
    const uint32_t n = height * width;
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x) {
        const uint32_t y = e / width, x = e - y * width;
        const uint32_t band = y / band_rows, owner = band % world;
        if (owner == rank) continue;                                   // this rank's own entries are already in place
        const uint32_t packed_row = (band / world) * band_rows + y % band_rows;
        const uint32_t* slot = recv + (size_t)owner * slot_words;
        for (uint32_t k = 0; k < segs.n; k++) {
            const ExchangeSeg& sg = segs.s[k];
            exchange_copy(sg.base + exchange_entry(x, y, width, sg.layout, block_row_pitch) * sg.words,
                          slot + sg.slot_offset_words + ((size_t)packed_row * width + x) * sg.words, sg.words);
        }
    }
}

}  // namespace rt2

using namespace rt2;

#define RT2_GROUP_BUFFERS 2
// One host thread per local GPU beyond the first (single-process groups): a frame is 60-270 kernel launches per GPU, and one thread
// issuing them for 8 GPUs in turn takes longer than a GPU needs to run its share (measured: C5 on 8 GPUs 7.4 ms per frame, of which
// ~4 ms of GPU work).  The caller's thread drives context 0 and waits for the others.
struct GroupWorker {
    std::thread th; std::mutex m; std::condition_variable cv;
    std::function<int()> job; bool has_job = false, quit = false; int rc = 0;
    void loop() {
        for (;;) {
            std::function<int()> j;
            { std::unique_lock<std::mutex> l(m); cv.wait(l, [&] { return has_job || quit; }); if (quit && !has_job) return; j = job; }
            const int r = j();
            { std::lock_guard<std::mutex> l(m); rc = r; has_job = false; }
            cv.notify_all();
        }
    }
    void start(std::function<int()> j) { { std::lock_guard<std::mutex> l(m); job = std::move(j); has_job = true; } cv.notify_all(); }
    int wait() { std::unique_lock<std::mutex> l(m); cv.wait(l, [&] { return !has_job; }); return rc; }
    void stop() { { std::lock_guard<std::mutex> l(m); quit = true; } cv.notify_all(); if (th.joinable()) th.join(); }
};
struct rt2_group {
    int world = 1, first_rank = 0;                    // this process owns ranks [first_rank, first_rank + ctxs.size())
    std::vector<rt2_ctx*> ctxs;
    std::vector<nccl_comm_t> comms;
    std::vector<cudaStream_t> comm_streams;
    // per local rank, double-buffered: packed bands, gathered slots, assembled frame; ready = bands packed (render stream),
    // done = assembled (communication stream)
    struct Lane {
        void* send[RT2_GROUP_BUFFERS] = {}; void* recv[RT2_GROUP_BUFFERS] = {}; void* assembled[RT2_GROUP_BUFFERS] = {};
        cudaEvent_t ready[RT2_GROUP_BUFFERS] = {}, done[RT2_GROUP_BUFFERS] = {}; cudaEvent_t tail = nullptr;
        size_t cap_slot = 0, cap_frame = 0;
        uint32_t* xsend = nullptr; uint32_t* xrecv = nullptr; size_t cap_xslot = 0;     // in-frame exchange staging (K6 / K7), on the render stream
    };
    std::vector<Lane> lanes;
    uint32_t width = 0, height = 0;
    uint64_t gathers = 0;                             // gathers issued; buffer of gather k is k % RT2_GROUP_BUFFERS
    uint64_t exchanges = 0;                           // in-frame exchanges issued (K6 / K7 inputs)
    std::vector<GroupWorker*> workers;                // workers[i] drives ctxs[i], i >= 1 (workers[0] is null)
    // fn(i) for every local context, in parallel; returns the first non-zero code (and its context index)
    int each(const std::function<int(size_t)>& fn, size_t* failed = nullptr) {
        const size_t n = ctxs.size();
        for (size_t i = 1; i < n; i++) workers[i]->start([&fn, i] { return fn(i); });
        int rc = fn(0); size_t who = 0;
        for (size_t i = 1; i < n; i++) { const int r = workers[i]->wait(); if (r != 0 && rc == 0) { rc = r; who = i; } }
        if (failed) *failed = who;
        return rc;
    }
    size_t frame_bytes = 0;                           // of the last gathered buffer
    std::string last_error;
};

static int gfail(rt2_group* g, int code, const std::string& msg) {
    if (g) g->last_error = msg;
    else { std::lock_guard<std::mutex> l(g_nccl_mutex); g_group_error = msg; }
    return code;
}
#define RT2_GCUDA(g, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return gfail((g), RT2_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)
#define RT2_GNCCL(g, call) do { int r_ = (call); if (r_ != 0) return gfail((g), RT2_E_CUDA, std::string(#call) + ": " + g_nccl.GetErrorString(r_)); } while (0)
// a per-context call failed: carry its message up
#define RT2_GCTX(g, c, call) do { int r_ = (call); if (r_ != RT2_OK) return gfail((g), r_, std::string(#call) + ": " + rt2_last_error(c)); } while (0)

static int group_finish_create(rt2_group* g) {
    const size_t n = g->ctxs.size();
    g->comm_streams.assign(n, nullptr);
    g->lanes.assign(n, rt2_group::Lane());
    g->workers.assign(n, nullptr);
    for (size_t i = 1; i < n; i++) { g->workers[i] = new GroupWorker(); GroupWorker* w = g->workers[i]; w->th = std::thread([w] { w->loop(); }); }
    for (size_t i = 0; i < n; i++) {
        RT2_GCUDA(g, cudaSetDevice(g->ctxs[i]->device));
        RT2_GCUDA(g, cudaStreamCreateWithFlags(&g->comm_streams[i], cudaStreamNonBlocking));
        for (int b = 0; b < RT2_GROUP_BUFFERS; b++) {
            RT2_GCUDA(g, cudaEventCreateWithFlags(&g->lanes[i].ready[b], cudaEventDisableTiming));
            RT2_GCUDA(g, cudaEventCreateWithFlags(&g->lanes[i].done[b], cudaEventDisableTiming));
        }
        RT2_GCUDA(g, cudaEventCreateWithFlags(&g->lanes[i].tail, cudaEventDisableTiming));
    }
    return RT2_OK;
}

extern "C" {

const char* rt2_group_last_error(rt2_group* g) {
    if (g) return g->last_error.c_str();
    std::lock_guard<std::mutex> l(g_nccl_mutex);
    static thread_local std::string copy;
    copy = g_group_error;
    return copy.c_str();
}

int rt2_group_destroy(rt2_group* g) {
    if (!g) return RT2_E_INVALID;
    for (GroupWorker* w : g->workers) if (w) { w->stop(); delete w; }
    g->workers.clear();
    for (size_t i = 0; i < g->ctxs.size(); i++) {
        if (!g->ctxs[i]) continue;
        cudaSetDevice(g->ctxs[i]->device);
        cudaStreamSynchronize(g->ctxs[i]->stream);
        if (i < g->comm_streams.size() && g->comm_streams[i]) cudaStreamSynchronize(g->comm_streams[i]);
    }
    for (size_t i = 0; i < g->comms.size(); i++) if (g->comms[i] && g_nccl.CommDestroy) g_nccl.CommDestroy(g->comms[i]);
    for (size_t i = 0; i < g->ctxs.size(); i++) {
        if (!g->ctxs[i]) continue;
        cudaSetDevice(g->ctxs[i]->device);
        if (i < g->lanes.size()) {
            rt2_group::Lane& L = g->lanes[i];
            for (int b = 0; b < RT2_GROUP_BUFFERS; b++) {
                if (L.send[b]) cudaFree(L.send[b]);
                if (L.recv[b]) cudaFree(L.recv[b]);
                if (L.assembled[b]) cudaFree(L.assembled[b]);
                if (L.ready[b]) cudaEventDestroy(L.ready[b]);
                if (L.done[b]) cudaEventDestroy(L.done[b]);
            }
            if (L.tail) cudaEventDestroy(L.tail);
            if (L.xsend) cudaFree(L.xsend);
            if (L.xrecv) cudaFree(L.xrecv);
        }
        if (i < g->comm_streams.size() && g->comm_streams[i]) cudaStreamDestroy(g->comm_streams[i]);
        rt2_destroy(g->ctxs[i]);
    }
    delete g;
    return RT2_OK;
}

int rt2_group_create(const int* cuda_devices, int n, rt2_group** out) {
    if (!out) return RT2_E_INVALID;
    *out = nullptr;
    if (!cuda_devices || n < 1 || n > 64) return gfail(nullptr, RT2_E_INVALID, "rt2_group_create: need 1..64 devices");
    for (int i = 0; i < n; i++) for (int j = 0; j < i; j++) if (cuda_devices[i] == cuda_devices[j]) return gfail(nullptr, RT2_E_INVALID, "rt2_group_create: a device appears twice");
    rt2_group* g = new (std::nothrow) rt2_group();
    if (!g) return gfail(nullptr, RT2_E_OOM, "out of host memory");
    g->world = n; g->first_rank = 0;
    g->ctxs.assign((size_t)n, nullptr);
    for (int i = 0; i < n; i++) {
        int rc = rt2_create(cuda_devices[i], &g->ctxs[(size_t)i]);
        if (rc != RT2_OK) { std::string m = rt2_last_error(nullptr); rt2_group_destroy(g); return gfail(nullptr, rc, "rt2_create: " + m); }
    }
    if (n > 1) {
        std::string err;
        if (!nccl_load(err)) { rt2_group_destroy(g); return gfail(nullptr, RT2_E_STATE, err); }
        g->comms.assign((size_t)n, nullptr);
        int r = g_nccl.CommInitAll(g->comms.data(), n, cuda_devices);
        if (r != 0) { std::string m = g_nccl.GetErrorString(r); rt2_group_destroy(g); return gfail(nullptr, RT2_E_CUDA, "ncclCommInitAll: " + m); }
    }
    int rc = group_finish_create(g);
    if (rc != RT2_OK) { std::string m = g->last_error; rt2_group_destroy(g); return gfail(nullptr, rc, m); }
    *out = g;
    return RT2_OK;
}

int rt2_group_unique_id(void* id128) {
    if (!id128) return RT2_E_INVALID;
    std::string err;
    if (!nccl_load(err)) return gfail(nullptr, RT2_E_STATE, err);
    nccl_unique_id id;
    int r = g_nccl.GetUniqueId(&id);
    if (r != 0) return gfail(nullptr, RT2_E_CUDA, std::string("ncclGetUniqueId: ") + g_nccl.GetErrorString(r));
    memcpy(id128, &id, RT2_UNIQUE_ID_BYTES);
    return RT2_OK;
}

int rt2_group_create_rank(int cuda_device, int rank, int world, const void* id128, rt2_group** out) {
    if (!out) return RT2_E_INVALID;
    *out = nullptr;
    if (world < 1 || rank < 0 || rank >= world) return gfail(nullptr, RT2_E_INVALID, "rt2_group_create_rank: rank must be in [0, world)");
    if (world > 1 && !id128) return gfail(nullptr, RT2_E_INVALID, "rt2_group_create_rank: id128 is NULL (rt2_group_unique_id on rank 0, distributed by the host)");
    rt2_group* g = new (std::nothrow) rt2_group();
    if (!g) return gfail(nullptr, RT2_E_OOM, "out of host memory");
    g->world = world; g->first_rank = rank;
    g->ctxs.assign(1, nullptr);
    int rc = rt2_create(cuda_device, &g->ctxs[0]);
    if (rc != RT2_OK) { std::string m = rt2_last_error(nullptr); rt2_group_destroy(g); return gfail(nullptr, rc, "rt2_create: " + m); }
    if (world > 1) {
        std::string err;
        if (!nccl_load(err)) { rt2_group_destroy(g); return gfail(nullptr, RT2_E_STATE, err); }
        nccl_unique_id id;
        memcpy(&id, id128, RT2_UNIQUE_ID_BYTES);
        g->comms.assign(1, nullptr);
        cudaSetDevice(cuda_device);
        int r = g_nccl.CommInitRank(&g->comms[0], world, id, rank);
        if (r != 0) { std::string m = g_nccl.GetErrorString(r); rt2_group_destroy(g); return gfail(nullptr, RT2_E_CUDA, "ncclCommInitRank: " + m); }
    }
    rc = group_finish_create(g);
    if (rc != RT2_OK) { std::string m = g->last_error; rt2_group_destroy(g); return gfail(nullptr, rc, m); }
    *out = g;
    return RT2_OK;
}

int rt2_group_world(rt2_group* g) { return g ? g->world : 0; }
int rt2_group_local_count(rt2_group* g) { return g ? (int)g->ctxs.size() : 0; }
rt2_ctx* rt2_group_ctx(rt2_group* g, int i) { return (g && i >= 0 && (size_t)i < g->ctxs.size()) ? g->ctxs[(size_t)i] : nullptr; }

}  // extern "C"

// ---- per-group passes: the same call on every local context (the scene is replicated) -----------------------------------
#define RT2_GROUP_EACH(g, expr) do { if (!(g)) return RT2_E_INVALID; for (size_t i_ = 0; i_ < (g)->ctxs.size(); i_++) { rt2_ctx* c = (g)->ctxs[i_]; RT2_GCTX((g), c, (expr)); } return RT2_OK; } while (0)

extern "C" {

int rt2_group_upload_scene(rt2_group* g, const rt2_scene_desc* scene) { RT2_GROUP_EACH(g, rt2_upload_scene(c, scene)); }
int rt2_group_build_accel(rt2_group* g) { RT2_GROUP_EACH(g, rt2_build_accel(c)); }
int rt2_group_set_constants(rt2_group* g, const void* gconst, size_t nbytes) { RT2_GROUP_EACH(g, rt2_set_constants(c, gconst, nbytes)); }
int rt2_group_prepare_lights(rt2_group* g) { RT2_GROUP_EACH(g, rt2_prepare_lights(c)); }
int rt2_group_generate_mips(rt2_group* g, int which) { RT2_GROUP_EACH(g, rt2_generate_mips(c, which)); }
int rt2_group_postprocess(rt2_group* g, uint32_t frame) {
    if (!g) return RT2_E_INVALID;
    for (size_t i = 0; i < g->ctxs.size(); i++) {
        rt2_ctx* c = g->ctxs[i];
        // K13 is per pixel: every rank tonemaps the rows of its own bands only (the others hold stale lighting anyway)
        c->post_band[0] = g->world > 1 ? RT2_GROUP_BAND_ROWS : 0u; c->post_band[1] = (uint32_t)g->first_rank + (uint32_t)i; c->post_band[2] = (uint32_t)g->world;
        const int rc = rt2_postprocess(c, frame);
        c->post_band[0] = 0u;
        if (rc != RT2_OK) return gfail(g, rc, std::string("rt2_postprocess: ") + rt2_last_error(c));
    }
    return RT2_OK;
}

int rt2_group_resize(rt2_group* g, uint32_t width, uint32_t height) {
    if (!g) return RT2_E_INVALID;
    for (rt2_ctx* c : g->ctxs) RT2_GCTX(g, c, rt2_resize(c, width, height));
    g->width = width; g->height = height;
    return RT2_OK;
}

int rt2_group_sync(rt2_group* g) {
    if (!g) return RT2_E_INVALID;
    for (size_t i = 0; i < g->ctxs.size(); i++) {
        RT2_GCTX(g, g->ctxs[i], rt2_sync(g->ctxs[i]));
        RT2_GCUDA(g, cudaStreamSynchronize(g->comm_streams[i]));
    }
    return RT2_OK;
}

}  // extern "C"

// One exchange: `segs(i)` lists local context i's buffers (same shapes on every rank).  Runs on the render streams: the passes
// behind it wait for it anyway.
template <typename SegsOf>
static int group_exchange(rt2_group* g, SegsOf segs_of) {
    const size_t n = g->ctxs.size();
    const uint32_t world = (uint32_t)g->world, W = g->width, H = g->height;
    BandPlanC plan; plan.height = H; plan.world = world;
    const uint32_t max_rows = plan.max_rows();
    std::vector<std::vector<ExchangeSeg>> all(n);
    size_t slot_words = 0;
    for (size_t i = 0; i < n; i++) {
        all[i] = segs_of(i);
        // 8-word entries first and slots padded to 16 bytes: they move as uint4 pairs
        std::stable_sort(all[i].begin(), all[i].end(), [](const ExchangeSeg& a, const ExchangeSeg& b) { return a.words > b.words; });
        size_t off = 0;
        for (ExchangeSeg& sg : all[i]) { sg.slot_offset_words = off; off += (size_t)max_rows * W * sg.words; }
        slot_words = (off + 3u) & ~(size_t)3u;
    }
    for (size_t i = 0; i < n; i++) {
        rt2_group::Lane& L = g->lanes[i];
        rt2_ctx* c = g->ctxs[i];
        RT2_GCUDA(g, cudaSetDevice(c->device));
        if (L.cap_xslot < slot_words) {
            RT2_GCUDA(g, cudaStreamSynchronize(c->stream));
            if (L.xsend) cudaFree(L.xsend); if (L.xrecv) cudaFree(L.xrecv);
            L.xsend = L.xrecv = nullptr;
            RT2_GCUDA(g, cudaMalloc((void**)&L.xsend, slot_words * 4)); RT2_GCUDA(g, cudaMalloc((void**)&L.xrecv, slot_words * 4 * world));
            RT2_GCUDA(g, cudaMemsetAsync(L.xsend, 0, slot_words * 4, c->stream));
            L.cap_xslot = slot_words;
        }
        const uint32_t rank = (uint32_t)g->first_rank + (uint32_t)i, rows = plan.rows(rank);
        if (all[i].size() > RT2_EXCHANGE_MAX_SEGS) return gfail(g, RT2_E_INVALID, "too many exchange segments");
        ExchangeSegs segs{}; segs.n = (uint32_t)all[i].size();
        for (size_t k = 0; k < all[i].size(); k++) segs.s[k] = all[i][k];
        {
            const uint64_t items = (uint64_t)rows * W;
            const int blocks = (int)std::min<uint64_t>((uint64_t)c->num_sms * 8u, (items + 255u) / 256u);
            k_exchange_pack<<<std::max(blocks, 1), 256, 0, c->stream>>>(segs, L.xsend, W, rows, RT2_GROUP_BAND_ROWS, rank, world, c->fb.reservoir_block_row_pitch);
        }
        RT2_GCUDA(g, cudaGetLastError());
    }
    RT2_GNCCL(g, g_nccl.GroupStart());
    for (size_t i = 0; i < n; i++) {
        int r = g_nccl.AllGather(g->lanes[i].xsend, g->lanes[i].xrecv, slot_words * 4, NCCL_INT8, g->comms[i], g->ctxs[i]->stream);
        if (r != 0) { g_nccl.GroupEnd(); return gfail(g, RT2_E_CUDA, std::string("ncclAllGather: ") + g_nccl.GetErrorString(r)); }
    }
    RT2_GNCCL(g, g_nccl.GroupEnd());
    for (size_t i = 0; i < n; i++) {
        rt2_ctx* c = g->ctxs[i];
        RT2_GCUDA(g, cudaSetDevice(c->device));
        const uint32_t rank = (uint32_t)g->first_rank + (uint32_t)i;
        ExchangeSegs segs{}; segs.n = (uint32_t)all[i].size();
        for (size_t k = 0; k < all[i].size(); k++) segs.s[k] = all[i][k];
        {
            const uint64_t items = (uint64_t)H * W;
            const int blocks = (int)std::min<uint64_t>((uint64_t)c->num_sms * 8u, (items + 255u) / 256u);
            k_exchange_unpack<<<std::max(blocks, 1), 256, 0, c->stream>>>(segs, g->lanes[i].xrecv, slot_words, W, H, RT2_GROUP_BAND_ROWS, rank, world, c->fb.reservoir_block_row_pitch);
        }
        RT2_GCUDA(g, cudaGetLastError());
    }
    g->exchanges++;
    return RT2_OK;
}

extern "C" {

int rt2_group_render(rt2_group* g, uint32_t frame, const rt2_launch_params* params, int gather_which) {
    if (!g) return RT2_E_INVALID;
    if (!g->height) return gfail(g, RT2_E_STATE, "rt2_group_render before rt2_group_resize");
    const size_t n = g->ctxs.size();
    const uint32_t world = (uint32_t)g->world;
    auto render_all = [&](uint32_t phase) -> int {
        size_t who = 0;
        const int rc = g->each([&](size_t i) {
            rt2_launch_params lp{};
            if (params) lp = *params;
            lp.band_rows = world > 1u ? RT2_GROUP_BAND_ROWS : 0u; lp.band_rank = (uint32_t)g->first_rank + (uint32_t)i; lp.band_world = world > 1u ? world : 0u;
            if (world <= 1u) lp.band_rank = 0u;
            g->ctxs[i]->restir_phase = phase;
            const int r = rt2_render(g->ctxs[i], frame, &lp);
            g->ctxs[i]->restir_phase = 0u;
            return r;
        }, &who);
        if (rc != RT2_OK) return gfail(g, rc, std::string("rt2_render: ") + rt2_last_error(g->ctxs[who]));
        return RT2_OK;
    };
    // GI temporal / spatial resampling (K6 / K7) under sharding: the frame runs in phases with the inputs those two passes read at
    // other ranks' pixels exchanged in between (every rank's GConst is the same: rt2_group_set_constants)
    const rt2_gconst& gc = g->ctxs[0]->gconst;
    const bool temporal = gc.enable_temporal_resampling == 1u, spatial = gc.enable_spatial_resampling == 1u;
    if (world > 1u && gc.refrence_mode != 1u && gc.enable_restir_gi == 1u && (temporal || spatial)) {
        const uint32_t cur = frame & 1u;                                // the G-buffer set K3 writes this frame (src/light_passes.rs:565-577)
        auto page = [&](size_t i, uint32_t index) {
            FrameBuffers& f = g->ctxs[i]->fb;
            return ExchangeSeg{reinterpret_cast<uint32_t*>(f.gi_reservoirs + 2 * (size_t)f.reservoir_array_pitch * index), 8u, 1u, 0};
        };
        int rc = render_all(1u);                                        // K3, K9, K4, K5 on the rank's bands
        if (rc) return rc;
        if (temporal) {                                                 // K6 gathers from the previous frame's reservoirs; its G-buffer was exchanged last frame
            const uint32_t tin = gc.restir_gi.buffer_indices.temporal_resampling_input_buffer_index;
            if ((rc = group_exchange(g, [&](size_t i) { return std::vector<ExchangeSeg>{page(i, tin)}; }))) return rc;
            if ((rc = render_all(2u))) return rc;
        }
        // the current G-buffer: K7's neighbours now, K6's reprojected surfaces next frame; + the page K7 gathers from
        const uint32_t sin = gc.restir_gi.buffer_indices.spatial_resampling_input_buffer_index;
        rc = group_exchange(g, [&](size_t i) {
            FrameBuffers& f = g->ctxs[i]->fb;
            std::vector<ExchangeSeg> v;
            void* planes[5] = {f.gb_depth[cur], f.gb_normals[cur], f.gb_geo_normals[cur], f.gb_diffuse_albedo[cur], f.gb_specular_rough[cur]};
            for (void* p : planes) v.push_back(ExchangeSeg{static_cast<uint32_t*>(p), 1u, 0u, 0});
            if (spatial) v.push_back(page(i, sin));
            return v;
        });
        if (rc) return rc;
        if ((rc = render_all(4u))) return rc;                           // K7, K8
    } else {
        const int rc = render_all(0u);
        if (rc) return rc;
    }
    if (gather_which < 0) return RT2_OK;
    return rt2_group_gather(g, gather_which);
}

int rt2_group_gather(rt2_group* g, int gather_which) {
    if (!g) return RT2_E_INVALID;
    if (!g->height) return gfail(g, RT2_E_STATE, "rt2_group_gather before rt2_group_resize");
    const size_t n = g->ctxs.size();
    const uint32_t world = (uint32_t)g->world;
    if (gather_which < 0 || gather_which > RT2_BUF_OUTPUT_RGBA8) return gfail(g, RT2_E_INVALID, "rt2_group_gather: not a W x H frame buffer");
    const size_t frame_bytes = rt2_buffer_size(g->ctxs[0], gather_which);
    if (!frame_bytes) return gfail(g, RT2_E_STATE, "rt2_group_gather: buffer is not allocated");
    const size_t row_bytes = frame_bytes / g->height;
    if (row_bytes % 16u) return gfail(g, RT2_E_INVALID, "rt2_group_gather: rows of the gathered buffer must be a multiple of 16 bytes");
    BandPlanC plan; plan.height = g->height; plan.world = world;
    const uint32_t max_rows = plan.max_rows();
    const size_t slot_bytes = (size_t)max_rows * row_bytes;
    const int b = (int)(g->gathers % RT2_GROUP_BUFFERS);
    for (size_t i = 0; i < n; i++) {
        rt2_group::Lane& L = g->lanes[i];
        rt2_ctx* c = g->ctxs[i];
        RT2_GCUDA(g, cudaSetDevice(c->device));
        if (L.cap_slot < slot_bytes || L.cap_frame < frame_bytes) {       // first gather of a size: allocate (steady state allocates nothing)
            RT2_GCUDA(g, cudaStreamSynchronize(c->stream)); RT2_GCUDA(g, cudaStreamSynchronize(g->comm_streams[i]));
            for (int k = 0; k < RT2_GROUP_BUFFERS; k++) {
                if (L.send[k]) cudaFree(L.send[k]); if (L.recv[k]) cudaFree(L.recv[k]); if (L.assembled[k]) cudaFree(L.assembled[k]);
                L.send[k] = L.recv[k] = L.assembled[k] = nullptr;
                RT2_GCUDA(g, cudaMalloc(&L.send[k], slot_bytes)); RT2_GCUDA(g, cudaMalloc(&L.recv[k], slot_bytes * world)); RT2_GCUDA(g, cudaMalloc(&L.assembled[k], frame_bytes));
                RT2_GCUDA(g, cudaMemsetAsync(L.send[k], 0, slot_bytes, c->stream));
            }
            L.cap_slot = slot_bytes; L.cap_frame = frame_bytes;
        }
        // buffer b was last used two gathers ago: its gather and un-permute (and any read of it) are ordered on the communication stream
        RT2_GCUDA(g, cudaStreamWaitEvent(c->stream, L.done[b], 0));
        if (world > 1u) {
            const uint32_t rank = (uint32_t)g->first_rank + (uint32_t)i;
            RT2_GCTX(g, c, rt2_pack_bands(c, gather_which, L.send[b], (size_t)plan.rows(rank) * row_bytes, RT2_GROUP_BAND_ROWS, rank, world));
        } else {
            RT2_GCTX(g, c, rt2_copy_to_device(c, gather_which, L.assembled[b], frame_bytes, 0, 0));
        }
        RT2_GCUDA(g, cudaEventRecord(L.ready[b], c->stream));
        RT2_GCUDA(g, cudaStreamWaitEvent(g->comm_streams[i], L.ready[b], 0));
    }
    if (world > 1u) {
        RT2_GNCCL(g, g_nccl.GroupStart());
        for (size_t i = 0; i < n; i++) {
            int r = g_nccl.AllGather(g->lanes[i].send[b], g->lanes[i].recv[b], slot_bytes, NCCL_INT8, g->comms[i], g->comm_streams[i]);
            if (r != 0) { g_nccl.GroupEnd(); return gfail(g, RT2_E_CUDA, std::string("ncclAllGather: ") + g_nccl.GetErrorString(r)); }
        }
        RT2_GNCCL(g, g_nccl.GroupEnd());
    }
    for (size_t i = 0; i < n; i++) {
        rt2_group::Lane& L = g->lanes[i];
        RT2_GCUDA(g, cudaSetDevice(g->ctxs[i]->device));
        if (world > 1u) {
            const uint32_t row_vec = (uint32_t)(row_bytes / 16u);
            const uint64_t vecs = (uint64_t)g->height * row_vec;
            const int blocks = (int)std::min<uint64_t>((uint64_t)g->ctxs[i]->num_sms * 8u, (vecs + 255u) / 256u);
            k_unpermute_bands<<<std::max(blocks, 1), 256, 0, g->comm_streams[i]>>>(static_cast<const uint4*>(L.recv[b]), static_cast<uint4*>(L.assembled[b]), g->height, row_vec,
                                                                                 RT2_GROUP_BAND_ROWS, world, max_rows);
            RT2_GCUDA(g, cudaGetLastError());
        }
        RT2_GCUDA(g, cudaEventRecord(L.done[b], g->comm_streams[i]));
    }
    g->gathers++;
    g->frame_bytes = frame_bytes;
    return RT2_OK;
}

void* rt2_group_assembled(rt2_group* g, int local_index, size_t* nbytes) {
    if (!g || local_index < 0 || (size_t)local_index >= g->lanes.size() || g->gathers == 0) return nullptr;
    if (nbytes) *nbytes = g->frame_bytes;
    return g->lanes[(size_t)local_index].assembled[(g->gathers - 1) % RT2_GROUP_BUFFERS];
}

static int group_read(rt2_group* g, void* dst, size_t nbytes, bool block) {
    if (!g) return RT2_E_INVALID;
    if (!dst) return gfail(g, RT2_E_INVALID, "dst is NULL");
    if (g->gathers == 0) return gfail(g, RT2_E_STATE, "no gathered frame: call rt2_group_render with gather_which >= 0 first");
    if (nbytes != g->frame_bytes) return gfail(g, RT2_E_INVALID, "size mismatch: the assembled frame is " + std::to_string(g->frame_bytes) + " bytes");
    RT2_GCUDA(g, cudaSetDevice(g->ctxs[0]->device));
    // on the communication stream, behind the un-permute that produced it and ahead of the next gather into the same buffer
    RT2_GCUDA(g, cudaMemcpyAsync(dst, g->lanes[0].assembled[(g->gathers - 1) % RT2_GROUP_BUFFERS], nbytes, cudaMemcpyDeviceToHost, g->comm_streams[0]));
    if (block) RT2_GCUDA(g, cudaStreamSynchronize(g->comm_streams[0]));
    return RT2_OK;
}
int rt2_group_read(rt2_group* g, void* dst, size_t nbytes) { return group_read(g, dst, nbytes, true); }
int rt2_group_read_async(rt2_group* g, void* dst, size_t nbytes) { return group_read(g, dst, nbytes, false); }

int rt2_group_wait_gather(rt2_group* g) {
    if (!g) return RT2_E_INVALID;
    for (size_t i = 0; i < g->ctxs.size(); i++) {
        RT2_GCUDA(g, cudaSetDevice(g->ctxs[i]->device));
        RT2_GCUDA(g, cudaEventRecord(g->lanes[i].tail, g->comm_streams[i]));
        RT2_GCUDA(g, cudaStreamWaitEvent(g->ctxs[i]->stream, g->lanes[i].tail, 0));
    }
    return RT2_OK;
}

}  // extern "C"
