// capi.cu -- the C ABI of include/svo_gpu.h: context, device buffers, and the host side of every entry point.
// No CPU implementation of any stage exists in this library: a call either enqueues the sm_100a kernels or fails.
#include <limits.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <new>
#include <vector>

#include "svo_internal.cuh"

using namespace svo;

namespace {

struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
};

struct Slot {
    DevBuf raw_own;                 // context-owned copy of host uploads
    const uint8_t* raw = nullptr;   // raw_own.p, or caller-owned device memory (mem_kind DEVICE: aliased, not copied)
    size_t raw_pitch = 0, raw_lane = 0;
    int w = 0, h = 0, batch = 0;
    bool uploaded = false;
    unsigned long long gen = 0;     // counts the images this slot has received (any upload path, the step's ingest included)
    DevBuf pyr;
    PyrGeom g{};
    bool pyr_built = false;
    int win_w = 0, win_h = 0;
};

// one submitted step (svo_gpu_step_submit .. svo_gpu_step_wait): its pinned read-back scratch, events and output views
struct StepTicket {
    bool busy = false;
    int* hs = nullptr;            // pinned, 6 * max_batch + 16 ints: tracked, keypoints, bucketed, bucket error flags, key-frame flags, survivors, LK iterations
    cudaEvent_t done = nullptr;
    cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t evk = nullptr;    // main stream, in front of LK (the pyramid may have been built on another stream, earlier)
    svo_gpu_step_outputs out{};
    bool has_out = false, prof = false, track = false, overlap = false;
    int batch = 0, kp_cap = 0;
};

// the captured chain of one step (capi_step.inc), valid for exactly the pointers and settings it was recorded with
struct StepGraph {
    cudaGraphExec_t exec = nullptr;
    svo_gpu_step_params P{};
    svo_gpu_step_outputs out{};
    bool has_out = false;
    int batch = 0, kernels = 0, eff = 0, kp_cap = 0;
    unsigned long long alloc_epoch = 0;   // ctx->alloc_epoch at capture: a graph holds device addresses by value
    const void* pyr_prev = nullptr;
    const void* pyr_cur = nullptr;
};

constexpr int kStepBase = SVO_GPU_NUM_SLOTS - 3;   // svo_gpu_step's ring of three slots
// No event orders the next frame's copy / pyramid build after the kernels of steps still in flight: the ring alone makes it
// safe, because the free slot's last reader belongs to a step that has been waited for.  That needs more slots than steps.
static_assert(SVO_GPU_STEP_DEPTH + 1 <= 3, "the step's slot ring must be larger than the number of outstanding steps");

thread_local char g_create_error[512] = "";

}  // namespace

struct svo_gpu_ctx {
    int device = 0;
    cudaStream_t st = nullptr;
    int max_w = 0, max_h = 0, max_levels = 0, max_points = 0, max_batch = 0;
    Slot slots[SVO_GPU_NUM_SLOTS];
    char err[512] = "";
    long long launches = 0;

    DevBuf bgr_stage;             // device copy of host BGR frames on their way to a slot (svo_gpu_upload_bgr*)

    // FAST
    FastBufs fb{};
    DevBuf mask;                  // [B][max_h * mask_pitch]
    size_t mask_pitch = 0, mask_lane = 0;
    // bucketing
    BucketBufs bb{};
    DevBuf cell_slots, cell_cnt, cell_last;
    float2* bk_in_xy = nullptr;   // [B][max_points] staging for the host-pointer bucketing call
    int* bk_in_age = nullptr;
    // tracks / LK
    float2* trk_xy = nullptr;     // [B][max_points] points LK starts from
    float2* nxt_xy = nullptr;
    uint8_t* trk_status = nullptr;
    float* trk_err = nullptr;
    int* trk_n = nullptr;         // [B]
    int* step_summary = nullptr;  // [6][B] tracked, keypoints, bucketed, bucket error flags, key-frame flags, survivors of the step (one
                                  // read-back) + [B] arrival counters of next_tracks_kernel (zero between launches)
    int* lane_kf = nullptr;       // [B] carry_tracks mode: the lane's frame being processed is a key-frame
    int* int_max = nullptr;       // [B] all INT_MAX: "n is given by n_max" for the host-pointer calls
    unsigned long long* d_iters = nullptr;
    long long last_iters = 0;
    // status / epipolar filter outputs (filter.cu)
    int* flt_index = nullptr;     // [B][max_points]
    float2* flt_a = nullptr;
    float2* flt_b = nullptr;
    float* flt_err = nullptr;
    int* flt_n = nullptr;         // [B]
    // pinned scratch for small read-backs
    int* h_small = nullptr;       // 6 * max_batch + 16 ints
    int fetch_hint = 0;           // keypoints svo_gpu_fast last returned (+ 25 %): copied back together with the count
    int bucket_fetch_hint = 0;    // same for svo_gpu_bucket
    uint8_t* h_klt = nullptr;     // pinned staging of svo_gpu_klt's results (13 bytes per point + the iteration count), allocated by its first call
    uint8_t* h_bucket = nullptr;  // pinned staging of svo_gpu_bucket's speculative read-back (bucket_stage_cap entries)
    int bucket_stage_cap = 0;

    // svo_gpu_step state
    bool step_has_prev = false;
    unsigned long long alloc_epoch = 0;   // bumped whenever ensure() replaces a device buffer
    int step_cur = kStepBase;
    svo_gpu_step_params step_params{};
    int step_batch = 0;
    bool profile = false;
    StepTicket tickets[SVO_GPU_STEP_DEPTH];
    int next_ticket = 0;
    int kp_hint = 0;              // keypoints per lane seen by the last completed step (+25 %), 0 = unknown
    cudaEvent_t ev_raw_free[3] = {nullptr, nullptr, nullptr};   // per step slot: the ingest that last read its raw frame is through
    // frame ingest pipelining: the next step's host frames are copied on a second stream while this step computes
    cudaStream_t copy_st = nullptr;
    cudaEvent_t ev_copy = nullptr;
    cudaStream_t out_st = nullptr;     // LK results go back to the host on this stream while FAST and bucketing still run
    cudaEvent_t ev_klt = nullptr, ev_out = nullptr;
    // detection beside tracking: FAST and bucketing run on aux_st while the LK kernel runs on st (small batches, see capi_step.inc)
    cudaStream_t aux_st = nullptr;
    // the step builds the new frame's pyramid on pyr_st (highest priority): with a step still in flight it runs beside that
    // step's kernels without waiting for any of them -- it writes the ring's free slot, which no step in flight reads
    // (three slots, at most SVO_GPU_STEP_DEPTH = 2 steps outstanding; see the static_assert at kStepBase)
    cudaStream_t pyr_st = nullptr;
    cudaEvent_t ev_pyr = nullptr;
    int pyr_async = 1;      // SVO_GPU_STEP_PYR_ASYNC (0 = pyramid on the main stream)
    int pyr_fused = -1;     // SVO_GPU_PYR_FUSED: -1 = per call (fused for small frame sets), 0 = ingest + one pyrDown launch per level, 1 = fused
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    int step_graph = 1;     // SVO_GPU_STEP_GRAPH: 1 = the step's chain is captured and replayed whenever possible (default), 0 = never, -1 = small steps only
    int steps_since_reset = 0;
    StepGraph graphs[3][SVO_GPU_STEP_DEPTH];
    bool graphs_precaptured = false;
    bool out_checked[SVO_GPU_STEP_DEPTH] = {}, out_checked_ok[SVO_GPU_STEP_DEPTH] = {};
    svo_gpu_step_outputs out_checked_key[SVO_GPU_STEP_DEPTH] = {};
    int overlap = 1;        // SVO_GPU_STEP_OVERLAP: 1 = detection beside LK (default: measured faster at every stream count), 0 = back to back, -1 = only when LK does not fill the device
    int fast_small = -1;    // SVO_GPU_FAST_SMALL: FAST CTA size in the step: -1 = by tile count (128 threads, 512 for few tiles), 1 = 128, 0 = 256, 2 = 512
    const uint8_t* pf_ptr = nullptr;   // host pointer of the prefetched frames (nullptr = none pending)
    int pf_slot = -1;
    float stage_ms[4] = {0, 0, 0, 0};
};

namespace {

int fail(svo_gpu_ctx* c, int code, const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(c ? c->err : g_create_error, 512, fmt, ap);
    va_end(ap);
    return code;
}

#define CK(call)                                                                                              \
    do {                                                                                                      \
        cudaError_t e_ = (call);                                                                              \
        if (e_ != cudaSuccess)                                                                                \
            return fail(ctx, SVO_GPU_E_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, \
                        __LINE__);                                                                            \
    } while (0)

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev)
    {
        cudaGetDevice(&prev);
        if (prev != dev) cudaSetDevice(dev);
        else prev = -1;
    }
    ~DeviceGuard()
    {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// grow-only device buffer; growing synchronises the stream first so that no kernel still uses the old storage
int ensure(svo_gpu_ctx* ctx, DevBuf& b, size_t bytes)
{
    if (bytes <= b.bytes) return 0;
    CK(cudaStreamSynchronize(ctx->st));
    ++ctx->alloc_epoch;   // captured step graphs hold the old address (bucket tables, mask ...): they must be recorded again
    if (b.p) CK(cudaFree(b.p));
    b.p = nullptr;
    b.bytes = 0;
    CK(cudaMalloc(&b.p, bytes));
    b.bytes = bytes;
    return 0;
}

template <typename T>
int dmalloc(svo_gpu_ctx* ctx, T*& p, size_t count)
{
    CK(cudaMalloc(reinterpret_cast<void**>(&p), sizeof(T) * (count ? count : 1)));
    return 0;
}

// effective maxLevel of cv::buildOpticalFlowPyramid: stop when the next level would be <= winSize in a dimension
int effective_levels(int w, int h, int win_w, int win_h, int max_level)
{
    for (int level = 0; level <= max_level; ++level) {
        w = (w + 1) / 2;
        h = (h + 1) / 2;
        if (w <= win_w || h <= win_h) return level;
    }
    return max_level;
}

int margin_for(int win_w, int win_h) { return (int)round_up((size_t)std::max(win_w, win_h) + 9, 16); }

void make_geometry(PyrGeom& g, int w, int h, int nlevels, int P)
{
    g.nlevels = nlevels;
    g.P = P;
    long long off = 0;
    for (int l = 0; l < nlevels; ++l) {
        LevelGeom& L = g.lv[l];
        L.w = w;
        L.h = h;
        L.pitch = (int)round_up((size_t)w + 2 * P, 128);
        L.pad_ = 0;
        L.off = off + (long long)P * L.pitch + P;
        off += (long long)L.pitch * (h + 2 * P);
        off = (long long)round_up((size_t)off, 256);
        w = (w + 1) / 2;
        h = (h + 1) / 2;
    }
    g.lane_bytes = off;
}

int check_slot(svo_gpu_ctx* ctx, int slot)
{
    if (slot < 0 || slot >= SVO_GPU_NUM_SLOTS) return fail(ctx, SVO_GPU_E_ARG, "slot %d out of range", slot);
    return 0;
}

}  // namespace

// ================================================================================================ context
extern "C" int svo_gpu_abi_version(void) { return 1; }

extern "C" const char* svo_gpu_last_error(const svo_gpu_ctx* ctx) { return ctx ? ctx->err : g_create_error; }

extern "C" void* svo_gpu_stream(svo_gpu_ctx* ctx) { return ctx ? (void*)ctx->st : nullptr; }

extern "C" long long svo_gpu_launch_count(const svo_gpu_ctx* ctx) { return ctx ? ctx->launches : 0; }

extern "C" long long svo_gpu_last_klt_iterations(svo_gpu_ctx* ctx) { return ctx ? ctx->last_iters : 0; }

static int create_buffers(svo_gpu_ctx* ctx)
{
    const size_t B = (size_t)ctx->max_batch, NP = (size_t)ctx->max_points;
    // FAST segments
    FastBufs& fb = ctx->fb;
    fb.rows_cap = ctx->max_h;
    fb.ntx_cap = (ctx->max_w + kFastTW - 1) / kFastTW;
    fb.kp_cap = ctx->max_points;
    if (int r = dmalloc(ctx, fb.rowtotal, B * fb.rows_cap)) return r;
    if (int r = dmalloc(ctx, fb.segcnt, B * fb.rows_cap * fb.ntx_cap)) return r;
    if (int r = dmalloc(ctx, fb.segkp, B * fb.rows_cap * fb.ntx_cap * kFastTW)) return r;
    if (int r = dmalloc(ctx, fb.kp_xy, B * NP)) return r;
    if (int r = dmalloc(ctx, fb.kp_resp, B * NP)) return r;
    if (int r = dmalloc(ctx, fb.kp_count, B)) return r;
    // bucketing
    BucketBufs& bb = ctx->bb;
    bb.out_cap = ctx->max_points;
    if (int r = dmalloc(ctx, bb.out_xy, B * NP)) return r;
    if (int r = dmalloc(ctx, bb.out_age, B * NP)) return r;
    if (int r = dmalloc(ctx, bb.out_count, B)) return r;
    if (int r = dmalloc(ctx, bb.err_flag, B)) return r;
    if (int r = dmalloc(ctx, ctx->bk_in_xy, B * NP)) return r;
    if (int r = dmalloc(ctx, ctx->bk_in_age, B * NP)) return r;
    // tracks
    if (int r = dmalloc(ctx, ctx->trk_xy, B * NP)) return r;
    if (int r = dmalloc(ctx, ctx->nxt_xy, B * NP)) return r;
    if (int r = dmalloc(ctx, ctx->trk_status, B * NP)) return r;
    if (int r = dmalloc(ctx, ctx->trk_err, B * NP)) return r;
    if (int r = dmalloc(ctx, ctx->trk_n, B)) return r;
    if (int r = dmalloc(ctx, ctx->int_max, B)) return r;
    if (int r = dmalloc(ctx, ctx->step_summary, 7 * B)) return r;
    if (int r = dmalloc(ctx, ctx->lane_kf, B)) return r;
    if (int r = dmalloc(ctx, ctx->d_iters, 2)) return r;   // [0] LK iterations, [1] LK work counter
    if (int r = dmalloc(ctx, ctx->flt_index, B * NP)) return r;
    if (int r = dmalloc(ctx, ctx->flt_a, B * NP)) return r;
    if (int r = dmalloc(ctx, ctx->flt_b, B * NP)) return r;
    if (int r = dmalloc(ctx, ctx->flt_err, B * NP)) return r;
    if (int r = dmalloc(ctx, ctx->flt_n, B)) return r;
    CK(cudaMemsetAsync(ctx->trk_n, 0, sizeof(int) * B, ctx->st));
    CK(cudaMemsetAsync(ctx->flt_n, 0, sizeof(int) * B, ctx->st));
    CK(cudaMemsetAsync(ctx->step_summary, 0, sizeof(int) * 7 * B, ctx->st));
    CK(cudaMemsetAsync(ctx->d_iters, 0, sizeof(unsigned long long), ctx->st));
    {
        std::vector<int> big(B, INT_MAX);
        CK(cudaMemcpy(ctx->int_max, big.data(), sizeof(int) * B, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(ctx->lane_kf, big.data(), sizeof(int) * B, cudaMemcpyHostToDevice));   // non-zero: first frames are key-frames
    }
    CK(cudaMallocHost(reinterpret_cast<void**>(&ctx->h_small), sizeof(int) * (6 * B + 16)));
    ctx->bucket_stage_cap = (int)std::min<size_t>(NP, 16384);
    CK(cudaMallocHost(reinterpret_cast<void**>(&ctx->h_bucket), (sizeof(float2) + sizeof(int)) * (size_t)ctx->bucket_stage_cap));
    for (auto& tk : ctx->tickets) {
        CK(cudaMallocHost(reinterpret_cast<void**>(&tk.hs), sizeof(int) * (6 * B + 16)));
        CK(cudaEventCreateWithFlags(&tk.done, cudaEventDisableTiming));
        for (auto& e : tk.ev) CK(cudaEventCreate(&e));
        CK(cudaEventCreate(&tk.evk));
    }
    for (auto& e : ctx->ev_raw_free) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    CK(cudaStreamCreateWithFlags(&ctx->copy_st, cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&ctx->ev_copy, cudaEventDisableTiming));
    CK(cudaStreamCreateWithFlags(&ctx->out_st, cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&ctx->ev_klt, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&ctx->ev_out, cudaEventDisableTiming));
    {   // lowest priority: detection CTAs only take what the LK kernel (on st, default priority) leaves free
        int lo = 0, hi = 0;
        CK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        CK(cudaStreamCreateWithPriority(&ctx->aux_st, cudaStreamNonBlocking, lo));
    }
    CK(cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming));
    {
        int lo = 0, hi = 0;
        CK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        CK(cudaStreamCreateWithPriority(&ctx->pyr_st, cudaStreamNonBlocking, hi));
        CK(cudaEventCreateWithFlags(&ctx->ev_pyr, cudaEventDisableTiming));
    }
    if (const char* v = getenv("SVO_GPU_STEP_PYR_ASYNC")) ctx->pyr_async = atoi(v);
    if (const char* v = getenv("SVO_GPU_PYR_FUSED")) ctx->pyr_fused = atoi(v);
    if (const char* v = getenv("SVO_GPU_BUCKET_MODE")) ctx->bb.mode = atoi(v);   // 1: global-memory kernels, 2: one CTA per lane
    if (const char* v = getenv("SVO_GPU_STEP_OVERLAP")) ctx->overlap = atoi(v);
    if (const char* v = getenv("SVO_GPU_STEP_GRAPH")) ctx->step_graph = atoi(v);
    if (const char* v = getenv("SVO_GPU_FAST_SMALL")) ctx->fast_small = atoi(v);
    CK(cudaStreamSynchronize(ctx->st));
    return 0;
}

extern "C" int svo_gpu_create(svo_gpu_ctx** out, int device, int max_w, int max_h, int max_levels, int max_points,
                              int max_batch)
{
    svo_gpu_ctx* ctx = nullptr;   // CK/fail write to g_create_error while ctx is null
    if (!out) return fail(ctx, SVO_GPU_E_ARG, "out is null");
    *out = nullptr;
    if (max_w < 1 || max_h < 1 || max_w > 65536 || max_h > 65536 || max_levels < 1 || max_levels > SVO_GPU_MAX_LEVELS ||
        max_points < 1 || max_batch < 1 || max_batch > 65535)
        return fail(ctx, SVO_GPU_E_ARG, "svo_gpu_create: argument out of range");
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail(ctx, SVO_GPU_E_ARG, "device %d not present (%d devices)", device, ndev);
    DeviceGuard guard(device);
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10)
        return fail(ctx, SVO_GPU_E_UNSUPPORTED, "device %d is sm_%d%d; this library carries sm_100a code only", device,
                    prop.major, prop.minor);
    svo_gpu_ctx* c = new (std::nothrow) svo_gpu_ctx();
    if (!c) return fail(ctx, SVO_GPU_E_ARG, "out of host memory");
    c->device = device;
    c->max_w = max_w; c->max_h = max_h; c->max_levels = max_levels; c->max_points = max_points; c->max_batch = max_batch;
    cudaError_t e = cudaStreamCreateWithFlags(&c->st, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        delete c;
        return fail(ctx, SVO_GPU_E_CUDA, "cudaStreamCreate failed: %s", cudaGetErrorString(e));
    }
    int r = create_buffers(c);
    if (r) {
        snprintf(g_create_error, sizeof g_create_error, "%s", c->err);
        svo_gpu_destroy(c);
        return r;
    }
    *out = c;
    return 0;
}

extern "C" int svo_gpu_destroy(svo_gpu_ctx* ctx)
{
    if (!ctx) return 0;
    DeviceGuard guard(ctx->device);
    if (ctx->st) cudaStreamSynchronize(ctx->st);
    if (ctx->copy_st) { cudaStreamSynchronize(ctx->copy_st); cudaStreamDestroy(ctx->copy_st); }
    if (ctx->ev_copy) cudaEventDestroy(ctx->ev_copy);
    if (ctx->out_st) { cudaStreamSynchronize(ctx->out_st); cudaStreamDestroy(ctx->out_st); }
    if (ctx->ev_klt) cudaEventDestroy(ctx->ev_klt);
    for (auto& row : ctx->graphs)
        for (auto& g : row)
            if (g.exec) cudaGraphExecDestroy(g.exec);
    if (ctx->ev_out) cudaEventDestroy(ctx->ev_out);
    if (ctx->aux_st) { cudaStreamSynchronize(ctx->aux_st); cudaStreamDestroy(ctx->aux_st); }
    if (ctx->pyr_st) { cudaStreamSynchronize(ctx->pyr_st); cudaStreamDestroy(ctx->pyr_st); }
    if (ctx->ev_pyr) cudaEventDestroy(ctx->ev_pyr);
    if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
    if (ctx->ev_join) cudaEventDestroy(ctx->ev_join);
    for (auto& tk : ctx->tickets) {
        if (tk.hs) cudaFreeHost(tk.hs);
        if (tk.done) cudaEventDestroy(tk.done);
        for (auto& e : tk.ev) if (e) cudaEventDestroy(e);
        if (tk.evk) cudaEventDestroy(tk.evk);
    }
    for (auto& e : ctx->ev_raw_free) if (e) cudaEventDestroy(e);
    for (auto& s : ctx->slots) {
        cudaFree(s.raw_own.p);
        cudaFree(s.pyr.p);
    }
    void* bufs[] = {ctx->fb.rowtotal, ctx->fb.segcnt, ctx->fb.segkp, ctx->fb.kp_xy, ctx->fb.kp_resp, ctx->fb.kp_count,
                    ctx->mask.p, ctx->bgr_stage.p, ctx->bb.out_xy, ctx->bb.out_age, ctx->bb.out_count, ctx->bb.err_flag, ctx->cell_slots.p,
                    ctx->cell_cnt.p, ctx->cell_last.p, ctx->bk_in_xy, ctx->bk_in_age, ctx->trk_xy, ctx->nxt_xy,
                    ctx->trk_status, ctx->trk_err, ctx->trk_n, ctx->int_max, ctx->step_summary, ctx->lane_kf, ctx->d_iters, ctx->flt_index, ctx->flt_a,
                    ctx->flt_b, ctx->flt_err, ctx->flt_n};
    for (void* p : bufs) cudaFree(p);
    if (ctx->h_small) cudaFreeHost(ctx->h_small);
    if (ctx->h_bucket) cudaFreeHost(ctx->h_bucket);
    if (ctx->h_klt) cudaFreeHost(ctx->h_klt);
    if (ctx->st) cudaStreamDestroy(ctx->st);
    delete ctx;
    return 0;
}

extern "C" int svo_gpu_host_alloc(void** out, size_t bytes)
{
    svo_gpu_ctx* ctx = nullptr;   // failures are reported like those of svo_gpu_create
    if (!out || bytes == 0) return fail(ctx, SVO_GPU_E_ARG, "svo_gpu_host_alloc: null pointer or zero size");
    *out = nullptr;
    CK(cudaMallocHost(out, bytes));
    return 0;
}

extern "C" int svo_gpu_host_free(void* p)
{
    svo_gpu_ctx* ctx = nullptr;
    if (p) CK(cudaFreeHost(p));
    return 0;
}

// ================================================================================================ images
// copies `batch` images into the slot on stream `cs` (enqueue only).  Host images that are tightly packed (stride == w,
// images back to back) go over PCIe as ONE contiguous copy and stay compact on the device (row pitch w: the ingest kernel
// reads any alignment); otherwise each image is a 2-D copy into 128 B-pitched rows.
static int upload_to_slot(svo_gpu_ctx* ctx, int slot, const uint8_t* images, int mem_kind, int w, int h, size_t stride_bytes,
                          size_t image_stride_bytes, int batch, cudaStream_t cs)
{
    if (int r = check_slot(ctx, slot)) return r;
    if (!images) return fail(ctx, SVO_GPU_E_ARG, "images is null");
    if (w < 1 || h < 1 || w > ctx->max_w || h > ctx->max_h)
        return fail(ctx, SVO_GPU_E_ARG, "image %dx%d outside the context's 1..%dx1..%d", w, h, ctx->max_w, ctx->max_h);
    if (batch < 1 || batch > ctx->max_batch) return fail(ctx, SVO_GPU_E_ARG, "batch %d outside 1..%d", batch, ctx->max_batch);
    if (stride_bytes < (size_t)w) return fail(ctx, SVO_GPU_E_ARG, "stride %zu < width %d", stride_bytes, w);
    if (batch > 1 && image_stride_bytes < stride_bytes * (size_t)h)
        return fail(ctx, SVO_GPU_E_ARG, "image stride smaller than one image");
    Slot& s = ctx->slots[slot];
    if (mem_kind == SVO_GPU_MEM_DEVICE) {
        s.raw = images;   // aliased: the caller keeps it alive until the consuming calls have returned
        s.raw_pitch = stride_bytes;
        s.raw_lane = image_stride_bytes;
    } else if (mem_kind == SVO_GPU_MEM_HOST) {
        const size_t pitched = round_up((size_t)ctx->max_w, 128) * (size_t)ctx->max_h;
        if (int r = ensure(ctx, s.raw_own, pitched * (size_t)ctx->max_batch + 64)) return r;
        uint8_t* d = static_cast<uint8_t*>(s.raw_own.p);
        const bool compact = stride_bytes == (size_t)w && (batch == 1 || image_stride_bytes == (size_t)w * h);
        if (compact) {
            CK(cudaMemcpyAsync(d, images, (size_t)w * h * batch, cudaMemcpyHostToDevice, cs));
            s.raw_pitch = (size_t)w;
            s.raw_lane = (size_t)w * h;
        } else {
            const size_t pitch = round_up((size_t)w, 128), lane = pitch * (size_t)h;
            for (int b = 0; b < batch; ++b)
                CK(cudaMemcpy2DAsync(d + b * lane, pitch, images + b * image_stride_bytes, stride_bytes, (size_t)w,
                                     (size_t)h, cudaMemcpyHostToDevice, cs));
            s.raw_pitch = pitch;
            s.raw_lane = lane;
        }
        s.raw = d;
    } else {
        return fail(ctx, SVO_GPU_E_ARG, "mem_kind %d unknown", mem_kind);
    }
    s.w = w; s.h = h; s.batch = batch;
    s.uploaded = true;
    ++s.gen;
    s.pyr_built = false;
    return 0;
}

extern "C" int svo_gpu_upload_batch(svo_gpu_ctx* ctx, int slot, const uint8_t* images, int mem_kind, int w, int h,
                                    size_t stride_bytes, size_t image_stride_bytes, int batch)
{
    if (!ctx) return SVO_GPU_E_ARG;
    DeviceGuard guard(ctx->device);
    return upload_to_slot(ctx, slot, images, mem_kind, w, h, stride_bytes, image_stride_bytes, batch, ctx->st);
}

extern "C" int svo_gpu_upload(svo_gpu_ctx* ctx, int slot, const uint8_t* host, int w, int h, size_t stride_bytes)
{
    return svo_gpu_upload_batch(ctx, slot, host, SVO_GPU_MEM_HOST, w, h, stride_bytes, stride_bytes * (size_t)(h > 0 ? h : 0), 1);
}

// Colour frames: interleaved BGR (cv::Mat CV_8UC3 as cv::imread returns it) -> gray in the slot, converted on the device
// with OpenCV's fixed-point COLOR_BGR2GRAY.  Replaces the host-side cv::cvtColor of include/async_image_loader.h:68-69.
extern "C" int svo_gpu_upload_bgr_batch(svo_gpu_ctx* ctx, int slot, const uint8_t* images, int mem_kind, int w, int h,
                                        size_t stride_bytes, size_t image_stride_bytes, int batch)
{
    if (!ctx) return SVO_GPU_E_ARG;
    if (int r = check_slot(ctx, slot)) return r;
    if (!images) return fail(ctx, SVO_GPU_E_ARG, "images is null");
    if (w < 1 || h < 1 || w > ctx->max_w || h > ctx->max_h)
        return fail(ctx, SVO_GPU_E_ARG, "image %dx%d outside the context's 1..%dx1..%d", w, h, ctx->max_w, ctx->max_h);
    if (batch < 1 || batch > ctx->max_batch) return fail(ctx, SVO_GPU_E_ARG, "batch %d outside 1..%d", batch, ctx->max_batch);
    if (stride_bytes < 3 * (size_t)w) return fail(ctx, SVO_GPU_E_ARG, "stride %zu < 3 * width %d", stride_bytes, w);
    if (batch > 1 && image_stride_bytes < stride_bytes * (size_t)h)
        return fail(ctx, SVO_GPU_E_ARG, "image stride smaller than one image");
    if (mem_kind != SVO_GPU_MEM_HOST && mem_kind != SVO_GPU_MEM_DEVICE) return fail(ctx, SVO_GPU_E_ARG, "mem_kind %d unknown", mem_kind);
    DeviceGuard guard(ctx->device);
    Slot& s = ctx->slots[slot];
    const size_t pitch = round_up((size_t)w, 128), lane = pitch * (size_t)h;
    if (int r = ensure(ctx, s.raw_own, round_up((size_t)ctx->max_w, 128) * (size_t)ctx->max_h * (size_t)ctx->max_batch + 64)) return r;
    const uint8_t* src = images;
    size_t src_pitch = stride_bytes, src_lane = image_stride_bytes;
    if (mem_kind == SVO_GPU_MEM_HOST) {
        const size_t row = 3 * (size_t)w;
        if (int r = ensure(ctx, ctx->bgr_stage, 3 * (size_t)ctx->max_w * (size_t)ctx->max_h * (size_t)ctx->max_batch + 64)) return r;
        uint8_t* d = static_cast<uint8_t*>(ctx->bgr_stage.p);
        if (stride_bytes == row && (batch == 1 || image_stride_bytes == row * h))
            CK(cudaMemcpyAsync(d, images, row * h * batch, cudaMemcpyHostToDevice, ctx->st));   // one contiguous copy
        else
            for (int b = 0; b < batch; ++b)
                CK(cudaMemcpy2DAsync(d + b * row * h, row, images + b * image_stride_bytes, stride_bytes, row, (size_t)h,
                                     cudaMemcpyHostToDevice, ctx->st));
        src = d; src_pitch = row; src_lane = row * h;
    }
    launch_bgr2gray(ctx->st, src, src_pitch, src_lane, static_cast<uint8_t*>(s.raw_own.p), pitch, lane, w, h, batch, &ctx->launches);
    CK(cudaGetLastError());
    if (mem_kind == SVO_GPU_MEM_DEVICE) CK(cudaStreamSynchronize(ctx->st));   // the caller's buffer is free again on return
    s.raw = static_cast<const uint8_t*>(s.raw_own.p);
    s.raw_pitch = pitch; s.raw_lane = lane;
    s.w = w; s.h = h; s.batch = batch;
    s.uploaded = true;
    ++s.gen;
    s.pyr_built = false;
    return 0;
}

extern "C" int svo_gpu_upload_bgr(svo_gpu_ctx* ctx, int slot, const uint8_t* host, int w, int h, size_t stride_bytes)
{
    return svo_gpu_upload_bgr_batch(ctx, slot, host, SVO_GPU_MEM_HOST, w, h, stride_bytes, stride_bytes * (size_t)(h > 0 ? h : 0), 1);
}

// the 8-bit gray image a slot holds (what the reference keeps as frame->leftImg() for its viewer and descriptors)
extern "C" int svo_gpu_download_image(svo_gpu_ctx* ctx, int slot, int lane, uint8_t* host, size_t stride_bytes)
{
    if (!ctx) return SVO_GPU_E_ARG;
    if (int r = check_slot(ctx, slot)) return r;
    Slot& s = ctx->slots[slot];
    if (!s.uploaded) return fail(ctx, SVO_GPU_E_STATE, "slot %d has no image", slot);
    if (lane < 0 || lane >= s.batch) return fail(ctx, SVO_GPU_E_ARG, "lane %d out of range", lane);
    if (!host || stride_bytes < (size_t)s.w) return fail(ctx, SVO_GPU_E_ARG, "host is null or stride %zu < width %d", stride_bytes, s.w);
    DeviceGuard guard(ctx->device);
    CK(cudaMemcpy2DAsync(host, stride_bytes, s.raw + (size_t)lane * s.raw_lane, s.raw_pitch, (size_t)s.w, (size_t)s.h,
                         cudaMemcpyDeviceToHost, ctx->st));
    CK(cudaStreamSynchronize(ctx->st));
    return 0;
}

// enqueue only (no synchronisation): ingest + pyrDown chain of every lane of the slot
static int enqueue_pyramid(svo_gpu_ctx* ctx, Slot& s, int win_w, int win_h, int max_level, int* levels_out,
                           cudaStream_t ps = nullptr)
{
    if (!ps) ps = ctx->st;
    if (!s.uploaded) return fail(ctx, SVO_GPU_E_STATE, "slot has no image");
    if (win_w <= 2 || win_h <= 2 || win_w > SVO_GPU_MAX_WIN || win_h > SVO_GPU_MAX_WIN)
        return fail(ctx, win_w <= 2 || win_h <= 2 ? SVO_GPU_E_ARG : SVO_GPU_E_UNSUPPORTED, "window %dx%d outside 3..%d",
                    win_w, win_h, SVO_GPU_MAX_WIN);
    if (max_level < 0) return fail(ctx, SVO_GPU_E_ARG, "maxLevel %d < 0", max_level);
    if (max_level > ctx->max_levels - 1) max_level = ctx->max_levels - 1;
    const int eff = effective_levels(s.w, s.h, win_w, win_h, max_level);
    make_geometry(s.g, s.w, s.h, eff + 1, margin_for(win_w, win_h));
    if (int r = ensure(ctx, s.pyr, (size_t)s.g.lane_bytes * (size_t)ctx->max_batch)) return r;
    uint8_t* pyr = static_cast<uint8_t*>(s.pyr.p);
    // One fused launch for levels 0..3 when the frame set is small (its CTAs fit the device about twice over: what counts
    // there is the chain of dependent launches), the streaming per-level kernels otherwise (measured: 1 lane 27.6 -> 21.5 us;
    // 4 lanes the step gets 6 % slower; 64 lanes 0.069 -> 0.117 ms: the fused CTAs spend their time between barriers).
    bool fused = ctx->pyr_fused > 0;
    if (ctx->pyr_fused < 0) {
        int sms = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
        fused = (long long)((s.w + 63) / 64) * ((s.h + 63) / 64) * s.batch <= 2LL * sms;
    }
    int built = fused ? launch_pyramid_fused(ps, s.raw, s.raw_pitch, s.raw_lane, pyr, s.g, s.batch, &ctx->launches) : 0;
    if (built == 0) {
        launch_ingest(ps, s.raw, s.raw_pitch, s.raw_lane, pyr, s.g, s.batch, &ctx->launches);
        built = 1;
    }
    for (int l = built; l <= eff; ++l) launch_pyrdown(ps, pyr, s.g, l, s.batch, &ctx->launches);
    CK(cudaGetLastError());
    s.pyr_built = true;
    s.win_w = win_w; s.win_h = win_h;
    if (levels_out) *levels_out = eff;
    return 0;
}

extern "C" int svo_gpu_pyramid_levels(int w, int h, int win_w, int win_h, int max_level)
{
    if (w < 1 || h < 1 || win_w < 1 || win_h < 1 || max_level < 0) return 0;
    return effective_levels(w, h, win_w, win_h, std::min(max_level, SVO_GPU_MAX_LEVELS - 1));
}

extern "C" int svo_gpu_pyramid_margin(int win_w, int win_h) { return margin_for(win_w, win_h); }

extern "C" int svo_gpu_slot_pyramid(const svo_gpu_ctx* ctx, int slot, int* levels_out, int* margin_out)
{
    if (!ctx || slot < 0 || slot >= SVO_GPU_NUM_SLOTS) return SVO_GPU_E_ARG;
    const Slot& s = ctx->slots[slot];
    if (levels_out) *levels_out = s.pyr_built ? s.g.nlevels : 0;
    if (margin_out) *margin_out = s.pyr_built ? s.g.P : 0;
    return 0;
}

extern "C" int svo_gpu_slot_generation(const svo_gpu_ctx* ctx, int slot, unsigned long long* generation_out)
{
    if (!ctx || slot < 0 || slot >= SVO_GPU_NUM_SLOTS || !generation_out) return SVO_GPU_E_ARG;
    *generation_out = ctx->slots[slot].gen;
    return 0;
}

extern "C" int svo_gpu_build_pyramid(svo_gpu_ctx* ctx, int slot, int win_w, int win_h, int max_level, int* levels_out)
{
    if (!ctx) return SVO_GPU_E_ARG;
    if (int r = check_slot(ctx, slot)) return r;
    DeviceGuard guard(ctx->device);
    if (int r = enqueue_pyramid(ctx, ctx->slots[slot], win_w, win_h, max_level, levels_out)) return r;
    CK(cudaStreamSynchronize(ctx->st));
    return 0;
}

extern "C" int svo_gpu_download_level(svo_gpu_ctx* ctx, int slot, int lane, int level, uint8_t* host, int capacity_bytes,
                                      int* w_out, int* h_out)
{
    if (!ctx) return SVO_GPU_E_ARG;
    if (int r = check_slot(ctx, slot)) return r;
    Slot& s = ctx->slots[slot];
    if (!s.pyr_built) return fail(ctx, SVO_GPU_E_STATE, "pyramid of slot %d not built", slot);
    if (level < 0 || level >= s.g.nlevels || lane < 0 || lane >= s.batch)
        return fail(ctx, SVO_GPU_E_ARG, "level %d / lane %d out of range", level, lane);
    const LevelGeom& L = s.g.lv[level];
    if (w_out) *w_out = L.w;
    if (h_out) *h_out = L.h;
    if (!host || capacity_bytes < L.w * L.h) return fail(ctx, SVO_GPU_E_CAPACITY, "need %d bytes", L.w * L.h);
    DeviceGuard guard(ctx->device);
    const uint8_t* src = static_cast<const uint8_t*>(s.pyr.p) + (size_t)lane * s.g.lane_bytes + L.off;
    CK(cudaMemcpy2DAsync(host, (size_t)L.w, src, (size_t)L.pitch, (size_t)L.w, (size_t)L.h, cudaMemcpyDeviceToHost, ctx->st));
    CK(cudaStreamSynchronize(ctx->st));
    return 0;
}

extern "C" int svo_gpu_download_level_padded(svo_gpu_ctx* ctx, int slot, int lane, int level, uint8_t* host,
                                             int capacity_bytes, int* w_out, int* h_out, int* margin_out)
{
    if (!ctx) return SVO_GPU_E_ARG;
    if (int r = check_slot(ctx, slot)) return r;
    Slot& s = ctx->slots[slot];
    if (!s.pyr_built) return fail(ctx, SVO_GPU_E_STATE, "pyramid of slot %d not built", slot);
    if (level < 0 || level >= s.g.nlevels || lane < 0 || lane >= s.batch)
        return fail(ctx, SVO_GPU_E_ARG, "level %d / lane %d out of range", level, lane);
    const LevelGeom& L = s.g.lv[level];
    const int P = s.g.P, pw = L.w + 2 * P, ph = L.h + 2 * P;
    if (w_out) *w_out = L.w;
    if (h_out) *h_out = L.h;
    if (margin_out) *margin_out = P;
    if (!host || capacity_bytes < pw * ph) return fail(ctx, SVO_GPU_E_CAPACITY, "need %d bytes", pw * ph);
    DeviceGuard guard(ctx->device);
    const uint8_t* src = static_cast<const uint8_t*>(s.pyr.p) + (size_t)lane * s.g.lane_bytes + L.off - (long long)P * L.pitch - P;
    CK(cudaMemcpy2DAsync(host, (size_t)pw, src, (size_t)L.pitch, (size_t)pw, (size_t)ph, cudaMemcpyDeviceToHost, ctx->st));
    CK(cudaStreamSynchronize(ctx->st));
    return 0;
}

#include "capi_ops.inc"
#include "capi_fused.inc"
#include "capi_step.inc"
