// bucket.cu -- K3: grid bucketing, bit-exact with the reference's FeatureSet::bucketingFeatures, quirks included.
//
// Reference: src/bucket.cpp:24-68 (bucketingFeatures), :72-101 (Bucket::add_feature), :103-106 (get_features),
// include/bucket.h:21-50; restated in SURVEY.md Appendix B and oracle/svo_oracle.c:orc_bucket.
//   Hc = rows / bs, Wc = cols / bs; (Hc+1)*(Wc+1) buckets addressed with stride Wc:  idx = int(y/bs)*Wc + int(x/bs)
//   a bucket keeps its first k arrivals (input order); once full, EVERY later arrival overwrites slot 0
//   (the age comparison at :91 tests the incoming age against itself), so a bucket that received m > k arrivals
//   a_0..a_{m-1} ends as [a_{m-1}, a_1, ..., a_{k-1}]
//   emission visits h in [0,Hc], w in [0,Wc] with idx = h*Wc + w, so aliased buckets are emitted twice.
//
// GPU formulation (order-independent, hence deterministic although it uses atomics):
//   per bucket we need  m = number of arrivals, the k SMALLEST input indices, and the LARGEST input index.
//   insert kernel: atomicAdd for m, atomicMax for the largest index, and a lock-free sorted insertion for the k
//   smallest: walk the bucket's k slots with atomicMin, carrying the displaced (larger) value onwards.  Every slot only
//   ever decreases and every atomicMin forwards the larger of (incoming, resident), so slot j ends up holding the
//   (j+1)-th smallest index whatever the interleaving.  A plain read of the last slot lets late arrivals skip the walk.
//   emit kernel: each CTA owns 256 consecutive entries of the emission sequence, recomputes the offset of its first
//   entry by summing min(m, k) over all earlier entries (a few thousand L2 reads), scans its own, and gathers.
#include <limits.h>
#include <stdarg.h>
#include <stdio.h>

#include <algorithm>

#include "svo_internal.cuh"

namespace svo {

namespace {

constexpr int kEmpty = 0x7f7f7f7f;   // cudaMemset(0x7f) pattern; larger than any input index

__global__ void __launch_bounds__(256) bucket_insert_kernel(const float2* __restrict__ in_xy, size_t in_lane,
                                                            const int* __restrict__ n_dev, int n_max, int Wc, int nb,
                                                            int bs, int k, BucketBufs bb)
{
    const int lane_id = blockIdx.y;
    const int n = min(n_dev[lane_id], n_max);
    int* slots = bb.cell_slots + (size_t)lane_id * bb.slots_cap;
    int* cnt = bb.cell_cnt + (size_t)lane_id * bb.cells_cap;
    int* last = bb.cell_last + (size_t)lane_id * bb.cells_cap;
    const float fbs = (float)bs;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float2 p = in_xy[(size_t)lane_id * in_lane + i];
        const float fy = __fdiv_rn(p.y, fbs), fx = __fdiv_rn(p.x, fbs);   // float / int -> float, src/bucket.cpp:50-51
        if (!(fabsf(fy) < 2.0e9f) || !(fabsf(fx) < 2.0e9f)) { bb.err_flag[lane_id] = 1; continue; }
        const long long idx = (long long)(int)fy * Wc + (int)fx;           // truncation toward zero, stride Wc (:52)
        if (idx < 0 || idx >= nb) { bb.err_flag[lane_id] = 1; continue; }   // out-of-bounds write in the reference (UB)
        atomicAdd(cnt + idx, 1);
        atomicMax(last + idx, i);
        int* s = slots + idx * k;
        if (*reinterpret_cast<volatile int*>(s + k - 1) < i) continue;     // k smaller indices are already in
        int v = i;
        for (int j = 0; j < k; ++j) {
            const int old = atomicMin(s + j, v);
            if (old == kEmpty) break;
            if (old > v) v = old;
        }
    }
}

// all per-call state of every lane in one launch: slots = empty, counts = 0, largest index = -1, error flag = 0
__global__ void __launch_bounds__(256) bucket_clear_kernel(BucketBufs bb, int batch)
{
    const size_t nslots = bb.slots_cap * (size_t)batch, ncells = bb.cells_cap * (size_t)batch;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nslots; i += stride) bb.cell_slots[i] = kEmpty;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < ncells; i += stride) {
        bb.cell_cnt[i] = 0;
        bb.cell_last[i] = -1;
    }
    if (blockIdx.x == 0 && threadIdx.x < batch) bb.err_flag[threadIdx.x] = 0;
    if (blockIdx.x == 0)
        for (int i = threadIdx.x + 256; i < batch; i += 256) bb.err_flag[i] = 0;
}

__device__ __forceinline__ int emit_count(const int* __restrict__ cnt, int e, int Wc, int k)
{
    const int hh = e / (Wc + 1), ww = e - hh * (Wc + 1);
    return min(cnt[hh * Wc + ww], k);
}

__global__ void __launch_bounds__(256) bucket_emit_kernel(const float2* __restrict__ in_xy, const int* __restrict__ in_age,
                                                          size_t in_lane, int Wc, int nemit, int k, BucketBufs bb)
{
    __shared__ int wsum[8];
    __shared__ int sbase;
    const int lane_id = blockIdx.y;
    const int tid = threadIdx.x, warp = tid >> 5, ln = tid & 31;
    const int* slots = bb.cell_slots + (size_t)lane_id * bb.slots_cap;
    const int* cnt = bb.cell_cnt + (size_t)lane_id * bb.cells_cap;
    const int* last = bb.cell_last + (size_t)lane_id * bb.cells_cap;
    const int e0 = blockIdx.x * 256;

    // offset of this CTA's first entry
    int part = 0;
    for (int e = tid; e < e0; e += 256) part += emit_count(cnt, e, Wc, k);
#pragma unroll
    for (int o = 16; o; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    if (ln == 0) wsum[warp] = part;
    __syncthreads();
    if (tid == 0) {
        int s = 0;
        for (int i = 0; i < 8; ++i) s += wsum[i];
        sbase = s;
    }
    __syncthreads();
    const int base = sbase;
    __syncthreads();

    // exclusive scan of this CTA's 256 entries
    const int e = e0 + tid;
    int idx = 0, m = 0;
    if (e < nemit) {
        const int hh = e / (Wc + 1), ww = e - hh * (Wc + 1);
        idx = hh * Wc + ww;
        m = min(cnt[idx], k);
    }
    int inc = m;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, inc, o);
        if (ln >= o) inc += v;
    }
    if (ln == 31) wsum[warp] = inc;
    __syncthreads();
    int woff = 0;
    for (int i = 0; i < warp; ++i) woff += wsum[i];
    const int off = base + woff + inc - m;

    if (m > 0) {
        const float2* ixy = in_xy + (size_t)lane_id * in_lane;
        const int* iag = in_age ? in_age + (size_t)lane_id * in_lane : nullptr;
        float2* oxy = bb.out_xy + (size_t)lane_id * bb.out_cap;
        int* oag = bb.out_age + (size_t)lane_id * bb.out_cap;
        const bool overflow = cnt[idx] > k;
        for (int s = 0; s < m; ++s) {
            const int src = (s == 0 && overflow) ? last[idx] : slots[(size_t)idx * k + s];
            const int o = off + s;
            if (o < bb.out_cap) {
                oxy[o] = ixy[src];
                oag[o] = iag ? iag[src] : 0;
            }
        }
    }
    if (e == nemit - 1) bb.out_count[lane_id] = off + m;   // required output length (may exceed out_cap)
}

// ---- one CTA per lane, the whole bucket table of the lane in shared memory: clear, insert and emit in one launch.
// Same formulation as the three kernels above (count, k smallest indices by atomicMin chains, largest index; emission over
// the aliased sequence), but the atomics are shared-memory atomics and nothing is cleared or re-read through L2.  Used when
// several lanes run at once (every lane gets an SM of its own) and the table fits.
constexpr int kFusedThreads = 1024;

__global__ void __launch_bounds__(kFusedThreads) bucket_lane_kernel(const float2* __restrict__ in_xy, const int* __restrict__ in_age,
                                                                    size_t in_lane, const int* __restrict__ n_dev, int n_max,
                                                                    int Wc, int nb, int bs, int k, BucketBufs bb)
{
    extern __shared__ int sm[];
    int* cnt = sm;               // [nb]
    int* last = sm + nb;         // [nb]
    int* slots = sm + 2 * nb;    // [nb * k]
    __shared__ int wsum[kFusedThreads / 32];
    __shared__ int s_err;
    const int lane_id = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, ln = tid & 31;
    const int n = min(n_dev[lane_id], n_max);
    const float2* ixy = in_xy + (size_t)lane_id * in_lane;

    for (int i = tid; i < nb; i += kFusedThreads) { cnt[i] = 0; last[i] = -1; }
    for (int i = tid; i < nb * k; i += kFusedThreads) slots[i] = kEmpty;
    if (tid == 0) s_err = 0;
    __syncthreads();

    const float fbs = (float)bs;
    float2 pnext = tid < n ? ixy[tid] : make_float2(0.f, 0.f);
    for (int i = tid; i < n; i += kFusedThreads) {
        const float2 p = pnext;
        if (i + kFusedThreads < n) pnext = ixy[i + kFusedThreads];        // in flight while this point is inserted
        const float fy = __fdiv_rn(p.y, fbs), fx = __fdiv_rn(p.x, fbs);   // float / int -> float, src/bucket.cpp:50-51
        if (!(fabsf(fy) < 2.0e9f) || !(fabsf(fx) < 2.0e9f)) { s_err = 1; continue; }
        const long long idx = (long long)(int)fy * Wc + (int)fx;           // truncation toward zero, stride Wc (:52)
        if (idx < 0 || idx >= nb) { s_err = 1; continue; }                  // out-of-bounds write in the reference (UB)
        atomicAdd(cnt + idx, 1);
        atomicMax(last + idx, i);
        int* s = slots + idx * k;
        if (*reinterpret_cast<volatile int*>(s + k - 1) < i) continue;     // k smaller indices are already in
        int v = i;
        for (int j = 0; j < k; ++j) {
            const int old = atomicMin(s + j, v);
            if (old == kEmpty) break;
            if (old > v) v = old;
        }
    }
    __syncthreads();
    if (tid == 0) bb.err_flag[lane_id] = s_err;

    // emission: entry e of the (Hc+1) x (Wc+1) sequence reads bucket hh * Wc + ww (aliased buckets come out twice)
    const int nemit = nb;
    const int* iag = in_age ? in_age + (size_t)lane_id * in_lane : nullptr;
    float2* oxy = bb.out_xy + (size_t)lane_id * bb.out_cap;
    int* oag = bb.out_age + (size_t)lane_id * bb.out_cap;
    int base = 0;
    for (int e0 = 0; e0 < nemit; e0 += kFusedThreads) {
        const int e = e0 + tid;
        int idx = 0, m = 0;
        if (e < nemit) {
            const int hh = e / (Wc + 1), ww = e - hh * (Wc + 1);
            idx = hh * Wc + ww;
            m = min(cnt[idx], k);
        }
        int inc = m;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, inc, o);
            if (ln >= o) inc += v;
        }
        if (ln == 31) wsum[warp] = inc;
        __syncthreads();
        int woff = 0, tot = 0;
        for (int i = 0; i < kFusedThreads / 32; ++i) {
            const int v = wsum[i];
            if (i < warp) woff += v;
            tot += v;
        }
        const int off = base + woff + inc - m;
        if (m > 0) {
            const bool overflow = cnt[idx] > k;
            for (int s = 0; s < m; ++s) {
                const int src = (s == 0 && overflow) ? last[idx] : slots[idx * k + s];
                const int o = off + s;
                if (o < bb.out_cap) {
                    oxy[o] = ixy[src];
                    oag[o] = iag ? iag[src] : 0;
                }
            }
        }
        base += tot;
        __syncthreads();   // wsum is rewritten by the next chunk
    }
    if (tid == 0) bb.out_count[lane_id] = base;   // required output length (may exceed out_cap)
}

}  // namespace

void launch_bucket(cudaStream_t st, const float2* in_xy, const int* in_age, size_t in_lane, const int* n_dev, int n_max,
                   int img_w, int img_h, int bucket_size, int per_bucket, const BucketBufs& bb, int batch,
                   long long* launches, int n_hint)
{
    const int Hc = img_h / bucket_size, Wc = img_w / bucket_size;
    const int nb = (Hc + 1) * (Wc + 1);
    const int k = per_bucket;
    static std::atomic<unsigned long long> seen{0};
    int dev = 0, smem_max = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    smem_max -= 1024;   // the kernel's static shared memory
    if (first_on_device(seen)) {
        prefer_max_shared(bucket_insert_kernel); prefer_max_shared(bucket_emit_kernel); prefer_max_shared(bucket_clear_kernel);
        cudaFuncSetAttribute(bucket_lane_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max);
    }
    // several lanes at once, a table that fits in shared memory and a list one CTA walks quickly (about 2 us per 1024 points;
    // measured crossover against the three kernels near 24 k points per lane): one CTA per lane does the whole call
    const size_t table = sizeof(int) * (size_t)nb * (size_t)(2 + k);
    const int n_expected = n_hint > 0 ? std::min(n_max, n_hint) : n_max;
    // (one or a few lanes: the single launch wins while one CTA walks the list in less time than two more launches cost,
    //  measured at one lane with 5 300 points: 22 -> 18 us)
    if (bb.mode != 1 && table <= (size_t)smem_max &&
        ((batch >= 4 && n_expected <= 24576) || (batch < 4 && n_expected <= 8192) || bb.mode == 2)) {
        bucket_lane_kernel<<<batch, kFusedThreads, table, st>>>(in_xy, in_age, in_lane, n_dev, n_max, Wc, nb, bucket_size, k, bb);
        if (launches) ++*launches;
        return;
    }
    // whole-batch clear (lanes are contiguous), one launch
    {
        const size_t n = bb.slots_cap * (size_t)batch;
        const int blocks = (int)std::min<size_t>((n + 1023) / 1024, 1184);   // 4 elements per thread, at most 148 x 8 CTAs
        bucket_clear_kernel<<<std::max(blocks, 1), 256, 0, st>>>(bb, batch);
    }
    // the grid covers the expected list length (n_hint, when the caller has one) -- the grid-stride loop of the kernel
    // makes any grid correct, a tighter one only spares CTAs that would find nothing to do
    const int n_grid = n_hint > 0 ? std::min(n_max, n_hint) : n_max;
    int nblk = (n_grid + 255) / 256;
    nblk = nblk < 1 ? 1 : (nblk > 592 ? 592 : nblk);   // 148 SMs x 4
    bucket_insert_kernel<<<dim3(nblk, batch), 256, 0, st>>>(in_xy, in_lane, n_dev, n_max, Wc, nb, bucket_size, k, bb);
    bucket_emit_kernel<<<dim3((nb + 255) / 256, batch), 256, 0, st>>>(in_xy, in_age, in_lane, Wc, nb, k, bb);
    if (launches) *launches += 3;
}

}  // namespace svo

// ================================================================================================ svo_gpu_bucket_static
// Bucketing for callers that are not allowed to allocate (include/svo_gpu.h has the user-facing contract).  It runs the very
// same clear / insert / emit kernels as launch_bucket's global-memory path; the only difference is where the per-bucket
// tables live: not in cudaMalloc'ed context buffers but in the file-scope device array below, which the CUDA module reserves
// once for every device that loads this library and never gives back.
namespace svo {
namespace {

constexpr int kStaticInts = SVO_GPU_BUCKET_STATIC_BYTES / (int)sizeof(int);
constexpr int kStaticHead = 64;   // ints [0] output length, [1] error flag, [2] "no limit" list length; tables start at 256 B
constexpr int kStaticMaxDevices = 64;

__device__ __align__(256) int g_bucket_static[kStaticInts];

// `g_bucket_static` above is what DEVICE code calls the array.  In host code that name is only the runtime's handle of the
// symbol; the address a kernel can dereference has to be asked for, once per device and with that device current.
std::atomic<int*> g_bucket_static_dev[kStaticMaxDevices];

int static_fail(char* buf, size_t len, int code, const char* fmt, ...)
{
    if (buf && len) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(buf, len, fmt, ap);
        va_end(ap);
    }
    return code;
}

bool on_device(const void* p)
{
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

}  // namespace
}  // namespace svo

extern "C" int svo_gpu_bucket_static(const float* xy_in_dev, const int* ages_in_dev_or_null, int n_in, int img_w, int img_h,
                                     int bucket_size, int per_bucket, float* xy_out_dev, int* ages_out_dev, int capacity,
                                     int* n_out, void* stream, char* errbuf_or_null, size_t errbuf_len)
{
    using namespace svo;
    char* eb = errbuf_or_null;
    const size_t el = errbuf_len;
    if (eb && el) eb[0] = 0;
    // ---- the same argument rules as svo_gpu_bucket
    if (!n_out || n_in < 0 || capacity < 0 || (n_in > 0 && !xy_in_dev) || (capacity > 0 && (!xy_out_dev || !ages_out_dev)))
        return static_fail(eb, el, SVO_GPU_E_ARG, "svo_gpu_bucket_static: null pointer or negative count");
    if (bucket_size < 1 || img_w < 0 || img_h < 0)
        return static_fail(eb, el, SVO_GPU_E_ARG, "bucket_size %d / image %dx%d invalid", bucket_size, img_w, img_h);
    if (per_bucket < 1 && n_in > 0) return static_fail(eb, el, SVO_GPU_E_ARG, "features_per_bucket %d < 1", per_bucket);
    if (per_bucket < 1) per_bucket = 1;
    if (((uintptr_t)xy_in_dev | (uintptr_t)xy_out_dev) & 7u)
        return static_fail(eb, el, SVO_GPU_E_ARG, "xy_in_dev / xy_out_dev must be 8-byte aligned (they are read and written as x,y pairs)");
    // ---- the limit of this entry point: the bucket tables have to fit the static array
    const long long Hc = img_h / bucket_size, Wc = img_w / bucket_size;
    const long long nb = (Hc + 1) * (Wc + 1);
    const long long need = kStaticHead + nb * (2 + (long long)per_bucket);
    if (need > kStaticInts)
        return static_fail(eb, el, SVO_GPU_E_CAPACITY,
                           "svo_gpu_bucket_static: %lld buckets x (2 + %d per bucket) need %lld bytes of table, the static scratch "
                           "array holds %d (SVO_GPU_BUCKET_STATIC_BYTES); use svo_gpu_bucket, which sizes its tables with cudaMalloc",
                           nb, per_bucket, need * 4, SVO_GPU_BUCKET_STATIC_BYTES);
    if ((n_in > 0 && !on_device(xy_in_dev)) || (ages_in_dev_or_null && !on_device(ages_in_dev_or_null)) ||
        (capacity > 0 && (!on_device(xy_out_dev) || !on_device(ages_out_dev))))
        return static_fail(eb, el, SVO_GPU_E_ARG, "svo_gpu_bucket_static: the point / age arrays must be device memory");

    // ---- set-up, once per device: where the static array lives on the device that is current now
    int dev = -1;
    cudaError_t ce = cudaGetDevice(&dev);
    if (ce != cudaSuccess) return static_fail(eb, el, SVO_GPU_E_CUDA, "cudaGetDevice: %s", cudaGetErrorString(ce));
    if (dev < 0 || dev >= kStaticMaxDevices) return static_fail(eb, el, SVO_GPU_E_ARG, "device ordinal %d not supported", dev);
    int* base = g_bucket_static_dev[dev].load(std::memory_order_acquire);
    if (!base) {
        ce = cudaGetSymbolAddress((void**)&base, g_bucket_static);
        if (ce != cudaSuccess || !base)
            return static_fail(eb, el, SVO_GPU_E_CUDA, "cudaGetSymbolAddress(static bucket scratch) on device %d: %s", dev,
                               cudaGetErrorString(ce));
        g_bucket_static_dev[dev].store(base, std::memory_order_release);   // every thread would store the same address
    }

    cudaStream_t st = (cudaStream_t)stream;
    BucketBufs bb{};
    bb.err_flag = base + 1;
    bb.cell_cnt = base + kStaticHead;
    bb.cell_last = bb.cell_cnt + nb;
    bb.cell_slots = bb.cell_last + nb;
    bb.cells_cap = (size_t)nb;                       // exactly this call's table: the clear kernel touches nothing else
    bb.slots_cap = (size_t)nb * (size_t)per_bucket;
    bb.out_xy = reinterpret_cast<float2*>(xy_out_dev);
    bb.out_age = ages_out_dev;
    bb.out_count = base;
    bb.out_cap = capacity;
    bb.mode = 1;

    // The array still holds whatever the previous call (any thread, any stream) left there: everything this call reads is
    // rewritten first, in stream order -- the list-length word by the memset, the tables and the error flag by the clear
    // kernel, the output length by the emit kernel.
    ce = cudaMemsetAsync(base + 2, 0x7f, sizeof(int), st);   // 0x7f7f7f7f: "as many points as n_in says"
    if (ce != cudaSuccess) return static_fail(eb, el, SVO_GPU_E_CUDA, "cudaMemsetAsync: %s", cudaGetErrorString(ce));
    {
        const size_t n = bb.slots_cap;
        const int blocks = (int)std::min<size_t>((n + 1023) / 1024, 1184);
        bucket_clear_kernel<<<std::max(blocks, 1), 256, 0, st>>>(bb, 1);
    }
    int nblk = (n_in + 255) / 256;
    nblk = nblk < 1 ? 1 : (nblk > 592 ? 592 : nblk);
    bucket_insert_kernel<<<dim3(nblk, 1), 256, 0, st>>>(reinterpret_cast<const float2*>(xy_in_dev), 0, base + 2, n_in, (int)Wc,
                                                        (int)nb, bucket_size, per_bucket, bb);
    bucket_emit_kernel<<<dim3((unsigned)((nb + 255) / 256), 1), 256, 0, st>>>(reinterpret_cast<const float2*>(xy_in_dev),
                                                                              ages_in_dev_or_null, 0, (int)Wc, (int)nb,
                                                                              per_bucket, bb);
    ce = cudaGetLastError();
    if (ce != cudaSuccess) return static_fail(eb, el, SVO_GPU_E_CUDA, "bucket kernels: %s", cudaGetErrorString(ce));
    int head[2] = {0, 0};   // output length, error flag: two words to the caller's stack, no staging buffer of ours
    ce = cudaMemcpyAsync(head, base, sizeof(head), cudaMemcpyDeviceToHost, st);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(st);
    if (ce != cudaSuccess) return static_fail(eb, el, SVO_GPU_E_CUDA, "svo_gpu_bucket_static: %s", cudaGetErrorString(ce));
    if (head[1])
        return static_fail(eb, el, SVO_GPU_E_ARG, "a point maps outside the bucket table (undefined behaviour in the reference)");
    *n_out = head[0];
    if (head[0] > capacity)
        return static_fail(eb, el, SVO_GPU_E_CAPACITY, "%d bucketed features exceed the capacity %d", head[0], capacity);
    return SVO_GPU_OK;
}
