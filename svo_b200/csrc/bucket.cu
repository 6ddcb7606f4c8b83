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
