// filter.cu -- K5: the status / epipolar filters that follow each LK call in the reference, as an order-preserving
// stream compaction on the device, so that only surviving tracks cross PCIe.
//
//   temporal  src/tracking.cpp:168-175   keep i iff status[i]
//   stereo    src/tracking.cpp:109-115   keep i iff status[i] && std::abs(right[i].y - left[i].y) < y_threshold
//                                        (float arithmetic: Point2f members and `float y_threshold`, include/config_reader.h:41)
//
// Survivors keep their input order (the reference pushes them back in index order), and `index` maps each survivor to its
// input position so that the host can carry Feature::mapPoint over (src/tracking.cpp:112,171) without a second pass.
//
// One CTA per lane (independent stream); the CTA walks the lane's list in chunks of 1024 with a running offset: ballot +
// popc inside a warp, one shared-memory scan over the 32 warp totals.  HBM-bound by bytes (21 B read, up to 20 B written
// per track), latency-bound at the reference's list sizes (2k-50k tracks).
#include "svo_internal.cuh"

namespace svo {

namespace {

constexpr int kFilterThreads = 1024;

__global__ void __launch_bounds__(kFilterThreads) filter_compact_kernel(FilterArgs a)
{
    __shared__ int warp_tot[32];
    __shared__ int chunk_base;
    const int lane_id = blockIdx.x;
    const int n = min(a.n_dev[lane_id], a.n_max);
    const size_t base = (size_t)lane_id * a.pts_lane;
    const float2* __restrict__ A = a.a_xy + base;
    const float2* __restrict__ B = a.b_xy + base;
    const uint8_t* __restrict__ S = a.status + base;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) chunk_base = 0;
    __syncthreads();
    for (int c0 = 0; c0 < n; c0 += kFilterThreads) {
        const int i = c0 + threadIdx.x;
        bool keep = false;
        float2 pa = make_float2(0.f, 0.f), pb = pa;
        if (i < n) {
            pa = A[i];
            pb = B[i];
            keep = S[i] != 0;
            if (a.use_y) keep = keep && fabsf(__fsub_rn(pb.y, pa.y)) < a.y_threshold;
        }
        const unsigned m = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) warp_tot[warp] = __popc(m);
        __syncthreads();
        const int start = chunk_base;
        int before = 0, total = 0;
        {
            // every warp scans the 32 warp totals redundantly (one value per lane)
            const int t = warp_tot[lane];
            int incl = t;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int u = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += u;
            }
            before = __shfl_sync(0xffffffffu, incl - t, warp);
            total = __shfl_sync(0xffffffffu, incl, 31);
        }
        if (keep) {
            const int o = start + before + __popc(m & ((1u << lane) - 1u));
            if (o < a.out_cap) {
                if (a.out_index) a.out_index[base + o] = i;
                if (a.out_a) a.out_a[base + o] = pa;
                if (a.out_b) a.out_b[base + o] = pb;
                if (a.out_err) a.out_err[base + o] = a.err[base + i];
            }
        }
        __syncthreads();
        if (threadIdx.x == 0) chunk_base = start + total;
        __syncthreads();
    }
    if (threadIdx.x == 0) a.out_n[lane_id] = chunk_base;   // may exceed out_cap: the caller reports the overflow
}

}  // namespace

void launch_filter(cudaStream_t st, const FilterArgs& a, int batch, long long* launches)
{
    if (batch <= 0) return;
    filter_compact_kernel<<<batch, kFilterThreads, 0, st>>>(a);
    if (launches) ++*launches;
}

}  // namespace svo
