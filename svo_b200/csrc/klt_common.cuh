// klt_common.cuh -- arithmetic shared by the two LK kernels (klt.cu: compiled windows, klt_generic.cu: any window).
// Every function here restates one step of OpenCV's LKTrackerInvoker (video/src/lkpyramid.cpp, un-vendored; SURVEY.md
// Appendix A.5) with its exact rounding; the float32 steps use __f*_rn so that no FMA contraction can reorder them.
#pragma once
#include <float.h>

#include "svo_internal.cuh"

namespace svo {
namespace lk {

constexpr unsigned kFull = 0xffffffffu;
constexpr int kWBits = 14;

__device__ __forceinline__ int descale(int v, int n) { return (v + (1 << (n - 1))) >> n; }

struct BiW { int w00, w01, w10, w11; };

// iw00 = cvRound((1-a)(1-b) 2^14), iw01 = cvRound(a (1-b) 2^14), iw10 = cvRound((1-a) b 2^14), iw11 = 2^14 - the rest
__device__ __forceinline__ BiW bilinear_weights(float a, float b)
{
    const float oa = __fsub_rn(1.f, a), ob = __fsub_rn(1.f, b);
    BiW w;
    w.w00 = __float2int_rn(__fmul_rn(__fmul_rn(oa, ob), 16384.f));
    w.w01 = __float2int_rn(__fmul_rn(__fmul_rn(a, ob), 16384.f));
    w.w10 = __float2int_rn(__fmul_rn(__fmul_rn(oa, b), 16384.f));
    w.w11 = (1 << kWBits) - w.w00 - w.w01 - w.w10;
    return w;
}

// exact warp sum of per-lane int32 partials (the total may need more than 32 bits): REDUX on the two 16-bit halves
__device__ __forceinline__ long long warp_sum_exact(int v)
{
    const unsigned lo = __reduce_add_sync(kFull, (unsigned)(v & 0xffff));
    const int hi = __reduce_add_sync(kFull, v >> 16);
    return ((long long)hi << 16) + (long long)lo;
}

// the same sum as a correctly rounded float32, i.e. (float)(exact 64-bit total): both partial sums are below 2^22 in
// magnitude, so hi * 65536 and lo are exact floats and the single rounding happens in the final addition
__device__ __forceinline__ float warp_sum_exact_f(int v)
{
    const unsigned lo = __reduce_add_sync(kFull, (unsigned)(v & 0xffff));
    const int hi = __reduce_add_sync(kFull, v >> 16);
    return __fadd_rn(__fmul_rn(__int2float_rn(hi), 65536.f), __uint2float_rn(lo));
}

// FLT_SCALE (2^-20) times that sum, one instruction shorter: hi 2^-4 and lo 2^-20 are exact, so the FMA rounds once, at the
// same place as the addition above (scaling by a power of two commutes with round-to-nearest; no underflow, |sum| >= 1 or 0)
__device__ __forceinline__ float warp_sum_exact_scaled(int v)
{
    const unsigned lo = __reduce_add_sync(kFull, (unsigned)(v & 0xffff));
    const int hi = __reduce_add_sync(kFull, v >> 16);
    return __fmaf_rn(__uint2float_rn(lo), 1.f / (1 << 20), __fmul_rn(__int2float_rn(hi), 1.f / 16.f));
}

// unrounded bilinear sample at byte pointer p (row pitch rp)
__device__ __forceinline__ int bilin(const uint8_t* __restrict__ p, int rp, const BiW& w)
{
    return (int)p[0] * w.w00 + (int)p[1] * w.w01 + (int)p[rp] * w.w10 + (int)p[rp + 1] * w.w11;
}

// minEig = (A22 + A11 - sqrt((A11-A22)^2 + 4 A12^2)) / (2 w h), OpenCV's operation order
__device__ __forceinline__ float min_eig(float A11, float A12, float A22, int npix)
{
    const float dA = __fsub_rn(A11, A22);
    return __fdiv_rn(__fsub_rn(__fadd_rn(A22, A11),
                               __fsqrt_rn(__fadd_rn(__fmul_rn(dA, dA), __fmul_rn(__fmul_rn(4.f, A12), A12)))),
                     (float)(2 * npix));
}

}  // namespace lk

// klt_generic.cu: any window up to SVO_GPU_MAX_WIN, patch in shared memory
void launch_klt_generic(cudaStream_t st, const KltArgs& a, int batch);

}  // namespace svo
