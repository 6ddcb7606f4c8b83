// klt_generic.cu -- K4, generic instantiation: pyramidal Lucas-Kanade for ANY window size (the compiled 21x21 / 11x11 /
// 31x31 windows take the faster kernel of klt.cu).  One warp per feature, all pyramid levels in one launch.
//
// Replaces cv::calcOpticalFlowPyrLK at the reference's two call sites:
//   temporal tracking  src/tracking.cpp:160-165  (21x21, maxLevel 3, {COUNT+EPS, 50, 0.001}, OPTFLOW_LK_GET_MIN_EIGENVALS)
//   stereo matching    src/tracking.cpp:101-105  (11x11, maxLevel 3, {COUNT+EPS, 30, 0.001}, flags 0 -> L1 residual err)
// The arithmetic is OpenCV's LKTrackerInvoker (video/src/lkpyramid.cpp, un-vendored), restated step by step in SURVEY.md
// Appendix A.3-A.5 and in oracle/svo_oracle.c:orc_klt_ex (accum_mode 0), which this kernel matches bit for bit:
//   - 14-bit integer bilinear weights, DESCALE roundings, int16-range patch values (I*32, Scharr Ix, Iy)
//   - window sums (A11, A12, A22, b1, b2, L1 error) accumulated EXACTLY in integers (per-lane int32 partials, warp
//     reduction with REDUX on 16-bit split halves), converted to float32 once
//   - float32 epilogue in OpenCV's operation order with FMA contraction disabled (the file is compiled with -fmad=false)
//   - status cleared only by level-0 outcomes; results are not clipped to the image.
//
// Design for sm_100a (the kernel is instruction-issue bound, not HBM bound: its footprint, two pyramids, lives in L2):
//   - the I patch, Ix, Iy of the feature live in shared memory as int16 (the compiled windows of klt.cu keep them in
//     registers)
//   - source regions are staged per warp in shared memory with aligned 32-bit loads from the REFLECT_101-margined
//     pyramid (no bounds logic), all loads of a region in flight before the first store
//   - the reference's int16x2 Scharr derivative images are never materialised: because bilinear interpolation and the
//     Scharr stencil are both integer-linear, sum_ij w_ij * Scharr(I)(p+ij) == Scharr(B)(p) with B = sum_ij w_ij * I(.+ij)
//     the UNROUNDED bilinear plane, so the patch gradients come from one (w+2)x(h+2) int32 plane in shared memory.
//     Windows that touch the image border (where OpenCV's derivative image is ZERO-padded, which breaks that identity)
//     take the literal path: Scharr plane with zeros outside the image, then bilinear.
//   - lanes never diverge: every per-feature scalar is warp-uniform.
#include "klt_common.cuh"

namespace svo {

namespace {

using namespace lk;

// Window geometry (runtime).
struct Geo {
    int ww_, wh_;
    __device__ __forceinline__ Geo(int w, int h) : ww_(w), wh_(h) {}
    __device__ __forceinline__ int ww() const { return ww_; }
    __device__ __forceinline__ int wh() const { return wh_; }
    __device__ __forceinline__ int npx() const { return ww() * wh(); }
    __device__ __forceinline__ int nwords() const { return (ww() + 9) >> 2; }      // words per staged row
    __device__ __forceinline__ int rpitch() const { return nwords() * 4; }          // bytes per staged row
    __device__ __forceinline__ int bp() const { return ww() + 2; }                  // B plane row pitch (words)
    __device__ __forceinline__ int region_words() const { return (wh() + 3) * nwords(); }
    __device__ __forceinline__ int plane_words() const { return (ww() + 2) * (wh() + 2); }
};

// Stages `nrows` rows of nwords aligned 32-bit words, starting at the word holding column x0 of row y0, into sm
// (row-major, nwords per row).  Lane (r0, c0) = (lane / nwords, lane % nwords) walks rows r0, r0 + rpi, ...
// walk its rows in turn.  Returns the byte offset (0..3) of column x0 inside the first word.
__device__ __forceinline__ int stage_region(uint32_t* __restrict__ sm, const uint8_t* __restrict__ origin, int pitch,
                                            int x0, int y0, int nrows, int nwords, int r0, int c0, int rpi)
{
    const int a = x0 & 3;
    const uint8_t* src = origin + (long long)(y0 + r0) * pitch + (x0 - a) + 4 * c0;
    const long long step = (long long)rpi * pitch;
    uint32_t* dst = sm + r0 * nwords + c0;
    const int dstep = rpi * nwords;
    for (int r = r0; r < nrows; r += rpi, src += step, dst += dstep) *dst = __ldg(reinterpret_cast<const uint32_t*>(src));
    return a;
}

template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32) klt_kernel(const KltArgs a)
{
    const Geo g(a.win_w, a.win_h);
    const int ww = g.ww(), wh = g.wh(), npx = g.npx();
    const int nwords = g.nwords(), rpitch = g.rpitch(), BP = g.bp();
    const int region_words = g.region_words(), plane_words = g.plane_words();

    extern __shared__ __align__(16) uint32_t smem_dyn[];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int lane_id = blockIdx.y;
    const int n = min(a.n_dev[lane_id], a.n_max);
    const int pt = blockIdx.x * WARPS + warp;
    if (pt >= n) return;

    const int per_warp_words = region_words + plane_words + (3 * npx + 1) / 2;
    uint32_t* region = smem_dyn + warp * per_warp_words;
    int* plane = reinterpret_cast<int*>(region + region_words);   // B plane (fast path) or packed Scharr plane (border path)
    short* pIw = reinterpret_cast<short*>(plane + plane_words);
    short* pIx = pIw + npx;
    short* pIy = pIx + npx;
    const uint8_t* rbytes = reinterpret_cast<const uint8_t*>(region);

    // staging map of this lane
    const int rpi = 32 / nwords;
    int st_r0 = lane / nwords;
    const int st_c0 = lane - st_r0 * nwords;
    if (st_r0 >= rpi) st_r0 = 1 << 20;   // spare lanes stage nothing
    // plane map of this lane: element lane + 32 i sits at (u, v), advanced incrementally (no divisions in the loops)
    const int pl_v0 = lane / BP, pl_u0 = lane - pl_v0 * BP;
    const int pl_du = 32 % BP, pl_dv = 32 / BP;

    const size_t pbase = (size_t)lane_id * a.pts_lane + pt;
    const float2 prev = a.prev_xy[pbase];
    const uint8_t* Ppyr = a.prev_pyr + (size_t)lane_id * a.g.lane_bytes;
    const uint8_t* Npyr = a.next_pyr + (size_t)lane_id * a.g.lane_bytes;

    const float halfx = (ww - 1) * 0.5f, halfy = (wh - 1) * 0.5f;
    const float FLT_SCALE = 1.f / (1 << 20);
    float ox = 0.f, oy = 0.f;   // nextPts[pt]
    int st = 1;
    float er = 0.f;
    unsigned iters = 0;

    for (int level = a.max_level; level >= 0; --level) {
        const LevelGeom L = a.g.lv[level];
        const float lscale = __int_as_float((127 - level) << 23);   // (float)(1. / (1 << level))
        float px = __fmul_rn(prev.x, lscale), py = __fmul_rn(prev.y, lscale);
        float nx, ny;
        if (level == a.max_level) { nx = px; ny = py; }
        else { nx = __fmul_rn(ox, 2.f); ny = __fmul_rn(oy, 2.f); }
        ox = nx; oy = ny;

        px = __fsub_rn(px, halfx); py = __fsub_rn(py, halfy);
        const int ipx = (int)floorf(px), ipy = (int)floorf(py);
        if (ipx < -ww || ipx >= L.w || ipy < -wh || ipy >= L.h) {
            if (level == 0) { st = 0; er = 0.f; }
            continue;
        }
        BiW w = bilinear_weights(__fsub_rn(px, (float)ipx), __fsub_rn(py, (float)ipy));

        // ---- stage the (ww+3) x (wh+3) region of I whose top-left is (ipx-1, ipy-1)
        __syncwarp();
        int ao = stage_region(region, Ppyr + L.off, L.pitch, ipx - 1, ipy - 1, wh + 3, nwords, st_r0, st_c0, rpi);
        __syncwarp();
        const uint8_t* rb = rbytes + ao;

        // all derivative taps (ipx..ipx+ww, ipy..ipy+wh) inside the image?  (warp-uniform)
        const bool interior = ipx >= 0 && ipy >= 0 && ipx + ww < L.w && ipy + wh < L.h;
        int sA11 = 0, sA12 = 0, sA22 = 0;
        if (interior) {
            // ---- B(u,v) = unrounded bilinear sample of I at (ipx-1+u, ipy-1+v), u in [0,ww+1], v in [0,wh+1]
            {
                int u = pl_u0, v = pl_v0;
                for (int d = lane; d < plane_words; d += 32) {
                    plane[d] = bilin(rb + v * rpitch + u, rpitch, w);
                    u += pl_du; v += pl_dv;
                    if (u >= BP) { u -= BP; ++v; }
                }
            }
            __syncwarp();
            // Iw = DESCALE(B, 9);  Ixw = DESCALE(Scharr_x(B), 14);  Iyw = DESCALE(Scharr_y(B), 14)   (integer linearity)
            auto fast_px = [&](int q, int& iv, int& ixv, int& iyv) {
                const int y = q / ww;
                const int* c = plane + q + 2 * y + BP + 1;   // (x+1, y+1) in the plane
                const int nw_ = c[-BP - 1], n_ = c[-BP], ne_ = c[-BP + 1], w_ = c[-1], e_ = c[1];
                const int sw_ = c[BP - 1], s_ = c[BP], se_ = c[BP + 1];
                iv = descale(c[0], kWBits - 5);
                ixv = descale(3 * ((ne_ - nw_) + (se_ - sw_)) + 10 * (e_ - w_), kWBits);
                iyv = descale(3 * ((sw_ - nw_) + (se_ - ne_)) + 10 * (s_ - n_), kWBits);
            };
            for (int q = lane; q < npx; q += 32) {
                int iv, ixv, iyv;
                fast_px(q, iv, ixv, iyv);
                pIw[q] = (short)iv; pIx[q] = (short)ixv; pIy[q] = (short)iyv;
                sA11 += ixv * ixv; sA12 += ixv * iyv; sA22 += iyv * iyv;
            }
        } else {
            // ---- border path: Scharr plane at (ipx..ipx+ww, ipy..ipy+wh), ZERO outside the image (calcScharrDeriv +
            //      BORDER_CONSTANT), then bilinear on I, Ix, Iy exactly as OpenCV does
            const int dpitch = ww + 1;
            const int deriv_words = dpitch * (wh + 1);
            {
                const int dv = 32 / dpitch, du = 32 - dv * dpitch;
                int v = lane / dpitch, u = lane - v * dpitch;
                for (int d = lane; d < deriv_words; d += 32) {
                    const int X = ipx + u, Y = ipy + v;
                    uint32_t packed = 0;
                    if (X >= 0 && X < L.w && Y >= 0 && Y < L.h) {
                        const uint8_t* r0 = rb + v * rpitch + u;   // row Y-1, column X-1
                        const uint8_t* r1 = r0 + rpitch;
                        const uint8_t* r2 = r1 + rpitch;
                        const int ix = 3 * ((int)r0[2] - (int)r0[0]) + 10 * ((int)r1[2] - (int)r1[0]) + 3 * ((int)r2[2] - (int)r2[0]);
                        const int iy = 3 * ((int)r2[0] - (int)r0[0]) + 10 * ((int)r2[1] - (int)r0[1]) + 3 * ((int)r2[2] - (int)r0[2]);
                        packed = ((uint32_t)ix & 0xffffu) | ((uint32_t)iy << 16);
                    }
                    plane[d] = (int)packed;
                    u += du; v += dv;
                    if (u >= dpitch) { u -= dpitch; ++v; }
                }
            }
            __syncwarp();
            auto slow_px = [&](int q, int& iv, int& ixv, int& iyv) {
                const int y = q / ww, x = q - y * ww;
                iv = descale(bilin(rb + (y + 1) * rpitch + x + 1, rpitch, w), kWBits - 5);
                const int* dp = plane + q + y;   // y * dpitch + x
                const int d00 = dp[0], d01 = dp[1], d10 = dp[dpitch], d11 = dp[dpitch + 1];
                ixv = descale((int)(short)(d00 & 0xffff) * w.w00 + (int)(short)(d01 & 0xffff) * w.w01 +
                                  (int)(short)(d10 & 0xffff) * w.w10 + (int)(short)(d11 & 0xffff) * w.w11, kWBits);
                iyv = descale((d00 >> 16) * w.w00 + (d01 >> 16) * w.w01 + (d10 >> 16) * w.w10 + (d11 >> 16) * w.w11, kWBits);
            };
            for (int q = lane; q < npx; q += 32) {
                int iv, ixv, iyv;
                slow_px(q, iv, ixv, iyv);
                pIw[q] = (short)iv; pIx[q] = (short)ixv; pIy[q] = (short)iyv;
                sA11 += ixv * ixv; sA12 += ixv * iyv; sA22 += iyv * iyv;
            }
        }
        const float A11 = __fmul_rn(__ll2float_rn(warp_sum_exact(sA11)), FLT_SCALE);
        const float A12 = __fmul_rn(__ll2float_rn(warp_sum_exact(sA12)), FLT_SCALE);
        const float A22 = __fmul_rn(__ll2float_rn(warp_sum_exact(sA22)), FLT_SCALE);

        float D = __fsub_rn(__fmul_rn(A11, A22), __fmul_rn(A12, A12));
        const float dA = __fsub_rn(A11, A22);
        const float minEig = __fdiv_rn(
            __fsub_rn(__fadd_rn(A22, A11),
                      __fsqrt_rn(__fadd_rn(__fmul_rn(dA, dA), __fmul_rn(__fmul_rn(4.f, A12), A12)))),
            (float)(2 * ww * wh));
        if (a.flags & SVO_GPU_LK_GET_MIN_EIGENVALS) er = minEig;
        if ((double)minEig < a.min_eig_threshold || D < FLT_EPSILON) {
            if (level == 0) st = 0;
            continue;
        }
        D = __fdiv_rn(1.f, D);

        // one pass over the window at the staged position of the next image: exact integer sums of
        // diff*Ix, diff*Iy (Newton step) or |diff| (L1 residual)
        auto window_pass = [&](const uint8_t* jb, int& s1, int& s2, bool l1) {
            s1 = 0; s2 = 0;
            for (int q = lane; q < npx; q += 32) {
                const int y = q / ww, x = q - y * ww;
                const int diff = descale(bilin(jb + y * rpitch + x, rpitch, w), kWBits - 5) - (int)pIw[q];
                if (l1) s1 += abs(diff);
                else { s1 += diff * (int)pIx[q]; s2 += diff * (int)pIy[q]; }
            }
        };

        nx = __fsub_rn(nx, halfx); ny = __fsub_rn(ny, halfy);
        float pdx = 0.f, pdy = 0.f;
        const uint8_t* Jorigin = Npyr + L.off;
        for (int j = 0; j < a.max_count; ++j) {
            const int inx = (int)floorf(nx), iny = (int)floorf(ny);
            if (inx < -ww || inx >= L.w || iny < -wh || iny >= L.h) {
                if (level == 0) st = 0;
                break;
            }
            ++iters;
            w = bilinear_weights(__fsub_rn(nx, (float)inx), __fsub_rn(ny, (float)iny));
            __syncwarp();
            ao = stage_region(region, Jorigin, L.pitch, inx, iny, wh + 1, nwords, st_r0, st_c0, rpi);
            __syncwarp();
            int sb1, sb2;
            window_pass(rbytes + ao, sb1, sb2, false);
            const float b1 = __fmul_rn(__ll2float_rn(warp_sum_exact(sb1)), FLT_SCALE);
            const float b2 = __fmul_rn(__ll2float_rn(warp_sum_exact(sb2)), FLT_SCALE);
            const float dx = __fmul_rn(__fsub_rn(__fmul_rn(A12, b2), __fmul_rn(A22, b1)), D);
            const float dy = __fmul_rn(__fsub_rn(__fmul_rn(A12, b1), __fmul_rn(A11, b2)), D);
            nx = __fadd_rn(nx, dx); ny = __fadd_rn(ny, dy);
            ox = __fadd_rn(nx, halfx); oy = __fadd_rn(ny, halfy);
            if (__dadd_rn(__dmul_rn((double)dx, (double)dx), __dmul_rn((double)dy, (double)dy)) <= a.eps2) break;
            if (j > 0 && (double)fabsf(__fadd_rn(dx, pdx)) < 0.01 && (double)fabsf(__fadd_rn(dy, pdy)) < 0.01) {
                ox = __fsub_rn(ox, __fmul_rn(dx, 0.5f));
                oy = __fsub_rn(oy, __fmul_rn(dy, 0.5f));
                break;
            }
            pdx = dx; pdy = dy;
        }

        // ---- L1 patch residual at the final position (flags without GET_MIN_EIGENVALS: the stereo call)
        if (st && level == 0 && (a.flags & SVO_GPU_LK_GET_MIN_EIGENVALS) == 0) {
            const float qx = __fsub_rn(ox, halfx), qy = __fsub_rn(oy, halfy);
            const int iqx = (int)floorf(qx), iqy = (int)floorf(qy);
            if (iqx < -ww || iqx >= L.w || iqy < -wh || iqy >= L.h) {
                st = 0;
                continue;
            }
            w = bilinear_weights(__fsub_rn(qx, (float)iqx), __fsub_rn(qy, (float)iqy));
            __syncwarp();
            ao = stage_region(region, Jorigin, L.pitch, iqx, iqy, wh + 1, nwords, st_r0, st_c0, rpi);
            __syncwarp();
            int serr, unused;
            window_pass(rbytes + ao, serr, unused, true);
            er = __fdiv_rn(__fmul_rn(__ll2float_rn(warp_sum_exact(serr)), 1.f), (float)(32 * ww * wh));
        }
    }

    if (lane == 0) {
        a.next_xy[pbase] = make_float2(ox, oy);
        a.status[pbase] = (uint8_t)st;
        a.err[pbase] = er;
        if (a.iters) atomicAdd(a.iters, (unsigned long long)iters);
    }
}

}  // namespace

// any window size: patch (I*32, Ix, Iy as int16) in dynamic shared memory
void launch_klt_generic(cudaStream_t st, const KltArgs& a, int batch)
{
    constexpr int WARPS = 4;
    const int ww = a.win_w, wh = a.win_h;
    const int words = (wh + 3) * ((ww + 9) >> 2) + (ww + 2) * (wh + 2) + (3 * ww * wh + 1) / 2;
    const size_t bytes = (size_t)WARPS * words * 4;
    cudaFuncSetAttribute(klt_kernel<WARPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    dim3 grid((a.n_max + WARPS - 1) / WARPS, batch);
    klt_kernel<WARPS><<<grid, WARPS * 32, bytes, st>>>(a);
}

}  // namespace svo
