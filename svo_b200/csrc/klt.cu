// klt.cu -- K4: pyramidal Lucas-Kanade, one warp per feature, all pyramid levels in one launch.
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
// Design for sm_100a: the I patch, Ix, Iy of the feature live in REGISTERS (ceil(w*h/32) pixels per lane) for the
// compiled window sizes (21x21, 11x11, 31x31); the (w+3)x(h+3) source region of each level is staged once per level in
// the warp's slice of shared memory with aligned 32-bit loads from the REFLECT_101-margined pyramid (no bounds logic),
// the Scharr derivative plane is computed there on the fly (the reference's int16x2 derivative images are never written
// to HBM), and each Newton iteration stages the (w+1)x(h+1) region of the next image the same way.  Lanes never
// diverge: all per-feature scalars are warp-uniform.  Other window sizes take the generic kernel (patch in shared memory).
#include <float.h>

#include "svo_internal.cuh"

namespace svo {

namespace {

constexpr unsigned kFull = 0xffffffffu;
constexpr int kWBits = 14;

__device__ __forceinline__ int descale(int v, int n) { return (v + (1 << (n - 1))) >> n; }

struct BiW { int w00, w01, w10, w11; };

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

// stage `nrows` rows of `nwords` aligned 32-bit words starting at the word that holds column x0 of row y0;
// returns the byte offset (0..3) of column x0 inside the first word
__device__ __forceinline__ int stage_region(uint32_t* __restrict__ sm, const uint8_t* __restrict__ origin, int pitch,
                                            int x0, int y0, int nrows, int nwords, int lane)
{
    const int a = x0 & 3;
    const uint8_t* src = origin + (long long)y0 * pitch + (x0 - a);
    const int total = nrows * nwords;
    for (int i = lane; i < total; i += 32) {
        const int r = i / nwords, c = i - r * nwords;
        sm[i] = __ldg(reinterpret_cast<const uint32_t*>(src + (long long)r * pitch) + c);
    }
    return a;
}

template <int WW, int WH>
struct Cfg {
    static constexpr bool kFixed = WW > 0;
    static constexpr int kNPL = kFixed ? (WW * WH + 31) / 32 : 1;   // patch pixels per lane (register-resident)
};

// WW, WH > 0: compiled window, patch in registers, static shared memory.  WW == 0: runtime window, patch in shared memory.
template <int WW, int WH, int WARPS>
__global__ void __launch_bounds__(WARPS * 32) klt_kernel(const KltArgs a)
{
    using C = Cfg<WW, WH>;
    constexpr bool FIXED = C::kFixed;
    constexpr int NPL = C::kNPL;
    const int ww = FIXED ? WW : a.win_w;
    const int wh = FIXED ? WH : a.win_h;
    const int npx = ww * wh;
    const int nwords = (ww + 9) >> 2;        // words per staged row
    const int rpitch = nwords * 4;           // bytes per staged row
    const int dpitch = ww + 1;               // derivative plane row pitch (words)
    const int region_words = (wh + 3) * nwords;
    const int deriv_words = (ww + 1) * (wh + 1);

    extern __shared__ __align__(16) uint32_t smem_dyn[];
    constexpr int kStaticWords =
        FIXED ? WARPS * ((WH + 3) * ((WW + 9) >> 2) + (WW + 1) * (WH + 1)) : 1;
    __shared__ __align__(16) uint32_t smem_static[kStaticWords];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int lane_id = blockIdx.y;
    const int n = min(a.n_dev[lane_id], a.n_max);
    const int pt = blockIdx.x * WARPS + warp;
    if (pt >= n) return;

    const int per_warp_words = region_words + deriv_words + (FIXED ? 0 : (3 * npx + 1) / 2);
    uint32_t* region = (FIXED ? smem_static : smem_dyn) + warp * per_warp_words;
    uint32_t* deriv = region + region_words;
    short* pIw = reinterpret_cast<short*>(deriv + deriv_words);   // generic only
    short* pIx = pIw + npx;
    short* pIy = pIx + npx;
    const uint8_t* rbytes = reinterpret_cast<const uint8_t*>(region);

    // per-lane pixel assignment q = lane + 32 k  ->  (x, y) in the window
    int offI[NPL], offD[NPL];
    int Iw[NPL], Ix[NPL], Iy[NPL];
    if constexpr (FIXED) {
#pragma unroll
        for (int k = 0; k < NPL; ++k) {
            int q = lane + 32 * k;
            if (q >= WW * WH) q = 0;   // padding pixel: reads pixel 0, contributes nothing (Ix = Iy = 0, see below)
            const int y = q / WW, x = q - y * WW;
            offI[k] = y * rpitch + x;
            offD[k] = y * dpitch + x;
        }
    }

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
        int ao = stage_region(region, Ppyr + L.off, L.pitch, ipx - 1, ipy - 1, wh + 3, nwords, lane);
        __syncwarp();

        // ---- Scharr derivative plane at (ipx..ipx+ww, ipy..ipy+wh): zero outside the image (calcScharrDeriv + BORDER_CONSTANT)
        for (int d = lane; d < deriv_words; d += 32) {
            const int dy = d / dpitch, dx = d - dy * dpitch;
            const int X = ipx + dx, Y = ipy + dy;
            uint32_t packed = 0;
            if (X >= 0 && X < L.w && Y >= 0 && Y < L.h) {
                const uint8_t* r0 = rbytes + dy * rpitch + ao + dx;   // row Y-1, column X-1
                const uint8_t* r1 = r0 + rpitch;
                const uint8_t* r2 = r1 + rpitch;
                const int ix = 3 * ((int)r0[2] - (int)r0[0]) + 10 * ((int)r1[2] - (int)r1[0]) + 3 * ((int)r2[2] - (int)r2[0]);
                const int iy = 3 * ((int)r2[0] - (int)r0[0]) + 10 * ((int)r2[1] - (int)r0[1]) + 3 * ((int)r2[2] - (int)r0[2]);
                packed = ((uint32_t)ix & 0xffffu) | ((uint32_t)iy << 16);
            }
            deriv[d] = packed;
        }
        __syncwarp();

        // ---- patch: Iw = I*32 (bilinear, DESCALE 9), Ixw, Iyw (bilinear, DESCALE 14); Gram sums in exact integers
        int sA11 = 0, sA12 = 0, sA22 = 0;
        auto patch_px = [&](int oI, int oD, int& ival, int& ixval, int& iyval) {
            const uint8_t* p = rbytes + oI + rpitch + 1 + ao;   // (ipx + x, ipy + y)
            ival = descale((int)p[0] * w.w00 + (int)p[1] * w.w01 + (int)p[rpitch] * w.w10 + (int)p[rpitch + 1] * w.w11,
                           kWBits - 5);
            const uint32_t d00 = deriv[oD], d01 = deriv[oD + 1], d10 = deriv[oD + dpitch], d11 = deriv[oD + dpitch + 1];
            ixval = descale((int)(short)(d00 & 0xffff) * w.w00 + (int)(short)(d01 & 0xffff) * w.w01 +
                                (int)(short)(d10 & 0xffff) * w.w10 + (int)(short)(d11 & 0xffff) * w.w11, kWBits);
            iyval = descale(((int)d00 >> 16) * w.w00 + ((int)d01 >> 16) * w.w01 + ((int)d10 >> 16) * w.w10 +
                                ((int)d11 >> 16) * w.w11, kWBits);
        };
        if constexpr (FIXED) {
#pragma unroll
            for (int k = 0; k < NPL; ++k) {
                int iv, ixv, iyv;
                patch_px(offI[k], offD[k], iv, ixv, iyv);
                if (lane + 32 * k >= WW * WH) { ixv = 0; iyv = 0; }
                Iw[k] = iv; Ix[k] = ixv; Iy[k] = iyv;
                sA11 += ixv * ixv; sA12 += ixv * iyv; sA22 += iyv * iyv;
            }
        } else {
            for (int q = lane; q < npx; q += 32) {
                const int y = q / ww, x = q - y * ww;
                int iv, ixv, iyv;
                patch_px(y * rpitch + x, y * dpitch + x, iv, ixv, iyv);
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
            ao = stage_region(region, Jorigin, L.pitch, inx, iny, wh + 1, nwords, lane);
            __syncwarp();
            int sb1 = 0, sb2 = 0;
            if constexpr (FIXED) {
#pragma unroll
                for (int k = 0; k < NPL; ++k) {
                    const uint8_t* p = rbytes + offI[k] + ao;
                    const int v = (int)p[0] * w.w00 + (int)p[1] * w.w01 + (int)p[rpitch] * w.w10 + (int)p[rpitch + 1] * w.w11;
                    const int diff = descale(v, kWBits - 5) - Iw[k];
                    sb1 += diff * Ix[k];
                    sb2 += diff * Iy[k];
                }
            } else {
                for (int q = lane; q < npx; q += 32) {
                    const int y = q / ww, x = q - y * ww;
                    const uint8_t* p = rbytes + y * rpitch + x + ao;
                    const int v = (int)p[0] * w.w00 + (int)p[1] * w.w01 + (int)p[rpitch] * w.w10 + (int)p[rpitch + 1] * w.w11;
                    const int diff = descale(v, kWBits - 5) - (int)pIw[q];
                    sb1 += diff * (int)pIx[q];
                    sb2 += diff * (int)pIy[q];
                }
            }
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
            ao = stage_region(region, Jorigin, L.pitch, iqx, iqy, wh + 1, nwords, lane);
            __syncwarp();
            int serr = 0;
            if constexpr (FIXED) {
#pragma unroll
                for (int k = 0; k < NPL; ++k) {
                    const uint8_t* p = rbytes + offI[k] + ao;
                    const int v = (int)p[0] * w.w00 + (int)p[1] * w.w01 + (int)p[rpitch] * w.w10 + (int)p[rpitch + 1] * w.w11;
                    const int diff = descale(v, kWBits - 5) - Iw[k];
                    if (lane + 32 * k < WW * WH) serr += abs(diff);
                }
            } else {
                for (int q = lane; q < npx; q += 32) {
                    const int y = q / ww, x = q - y * ww;
                    const uint8_t* p = rbytes + y * rpitch + x + ao;
                    const int v = (int)p[0] * w.w00 + (int)p[1] * w.w01 + (int)p[rpitch] * w.w10 + (int)p[rpitch + 1] * w.w11;
                    serr += abs(descale(v, kWBits - 5) - (int)pIw[q]);
                }
            }
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

template <int WW, int WH, int WARPS>
void launch_fixed(cudaStream_t st, const KltArgs& a, int batch)
{
    dim3 grid((a.n_max + WARPS - 1) / WARPS, batch);
    klt_kernel<WW, WH, WARPS><<<grid, WARPS * 32, 0, st>>>(a);
}

}  // namespace

void launch_klt(cudaStream_t st, const KltArgs& a, int batch, long long* launches)
{
    if (a.n_max <= 0) return;
    if (a.win_w == 21 && a.win_h == 21) launch_fixed<21, 21, 8>(st, a, batch);
    else if (a.win_w == 11 && a.win_h == 11) launch_fixed<11, 11, 8>(st, a, batch);
    else if (a.win_w == 31 && a.win_h == 31) launch_fixed<31, 31, 8>(st, a, batch);
    else {
        constexpr int WARPS = 4;
        const int ww = a.win_w, wh = a.win_h;
        const int words = (wh + 3) * ((ww + 9) >> 2) + (ww + 1) * (wh + 1) + (3 * ww * wh + 1) / 2;
        const size_t bytes = (size_t)WARPS * words * 4;
        static bool attr_set = false;
        if (!attr_set) {
            cudaFuncSetAttribute(klt_kernel<0, 0, WARPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
            attr_set = true;
        }
        dim3 grid((a.n_max + WARPS - 1) / WARPS, batch);
        klt_kernel<0, 0, WARPS><<<grid, WARPS * 32, bytes, st>>>(a);
    }
    if (launches) ++*launches;
}

}  // namespace svo
