// fast.cu -- K2: FAST-9/16 segment test + corner score + strict 3x3 NMS + pixel mask + raster-ordered compaction.
//
// Replaces mDetector->detect(frame->leftImg(), keypoints, mask) (reference src/tracking.cpp:82; detector created at
// src/tracking.cpp:54-57 as cv::FastFeatureDetector::create(threshold, nonMaxSuppression); defaults
// include/config_reader.h:35-38).  The arithmetic is OpenCV's FAST_t<16> + cornerScore<16> + KeyPointsFilter::
// runByPixelsMask (features2d, un-vendored), restated in SURVEY.md Appendix A.1/A.2 and oracle/svo_oracle.c:orc_fast:
//   d_k = I(p) - I(p + ring_k);  best = max over the 16 arcs of 9 contiguous k of max(min d_k, min -d_k)
//   corner iff best > t;  response = best - 1;  NMS keeps a corner iff its score is strictly greater than all 8 neighbours';
//   keypoints come out in raster order (y, then x); points whose mask pixel is 0 are dropped afterwards.
//
// Two kernels, both deterministic (bit-exact list equality needs ordered compaction, so no atomics decide positions):
//   fast_tile_kernel   one CTA per 128x32 tile: image tile + 4 px halo staged in shared memory, packed pre-test, segment test on the
//                      candidates over the (128+2)x(32+2) score region, dense pass over the surviving corners for the score, NMS, and
//                      a per-(row, tile) segment of kept keypoints written in x order via warp ballots.
//   fast_gather_kernel one CTA per image row: offset = keypoints in all earlier rows (sum of the per-row counters) +
//                      segments to the left; copies the segments to the final (x, y, response) list.
#include "svo_internal.cuh"

namespace svo {

namespace {

constexpr int TW = kFastTW, TH = kFastTH;
constexpr int RW = TW + 8, RH = TH + 8;   // staged image region (halo 3 for the ring + 1 for the NMS neighbours), origin (x0-4, y0-4)
constexpr int RWW = RW / 4;               // words per staged row
constexpr int SW = TW + 2, SH = TH + 2;   // score region, origin (x0-1, y0-1) = staged (3, 3)
constexpr int SWP = SW + 2;               // padded score row
constexpr int NG = RWW;                   // 4-pixel groups per score row (staged columns 4g .. 4g+3)

// ring offsets in the staged tile (row pitch RW), order of OpenCV's pixel[] table
__device__ __forceinline__ constexpr int ring_off(int k)
{
    constexpr int dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
    constexpr int dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
    return dy[k] * RW + dx[k];
}

// One multiply classifies a ring pixel v against both thresholds: with C = (0x8000 - lo) | ((hi + 0x8000) << 16),
// C + v * 0xFFFF0001 has bit 15 = (v >= lo) = "not darker" and bit 31 = (v <= hi) = "not brighter" (no carry between halves).
__device__ __forceinline__ uint32_t classify(uint32_t v, uint32_t C) { return C + v * 0xFFFF0001u; }

// true iff the 16-bit circular mask m has 9 contiguous set bits
__device__ __forceinline__ bool run9(uint32_t m)
{
    uint32_t dd = m | (m << 16);
    uint32_t a = dd & (dd >> 1);
    uint32_t b = a & (a >> 2);
    uint32_t c = b & (b >> 4);
    uint32_t d = c & (dd >> 8);
    return (d & 0xFFFFu) != 0;
}

__device__ __forceinline__ int min3(int a, int b, int c) { return min(min(a, b), c); }
__device__ __forceinline__ int max3(int a, int b, int c) { return max(max(a, b), c); }

// ---- packed pre-test: four horizontally adjacent pixels per thread, two per 32-bit register (16-bit fields).
// For a field holding ring value r and centre c:  (0x8000 - (c+t+1)) + r has bit 15 set iff r > c + t  (brighter),
//                                                  (0x8000 + c-t-1) - r has bit 15 set iff r < c - t  (darker);
// every field stays inside [0x7DFF, 0x80FE], so no carry or borrow crosses the fields.
struct PreAcc {
    uint32_t be, bo, de, dout;   // running AND over opposite pairs of (flag_k | flag_k+8): bright/dark x even/odd pixels
};

__device__ __forceinline__ void pre_pair(PreAcc& acc, uint32_t wa, uint32_t wb, uint32_t k1e, uint32_t k1o, uint32_t k2e,
                                         uint32_t k2o, bool first)
{
    const uint32_t ae = __byte_perm(wa, 0, 0x4240), ao = __byte_perm(wa, 0, 0x4341);   // bytes (0,2) / (1,3) as 16-bit fields
    const uint32_t be = __byte_perm(wb, 0, 0x4240), bo = __byte_perm(wb, 0, 0x4341);
    const uint32_t bre = (ae + k1e) | (be + k1e), bro = (ao + k1o) | (bo + k1o);
    const uint32_t dke = (k2e - ae) | (k2e - be), dko = (k2o - ao) | (k2o - bo);
    if (first) { acc.be = bre; acc.bo = bro; acc.de = dke; acc.dout = dko; }
    else { acc.be &= bre; acc.bo &= bro; acc.de &= dke; acc.dout &= dko; }
}

// cornerScore<16>: best = max over arcs of min(d) and min(-d).
// Both polarities are computed as minima (of d and of e = -d).  Writing the bright side as -max(d) made nvcc 12.9 emit
// VIMNMX3 with the negation dropped for sm_100a (wrong scores on the B200; caught by the parity tests) -- keep it this way.
__device__ __forceinline__ int corner_best(const uint8_t* __restrict__ p)
{
    const int c = p[0];
    int d[16], e[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const int v = p[ring_off(k)];
        d[k] = c - v;
        e[k] = v - c;
    }
    int d3[16], e3[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        d3[k] = min3(d[k], d[(k + 1) & 15], d[(k + 2) & 15]);
        e3[k] = min3(e[k], e[(k + 1) & 15], e[(k + 2) & 15]);
    }
    int best = -256;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const int d9 = min3(d3[k], d3[(k + 3) & 15], d3[(k + 6) & 15]);
        const int e9 = min3(e3[k], e3[(k + 3) & 15], e3[(k + 6) & 15]);
        best = max3(best, d9, e9);
    }
    return best;
}

__global__ void __launch_bounds__(256) fast_tile_kernel(const uint8_t* __restrict__ raw, size_t raw_pitch, size_t raw_lane,
                                                        int w, int h, int t, int nms, const uint8_t* __restrict__ mask,
                                                        size_t mask_pitch, size_t mask_lane, FastBufs fb, int ntx)
{
    __shared__ __align__(16) uint32_t simg_pad[RH * RWW + 2];   // one pad word on each side: the pre-test reads words g-1, g+1
    __shared__ uint8_t best[SH * SWP];
    __shared__ uint16_t cand[SH * SW];     // pixels that pass the packed pre-test
    __shared__ uint16_t corner[SH * SW];   // pixels that pass the full segment test
    __shared__ int ncand, ncorner;

    uint32_t* simg = simg_pad + 1;
    const uint8_t* img = reinterpret_cast<const uint8_t*>(simg);
    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * TW, y0 = blockIdx.y * TH;
    const int lane_id = blockIdx.z;
    const uint8_t* src = raw + (size_t)lane_id * raw_lane;

    if (tid == 0) { ncand = 0; ncorner = 0; simg_pad[0] = 0; simg_pad[RH * RWW + 1] = 0; }

    // ---- stage the (TW+8) x (TH+8) region; coordinates are clamped to the image (clamped pixels are never used by a
    //      valid candidate: those sit at least 3 px inside the image)
    const bool aligned = ((reinterpret_cast<uintptr_t>(src) | raw_pitch) & 3) == 0;
    for (int i = tid; i < RH * RWW; i += 256) {
        const int r = i / RWW, cw = i - r * RWW;
        const int gy = min(max(y0 - 4 + r, 0), h - 1);
        const int gx = x0 - 4 + cw * 4;
        const uint8_t* row = src + (size_t)gy * raw_pitch;
        uint32_t v;
        if (aligned && gx >= 0 && gx + 3 < w) {
            v = __ldg(reinterpret_cast<const uint32_t*>(row + gx));
        } else {
            v = 0;
#pragma unroll
            for (int b = 0; b < 4; ++b) v |= (uint32_t)__ldg(row + min(max(gx + b, 0), w - 1)) << (8 * b);
        }
        simg[i] = v;
    }
    for (int i = tid; i < SH * SWP; i += 256) best[i] = 0;
    __syncthreads();

    // ---- pass 0: packed pre-test on 4-pixel groups.  A 9-arc holds at least one pixel of every opposite pair (k, k+8), so
    //      a corner needs, for each of the pairs (0,8) (4,12) (2,10) (6,14), a brighter member (bright corner) or, for each,
    //      a darker member (dark corner).  About one pixel in ten survives on textured frames.
    const uint32_t t1 = (uint32_t)(t + 1) * 0x00010001u;
    for (int idx = tid; idx < SH * NG; idx += 256) {
        const int r = idx / NG, g = idx - r * NG;
        const uint32_t* c = simg + (r + 3) * RWW + g;   // centre row of score row r
        const uint32_t wc = c[0];
        const uint32_t ce = __byte_perm(wc, 0, 0x4240), co = __byte_perm(wc, 0, 0x4341);
        const uint32_t k1e = 0x80008000u - t1 - ce, k1o = 0x80008000u - t1 - co;   // + ring: bit 15 of a field = brighter
        const uint32_t k2e = 0x80008000u - t1 + ce, k2o = 0x80008000u - t1 + co;   // - ring: bit 15 of a field = darker
        PreAcc acc;
        pre_pair(acc, c[3 * RWW], c[-3 * RWW], k1e, k1o, k2e, k2o, true);                                    // (0,+3) (0,-3)
        uint32_t flag = ((acc.be | acc.de) | (acc.bo | acc.dout)) & 0x80008000u;
        if (flag) {
            pre_pair(acc, __funnelshift_r(c[0], c[1], 24), __funnelshift_r(c[-1], c[0], 8), k1e, k1o, k2e, k2o, false);   // (+3,0) (-3,0)
            pre_pair(acc, __funnelshift_r(c[2 * RWW], c[2 * RWW + 1], 16), __funnelshift_r(c[-2 * RWW - 1], c[-2 * RWW], 16),
                     k1e, k1o, k2e, k2o, false);                                                              // (+2,+2) (-2,-2)
            pre_pair(acc, __funnelshift_r(c[-2 * RWW], c[-2 * RWW + 1], 16), __funnelshift_r(c[2 * RWW - 1], c[2 * RWW], 16),
                     k1e, k1o, k2e, k2o, false);                                                              // (+2,-2) (-2,+2)
            const uint32_t fe = (acc.be | acc.de) & 0x80008000u, fo = (acc.bo | acc.dout) & 0x80008000u;
            // pixel i of the group: i = 0 -> fe bit 15, 1 -> fo bit 15, 2 -> fe bit 31, 3 -> fo bit 31
            const uint32_t bits = ((fe >> 15) & 1u) | ((fo >> 14) & 2u) | ((fe >> 29) & 4u) | ((fo >> 28) & 8u);
            const int gy = y0 - 1 + r;
            if (bits && gy >= 3 && gy <= h - 4) {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int sx = 4 * g + i - 3;
                    const int gx = x0 - 1 + sx;
                    if ((bits >> i & 1u) && sx >= 0 && sx < SW && gx >= 3 && gx <= w - 4)
                        cand[atomicAdd(&ncand, 1)] = (uint16_t)(r * SW + sx);
                }
            }
        }
    }
    __syncthreads();

    // ---- pass 1: full segment test, dense over the candidate list (order irrelevant: scores are per pixel)
    const uint32_t tt = (uint32_t)t;
    const int nc0 = ncand;
    for (int i = tid; i < nc0; i += 256) {
        const int p = cand[i];
        const int sy = p / SW, sx = p - sy * SW;
        const uint8_t* c = img + (sy + 3) * RW + sx + 3;
        const uint32_t cv = c[0];
        const uint32_t C = (0x8000u - (cv - tt)) + ((cv + tt + 0x8000u) << 16);
        uint32_t acc = 0;
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            const uint32_t f = classify(c[ring_off(k)], C);
            acc |= k == 15 ? (f & (0x00010001u << 15)) : ((f >> (15 - k)) & (0x00010001u << k));
        }
        const uint32_t inv = ~acc;   // low half: darker bits, high half: brighter bits
        if (run9(inv & 0xFFFFu) || run9(inv >> 16)) corner[atomicAdd(&ncorner, 1)] = (uint16_t)p;
    }
    __syncthreads();

    // ---- pass 2: corner score for the survivors (dense over the corner list)
    const int nc = ncorner;
    for (int i = tid; i < nc; i += 256) {
        const int p = corner[i];
        const int sy = p / SW, sx = p - sy * SW;
        best[sy * SWP + sx] = (uint8_t)corner_best(img + (sy + 3) * RW + sx + 3);   // 1..255 for a corner
    }
    __syncthreads();

    // ---- pass 3: NMS + mask, then one ordered segment per (row, tile): warp w owns rows w, w + 8, ...
    const int warp = tid >> 5, ln = tid & 31;
    const size_t seg_lane = (size_t)fb.rows_cap * fb.ntx_cap;
    const uint32_t lt = (1u << ln) - 1u;
    for (int r = warp; r < TH; r += 8) {
        const int gy = y0 + r;
        if (gy >= h) break;   // warp-uniform
        const size_t seg = (size_t)lane_id * seg_lane + (size_t)gy * ntx + blockIdx.x;
        uint16_t* out = fb.segkp + seg * TW;
        int total = 0;
#pragma unroll
        for (int q = 0; q < TW / 32; ++q) {
            const int x = ln + 32 * q;
            const uint8_t* b = best + (r + 1) * SWP + x + 1;
            const int s = b[0];
            bool keep = s != 0;
            if (keep && nms) {
                // stored value = best = score + 1 (0 = no corner); OpenCV compares scores, and a corner of score 0
                // (only possible at threshold 0) never beats a non-corner neighbour's 0: hence s > 1
                keep = s > 1 && s > b[-1] && s > b[1] && s > b[-SWP - 1] && s > b[-SWP] && s > b[-SWP + 1] && s > b[SWP - 1] &&
                       s > b[SWP] && s > b[SWP + 1];
            }
            if (keep && mask) keep = mask[(size_t)lane_id * mask_lane + (size_t)gy * mask_pitch + x0 + x] != 0;
            const uint32_t bal = __ballot_sync(0xffffffffu, keep);
            if (keep) out[total + __popc(bal & lt)] = (uint16_t)(x | (s << 8));
            total += __popc(bal);
        }
        if (ln == 0) {
            fb.segcnt[seg] = (uint8_t)total;
            if (total) atomicAdd(fb.rowtotal + (size_t)lane_id * fb.rows_cap + gy, total);
        }
    }
}

// grid = (h, batch), 128 threads
__global__ void __launch_bounds__(128) fast_gather_kernel(int w, int h, int nms, FastBufs fb, int ntx)
{
    __shared__ int red[4];
    __shared__ int segoff[1024];
    const int y = blockIdx.x, lane_id = blockIdx.y;
    const int tid = threadIdx.x, warp = tid >> 5, ln = tid & 31;
    const int* rt = fb.rowtotal + (size_t)lane_id * fb.rows_cap;

    int part = 0;
    for (int i = tid; i < y; i += 128) part += rt[i];
#pragma unroll
    for (int o = 16; o; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    if (ln == 0) red[warp] = part;
    __syncthreads();
    const int base = red[0] + red[1] + red[2] + red[3];
    const int mine = rt[y];
    if (y == h - 1 && tid == 0) fb.kp_count[lane_id] = base + mine;
    if (mine == 0) return;

    const size_t seg0 = (size_t)lane_id * fb.rows_cap * fb.ntx_cap + (size_t)y * ntx;
    if (warp == 0) {   // exclusive scan of the row's segment counts
        int carry = 0;
        for (int s0 = 0; s0 < ntx; s0 += 32) {
            const int s = s0 + ln;
            const int c = s < ntx ? (int)fb.segcnt[seg0 + s] : 0;
            int inc = c;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, inc, o);
                if (ln >= o) inc += v;
            }
            if (s < ntx) segoff[s] = carry + inc - c;
            carry += __shfl_sync(0xffffffffu, inc, 31);
        }
    }
    __syncthreads();

    float2* oxy = fb.kp_xy + (size_t)lane_id * fb.kp_cap;
    float* orp = fb.kp_resp + (size_t)lane_id * fb.kp_cap;
    for (int s = warp; s < ntx; s += 4) {
        const int c = fb.segcnt[seg0 + s];
        for (int j = ln; j < c; j += 32) {
            const uint32_t e = fb.segkp[(seg0 + s) * TW + j];
            const int idx = base + segoff[s] + j;
            if (idx < fb.kp_cap) {
                oxy[idx] = make_float2((float)(s * TW + (int)(e & 255u)), (float)y);
                orp[idx] = nms ? (float)((int)(e >> 8) - 1) : 0.f;
            }
        }
    }
}

// one warp per point: zero the rounded inclusive box (src/tracking.cpp:76-79)
__global__ void __launch_bounds__(128) mask_boxes_kernel(uint8_t* __restrict__ mask, size_t mask_pitch, size_t mask_lane,
                                                         int w, int h, const float2* __restrict__ pts, size_t pts_lane,
                                                         const int* __restrict__ n_dev, int n_max, float half)
{
    const int lane_id = blockIdx.y;
    const int i = blockIdx.x * 4 + (threadIdx.x >> 5);
    const int n = min(n_dev[lane_id], n_max);
    if (i >= n) return;
    const float2 p = pts[(size_t)lane_id * pts_lane + i];
    // float subtraction first, then cv::Point2f -> cv::Point (saturate_cast<int> = round half to even)
    int xa = __float2int_rn(__fsub_rn(p.x, half)), xb = __float2int_rn(__fadd_rn(p.x, half));
    int ya = __float2int_rn(__fsub_rn(p.y, half)), yb = __float2int_rn(__fadd_rn(p.y, half));
    xa = max(xa, 0); ya = max(ya, 0); xb = min(xb, w - 1); yb = min(yb, h - 1);
    uint8_t* m = mask + (size_t)lane_id * mask_lane;
    for (int y = ya; y <= yb; ++y)
        for (int x = xa + (int)(threadIdx.x & 31); x <= xb; x += 32) m[(size_t)y * mask_pitch + x] = 0;
}

}  // namespace

void launch_fast(cudaStream_t st, const uint8_t* raw, size_t raw_pitch, size_t raw_lane, int w, int h, int threshold,
                 int nms, const uint8_t* mask, size_t mask_pitch, size_t mask_lane, const FastBufs& fb, int batch,
                 long long* launches)
{
    threshold = threshold < 0 ? 0 : (threshold > 255 ? 255 : threshold);
    cudaMemsetAsync(fb.rowtotal, 0, sizeof(int) * (size_t)fb.rows_cap * batch, st);
    if (w < 7 || h < 7) {   // no candidate pixel at all (OpenCV returns an empty list)
        cudaMemsetAsync(fb.kp_count, 0, sizeof(int) * batch, st);
        return;
    }
    const int ntx = (w + TW - 1) / TW;
    dim3 grid(ntx, (h + TH - 1) / TH, batch);
    fast_tile_kernel<<<grid, 256, 0, st>>>(raw, raw_pitch, raw_lane, w, h, threshold, nms, mask, mask_pitch, mask_lane,
                                           fb, ntx);
    fast_gather_kernel<<<dim3(h, batch), 128, 0, st>>>(w, h, nms, fb, ntx);
    if (launches) *launches += 2;
}

void launch_build_mask(cudaStream_t st, uint8_t* mask, size_t mask_pitch, size_t mask_lane, int w, int h,
                       const float2* pts, size_t pts_lane, const int* n_dev, int n_max, float half, int batch,
                       long long* launches)
{
    for (int b = 0; b < batch; ++b)
        cudaMemset2DAsync(mask + (size_t)b * mask_lane, mask_pitch, 255, (size_t)w, (size_t)h, st);
    if (n_max <= 0) return;
    mask_boxes_kernel<<<dim3((n_max + 3) / 4, batch), 128, 0, st>>>(mask, mask_pitch, mask_lane, w, h, pts, pts_lane,
                                                                  n_dev, n_max, half);
    if (launches) ++*launches;
}

}  // namespace svo
