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
//   fast_gather_kernel one warp per image row: offset = keypoints in all earlier rows (sum of the per-row counters) +
//                      segments to the left; copies the segments to the final (x, y, response) list.
#include <type_traits>

#include "svo_internal.cuh"

namespace svo {

namespace {

constexpr int TW = kFastTW, TH = kFastTH;
constexpr int RW = TW + 8, RH = TH + 8;   // staged image region (halo 3 for the ring + 1 for the NMS neighbours), origin (x0-4, y0-4)
constexpr int RWW = RW / 4;               // words per staged row
constexpr int SH = TH + 2;   // score region, origin (x0-1, y0-1) = staged (3, 3)
constexpr int NG = RWW;                   // 4-pixel groups per score row (staged columns 4g .. 4g+3)
constexpr int MX = 32, MY = 2;            // margin of the tile's mask bit plane (columns, rows)
constexpr int MW = TW / 32 + 2, MH = TH + 2 * MY;   // its size in words per row and rows
static_assert(SH * NG < 65536 && TW % 32 == 0, "group indices are stored as 16 bits");

// ---- packed arithmetic: two pixels per 32-bit register (16-bit fields); a 4-pixel group = an "even" register (pixels
// 0, 2) and an "odd" one (pixels 1, 3).  sm_100a has 3-input packed min/max (VIMNMX3.U16x2), which ptxas forms from
// nested min.u16x2 / max.u16x2.
__device__ __forceinline__ uint32_t pmin(uint32_t a, uint32_t b)
{
    uint32_t d;
    asm("min.u16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
    return d;
}
__device__ __forceinline__ uint32_t pmax(uint32_t a, uint32_t b)
{
    uint32_t d;
    asm("max.u16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
    return d;
}
__device__ __forceinline__ uint32_t pmin3(uint32_t a, uint32_t b, uint32_t c) { return pmin(pmin(a, b), c); }
__device__ __forceinline__ uint32_t pmax3(uint32_t a, uint32_t b, uint32_t c) { return pmax(pmax(a, b), c); }
__device__ __forceinline__ uint32_t even_px(uint32_t w) { return __byte_perm(w, 0, 0x4240); }   // bytes 0, 2 -> fields
__device__ __forceinline__ uint32_t odd_px(uint32_t w) { return __byte_perm(w, 0, 0x4341); }    // bytes 1, 3 -> fields

// the staged word holding bytes (4g + dx .. 4g + dx + 3) of the row at c (c points at word g)
template <int DX>
__device__ __forceinline__ uint32_t ring_word(const uint32_t* __restrict__ c)
{
    if constexpr (DX == 0) return c[0];
    else if constexpr (DX > 0) return __funnelshift_r(c[0], c[1], 8 * DX);
    else return __funnelshift_r(c[-1], c[0], 32 + 8 * DX);
}

// cornerScore<16> for two pixels at once.  D[k] = centre + 256 - ring_k (1..511 per field).  With d = D - 256:
//   best = max over the 16 arcs of 9 of max(min d, min -d) = max(max_arcs min9 D, 512 - min_arcs max9 D) - 256.
// Returns best + 256 per field.
__device__ __forceinline__ uint32_t packed_best(const uint32_t (&D)[16])
{
    uint32_t mn3[16], mx3[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        mn3[k] = pmin3(D[k], D[(k + 1) & 15], D[(k + 2) & 15]);
        mx3[k] = pmax3(D[k], D[(k + 1) & 15], D[(k + 2) & 15]);
    }
    uint32_t hi = 0u, lo = 0xffffffffu;
#pragma unroll
    for (int k = 0; k < 16; k += 2) {
        const uint32_t a = pmin3(mn3[k], mn3[(k + 3) & 15], mn3[(k + 6) & 15]);
        const uint32_t b = pmin3(mn3[k + 1], mn3[(k + 4) & 15], mn3[(k + 7) & 15]);
        hi = pmax3(hi, a, b);
        const uint32_t c = pmax3(mx3[k], mx3[(k + 3) & 15], mx3[(k + 6) & 15]);
        const uint32_t d = pmax3(mx3[k + 1], mx3[(k + 4) & 15], mx3[(k + 7) & 15]);
        lo = pmin3(lo, c, d);
    }
    return pmax(hi, 0x02000200u - lo);
}

// scalar cornerScore<16> kept for reference in comments of the tests: see oracle/svo_oracle.c:orc_corner_score

// NT threads per CTA: 128 (8 192 registers per CTA, eight CTAs per SM) or 256
template <int NT>
__global__ void __launch_bounds__(NT, NT == 128 ? 8 : 1) fast_tile_kernel(const uint8_t* __restrict__ raw, size_t raw_pitch, size_t raw_lane,
                                                       int w, int h, int t, int nms, const uint8_t* __restrict__ mask,
                                                       size_t mask_pitch, size_t mask_lane, FastBufs fb, int ntx,
                                                       FastMaskPts mp)
{
    if (mp.lane_active && mp.lane_active[blockIdx.z] == 0) return;   // this lane detects nothing in this call (CTA-uniform)
    __shared__ __align__(16) uint32_t simg_pad[RH * RWW + 2];   // one pad word on each side: ring_word reads words g-1, g+1
    __shared__ __align__(16) uint32_t sbest_pad[SH * RWW + 2];  // excess of the score over the threshold, one byte per pixel,
                                                                // same column layout as the staged image (0 = no corner)
    // mask given as points (src/tracking.cpp:76-79), as a bit plane over the tile and a margin of 32 columns / 2 rows around
    // it: bit (x - (x0 - 32), y - (y0 - 2)) = mask pixel (x, y) != 0.  Bits outside the frame stay set.
    __shared__ uint32_t smask[MH * MW];
    __shared__ uint16_t slist[SH * NG];   // the 4-pixel groups of the score region that need a score (see below)
    __shared__ int scount;
    uint32_t* simg = simg_pad + 1;
    uint32_t* sbest = sbest_pad + 1;
    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * TW, y0 = blockIdx.y * TH;
    const int lane_id = blockIdx.z;
    const uint8_t* src = raw + (size_t)lane_id * raw_lane;

    if (tid == 0) { simg_pad[0] = 0; simg_pad[RH * RWW + 1] = 0; sbest_pad[0] = 0; sbest_pad[SH * RWW + 1] = 0; }
    if (mp.pts) {
        for (int i = tid; i < MH * MW; i += NT) smask[i] = 0xffffffffu;
        if (tid == 0) scount = 0;
    }

    // ---- stage the (TW+8) x (TH+8) region; coordinates are clamped to the image (clamped pixels are never used by a
    //      valid candidate: those sit at least 3 px inside the image)
    //      Four words per thread and round, every load of the round issued before the first use (the kernel's first phase is
    //      all latency: one CTA has nothing else to run).  Words at the frame border (rare) are assembled byte by byte.
    constexpr int SU = 4;
    for (int i0 = tid; i0 < RH * RWW; i0 += SU * NT) {
        uint32_t lo[SU], hi[SU];
        int sh[SU];
#pragma unroll
        for (int k = 0; k < SU; ++k) {
            const int i = min(i0 + k * NT, RH * RWW - 1);
            const int r = i / RWW, cw = i - r * RWW;
            const int gy = min(max(y0 - 4 + r, 0), h - 1);
            const int gx = x0 - 4 + cw * 4;
            const uint8_t* row = src + (size_t)gy * raw_pitch;
            if (gx >= 4 && gx + 7 < w) {
                // two aligned words + funnel shift: any row pitch / alignment; the eight bytes read lie inside this row
                const uintptr_t addr = reinterpret_cast<uintptr_t>(row + gx);
                const uint32_t* wp = reinterpret_cast<const uint32_t*>(addr & ~(uintptr_t)3);
                lo[k] = __ldg(wp); hi[k] = __ldg(wp + 1);
                sh[k] = 8 * (int)(addr & 3);
            } else {
                uint32_t v = 0;
#pragma unroll
                for (int b = 0; b < 4; ++b) v |= (uint32_t)__ldg(row + min(max(gx + b, 0), w - 1)) << (8 * b);
                lo[k] = v; hi[k] = 0; sh[k] = 0;
            }
        }
#pragma unroll
        for (int k = 0; k < SU; ++k)
            if (i0 + k * NT < RH * RWW) simg[i0 + k * NT] = __funnelshift_r(lo[k], hi[k], sh[k]);
    }
    __syncthreads();

    // ---- the tile's part of the reference's mask, straight from the points it is drawn around: every thread walks its share
    //      of the lane's points and clears the bits of the boxes that reach into this tile (the full-frame mask plane,
    //      its 255-fill and the box kernel are never materialised).  Box = [cvRound(x - half), cvRound(x + half)] x
    //      [cvRound(y - half), cvRound(y + half)], float subtraction first, round half to even (cv::Point2f -> cv::Point).
    if (mp.pts) {
        const int n = min(mp.n_dev[lane_id], mp.n_max);
        const float2* pts = mp.pts + (size_t)lane_id * mp.pts_lane;
        // most points are nowhere near this tile: one float comparison per axis (with a pixel of slack for the rounding below)
        // turns them away before any float -> int conversion (quarter-rate pipe); NaN coordinates fall through to the exact test
        const float cy = (float)(y0 - MY) + 0.5f * (float)(MH - 1), ry = fabsf(mp.half) + 0.5f * (float)(MH - 1) + 1.f;
        const float cx = (float)(x0 - MX) + 0.5f * (float)(32 * MW - 1), rx = fabsf(mp.half) + 0.5f * (float)(32 * MW - 1) + 1.f;
        constexpr int PU = 8;   // points per thread and round, all loaded before the first is looked at
        for (int i0 = tid; i0 < n; i0 += PU * NT) {
          float2 pk[PU];
#pragma unroll
          for (int k = 0; k < PU; ++k) pk[k] = __ldg(pts + min(i0 + k * NT, n - 1));
#pragma unroll
          for (int k = 0; k < PU; ++k) {
            if (i0 + k * NT >= n) break;
            const float2 p = pk[k];
            if (fabsf(p.y - cy) > ry || fabsf(p.x - cx) > rx) continue;
            // clamped to the plane before its origin is subtracted (the conversions saturate at INT_MIN / INT_MAX)
            const int my0 = y0 - MY, mx0 = x0 - MX;
            const int ya = max(__float2int_rn(__fsub_rn(p.y, mp.half)), my0) - my0, yb = min(__float2int_rn(__fadd_rn(p.y, mp.half)), my0 + MH - 1) - my0;
            if (ya > yb) continue;
            const int xa = max(__float2int_rn(__fsub_rn(p.x, mp.half)), mx0) - mx0, xb = min(__float2int_rn(__fadd_rn(p.x, mp.half)), mx0 + 32 * MW - 1) - mx0;
            if (xa > xb) continue;
            for (int wd = xa >> 5; wd <= xb >> 5; ++wd) {
                const int lo = max(xa - 32 * wd, 0), hi = min(xb - 32 * wd, 31);
                const uint32_t clr = ~((0xffffffffu >> (31 - hi)) & (0xffffffffu << lo));
                for (int y = ya; y <= yb; ++y) atomicAnd(&smask[y * MW + wd], clr);
            }
          }
        }
    }

    // ---- which groups need a score at all.  The reference masks two thirds of the frame (2 000 boxes of 21 x 21 px), and a
    //      pixel's score can only reach the output as its own response (if it survives the mask) or through the NMS of one of
    //      its 8 neighbours (if THAT one survives the mask): a pixel that is masked together with all its neighbours needs no
    //      score.  A group of four such pixels (its 6 x 3 neighbourhood fully masked: ~60 % of the groups on the bench frames)
    //      gets 0 and the rest are compacted into a list, so that the warps of pass 1 stay full.
    const bool use_list = mp.pts != nullptr;
    int ngroups = SH * NG;
    if (use_list) {
        __syncthreads();   // the bit plane is complete
        const uint32_t lt_ = (1u << (tid & 31)) - 1u;
        int r = tid / NG, g = tid - r * NG;   // (row, group) of idx, advanced without divisions
        for (int base = 0; base < SH * NG; base += NT, r += NT / NG, g += NT % NG) {
            const int idx = base + tid;
            if (g >= NG) { g -= NG; ++r; }
            bool work = false;
            if (idx < SH * NG) {
                // score row r = image row y0 - 1 + r -> plane rows r .. r + 2; group g = image columns x0 - 4 + 4 g .. + 3,
                // neighbourhood from x0 - 5 + 4 g -> plane bit 4 g + MX - 5, six bits
                const int bx = 4 * g + MX - 5;
                const uint32_t* m0 = smask + r * MW + (bx >> 5);
                const uint32_t bits = __funnelshift_r(m0[0], m0[1], bx & 31) | __funnelshift_r(m0[MW], m0[MW + 1], bx & 31) |
                                      __funnelshift_r(m0[2 * MW], m0[2 * MW + 1], bx & 31);
                work = (bits & 0x3fu) != 0;
                if (!work) sbest[r * RWW + g] = 0;
            }
            const unsigned m = __ballot_sync(0xffffffffu, work);
            int wbase = 0;
            if ((tid & 31) == 0 && m) wbase = atomicAdd(&scount, __popc(m));
            wbase = __shfl_sync(0xffffffffu, wbase, 0);
            if (work) slist[wbase + __popc(m & lt_)] = (uint16_t)idx;
        }
        __syncthreads();
        ngroups = scount;
    }

    // ---- pass 1: dense packed corner score of every 4-pixel group of the score region (rows y0-1 .. y0+TH, staged
    //      columns 0 .. RW-1): sbest byte = max(best - t, 0), where best = cornerScore + 1.
    // does the score region (x0-4 .. x0+TW+3, y0-1 .. y0+TH) reach into the 3-px frame?  (CTA-uniform)
    const bool edge_tile = x0 - 4 < 3 || x0 + TW + 3 > w - 4 || y0 - 1 < 3 || y0 + TH > h - 4;
    const uint32_t bias = 0x01000100u;
    const uint32_t tt = (uint32_t)t * 0x00010001u;
    for (int it = tid; it < ngroups; it += NT) {
        const int idx = use_list ? (int)slist[it] : it;
        const int r = idx / NG, g = idx - r * NG;
        const uint32_t* c = simg + (r + 3) * RWW + g;   // centre row of score row r
        const uint32_t wc = c[0];
        const uint32_t be = even_px(wc) + bias, bo = odd_px(wc) + bias;   // D = b - ring
        uint32_t De[16], Do[16];
        auto ring = [&](auto kk, auto dxx, auto dyy) {
            constexpr int K = decltype(kk)::value, DX = decltype(dxx)::value, DY = decltype(dyy)::value;
            const uint32_t v = ring_word<DX>(c + DY * RWW);
            De[K] = be - even_px(v);
            Do[K] = bo - odd_px(v);
        };
#define RING(K, DX, DY) ring(std::integral_constant<int, K>{}, std::integral_constant<int, DX>{}, std::integral_constant<int, DY>{})
        // the compass points first: a 9-arc holds a pixel of every opposite pair, so a corner needs a darker member in both
        // (0,8) and (4,12), or a brighter member in both.  darker: ring < c - t <=> D >= 257 + t; brighter: D <= 255 - t.
        RING(0, 0, 3); RING(8, 0, -3); RING(4, 3, 0); RING(12, -3, 0);
        {
            const uint32_t kd = 0x7eff7effu - tt;   // D + kd: bit 15 of a field set <=> D >= 257 + t
            const uint32_t kb = 0x80ff80ffu - tt;   // kb - D: bit 15 of a field set <=> D <= 255 - t
            const uint32_t fe = (((De[0] + kd) | (De[8] + kd)) & ((De[4] + kd) | (De[12] + kd))) |
                                (((kb - De[0]) | (kb - De[8])) & ((kb - De[4]) | (kb - De[12])));
            const uint32_t fo = (((Do[0] + kd) | (Do[8] + kd)) & ((Do[4] + kd) | (Do[12] + kd))) |
                                (((kb - Do[0]) | (kb - Do[8])) & ((kb - Do[4]) | (kb - Do[12])));
            if (((fe | fo) & 0x80008000u) == 0) {   // no pixel of the group can be a corner
                sbest[r * RWW + g] = 0;
                continue;
            }
        }
        RING(1, 1, 3); RING(2, 2, 2); RING(3, 3, 1); RING(5, 3, -1); RING(6, 2, -2); RING(7, 1, -3);
        RING(9, -1, -3); RING(10, -2, -2); RING(11, -3, -1); RING(13, -3, 1); RING(14, -2, 2); RING(15, -1, 3);
#undef RING
        const uint32_t floor_ = 0x01000100u + tt;                     // best + 256 <= 256 + t  ->  not a corner
        const uint32_t se = pmax(packed_best(De), floor_) - floor_;   // best - t, 0 if not a corner
        const uint32_t so = pmax(packed_best(Do), floor_) - floor_;
        uint32_t out = __byte_perm(se, so, 0x6240);                   // bytes: pixel 0, 1, 2, 3
        // pixels outside the 3-px frame of the image are not candidates (only tiles at the image edge hold any)
        if (edge_tile) {
            const int gy = y0 - 1 + r, gx = x0 - 4 + 4 * g;
            if (gy < 3 || gy > h - 4) out = 0;
            else if (gx < 3 || gx + 3 > w - 4) {
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    if (gx + i < 3 || gx + i > w - 4) out &= ~(0xffu << (8 * i));
            }
        }
        sbest[r * RWW + g] = out;
    }
    __syncthreads();

    // ---- pass 2: strict 3x3 NMS + mask, then one ordered segment per (row, tile): a warp owns a row, a lane a 4-pixel group
    const int warp = tid >> 5, ln = tid & 31;
    const size_t seg_lane = (size_t)fb.rows_cap * fb.ntx_cap;
    const uint32_t lt = (1u << ln) - 1u;
    // OpenCV compares scores (= best - 1); a corner of score 0 (threshold 0 only) never beats a non-corner's 0
    const uint32_t nms_floor = (nms && t == 0) ? 0x00010001u : 0u;
    for (int r = warp; r < TH; r += NT / 32) {
        const int gy = y0 + r;
        if (gy >= h) break;   // warp-uniform
        const uint32_t* b = sbest + (r + 1) * RWW + ln + 1;   // tile pixel x = 4 ln + i sits in staged group ln + 1
        const uint32_t wc = b[0];
        uint32_t bits = 0;
        if (wc) {
            uint32_t me = nms_floor, mo = nms_floor;
            if (nms) {
#pragma unroll
                for (int dr = -1; dr <= 1; ++dr) {
                    const uint32_t* q = b + dr * RWW;
                    const uint32_t L = __funnelshift_r(q[-1], q[0], 24), R = __funnelshift_r(q[0], q[1], 8);
                    if (dr == 0) { me = pmax3(me, even_px(L), even_px(R)); mo = pmax3(mo, odd_px(L), odd_px(R)); }
                    else {
                        me = pmax(me, pmax3(even_px(L), even_px(q[0]), even_px(R)));
                        mo = pmax(mo, pmax3(odd_px(L), odd_px(q[0]), odd_px(R)));
                    }
                }
            }
            // keep <=> s > m <=> (s + 0x7fff) - m has bit 15 set (fields stay inside [0x7f00, 0x80fe])
            const uint32_t ke = (even_px(wc) + 0x7fff7fffu) - me, ko = (odd_px(wc) + 0x7fff7fffu) - mo;
            bits = ((ke >> 15) & 1u) | ((ko >> 14) & 2u) | ((ke >> 29) & 4u) | ((ko >> 28) & 8u);
            if (bits && mp.pts) bits &= smask[(r + MY) * MW + MX / 32 + (ln >> 3)] >> (4 * (ln & 7));
            if (bits && mask) {
                const uint8_t* mrow = mask + (size_t)lane_id * mask_lane + (size_t)gy * mask_pitch + x0 + 4 * ln;
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    if ((bits >> i & 1u) && mrow[i] == 0) bits &= ~(1u << i);
            }
        }
        const uint32_t b0 = __ballot_sync(0xffffffffu, bits & 1u), b1 = __ballot_sync(0xffffffffu, bits & 2u);
        const uint32_t b2 = __ballot_sync(0xffffffffu, bits & 4u), b3 = __ballot_sync(0xffffffffu, bits & 8u);
        const int total = __popc(b0) + __popc(b1) + __popc(b2) + __popc(b3);
        const size_t seg = (size_t)lane_id * seg_lane + (size_t)gy * ntx + blockIdx.x;
        if (bits) {   // mostly one keypoint per group: walk the set bits instead of testing all four
            uint16_t* out = fb.segkp + seg * TW + __popc(b0 & lt) + __popc(b1 & lt) + __popc(b2 & lt) + __popc(b3 & lt);
            for (uint32_t rest = bits; rest; rest &= rest - 1u) {
                const int i = __ffs((int)rest) - 1;
                *out++ = (uint16_t)((4 * ln + i) | (((wc >> (8 * i)) & 0xffu) << 8));
            }
        }
        if (ln == 0) {
            fb.segcnt[seg] = (uint8_t)total;
            if (total) atomicAdd(fb.rowtotal + (size_t)lane_id * fb.rows_cap + gy, total);
        }
    }
}

// grid = (ceil(h / 8), batch), 256 threads: a warp owns an image row and never meets the other warps (no barrier, no shared
// memory).  Row offset = keypoints in all earlier rows; segment offsets by a warp scan of the row's per-tile counts.
constexpr int kGatherRows = 8;
__global__ void __launch_bounds__(kGatherRows * 32) fast_gather_kernel(int w, int h, int t, int nms, FastBufs fb, int ntx)
{
    const int ln = threadIdx.x & 31;
    const int y = blockIdx.x * kGatherRows + (threadIdx.x >> 5), lane_id = blockIdx.y;
    if (y >= h) return;
    const int* rt = fb.rowtotal + (size_t)lane_id * fb.rows_cap;

    int base = 0;
    for (int i = ln; i < y; i += 32) base += rt[i];
#pragma unroll
    for (int o = 16; o; o >>= 1) base += __shfl_xor_sync(0xffffffffu, base, o);
    const int mine = rt[y];
    if (y == h - 1 && ln == 0) fb.kp_count[lane_id] = base + mine;
    if (mine == 0) return;

    const size_t seg0 = (size_t)lane_id * fb.rows_cap * fb.ntx_cap + (size_t)y * ntx;
    float2* oxy = fb.kp_xy + (size_t)lane_id * fb.kp_cap;
    float* orp = fb.kp_resp + (size_t)lane_id * fb.kp_cap;
    const float fy = (float)y;
    for (int s0 = 0; s0 < ntx; s0 += 32) {
        const int s = s0 + ln;
        const int c = s < ntx ? (int)fb.segcnt[seg0 + s] : 0;
        int inc = c;   // inclusive scan of the counts of these 32 segments
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, inc, o);
            if (ln >= o) inc += v;
        }
        // the entries of these segments, flattened: output position p of the chunk lies in the first segment whose
        // inclusive count exceeds p (binary search over the lanes' scan values), so all loads of a round are independent
        const int tot = __shfl_sync(0xffffffffu, inc, 31);
        for (int p0 = 0; p0 < tot; p0 += 32) {
            const int p = p0 + ln;
            int lo = 0, hi = 31;
#pragma unroll
            for (int it = 0; it < 5; ++it) {
                const int mid = (lo + hi) >> 1;
                if (__shfl_sync(0xffffffffu, inc, mid) > p) hi = mid;
                else lo = mid + 1;
            }
            const int j = p - (__shfl_sync(0xffffffffu, inc, lo) - __shfl_sync(0xffffffffu, c, lo));
            if (p < tot) {
                const uint32_t e = fb.segkp[(seg0 + s0 + lo) * TW + j];
                const int idx = base + p;
                if (idx < fb.kp_cap) {
                    oxy[idx] = make_float2((float)((s0 + lo) * TW + (int)(e & 255u)), fy);
                    orp[idx] = nms ? (float)((int)(e >> 8) + t - 1) : 0.f;   // stored byte = best - t, score = best - 1
                }
            }
        }
        base += __shfl_sync(0xffffffffu, inc, 31);
    }
}

}  // namespace

void launch_fast(cudaStream_t st, const uint8_t* raw, size_t raw_pitch, size_t raw_lane, int w, int h, int threshold,
                 int nms, const uint8_t* mask, size_t mask_pitch, size_t mask_lane, const FastBufs& fb, int batch,
                 long long* launches, int small_ctas, const FastMaskPts* mask_pts)
{
    FastMaskPts mp = mask_pts ? *mask_pts : FastMaskPts{};
    if (!mp.pts || mp.n_max <= 0) { mp.pts = nullptr; mp.n_max = 0; }
    threshold = threshold < 0 ? 0 : (threshold > 255 ? 255 : threshold);
    cudaMemsetAsync(fb.rowtotal, 0, sizeof(int) * (size_t)fb.rows_cap * batch, st);
    if (w < 7 || h < 7) {   // no candidate pixel at all (OpenCV returns an empty list)
        cudaMemsetAsync(fb.kp_count, 0, sizeof(int) * batch, st);
        return;
    }
    static std::atomic<unsigned long long> seen{0};
    if (first_on_device(seen)) {
        prefer_max_shared(fast_tile_kernel<128>); prefer_max_shared(fast_tile_kernel<256>); prefer_max_shared(fast_tile_kernel<512>);
        prefer_max_shared(fast_gather_kernel);
    }
    const int ntx = (w + TW - 1) / TW;
    dim3 grid(ntx, (h + TH - 1) / TH, batch);
    // CTA size: 128 threads when the tiles fill the device several times over (measured 2 % faster than 256, and 64 slower:
    // gpurun_out/s11, s12); with few tiles (one stream = 120) the kernel's time is one CTA's, so the tile is spread over 512
    // threads instead (small_ctas: 1 = 128, 0 = 256, 2 = 512, -1 = by tile count)
    if (small_ctas < 0) {
        int dev = 0, sms = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        small_ctas = (long long)grid.x * grid.y * grid.z <= 2LL * sms ? 2 : 1;
    }
    if (small_ctas == 1)
        fast_tile_kernel<128><<<grid, 128, 0, st>>>(raw, raw_pitch, raw_lane, w, h, threshold, nms, mask, mask_pitch, mask_lane,
                                                    fb, ntx, mp);
    else if (small_ctas == 2)
        fast_tile_kernel<512><<<grid, 512, 0, st>>>(raw, raw_pitch, raw_lane, w, h, threshold, nms, mask, mask_pitch, mask_lane,
                                                    fb, ntx, mp);
    else
        fast_tile_kernel<256><<<grid, 256, 0, st>>>(raw, raw_pitch, raw_lane, w, h, threshold, nms, mask, mask_pitch, mask_lane,
                                                    fb, ntx, mp);
    fast_gather_kernel<<<dim3((h + kGatherRows - 1) / kGatherRows, batch), kGatherRows * 32, 0, st>>>(w, h, threshold, nms, fb, ntx);
    if (launches) *launches += 2;
}

}  // namespace svo
