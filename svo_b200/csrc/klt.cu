// klt.cu -- K4: pyramidal Lucas-Kanade for the compiled windows (21x21, 11x11, 31x31), one warp per feature, all
// pyramid levels in one launch.  Other window sizes take klt_generic.cu.
//
// Replaces cv::calcOpticalFlowPyrLK at the reference's two call sites:
//   temporal tracking  src/tracking.cpp:160-165  (21x21, maxLevel 3, {COUNT+EPS, 50, 0.001}, OPTFLOW_LK_GET_MIN_EIGENVALS)
//   stereo matching    src/tracking.cpp:101-105  (11x11, maxLevel 3, {COUNT+EPS, 30, 0.001}, flags 0 -> L1 residual err)
// The arithmetic is OpenCV's LKTrackerInvoker (video/src/lkpyramid.cpp, un-vendored), restated in SURVEY.md Appendix
// A.3-A.5 and oracle/svo_oracle.c:orc_klt_ex (accum_mode 0), which this kernel matches bit for bit: 14-bit integer
// bilinear weights, DESCALE roundings, window sums accumulated exactly in integers, float32 epilogue in OpenCV's
// operation order (-fmad=false), status cleared only by level-0 outcomes, results not clipped to the image.
//
// Design for sm_100a.  The kernel is issue-bound (its footprint, two pyramids, sits in L2), so the layout is chosen to
// minimise instructions per window pixel:
//   - a window row is cut into RUNS of 8 pixels; a lane owns RPL runs.  The 9 source bytes a run needs from each of its
//     two image rows arrive as one 64-bit + one 32-bit shared-memory load, and the bilinear sample of a pixel is two
//     DP2A instructions (u16 weight pair x u8 pixel pair, accumulated), instead of four byte loads and four multiply-adds.
//   - source regions are staged per warp in shared memory ALREADY SHIFTED to the window's first column (aligned 32-bit
//     global loads + funnel shift), rows 8*NSEG bytes apart, so run r of the window starts at byte 8 r: the lanes'
//     64-bit loads are contiguous, hence bank-conflict free, and no per-pixel address arithmetic remains.
//   - the patch (I*32, Ix, Iy) lives in registers for the whole level.
//   - the reference's int16x2 Scharr derivative images are never materialised: bilinear interpolation and the Scharr
//     stencil are both integer-linear, so sum_ij w_ij Scharr(I)(p+ij) == Scharr(B)(p) with B the UNROUNDED bilinear plane
//     ((w+2)x(h+2) int32, built with the same run/DP2A scheme); a lane reads the 3x10 plane values of a run with 128-bit
//     loads and forms Ix, Iy separably.  Windows that touch the image border (where OpenCV's derivative image is
//     ZERO-padded, which breaks that identity) take the literal path: Scharr plane with zeros outside, then bilinear.
//   - lanes never diverge: every per-feature scalar is warp-uniform; reductions are REDUX on 16-bit halves (exact).
#include <stdio.h>

#include <algorithm>
#include <type_traits>

#include "klt_common.cuh"

// epilogue shortcuts of the Newton loop, individually switchable for the variant sweeps (scripts/build_variants.sh):
//   1 scaled warp sum (one FMA instead of add + multiply)      2 float pre-test in front of the double convergence test
//   4 weight pairs packed by PRMT                               8 level-bounds test only when the window leaves the staged region
// Measured (gpurun_out/s5): 1 and 4 each save ~1 %, 2 and 8 are slower (extra branch / registers) and stay off.
#ifndef KLT_OPTS
#define KLT_OPTS 5
#endif

namespace svo {

namespace {

using namespace lk;

__device__ __forceinline__ float warp_sum_scaled(int v)
{
    if constexpr (KLT_OPTS & 1) return warp_sum_exact_scaled(v);
    else return __fmul_rn(warp_sum_exact_f(v), 1.f / (1 << 20));
}

template <int WW, int WH>
struct RunGeo {
    static_assert(WW % 8 != 0 && (WW + 2) % 8 != 0, "window widths that are 0 or 6 mod 8 take the generic kernel");
    static constexpr int NSEG = WW / 8 + 1;            // 8-pixel runs per window row (the last one partial)
    static constexpr int RUNS = WH * NSEG;
    static constexpr int RPL = (RUNS + 31) / 32;       // runs per lane
    // next-image region: the (WW+1) x (WH+1) window plus a margin of MJ pixels on every side, staged UNSHIFTED (word
    // aligned) once per level; the Newton iterations move the window inside it and only re-stage when they leave it
    static constexpr int MJ = 3;
    // region rows start at a word boundary of the image row, so the window's first column sits MJ .. MJ + 3 bytes into a
    // region row; a run reads three whole words from the word holding its first byte.  OWJ words of every row are copied;
    // rows are PJW words apart, PJW picked (by enumeration) so that the lanes' 32-bit loads of one instruction -- word
    // index PJW y + 2 seg for the run (y, seg) -- spread over the banks: 21x21 3 wavefronts for the two rounds of 32 runs
    // (a 12-word pitch needs 6), 11x11 1, 31x31 8 for four rounds
    static constexpr int OWJ = ((2 * MJ + 3) & ~3) / 4 + 2 * (NSEG - 1) + 3;
    static constexpr int PJW = NSEG == 3 ? 19 : NSEG == 2 ? 9 : NSEG == 4 ? 13 : (OWJ | 1);
    static_assert(PJW >= OWJ, "row pitch shorter than the copied words");
    static constexpr int PJ = 4 * PJW;                                  // bytes per region row
    static constexpr int DXMAX = 4 * OWJ - 1 - WW, DYMAX = 2 * MJ;      // legal window offsets inside the region
    static constexpr int JROWS = WH + 1 + 2 * MJ;
    static constexpr int RPPJ = 32 / OWJ;                               // rows per staging pass
    static constexpr int PASSJ = (JROWS + RPPJ - 1) / RPPJ;
    static constexpr int JREGION_WORDS = (PASSJ * RPPJ * PJW + 4 + 3) & ~3;
    static constexpr int NSEGB = (WW + 2) / 8 + 1;     // runs per row of the B plane (WW + 2 values per row)
    static constexpr int RPI = 8 * NSEGB;              // bytes per staged row of the previous image (needs WW + 3)
    static constexpr int IROWS = WH + 3;
    // a plane run (8 values) occupies a block of 12 words: lanes 48 bytes apart make every 128-bit access of a warp
    // conflict-free (8 consecutive lanes cover the 32 banks), where 32-byte blocks would only ever touch half the banks
    static constexpr int BBLK = 12;
    static constexpr int BP = BBLK * NSEGB;            // plane pitch in words
    static constexpr int PROWS = WH + 2;
    static constexpr int RUNSB = PROWS * NSEGB;
    static constexpr int RPLB = (RUNSB + 31) / 32;
    // staging works in whole passes of 32 / (RP/4) rows (no predicates): the region holds the rounded-up row counts
    static constexpr int IROWS_ST = (IROWS + 32 / (RPI / 4) - 1) / (32 / (RPI / 4)) * (32 / (RPI / 4));
    static constexpr int REGION_WORDS = (IROWS_ST * RPI / 4 + 4 + 3) & ~3;
    static constexpr int PLANE_WORDS = (PROWS * BP + 8 + 3) & ~3;
    // border windows: packed (Ix | Iy << 16) of the few window pixels whose bilinear taps straddle the image edge
    static constexpr int FIX_WORDS = (2 * (WW + WH) + 3) & ~3;   // left, right columns (by y), top, bottom rows (by x)
    static constexpr int WARP_WORDS = REGION_WORDS + PLANE_WORDS + JREGION_WORDS + FIX_WORDS;
};

// The eight (byte i, byte i+1) pairs of a run, as DP2A operands: with w0 = bytes 0..3, w1 = 4..7, w2 = 8.. of the run,
// pixel 0 = lo(w0), 1 = lo(s0), 2 = hi(w0), 3 = hi(s0), 4 = lo(w1), 5 = lo(s1), 6 = hi(w1), 7 = hi(s1),
// where s0 = bytes 1..4 and s1 = bytes 5..8.
struct RunRow {
    uint32_t w0, w1, s0, s1;
};

// shared-memory accesses by 32-bit shared-window address (keeps the address arithmetic out of the inner loops)
__device__ __forceinline__ void sts32(uint32_t addr, uint32_t v)
{
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ void sts128(uint32_t addr, int4 v)
{
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ uint2 lds64(uint32_t addr)
{
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ int4 lds128(uint32_t addr)
{
    int4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
    return v;
}

__device__ __forceinline__ RunRow load_run_row(uint32_t addr)
{
    const uint2 a = lds64(addr);
    const uint32_t b = lds32(addr + 8);
    RunRow r;
    r.w0 = a.x;
    r.w1 = a.y;
    r.s0 = __funnelshift_r(a.x, a.y, 8);
    r.s1 = __funnelshift_r(a.y, b, 8);
    return r;
}

// same for a run that starts `sh / 8` bytes into the word at addr (sh = 0, 8, 16, 24; warp-uniform)
__device__ __forceinline__ RunRow load_run_row_unaligned(uint32_t addr, int sh)
{
    const uint32_t a = lds32(addr), b = lds32(addr + 4), c = lds32(addr + 8);
    RunRow r;
    r.w0 = __funnelshift_r(a, b, sh);
    r.w1 = __funnelshift_r(b, c, sh);
    r.s0 = __funnelshift_rc(a, b, sh + 8);
    r.s1 = __funnelshift_rc(b, c, sh + 8);
    return r;
}

__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// d = c + a.lo16 * b.byte[0 or 2] + a.hi16 * b.byte[1 or 3]: SIGNED 16-bit weights (iw11 = 2^14 - iw00 - iw01 - iw10 can
// come out as -1 when the three roundings all go up), UNSIGNED 8-bit pixels.
__device__ __forceinline__ int dp2a_lo_su(uint32_t a, uint32_t b, int c)
{
    int d;
    asm("dp2a.lo.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ int dp2a_hi_su(uint32_t a, uint32_t b, int c)
{
    int d;
    asm("dp2a.hi.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ uint32_t pack_weights(int lo, int hi)
{
    if constexpr (KLT_OPTS & 4) return __byte_perm((uint32_t)lo, (uint32_t)hi, 0x5410);
    else return ((uint32_t)lo & 0xffffu) | ((uint32_t)hi << 16);
}

template <int I>
__device__ __forceinline__ int run_px(const RunRow& top, const RunRow& bot, uint32_t wt, uint32_t wb, int c)
{
    const uint32_t t = (I & 4) ? ((I & 1) ? top.s1 : top.w1) : ((I & 1) ? top.s0 : top.w0);
    const uint32_t b = (I & 4) ? ((I & 1) ? bot.s1 : bot.w1) : ((I & 1) ? bot.s0 : bot.w0);
    if (I & 2) return dp2a_hi_su(wb, b, dp2a_hi_su(wt, t, c));
    return dp2a_lo_su(wb, b, dp2a_lo_su(wt, t, c));
}

// Checked build (make EXTRA=-DSVO_BOUNDS_CHECK, see profiles/r03_bounds_check.txt): the two staging routines below read
// without predicates and rely on the margin arithmetic of DESIGN.md section 3.  In the checked build every one of those reads
// is tested against the padded extent of its pyramid level -- rows [-P, h + P), columns [-P, w + P) around the interior --
// which is tighter than the allocation: a read outside it would be memory-safe at best and would see bytes that are not the
// BORDER_REFLECT_101 margin.  A violation prints the site and traps (the call returns SVO_GPU_E_CUDA).
#ifdef SVO_BOUNDS_CHECK
#ifndef SVO_BOUNDS_CHECK_SHRINK
#define SVO_BOUNDS_CHECK_SHRINK 0   // > 0: pretend the margin is that much narrower -- the control run that shows the check fires
#endif
__device__ __noinline__ void check_level_read(const uint8_t* p, int bytes, const uint8_t* origin, int w, int h, int pitch, int P,
                                              int site)
{
    const long long d = p - origin;   // byte offset from interior pixel (0, 0)
    long long row = d / pitch, col = d - row * pitch;
    if (col < -P) { col += pitch; --row; }
    else if (col >= pitch - P) { col -= pitch; ++row; }
    if (row < -P || row >= h + P || col < -P || col + bytes > w + P) {
        printf("SVO_BOUNDS_CHECK site %d: rows %lld cols %lld..%lld outside [-%d, %d) x [-%d, %d)\n", site, row, col, col + bytes,
               P, h + P, P, w + P);
        __trap();
    }
}
#define SVO_CHECK_READ(p, bytes, origin, L, P, site) \
    check_level_read((p), (bytes), (origin), (L).w, (L).h, (L).pitch, (P) - SVO_BOUNDS_CHECK_SHRINK, (site))
#else
#define SVO_CHECK_READ(p, bytes, origin, L, P, site) ((void)0)
#endif

// Stages rows of RP bytes whose first byte is column x0 of row y0 (any alignment) into shared memory, rows RP bytes
// apart: two aligned 32-bit loads + one funnel shift per output word, all loads issued before the first store.
// Works in PASSES whole passes of 32 / (RP/4) rows, so ROWS is rounded up (the pyramid margin covers the extra rows) and
// no lane needs a predicate; lanes beyond the last full row of a pass repeat the work of the last row (same values).
// src0 = address of the word holding (x0, y0 + r0) + 4 c0 for this lane's (r0, c0); dst0 = shared address of (r0, c0).
template <int ROWS, int RP>
__device__ __forceinline__ void stage_shifted(uint32_t dst0, const uint8_t* __restrict__ src0, int pitch, int x0)
{
    constexpr int OW = RP / 4, RPP = 32 / OW, PASSES = (ROWS + RPP - 1) / RPP;
    const int sh = 8 * (x0 & 3);
    uint32_t lo[PASSES], hi[PASSES];
#pragma unroll
    for (int k = 0; k < PASSES; ++k) {
        const uint8_t* p = src0 + (long long)(k * RPP) * pitch;
        lo[k] = __ldg(reinterpret_cast<const uint32_t*>(p));
        hi[k] = __ldg(reinterpret_cast<const uint32_t*>(p + 4));
    }
#pragma unroll
    for (int k = 0; k < PASSES; ++k) sts32(dst0 + k * RPP * RP, __funnelshift_r(lo[k], hi[k], sh));
}

template <int WW, int WH, int WARPS, bool PACKG, int MINB>
__global__ void __launch_bounds__(WARPS * 32, MINB) klt_run_kernel(const KltArgs a)
{
    using G = RunGeo<WW, WH>;
    constexpr int RPL = G::RPL, BP = G::BP;
    constexpr int NG = PACKG ? 1 : 2;   // registers per pixel for the gradients

    __shared__ __align__(16) uint32_t smem[WARPS * G::WARP_WORDS];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    static_assert(WARPS == 1, "a CTA is one warp: it fetches one track at a time");

    uint32_t* region = smem + warp * G::WARP_WORDS;
    int* plane = reinterpret_cast<int*>(region + G::REGION_WORDS);
    const uint8_t* rbytes = reinterpret_cast<const uint8_t*>(region);

    // ---- per-lane constants: staging maps, and for each owned run its offsets and number of valid pixels
    constexpr int OWI = G::RPI / 4;
    const uint32_t region_sa = (uint32_t)__cvta_generic_to_shared(region);   // 32-bit shared-window addresses
    const uint32_t plane_sa = (uint32_t)__cvta_generic_to_shared(plane);
    const int sI_r = min(lane / OWI, 32 / OWI - 1), sI_c = lane % OWI;       // staging maps: lane -> (row, word) of a pass;
                                                                             // spare lanes repeat a word of the last row
    const uint32_t jregion_sa = plane_sa + 4 * G::PLANE_WORDS;
    const uint32_t sI_dst = region_sa + 4 * (sI_r * OWI + sI_c);
    // next-image region: lane -> (row, word) of a staging pass; spare lanes repeat a word of the pass's last row
    const int sJ_r = min(lane / G::OWJ, G::RPPJ - 1), sJ_c = lane % G::OWJ;
    const uint32_t sJ_dst = jregion_sa + 4 * (sJ_r * G::PJW + sJ_c);
    uint32_t jo[RPL], po[RPL];
    int nval[RPL];
#pragma unroll
    for (int j = 0; j < RPL; ++j) {
        const int r = lane + 32 * j;
        const int rr = min(r, G::RUNS - 1);
        const int y = rr / G::NSEG, seg = rr - y * G::NSEG;
        jo[j] = jregion_sa + y * G::PJ + 8 * seg;    // shared address of the run when the window sits at (0, 0) of the region
        po[j] = plane_sa + 4 * (y * BP + G::BBLK * seg);   // shared address of plane element (x, y), x = the run's first pixel
        nval[j] = r < G::RUNS ? min(8, WW - 8 * seg) : 0;
    }

    int Iw[RPL][8], Gx[RPL][8 * NG];   // Gx: Ix then Iy (NG == 2) or (Ix & 0xffff) | (Iy << 16) (PACKG)

    const float halfx = (WW - 1) * 0.5f, halfy = (WH - 1) * 0.5f;
    // float brackets of eps^2 for the convergence test: a float sum of squares outside [eps_lo, eps_hi] decides the double test
    const float eps_lo = __double2float_rd(a.eps2 * 0.99999), eps_hi = __double2float_ru(a.eps2 * 1.00001);
    // Persistent CTAs (one warp each, as many as fit on the device): tracks are handed out one at a time through a global
    // counter, so a CTA that finishes a short track fetches the next one at once (tracks differ a lot in iteration count)
    // and the per-lane constants above are set up once per CTA instead of once per track.
    const unsigned total = (unsigned)a.n_max * gridDim.y;   // gridDim.y carries the batch
    unsigned long long iters_cta = 0;
    for (;;) {
    unsigned t = 0;
    if (lane == 0) t = atomicAdd(a.work, 1u);
    t = __shfl_sync(kFull, t, 0);
    if (t >= total) break;
    const int lane_id = (int)(t / (unsigned)a.n_max), pt = (int)(t - (unsigned)lane_id * (unsigned)a.n_max);
    if (pt >= min(a.n_dev[lane_id], a.n_max)) continue;
    const size_t pbase = (size_t)lane_id * a.pts_lane + pt;
    const float2 prev = a.prev_xy[pbase];
    const uint8_t* Ppyr = a.prev_pyr + (size_t)lane_id * a.g.lane_bytes;
    const uint8_t* Npyr = a.next_pyr + (size_t)lane_id * a.g.lane_bytes;

    float ox = 0.f, oy = 0.f;   // nextPts[pt]
    int st = 1;
    float er = 0.f;
    unsigned iters = 0;

    auto set_px = [&](int j, int i, int iv, int ixv, int iyv) {
        Iw[j][i] = iv;
        if constexpr (PACKG) Gx[j][i] = (ixv & 0xffff) | (iyv << 16);
        else { Gx[j][i] = ixv; Gx[j][8 + i] = iyv; }
    };
    auto get_ix = [&](int j, int i) { if constexpr (PACKG) return (int)(short)(Gx[j][i] & 0xffff); else return Gx[j][i]; };
    auto get_iy = [&](int j, int i) { if constexpr (PACKG) return Gx[j][i] >> 16; else return Gx[j][8 + i]; };

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
        if (ipx < -WW || ipx >= L.w || ipy < -WH || ipy >= L.h) {
            if (level == 0) { st = 0; er = 0.f; }
            continue;
        }
        BiW w = bilinear_weights(__fsub_rn(px, (float)ipx), __fsub_rn(py, (float)ipy));

        // ---- next image: start the asynchronous copy of the window region (+ margin) around the initial guess now, so
        //      that its L2 latency hides behind the patch set-up below
        nx = __fsub_rn(nx, halfx); ny = __fsub_rn(ny, halfy);
        const uint8_t* Jorigin = Npyr + L.off;
        int sx0 = 0, sy0 = 0;      // image coordinates of region byte (0, 0)
        bool jstaged = false;
        bool rsafe = false;        // staged, and every window origin inside the region passes the level-bounds test
        auto stage_j = [&](int inx, int iny) {
            sx0 = (inx - G::MJ) & ~3;
            sy0 = iny - G::MJ;
            const uint8_t* src = Jorigin + (long long)(sy0 + sJ_r) * L.pitch + (sx0 + 4 * sJ_c);
#pragma unroll
            for (int k = 0; k < G::PASSJ; ++k) {
                SVO_CHECK_READ(src + (long long)(k * G::RPPJ) * L.pitch, 4, Jorigin, L, a.g.P, 2);
                cp_async4(sJ_dst + k * G::RPPJ * G::PJ, src + (long long)(k * G::RPPJ) * L.pitch);
            }
            jstaged = true;
            rsafe = sx0 >= -WW && sx0 + G::DXMAX < L.w && sy0 >= -WH && sy0 + G::DYMAX < L.h;
        };
        cp_async_wait_all();   // a copy started at a level that was then skipped may still be in flight
        __syncwarp();          // and every lane is done reading the previous level's regions
        {
            const int inx = (int)floorf(nx), iny = (int)floorf(ny);
            if (inx >= -WW && inx < L.w && iny >= -WH && iny < L.h) stage_j(inx, iny);
        }

        // ---- stage the (WW+3) x (WH+3) region of I whose top-left is (ipx-1, ipy-1); staged byte (u, v) = I(ipx-1+u, ipy-1+v)
        {
            const int x0 = ipx - 1;
            const uint8_t* src0 = Ppyr + L.off + (long long)(ipy - 1 + sI_r) * L.pitch + ((x0 & ~3) + 4 * sI_c);
#ifdef SVO_BOUNDS_CHECK
            {
                constexpr int RPPI = 32 / (G::RPI / 4), PASSI = (G::IROWS + RPPI - 1) / RPPI;   // as in stage_shifted
                for (int k = 0; k < PASSI; ++k) SVO_CHECK_READ(src0 + (long long)(k * RPPI) * L.pitch, 8, Ppyr + L.off, L, a.g.P, 1);
            }
#endif
            stage_shifted<G::IROWS, G::RPI>(sI_dst, src0, L.pitch, x0);
        }
        __syncwarp();

        // all derivative taps (ipx..ipx+WW, ipy..ipy+WH) inside the image?  (warp-uniform)
        const bool interior = ipx >= 0 && ipy >= 0 && ipx + WW < L.w && ipy + WH < L.h;
        int sA11 = 0, sA12 = 0, sA22 = 0;
        // ---- B(u,v) = unrounded bilinear sample of I at (ipx-1+u, ipy-1+v): plane run rb covers u = 8 seg .. 8 seg + 7
        //      of row v, and sits in block rb of the plane (BBLK words each) / at byte 8 rb of the region (NSEGB runs per row)
        {
            const uint32_t wt = pack_weights(w.w00, w.w01), wb = pack_weights(w.w10, w.w11);
#pragma unroll
            for (int j = 0; j < G::RPLB; ++j) {
                const int rb = lane + 32 * j;
                if (rb < G::RUNSB) {
                    const RunRow top = load_run_row(region_sa + 8 * rb), bot = load_run_row(region_sa + 8 * rb + G::RPI);
                    int4 lo4, hi4;
                    lo4.x = run_px<0>(top, bot, wt, wb, 0); lo4.y = run_px<1>(top, bot, wt, wb, 0);
                    lo4.z = run_px<2>(top, bot, wt, wb, 0); lo4.w = run_px<3>(top, bot, wt, wb, 0);
                    hi4.x = run_px<4>(top, bot, wt, wb, 0); hi4.y = run_px<5>(top, bot, wt, wb, 0);
                    hi4.z = run_px<6>(top, bot, wt, wb, 0); hi4.w = run_px<7>(top, bot, wt, wb, 0);
                    sts128(plane_sa + 4 * G::BBLK * rb, lo4);
                    sts128(plane_sa + 4 * G::BBLK * rb + 16, hi4);
                }
            }
        }
        // ---- Iw = DESCALE(B, 9);  Ixw = DESCALE(Scharr_x(B), 14);  Iyw = DESCALE(Scharr_y(B), 14)   (integer linearity)
        //      separably: cs(u) = 3 B(u,v-1) + 10 B(u,v) + 3 B(u,v+1), rd(u) = B(u,v+1) - B(u,v-1);
        //      Scharr_x = cs(u+1) - cs(u-1), Scharr_y = 3 rd(u-1) + 10 rd(u) + 3 rd(u+1)
        auto derive_run = [&](int j, int* iv, int* ixv, int* iyv) {
            int cs[10], rd[10], mid[10];
            {
                int r0[10], r2[10];
                auto load10 = [](uint32_t sa, int* d) {
                    const int4 x = lds128(sa), y = lds128(sa + 16);
                    const uint2 z = lds64(sa + 4 * G::BBLK);   // values 8, 9 open the next block
                    d[0] = x.x; d[1] = x.y; d[2] = x.z; d[3] = x.w; d[4] = y.x; d[5] = y.y; d[6] = y.z; d[7] = y.w;
                    d[8] = (int)z.x; d[9] = (int)z.y;
                };
                load10(po[j], r0);
                load10(po[j] + 4 * BP, mid);
                load10(po[j] + 8 * BP, r2);
#pragma unroll
                for (int u = 0; u < 10; ++u) {
                    cs[u] = 3 * (r0[u] + r2[u]) + 10 * mid[u];
                    rd[u] = r2[u] - r0[u];
                }
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                iv[i] = descale(mid[i + 1], kWBits - 5);
                ixv[i] = descale(cs[i + 2] - cs[i], kWBits);
                iyv[i] = descale(3 * (rd[i] + rd[i + 2]) + 10 * rd[i + 1], kWBits);
            }
        };
        if (interior) {
            __syncwarp();
#pragma unroll
            for (int j = 0; j < RPL; ++j) {
                int iv[8], ixv[8], iyv[8];
                derive_run(j, iv, ixv, iyv);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    if (i >= nval[j]) { ixv[i] = 0; iyv[i] = 0; }
                    set_px(j, i, iv[i], ixv[i], iyv[i]);
                    sA11 += ixv[i] * ixv[i]; sA12 += ixv[i] * iyv[i]; sA22 += iyv[i] * iyv[i];
                }
            }
        } else {
            // ---- border window.  OpenCV's derivative image is ZERO outside the level (calcScharrDeriv writes the interior,
            //      the winSize border is BORDER_CONSTANT) while the level itself is REFLECT_101-padded, so a window pixel at
            //      image position (X, Y) has taps (X+i, Y+j), i, j in {0, 1}, of which only those inside the level count:
            //        all four inside            -> the identity above holds (the margin of the stored level IS the
            //                                      REFLECT_101 border calcScharrDeriv computes with)
            //        none inside                -> Ix = Iy = 0
            //        some inside ("straddling": X == -1, X == w-1, Y == -1 or Y == h-1; at most two columns and two rows
            //                      of the window) -> the literal masked sum, computed here one pixel per lane into `fix`
            const uint32_t fix_sa = jregion_sa + 4 * G::JREGION_WORDS;
            const int xlo = max(0, -1 - ipx), xhi = min(WW, L.w - ipx);   // columns / rows with at least one tap inside
            const int ylo = max(0, -1 - ipy), yhi = min(WH, L.h - ipy);
            const int xsL = -1 - ipx, xsR = L.w - 1 - ipx, ysT = -1 - ipy, ysB = L.h - 1 - ipy;   // straddling ones
            const bool hL = xsL >= 0 && xsL < WW, hR = xsR >= 0 && xsR < WW && xsR != xsL;
            const bool hT = ysT >= 0 && ysT < WH, hB = ysB >= 0 && ysB < WH && ysB != ysT;
            const int nL = hL ? WH : 0, nR = hR ? WH : 0, nT = hT ? WW : 0, nB = hB ? WW : 0;
            const int nfix = nL + nR + nT + nB;
            // list position k (compacted over the lists that exist) -> fixed slot: L at 0, R at WH, T at 2 WH, B at 2 WH + WW
            auto k2slot = [&](int k) {
                int t = k;
                if (t < nL) return t;
                if ((t -= nL) < nR) return WH + t;
                if ((t -= nR) < nT) return 2 * WH + t;
                return 2 * WH + WW + t - nT;
            };
            for (int k = lane; k < nfix; k += 32) {
                int x, y, t = k;
                if (t < nL) { x = xsL; y = t; }
                else if ((t -= nL) < nR) { x = xsR; y = t; }
                else if ((t -= nR) < nT) { x = t; y = ysT; }
                else { x = t - nT; y = ysB; }
                int ax = 0, ay = 0;
#pragma unroll
                for (int tj = 0; tj < 2; ++tj)
#pragma unroll
                    for (int ti = 0; ti < 2; ++ti) {
                        const int X = ipx + x + ti, Y = ipy + y + tj;
                        if (X >= 0 && X < L.w && Y >= 0 && Y < L.h) {
                            const uint8_t* r0 = rbytes + (y + tj) * G::RPI + (x + ti);   // row Y-1, column X-1
                            const uint8_t* r1 = r0 + G::RPI;
                            const uint8_t* r2 = r1 + G::RPI;
                            const int sx = 3 * ((int)r0[2] - (int)r0[0]) + 10 * ((int)r1[2] - (int)r1[0]) + 3 * ((int)r2[2] - (int)r2[0]);
                            const int sy = 3 * ((int)r2[0] - (int)r0[0]) + 10 * ((int)r2[1] - (int)r0[1]) + 3 * ((int)r2[2] - (int)r0[2]);
                            const int wk = tj ? (ti ? w.w11 : w.w10) : (ti ? w.w01 : w.w00);
                            ax += sx * wk; ay += sy * wk;
                        }
                    }
                sts32(fix_sa + 4 * k2slot(k), ((uint32_t)descale(ax, kWBits) & 0xffffu) | ((uint32_t)descale(ay, kWBits) << 16));
            }
            __syncwarp();
#pragma unroll
            for (int j = 0; j < RPL; ++j) {
                int iv[8], ixv[8], iyv[8];
                derive_run(j, iv, ixv, iyv);
                const int r = min(lane + 32 * j, G::RUNS - 1);
                const int y = r / G::NSEG, x0 = 8 * (r - y * G::NSEG);
                // bit i of zm: pixel i of the run has no tap inside (or is padding); of fm: it straddles -> take `fix`
                const int lo = min(max(xlo - x0, 0), 8), hi = min(max(min(xhi, x0 + nval[j]) - x0, 0), 8);
                unsigned keep = (y >= ylo && y < yhi) ? (((1u << hi) - 1u) & ~((1u << lo) - 1u)) : 0u;
                unsigned fm = 0;
                if ((hT && y == ysT) || (hB && y == ysB)) fm = keep;
                else {
                    if (hL && (unsigned)(xsL - x0) < 8u) fm |= 1u << (xsL - x0);
                    if (hR && (unsigned)(xsR - x0) < 8u) fm |= 1u << (xsR - x0);
                    fm &= keep;
                }
                if ((keep & ~fm) != 0xffu) {   // rare lanes: something to zero or to patch
                    const bool rowT = hT && y == ysT, rowB = hB && y == ysB;
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        if (!((keep >> i) & 1u)) { ixv[i] = 0; iyv[i] = 0; }
                        else if ((fm >> i) & 1u) {
                            const int slot = rowT ? 2 * WH + x0 + i : rowB ? 2 * WH + WW + x0 + i : (x0 + i == xsL && hL) ? y : WH + y;
                            const uint32_t v = lds32(fix_sa + 4 * slot);
                            ixv[i] = (int)(short)(v & 0xffffu);
                            iyv[i] = (int)v >> 16;
                        }
                    }
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    set_px(j, i, iv[i], ixv[i], iyv[i]);
                    sA11 += ixv[i] * ixv[i]; sA12 += ixv[i] * iyv[i]; sA22 += iyv[i] * iyv[i];
                }
            }
        }
        const float A11 = warp_sum_scaled(sA11), A12 = warp_sum_scaled(sA12), A22 = warp_sum_scaled(sA22);   // x FLT_SCALE

        float D = __fsub_rn(__fmul_rn(A11, A22), __fmul_rn(A12, A12));
        const float minEig = min_eig(A11, A12, A22, WW * WH);
        if (a.flags & SVO_GPU_LK_GET_MIN_EIGENVALS) er = minEig;
        if ((double)minEig < a.min_eig_threshold || D < FLT_EPSILON) {
            if (level == 0) st = 0;
            continue;
        }
        D = __fdiv_rn(1.f, D);

        // one pass over the window at the staged position of the next image: exact integer sums of diff*Ix, diff*Iy
        // (Newton step) or |diff| (L1 residual); diff = DESCALE(bilinear J, 9) - Iw, the rounding constant rides in the
        // DP2A accumulator
        auto window_pass = [&](int dx, int dy, int& s1, int& s2, auto l1_tag) {
            constexpr bool L1 = decltype(l1_tag)::value;
            const uint32_t uoff = (uint32_t)(dy * G::PJ + (dx & ~3));   // warp-uniform: window origin inside the region
            const int sh = 8 * (dx & 3);
            const uint32_t wt = pack_weights(w.w00, w.w01), wb = pack_weights(w.w10, w.w11);
            constexpr int RND = 1 << (kWBits - 6);
            s1 = 0; s2 = 0;
#pragma unroll
            for (int j = 0; j < RPL; ++j) {
                const RunRow top = load_run_row_unaligned(jo[j] + uoff, sh), bot = load_run_row_unaligned(jo[j] + uoff + G::PJ, sh);
                int v[8];
                v[0] = run_px<0>(top, bot, wt, wb, RND) >> (kWBits - 5); v[1] = run_px<1>(top, bot, wt, wb, RND) >> (kWBits - 5);
                v[2] = run_px<2>(top, bot, wt, wb, RND) >> (kWBits - 5); v[3] = run_px<3>(top, bot, wt, wb, RND) >> (kWBits - 5);
                v[4] = run_px<4>(top, bot, wt, wb, RND) >> (kWBits - 5); v[5] = run_px<5>(top, bot, wt, wb, RND) >> (kWBits - 5);
                v[6] = run_px<6>(top, bot, wt, wb, RND) >> (kWBits - 5); v[7] = run_px<7>(top, bot, wt, wb, RND) >> (kWBits - 5);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int diff = v[i] - Iw[j][i];
                    if constexpr (L1) { if (i < nval[j]) s1 += abs(diff); }
                    else { s1 += diff * get_ix(j, i); s2 += diff * get_iy(j, i); }   // invalid slots hold Ix = Iy = 0
                }
            }
        };

        // window origin inside the staged region; re-stages (rare) when the window has left it
        auto locate = [&](int inx, int iny, int& dx, int& dy) {
            dx = inx - sx0; dy = iny - sy0;
            if (!jstaged || (unsigned)dx > (unsigned)G::DXMAX || (unsigned)dy > (unsigned)G::DYMAX) {   // warp-uniform
                __syncwarp();
                cp_async_wait_all();
                stage_j(inx, iny);
                cp_async_wait_all();
                __syncwarp();
                dx = inx - sx0; dy = iny - sy0;
            }
        };
        cp_async_wait_all();
        __syncwarp();
        float pdx = 0.f, pdy = 0.f;
        for (int j = 0; j < a.max_count; ++j) {
            const int inx = (int)floorf(nx), iny = (int)floorf(ny);
            int wdx = inx - sx0, wdy = iny - sy0, sb1, sb2;
            if (!(KLT_OPTS & 8) || !rsafe || (unsigned)wdx > (unsigned)G::DXMAX || (unsigned)wdy > (unsigned)G::DYMAX) {
                if (inx < -WW || inx >= L.w || iny < -WH || iny >= L.h) {
                    if (level == 0) st = 0;
                    break;
                }
                locate(inx, iny, wdx, wdy);
            }
            ++iters;
            w = bilinear_weights(__fsub_rn(nx, (float)inx), __fsub_rn(ny, (float)iny));
            window_pass(wdx, wdy, sb1, sb2, std::false_type{});
            const float b1 = warp_sum_scaled(sb1), b2 = warp_sum_scaled(sb2);   // x FLT_SCALE
            const float dx = __fmul_rn(__fsub_rn(__fmul_rn(A12, b2), __fmul_rn(A22, b1)), D);
            const float dy = __fmul_rn(__fsub_rn(__fmul_rn(A12, b1), __fmul_rn(A11, b2)), D);
            nx = __fadd_rn(nx, dx); ny = __fadd_rn(ny, dy);
            ox = __fadd_rn(nx, halfx); oy = __fadd_rn(ny, halfy);
            // OpenCV tests delta.ddot(delta) <= eps^2 in double.  The float sum of squares is within 2e-7 of it, so outside the
            // bracket (1 -+ 1e-5) eps^2 it decides the test; NaNs and the bracket itself take the double path.
            if constexpr (KLT_OPTS & 2) {
                const float ss = __fmaf_rn(dx, dx, __fmul_rn(dy, dy));
                if (ss < eps_lo) break;
                if (!(ss > eps_hi) &&
                    __dadd_rn(__dmul_rn((double)dx, (double)dx), __dmul_rn((double)dy, (double)dy)) <= a.eps2) break;
            } else {
                if (__dadd_rn(__dmul_rn((double)dx, (double)dx), __dmul_rn((double)dy, (double)dy)) <= a.eps2) break;
            }
            // |d + dprev| < 0.01 (a double in OpenCV): 0.01f is the largest float below 0.01, so the float test `<= 0.01f` is the same
            if (j > 0 && fabsf(__fadd_rn(dx, pdx)) <= 0.01f && fabsf(__fadd_rn(dy, pdy)) <= 0.01f) {
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
            if (iqx < -WW || iqx >= L.w || iqy < -WH || iqy >= L.h) {
                st = 0;
                continue;
            }
            w = bilinear_weights(__fsub_rn(qx, (float)iqx), __fsub_rn(qy, (float)iqy));
            int wdx, wdy, serr, unused;
            locate(iqx, iqy, wdx, wdy);
            window_pass(wdx, wdy, serr, unused, std::true_type{});
            er = __fdiv_rn(__fmul_rn(warp_sum_exact_f(serr), 1.f), (float)(32 * WW * WH));
        }
    }

    if (lane == 0) {
        a.next_xy[pbase] = make_float2(ox, oy);
        a.status[pbase] = (uint8_t)st;
        a.err[pbase] = er;
    }
    iters_cta += iters;
    cp_async_wait_all();
    __syncwarp();
    }   // next track
    if (lane == 0 && a.iters) atomicAdd(a.iters, iters_cta);
}

// One launch with every resident CTA slot of the device (or one CTA per track when there are fewer tracks).
template <int WW, int WH, int WARPS, bool PACKG, int MINB>
void launch_run(cudaStream_t st, const KltArgs& a, int batch)
{
    static std::atomic<unsigned long long> seen{0};
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    // every kernel that may share an SM with this one asks for the same (largest) shared-memory carve-out
    if (first_on_device(seen)) prefer_max_shared(klt_run_kernel<WW, WH, WARPS, PACKG, MINB>);
    const long long total = (long long)a.n_max * batch;
    dim3 grid((unsigned)std::min<long long>(total, (long long)sms * MINB), batch);
    grid.x = (grid.x + batch - 1) / batch;   // gridDim.y = batch only tells the kernel the batch size
    cudaMemsetAsync(a.work, 0, sizeof(unsigned), st);
    klt_run_kernel<WW, WH, WARPS, PACKG, MINB><<<grid, WARPS * 32, 0, st>>>(a);
}

// tuning knobs of the 21x21 instantiation (warps per CTA, packed gradients, minimum CTAs per SM -> register cap)
#ifndef KLT21_WARPS
#define KLT21_WARPS 1   // one track per CTA: a slot is refilled as soon as its track converges (no intra-CTA tail)
#endif
#ifndef KLT21_PACKG
#define KLT21_PACKG false
#endif
#ifndef KLT21_MINB
#define KLT21_MINB 20
#endif
#ifndef KLT11_MINB
#define KLT11_MINB 32   // one run per lane, 64 registers: 16 -> 32 CTAs per SM, 0.757 -> 0.667 ms per 128 k tracks (gpurun_out: w11 sweep)
#endif
#ifndef KLT31_PACKG
#define KLT31_PACKG true
#endif
#ifndef KLT31_MINB
#define KLT31_MINB 12   // 3 warps per scheduler: 6.14 -> 5.64 ms on the 4K / 50 k-track workload (8 and 10 CTAs per SM measured equal)
#endif

}  // namespace

int klt_resident_tracks(int win_w, int win_h)
{
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int minb = (win_w == 21 && win_h == 21) ? KLT21_MINB : (win_w == 11 && win_h == 11) ? KLT11_MINB
                   : (win_w == 31 && win_h == 31) ? KLT31_MINB : 16;
    return sms * minb;
}

void launch_klt(cudaStream_t st, const KltArgs& a, int batch, long long* launches)
{
    if (a.n_max <= 0) return;
    if (a.win_w == 21 && a.win_h == 21) launch_run<21, 21, KLT21_WARPS, KLT21_PACKG, KLT21_MINB>(st, a, batch);
    else if (a.win_w == 11 && a.win_h == 11) launch_run<11, 11, 1, false, KLT11_MINB>(st, a, batch);
    else if (a.win_w == 31 && a.win_h == 31) launch_run<31, 31, 1, KLT31_PACKG, KLT31_MINB>(st, a, batch);
    else launch_klt_generic(st, a, batch);
    if (launches) ++*launches;
}

}  // namespace svo
