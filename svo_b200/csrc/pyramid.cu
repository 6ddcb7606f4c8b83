// pyramid.cu -- K1: image ingest (REFLECT_101 margin) and 5x5 Gaussian pyrDown chain for sm_100a.
//
// Same arithmetic as cv::buildOpticalFlowPyramid(img, pyr, winSize, maxLevel, withDerivatives=false), which
// cv::calcOpticalFlowPyrLK runs on both images at every call (reference call sites src/tracking.cpp:101-105, :160-165):
//   level l = pyrDown(level l-1):  dst(x,y) = (sum_{i,j=-2..2} w_i w_j src(R(2x+i), R(2y+j)) + 128) >> 8, w = [1 4 6 4 1],
//   R = BORDER_REFLECT_101, dst size ((w+1)/2, (h+1)/2); every level carries a REFLECT_101 margin.
// HBM-bound streaming stencils: each thread produces 4 output bytes from aligned word loads, once, and scatters them to
// the margin positions that reflect onto them.
#include <algorithm>

#include "svo_internal.cuh"

namespace svo {

// ---------------------------------------------------------------------------------------------------------------
// raw -> padded level 0, in 16-byte chunks of a padded row (one aligned 128-bit store each; the margin is a multiple of 16, so
// chunk boundaries fall on column 0).  Chunks whose 16 + 4 source bytes lie inside the raw row are copied by
// ingest_kernel: five aligned words + four funnel shifts, rows of any pitch / alignment (compact uploads have pitch = w).
// The few chunks per row that touch the left / right margin go to ingest_edge_kernel, whose warps hold 32 consecutive ROWS
// of one edge chunk, so that the per-byte BORDER_REFLECT_101 path never shares a warp with the streaming path.
// grid = (ceil(nfast (h+2P) / 256), 1, batch): (row, chunk) pairs flattened, so that every thread of a CTA has a chunk whatever
// the row length  and  (ceil((h+2P) / 128), nedge, batch)
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ingest_kernel(const uint8_t* __restrict__ raw, size_t raw_pitch, size_t raw_lane,
                                                     uint8_t* __restrict__ pyr, long long lane_bytes, LevelGeom L, int P,
                                                     int c_begin, int c_end)
{
    const int nfast = c_end - c_begin;
    const int item = blockIdx.x * blockDim.x + threadIdx.x;
    const int row = item / nfast;
    if (row >= L.h + 2 * P) return;
    const int c = c_begin + (item - row * nfast);
    const int x0 = 16 * c - P;
    const int py = row - P;
    const int lane = blockIdx.z;
    const uint8_t* srow = raw + (size_t)lane * raw_lane + (size_t)reflect101(py, L.h) * raw_pitch;
    const uintptr_t addr = reinterpret_cast<uintptr_t>(srow + x0);
    const uint32_t* wp = reinterpret_cast<const uint32_t*>(addr & ~(uintptr_t)3);
    const int sh = 8 * (int)(addr & 3);
    const uint32_t w0 = __ldg(wp), w1 = __ldg(wp + 1), w2 = __ldg(wp + 2), w3 = __ldg(wp + 3), w4 = __ldg(wp + 4);
    uint4 v;
    v.x = __funnelshift_r(w0, w1, sh);
    v.y = __funnelshift_r(w1, w2, sh);
    v.z = __funnelshift_r(w2, w3, sh);
    v.w = __funnelshift_r(w3, w4, sh);
    uint8_t* drow = pyr + (size_t)lane * lane_bytes + L.off + (long long)py * L.pitch;
    *reinterpret_cast<uint4*>(drow + x0) = v;
}

__global__ void __launch_bounds__(128) ingest_edge_kernel(const uint8_t* __restrict__ raw, size_t raw_pitch, size_t raw_lane,
                                                          uint8_t* __restrict__ pyr, long long lane_bytes, LevelGeom L, int P,
                                                          int c_begin, int c_end, int n_chunks)
{
    const int row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= L.h + 2 * P) return;
    // edge chunk e: the chunks before c_begin, then those from c_end on
    const int e = blockIdx.y;
    const int c = e < c_begin ? e : c_end + (e - c_begin);
    if (c >= n_chunks) return;
    const int x0 = 16 * c - P;
    const int py = row - P;
    const int lane = blockIdx.z;
    const uint8_t* srow = raw + (size_t)lane * raw_lane + (size_t)reflect101(py, L.h) * raw_pitch;
    uint32_t v[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        v[q] = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) v[q] |= (uint32_t)__ldg(srow + reflect101(x0 + 4 * q + i, L.w)) << (8 * i);
    }
    uint8_t* drow = pyr + (size_t)lane * lane_bytes + L.off + (long long)py * L.pitch;
    *reinterpret_cast<uint4*>(drow + x0) = make_uint4(v[0], v[1], v[2], v[3]);
}

// ---------------------------------------------------------------------------------------------------------------
// level l-1 -> level l.  One thread = 4 consecutive dst pixels of one dst row, each output computed ONCE: the thread then
// also stores its pixels at every margin position that BORDER_REFLECT_101 maps onto them (rows: the same quad, aligned;
// columns: mirrored, byte by byte), so the level leaves the kernel with its complete margin and the next level -- and LK --
// read it without border logic.  The source level carries its own margin (ingest / the previous launch), hence no thread
// needs border logic for its reads either: 5 rows x 16 bytes as aligned 32/64-bit loads, and the separable [1 4 6 4 1]
// filter as DP4A: the horizontal taps of output k sit in two ALIGNED words, so with the vertical weight folded in
//   out0 = w0 . (0,0,1,4)K + w1 . (6,4,1,0)K     out1 = w1 . (1,4,6,4)K + w2 . (1,0,0,0)K
//   out2 = w1 . (0,0,1,4)K + w2 . (6,4,1,0)K     out3 = w2 . (1,4,6,4)K + w3 . (1,0,0,0)K      (K = 1, 4, 6, 4, 1 per row)
// i.e. 8 DP4A per source row and no shifts or byte extraction.
// grid = (ceil(ceil(dw/4) dh / 256), 1, batch)
// ---------------------------------------------------------------------------------------------------------------
// smallest value >= lo that is congruent to v modulo period
__device__ __forceinline__ int first_at_least(int v, int period, int lo)
{
    int m = (v - lo) % period;
    if (m < 0) m += period;
    return lo + m;
}

__device__ __forceinline__ uint32_t pack4(uint32_t a, uint32_t b, uint32_t c, uint32_t d)
{
    return a | (b << 8) | (c << 16) | (d << 24);
}

__global__ void __launch_bounds__(256) pyrdown_kernel(uint8_t* __restrict__ pyr, long long lane_bytes, LevelGeom S,
                                                      LevelGeom D, int P, int quads)
{
    // (row, quad) pairs flattened: every thread of a CTA has work whatever the row length
    const int item = blockIdx.x * blockDim.x + threadIdx.x;
    const int py = item / quads;
    if (py >= D.h) return;
    const int px = 4 * (item - py * quads);
    uint8_t* base = pyr + (size_t)blockIdx.z * lane_bytes;
    // source columns 2px-4 .. 2px+11 (bytes 2..12 are used), rows 2py-2 .. 2py+2; 2px-4 is 4 mod 8: one 32-bit, one 64-bit,
    // one 32-bit load per row
    uint32_t acc0 = 128, acc1 = 128, acc2 = 128, acc3 = 128;
    const uint8_t* r0 = base + S.off + (long long)(2 * py - 2) * S.pitch + 2 * px - 4;
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        const uint32_t k = j == 0 || j == 4 ? 1 : j == 2 ? 6 : 4;
        const uint8_t* r = r0 + (long long)j * S.pitch;
        const uint2 mid = __ldg(reinterpret_cast<const uint2*>(r + 4));
        const uint32_t w0 = __ldg(reinterpret_cast<const uint32_t*>(r)), w1 = mid.x, w2 = mid.y;
        const uint32_t w3 = __ldg(reinterpret_cast<const uint32_t*>(r + 12));
        const uint32_t A = (k << 16) | ((4 * k) << 24);                            // (0, 0, 1, 4) k
        const uint32_t B = (6 * k) | ((4 * k) << 8) | (k << 16);                   // (6, 4, 1, 0) k
        const uint32_t Cw = k | ((4 * k) << 8) | ((6 * k) << 16) | ((4 * k) << 24);   // (1, 4, 6, 4) k
        const uint32_t E = k;                                                      // (1, 0, 0, 0) k
        acc0 = __dp4a(w0, A, acc0); acc0 = __dp4a(w1, B, acc0);
        acc1 = __dp4a(w1, Cw, acc1); acc1 = __dp4a(w2, E, acc1);
        acc2 = __dp4a(w1, A, acc2); acc2 = __dp4a(w2, B, acc2);
        acc3 = __dp4a(w2, Cw, acc3); acc3 = __dp4a(w3, E, acc3);
    }
    const uint32_t out = pack4(acc0 >> 8, acc1 >> 8, acc2 >> 8, acc3 >> 8);
    const int nv = min(4, D.w - px);                    // valid pixels of the quad (the rest would be margin positions)

    uint8_t* lv = base + D.off;
    // does any pixel of the quad have mirror images among the margin columns?  (x <= P on the left, x >= w-1-P on the right)
    const bool colwork = px <= P || px + nv - 1 >= D.w - 1 - P;
    if (P <= D.w - 1 && P <= D.h - 1) {
        // the usual case: the margin is narrower than the level, every pixel has at most one mirror image per side
        auto put_row = [&](int t) {
            uint8_t* drow = lv + (long long)t * D.pitch;
            if (nv == 4) *reinterpret_cast<uint32_t*>(drow + px) = out;
            else
                for (int i = 0; i < nv; ++i) drow[px + i] = (uint8_t)(out >> (8 * i));
            if (colwork) {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int x = px + i;
                    const uint8_t v = (uint8_t)(out >> (8 * i));
                    if (x >= 1 && x <= P) drow[-x] = v;
                    if (x >= D.w - 1 - P && x <= D.w - 2) drow[2 * (D.w - 1) - x] = v;
                }
            }
        };
        put_row(py);
        if (py >= 1 && py <= P) put_row(-py);
        if (py >= D.h - 1 - P && py <= D.h - 2) put_row(2 * (D.h - 1) - py);
        return;
    }
    // tiny levels: all rows t in [-P, h+P) with reflect101(t) == py are t = +-py + k (2h - 2); the same for columns
    const int perh = 2 * D.h - 2, perw = 2 * D.w - 2;
#pragma unroll 1
    for (int sgn = 0; sgn < 2; ++sgn) {
        if (sgn == 1 && (py == 0 || py == D.h - 1)) break;          // -py is the same residue as py
#pragma unroll 1
        for (int t = first_at_least(sgn ? -py : py, perh, -P); t < D.h + P; t += perh) {
            uint8_t* drow = lv + (long long)t * D.pitch;
            for (int i = 0; i < nv; ++i) {
                const int x = px + i;
                const uint8_t v = (uint8_t)(out >> (8 * i));
                for (int sg = 0; sg < 2; ++sg) {
                    if (sg == 1 && (x == 0 || x == D.w - 1)) break;
                    for (int c = first_at_least(sg ? -x : x, perw, -P); c < D.w + P; c += perw) drow[c] = v;
                }
            }
        }
    }
}


// ---------------------------------------------------------------------------------------------------------------
// Fused build of levels 0 .. NL-1 (NL <= 4) in ONE launch: a CTA owns a 64x64 tile of level 0 and the 32x32 / 16x16 / 8x8
// tiles of levels 1 / 2 / 3 below it.  It stages the tile of the raw frame with the halo the whole chain needs (14 px left /
// top, 7 px right / bottom for four levels: 85x85 bytes, BORDER_REFLECT_101 applied while loading), writes its part of the
// padded level 0, and produces each further level from the previous one in shared memory (same DP4A formulation as
// pyrdown_kernel), writing the owned tile of every level together with the margin positions that reflect onto it.  Replaces
// ingest_kernel + ingest_edge_kernel + up to three pyrdown_kernel launches: the raw frame is read once (plus halo, from L2),
// every level is written once, and the five dependent sub-wave launches of a frame become one.
// Halo slots that lie outside a level hold that level's BORDER_REFLECT_101 value: computing them like inside pixels is only
// right when the level's width is odd, so they are overwritten with the value of the inside pixel they mirror.
// ---------------------------------------------------------------------------------------------------------------
struct FusedPyrArgs {
    const uint8_t* raw;
    size_t raw_pitch, raw_lane;
    uint8_t* pyr;
    long long lane_bytes;
    LevelGeom lv[4];
    int P;
};

constexpr int kFT = 64;   // level-0 tile edge of the fused kernel

template <int NL>
struct FusedGeo {
    // halo of level l: left / top HL, right / bottom HR (level l+1's region [a, b] needs [2a - 2, 2b + 2] of level l)
    __host__ __device__ static constexpr int hl(int l) { return l >= NL - 1 ? 0 : 2 * hl(l + 1) + 2; }
    __host__ __device__ static constexpr int hr(int l) { return l >= NL - 1 ? 0 : 2 * hr(l + 1) + 1; }
    __host__ __device__ static constexpr int n(int l) { return kFT >> l; }
    __host__ __device__ static constexpr int rw(int l) { return n(l) + hl(l) + hr(l); }                  // region edge (square)
    __host__ __device__ static constexpr int rq(int l) { return (rw(l) + 3) / 4; }                       // output quads per region row
    // row pitch in bytes: the rounded-up quads of this level, and the words the next level's last quad reads
    __host__ __device__ static constexpr int pb(int l) { return l >= NL - 1 ? 4 * rq(l) : (4 * rq(l) > 8 * rq(l + 1) + 4 ? 4 * rq(l) : 8 * rq(l + 1) + 4); }
    __host__ __device__ static constexpr int words(int l) { return rw(l) * pb(l) / 4; }
    __host__ __device__ static constexpr int off(int l) { return l == 0 ? 0 : off(l - 1) + words(l - 1); }
    static constexpr int total_words = off(NL - 1) + words(NL - 1);
};

// owned tile of level D (n x n, origin (x0, y0)) from its shared-memory region -> the padded level in global memory, every
// pixel also at the margin positions that mirror it (the usual case of pyrdown_kernel: margin narrower than the level)
template <int N, int HL, int PB>
__device__ __forceinline__ void fused_emit(const uint32_t* __restrict__ S, uint8_t* __restrict__ lv, const LevelGeom D, int P,
                                           int x0, int y0, int tid)
{
    constexpr int QW = N / 4;
    for (int idx = tid; idx < N * QW; idx += 256) {
        const int v = idx / QW, u = idx - v * QW;
        const int px = x0 + 4 * u, py = y0 + v;
        if (py >= D.h || px >= D.w) continue;
        const uint32_t* r = S + (HL + v) * (PB / 4) + (HL + 4 * u) / 4;
        const uint32_t out = (HL & 3) ? __funnelshift_r(r[0], r[1], 8 * (HL & 3)) : r[0];
        const int nv = min(4, D.w - px);
        const bool colwork = px <= P || px + nv - 1 >= D.w - 1 - P;
        auto put_row = [&](int t) {
            uint8_t* drow = lv + (long long)t * D.pitch;
            if (nv == 4) *reinterpret_cast<uint32_t*>(drow + px) = out;
            else
                for (int i = 0; i < nv; ++i) drow[px + i] = (uint8_t)(out >> (8 * i));
            if (colwork) {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int x = px + i;
                    const uint8_t b = (uint8_t)(out >> (8 * i));
                    if (i < nv && x >= 1 && x <= P) drow[-x] = b;
                    if (i < nv && x >= D.w - 1 - P && x <= D.w - 2) drow[2 * (D.w - 1) - x] = b;
                }
            }
        };
        put_row(py);
        if (py >= 1 && py <= P) put_row(-py);
        if (py >= D.h - 1 - P && py <= D.h - 2) put_row(2 * (D.h - 1) - py);
    }
}

// level l (region in S, pitch PBS bytes) -> level l+1 region (RWN x RWN outputs, pitch PBD bytes): slot (i, j) of the next
// level takes rows 2i .. 2i+4 and columns 2j .. 2j+4 of this one
template <int RWN, int PBS, int PBD>
__device__ __forceinline__ void fused_down(const uint32_t* __restrict__ S, uint32_t* __restrict__ Dn, int tid)
{
    constexpr int Q = (RWN + 3) / 4;
    for (int idx = tid; idx < RWN * Q; idx += 256) {
        const int i = idx / Q, q = idx - i * Q;
        uint32_t acc0 = 128, acc1 = 128, acc2 = 128, acc3 = 128;
        const uint32_t* r0 = S + (2 * i) * (PBS / 4) + 2 * q;
#pragma unroll
        for (int j = 0; j < 5; ++j) {
            const uint32_t k = j == 0 || j == 4 ? 1 : j == 2 ? 6 : 4;
            const uint32_t* r = r0 + j * (PBS / 4);
            const uint32_t w0 = r[0], w1 = r[1], w2 = r[2];
            const uint32_t Cw = k | ((4 * k) << 8) | ((6 * k) << 16) | ((4 * k) << 24);   // (1, 4, 6, 4) k
            const uint32_t E = k;                                                      // (1, 0, 0, 0) k
            const uint32_t A = (k << 16) | ((4 * k) << 24);                            // (0, 0, 1, 4) k
            const uint32_t B = (6 * k) | ((4 * k) << 8) | (k << 16);                   // (6, 4, 1, 0) k
            acc0 = __dp4a(w0, Cw, acc0); acc0 = __dp4a(w1, E, acc0);
            acc1 = __dp4a(w0, A, acc1); acc1 = __dp4a(w1, B, acc1);
            acc2 = __dp4a(w1, Cw, acc2); acc2 = __dp4a(w2, E, acc2);
            acc3 = __dp4a(w1, A, acc3); acc3 = __dp4a(w2, B, acc3);
        }
        Dn[i * (PBD / 4) + q] = pack4(acc0 >> 8, acc1 >> 8, acc2 >> 8, acc3 >> 8);
    }
}

// slots of a region (origin (lox, loy), RW x RW, pitch PB) that lie outside the level take the value of the inside pixel
// BORDER_REFLECT_101 maps them to: first the outside columns of every row, then (after a barrier) the outside rows as copies
// of whole rows, which brings the corners along.  Only CTAs at the level's border have any (CTA-uniform test).
template <int RW, int PB>
__device__ __forceinline__ void fused_reflect(uint32_t* __restrict__ S, const LevelGeom L, int lox, int loy, int tid)
{
    const int nl = max(-lox, 0), nr = max(lox + RW - L.w, 0);      // outside columns on the left / right
    const int nt = max(-loy, 0), nb = max(loy + RW - L.h, 0);      // outside rows above / below
    if ((nl | nr | nt | nb) == 0) return;
    uint8_t* B = reinterpret_cast<uint8_t*>(S);
    const int tx = tid & 15, ty = tid >> 4;
    if (nl | nr) {
        for (int i = ty; i < RW; i += 16)
            for (int c = tx; c < nl + nr; c += 16) {
                const int j = c < nl ? c : RW - nr + (c - nl);
                const int rj = min(max(reflect101(lox + j, L.w) - lox, 0), RW - 1);
                B[i * PB + j] = B[i * PB + rj];
            }
        __syncthreads();
    }
    if (nt | nb) {
        for (int r = ty; r < nt + nb; r += 16) {
            const int i = r < nt ? r : RW - nb + (r - nt);
            const int ri = min(max(reflect101(loy + i, L.h) - loy, 0), RW - 1);
            for (int j = tx; j < RW; j += 16) B[i * PB + j] = B[ri * PB + j];
        }
    }
}

// owned tile (N x N, N a multiple of 16) -> the padded level, 16 bytes per thread and row: one aligned 128-bit store for
// the interior (tile origins and level rows are 16 B-aligned), bytes only at a ragged right edge and for the mirror columns
template <int N, int HL, int PB>
__device__ __forceinline__ void fused_emit16(const uint32_t* __restrict__ S, uint8_t* __restrict__ lv, const LevelGeom D, int P,
                                             int x0, int y0, int tid)
{
    constexpr int CW = N / 16;
    for (int idx = tid; idx < N * CW; idx += 256) {
        const int v = idx / CW, c = idx - v * CW;
        const int px = x0 + 16 * c, py = y0 + v;
        if (py >= D.h || px >= D.w) continue;
        const uint32_t* r = S + (HL + v) * (PB / 4) + HL / 4 + 4 * c;
        uint32_t o[4];
        if (HL & 3) {
            const uint32_t w0 = r[0], w1 = r[1], w2 = r[2], w3 = r[3], w4 = r[4];
            o[0] = __funnelshift_r(w0, w1, 8 * (HL & 3)); o[1] = __funnelshift_r(w1, w2, 8 * (HL & 3));
            o[2] = __funnelshift_r(w2, w3, 8 * (HL & 3)); o[3] = __funnelshift_r(w3, w4, 8 * (HL & 3));
        } else { o[0] = r[0]; o[1] = r[1]; o[2] = r[2]; o[3] = r[3]; }
        const int nv = min(16, D.w - px);
        const bool colwork = px <= P || px + nv - 1 >= D.w - 1 - P;
        auto put_row = [&](int t) {
            uint8_t* drow = lv + (long long)t * D.pitch;
            if (nv == 16) *reinterpret_cast<uint4*>(drow + px) = make_uint4(o[0], o[1], o[2], o[3]);
            else
                for (int i = 0; i < nv; ++i) drow[px + i] = (uint8_t)(o[i >> 2] >> (8 * (i & 3)));
            if (colwork) {
                for (int i = 0; i < nv; ++i) {
                    const int x = px + i;
                    const uint8_t b = (uint8_t)(o[i >> 2] >> (8 * (i & 3)));
                    if (x >= 1 && x <= P) drow[-x] = b;
                    if (x >= D.w - 1 - P && x <= D.w - 2) drow[2 * (D.w - 1) - x] = b;
                }
            }
        };
        put_row(py);
        if (py >= 1 && py <= P) put_row(-py);
        if (py >= D.h - 1 - P && py <= D.h - 2) put_row(2 * (D.h - 1) - py);
    }
}

template <int NL>
__global__ void __launch_bounds__(256) pyramid_fused_kernel(const FusedPyrArgs a)
{
    using G = FusedGeo<NL>;
    __shared__ __align__(16) uint32_t smem[G::total_words];
    const int tid = threadIdx.x;
    const int lane = blockIdx.z;
    const uint8_t* src = a.raw + (size_t)lane * a.raw_lane;
    uint8_t* base = a.pyr + (size_t)lane * a.lane_bytes;
    const int P = a.P;

    // ---- level-0 region, BORDER_REFLECT_101 applied while loading.  A warp takes a region row at a time (lanes = words);
    //      CTAs whose region lies inside the frame (CTA-uniform) load two aligned words + a funnel shift per word, nothing else
    {
        const LevelGeom L = a.lv[0];
        constexpr int PW = G::pb(0) / 4;
        static_assert(PW < 32, "a region row (+ one word) is one warp pass");
        const int lox = blockIdx.x * kFT - G::hl(0), loy = blockIdx.y * kFT - G::hl(0);
        const int warp = tid >> 5, ln = tid & 31;
        const bool inside = lox >= 0 && loy >= 0 && lox + 4 * PW + 4 <= L.w && loy + G::rw(0) <= L.h;
        {
            if (inside) {
                // one aligned word per lane (lanes 0 .. PW: the row's PW words + the one the last funnel shift needs), all
                // rows of this warp in flight before the first use; the neighbour's word comes by shuffle
                constexpr int RMAX = (G::rw(0) + 7) / 8;
                uint32_t v[RMAX];
#pragma unroll
                for (int k = 0; k < RMAX; ++k) {
                    const int r = min(warp + 8 * k, G::rw(0) - 1);
                    const uintptr_t addr = reinterpret_cast<uintptr_t>(src + (size_t)(loy + r) * a.raw_pitch + lox);
                    v[k] = ln <= PW ? __ldg(reinterpret_cast<const uint32_t*>(addr & ~(uintptr_t)3) + ln) : 0u;
                }
#pragma unroll
                for (int k = 0; k < RMAX; ++k) {
                    const int r = warp + 8 * k;
                    const int sh = 8 * (int)(reinterpret_cast<uintptr_t>(src + (size_t)(loy + min(r, G::rw(0) - 1)) * a.raw_pitch + lox) & 3);
                    const uint32_t nxt = __shfl_down_sync(0xffffffffu, v[k], 1);
                    if (r < G::rw(0) && ln < PW) smem[r * PW + ln] = __funnelshift_r(v[k], nxt, sh);
                }
            } else if (ln < PW) {
                // CTA at the frame border.  Whether a word needs BORDER_REFLECT_101 depends on its columns only, so a lane
                // is "inside" or "border" for all its rows: inside lanes load two aligned words per row, border lanes four
                // reflected bytes; either way every load of the lane is in flight before the first use (rows in two halves)
                constexpr int RMAX = (G::rw(0) + 7) / 8, HALF = (RMAX + 1) / 2;
                const int gx = lox + 4 * ln;
                const bool word_inside = gx >= 0 && gx + 7 < L.w;
                int cx[4];
#pragma unroll
                for (int b = 0; b < 4; ++b) cx[b] = reflect101(gx + b, L.w);
#pragma unroll
                for (int h0 = 0; h0 < RMAX; h0 += HALF) {
                    uint32_t v[HALF];
                    if (word_inside) {
                        uint32_t lo[HALF], hi[HALF];
#pragma unroll
                        for (int k = 0; k < HALF; ++k) {
                            const int r = min(warp + 8 * (h0 + k), G::rw(0) - 1);
                            const uintptr_t addr = reinterpret_cast<uintptr_t>(src + (size_t)reflect101(loy + r, L.h) * a.raw_pitch + gx);
                            const uint32_t* wp = reinterpret_cast<const uint32_t*>(addr & ~(uintptr_t)3);
                            lo[k] = __ldg(wp);
                            hi[k] = __ldg(wp + 1);
                        }
#pragma unroll
                        for (int k = 0; k < HALF; ++k) {
                            const int r = min(warp + 8 * (h0 + k), G::rw(0) - 1);
                            const uintptr_t addr = reinterpret_cast<uintptr_t>(src + (size_t)reflect101(loy + r, L.h) * a.raw_pitch + gx);
                            v[k] = __funnelshift_r(lo[k], hi[k], 8 * (int)(addr & 3));
                        }
                    } else {
                        uint32_t bb[HALF][4];
#pragma unroll
                        for (int k = 0; k < HALF; ++k) {
                            const int r = min(warp + 8 * (h0 + k), G::rw(0) - 1);
                            const uint8_t* row = src + (size_t)reflect101(loy + r, L.h) * a.raw_pitch;
#pragma unroll
                            for (int b = 0; b < 4; ++b) bb[k][b] = __ldg(row + cx[b]);
                        }
#pragma unroll
                        for (int k = 0; k < HALF; ++k) v[k] = bb[k][0] | (bb[k][1] << 8) | (bb[k][2] << 16) | (bb[k][3] << 24);
                    }
#pragma unroll
                    for (int k = 0; k < HALF; ++k) {
                        const int r = warp + 8 * (h0 + k);
                        if (h0 + k < RMAX && r < G::rw(0)) smem[r * PW + ln] = v[k];
                    }
                }
            }
        }
    }
    __syncthreads();
    fused_emit16<G::n(0), G::hl(0), G::pb(0)>(smem, base + a.lv[0].off, a.lv[0], P, blockIdx.x * kFT, blockIdx.y * kFT, tid);
    if constexpr (NL >= 2) {
        fused_down<G::rw(1), G::pb(0), G::pb(1)>(smem + G::off(0), smem + G::off(1), tid);
        __syncthreads();
        fused_reflect<G::rw(1), G::pb(1)>(smem + G::off(1), a.lv[1], blockIdx.x * G::n(1) - G::hl(1), blockIdx.y * G::n(1) - G::hl(1), tid);
        __syncthreads();
        fused_emit16<G::n(1), G::hl(1), G::pb(1)>(smem + G::off(1), base + a.lv[1].off, a.lv[1], P, blockIdx.x * G::n(1), blockIdx.y * G::n(1), tid);
    }
    if constexpr (NL >= 3) {
        fused_down<G::rw(2), G::pb(1), G::pb(2)>(smem + G::off(1), smem + G::off(2), tid);
        __syncthreads();
        fused_reflect<G::rw(2), G::pb(2)>(smem + G::off(2), a.lv[2], blockIdx.x * G::n(2) - G::hl(2), blockIdx.y * G::n(2) - G::hl(2), tid);
        __syncthreads();
        fused_emit<G::n(2), G::hl(2), G::pb(2)>(smem + G::off(2), base + a.lv[2].off, a.lv[2], P, blockIdx.x * G::n(2), blockIdx.y * G::n(2), tid);
    }
    if constexpr (NL >= 4) {
        fused_down<G::rw(3), G::pb(2), G::pb(3)>(smem + G::off(2), smem + G::off(3), tid);
        __syncthreads();
        fused_emit<G::n(3), G::hl(3), G::pb(3)>(smem + G::off(3), base + a.lv[3].off, a.lv[3], P, blockIdx.x * G::n(3), blockIdx.y * G::n(3), tid);
    }
}

void launch_ingest(cudaStream_t st, const uint8_t* raw, size_t raw_pitch, size_t raw_lane, uint8_t* pyr,
                   const PyrGeom& g, int batch, long long* launches)
{
    const LevelGeom& L = g.lv[0];
    const int P = g.P;                                        // a multiple of 16
    const int n_chunks = (L.w + 2 * P + 15) / 16;             // the row pitch (a multiple of 128) covers the last chunk
    // streaming chunks: x0 = 16 c - P >= 0 and x0 + 20 <= w
    const int c_begin = P / 16;
    const int c_end = L.w >= 20 ? std::max(c_begin, (L.w - 20 + P) / 16 + 1) : c_begin;
    const int rows = L.h + 2 * P;
    if (c_end > c_begin) {
        dim3 grid(((c_end - c_begin) * rows + 255) / 256, 1, batch);
        ingest_kernel<<<grid, 256, 0, st>>>(raw, raw_pitch, raw_lane, pyr, g.lane_bytes, L, P, c_begin, c_end);
        if (launches) ++*launches;
    }
    const int n_edge = c_begin + (n_chunks - c_end);
    dim3 egrid((rows + 127) / 128, n_edge, batch);
    ingest_edge_kernel<<<egrid, 128, 0, st>>>(raw, raw_pitch, raw_lane, pyr, g.lane_bytes, L, P, c_begin, c_end, n_chunks);
    if (launches) ++*launches;
}

// ---- BGR -> gray (frame ingest of a colour sequence).  The reference's loader does cv::imread + cv::cvtColor(...,
// COLOR_BGR2GRAY) on the host (include/async_image_loader.h:66-69); here the interleaved CV_8UC3 image goes over PCIe as it
// is and the conversion runs on the way into the slot.  OpenCV's 8-bit fixed point, bit for bit:
//   gray = (B 3735 + G 19235 + R 9798 + 2^14) >> 15            (oracle/svo_oracle.c:orc_bgr2gray, pinned against cv2)
// HBM-bound: 3 bytes read + 1 written per pixel.  A thread makes 4 gray bytes from 12 source bytes: aligned 32-bit loads
// + funnel shifts (any source alignment), consecutive lanes read consecutive 12-byte groups.
__device__ __forceinline__ uint32_t gray15(uint32_t b, uint32_t g, uint32_t r)
{
    return (b * 3735u + g * 19235u + r * 9798u + (1u << 14)) >> 15;
}

__global__ void __launch_bounds__(128) bgr2gray_kernel(const uint8_t* __restrict__ src, size_t src_pitch, size_t src_lane,
                                                       uint8_t* __restrict__ dst, size_t dst_pitch, size_t dst_lane, int w)
{
    const int q = blockIdx.x * 128 + threadIdx.x;   // 4-pixel group of the row
    const int x = 4 * q;
    if (x >= w) return;
    const uint8_t* s = src + (size_t)blockIdx.z * src_lane + (size_t)blockIdx.y * src_pitch + 3 * (size_t)x;
    uint32_t* d = reinterpret_cast<uint32_t*>(dst + (size_t)blockIdx.z * dst_lane + (size_t)blockIdx.y * dst_pitch) + q;
    uint32_t out;
    if (x + 4 <= w) {
        const uintptr_t addr = reinterpret_cast<uintptr_t>(s);
        const uint32_t* wp = reinterpret_cast<const uint32_t*>(addr & ~(uintptr_t)3);
        const int sh = 8 * (int)(addr & 3);
        const uint32_t w0 = __ldg(wp), w1 = __ldg(wp + 1), w2 = __ldg(wp + 2);
        const uint32_t w3 = sh ? __ldg(wp + 3) : 0u;   // holds a needed byte exactly when the group is not word aligned
        const uint32_t a = __funnelshift_r(w0, w1, sh), b = __funnelshift_r(w1, w2, sh), c = __funnelshift_r(w2, w3, sh);
        out = gray15(a & 255u, (a >> 8) & 255u, (a >> 16) & 255u) |
              gray15(a >> 24, b & 255u, (b >> 8) & 255u) << 8 |
              gray15((b >> 16) & 255u, b >> 24, c & 255u) << 16 |
              gray15((c >> 8) & 255u, (c >> 16) & 255u, c >> 24) << 24;
    } else {   // ragged end of the row: byte loads, nothing beyond the last pixel is touched
        out = 0;
        for (int i = 0; x + i < w; ++i) out |= gray15(s[3 * i], s[3 * i + 1], s[3 * i + 2]) << (8 * i);
    }
    *d = out;   // the destination pitch is a multiple of 4 >= w rounded up, so the padding bytes of the last word exist
}

void launch_bgr2gray(cudaStream_t st, const uint8_t* src, size_t src_pitch, size_t src_lane, uint8_t* dst, size_t dst_pitch,
                     size_t dst_lane, int w, int h, int batch, long long* launches)
{
    dim3 grid(((w + 3) / 4 + 127) / 128, h, batch);
    bgr2gray_kernel<<<grid, 128, 0, st>>>(src, src_pitch, src_lane, dst, dst_pitch, dst_lane, w);
    if (launches) ++*launches;
}

int launch_pyramid_fused(cudaStream_t st, const uint8_t* raw, size_t raw_pitch, size_t raw_lane, uint8_t* pyr,
                         const PyrGeom& g, int batch, long long* launches)
{
    const int nl = std::min(g.nlevels, 4);
    for (int l = 0; l < nl; ++l)   // the margin must be narrower than every fused level (one mirror image per side)
        if (g.P > g.lv[l].w - 1 || g.P > g.lv[l].h - 1) return 0;
    FusedPyrArgs a{};
    a.raw = raw; a.raw_pitch = raw_pitch; a.raw_lane = raw_lane; a.pyr = pyr; a.lane_bytes = g.lane_bytes; a.P = g.P;
    for (int l = 0; l < nl; ++l) a.lv[l] = g.lv[l];
    dim3 grid((g.lv[0].w + kFT - 1) / kFT, (g.lv[0].h + kFT - 1) / kFT, batch);
    switch (nl) {
    case 1: pyramid_fused_kernel<1><<<grid, 256, 0, st>>>(a); break;
    case 2: pyramid_fused_kernel<2><<<grid, 256, 0, st>>>(a); break;
    case 3: pyramid_fused_kernel<3><<<grid, 256, 0, st>>>(a); break;
    default: pyramid_fused_kernel<4><<<grid, 256, 0, st>>>(a); break;
    }
    if (launches) ++*launches;
    return nl;
}

void launch_pyrdown(cudaStream_t st, uint8_t* pyr, const PyrGeom& g, int level, int batch, long long* launches)
{
    const LevelGeom& S = g.lv[level - 1];
    const LevelGeom& D = g.lv[level];
    const int quads = (D.w + 3) / 4;
    dim3 grid((quads * D.h + 255) / 256, 1, batch);
    pyrdown_kernel<<<grid, 256, 0, st>>>(pyr, g.lane_bytes, S, D, g.P, quads);
    if (launches) ++*launches;
}

}  // namespace svo
