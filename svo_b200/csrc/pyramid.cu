// pyramid.cu -- K1: image ingest (REFLECT_101 margin) and 5x5 Gaussian pyrDown chain for sm_100a.
//
// Same arithmetic as cv::buildOpticalFlowPyramid(img, pyr, winSize, maxLevel, withDerivatives=false), which
// cv::calcOpticalFlowPyrLK runs on both images at every call (reference call sites src/tracking.cpp:101-105, :160-165):
//   level l = pyrDown(level l-1):  dst(x,y) = (sum_{i,j=-2..2} w_i w_j src(R(2x+i), R(2y+j)) + 128) >> 8, w = [1 4 6 4 1],
//   R = BORDER_REFLECT_101, dst size ((w+1)/2, (h+1)/2); every level carries a REFLECT_101 margin.
// HBM-bound streaming stencils: each thread produces 4 output bytes from aligned 32-bit loads; because the source level
// already carries its reflected margin, no thread needs border logic for its reads.
#include "svo_internal.cuh"

namespace svo {

// ---------------------------------------------------------------------------------------------------------------
// raw -> padded level 0.  One thread = 4 consecutive bytes of one padded row (one aligned 32-bit store).
// grid = (ceil((w+2P)/4 / 128), h+2P, batch)
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) ingest_kernel(const uint8_t* __restrict__ raw, size_t raw_pitch, size_t raw_lane,
                                                     uint8_t* __restrict__ pyr, long long lane_bytes, LevelGeom L, int P)
{
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    const int px = q * 4 - P;
    if (px >= L.w + P) return;
    const int py = (int)blockIdx.y - P;
    const int lane = blockIdx.z;
    const int sy = reflect101(py, L.h);
    const uint8_t* srow = raw + (size_t)lane * raw_lane + (size_t)sy * raw_pitch;
    uint32_t v;
    if (px >= 4 && px + 7 < L.w) {
        // two aligned words + funnel shift: rows of any pitch / alignment (compact uploads have pitch = width); the eight
        // bytes read lie inside this row
        const uintptr_t addr = reinterpret_cast<uintptr_t>(srow + px);
        const uint32_t* wp = reinterpret_cast<const uint32_t*>(addr & ~(uintptr_t)3);
        v = __funnelshift_r(__ldg(wp), __ldg(wp + 1), 8 * (int)(addr & 3));
    } else {
        v = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) v |= (uint32_t)__ldg(srow + reflect101(px + i, L.w)) << (8 * i);
    }
    uint8_t* drow = pyr + (size_t)lane * lane_bytes + L.off + (long long)py * L.pitch;
    *reinterpret_cast<uint32_t*>(drow + px) = v;
}

// ---------------------------------------------------------------------------------------------------------------
// level l-1 -> level l, margins included.  One thread = 4 consecutive dst pixels of one padded dst row.
// Interior quads read 5 rows x 16 source bytes as aligned words; the margin quads (left/right only: rows are handled by
// reflecting the row index) fall back to per-pixel reflected reads.
// grid = (ceil((dw+2P)/4 / 128), dh+2P, batch)
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ int pyr_px(const uint8_t* __restrict__ s, int pitch, int cx, int cy)
{
    // 5x5 around (cx, cy) in a padded source level (margin >= 2 guaranteed)
    const int K[5] = {1, 4, 6, 4, 1};
    int acc = 0;
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        const uint8_t* r = s + (long long)(cy + j - 2) * pitch + cx - 2;
        int hsum = (int)r[0] + 4 * (int)r[1] + 6 * (int)r[2] + 4 * (int)r[3] + (int)r[4];
        acc += K[j] * hsum;
    }
    return (acc + 128) >> 8;
}

__global__ void __launch_bounds__(128) pyrdown_kernel(uint8_t* __restrict__ pyr, long long lane_bytes, LevelGeom S,
                                                      LevelGeom D, int P)
{
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    const int px = q * 4 - P;
    if (px >= D.w + P) return;
    const int py = (int)blockIdx.y - P;
    const int lane = blockIdx.z;
    uint8_t* base = pyr + (size_t)lane * lane_bytes;
    const uint8_t* s = base + S.off;
    const int ry = reflect101(py, D.h);
    uint32_t out;
    if (px >= 0 && px + 3 < D.w) {
        // source columns 2px-4 .. 2px+11 (aligned words), rows 2ry-2 .. 2ry+2
        int acc0 = 0, acc1 = 0, acc2 = 0, acc3 = 0;
        const int K[5] = {1, 4, 6, 4, 1};
#pragma unroll
        for (int j = 0; j < 5; ++j) {
            const uint32_t* r =
                reinterpret_cast<const uint32_t*>(s + (long long)(2 * ry + j - 2) * S.pitch + 2 * px - 4);
            const uint32_t w0 = r[0], w1 = r[1], w2 = r[2], w3 = r[3];
            // bytes b[i] = column 2px-4+i
            const int b2 = (w0 >> 16) & 255, b3 = w0 >> 24;
            const int b4 = w1 & 255, b5 = (w1 >> 8) & 255, b6 = (w1 >> 16) & 255, b7 = w1 >> 24;
            const int b8 = w2 & 255, b9 = (w2 >> 8) & 255, b10 = (w2 >> 16) & 255, b11 = w2 >> 24;
            const int b12 = w3 & 255;
            acc0 += K[j] * (b2 + 4 * b3 + 6 * b4 + 4 * b5 + b6);
            acc1 += K[j] * (b4 + 4 * b5 + 6 * b6 + 4 * b7 + b8);
            acc2 += K[j] * (b6 + 4 * b7 + 6 * b8 + 4 * b9 + b10);
            acc3 += K[j] * (b8 + 4 * b9 + 6 * b10 + 4 * b11 + b12);
        }
        out = (uint32_t)((acc0 + 128) >> 8) | ((uint32_t)((acc1 + 128) >> 8) << 8) |
              ((uint32_t)((acc2 + 128) >> 8) << 16) | ((uint32_t)((acc3 + 128) >> 8) << 24);
    } else {
        out = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int rx = reflect101(px + i, D.w);
            out |= (uint32_t)pyr_px(s, S.pitch, 2 * rx, 2 * ry) << (8 * i);
        }
    }
    uint8_t* drow = base + D.off + (long long)py * D.pitch;
    *reinterpret_cast<uint32_t*>(drow + px) = out;
}

void launch_ingest(cudaStream_t st, const uint8_t* raw, size_t raw_pitch, size_t raw_lane, uint8_t* pyr,
                   const PyrGeom& g, int batch, long long* launches)
{
    const LevelGeom& L = g.lv[0];
    const int quads = (L.w + 2 * g.P + 3) / 4;
    dim3 grid((quads + 127) / 128, L.h + 2 * g.P, batch);
    ingest_kernel<<<grid, 128, 0, st>>>(raw, raw_pitch, raw_lane, pyr, g.lane_bytes, L, g.P);
    if (launches) ++*launches;
}

void launch_pyrdown(cudaStream_t st, uint8_t* pyr, const PyrGeom& g, int level, int batch, long long* launches)
{
    const LevelGeom& S = g.lv[level - 1];
    const LevelGeom& D = g.lv[level];
    const int quads = (D.w + 2 * g.P + 3) / 4;
    dim3 grid((quads + 127) / 128, D.h + 2 * g.P, batch);
    pyrdown_kernel<<<grid, 128, 0, st>>>(pyr, g.lane_bytes, S, D, g.P);
    if (launches) ++*launches;
}

}  // namespace svo
