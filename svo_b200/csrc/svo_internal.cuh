// svo_internal.cuh -- shared declarations of the sm_100a front-end kernels (not part of the public ABI).
//
// Layout of one image "slot" in HBM (per lane, i.e. per independent stream):
//   raw      : the uploaded 8-bit image, row pitch raw_pitch (or an alias of caller-owned device memory)
//   pyramid  : levels 0..L, each stored with a BORDER_REFLECT_101 margin of P pixels on every side so that the LK
//              kernel reads (win+3)x(win+3) regions with unconditional aligned 32-bit loads.  Rows are 128 B-pitched;
//              P is a multiple of 16, so interior row starts are 16 B aligned.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>

#include "svo_gpu.h"

namespace svo {

constexpr int kMaxLevels = SVO_GPU_MAX_LEVELS;

struct LevelGeom {
    int w, h;        // interior size
    int pitch;       // bytes per padded row
    int pad_;
    long long off;   // byte offset of interior pixel (0,0) from the lane's pyramid base
};

struct PyrGeom {
    int nlevels;     // number of valid levels (effective maxLevel + 1)
    int P;           // margin in pixels
    LevelGeom lv[kMaxLevels];
    long long lane_bytes;   // stride between lanes
};

// Function attributes (carve-out, dynamic shared-memory limit) are per device: true the first time a launch site runs on
// the calling thread's current device (one host thread per GPU may share this library).
inline bool first_on_device(std::atomic<unsigned long long>& seen)
{
    int dev = 0;
    cudaGetDevice(&dev);
    const unsigned long long bit = 1ull << (dev & 63);
    return (seen.fetch_or(bit) & bit) == 0;
}

// Kernels that may share an SM with the persistent LK kernel (the step runs FAST and bucketing beside it) all ask for
// the same, largest shared-memory carve-out: CTAs of kernels with different carve-outs cannot be co-resident on an SM.
template <class Kernel>
inline void prefer_max_shared(Kernel k)
{
    cudaFuncSetAttribute(k, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
}

// REFLECT_101 index map (cv::BORDER_REFLECT_101), valid for any i when n >= 2 (loops for tiny n), n == 1 -> 0
__host__ __device__ inline int reflect101(int i, int n)
{
    if (n == 1) return 0;
    while (i < 0 || i >= n) i = i < 0 ? -i : 2 * n - 2 - i;
    return i;
}

// ------------------------------------------------------------------------------------------------ pyramid.cu
// raw image (pitch raw_pitch, lane stride raw_lane) -> padded level 0 of every lane
void launch_ingest(cudaStream_t st, const uint8_t* raw, size_t raw_pitch, size_t raw_lane, uint8_t* pyr,
                   const PyrGeom& g, int batch, long long* launches);
// raw image -> padded levels 0 .. min(nlevels, 4) - 1 of every lane in one launch; returns the number of levels it built
// (0: geometry not covered -- a margin wider than a level -- use launch_ingest + launch_pyrdown)
int launch_pyramid_fused(cudaStream_t st, const uint8_t* raw, size_t raw_pitch, size_t raw_lane, uint8_t* pyr,
                         const PyrGeom& g, int batch, long long* launches);
// interleaved BGR (CV_8UC3) -> gray rows of pitch dst_pitch (a multiple of 4, >= w rounded up to 4), every lane
void launch_bgr2gray(cudaStream_t st, const uint8_t* src, size_t src_pitch, size_t src_lane, uint8_t* dst, size_t dst_pitch,
                     size_t dst_lane, int w, int h, int batch, long long* launches);
// level l-1 -> level l (5x5 Gaussian pyrDown, margins included), every lane
void launch_pyrdown(cudaStream_t st, uint8_t* pyr, const PyrGeom& g, int level, int batch, long long* launches);

// ------------------------------------------------------------------------------------------------ fast.cu
constexpr int kFastTW = 128;  // tile width: one (row, tile) segment holds at most 128 keypoints
#ifndef SVO_FAST_TH
#define SVO_FAST_TH 32   // tile height: 64 measured against 32 in profiles/r02_s_fast_tiles.txt
#endif
constexpr int kFastTH = SVO_FAST_TH;

struct FastBufs {
    // per lane strides are in elements of the respective type
    int* rowtotal;        // [B][rows_cap]           keypoints per image row (atomic counters, zeroed per call)
    uint8_t* segcnt;      // [B][rows_cap * ntx_cap] keypoints per (row, tile) segment
    uint16_t* segkp;      // [B][rows_cap * ntx_cap * kFastTW] x_local | best << 8, in x order inside a segment
    int rows_cap, ntx_cap;
    float2* kp_xy;        // [B][kp_cap]
    float* kp_resp;       // [B][kp_cap]
    int* kp_count;        // [B]  total keypoints found (may exceed kp_cap; only the first kp_cap are stored)
    int kp_cap;
};

// The reference's mask given by the points it is drawn around (src/tracking.cpp:76-79: 255 everywhere, 0 in the rounded
// inclusive box of +-half px around each point): evaluated per tile inside the FAST kernel, never materialised.
// pts: [B][pts_lane] float2, n per lane = min(n_dev[lane], n_max); pts == nullptr or n_max <= 0: no mask.
struct FastMaskPts {
    const float2* pts = nullptr;
    size_t pts_lane = 0;
    const int* n_dev = nullptr;
    int n_max = 0;
    float half = 0.f;
    const int* lane_active = nullptr;   // [B] or nullptr: lanes whose flag is 0 are skipped altogether (they report 0 keypoints)
};

// mask: device u8 plane per lane (pitch mask_pitch, lane stride mask_lane) or nullptr; mask_pts: see above (either, or none)
void launch_fast(cudaStream_t st, const uint8_t* raw, size_t raw_pitch, size_t raw_lane, int w, int h, int threshold,
                 int nms, const uint8_t* mask, size_t mask_pitch, size_t mask_lane, const FastBufs& fb, int batch,
                 long long* launches, int small_ctas = -1, const FastMaskPts* mask_pts = nullptr);

// ------------------------------------------------------------------------------------------------ bucket.cu
struct BucketBufs {
    int* cell_slots;      // [B][cells_cap * k_cap]  the k smallest input indices per bucket (INT_MAX = empty)
    int* cell_cnt;        // [B][cells_cap]          arrivals per bucket
    int* cell_last;       // [B][cells_cap]          largest input index per bucket
    int* err_flag;        // [B]                     set when a point maps outside the bucket table
    size_t cells_cap;     // per-lane capacity of cell_cnt / cell_last
    size_t slots_cap;     // per-lane capacity of cell_slots
    float2* out_xy;       // [B][out_cap]
    int* out_age;         // [B][out_cap]
    int* out_count;       // [B]  required output count (may exceed out_cap)
    int out_cap;
    int mode;             // 0: pick per call, 1: always the three global-memory kernels, 2: the one-CTA-per-lane kernel whenever its table fits
};

// in_xy: [B][in_lane] float2; in_age may be nullptr (all ages 0, FeatureSet::appendNewFeatures, src/bucket.cpp:18-22);
// n per lane = min(n_dev[lane], n_max)
void launch_bucket(cudaStream_t st, const float2* in_xy, const int* in_age, size_t in_lane, const int* n_dev, int n_max,
                   int img_w, int img_h, int bucket_size, int per_bucket, const BucketBufs& bb, int batch,
                   long long* launches, int n_hint = 0);

// ------------------------------------------------------------------------------------------------ klt.cu
struct KltArgs {
    const uint8_t* prev_pyr;   // lane 0 base of the prev slot's pyramid
    const uint8_t* next_pyr;
    PyrGeom g;                 // same geometry for both slots
    const float2* prev_xy;     // [B][pts_lane]
    float2* next_xy;           // [B][pts_lane]
    uint8_t* status;           // [B][pts_lane]
    float* err;                // [B][pts_lane]
    size_t pts_lane;
    const int* n_dev;          // [B] points per lane (device), clamped to n_max
    int n_max;
    int win_w, win_h;
    int max_level;             // effective
    int max_count;             // already clamped like OpenCV
    double eps2;               // epsilon^2, already clamped
    int flags;
    double min_eig_threshold;
    unsigned long long* iters; // optional global iteration counter
    unsigned* work;            // track hand-out counter of the persistent variant (zeroed by the launcher)
};
void launch_klt(cudaStream_t st, const KltArgs& a, int batch, long long* launches);
// tracks the LK kernel of this window keeps resident at once on the current device (one one-warp CTA per track)
int klt_resident_tracks(int win_w, int win_h);

// ------------------------------------------------------------------------------------------------ filter.cu
struct FilterArgs {
    const float2* a_xy;        // [B][pts_lane] points LK started from (left / previous)
    const float2* b_xy;        // [B][pts_lane] points LK found (right / current)
    const uint8_t* status;     // [B][pts_lane]
    size_t pts_lane;
    const int* n_dev;          // [B] list length per lane (device), clamped to n_max
    int n_max;
    int use_y;                 // 1: also require |b.y - a.y| < y_threshold (src/tracking.cpp:110)
    float y_threshold;
    const float* err;          // [B][pts_lane] LK err (may be null)
    float* out_err;            // [B][pts_lane] err of each survivor (may be null)
    int* out_index;            // [B][pts_lane] input position of each survivor (may be null)
    float2* out_a;             // [B][pts_lane] (may be null)
    float2* out_b;             // [B][pts_lane] (may be null)
    int out_cap;
    int* out_n;                // [B] survivors per lane
};
void launch_filter(cudaStream_t st, const FilterArgs& a, int batch, long long* launches);

}  // namespace svo
