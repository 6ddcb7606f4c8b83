// feature.cpp -- the detection / tracking entry points of the front-end with the reference's call signatures
// (declared in svo_frontend.hpp), implemented on the C ABI of include/svo_gpu.h.  The reference's own src/feature.cpp
// holds only the Feature value struct (src/feature.cpp:6-16); its FAST and LK arithmetic are OpenCV calls made from
// src/tracking.cpp:82, :101-105 and :160-165 -- those are what this file stands in for.  No CPU fallback.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>

#include "svo_frontend.hpp"

namespace svo {

namespace {

// What an image slot of the thread's context holds.  The reference hands the SAME frame to the front-end up to three times
// -- as `next` of trackFrames (src/tracking.cpp:160-165), as `prev` of the following frame's trackFrames, and (a deep copy of
// it, :248-249) to detect and to the stereo LK (:82, :101-105) -- and each cv::calcOpticalFlowPyrLK call rebuilds both
// pyramids.  Here a frame is uploaded, and its pyramid built, once: slots are found again by CONTENT (size, type and a 64-bit
// checksum of the pixels), not by address, so a recycled cv::Mat buffer can never alias a stale frame and a deep copy hits.
struct SlotTag {
    bool valid = false;
    int rows = 0, cols = 0, type = 0;
    uint64_t sum = 0;
    // second key: the first and last (up to) 64 bytes of the frame, compared verbatim on a checksum match -- the checksum is
    // not cryptographic, and a collision would track against a stale pyramid without any sign of it
    unsigned char head[64] = {}, tail[64] = {};
    unsigned long long gen = 0;   // svo_gpu_slot_generation right after OUR upload: anyone else's upload into the slot changes it
    uint64_t used = 0;     // LRU stamp
};

struct CtxEntry {
    svo_gpu_ctx* ctx = nullptr;
    int max_w = 0, max_h = 0, max_points = 0;
    SlotTag slot[SVO_GPU_NUM_SLOTS];
    uint64_t tick = 0;
    // page-locked result staging of detect(): sized once per context, never zero-filled
    float* pin_xy = nullptr;
    float* pin_resp = nullptr;
    FrameCacheStats stats;
    void release()
    {
        if (pin_xy) svo_gpu_host_free(pin_xy);
        if (pin_resp) svo_gpu_host_free(pin_resp);
        if (ctx) svo_gpu_destroy(ctx);
        *this = CtxEntry();
    }
};

struct CtxTable {
    std::map<int, CtxEntry> by_device;
    int device = 0;
    bool cache = !(getenv("SVO_FRAME_CACHE") && atoi(getenv("SVO_FRAME_CACHE")) == 0);   // setFrameCache() overrides
    ~CtxTable()
    {
        for (auto& kv : by_device) kv.second.release();
    }
};

thread_local CtxTable g_tab;

constexpr int kMaxLevels = SVO_GPU_MAX_LEVELS;

[[noreturn]] void raise(svo_gpu_ctx* ctx, const char* where, int rc)
{
    SVO_THROW(std::string(where) + ": svo_gpu error " + std::to_string(rc) + ": " + svo_gpu_last_error(ctx));
}

// gray frames as the reference passes them, or BGR frames straight from cv::imread: the loader's cvtColor
// (include/async_image_loader.h:68-69) then happens on the device, bit-identical to cv::COLOR_BGR2GRAY
void check_image(const cv::Mat& m, const char* what)
{
    if (m.type() != CV_8UC1 && m.type() != CV_8UC3) SVO_THROW(std::string(what) + " must be CV_8UC1 or CV_8UC3 (BGR)");
}

int upload_image(svo_gpu_ctx* ctx, int slot, const cv::Mat& m)
{
    if (m.type() == CV_8UC3) return svo_gpu_upload_bgr(ctx, slot, m.data, m.cols, m.rows, m.step);
    return svo_gpu_upload(ctx, slot, m.data, m.cols, m.rows, m.step);
}

// 64-bit checksum of the pixels (rows may be strided): four independent multiply-xor lanes over 32-byte blocks, folded with
// a final avalanche.  ~20 us for a 1241x376 frame, against ~100 us for the upload + pyramid build it saves on a hit.
uint64_t image_checksum(const cv::Mat& m)
{
    const size_t row = (size_t)m.cols * m.elemSize();
    uint64_t h0 = 0x9e3779b97f4a7c15ull, h1 = 0xc2b2ae3d27d4eb4full, h2 = 0x165667b19e3779f9ull, h3 = 0x27d4eb2f165667c5ull;
    const uint64_t k0 = 0xff51afd7ed558ccdull, k1 = 0xc4ceb9fe1a85ec53ull, k2 = 0x9fb21c651e98df25ull, k3 = 0xd6e8feb86659fd93ull;
    for (int r = 0; r < m.rows; ++r) {
        const uchar* p = m.data + (size_t)r * m.step;
        size_t i = 0;
        for (; i + 32 <= row; i += 32) {
            uint64_t w[4];
            memcpy(w, p + i, 32);
            h0 = (h0 ^ w[0]) * k0; h1 = (h1 ^ w[1]) * k1; h2 = (h2 ^ w[2]) * k2; h3 = (h3 ^ w[3]) * k3;
            h0 ^= h0 >> 29; h1 ^= h1 >> 31; h2 ^= h2 >> 30; h3 ^= h3 >> 28;
        }
        if (i < row) {
            uint64_t w[4] = {0, 0, 0, 0};
            memcpy(w, p + i, row - i);
            h0 = (h0 ^ w[0]) * k0; h1 = (h1 ^ w[1]) * k1; h2 = (h2 ^ w[2]) * k2; h3 = (h3 ^ w[3]) * k3;
            h0 ^= h0 >> 29; h1 ^= h1 >> 31; h2 ^= h2 >> 30; h3 ^= h3 >> 28;
        }
    }
    uint64_t h = h0 ^ (h1 * k1) ^ (h2 * k2) ^ (h3 * k3) ^ ((uint64_t)m.rows << 40) ^ ((uint64_t)m.cols << 16) ^ (uint64_t)m.type();
    h ^= h >> 33; h *= k0; h ^= h >> 33; h *= k1; h ^= h >> 33;
    return h;
}

size_t edge_bytes(const cv::Mat& m) { return std::min<size_t>(64, (size_t)m.cols * m.elemSize()); }
const unsigned char* last_row_tail(const cv::Mat& m)
{
    return m.data + (size_t)(m.rows - 1) * m.step + (size_t)m.cols * m.elemSize() - edge_bytes(m);
}

// The slot of the thread's context that holds `m` (uploading it, into the least recently used slot other than `keep`, if no
// slot does).  Every image argument of the wrappers goes through here.
int slot_of(CtxEntry& e, const cv::Mat& m, int keep = -1)
{
    const uint64_t sum = g_tab.cache ? image_checksum(m) : 0;
    if (g_tab.cache)
        for (int s = 0; s < SVO_GPU_NUM_SLOTS; ++s) {
            SlotTag& t = e.slot[s];
            if (t.valid && s != keep && t.rows == m.rows && t.cols == m.cols && t.type == m.type() && t.sum == sum &&
                memcmp(t.head, m.data, edge_bytes(m)) == 0 && memcmp(t.tail, last_row_tail(m), edge_bytes(m)) == 0) {
                // the context is public (svo::defaultContext()): a direct svo_gpu_upload / svo_gpu_step on it may have
                // replaced the slot's image behind the tag's back
                unsigned long long gen = 0;
                if (svo_gpu_slot_generation(e.ctx, s, &gen) != SVO_GPU_OK || gen != t.gen) {
                    t.valid = false;
                    continue;
                }
                t.used = ++e.tick;
                ++e.stats.hits;
                return s;
            }
        }
    int victim = -1;
    for (int s = 0; s < SVO_GPU_NUM_SLOTS; ++s) {
        if (s == keep) continue;
        if (victim < 0 || !e.slot[s].valid || (e.slot[victim].valid && e.slot[s].used < e.slot[victim].used)) victim = s;
        if (!e.slot[s].valid) break;
    }
    e.slot[victim].valid = false;
    const int rc = upload_image(e.ctx, victim, m);
    if (rc != SVO_GPU_OK) raise(e.ctx, "svo_gpu_upload", rc);
    SlotTag& t = e.slot[victim];
    t.valid = g_tab.cache;
    t.rows = m.rows; t.cols = m.cols; t.type = m.type(); t.sum = sum; t.used = ++e.tick;
    svo_gpu_slot_generation(e.ctx, victim, &t.gen);
    memcpy(t.head, m.data, edge_bytes(m));
    memcpy(t.tail, last_row_tail(m), edge_bytes(m));
    ++e.stats.uploads;
    return victim;
}

// Makes sure `slot` holds a pyramid that serves an LK call with this window and maxLevel (see svo_gpu_slot_pyramid): built
// only if the slot has none, one with another margin, or too few levels.
void need_pyramid(CtxEntry& e, int slot, const cv::Mat& m, cv::Size win, int max_level)
{
    int have = 0, margin = 0;
    svo_gpu_slot_pyramid(e.ctx, slot, &have, &margin);
    const int want = svo_gpu_pyramid_levels(m.cols, m.rows, win.width, win.height, max_level) + 1;
    if (have >= want && margin == svo_gpu_pyramid_margin(win.width, win.height)) return;
    const int rc = svo_gpu_build_pyramid(e.ctx, slot, win.width, win.height, max_level, nullptr);
    if (rc != SVO_GPU_OK) raise(e.ctx, "svo_gpu_build_pyramid", rc);
    ++e.stats.pyramids;
}

CtxEntry& entry_for(int min_w, int min_h, int min_points)
{
    defaultContext(g_tab.device, min_w, min_h, min_points);
    return g_tab.by_device[g_tab.device];
}

}  // namespace

void setDevice(int device) { g_tab.device = device; }
int currentDevice() { return g_tab.device; }
void setFrameCache(bool enabled)
{
    g_tab.cache = enabled;
    for (auto& kv : g_tab.by_device)
        for (auto& t : kv.second.slot) t.valid = false;
}
FrameCacheStats frameCacheStats()
{
    auto it = g_tab.by_device.find(g_tab.device);
    return it == g_tab.by_device.end() ? FrameCacheStats() : it->second.stats;
}

void releaseContexts()
{
    for (auto& kv : g_tab.by_device) kv.second.release();
    g_tab.by_device.clear();
}

svo_gpu_ctx* defaultContext(int device, int min_w, int min_h, int min_points)
{
    CtxEntry& e = g_tab.by_device[device];
    if (e.ctx && min_w <= e.max_w && min_h <= e.max_h && min_points <= e.max_points) return e.ctx;
    const int w = std::max({e.max_w, min_w, 64}), h = std::max({e.max_h, min_h, 64});
    // room for every pixel of a quarter-density detection; grown on SVO_GPU_E_CAPACITY
    const int pts = std::max({e.max_points, min_points, 1 << 16, w * h / 8});
    const FrameCacheStats stats = e.stats;
    e.release();   // the slots (and with them the frame cache) go with the old context
    svo_gpu_ctx* c = nullptr;
    int rc = svo_gpu_create(&c, device, w, h, kMaxLevels, pts, 1);
    if (rc != SVO_GPU_OK) raise(nullptr, "svo_gpu_create", rc);
    e.ctx = c; e.max_w = w; e.max_h = h; e.max_points = pts; e.stats = stats;
    void* a = nullptr;
    void* b = nullptr;
    if ((rc = svo_gpu_host_alloc(&a, sizeof(float) * 2 * (size_t)pts)) || (rc = svo_gpu_host_alloc(&b, sizeof(float) * (size_t)pts))) {
        if (a) svo_gpu_host_free(a);
        e.release();
        raise(nullptr, "svo_gpu_host_alloc", rc);
    }
    e.pin_xy = static_cast<float*>(a);
    e.pin_resp = static_cast<float*>(b);
    return c;
}

// ------------------------------------------------------------------------------------------------ FAST
namespace {

template <typename Call>
void run_fast(const cv::Mat& image, std::vector<cv::KeyPoint>& keypoints, Call&& call)
{
    keypoints.clear();
    if (image.empty()) return;   // cv::Feature2D::detect on an empty image clears the keypoints
    check_image(image, "image");
    int min_points = 0;
    for (int attempt = 0; attempt < 2; ++attempt) {
        CtxEntry& e = entry_for(image.cols, image.rows, min_points);
        const int slot = slot_of(e, image);
        const float* xy = e.pin_xy;
        const float* resp = e.pin_resp;
        int count = 0;
        const int rc = call(e.ctx, slot, e.pin_xy, e.pin_resp, e.max_points, &count);
        if (rc == SVO_GPU_E_CAPACITY && attempt == 0) {   // more corners than the context was sized for: grow once
            min_points = count + count / 4;
            continue;
        }
        if (rc != SVO_GPU_OK) raise(e.ctx, "svo_gpu_fast", rc);
        keypoints.resize((size_t)count);
        for (int i = 0; i < count; ++i)
            keypoints[i] = cv::KeyPoint(cv::Point2f(xy[2 * i], xy[2 * i + 1]), 7.f, -1.f, resp[i], 0, -1);
        return;
    }
}

}  // namespace

void cvtColorBGR2GRAY(const cv::Mat& bgr, cv::Mat& gray)
{
    if (bgr.empty()) { gray = cv::Mat(); return; }
    if (bgr.type() != CV_8UC3) SVO_THROW("cvtColorBGR2GRAY: source must be CV_8UC3");
    CtxEntry& e = entry_for(bgr.cols, bgr.rows, 0);
    const int slot = slot_of(e, bgr);
    cv::Mat out(bgr.rows, bgr.cols, CV_8UC1);   // a fresh buffer: `gray` may be `bgr` itself, as in the loader
    const int rc = svo_gpu_download_image(e.ctx, slot, 0, out.data, out.step);
    if (rc != SVO_GPU_OK) raise(e.ctx, "svo_gpu_download_image", rc);
    gray = out;
}

void detectFAST(const cv::Mat& image, std::vector<cv::KeyPoint>& keypoints, const cv::Mat& mask, int threshold, bool nms)
{
    if (!mask.empty()) {
        if (mask.type() != CV_8UC1) SVO_THROW("mask must be CV_8UC1");
        if (mask.rows != image.rows || mask.cols != image.cols) SVO_THROW("mask size differs from the image size");
    }
    run_fast(image, keypoints, [&](svo_gpu_ctx* ctx, int slot, float* xy, float* resp, int cap, int* count) {
        return svo_gpu_fast(ctx, slot, threshold, nms ? 1 : 0, mask.empty() ? nullptr : mask.data, mask.step, xy, resp, cap, count);
    });
}

void detectFASTAvoiding(const cv::Mat& image, const std::vector<cv::Point2f>& tracked, float half,
                        std::vector<cv::KeyPoint>& keypoints, int threshold, bool nms)
{
    static_assert(sizeof(cv::Point2f) == 2 * sizeof(float), "Point2f must be two packed floats");
    run_fast(image, keypoints, [&](svo_gpu_ctx* ctx, int slot, float* xy, float* resp, int cap, int* count) {
        return svo_gpu_fast_masked_by_points(ctx, slot, threshold, nms ? 1 : 0,
                                             tracked.empty() ? nullptr : reinterpret_cast<const float*>(tracked.data()),
                                             (int)tracked.size(), half, xy, resp, cap, count);
    });
}

cv::Ptr<GpuFastDetector> GpuFastDetector::create(int threshold, bool nonmaxSuppression)
{
    return cv::makePtr<GpuFastDetector>(threshold, nonmaxSuppression);
}

void GpuFastDetector::detect(cv::InputArray image, std::vector<cv::KeyPoint>& keypoints, cv::InputArray mask)
{
    detectFAST(image.getMat(), keypoints, mask.empty() ? cv::Mat() : mask.getMat(), threshold_, nms_);
}

// ------------------------------------------------------------------------------------------------ LK
void calcOpticalFlowPyrLK(const cv::Mat& prevImg, const cv::Mat& nextImg, const std::vector<cv::Point2f>& prevPts,
                          std::vector<cv::Point2f>& nextPts, std::vector<uchar>& status, cv::Mat& err, cv::Size winSize,
                          int maxLevel, cv::TermCriteria criteria, int flags, double minEigThreshold)
{
    const int n = (int)prevPts.size();
    if (n == 0) {   // OpenCV: empty prevPts -> all outputs released
        nextPts.clear();
        status.clear();
        err = cv::Mat();
        return;
    }
    check_image(prevImg, "prevImg");
    check_image(nextImg, "nextImg");
    if (prevImg.empty() || prevImg.rows != nextImg.rows || prevImg.cols != nextImg.cols)
        SVO_THROW("calcOpticalFlowPyrLK: prevImg and nextImg must be non-empty and of the same size");
    if (maxLevel < 0 || winSize.width <= 2 || winSize.height <= 2)
        SVO_THROW("calcOpticalFlowPyrLK: maxLevel >= 0 && winSize.width > 2 && winSize.height > 2");
    CtxEntry& e = entry_for(prevImg.cols, prevImg.rows, n);
    svo_gpu_ctx* ctx = e.ctx;
    const int ml = std::min(maxLevel, kMaxLevels - 1);
    // the previous image is normally the `next` of the call before: found by content, its pyramid still in place
    const int sp = slot_of(e, prevImg), sn = slot_of(e, nextImg, sp);
    need_pyramid(e, sp, prevImg, winSize, ml);
    need_pyramid(e, sn, nextImg, winSize, ml);
    nextPts.resize((size_t)n);
    status.resize((size_t)n);
    err.create(n, 1, CV_32FC1);
    const int rc = svo_gpu_klt(ctx, sp, sn, reinterpret_cast<const float*>(prevPts.data()), reinterpret_cast<float*>(nextPts.data()),
                     status.data(), err.ptr<float>(), n, winSize.width, winSize.height, ml, criteria.maxCount,
                     criteria.epsilon, criteria.type, flags, minEigThreshold);
    if (rc != SVO_GPU_OK) raise(ctx, "svo_gpu_klt", rc);
}

// ------------------------------------------------------------------------------------------------ fused chains
namespace {

void upload_pair(CtxEntry& e, const cv::Mat& a, const cv::Mat& b, int win, int max_level, int& sa, int& sb)
{
    sa = slot_of(e, a);
    sb = slot_of(e, b, sa);
    need_pyramid(e, sa, a, cv::Size(win, win), max_level);
    need_pyramid(e, sb, b, cv::Size(win, win), max_level);
}

void check_pair(const cv::Mat& a, const cv::Mat& b, const char* who)
{
    check_image(a, who);
    check_image(b, who);
    if (a.empty() || a.rows != b.rows || a.cols != b.cols)
        SVO_THROW(std::string(who) + ": the two images must be non-empty and of the same size");
}

}  // namespace

void trackFrames(const cv::Mat& prevImg, const cv::Mat& currImg, const std::vector<cv::Point2f>& prevPts,
                 std::vector<cv::Point2f>& currPts, std::vector<int>& indices)
{
    const int n = (int)prevPts.size();
    currPts.clear();
    indices.clear();
    if (n == 0) return;
    check_pair(prevImg, currImg, "trackFrames");
    CtxEntry& e = entry_for(prevImg.cols, prevImg.rows, n);
    svo_gpu_ctx* ctx = e.ctx;
    int sp, sn;
    upload_pair(e, prevImg, currImg, 21, 3, sp, sn);
    currPts.resize((size_t)n);
    indices.resize((size_t)n);
    int m = 0;
    const int rc = svo_gpu_track_frames(ctx, sp, sn, reinterpret_cast<const float*>(prevPts.data()), n, 21, 21, 3, 50, 0.001,
                                        SVO_GPU_TC_COUNT | SVO_GPU_TC_EPS, SVO_GPU_LK_GET_MIN_EIGENVALS, 1e-4, indices.data(),
                                        reinterpret_cast<float*>(currPts.data()), nullptr, n, &m);
    if (rc != SVO_GPU_OK) raise(ctx, "svo_gpu_track_frames", rc);
    currPts.resize((size_t)m);
    indices.resize((size_t)m);
}

void extractStereoFeatures(const cv::Mat& leftImg, const cv::Mat& rightImg, const std::vector<cv::Point2f>& tracked,
                           int threshold, bool nms, float yThreshold, std::vector<cv::Point2f>& leftPts,
                           std::vector<cv::Point2f>& rightPts, int bucketSize, int featuresPerBucket)
{
    check_pair(leftImg, rightImg, "extractStereoFeatures");
    // every pixel can be a keypoint without NMS; with it at most one in four
    const int cap = std::max(4096, leftImg.rows * leftImg.cols / 4);
    CtxEntry& e = entry_for(leftImg.cols, leftImg.rows, std::max(cap, (int)tracked.size()));
    svo_gpu_ctx* ctx = e.ctx;
    int sl, sr;
    upload_pair(e, leftImg, rightImg, 11, 3, sl, sr);
    svo_gpu_keyframe_params P{};
    P.fast_threshold = threshold; P.fast_nms = nms ? 1 : 0; P.mask_half = 10.f;
    P.bucket_size = bucketSize; P.per_bucket = featuresPerBucket;
    P.win_w = 11; P.win_h = 11; P.max_level = 3; P.max_count = 30; P.epsilon = 0.001;
    P.criteria_type = SVO_GPU_TC_COUNT | SVO_GPU_TC_EPS; P.flags = 0; P.min_eig_threshold = 1e-4;
    P.y_threshold = yThreshold;
    leftPts.resize((size_t)cap);
    rightPts.resize((size_t)cap);
    int nd = 0, m = 0;
    const int rc = svo_gpu_keyframe(ctx, sl, sr, tracked.empty() ? nullptr : reinterpret_cast<const float*>(tracked.data()),
                                    (int)tracked.size(), &P, reinterpret_cast<float*>(leftPts.data()),
                                    reinterpret_cast<float*>(rightPts.data()), cap, &nd, &m);
    if (rc != SVO_GPU_OK) {
        leftPts.clear();
        rightPts.clear();
        raise(ctx, "svo_gpu_keyframe", rc);
    }
    leftPts.resize((size_t)m);
    rightPts.resize((size_t)m);
}

void filterTracks(const std::vector<cv::Point2f>& a, const std::vector<cv::Point2f>& b, const std::vector<uchar>& status,
                  float yThreshold, std::vector<int>& indices)
{
    const int n = (int)a.size();
    if (b.size() != a.size() || status.size() != a.size()) SVO_THROW("filterTracks: a, b and status differ in length");
    indices.clear();
    if (n == 0) return;
    svo_gpu_ctx* ctx = defaultContext(g_tab.device, 0, 0, n);
    indices.resize((size_t)n);
    int m = 0;
    const int rc = svo_gpu_filter_tracks(ctx, reinterpret_cast<const float*>(a.data()), reinterpret_cast<const float*>(b.data()),
                                         status.data(), n, yThreshold, indices.data(), nullptr, nullptr, n, &m);
    if (rc != SVO_GPU_OK) raise(ctx, "svo_gpu_filter_tracks", rc);
    indices.resize((size_t)m);
}

}  // namespace svo
