// feature.cpp -- the detection / tracking entry points of the front-end with the reference's call signatures
// (declared in svo_frontend.hpp), implemented on the C ABI of include/svo_gpu.h.  The reference's own src/feature.cpp
// holds only the Feature value struct (src/feature.cpp:6-16); its FAST and LK arithmetic are OpenCV calls made from
// src/tracking.cpp:82, :101-105 and :160-165 -- those are what this file stands in for.  No CPU fallback.
#include <algorithm>
#include <map>
#include <string>

#include "svo_frontend.hpp"

namespace svo {

namespace {

struct CtxEntry {
    svo_gpu_ctx* ctx = nullptr;
    int max_w = 0, max_h = 0, max_points = 0;
};

struct CtxTable {
    std::map<int, CtxEntry> by_device;
    int device = 0;
    ~CtxTable()
    {
        for (auto& kv : by_device) svo_gpu_destroy(kv.second.ctx);
    }
};

thread_local CtxTable g_tab;

constexpr int kMaxLevels = SVO_GPU_MAX_LEVELS;

[[noreturn]] void raise(svo_gpu_ctx* ctx, const char* where, int rc)
{
    throw cv::Exception(std::string(where) + ": svo_gpu error " + std::to_string(rc) + ": " + svo_gpu_last_error(ctx));
}

// gray frames as the reference passes them, or BGR frames straight from cv::imread: the loader's cvtColor
// (include/async_image_loader.h:68-69) then happens on the device, bit-identical to cv::COLOR_BGR2GRAY
void check_image(const cv::Mat& m, const char* what)
{
    if (m.type() != CV_8UC1 && m.type() != CV_8UC3) throw cv::Exception(std::string(what) + " must be CV_8UC1 or CV_8UC3 (BGR)");
}

int upload_image(svo_gpu_ctx* ctx, int slot, const cv::Mat& m)
{
    if (m.type() == CV_8UC3) return svo_gpu_upload_bgr(ctx, slot, m.data, m.cols, m.rows, m.step);
    return svo_gpu_upload(ctx, slot, m.data, m.cols, m.rows, m.step);
}

}  // namespace

void setDevice(int device) { g_tab.device = device; }
int currentDevice() { return g_tab.device; }

void releaseContexts()
{
    for (auto& kv : g_tab.by_device) svo_gpu_destroy(kv.second.ctx);
    g_tab.by_device.clear();
}

svo_gpu_ctx* defaultContext(int device, int min_w, int min_h, int min_points)
{
    CtxEntry& e = g_tab.by_device[device];
    if (e.ctx && min_w <= e.max_w && min_h <= e.max_h && min_points <= e.max_points) return e.ctx;
    const int w = std::max({e.max_w, min_w, 64}), h = std::max({e.max_h, min_h, 64});
    // room for every pixel of a quarter-density detection; grown on SVO_GPU_E_CAPACITY
    const int pts = std::max({e.max_points, min_points, 1 << 16, w * h / 8});
    if (e.ctx) svo_gpu_destroy(e.ctx);
    e = CtxEntry();
    svo_gpu_ctx* c = nullptr;
    const int rc = svo_gpu_create(&c, device, w, h, kMaxLevels, pts, 1);
    if (rc != SVO_GPU_OK) raise(nullptr, "svo_gpu_create", rc);
    e.ctx = c; e.max_w = w; e.max_h = h; e.max_points = pts;
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
    std::vector<float> xy, resp;
    int min_points = 0;
    for (int attempt = 0; attempt < 2; ++attempt) {
        svo_gpu_ctx* ctx = defaultContext(g_tab.device, image.cols, image.rows, min_points);
        int rc = upload_image(ctx, 0, image);
        if (rc != SVO_GPU_OK) raise(ctx, "svo_gpu_upload", rc);
        const int cap = g_tab.by_device[g_tab.device].max_points;
        xy.resize(2 * (size_t)cap);
        resp.resize((size_t)cap);
        int count = 0;
        rc = call(ctx, xy.data(), resp.data(), cap, &count);
        if (rc == SVO_GPU_E_CAPACITY && attempt == 0) {   // more corners than the context was sized for: grow once
            min_points = count + count / 4;
            continue;
        }
        if (rc != SVO_GPU_OK) raise(ctx, "svo_gpu_fast", rc);
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
    if (bgr.type() != CV_8UC3) throw cv::Exception("cvtColorBGR2GRAY: source must be CV_8UC3");
    svo_gpu_ctx* ctx = defaultContext(g_tab.device, bgr.cols, bgr.rows, 0);
    int rc = svo_gpu_upload_bgr(ctx, 0, bgr.data, bgr.cols, bgr.rows, bgr.step);
    if (rc != SVO_GPU_OK) raise(ctx, "svo_gpu_upload_bgr", rc);
    cv::Mat out(bgr.rows, bgr.cols, CV_8UC1);   // a fresh buffer: `gray` may be `bgr` itself, as in the loader
    rc = svo_gpu_download_image(ctx, 0, 0, out.data, out.step);
    if (rc != SVO_GPU_OK) raise(ctx, "svo_gpu_download_image", rc);
    gray = out;
}

void detectFAST(const cv::Mat& image, std::vector<cv::KeyPoint>& keypoints, const cv::Mat& mask, int threshold, bool nms)
{
    if (!mask.empty()) {
        if (mask.type() != CV_8UC1) throw cv::Exception("mask must be CV_8UC1");
        if (mask.rows != image.rows || mask.cols != image.cols) throw cv::Exception("mask size differs from the image size");
    }
    run_fast(image, keypoints, [&](svo_gpu_ctx* ctx, float* xy, float* resp, int cap, int* count) {
        return svo_gpu_fast(ctx, 0, threshold, nms ? 1 : 0, mask.empty() ? nullptr : mask.data, mask.step, xy, resp, cap, count);
    });
}

void detectFASTAvoiding(const cv::Mat& image, const std::vector<cv::Point2f>& tracked, float half,
                        std::vector<cv::KeyPoint>& keypoints, int threshold, bool nms)
{
    static_assert(sizeof(cv::Point2f) == 2 * sizeof(float), "Point2f must be two packed floats");
    run_fast(image, keypoints, [&](svo_gpu_ctx* ctx, float* xy, float* resp, int cap, int* count) {
        return svo_gpu_fast_masked_by_points(ctx, 0, threshold, nms ? 1 : 0,
                                             tracked.empty() ? nullptr : reinterpret_cast<const float*>(tracked.data()),
                                             (int)tracked.size(), half, xy, resp, cap, count);
    });
}

cv::Ptr<GpuFastDetector> GpuFastDetector::create(int threshold, bool nonmaxSuppression)
{
    return cv::makePtr<GpuFastDetector>(threshold, nonmaxSuppression);
}

#ifdef SVO_COMPAT_OPENCV
void GpuFastDetector::detect(const cv::Mat& image, std::vector<cv::KeyPoint>& keypoints, const cv::Mat& mask)
{
    detectFAST(image, keypoints, mask, threshold_, nms_);
}
#else
void GpuFastDetector::detect(cv::InputArray image, std::vector<cv::KeyPoint>& keypoints, cv::InputArray mask)
{
    detectFAST(image.getMat(), keypoints, mask.empty() ? cv::Mat() : mask.getMat(), threshold_, nms_);
}
#endif

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
        throw cv::Exception("calcOpticalFlowPyrLK: prevImg and nextImg must be non-empty and of the same size");
    if (maxLevel < 0 || winSize.width <= 2 || winSize.height <= 2)
        throw cv::Exception("calcOpticalFlowPyrLK: maxLevel >= 0 && winSize.width > 2 && winSize.height > 2");
    svo_gpu_ctx* ctx = defaultContext(g_tab.device, prevImg.cols, prevImg.rows, n);
    int rc;
    if ((rc = upload_image(ctx, 0, prevImg)) ||
        (rc = upload_image(ctx, 1, nextImg)))
        raise(ctx, "svo_gpu_upload", rc);
    const int ml = std::min(maxLevel, kMaxLevels - 1);
    if ((rc = svo_gpu_build_pyramid(ctx, 0, winSize.width, winSize.height, ml, nullptr)) ||
        (rc = svo_gpu_build_pyramid(ctx, 1, winSize.width, winSize.height, ml, nullptr)))
        raise(ctx, "svo_gpu_build_pyramid", rc);
    nextPts.resize((size_t)n);
    status.resize((size_t)n);
    err.create(n, 1, CV_32FC1);
    rc = svo_gpu_klt(ctx, 0, 1, reinterpret_cast<const float*>(prevPts.data()), reinterpret_cast<float*>(nextPts.data()),
                     status.data(), err.ptr<float>(), n, winSize.width, winSize.height, ml, criteria.maxCount,
                     criteria.epsilon, criteria.type, flags, minEigThreshold);
    if (rc != SVO_GPU_OK) raise(ctx, "svo_gpu_klt", rc);
}

// ------------------------------------------------------------------------------------------------ fused chains
namespace {

void upload_pair(svo_gpu_ctx* ctx, const cv::Mat& a, const cv::Mat& b, int win, int max_level)
{
    int rc;
    if ((rc = upload_image(ctx, 0, a)) || (rc = upload_image(ctx, 1, b)))
        raise(ctx, "svo_gpu_upload", rc);
    if ((rc = svo_gpu_build_pyramid(ctx, 0, win, win, max_level, nullptr)) || (rc = svo_gpu_build_pyramid(ctx, 1, win, win, max_level, nullptr)))
        raise(ctx, "svo_gpu_build_pyramid", rc);
}

void check_pair(const cv::Mat& a, const cv::Mat& b, const char* who)
{
    check_image(a, who);
    check_image(b, who);
    if (a.empty() || a.rows != b.rows || a.cols != b.cols)
        throw cv::Exception(std::string(who) + ": the two images must be non-empty and of the same size");
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
    svo_gpu_ctx* ctx = defaultContext(g_tab.device, prevImg.cols, prevImg.rows, n);
    upload_pair(ctx, prevImg, currImg, 21, 3);
    currPts.resize((size_t)n);
    indices.resize((size_t)n);
    int m = 0;
    const int rc = svo_gpu_track_frames(ctx, 0, 1, reinterpret_cast<const float*>(prevPts.data()), n, 21, 21, 3, 50, 0.001,
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
    svo_gpu_ctx* ctx = defaultContext(g_tab.device, leftImg.cols, leftImg.rows, std::max(cap, (int)tracked.size()));
    upload_pair(ctx, leftImg, rightImg, 11, 3);
    svo_gpu_keyframe_params P{};
    P.fast_threshold = threshold; P.fast_nms = nms ? 1 : 0; P.mask_half = 10.f;
    P.bucket_size = bucketSize; P.per_bucket = featuresPerBucket;
    P.win_w = 11; P.win_h = 11; P.max_level = 3; P.max_count = 30; P.epsilon = 0.001;
    P.criteria_type = SVO_GPU_TC_COUNT | SVO_GPU_TC_EPS; P.flags = 0; P.min_eig_threshold = 1e-4;
    P.y_threshold = yThreshold;
    leftPts.resize((size_t)cap);
    rightPts.resize((size_t)cap);
    int nd = 0, m = 0;
    const int rc = svo_gpu_keyframe(ctx, 0, 1, tracked.empty() ? nullptr : reinterpret_cast<const float*>(tracked.data()),
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
    if (b.size() != a.size() || status.size() != a.size()) throw cv::Exception("filterTracks: a, b and status differ in length");
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
