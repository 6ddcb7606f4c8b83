// svo_frontend.hpp -- host C++ face of the B200 feature front-end, with the signatures the reference's tracking.cpp
// already calls, so the three call sites change by an identifier each:
//
//   src/tracking.cpp:54    mDetector = cv::FastFeatureDetector::create(thr, nms)   ->  svo::GpuFastDetector::create(thr, nms)
//   src/tracking.cpp:82    mDetector->detect(frame->leftImg(), keypoints, mask)    ->  unchanged (virtual cv::Feature2D::detect)
//   src/tracking.cpp:101   cv::calcOpticalFlowPyrLK(... Size(11,11), 3, tc)        ->  svo::calcOpticalFlowPyrLK(same arguments)
//   src/tracking.cpp:160   cv::calcOpticalFlowPyrLK(... Size(21,21), 3, tc, flags) ->  svo::calcOpticalFlowPyrLK(same arguments)
//   src/tracking.cpp:88    "// todo feature bucketing ???"                         ->  FeatureSet::bucketingFeatures (bucket.h)
//
// Everything numeric happens behind the C ABI of include/svo_gpu.h (libsvo_gpu.so, sm_100a kernels).  There is no CPU
// fallback: a failing call throws cv::Exception carrying svo_gpu_last_error(), the way a failing OpenCV call would.
#pragma once

#if __has_include(<opencv2/core.hpp>) && !defined(SVO_FORCE_COMPAT_OPENCV)
#include <opencv2/core.hpp>
#include <opencv2/features2d.hpp>
#include <opencv2/video/tracking.hpp>
#else
#include "compat/opencv2/core.hpp"
#endif

#include <vector>

#include "svo_gpu.h"

// failures surface the way a failing OpenCV call does: cv::Exception(code, err, func, file, line)
#define SVO_THROW(msg) throw cv::Exception(cv::Error::StsError, (msg), __func__, __FILE__, __LINE__)

namespace svo {

// The calling thread's context on `device` (created on first use, grown when a larger image or more points arrive).
// Tracking::trackFrames is a static member (include/tracking.h:34) with no object to hang state on, hence thread-local.
svo_gpu_ctx* defaultContext(int device = 0, int min_w = 0, int min_h = 0, int min_points = 0);
// Selects the CUDA device the wrappers below use on this thread (default 0): one host thread per GPU.
void setDevice(int device);
int currentDevice();
// Releases this thread's contexts (otherwise released at thread exit).
void releaseContexts();

// Frame cache of the wrappers below.  The reference passes plain cv::Mat images, so cv::calcOpticalFlowPyrLK rebuilds both
// pyramids on every call: a frame's pyramid is built as `next` (src/tracking.cpp:160-165), again as `prev` of the following
// frame, and a third time for the stereo call on a key-frame (:101-105).  The wrappers keep the last SVO_GPU_NUM_SLOTS
// frames on the device, recognised by CONTENT (size, type, 64-bit checksum of the pixels; a deep copy hits, a recycled buffer
// with new pixels cannot), so each frame is uploaded and its pyramid built once.  Results are identical either way.
void setFrameCache(bool enabled);   // default: enabled; per thread
struct FrameCacheStats {
    long long hits = 0;       // image arguments found on the device
    long long uploads = 0;    // image arguments copied to the device
    long long pyramids = 0;   // pyramid builds
};
FrameCacheStats frameCacheStats();   // of the calling thread's context on its current device

// cv::FastFeatureDetector stand-in: FAST-9/16, threshold, optional 3x3 NMS, optional mask; keypoints in raster order with
// size 7, angle -1, response = score, octave 0, class_id -1, exactly as cv::FastFeatureDetector::detect returns them.
class GpuFastDetector : public cv::Feature2D {
public:
    static cv::Ptr<GpuFastDetector> create(int threshold = 10, bool nonmaxSuppression = true);
    GpuFastDetector(int threshold, bool nonmaxSuppression) : threshold_(threshold), nms_(nonmaxSuppression) {}
    // cv::Feature2D::detect's own signature (features2d.hpp), with real OpenCV headers and with the compat ones alike
    void detect(cv::InputArray image, std::vector<cv::KeyPoint>& keypoints, cv::InputArray mask = cv::noArray()) override;
    void setThreshold(int t) { threshold_ = t; }
    int getThreshold() const { return threshold_; }
    void setNonmaxSuppression(bool f) { nms_ = f; }
    bool getNonmaxSuppression() const { return nms_; }

private:
    int threshold_;
    bool nms_;
};

// Free-function form of the same call.
void detectFAST(const cv::Mat& image, std::vector<cv::KeyPoint>& keypoints, const cv::Mat& mask, int threshold, bool nms);

// The reference's mask + detect pair in one device pass (src/tracking.cpp:76-82): the 255/0 mask around `tracked` points is
// built on the GPU, so no w*h-byte mask crosses PCIe.  Same keypoints as building the mask with cv::rectangle.
void detectFASTAvoiding(const cv::Mat& image, const std::vector<cv::Point2f>& tracked, float half,
                        std::vector<cv::KeyPoint>& keypoints, int threshold, bool nms);

// cv::cvtColor(bgr, gray, cv::COLOR_BGR2GRAY) for CV_8UC3 frames, computed on the device, bit-identical to OpenCV's 8-bit
// fixed point: the conversion the reference's loader applies to the colour sequence (include/async_image_loader.h:68-69).
// Every image argument below also accepts the CV_8UC3 frame itself (converted on the way in), which saves this round trip.
void cvtColorBGR2GRAY(const cv::Mat& bgr, cv::Mat& gray);

// cv::calcOpticalFlowPyrLK with its full parameter list and defaults.  status: 0/1 per point; err: CV_32F n x 1.
// OPTFLOW_USE_INITIAL_FLOW is not supported (the reference never sets it) and throws.
void calcOpticalFlowPyrLK(const cv::Mat& prevImg, const cv::Mat& nextImg, const std::vector<cv::Point2f>& prevPts,
                          std::vector<cv::Point2f>& nextPts, std::vector<uchar>& status, cv::Mat& err,
                          cv::Size winSize = cv::Size(21, 21), int maxLevel = 3,
                          cv::TermCriteria criteria = cv::TermCriteria(cv::TermCriteria::COUNT + cv::TermCriteria::EPS, 30, 0.01),
                          int flags = 0, double minEigThreshold = 1e-4);


// ---- the reference's call chains with their filter loops, fused on the device ----------------

// Tracking::trackFrames (src/tracking.cpp:154-179): LK 21x21 / maxLevel 3 / {COUNT+EPS, 50, 0.001} /
// OPTFLOW_LK_GET_MIN_EIGENVALS from prevImg to currImg, then `if (status[i])`.  currPts = positions of the surviving
// tracks in input order, indices[k] = position in prevPts of survivor k (to carry Feature::mapPoint over, :171).
void trackFrames(const cv::Mat& prevImg, const cv::Mat& currImg, const std::vector<cv::Point2f>& prevPts,
                 std::vector<cv::Point2f>& currPts, std::vector<int>& indices);

// Tracking::extractFeatures + findLeftFeaturesInRight (src/tracking.cpp:74-122): mask around `tracked` (+-10 px), FAST
// (threshold, nms) on leftImg, optional bucketing (bucketSize > 0: the "todo" of :88), LK 11x11 / maxLevel 3 /
// {COUNT+EPS, 30, 0.001} into rightImg, then `status[i] && std::abs(right.y - left.y) < yThreshold`.
// leftPts / rightPts = the surviving stereo pairs in detection order.
void extractStereoFeatures(const cv::Mat& leftImg, const cv::Mat& rightImg, const std::vector<cv::Point2f>& tracked,
                           int threshold, bool nms, float yThreshold, std::vector<cv::Point2f>& leftPts,
                           std::vector<cv::Point2f>& rightPts, int bucketSize = 0, int featuresPerBucket = 0);

// ---- many independent sequences on one GPU (BASELINE config 5) --------------------------------------------
// The loop body of Tracking::startStereo (src/tracking.cpp:240-257) for `streams` sequences at once: per sequence, LK
// (trackFrames, :154-179) from its previous frame on its current tracks, then FAST around a mask of those tracks and
// bucketing (extractFeatures, :74-92); the first maxTracks bucketed detections become the sequence's tracks for the next
// frame.  One object = one device context; sequences never interact.  Frames are submitted and waited for separately,
// with up to two frame sets outstanding, so that the host (pose estimation, mapping) works on frame t while the GPU
// already runs frame t + 1 -- the overlap the reference gets from its loader thread (include/async_image_loader.h:27-33).
struct FrontendParams {
    int fastThreshold = 20;           // include/config_reader.h:36
    bool nonmaxSuppression = true;    // include/config_reader.h:37
    float maskHalf = 10.f;            // src/tracking.cpp:78; <= 0: no mask
    int bucketSize = 32, featuresPerBucket = 5;
    int maxTracks = 2000;
    cv::Size winSize = cv::Size(21, 21);   // src/tracking.cpp:163
    int maxLevel = 3;
    cv::TermCriteria criteria = cv::TermCriteria(cv::TermCriteria::COUNT + cv::TermCriteria::EPS, 50, 0.001);   // :157
    int flags = cv::OPTFLOW_LK_GET_MIN_EIGENVALS;                                                              // :164
    double minEigThreshold = 1e-4;
    // false: every frame is a key-frame and a sequence's tracks are re-seeded from its bucketed detections (fixed load).
    // true: the reference's own policy -- LK survivors carry over (src/tracking.cpp:168-177), detection runs only on
    // key-frames (a sequence's first frame, or after a non-key-frame that ended below featuresToTrack, :69-70), new features
    // are appended after the survivors (:254).  Decided per sequence on the device.
    bool carryTracks = false;
    int featuresToTrack = 70;         // configs/config.yaml:15
};

struct SequenceResult {
    std::vector<cv::Point2f> tracked;   // LK result for each track carried into this frame (empty on a sequence's first frame)
    std::vector<uchar> status;
    std::vector<float> err;
    std::vector<cv::Point2f> features;  // bucketed detections of this frame; the first maxTracks are the next frame's tracks
    int keypoints = 0;                  // FAST keypoints found before bucketing
    // carryTracks only: was this frame a key-frame (did detection run), and for each LK survivor its position in `tracked`
    // (= in the track list the frame started with), so that Feature::mapPoint bindings carry over (src/tracking.cpp:171);
    // the next frame's track list is the survivors' `tracked` entries followed by the head of `features`
    bool keyframe = true;
    std::vector<int> survivors;
};

class SequenceBatch {
public:
    SequenceBatch(int device, int width, int height, int streams, const FrontendParams& params = FrontendParams());
    ~SequenceBatch();
    SequenceBatch(const SequenceBatch&) = delete;
    SequenceBatch& operator=(const SequenceBatch&) = delete;

    // One CV_8UC1 frame of the configured size per sequence.  Copies them to page-locked staging and enqueues the step.
    void submit(const std::vector<cv::Mat>& frames);
    // The same without the staging copy: `streams` images back to back with compact rows (width bytes each), in memory the
    // caller keeps unchanged until the matching wait() -- page-locked (svo_gpu_host_alloc) if the copy is to overlap kernels.
    void submit(const uint8_t* frames);
    // Waits for the oldest outstanding frame set; results[s] belongs to sequence s.
    void wait(std::vector<SequenceResult>& results);
    int outstanding() const { return submitted_ - waited_; }
    void reset();   // forget every sequence's previous frame and tracks

private:
    struct Set;
    svo_gpu_ctx* ctx_ = nullptr;
    int w_, h_, streams_, bucket_cap_;
    FrontendParams p_;
    Set* sets_[SVO_GPU_STEP_DEPTH] = {};
    int submitted_ = 0, waited_ = 0;
};

// The filter loop alone (src/tracking.cpp:109-115; yThreshold < 0: :168-175) for callers that ran LK themselves.
void filterTracks(const std::vector<cv::Point2f>& a, const std::vector<cv::Point2f>& b, const std::vector<uchar>& status,
                  float yThreshold, std::vector<int>& indices);

}  // namespace svo
