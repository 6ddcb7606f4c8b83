// batch.cpp -- svo::SequenceBatch: the reference's per-frame loop body (Tracking::startStereo, src/tracking.cpp:240-257)
// for many independent sequences on one device, on top of svo_gpu_step_submit / svo_gpu_step_wait.  Host side only: frame
// staging in page-locked memory and result marshalling into the reference's containers (std::vector<cv::Point2f>, ...).
#include <algorithm>
#include <cstring>
#include <string>

#include "svo_frontend.hpp"

namespace svo {

namespace {

[[noreturn]] void raise(svo_gpu_ctx* ctx, const char* where, int rc)
{
    SVO_THROW(std::string(where) + ": svo_gpu error " + std::to_string(rc) + ": " + svo_gpu_last_error(ctx));
}

template <typename T>
T* pinned(size_t n)
{
    void* p = nullptr;
    const int rc = svo_gpu_host_alloc(&p, sizeof(T) * (n ? n : 1));
    if (rc != SVO_GPU_OK) raise(nullptr, "svo_gpu_host_alloc", rc);
    return static_cast<T*>(p);
}

}  // namespace

// page-locked frame staging and result arrays of one outstanding frame set
struct SequenceBatch::Set {
    uint8_t* frames = nullptr;
    int *n_tracked = nullptr, *n_keypoints = nullptr, *n_bucketed = nullptr, *keyframe = nullptr, *n_survived = nullptr, *survivor_index = nullptr;
    float *next_xy = nullptr, *err = nullptr, *bucket_xy = nullptr;
    uint8_t* status = nullptr;
    svo_gpu_step_outputs out{};
    int ticket = -1;
    ~Set()
    {
        for (void* p : {(void*)frames, (void*)n_tracked, (void*)n_keypoints, (void*)n_bucketed, (void*)next_xy, (void*)err,
                        (void*)bucket_xy, (void*)status, (void*)keyframe, (void*)n_survived, (void*)survivor_index})
            svo_gpu_host_free(p);
    }
};

SequenceBatch::SequenceBatch(int device, int width, int height, int streams, const FrontendParams& params)
    : w_(width), h_(height), streams_(streams), p_(params)
{
    if (width < 1 || height < 1 || streams < 1 || params.bucketSize < 1 || params.featuresPerBucket < 1 || params.maxTracks < 0)
        SVO_THROW("SequenceBatch: bad geometry or parameters");
    bucket_cap_ = (height / params.bucketSize + 1) * (width / params.bucketSize + 1) * params.featuresPerBucket;
    const int max_points = std::max(std::max(32768, width * height / 16), std::max(bucket_cap_, params.maxTracks));
    int rc = svo_gpu_create(&ctx_, device, width, height, std::min(params.maxLevel + 1, SVO_GPU_MAX_LEVELS), max_points, streams);
    if (rc != SVO_GPU_OK) raise(nullptr, "svo_gpu_create", rc);
    try {
        const size_t S = (size_t)streams, T = (size_t)std::max(params.maxTracks, 1), Bc = (size_t)std::max(bucket_cap_, 1);
        for (auto& sp : sets_) {
            Set* s = sp = new Set();
            s->frames = pinned<uint8_t>(S * width * height);
            s->n_tracked = pinned<int>(S); s->n_keypoints = pinned<int>(S); s->n_bucketed = pinned<int>(S);
            s->next_xy = pinned<float>(S * T * 2); s->status = pinned<uint8_t>(S * T); s->err = pinned<float>(S * T);
            s->bucket_xy = pinned<float>(S * Bc * 2);
            s->keyframe = pinned<int>(S); s->n_survived = pinned<int>(S); s->survivor_index = pinned<int>(S * T);
            s->out.keyframe = s->keyframe; s->out.n_survived = s->n_survived; s->out.survivor_index = s->survivor_index;
            s->out.n_tracked = s->n_tracked; s->out.n_keypoints = s->n_keypoints; s->out.n_bucketed = s->n_bucketed;
            s->out.next_xy = s->next_xy; s->out.status = s->status; s->out.err = s->err;
            s->out.track_capacity = (int)T;
            s->out.bucket_xy = s->bucket_xy; s->out.bucket_capacity = (int)Bc;
            s->out.kp_capacity = 0;   // the raw keypoint lists stay on the device; only their counts come back
        }
    } catch (...) {
        for (auto& sp : sets_) { delete sp; sp = nullptr; }
        svo_gpu_destroy(ctx_);
        throw;
    }
}

SequenceBatch::~SequenceBatch()
{
    svo_gpu_destroy(ctx_);   // drains the streams first: nothing still reads or writes the pinned buffers
    for (auto& sp : sets_) delete sp;
}

void SequenceBatch::reset()
{
    const int rc = svo_gpu_step_reset(ctx_);
    if (rc != SVO_GPU_OK) raise(ctx_, "svo_gpu_step_reset", rc);
    waited_ = submitted_;
}

void SequenceBatch::submit(const std::vector<cv::Mat>& frames)
{
    if ((int)frames.size() != streams_) SVO_THROW("SequenceBatch::submit: one frame per sequence expected");
    if (outstanding() >= SVO_GPU_STEP_DEPTH) SVO_THROW("SequenceBatch::submit: wait for the oldest frame set first");
    Set& s = *sets_[submitted_ % SVO_GPU_STEP_DEPTH];
    for (int b = 0; b < streams_; ++b) {
        const cv::Mat& m = frames[(size_t)b];
        if (m.type() != CV_8UC1 || m.cols != w_ || m.rows != h_) SVO_THROW("SequenceBatch::submit: frames must be CV_8UC1 of the configured size");
        uint8_t* dst = s.frames + (size_t)b * w_ * h_;
        for (int y = 0; y < h_; ++y) memcpy(dst + (size_t)y * w_, m.data + (size_t)y * m.step, (size_t)w_);
    }
    submit(s.frames);
}

void SequenceBatch::submit(const uint8_t* frames)
{
    if (!frames) SVO_THROW("SequenceBatch::submit: frames is null");
    if (outstanding() >= SVO_GPU_STEP_DEPTH) SVO_THROW("SequenceBatch::submit: wait for the oldest frame set first");
    Set& s = *sets_[submitted_ % SVO_GPU_STEP_DEPTH];
    svo_gpu_step_params P{};
    P.w = w_; P.h = h_;
    P.fast_threshold = p_.fastThreshold; P.fast_nms = p_.nonmaxSuppression ? 1 : 0; P.mask_half = p_.maskHalf;
    P.bucket_size = p_.bucketSize; P.per_bucket = p_.featuresPerBucket; P.max_tracks = p_.maxTracks;
    P.win_w = p_.winSize.width; P.win_h = p_.winSize.height; P.max_level = p_.maxLevel;
    P.max_count = p_.criteria.maxCount; P.epsilon = p_.criteria.epsilon; P.criteria_type = p_.criteria.type;
    P.flags = p_.flags; P.min_eig_threshold = p_.minEigThreshold;
    P.carry_tracks = p_.carryTracks ? 1 : 0; P.features_to_track = p_.featuresToTrack;
    const int rc = svo_gpu_step_submit(ctx_, frames, nullptr, SVO_GPU_MEM_HOST, (size_t)w_, (size_t)w_ * h_, streams_, &P,
                                       &s.out, &s.ticket);
    if (rc != SVO_GPU_OK) raise(ctx_, "svo_gpu_step_submit", rc);
    ++submitted_;
}

void SequenceBatch::wait(std::vector<SequenceResult>& results)
{
    if (outstanding() <= 0) SVO_THROW("SequenceBatch::wait: nothing outstanding");
    Set& s = *sets_[waited_ % SVO_GPU_STEP_DEPTH];
    ++waited_;
    const int rc = svo_gpu_step_wait(ctx_, s.ticket);
    if (rc != SVO_GPU_OK) raise(ctx_, "svo_gpu_step_wait", rc);
    results.resize((size_t)streams_);
    const size_t T = (size_t)s.out.track_capacity, Bc = (size_t)s.out.bucket_capacity;
    for (int b = 0; b < streams_; ++b) {
        SequenceResult& r = results[(size_t)b];
        const size_t nt = (size_t)s.n_tracked[b], nb = (size_t)s.n_bucketed[b];
        const cv::Point2f* nx = reinterpret_cast<const cv::Point2f*>(s.next_xy) + (size_t)b * T;
        r.tracked.assign(nx, nx + nt);
        r.status.assign(s.status + (size_t)b * T, s.status + (size_t)b * T + nt);
        r.err.assign(s.err + (size_t)b * T, s.err + (size_t)b * T + nt);
        const cv::Point2f* bx = reinterpret_cast<const cv::Point2f*>(s.bucket_xy) + (size_t)b * Bc;
        r.features.assign(bx, bx + nb);
        r.keypoints = s.n_keypoints[b];
        r.keyframe = !p_.carryTracks || s.keyframe[b] != 0;
        r.survivors.clear();
        if (p_.carryTracks) r.survivors.assign(s.survivor_index + (size_t)b * T, s.survivor_index + (size_t)b * T + (size_t)s.n_survived[b]);
    }
}

}  // namespace svo
