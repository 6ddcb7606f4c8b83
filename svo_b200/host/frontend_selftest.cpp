// frontend_selftest.cpp -- drives the drop-in wrappers the way the reference's Tracking does and dumps the results, so
// that tests/test_host_wrappers.py can compare them with the oracle.  Mirrors, in order:
//   Tracking::extractFeatures        src/tracking.cpp:74-92   mask (cv::rectangle equivalent) -> detect -> KeyPoint::convert
//   "todo feature bucketing"         src/tracking.cpp:88      FeatureSet::appendNewFeatures + bucketingFeatures
//   Tracking::findLeftFeaturesInRight src/tracking.cpp:94-118 LK 11x11, 30 it, flags 0
//   Tracking::trackFrames            src/tracking.cpp:154-179 LK 21x21, 50 it, OPTFLOW_LK_GET_MIN_EIGENVALS
// usage: frontend_selftest <in.bin> <out.bin>     in.bin = int32 {w, h, n_prev, bucket_size, per_bucket}, image a, image b,
//                                                  n_prev x float2 previously tracked points
//        frontend_selftest --bench <frames.bin> <n_frames>   one sequence through the wrappers, timed (see bench_main)
// Without a CUDA device every call must throw (no CPU fallback): `frontend_selftest --expect-no-gpu` checks exactly that.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "bucket.h"

static void put(FILE* f, const void* p, size_t bytes) { fwrite(p, 1, bytes, f); }
static void put_i(FILE* f, int v) { put(f, &v, sizeof v); }

static void fill_box(cv::Mat& m, cv::Point2f a, cv::Point2f b)   // cv::rectangle(mask, a, b, 0, FILLED) for the compat Mat
{
    auto rnd = [](float v) { return (int)lrintf(v); };
    int xa = rnd(a.x), ya = rnd(a.y), xb = rnd(b.x), yb = rnd(b.y);
    xa = xa < 0 ? 0 : xa; ya = ya < 0 ? 0 : ya;
    xb = xb >= m.cols ? m.cols - 1 : xb; yb = yb >= m.rows ? m.rows - 1 : yb;
    if (xa > xb) return;
    for (int y = ya; y <= yb; ++y) memset(m.ptr<uchar>(y) + xa, 0, (size_t)(xb - xa + 1));   // row fills, as cv::rectangle does
}

// BASELINE config 1 through the drop-in wrappers: ONE 1241x376 sequence, the loop body of Tracking::startStereo on a
// key-frame (src/tracking.cpp:240-257) written with the reference's own calls and host containers:
//   svo::calcOpticalFlowPyrLK(prev, cur, tracks, ...)  (:160-165)   host cv::Mat frames, std::vector<cv::Point2f> tracks
//   mask of +-10 px boxes on the host (:76-79)  ->  mDetector->detect(cur, keypoints, mask)  (:82)  ->  KeyPoint::convert (:86)
//   FeatureSet::appendNewFeatures + bucketingFeatures  (:88)  ->  the first max_tracks features are the next frame's tracks
// Pass "host_mask": exactly that (the three-identifier patch of INTEGRATION.md; the w*h mask crosses PCIe every frame).
// Pass "device_mask": detect replaced by svo::detectFASTAvoiding(cur, tracks, 10, ...) (one more changed line).
// frames.bin = int32 {w, h, ring, max_tracks, bucket_size, per_bucket, fast_threshold} + ring images; frame t = image t % ring
// with its first two pixels stamped with t, so that every frame is new to the wrappers' frame cache, as camera frames are.
static int bench_main(const char* path, int n_frames)
{
    using clk = std::chrono::steady_clock;
    auto ms = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
    FILE* fi = fopen(path, "rb");
    if (!fi) return 2;
    int hdr[7];
    if (fread(hdr, sizeof(int), 7, fi) != 7) return 2;
    const int w = hdr[0], h = hdr[1], ring = hdr[2], max_tracks = hdr[3], bs = hdr[4], k = hdr[5], thr = hdr[6];
    std::vector<cv::Mat> src((size_t)ring);
    for (auto& m : src) {
        m.create(h, w, CV_8UC1);
        if (fread(m.data, 1, (size_t)w * h, fi) != (size_t)w * h) return 2;
    }
    fclose(fi);
    try {
        cv::Ptr<cv::FeatureDetector> mDetector = svo::GpuFastDetector::create(thr, true);
        const cv::TermCriteria tc(cv::TermCriteria::COUNT + cv::TermCriteria::EPS, 50, 0.001);
        printf("{\"workload\": \"one %dx%d sequence, %d tracks, mask + FAST %d + NMS, bucketing %d/%d, LK 21x21 4 levels 50 it\", \"frames\": %d",
               w, h, max_tracks, thr, bs, k, n_frames);
        for (int pass = 0; pass < 2; ++pass) {
            const bool device_mask = pass == 1;
            double t_lk = 0, t_mask = 0, t_det = 0, t_conv = 0, t_bucket = 0, t_all = 0;
            long long ok = 0, ntr = 0, nkp = 0;
            std::vector<cv::Point2f> tracks, nextPts, newPoints;
            std::vector<uchar> status;
            std::vector<cv::KeyPoint> keypoints;
            cv::Mat error, prev;
            FeatureSet fs;
            const svo::FrameCacheStats c0 = svo::frameCacheStats();
            long long launches0 = 0;
            const int warm = 10;
            for (int t = 0; t < warm + n_frames; ++t) {
                cv::Mat cur(h, w, CV_8UC1);                      // what the loader hands over: a new frame in its own buffer
                memcpy(cur.data, src[(size_t)(t % ring)].data, (size_t)w * h);
                const int stamp = pass * 100000 + t;
                cur.data[0] = (uchar)(stamp & 255); cur.data[1] = (uchar)((stamp >> 8) & 255); cur.data[2] = (uchar)(stamp >> 16);
                if (t == warm) {
                    t_lk = t_mask = t_det = t_conv = t_bucket = t_all = 0;
                    ok = ntr = nkp = 0;
                    launches0 = svo_gpu_launch_count(svo::defaultContext(svo::currentDevice()));
                }
                const auto a0 = clk::now();
                if (!prev.empty() && !tracks.empty()) {
                    svo::calcOpticalFlowPyrLK(prev, cur, tracks, nextPts, status, error, cv::Size(21, 21), 3, tc,
                                              cv::OPTFLOW_LK_GET_MIN_EIGENVALS);
                    for (uchar s : status) ok += s;
                    ntr += (long long)status.size();
                }
                const auto a1 = clk::now();
                auto a2 = a1;
                if (device_mask) {
                    svo::detectFASTAvoiding(cur, tracks, 10.f, keypoints, thr, true);
                } else {
                    cv::Mat mask(cur.size(), CV_8UC1, 255);
                    for (const auto& p : tracks) fill_box(mask, p - cv::Point2f(10, 10), p + cv::Point2f(10, 10));
                    a2 = clk::now();
                    mDetector->detect(cur, keypoints, mask);
                }
                const auto a3 = clk::now();
                cv::KeyPoint::convert(keypoints, newPoints);
                const auto a4 = clk::now();
                fs.clear();
                fs.appendNewFeatures(newPoints);
                fs.bucketingFeatures(cur, bs, k);
                tracks.assign(fs.points.begin(), fs.points.begin() + std::min<size_t>(fs.points.size(), (size_t)max_tracks));
                const auto a5 = clk::now();
                prev = cur;
                nkp += (long long)keypoints.size();
                t_lk += ms(a0, a1); t_mask += ms(a1, a2); t_det += ms(a2, a3); t_conv += ms(a3, a4); t_bucket += ms(a4, a5);
                t_all += ms(a0, a5);
            }
            const svo::FrameCacheStats c1 = svo::frameCacheStats();
            const long long launches = svo_gpu_launch_count(svo::defaultContext(svo::currentDevice())) - launches0;
            const double n = n_frames;
            printf(", \"%s\": {\"ms_per_frame\": %.4f, \"frames_per_s\": %.1f, \"stage_ms\": {\"lk\": %.4f, \"mask_host\": %.4f, "
                   "\"detect\": %.4f, \"convert\": %.4f, \"bucket\": %.4f}, \"tracks_per_frame\": %.1f, \"status_ok_frac\": %.5f, "
                   "\"keypoints_per_frame\": %.1f, \"launches_per_frame\": %.2f, \"cache\": {\"hits\": %lld, \"uploads\": %lld, "
                   "\"pyramids\": %lld}}",
                   device_mask ? "device_mask" : "host_mask", t_all / n, 1e3 * n / t_all, t_lk / n, t_mask / n, t_det / n, t_conv / n,
                   t_bucket / n, ntr / n, ntr ? (double)ok / (double)ntr : 0., nkp / n, launches / n, c1.hits - c0.hits,
                   c1.uploads - c0.uploads, c1.pyramids - c0.pyramids);
        }
        // the same frame pair again: no upload, no pyramid, exactly one launch (the LK kernel)
        {
            cv::Mat a = src[0], b = src[1 % ring];
            std::vector<cv::Point2f> pts(100), out;
            for (int i = 0; i < 100; ++i) pts[(size_t)i] = cv::Point2f(40.f + 7.f * i, 50.f + (i % 9) * 20.f);
            std::vector<uchar> st;
            cv::Mat er;
            svo::calcOpticalFlowPyrLK(a, b, pts, out, st, er, cv::Size(21, 21), 3, tc, cv::OPTFLOW_LK_GET_MIN_EIGENVALS);
            const svo::FrameCacheStats c0 = svo::frameCacheStats();
            const long long l0 = svo_gpu_launch_count(svo::defaultContext(svo::currentDevice()));
            svo::calcOpticalFlowPyrLK(a, b, pts, out, st, er, cv::Size(21, 21), 3, tc, cv::OPTFLOW_LK_GET_MIN_EIGENVALS);
            svo::calcOpticalFlowPyrLK(a, b, pts, out, st, er, cv::Size(11, 11), 3, tc);   // the stereo window on the same pyramids
            const svo::FrameCacheStats c1 = svo::frameCacheStats();
            printf(", \"repeat_call\": {\"launches\": %lld, \"uploads\": %lld, \"pyramids\": %lld, \"hits\": %lld}",
                   svo_gpu_launch_count(svo::defaultContext(svo::currentDevice())) - l0, c1.uploads - c0.uploads,
                   c1.pyramids - c0.pyramids, c1.hits - c0.hits);
        }
        printf("}\n");
    } catch (const cv::Exception& e) {
        fprintf(stderr, "frontend_selftest --bench: %s\n", e.what());
        return 1;
    }
    svo::releaseContexts();
    return 0;
}

int main(int argc, char** argv)
{
    if (argc == 4 && !strcmp(argv[1], "--bench")) return bench_main(argv[2], atoi(argv[3]));
    if (argc == 2 && !strcmp(argv[1], "--expect-no-gpu")) {
        cv::Mat img(cv::Size(64, 64), CV_8UC1, 7);
        std::vector<cv::KeyPoint> kps;
        try {
            svo::GpuFastDetector::create(20, true)->detect(img, kps);
        } catch (const cv::Exception& e) {
            printf("threw as expected: %s\n", e.what());
            return 0;
        }
        fprintf(stderr, "detect() returned without a GPU: a CPU fallback must not exist\n");
        return 1;
    }
    if (argc == 2 && !strcmp(argv[1], "--stale-slot")) {
        // The wrappers remember which frame each slot of the thread's context holds; the context itself is public
        // (svo::defaultContext()).  Someone who uploads into its slots directly must not make the wrappers compute on the
        // wrong pixels: a cached frame whose slot has been overwritten is uploaded again.
        const int w = 160, h = 120;
        cv::Mat a(cv::Size(w, h), CV_8UC1, 0), b(cv::Size(w, h), CV_8UC1, 0);
        unsigned x = 12345u, y = 777u;
        for (int r = 0; r < h; ++r)
            for (int c = 0; c < w; ++c) {
                x = x * 1664525u + 1013904223u;
                y = y * 22695477u + 1u;
                a.data[r * a.step + c] = (unsigned char)(y >> 24);   // two different noise images: FAST finds many corners in each
                b.data[r * b.step + c] = (unsigned char)(x >> 24);
            }
        auto det = svo::GpuFastDetector::create(20, true);
        std::vector<cv::KeyPoint> k0, k1, kb;
        det->detect(a, k0);
        det->detect(b, kb);
        svo_gpu_ctx* ctx = svo::defaultContext(svo::currentDevice());
        for (int s = 0; s < SVO_GPU_NUM_SLOTS; ++s)       // overwrite every slot behind the wrappers' back
            if (svo_gpu_upload(ctx, s, b.data, w, h, b.step) != SVO_GPU_OK) { fprintf(stderr, "upload failed\n"); return 1; }
        const svo::FrameCacheStats c0 = svo::frameCacheStats();
        det->detect(a, k1);
        const svo::FrameCacheStats c1 = svo::frameCacheStats();
        auto equal = [](const std::vector<cv::KeyPoint>& p, const std::vector<cv::KeyPoint>& q) {
            if (p.size() != q.size()) return false;
            for (size_t i = 0; i < p.size(); ++i)
                if (p[i].pt.x != q[i].pt.x || p[i].pt.y != q[i].pt.y || p[i].response != q[i].response) return false;
            return true;
        };
        const bool same = !k0.empty() && equal(k0, k1) && !equal(k0, kb);
        printf("stale-slot: %zu keypoints before, %zu after, %zu on the other frame, uploads +%lld hits +%lld -> %s\n", k0.size(),
               k1.size(), kb.size(), c1.uploads - c0.uploads, c1.hits - c0.hits, same && c1.uploads - c0.uploads == 1 ? "ok" : "FAILED");
        return same && c1.uploads - c0.uploads == 1 ? 0 : 1;
    }
    if ((argc == 4 || argc == 5) && !strcmp(argv[1], "--batch")) {
        // optional 4th argument: features_to_track -> the reference's own track policy (FrontendParams::carryTracks); each
        // stream's record then ends with its key-frame flag and its survivor indices
        // svo::SequenceBatch the way a many-sequence driver uses it: frame set t + 1 is submitted before set t is waited for.
        // in.bin = int32 {w, h, streams, frames, max_tracks}, then frames x streams images; out.bin = per frame, per stream:
        // n_tracked, tracked xy, status, err, n_features, features xy, keypoints
        FILE* fi = fopen(argv[2], "rb");
        if (!fi) return 2;
        int hdr[5];
        if (fread(hdr, sizeof(int), 5, fi) != 5) return 2;
        const int w = hdr[0], h = hdr[1], S = hdr[2], T = hdr[3];
        std::vector<std::vector<cv::Mat>> seq((size_t)T, std::vector<cv::Mat>((size_t)S));
        for (auto& fs : seq)
            for (auto& m : fs) {
                m.create(h, w, CV_8UC1);
                if (fread(m.data, 1, (size_t)w * h, fi) != (size_t)w * h) return 2;
            }
        fclose(fi);
        FILE* fo = fopen(argv[3], "wb");
        try {
            svo::FrontendParams fp;
            fp.maxTracks = hdr[4];
            if (argc == 5) { fp.carryTracks = true; fp.featuresToTrack = atoi(argv[4]); }
            svo::SequenceBatch batch(0, w, h, S, fp);
            std::vector<svo::SequenceResult> res;
            auto dump = [&]() {
                batch.wait(res);
                for (const auto& r : res) {
                    put_i(fo, (int)r.tracked.size());
                    put(fo, r.tracked.data(), 8 * r.tracked.size());
                    put(fo, r.status.data(), r.status.size());
                    put(fo, r.err.data(), 4 * r.err.size());
                    put_i(fo, (int)r.features.size());
                    put(fo, r.features.data(), 8 * r.features.size());
                    put_i(fo, r.keypoints);
                    if (argc == 5) {
                        put_i(fo, r.keyframe ? 1 : 0);
                        put_i(fo, (int)r.survivors.size());
                        put(fo, r.survivors.data(), 4 * r.survivors.size());
                    }
                }
            };
            batch.submit(seq[0]);
            for (int t = 1; t < T; ++t) {
                batch.submit(seq[(size_t)t]);
                dump();   // frame set t - 1, while set t runs
            }
            dump();
            bool threw = false;
            try { batch.wait(res); } catch (const cv::Exception&) { threw = true; }   // nothing outstanding
            put_i(fo, threw ? 1 : 0);
        } catch (const cv::Exception& e) {
            fprintf(stderr, "frontend_selftest --batch: %s\n", e.what());
            fclose(fo);
            return 1;
        }
        fclose(fo);
        return 0;
    }
    if (argc != 3) return 2;
    FILE* fi = fopen(argv[1], "rb");
    if (!fi) return 2;
    int hdr[5];
    if (fread(hdr, sizeof(int), 5, fi) != 5) return 2;
    const int w = hdr[0], h = hdr[1], n_prev = hdr[2], bs = hdr[3], k = hdr[4];
    cv::Mat a(h, w, CV_8UC1), b(h, w, CV_8UC1);
    std::vector<cv::Point2f> prevPts((size_t)n_prev);
    if (fread(a.data, 1, (size_t)w * h, fi) != (size_t)w * h || fread(b.data, 1, (size_t)w * h, fi) != (size_t)w * h) return 2;
    if (n_prev && fread(prevPts.data(), sizeof(cv::Point2f), (size_t)n_prev, fi) != (size_t)n_prev) return 2;
    fclose(fi);
    FILE* fo = fopen(argv[2], "wb");
    try {
        // ---- extractFeatures: mask around the previous features, detect through the cv::Feature2D pointer
        cv::Ptr<cv::FeatureDetector> mDetector = svo::GpuFastDetector::create(20, true);
        cv::Mat mask(a.size(), CV_8UC1, 255);
        for (const auto& p : prevPts) fill_box(mask, p - cv::Point2f(10, 10), p + cv::Point2f(10, 10));
        std::vector<cv::KeyPoint> keypoints;
        mDetector->detect(a, keypoints, mask);
        std::vector<cv::Point2f> newPoints;
        cv::KeyPoint::convert(keypoints, newPoints);
        put_i(fo, (int)keypoints.size());
        for (const auto& kp : keypoints) { put(fo, &kp.pt, 8); put(fo, &kp.response, 4); }
        // the same keypoints with the mask built on the device
        std::vector<cv::KeyPoint> kp2;
        svo::detectFASTAvoiding(a, prevPts, 10.f, kp2, 20, true);
        int same = kp2.size() == keypoints.size();
        for (size_t i = 0; same && i < kp2.size(); ++i)
            same = kp2[i].pt.x == keypoints[i].pt.x && kp2[i].pt.y == keypoints[i].pt.y && kp2[i].response == keypoints[i].response;
        put_i(fo, same);
        // ---- bucketing at the "todo" of :88
        FeatureSet fs;
        fs.appendNewFeatures(newPoints);
        fs.bucketingFeatures(a, bs, k);
        put_i(fo, fs.size());
        put(fo, fs.points.data(), 8 * (size_t)fs.size());
        put(fo, fs.ages.data(), 4 * (size_t)fs.size());
        // ---- stereo-style LK (:101-105) and temporal LK (:160-165) on the bucketed points
        for (int mode = 0; mode < 2; ++mode) {
            std::vector<cv::Point2f> nextPts;
            std::vector<uchar> status;
            cv::Mat error;
            if (mode == 0) {
                cv::TermCriteria tc(cv::TermCriteria::COUNT + cv::TermCriteria::EPS, 30, 0.001);
                svo::calcOpticalFlowPyrLK(a, b, fs.points, nextPts, status, error, cv::Size(11, 11), 3, tc);
            } else {
                cv::TermCriteria tc(cv::TermCriteria::COUNT + cv::TermCriteria::EPS, 50, 0.001);
                svo::calcOpticalFlowPyrLK(a, b, fs.points, nextPts, status, error, cv::Size(21, 21), 3, tc,
                                          cv::OPTFLOW_LK_GET_MIN_EIGENVALS);
            }
            put_i(fo, (int)nextPts.size());
            put(fo, nextPts.data(), 8 * nextPts.size());
            put(fo, status.data(), status.size());
            put(fo, error.data, 4 * (size_t)error.rows);
        }
        // ---- the fused chains: trackFrames (:154-179) and extractFeatures + findLeftFeaturesInRight (:74-122, y_threshold 0.9)
        {
            std::vector<cv::Point2f> cur, lp, rp;
            std::vector<int> idx;
            svo::trackFrames(a, b, fs.points, cur, idx);
            put_i(fo, (int)cur.size());
            put(fo, idx.data(), 4 * idx.size());
            put(fo, cur.data(), 8 * cur.size());
            svo::extractStereoFeatures(a, b, prevPts, 20, true, 0.9f, lp, rp);
            put_i(fo, (int)lp.size());
            put(fo, lp.data(), 8 * lp.size());
            put(fo, rp.data(), 8 * rp.size());
        }
        // ---- empty inputs behave like OpenCV
        std::vector<cv::Point2f> none, out;
        std::vector<uchar> st;
        cv::Mat er;
        svo::calcOpticalFlowPyrLK(a, b, none, out, st, er);
        std::vector<cv::KeyPoint> kp3(3);
        mDetector->detect(cv::Mat(), kp3);
        put_i(fo, (int)(out.size() + st.size() + kp3.size() + (er.empty() ? 0 : 1)));
        // ---- colour frames (include/async_image_loader.h:66-69): BGR = (a, b, 255 - a); the device-side cvtColor, and the
        //      detector fed the CV_8UC3 frame directly, must agree with the gray path
        {
            cv::Mat bgr(h, w, CV_8UC3), gray;
            for (int y = 0; y < h; ++y)
                for (int x = 0; x < w; ++x) {
                    uchar* p = bgr.ptr<uchar>(y) + 3 * x;
                    p[0] = a.at<uchar>(y, x); p[1] = b.at<uchar>(y, x); p[2] = (uchar)(255 - a.at<uchar>(y, x));
                }
            svo::cvtColorBGR2GRAY(bgr, gray);
            put(fo, gray.data, (size_t)w * h);
            std::vector<cv::KeyPoint> kg, kc;
            mDetector->detect(gray, kg);
            mDetector->detect(bgr, kc);
            int same_c = kg.size() == kc.size();
            for (size_t i = 0; same_c && i < kg.size(); ++i)
                same_c = kg[i].pt.x == kc[i].pt.x && kg[i].pt.y == kc[i].pt.y && kg[i].response == kc[i].response;
            put_i(fo, same_c);
            put_i(fo, (int)kg.size());
        }
    } catch (const cv::Exception& e) {
        fprintf(stderr, "frontend_selftest: %s\n", e.what());
        fclose(fo);
        return 1;
    }
    fclose(fo);
    svo::releaseContexts();
    return 0;
}
