// TEST INFRASTRUCTURE ONLY (oracle/).  C entry point around the reference's own FeatureSet::bucketingFeatures
// (/root/reference/src/bucket.cpp:24-68, include/bucket.h:21-32), which is compiled UNMODIFIED from where it lies by
// oracle/Makefile into oracle/_ref/libref_bucket.so.  Used by tests/ to pin oracle/svo_oracle.c:orc_bucket and the CUDA
// bucketing kernel, and by bench.py's CPU-baseline leg.  No reference source is copied into this repo.
#include <opencv2/core/mat.hpp>

#include <cstring>

#include "bucket.h"

extern "C" __attribute__((visibility("default")))
int ref_bucketing(float* xy, int* ages, int n_in, int img_w, int img_h, int bucket_size, int per_bucket, int capacity,
                  int* n_out)
{
    FeatureSet fs;
    fs.points.reserve(n_in);
    fs.ages.reserve(n_in);
    for (int i = 0; i < n_in; ++i) {
        fs.points.emplace_back(xy[2 * i], xy[2 * i + 1]);
        fs.ages.push_back(ages[i]);
    }
    cv::Mat image(img_h, img_w);
    fs.bucketingFeatures(image, bucket_size, per_bucket);
    *n_out = fs.size();
    if (fs.size() > capacity) return -2;
    for (int i = 0; i < fs.size(); ++i) {
        xy[2 * i] = fs.points[i].x;
        xy[2 * i + 1] = fs.points[i].y;
        ages[i] = fs.ages[i];
    }
    return 0;
}

// FeatureSet::appendNewFeatures (src/bucket.cpp:18-22): appended points get age 0.
extern "C" __attribute__((visibility("default")))
int ref_append_then_size(const float* xy, int n)
{
    FeatureSet fs;
    std::vector<cv::Point2f> pts;
    for (int i = 0; i < n; ++i) pts.emplace_back(xy[2 * i], xy[2 * i + 1]);
    fs.appendNewFeatures(pts);
    for (int a : fs.ages)
        if (a != 0) return -1;
    return fs.size();
}
