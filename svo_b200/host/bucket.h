// bucket.h -- FeatureSet with the reference's public surface (include/bucket.h:21-32: size, clear, appendNewFeatures,
// bucketingFeatures, public members points / ages), its bucketing running on the GPU through svo_gpu_bucket.
// The reference's helper types Bucket (include/bucket.h:35-50) and FeaturePoint (:9-19) have no counterpart here: the
// per-cell containers exist only as device arrays inside libsvo_gpu.so.
#pragma once
#include "svo_frontend.hpp"

struct FeatureSet {
    std::vector<cv::Point2f> points;   // same order semantics as the reference: bucketing rewrites both vectors in place
    std::vector<int> ages;

    [[nodiscard]] int size() const { return static_cast<int>(points.size()); }
    void clear()
    {
        points.clear();
        ages.clear();
    }
    // src/bucket.cpp:18-22: appended features start with age 0
    void appendNewFeatures(const std::vector<cv::Point2f>& features)
    {
        points.insert(points.end(), features.begin(), features.end());
        ages.insert(ages.end(), features.size(), 0);
    }
    // src/bucket.cpp:24-68, bit-exact including the slot-0 overwrite and the duplicate emission of aliased cells
    // (SURVEY.md Appendix B).  `image` is used for rows / cols only, as in the reference (:29-30).
    // A point outside the bucket table (undefined behaviour in the reference) throws cv::Exception.
    void bucketingFeatures(const cv::Mat& image, int bucket_size, int features_per_bucket);
};
