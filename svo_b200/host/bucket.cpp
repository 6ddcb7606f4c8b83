// bucket.cpp -- FeatureSet::bucketingFeatures on the GPU (reference src/bucket.cpp:24-68 via svo_gpu_bucket).
#include "bucket.h"

#include <algorithm>
#include <string>

void FeatureSet::bucketingFeatures(const cv::Mat& image, int bucket_size, int features_per_bucket)
{
    static_assert(sizeof(cv::Point2f) == 2 * sizeof(float), "Point2f must be two packed floats");
    if (points.size() != ages.size()) SVO_THROW("FeatureSet: points and ages differ in length");
    if (bucket_size < 1) SVO_THROW("FeatureSet::bucketingFeatures: bucket_size must be >= 1");
    const int n = size();
    // the emission can be longer than the input (aliased cells are emitted twice): size for the worst case
    const long long cells = (long long)(image.rows / bucket_size + 1) * (image.cols / bucket_size + 1);
    const long long worst = std::max<long long>(n, std::min<long long>(cells * std::max(features_per_bucket, 1), 2LL * n));
    svo_gpu_ctx* ctx = svo::defaultContext(svo::currentDevice(), 0, 0, (int)worst);
    points.resize((size_t)worst);
    ages.resize((size_t)worst);
    int n_out = 0;
    const int rc = svo_gpu_bucket(ctx, reinterpret_cast<float*>(points.data()), ages.data(), n, image.cols, image.rows,
                                  bucket_size, features_per_bucket, (int)worst, &n_out);
    if (rc != SVO_GPU_OK) {
        points.resize((size_t)n);
        ages.resize((size_t)n);
        SVO_THROW(std::string("svo_gpu_bucket: error ") + std::to_string(rc) + ": " + svo_gpu_last_error(ctx));
    }
    points.resize((size_t)n_out);
    ages.resize((size_t)n_out);
}
