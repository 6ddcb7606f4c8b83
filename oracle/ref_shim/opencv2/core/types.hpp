// TEST INFRASTRUCTURE ONLY (oracle/).  Minimal stand-in for <opencv2/core/types.hpp> so that the reference's own
// src/bucket.cpp + include/bucket.h compile UNMODIFIED from /root/reference (C++ OpenCV is not installed in this image).
// bucket.cpp only uses cv::Point2f{x,y} (include/bucket.h:4,31) and cv::Mat::{rows,cols} (src/bucket.cpp:29-30).
#pragma once
namespace cv {
struct Point2f {
    float x, y;
    Point2f() : x(0), y(0) {}
    Point2f(float x_, float y_) : x(x_), y(y_) {}
};
}  // namespace cv
