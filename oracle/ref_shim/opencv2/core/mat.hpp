// TEST INFRASTRUCTURE ONLY (oracle/).  Minimal stand-in for <opencv2/core/mat.hpp>; see types.hpp next to this file.
#pragma once
#include <opencv2/core/types.hpp>
namespace cv {
struct Mat {
    int rows, cols;
    Mat() : rows(0), cols(0) {}
    Mat(int r, int c) : rows(r), cols(c) {}
};
}  // namespace cv
