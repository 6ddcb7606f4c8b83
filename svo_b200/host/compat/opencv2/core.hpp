// Minimal stand-in for the OpenCV types that cross the front-end boundary, used ONLY when the real OpenCV headers are
// absent (this build container has no C++ OpenCV; see svo_frontend.hpp, which picks real OpenCV when it is installed).
// Field names, constants and default values follow OpenCV 4.x so that code written against it compiles unchanged:
// the reference passes cv::Mat (CV_8UC1), std::vector<cv::Point2f>, std::vector<cv::KeyPoint>, std::vector<uchar>,
// cv::Size, cv::TermCriteria (src/tracking.cpp:76-86, :96-105, :155-165).
#pragma once
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#define SVO_COMPAT_OPENCV 1
#define CV_8U 0
#define CV_32F 5
#define CV_MAKETYPE(depth, cn) ((depth) + (((cn) - 1) << 3))
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_8UC3 CV_MAKETYPE(CV_8U, 3)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)

typedef unsigned char uchar;

namespace cv {

template <typename T> using Ptr = std::shared_ptr<T>;
template <typename T, typename... A> Ptr<T> makePtr(A&&... a) { return std::make_shared<T>(std::forward<A>(a)...); }

typedef std::string String;

// cv::Exception with OpenCV's constructor (code, err, func, file, line); what() carries the message
class Exception : public std::runtime_error {
public:
    Exception(int code_, const String& err_, const String& func_, const String& file_, int line_)
        : std::runtime_error(err_ + " (" + func_ + ", " + file_ + ":" + std::to_string(line_) + ")"), code(code_), err(err_),
          func(func_), file(file_), line(line_) {}
    int code;
    String err, func, file;
    int line;
};
namespace Error { enum Code { StsError = -2, StsBadArg = -5 }; }

template <typename T> struct Point_ {
    T x{}, y{};
    Point_() = default;
    Point_(T x_, T y_) : x(x_), y(y_) {}
};
typedef Point_<float> Point2f;
typedef Point_<int> Point;
template <typename T> Point_<T> operator+(const Point_<T>& a, const Point_<T>& b) { return {T(a.x + b.x), T(a.y + b.y)}; }
template <typename T> Point_<T> operator-(const Point_<T>& a, const Point_<T>& b) { return {T(a.x - b.x), T(a.y - b.y)}; }

struct Size {
    int width = 0, height = 0;
    Size() = default;
    Size(int w, int h) : width(w), height(h) {}
};

struct TermCriteria {
    enum Type { COUNT = 1, MAX_ITER = COUNT, EPS = 2 };
    int type = 0, maxCount = 0;
    double epsilon = 0;
    TermCriteria() = default;
    TermCriteria(int t, int c, double e) : type(t), maxCount(c), epsilon(e) {}
};

enum { OPTFLOW_USE_INITIAL_FLOW = 4, OPTFLOW_LK_GET_MIN_EIGENVALS = 8 };
enum { FILLED = -1 };

struct KeyPoint {
    Point2f pt;
    float size = 0, angle = -1, response = 0;
    int octave = 0, class_id = -1;
    KeyPoint() = default;
    KeyPoint(Point2f p, float s, float a = -1, float r = 0, int o = 0, int c = -1)
        : pt(p), size(s), angle(a), response(r), octave(o), class_id(c) {}
    static void convert(const std::vector<KeyPoint>& kps, std::vector<Point2f>& pts)
    {
        pts.resize(kps.size());
        for (size_t i = 0; i < kps.size(); ++i) pts[i] = kps[i].pt;
    }
};

// Dense 2-D array with shared ownership of its buffer: just what cv::Mat needs to be at this boundary.
class Mat {
public:
    int rows = 0, cols = 0;
    uchar* data = nullptr;
    size_t step = 0;   // bytes per row

    Mat() = default;
    Mat(int r, int c, int type) { create(r, c, type); }
    Mat(Size s, int type) { create(s.height, s.width, type); }
    Mat(Size s, int type, int fill) { create(s.height, s.width, type); if (data) memset(data, fill, step * rows); }
    Mat(int r, int c, int type, void* ext, size_t step_ = 0)   // external data, not owned
        : rows(r), cols(c), data(static_cast<uchar*>(ext)), type_(type)
    {
        step = step_ ? step_ : (size_t)c * elemSize();
    }
    void create(int r, int c, int type)
    {
        if (r == rows && c == cols && type == type_ && data) return;
        rows = r; cols = c; type_ = type;
        step = (size_t)c * elemSize();
        own_.reset(new uchar[step * (size_t)(r > 0 ? r : 0) + 1], std::default_delete<uchar[]>());
        data = own_.get();
    }
    int type() const { return type_; }
    int channels() const { return (type_ >> 3) + 1; }
    size_t elemSize() const { return ((type_ & 7) == CV_32F ? 4 : 1) * (size_t)channels(); }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    Size size() const { return Size(cols, rows); }
    bool isContinuous() const { return step == (size_t)cols * elemSize(); }
    template <typename T> T* ptr(int r = 0) { return reinterpret_cast<T*>(data + step * r); }
    template <typename T> const T* ptr(int r = 0) const { return reinterpret_cast<const T*>(data + step * r); }
    template <typename T> T& at(int r, int c = 0) { return ptr<T>(r)[c]; }
    template <typename T> const T& at(int r, int c = 0) const { return ptr<T>(r)[c]; }

private:
    int type_ = CV_8UC1;
    std::shared_ptr<uchar> own_;
};

// cv::InputArray as far as this boundary needs it: a read-only proxy a cv::Mat converts to implicitly, so that
// cv::Feature2D::detect has OpenCV's exact signature here too and the wrappers are ONE source for both header sets.
class _InputArray {
public:
    _InputArray() = default;
    _InputArray(const Mat& m) : m_(&m) {}   // implicit, as in OpenCV
    Mat getMat(int = -1) const { return m_ ? *m_ : Mat(); }
    bool empty() const { return !m_ || m_->empty(); }

private:
    const Mat* m_ = nullptr;
};
typedef const _InputArray& InputArray;
inline InputArray noArray()
{
    static const _InputArray none;
    return none;
}

// cv::Feature2D as far as the reference uses it: held as cv::Ptr<cv::FeatureDetector> mDetector (include/tracking.h:50),
// called as mDetector->detect(image, keypoints, mask) (src/tracking.cpp:82).  Same virtual as OpenCV's features2d.hpp.
class Feature2D {
public:
    virtual ~Feature2D() = default;
    virtual void detect(InputArray image, std::vector<KeyPoint>& keypoints, InputArray mask = noArray()) = 0;
};
typedef Feature2D FeatureDetector;

}  // namespace cv
