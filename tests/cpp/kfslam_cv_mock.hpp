// kfslam_cv_mock.hpp — the few OpenCV types the reference's hot-path call sites use
// (/root/reference/src/lib/VisualOdometry.cpp:21-62), for boxes WITHOUT OpenCV headers (this build
// container has none).  Layouts match OpenCV 4.x: cv::KeyPoint 28 bytes, cv::DMatch 16 bytes.  When the real
// <opencv2/core.hpp> is available kfslam_frontend_cv.hpp uses it instead and this file is not included.
#pragma once
#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace cv {

struct Point2f { float x = 0, y = 0; };
struct Point2d { double x = 0, y = 0; Point2d() = default; Point2d(double x_, double y_) : x(x_), y(y_) {} Point2d(const Point2f &p) : x(p.x), y(p.y) {} };
struct Point3d { double x = 0, y = 0, z = 0; Point3d() = default; Point3d(double x_, double y_, double z_) : x(x_), y(y_), z(z_) {} };

struct KeyPoint {
    Point2f pt;
    float size = 0, angle = -1, response = 0;
    int octave = 0, class_id = -1;
};

struct DMatch {
    int queryIdx = -1, trainIdx = -1, imgIdx = -1;
    float distance = 3.402823466e+38f;
};

static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout");
static_assert(sizeof(DMatch) == 16, "cv::DMatch layout");

enum { CV_8U = 0, CV_8UC1 = 0, CV_16U = 2, CV_16UC1 = 2, CV_8UC3 = 16 };
enum NormTypes { NORM_L2 = 4, NORM_HAMMING = 6 };

class Exception : public std::runtime_error {
public:
    int code;
    Exception(int c, const std::string &m) : std::runtime_error(m), code(c) {}
};

// Minimal row-major matrix for the element types the path uses (CV_8UC1 gray / mask / descriptors, CV_8UC3 colour,
// CV_16UC1 depth): owns its buffer or wraps external memory (like cv::Mat(rows, cols, type, data, step)).  step in bytes.
class Mat {
public:
    int rows = 0, cols = 0;
    size_t step = 0;
    uint8_t *data = nullptr;
    Mat() = default;
    Mat(int r, int c, int type) { create(r, c, type); }
    Mat(int r, int c, int type, void *ext, size_t st = 0) : rows(r), cols(c), step(st ? st : (size_t)c * esz(type)), data((uint8_t *)ext), type_(type) {}
    void create(int r, int c, int type)
    {
        if (r == rows && c == cols && type == type_ && own_) return;
        own_ = std::shared_ptr<uint8_t>(new uint8_t[(size_t)r * c * esz(type) + 1], std::default_delete<uint8_t[]>());
        rows = r; cols = c; type_ = type; step = (size_t)c * esz(type); data = own_.get();
    }
    void release() { own_.reset(); rows = cols = 0; step = 0; data = nullptr; }
    bool empty() const { return rows == 0 || cols == 0 || data == nullptr; }
    bool isContinuous() const { return step == (size_t)cols * elemSize(); }
    int type() const { return type_; }
    int channels() const { return type_ == CV_8UC3 ? 3 : 1; }
    size_t elemSize() const { return esz(type_); }
    uint8_t *ptr(int r = 0) { return data + (size_t)r * step; }
    const uint8_t *ptr(int r = 0) const { return data + (size_t)r * step; }
private:
    static size_t esz(int type) { return type == CV_8UC3 ? 3 : (type == CV_16UC1 ? 2 : 1); }
    std::shared_ptr<uint8_t> own_;
    int type_ = CV_8UC1;
};

using InputArray = const Mat &;
using OutputArray = Mat &;
inline const Mat &noArray() { static Mat m; return m; }

template <typename T> using Ptr = std::shared_ptr<T>;

} // namespace cv
