// Minimal stand-in for the parts of <opencv2/core.hpp> that include/b2orb_shim/ORBextractor.h touches.
// TEST INFRASTRUCTURE: this image has no OpenCV C++ headers, so the shim is compile- and run-tested against this.
#pragma once
#include <cassert>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>
#define CV_8U 0
#define CV_8UC1 0
#define CV_64F 6
#define CV_Assert(x) assert(x)
namespace cv {
struct Point2f { float x, y; };
struct Point2d {
    double x, y;
    Point2d() : x(0), y(0) {}
    Point2d(double x_, double y_) : x(x_), y(y_) {}
    Point2d(const Point2f& p) : x(p.x), y(p.y) {}   // the widening _curPixelsL.push_back(keypoint.pt) performs
};
struct KeyPoint {
    Point2f pt; float size, angle, response; int octave, class_id;
    KeyPoint() : pt{0, 0}, size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float s, float a = -1, float r = 0, int o = 0, int c = -1)
        : pt{x, y}, size(s), angle(a), response(r), octave(o), class_id(c) {}
};
struct DMatch {
    int queryIdx, trainIdx, imgIdx; float distance;
    DMatch(int q, int t, int i, float d) : queryIdx(q), trainIdx(t), imgIdx(i), distance(d) {}
};
class Mat {
public:
    int rows = 0, cols = 0; size_t step = 0; unsigned char* data = nullptr;
    Mat() {}
    Mat(int r, int c, int type) { create(r, c, type); }
    Mat(int r, int c, int type, void* d, size_t s) : rows(r), cols(c), step(s), data((unsigned char*)d), type_(type) {}
    void create(int r, int c, int type) {
        rows = r; cols = c; type_ = type; step = (size_t)c * elemSize();
        buf_ = std::make_shared<std::vector<unsigned char>>((size_t)r * step, 0); data = buf_->data();
    }
    void release() { rows = cols = 0; buf_.reset(); data = nullptr; }
    bool empty() const { return rows == 0 || cols == 0; }
    int type() const { return type_; }
    size_t elemSize() const { return type_ == CV_64F ? 8 : 1; }
    bool isContinuous() const { return step == (size_t)cols * elemSize(); }
    Mat clone() const { Mat m; m.create(rows, cols, type_); for (int i = 0; i < rows; ++i) std::memcpy(m.ptr(i), ptr(i), cols * elemSize()); return m; }
    template <typename T> T& at(int r, int c) { return ((T*)(data + (size_t)r * step))[c]; }
    template <typename T> const T& at(int r, int c) const { return ((const T*)(data + (size_t)r * step))[c]; }
    unsigned char* ptr(int r) { return data + (size_t)r * step; }
    const unsigned char* ptr(int r) const { return data + (size_t)r * step; }
    Mat& getMat() { return *this; }
    const Mat& getMat() const { return *this; }
private:
    std::shared_ptr<std::vector<unsigned char>> buf_;   // copies share the pixels, like cv::Mat's reference counting
    int type_ = CV_8U;
};
typedef const Mat& InputArray;
typedef Mat& OutputArray;
}  // namespace cv
