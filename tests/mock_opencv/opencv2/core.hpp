// Minimal stand-in for the parts of <opencv2/core.hpp> that include/b2orb_shim/ORBextractor.h touches.
// TEST INFRASTRUCTURE: this image has no OpenCV C++ headers, so the shim is compile- and run-tested against this.
#pragma once
#include <cassert>
#include <cstdint>
#include <cstring>
#include <vector>
#define CV_8U 0
#define CV_8UC1 0
#define CV_Assert(x) assert(x)
namespace cv {
struct Point2f { float x, y; };
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
    Mat(int r, int c, int /*type*/) { create(r, c, CV_8U); }
    Mat(int r, int c, int /*type*/, void* d, size_t s) : rows(r), cols(c), step(s), data((unsigned char*)d) {}
    void create(int r, int c, int /*type*/) { rows = r; cols = c; step = c; buf_.assign((size_t)r * c, 0); data = buf_.data(); }
    void release() { rows = cols = 0; buf_.clear(); data = nullptr; }
    bool empty() const { return rows == 0 || cols == 0; }
    int type() const { return CV_8U; }
    bool isContinuous() const { return step == (size_t)cols; }
    Mat clone() const { Mat m; m.create(rows, cols, CV_8U); for (int i = 0; i < rows; ++i) std::memcpy(m.ptr(i), ptr(i), cols); return m; }
    unsigned char* ptr(int r) { return data + (size_t)r * step; }
    const unsigned char* ptr(int r) const { return data + (size_t)r * step; }
    Mat& getMat() { return *this; }
    const Mat& getMat() const { return *this; }
private:
    std::vector<unsigned char> buf_;
};
typedef const Mat& InputArray;
typedef Mat& OutputArray;
}  // namespace cv
