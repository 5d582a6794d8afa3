// Minimal but real stand-in for the OpenCV C++ API surface that
//   (1) include/b2orb_shim/ORBextractor.h (the drop-in shim) and
//   (2) the reference's UNMODIFIED /root/reference/src/ORBextractor.cc + include/ORBextractor.h
// touch.  TEST INFRASTRUCTURE: this image has no OpenCV C++ headers or libraries.  (1) is compile- and
// run-tested against it (tests/test_cpp_shim.py); (2) is compiled against it into oracle/_ref/libref_orb.so
// (oracle/ref/Makefile), the image primitives (FAST, resize, GaussianBlur, fastAtan2, copyMakeBorder) being
// forwarded by oracle/ref/ref_glue.cpp to the cv2-4.13-pinned restatements of oracle/orb_oracle.cpp.
//
// cv::Mat here is a reference-counted 2-D byte/double matrix with ROI views that share the parent's pixels
// (rowRange / colRange / operator()(Rect)), which is what ComputePyramid relies on
// (mvImagePyramid[l] is a view into a bordered `temp`, src/ORBextractor.cc:1114-1115).  Pixel storage comes from
// malloc, not operator new, so a build that replaces operator new (oracle/ref: monotonic arena) leaves it alone.
#pragma once
#include <algorithm>   // the real <opencv2/core.hpp> pulls these in; src/ORBextractor.cc relies on it (std::sort, :684)
#include <cassert>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <iterator>
#include <vector>

#define CV_8U 0
#define CV_8UC1 0
#define CV_64F 6
#define CV_PI 3.1415926535897932384626433832795
#define CV_Assert(x) assert(x)

typedef unsigned char uchar;

// cvRound: round half to even (lrint under the default rounding mode == cvtsd2si), all overloads OpenCV >= 3 has.
static inline int cvRound(double v) { return (int)std::lrint(v); }
static inline int cvRound(float v) { return (int)std::lrintf(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { int i = (int)v; return i - (i > v); }
static inline int cvFloor(float v) { int i = (int)v; return i - (i > v); }
static inline int cvCeil(double v) { int i = (int)v; return i + (i < v); }
static inline int cvCeil(float v) { int i = (int)v; return i + (i < v); }

namespace cv {

enum { BORDER_CONSTANT = 0, BORDER_REPLICATE = 1, BORDER_REFLECT = 2, BORDER_WRAP = 3, BORDER_REFLECT_101 = 4,
       BORDER_REFLECT101 = 4, BORDER_DEFAULT = 4, BORDER_ISOLATED = 16 };
enum { INTER_NEAREST = 0, INTER_LINEAR = 1 };
enum { NORM_HAMMING = 6 };

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T x_, T y_) : x(x_), y(y_) {}
    template <typename U> Point_(const Point_<U>& p) : x((T)p.x), y((T)p.y) {}   // e.g. Point2d(keypoint.pt)
    template <typename S> Point_& operator*=(S s) { x = (T)(x * s); y = (T)(y * s); return *this; }
    Point_& operator+=(const Point_& o) { x += o.x; y += o.y; return *this; }
};
typedef Point_<int> Point2i;
typedef Point_<int> Point;
typedef Point_<float> Point2f;
typedef Point_<double> Point2d;

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};
struct Rect {
    int x, y, width, height;
    Rect() : x(0), y(0), width(0), height(0) {}
    Rect(int x_, int y_, int w, int h) : x(x_), y(y_), width(w), height(h) {}
};

struct KeyPoint {
    Point2f pt; float size, angle, response; int octave, class_id;
    KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(Point2f p, float s, float a = -1, float r = 0, int o = 0, int c = -1)
        : pt(p), size(s), angle(a), response(r), octave(o), class_id(c) {}
    KeyPoint(float x, float y, float s, float a = -1, float r = 0, int o = 0, int c = -1)
        : pt(x, y), size(s), angle(a), response(r), octave(o), class_id(c) {}
};
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout");

struct DMatch {
    int queryIdx, trainIdx, imgIdx; float distance;
    DMatch() : queryIdx(-1), trainIdx(-1), imgIdx(-1), distance(0) {}
    DMatch(int q, int t, int i, float d) : queryIdx(q), trainIdx(t), imgIdx(i), distance(d) {}
};

class Mat;
// What Mat::zeros returns: assigning it to a Mat of the same size and type fills that Mat IN PLACE
// (Mat::create is a no-op then) — computeDescriptors zeroes a rowRange view of the output this way
// (src/ORBextractor.cc:1037, 1089).
struct MatExpr { int rows, cols, type; };

class Mat {
public:
    int rows = 0, cols = 0; size_t step = 0; unsigned char* data = nullptr;
    Mat() {}
    Mat(int r, int c, int type) { create(r, c, type); }
    Mat(Size sz, int type) { create(sz.height, sz.width, type); }
    Mat(int r, int c, int type, void* d, size_t s = 0)
        : rows(r), cols(c), step(s ? s : (size_t)c * esz(type)), data((unsigned char*)d), type_(type) {}
    Mat(const Mat& m) : rows(m.rows), cols(m.cols), step(m.step), data(m.data), blk_(m.blk_), type_(m.type_) { ref(); }
    Mat(const MatExpr& e) { *this = e; }
    ~Mat() { unref(); }
    Mat& operator=(const Mat& m) {
        if (this != &m) {
            m.ref(); unref();
            rows = m.rows; cols = m.cols; step = m.step; data = m.data; blk_ = m.blk_; type_ = m.type_;
        }
        return *this;
    }
    Mat& operator=(const MatExpr& e) {
        create(e.rows, e.cols, e.type);
        for (int i = 0; i < rows; ++i) std::memset(ptr(i), 0, (size_t)cols * elemSize());
        return *this;
    }
    static MatExpr zeros(int r, int c, int type) { return MatExpr{r, c, type}; }
    void create(int r, int c, int type) {
        if (data && rows == r && cols == c && type_ == type) return;   // OpenCV: nothing to do
        unref();
        rows = r; cols = c; type_ = type; step = (size_t)c * elemSize();
        blk_ = (int*)std::calloc(1, 16 + (size_t)r * step + 1);
        *blk_ = 1;
        data = (unsigned char*)blk_ + 16;
    }
    void create(Size sz, int type) { create(sz.height, sz.width, type); }
    void release() { unref(); rows = cols = 0; step = 0; data = nullptr; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int type() const { return type_; }
    size_t elemSize() const { return esz(type_); }
    size_t step1() const { return step / elemSize(); }
    Size size() const { return Size(cols, rows); }
    bool isContinuous() const { return step == (size_t)cols * elemSize(); }
    Mat clone() const {
        Mat m; m.create(rows, cols, type_);
        for (int i = 0; i < rows; ++i) std::memcpy(m.ptr(i), ptr(i), (size_t)cols * elemSize());
        return m;
    }
    void copyTo(Mat& m) const {
        m.create(rows, cols, type_);
        for (int i = 0; i < rows; ++i) std::memmove(m.ptr(i), ptr(i), (size_t)cols * elemSize());
    }
    Mat rowRange(int r0, int r1) const { return (*this)(Rect(0, r0, cols, r1 - r0)); }
    Mat colRange(int c0, int c1) const { return (*this)(Rect(c0, 0, c1 - c0, rows)); }
    Mat operator()(const Rect& r) const {
        assert(r.x >= 0 && r.y >= 0 && r.x + r.width <= cols && r.y + r.height <= rows);
        Mat m(*this);
        m.data = data + (size_t)r.y * step + (size_t)r.x * elemSize();
        m.rows = r.height; m.cols = r.width;
        return m;
    }
    template <typename T> T& at(int r, int c) { return ((T*)(data + (size_t)r * step))[c]; }
    template <typename T> const T& at(int r, int c) const { return ((const T*)(data + (size_t)r * step))[c]; }
    unsigned char* ptr(int r = 0) { return data + (size_t)r * step; }
    const unsigned char* ptr(int r = 0) const { return data + (size_t)r * step; }
    template <typename T> T* ptr(int r = 0) { return (T*)(data + (size_t)r * step); }
    template <typename T> const T* ptr(int r = 0) const { return (const T*)(data + (size_t)r * step); }
    Mat& getMat() { return *this; }
    const Mat& getMat() const { return *this; }
private:
    static size_t esz(int type) { return type == CV_64F ? 8 : 1; }
    void ref() const { if (blk_) __atomic_add_fetch(blk_, 1, __ATOMIC_RELAXED); }
    void unref() {
        if (blk_ && __atomic_sub_fetch(blk_, 1, __ATOMIC_ACQ_REL) == 0) std::free(blk_);
        blk_ = nullptr;
    }
    int* blk_ = nullptr;   // malloc'd block: refcount word, 16-byte header, pixels; null for user-data Mats
    int type_ = CV_8U;
};
typedef const Mat& InputArray;
typedef Mat& OutputArray;

// The image primitives of <opencv2/imgproc.hpp> / <opencv2/features2d.hpp> that src/ORBextractor.cc calls.
// Declared here, defined by whoever links (oracle/ref/ref_glue.cpp forwards them to the cv2-pinned restatements).
void FAST(InputArray image, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression = true);
void resize(InputArray src, OutputArray dst, Size dsize, double fx = 0, double fy = 0, int interpolation = INTER_LINEAR);
void copyMakeBorder(InputArray src, OutputArray dst, int top, int bottom, int left, int right, int borderType);
void GaussianBlur(InputArray src, OutputArray dst, Size ksize, double sigmaX, double sigmaY = 0,
                  int borderType = BORDER_DEFAULT);
float fastAtan2(float y, float x);
struct KeyPointsFilter { static void retainBest(std::vector<KeyPoint>& keypoints, int npoints); };

}  // namespace cv
