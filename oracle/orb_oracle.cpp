// orb_oracle.cpp — CPU ORACLE (test infrastructure, NOT product code).
//
// Dependency-free C++17 restatement of the reference's stereo visual front-end hot path:
//   * ORB_SLAM2::ORBextractor            (/root/reference/src/ORBextractor.cc)
//   * cfsd::FeatureTracker matching parts (/root/reference/src/feature-tracker.cpp:140-328)
// together with the OpenCV primitives those files delegate to (resize, FAST, GaussianBlur,
// fastAtan2, cvRound, BFMatcher) and glibc's cosf/sinf, restated from their published
// algorithms.  OpenCV is an un-vendored system dependency of the reference (CMakeLists.txt:136;
// "opencv 3.2.0" per Dockerfile.amd64:20) and cannot be built here.
//
// HOW THIS ORACLE IS PINNED
//   1. To the reference's own code: /root/reference/src/ORBextractor.cc is compiled UNMODIFIED against a mock of the
//      OpenCV headers into oracle/_ref/libref_orb.so (oracle/ref/Makefile, ref_glue.cpp) and run next to this file on
//      seven synthetic shapes, six real camera frames shipped with the reference, odd shapes / strided inputs / other
//      pyramid parameters, under both heap-address orders that the quadtree's pointer tie-break (:684) can see:
//      keypoints, order, descriptors and pyramid levels are bit-identical (tests/test_ref_parity.py).  Its outputs are
//      committed as tests/golden/ref/*.npz (tools/gen_ref_golden.py) so the check also runs where the tree is absent.
//   2. The OpenCV primitives under both (resize, FAST, GaussianBlur, fastAtan2, BFMatcher, remap, triangulatePoints)
//      to the real OpenCV 4.13 kernels of the cv2 wheel of this image (tests/test_oracle_vs_cv2.py), and the whole
//      pipeline to cv2-assembled golden vectors (tools/gen_golden.py -> tests/golden/*.npz).
//   The reference has no tests or golden vectors of its own (test/vio-test.cpp is empty).  What stays outside any pin
//   on this box: OpenCV 3.2's Gaussian taps (the version the Dockerfile names; 4.x taps are the default, 3.2's are a
//   runtime option restated from the published 3.2 algorithm) and musl's cosf/sinf (glibc 2.39 is what is here).
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library.  The product (pose-estimation_b200/csrc) never links or calls it.
//
// Build: see oracle/Makefile (g++ -O2 -ffp-contract=off, no -march → no FMA contraction, which
// mirrors the reference build, CMakeLists.txt:53 / Dockerfile.amd64:50).

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <list>
#include <thread>
#include <utility>
#include <vector>

namespace orc {

// ------------------------------------------------------------------------------------------
// Plain image view (8-bit, single channel), standing in for cv::Mat CV_8UC1.
// ------------------------------------------------------------------------------------------
struct Img {
    int w = 0, h = 0;
    std::vector<uint8_t> d;
    Img() {}
    Img(int w_, int h_) : w(w_), h(h_), d((size_t)w_ * h_) {}
    inline const uint8_t* row(int y) const { return d.data() + (size_t)y * w; }
    inline uint8_t* row(int y) { return d.data() + (size_t)y * w; }
    inline uint8_t at(int y, int x) const { return d[(size_t)y * w + x]; }
};

// cv::KeyPoint layout (28 bytes): pt.x, pt.y, size, angle, response, octave, class_id.
struct KeyPoint {
    float x, y, size, angle, response;
    int octave, class_id;
};

// cvRound: round-half-to-even (lrint / cvtsd2si semantics).  Used at
// src/ORBextractor.cc:81,115,119-120,442,460,1112.
static inline int cv_round(double v) { return (int)std::nearbyint(v); }
static inline int cv_floor(double v) { return (int)std::floor(v); }
static inline int cv_ceil(double v) { return (int)std::ceil(v); }

// ------------------------------------------------------------------------------------------
// cv::resize(src, dst, sz, 0, 0, INTER_LINEAR) for CV_8UC1 (call site src/ORBextractor.cc:1120).
// Published algorithm (OpenCV imgproc/resize.cpp): per-axis source index/fraction tables computed
// in float from a double scale, Q11 integer weights, horizontal pass in int32, vertical pass
//   out = ( ((b0*(S0>>4))>>16) + ((b1*(S1>>4))>>16) + 2 ) >> 2.
// ------------------------------------------------------------------------------------------
static void resize_axis_table(int src, int dst, std::vector<int>& ofs, std::vector<short>& w) {
    ofs.resize(dst);
    w.resize((size_t)2 * dst);
    const double scale = (double)src / dst;
    for (int d = 0; d < dst; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = cv_floor(f);
        f -= s;
        if (s < 0) { s = 0; f = 0.f; }
        if (s >= src - 1) { s = src - 1; f = 0.f; }
        ofs[d] = s;
        w[2 * d] = (short)cv_round((1.f - f) * 2048.f);
        w[2 * d + 1] = (short)cv_round(f * 2048.f);
    }
}

static void resize_linear(const Img& src, Img& dst) {
    std::vector<int> xo, yo;
    std::vector<short> xw, yw;
    resize_axis_table(src.w, dst.w, xo, xw);
    resize_axis_table(src.h, dst.h, yo, yw);
    std::vector<int> r0(dst.w), r1(dst.w);
    auto hrow = [&](int sy, std::vector<int>& out) {
        const uint8_t* S = src.row(sy);
        for (int dx = 0; dx < dst.w; ++dx) {
            int sx = xo[dx];
            int sx1 = std::min(sx + 1, src.w - 1);
            out[dx] = S[sx] * xw[2 * dx] + S[sx1] * xw[2 * dx + 1];
        }
    };
    for (int dy = 0; dy < dst.h; ++dy) {
        int sy = yo[dy];
        int sy1 = std::min(sy + 1, src.h - 1);
        hrow(sy, r0);
        hrow(sy1, r1);
        int b0 = yw[2 * dy], b1 = yw[2 * dy + 1];
        uint8_t* D = dst.row(dy);
        for (int dx = 0; dx < dst.w; ++dx) {
            int v = (((b0 * (r0[dx] >> 4)) >> 16) + ((b1 * (r1[dx] >> 4)) >> 16) + 2) >> 2;
            D[dx] = (uint8_t)std::min(std::max(v, 0), 255);
        }
    }
}

// ------------------------------------------------------------------------------------------
// cv::GaussianBlur(img, img, Size(7,7), 2, 2, BORDER_REFLECT_101) on CV_8UC1
// (call site src/ORBextractor.cc:1086).  OpenCV 4.x fixed-point path: separable integer taps
// (sum 256), horizontal pass exact in Q8, vertical accumulate in Q16, (acc + 32768) >> 16.
// Taps are a parameter because OpenCV 3.2 would round the float kernel to
// {18,34,49,55,49,34,18}; default = the 4.13 set, the one that can be checked on this box.
// ------------------------------------------------------------------------------------------
static const int kTaps413[7] = {18, 34, 48, 56, 48, 34, 18};

static inline int reflect101(int i, int n) {
    if (n == 1) return 0;
    while (i < 0 || i >= n) {
        if (i < 0) i = -i;
        else i = 2 * (n - 1) - i;
    }
    return i;
}

static void gaussian_blur7(const Img& src, Img& dst, const int taps[7]) {
    const int w = src.w, h = src.h;
    dst = Img(w, h);
    std::vector<uint16_t> H((size_t)w * h);
    for (int y = 0; y < h; ++y) {
        const uint8_t* S = src.row(y);
        for (int x = 0; x < w; ++x) {
            int acc = 0;
            for (int k = -3; k <= 3; ++k) acc += taps[k + 3] * S[reflect101(x + k, w)];
            H[(size_t)y * w + x] = (uint16_t)acc;
        }
    }
    for (int y = 0; y < h; ++y) {
        uint8_t* D = dst.row(y);
        const uint16_t* R[7];
        for (int k = -3; k <= 3; ++k) R[k + 3] = &H[(size_t)reflect101(y + k, h) * w];
        for (int x = 0; x < w; ++x) {
            unsigned acc = 0;
            for (int k = 0; k < 7; ++k) acc += (unsigned)taps[k] * R[k][x];
            unsigned v = (acc + 32768u) >> 16;
            D[x] = (uint8_t)std::min(v, 255u);
        }
    }
}

// ------------------------------------------------------------------------------------------
// cv::FAST(roi, keypoints, threshold, nonmaxSuppression=true), TYPE_9_16, on a sub-rectangle of
// a level image (call sites src/ORBextractor.cc:809-815).  Published algorithm (OpenCV
// features2d/fast.cpp, fast_score.cpp): 16-pixel Bresenham circle of radius 3; a pixel is a
// corner iff 9 contiguous circle pixels are all darker than v-t or all brighter than v+t;
// score = largest t for which it stays a corner; 3x3 strict non-max suppression over rolling
// score rows (non-corners and pixels outside the 3-px inset of the ROI score 0); row-major output.
// The ROI is [x0,x1) x [y0,y1) of `img`; like cv::FAST on a submatrix, nothing outside the ROI
// is read.  Returned coordinates are relative to the ROI origin.
// ------------------------------------------------------------------------------------------
static const int kCircle[16][2] = {{0, 3},  {1, 3},   {2, 2},   {3, 1},   {3, 0},  {3, -1},
                                   {2, -2}, {1, -3},  {0, -3},  {-1, -3}, {-2, -2}, {-3, -1},
                                   {-3, 0}, {-3, 1},  {-2, 2},  {-1, 3}};

// b(p) = max over the 16 arcs of 9 of max(min_k d_k, min_k(-d_k)), d_k = I(p) - I(p + o_k).
// "corner at threshold t" <=> b > t, and cornerScore() == b - 1 for every detected corner.
static inline int fast_b(const uint8_t* p, const int* off) {
    int d[25];
    const int v = p[0];
    for (int k = 0; k < 16; ++k) d[k] = v - p[off[k]];
    for (int k = 16; k < 25; ++k) d[k] = d[k - 16];
    int best = 0;
    for (int s = 0; s < 16; ++s) {
        int mn = d[s], mx = d[s];
        for (int k = 1; k < 9; ++k) {
            mn = std::min(mn, d[s + k]);
            mx = std::max(mx, d[s + k]);
        }
        best = std::max(best, std::max(mn, -mx));
    }
    return best;
}

static void fast9_16_roi(const Img& img, int x0, int y0, int x1, int y1, int threshold,
                         std::vector<KeyPoint>& out) {
    out.clear();
    const int cols = x1 - x0, rows = y1 - y0, step = img.w;
    if (cols < 7 || rows < 7) return;
    int off[16];
    for (int k = 0; k < 16; ++k) off[k] = kCircle[k][0] + kCircle[k][1] * step;
    std::vector<uint8_t> sbuf((size_t)3 * cols, 0);
    std::vector<int> cbuf((size_t)3 * (cols + 1), 0);
    uint8_t* buf[3] = {sbuf.data(), sbuf.data() + cols, sbuf.data() + 2 * cols};
    int* cpbuf[3] = {cbuf.data(), cbuf.data() + (cols + 1), cbuf.data() + 2 * (cols + 1)};
    for (int i = 3; i < rows - 2; ++i) {
        const uint8_t* ptr = img.row(y0 + i) + x0 + 3;
        uint8_t* curr = buf[(i - 3) % 3];
        int* cornerpos = cpbuf[(i - 3) % 3] + 1;
        std::memset(curr, 0, cols);
        int ncorners = 0;
        if (i < rows - 3) {
            for (int j = 3; j < cols - 3; ++j, ++ptr) {
                const int v = ptr[0];
                const int lo = v - threshold, hi = v + threshold;
                // high-speed rejection on the 4 compass points (any 9-arc holds >= 2 adjacent ones)
                int a = ptr[off[0]], b = ptr[off[8]];
                int nd = (a < lo) + (b < lo), nb = (a > hi) + (b > hi);
                if (nd == 0 && nb == 0) continue;
                a = ptr[off[4]];
                b = ptr[off[12]];
                nd += (a < lo) + (b < lo);
                nb += (a > hi) + (b > hi);
                if (nd < 2 && nb < 2) continue;
                // 16-bit darker / brighter ring masks; 9 contiguous set bits on the doubled ring
                unsigned md = 0, mb = 0;
                for (int k = 0; k < 16; ++k) {
                    const int x = ptr[off[k]];
                    md |= (unsigned)(x < lo) << k;
                    mb |= (unsigned)(x > hi) << k;
                }
                auto run9 = [](unsigned m) {
                    m |= m << 16;
                    unsigned c = m & (m >> 1);
                    c &= c >> 2;
                    c &= c >> 4;
                    c &= m >> 8;
                    return c != 0;
                };
                const bool corner = run9(md) || run9(mb);
                if (corner) {
                    cornerpos[ncorners++] = j;
                    curr[j] = (uint8_t)(fast_b(ptr, off) - 1);
                }
            }
        }
        cornerpos[-1] = ncorners;
        if (i == 3) continue;
        const uint8_t* prev = buf[(i - 4 + 3) % 3];
        const uint8_t* pprev = buf[(i - 5 + 3) % 3];
        cornerpos = cpbuf[(i - 4 + 3) % 3] + 1;
        ncorners = cornerpos[-1];
        for (int k = 0; k < ncorners; ++k) {
            const int j = cornerpos[k];
            const int score = prev[j];
            if (score > prev[j + 1] && score > prev[j - 1] && score > pprev[j - 1] &&
                score > pprev[j] && score > pprev[j + 1] && score > curr[j - 1] &&
                score > curr[j] && score > curr[j + 1]) {
                KeyPoint kp{(float)j, (float)(i - 1), 7.f, -1.f, (float)score, 0, -1};
                out.push_back(kp);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------
// cv::fastAtan2(y, x) — degrees in [0,360), OpenCV core/mathfuncs_core: fp32 7th-order odd
// polynomial, evaluated WITHOUT fused multiply-add (call site src/ORBextractor.cc:103).
// ------------------------------------------------------------------------------------------
static float fast_atan2_deg(float y, float x) {
    const float scale = (float)(180.0 / 3.14159265358979323846);
    const float p1 = 0.9997878412794807f * scale, p3 = -0.3258083974640975f * scale,
                p5 = 0.1555786518463281f * scale, p7 = -0.04432655554792128f * scale;
    const float eps = (float)2.2204460492503131e-16;
    float ax = std::fabs(x), ay = std::fabs(y);
    float a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + eps);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + eps);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

// ------------------------------------------------------------------------------------------
// glibc 2.39 cosf/sinf (sysdeps/ieee754/flt-32/s_cosf.c, s_sinf.c, sincosf.h, sincosf_data.c —
// the ARM optimized-routines implementation): double-precision range reduction by pi/2 and
// degree-8/7 polynomials.  Restated for the ranges the path can reach (|y| < 120); verified
// bit-identical to the libm of this image for ALL floats in [0, 6.2832] (tests + tools).
// The reference calls them at src/ORBextractor.cc:113 (`using namespace std` picks the float
// overloads).  Restating instead of calling libm keeps the oracle identical on every host.
// ------------------------------------------------------------------------------------------
static const double kHpiInv = 0x1.45F306DC9C883p+23, kHpi = 0x1.921FB54442D18p0;
static const double kC0 = 0x1p0, kC1 = -0x1.ffffffd0c621cp-2, kC2 = 0x1.55553e1068f19p-5,
                    kC3 = -0x1.6c087e89a359dp-10, kC4 = 0x1.99343027bf8c3p-16;
static const double kS1 = -0x1.555545995a603p-3, kS2 = 0x1.1107605230bc4p-7,
                    kS3 = -0x1.994eb3774cf24p-13;

static inline uint32_t abstop12(float f) {
    uint32_t u;
    std::memcpy(&u, &f, 4);
    return (u >> 20) & 0x7ff;
}
// n even -> sine polynomial, n odd -> cosine polynomial (negated when `neg`).
static inline float sincos_poly(double x, double x2, int n, bool neg) {
    if ((n & 1) == 0) {
        double x3 = x * x2;
        double s1 = kS2 + x2 * kS3;
        double x7 = x3 * x2;
        double s = x + x3 * kS1;
        return (float)(s + x7 * s1);
    }
    const double sg = neg ? -1.0 : 1.0;
    double x4 = x2 * x2;
    double c2 = sg * kC3 + x2 * (sg * kC4);
    double c1 = sg * kC0 + x2 * (sg * kC1);
    double x6 = x4 * x2;
    double c = c1 + x4 * (sg * kC2);
    return (float)(c + x6 * c2);
}
static inline double reduce_fast(double x, int* np) {
    double r = x * kHpiInv;
    int n = ((int32_t)r + 0x800000) >> 24;
    *np = n;
    return x - n * kHpi;
}
static const double kSign[4] = {1.0, -1.0, -1.0, 1.0};

static float ref_cosf(float y) {
    double x = y;
    int n;
    if (abstop12(y) < abstop12(0x1.921FB6p-1f)) {
        if (abstop12(y) < abstop12(0x1p-12f)) return 1.0f;
        return sincos_poly(x, x * x, 1, false);
    }
    x = reduce_fast(x, &n);
    return sincos_poly(x * kSign[n & 3], x * x, n ^ 1, (n & 2) != 0);
}
static float ref_sinf(float y) {
    double x = y;
    int n;
    if (abstop12(y) < abstop12(0x1.921FB6p-1f)) {
        if (abstop12(y) < abstop12(0x1p-12f)) return y;
        return sincos_poly(x, x * x, 0, false);
    }
    x = reduce_fast(x, &n);
    return sincos_poly(x * kSign[n & 3], x * x, n, (n & 2) != 0);
}

// ------------------------------------------------------------------------------------------
// ExtractorNode / DistributeOctTree — src/ORBextractor.cc:481-763, include/ORBextractor.h:32-43.
// Literal std::list restatement.  The reference sorts pair<int, ExtractorNode*> at :684, i.e.
// breaks ties in point count by heap ADDRESS (non-deterministic).  Documented deterministic rule
// used here and by the CUDA path: every node gets a creation sequence number
// and stands in for the pointer — among equal counts the node created LATER is expanded first.
// ------------------------------------------------------------------------------------------
struct Node {
    int ulx, uly, urx, ury, blx, bly, brx, bry;  // UL, UR, BL, BR (cv::Point2i)
    std::vector<KeyPoint> keys;
    std::list<Node>::iterator lit;
    bool no_more = false;
    long seq = 0;
};

// ExtractorNode::DivideNode, src/ORBextractor.cc:481-537.
static void divide_node(const Node& p, Node& n1, Node& n2, Node& n3, Node& n4) {
    const int halfX = (int)std::ceil((float)(p.urx - p.ulx) / 2);
    const int halfY = (int)std::ceil((float)(p.bry - p.uly) / 2);
    n1.ulx = p.ulx; n1.uly = p.uly;
    n1.urx = p.ulx + halfX; n1.ury = p.uly;
    n1.blx = p.ulx; n1.bly = p.uly + halfY;
    n1.brx = p.ulx + halfX; n1.bry = p.uly + halfY;
    n2.ulx = n1.urx; n2.uly = n1.ury;
    n2.urx = p.urx; n2.ury = p.ury;
    n2.blx = n1.brx; n2.bly = n1.bry;
    n2.brx = p.urx; n2.bry = p.uly + halfY;
    n3.ulx = n1.blx; n3.uly = n1.bly;
    n3.urx = n1.brx; n3.ury = n1.bry;
    n3.blx = p.blx; n3.bly = p.bly;
    n3.brx = n1.brx; n3.bry = p.bly;
    n4.ulx = n3.urx; n4.uly = n3.ury;
    n4.urx = n2.brx; n4.ury = n2.bry;
    n4.blx = n3.brx; n4.bly = n3.bry;
    n4.brx = p.brx; n4.bry = p.bry;
    for (const KeyPoint& kp : p.keys) {
        if (kp.x < n1.urx) {
            if (kp.y < n1.bry) n1.keys.push_back(kp);
            else n3.keys.push_back(kp);
        } else if (kp.y < n1.bry) n2.keys.push_back(kp);
        else n4.keys.push_back(kp);
    }
    if (n1.keys.size() == 1) n1.no_more = true;
    if (n2.keys.size() == 1) n2.no_more = true;
    if (n3.keys.size() == 1) n3.no_more = true;
    if (n4.keys.size() == 1) n4.no_more = true;
}

// DistributeOctTree, src/ORBextractor.cc:539-763.
// tie_mode: +1 = among equal point counts the node created LATER is expanded first (default; models heap addresses
// that grow with allocation order), -1 = the node created EARLIER first (addresses that shrink).  Both orders are
// checked against the reference's own code under the matching allocator (oracle/ref/ref_glue.cpp).
// *ties (optional) counts the equal-count neighbours in the sorted vector that the expansion loop actually
// consumed or stopped at, i.e. whether the pointer comparison could influence this call's result at all.
static std::vector<KeyPoint> distribute_octtree(const std::vector<KeyPoint>& in, int minX, int maxX,
                                                int minY, int maxY, int N, int tie_mode = 1, int* ties = nullptr) {
    if (ties) *ties = 0;
    std::vector<KeyPoint> result;
    const int nIni = (int)std::round((float)(maxX - minX) / (maxY - minY));
    if (nIni < 1 || in.empty()) return result;  // reference divides by zero / indexes an empty vector here
    const float hX = (float)(maxX - minX) / nIni;
    long seq = 0;
    std::list<Node> nodes;
    std::vector<Node*> ini(nIni);
    for (int i = 0; i < nIni; ++i) {
        Node ni;
        ni.ulx = (int)(hX * (float)i); ni.uly = 0;
        ni.urx = (int)(hX * (float)(i + 1)); ni.ury = 0;
        ni.blx = ni.ulx; ni.bly = maxY - minY;
        ni.brx = ni.urx; ni.bry = maxY - minY;
        ni.seq = seq++;
        nodes.push_back(ni);
        ini[i] = &nodes.back();
    }
    for (const KeyPoint& kp : in) {
        int r = (int)(kp.x / hX);
        if (r >= nIni) r = nIni - 1;  // cannot happen for in-range x; the reference would index out of bounds
        ini[r]->keys.push_back(kp);
    }
    for (auto lit = nodes.begin(); lit != nodes.end();) {
        if (lit->keys.size() == 1) { lit->no_more = true; ++lit; }
        else if (lit->keys.empty()) lit = nodes.erase(lit);
        else ++lit;
    }

    typedef std::pair<int, long> SizeSeq;  // (point count, creation seq) — seq models the pointer
    std::vector<std::pair<SizeSeq, Node*>> sizeAndNode;
    bool finish = false;
    auto push_children = [&](Node* kids[4], int& nToExpand) {
        for (int c = 0; c < 4; ++c) {
            Node& n = *kids[c];
            if (n.keys.empty()) continue;
            n.seq = seq++;
            nodes.push_front(n);
            if (n.keys.size() > 1) {
                ++nToExpand;
                sizeAndNode.push_back({{(int)n.keys.size(), tie_mode * nodes.front().seq}, &nodes.front()});
                nodes.front().lit = nodes.begin();
            }
        }
    };
    while (!finish) {
        int prevSize = (int)nodes.size();
        int nToExpand = 0;
        sizeAndNode.clear();
        for (auto lit = nodes.begin(); lit != nodes.end();) {
            if (lit->no_more) { ++lit; continue; }
            Node n1, n2, n3, n4;
            divide_node(*lit, n1, n2, n3, n4);
            Node* kids[4] = {&n1, &n2, &n3, &n4};
            push_children(kids, nToExpand);
            lit = nodes.erase(lit);
        }
        if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) {
            finish = true;
        } else if ((int)nodes.size() + nToExpand * 3 > N) {
            while (!finish) {
                prevSize = (int)nodes.size();
                std::vector<std::pair<SizeSeq, Node*>> prev = sizeAndNode;
                sizeAndNode.clear();
                std::sort(prev.begin(), prev.end(),
                          [](const std::pair<SizeSeq, Node*>& a, const std::pair<SizeSeq, Node*>& b) {
                              return a.first < b.first;
                          });
                for (int j = (int)prev.size() - 1; j >= 0; --j) {
                    if (ties && j > 0 && prev[j].first.first == prev[j - 1].first.first) ++*ties;
                    Node n1, n2, n3, n4;
                    divide_node(*prev[j].second, n1, n2, n3, n4);
                    Node* kids[4] = {&n1, &n2, &n3, &n4};
                    int dummy = 0;
                    push_children(kids, dummy);
                    nodes.erase(prev[j].second->lit);
                    if ((int)nodes.size() >= N) break;
                }
                if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) finish = true;
            }
        }
    }
    result.reserve(nodes.size());
    for (const Node& n : nodes) {
        const KeyPoint* best = &n.keys[0];
        float maxResponse = best->response;
        for (size_t k = 1; k < n.keys.size(); ++k)
            if (n.keys[k].response > maxResponse) { best = &n.keys[k]; maxResponse = n.keys[k].response; }
        result.push_back(*best);
    }
    return result;
}

// ------------------------------------------------------------------------------------------
// ORBextractor — ctor tables (src/ORBextractor.cc:410-470), ComputePyramid (:1107-1132),
// ComputeKeyPointsOctTree (:765-853), IC_Angle (:77-104), computeOrbDescriptor (:108-147),
// operator() (:1043-1105).
// ------------------------------------------------------------------------------------------
static const int PATCH_SIZE = 31, HALF_PATCH_SIZE = 15, EDGE_THRESHOLD = 19;
static const int8_t kPattern[1024] = {
#include "orb_pattern.inc"
};

struct Extractor {
    int nfeatures, nlevels, iniThFAST, minThFAST;
    double scaleFactor;  // include/ORBextractor.h:98 declares it double; ctor takes float
    std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;
    std::vector<int> mnFeaturesPerLevel, umax;
    int taps[7];
    // per-stage wall-clock seconds accumulated over calls (bench.py's CPU legs print stage means like the reference's
    // benchmark programs do, benchmark/bundleAdjustment/bundleAdjustment.cpp:174-318): pyramid, FAST cells, quadtree,
    // orientation, blur, descriptors
    double stage_s[6] = {0, 0, 0, 0, 0, 0};
    static double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
    int tie_mode = 1;            // see distribute_octtree
    std::vector<int> ties;       // per level: equal-count comparisons that mattered in the last extract
    // intermediates kept for tests (the reference keeps mvImagePyramid as a public member too)
    std::vector<Img> pyramid, blurred;
    std::vector<std::vector<KeyPoint>> raw;  // vToDistributeKeys per level (coords rel. to border)

    Extractor(int nf, float sf, int nl, int ini, int mn)
        : nfeatures(nf), nlevels(nl), iniThFAST(ini), minThFAST(mn), scaleFactor(sf) {
        std::memcpy(taps, kTaps413, sizeof(taps));
        mvScaleFactor.resize(nlevels);
        mvLevelSigma2.resize(nlevels);
        mvScaleFactor[0] = 1.0f;
        mvLevelSigma2[0] = 1.0f;
        for (int i = 1; i < nlevels; ++i) {
            mvScaleFactor[i] = (float)(mvScaleFactor[i - 1] * scaleFactor);
            mvLevelSigma2[i] = mvScaleFactor[i] * mvScaleFactor[i];
        }
        mvInvScaleFactor.resize(nlevels);
        mvInvLevelSigma2.resize(nlevels);
        for (int i = 0; i < nlevels; ++i) {
            mvInvScaleFactor[i] = 1.0f / mvScaleFactor[i];
            mvInvLevelSigma2[i] = 1.0f / mvLevelSigma2[i];
        }
        mnFeaturesPerLevel.resize(nlevels);
        float factor = (float)(1.0f / scaleFactor);
        float nDesired = (float)(nfeatures * (1 - factor) /
                                 (1 - (float)std::pow((double)factor, (double)nlevels)));
        int sum = 0;
        for (int level = 0; level < nlevels - 1; ++level) {
            mnFeaturesPerLevel[level] = cv_round(nDesired);
            sum += mnFeaturesPerLevel[level];
            nDesired *= factor;
        }
        mnFeaturesPerLevel[nlevels - 1] = std::max(nfeatures - sum, 0);

        umax.resize(HALF_PATCH_SIZE + 1);
        int v, v0, vmax = cv_floor(HALF_PATCH_SIZE * std::sqrt(2.f) / 2 + 1);
        int vmin = cv_ceil(HALF_PATCH_SIZE * std::sqrt(2.f) / 2);
        const double hp2 = HALF_PATCH_SIZE * HALF_PATCH_SIZE;
        for (v = 0; v <= vmax; ++v) umax[v] = cv_round(std::sqrt(hp2 - v * v));
        for (v = HALF_PATCH_SIZE, v0 = 0; v >= vmin; --v) {
            while (umax[v0] == umax[v0 + 1]) ++v0;
            umax[v] = v0;
            ++v0;
        }
    }

    // ComputePyramid.  The 19-px reflect-101 apron the reference adds with copyMakeBorder is
    // never read by any later stage (FAST stays >=16 px inside, IC_Angle >=4, the descriptor reads
    // a blurred clone with its own reflect borders), so levels are stored un-padded.
    void compute_pyramid(const Img& image) {
        pyramid.assign(nlevels, Img());
        for (int level = 0; level < nlevels; ++level) {
            float scale = mvInvScaleFactor[level];
            int w = cv_round((float)image.w * scale), h = cv_round((float)image.h * scale);
            if (level == 0) pyramid[0] = image;
            else {
                pyramid[level] = Img(w, h);
                resize_linear(pyramid[level - 1], pyramid[level]);
            }
        }
    }

    // IC_Angle, src/ORBextractor.cc:77-104.
    float ic_angle(const Img& image, float ptx, float pty) const {
        int m_01 = 0, m_10 = 0;
        const int step = image.w;
        const uint8_t* center = image.row(cv_round(pty)) + cv_round(ptx);
        for (int u = -HALF_PATCH_SIZE; u <= HALF_PATCH_SIZE; ++u) m_10 += u * center[u];
        for (int v = 1; v <= HALF_PATCH_SIZE; ++v) {
            int v_sum = 0;
            int d = umax[v];
            for (int u = -d; u <= d; ++u) {
                int val_plus = center[u + v * step], val_minus = center[u - v * step];
                v_sum += (val_plus - val_minus);
                m_10 += u * (val_plus + val_minus);
            }
            m_01 += v * v_sum;
        }
        return fast_atan2_deg((float)m_01, (float)m_10);
    }

    // ComputeKeyPointsOctTree, src/ORBextractor.cc:765-853.
    void compute_keypoints(std::vector<std::vector<KeyPoint>>& all) {
        all.assign(nlevels, {});
        raw.assign(nlevels, {});
        ties.assign(nlevels, 0);
        const float W = 30;
        for (int level = 0; level < nlevels; ++level) {
            const Img& im = pyramid[level];
            const int minBorderX = EDGE_THRESHOLD - 3, minBorderY = minBorderX;
            const int maxBorderX = im.w - EDGE_THRESHOLD + 3, maxBorderY = im.h - EDGE_THRESHOLD + 3;
            std::vector<KeyPoint>& toDistribute = raw[level];
            const float width = (float)(maxBorderX - minBorderX), height = (float)(maxBorderY - minBorderY);
            const int nCols = (int)(width / W), nRows = (int)(height / W);
            if (nCols < 1 || nRows < 1) continue;  // the reference divides by zero here
            const int wCell = (int)std::ceil(width / nCols), hCell = (int)std::ceil(height / nRows);
            std::vector<KeyPoint> cell;
            double t0 = now_s();
            for (int i = 0; i < nRows; ++i) {
                const float iniY = (float)(minBorderY + i * hCell);
                float maxY = iniY + hCell + 6;
                if (iniY >= maxBorderY - 3) continue;
                if (maxY > maxBorderY) maxY = (float)maxBorderY;
                for (int j = 0; j < nCols; ++j) {
                    const float iniX = (float)(minBorderX + j * wCell);
                    float maxX = iniX + wCell + 6;
                    if (iniX >= maxBorderX - 6) continue;
                    if (maxX > maxBorderX) maxX = (float)maxBorderX;
                    fast9_16_roi(im, (int)iniX, (int)iniY, (int)maxX, (int)maxY, iniThFAST, cell);
                    if (cell.empty())
                        fast9_16_roi(im, (int)iniX, (int)iniY, (int)maxX, (int)maxY, minThFAST, cell);
                    for (KeyPoint& kp : cell) {
                        kp.x += j * wCell;
                        kp.y += i * hCell;
                        toDistribute.push_back(kp);
                    }
                }
            }
            stage_s[1] += now_s() - t0;
            t0 = now_s();
            std::vector<KeyPoint>& kps = all[level];
            kps = distribute_octtree(toDistribute, minBorderX, maxBorderX, minBorderY, maxBorderY,
                                     mnFeaturesPerLevel[level], tie_mode, &ties[level]);
            const int scaledPatchSize = (int)(PATCH_SIZE * mvScaleFactor[level]);
            for (KeyPoint& kp : kps) {
                kp.x += minBorderX;
                kp.y += minBorderY;
                kp.octave = level;
                kp.size = (float)scaledPatchSize;
            }
            stage_s[2] += now_s() - t0;
        }
        const double t1 = now_s();
        for (int level = 0; level < nlevels; ++level)
            for (KeyPoint& kp : all[level]) kp.angle = ic_angle(pyramid[level], kp.x, kp.y);
        stage_s[3] += now_s() - t1;
    }

    // computeOrbDescriptor, src/ORBextractor.cc:108-147.  fp32 products and sum are rounded
    // separately (no FMA), then cvRound (half-even).
    static void orb_descriptor(const KeyPoint& kpt, const Img& img, uint8_t* desc) {
        const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
        float angle = (float)kpt.angle * factorPI;
        float a = ref_cosf(angle), b = ref_sinf(angle);
        const int step = img.w;
        const uint8_t* center = img.row(cv_round(kpt.y)) + cv_round(kpt.x);
        const int8_t* pat = kPattern;
        auto get = [&](int idx) -> int {
            float px = (float)pat[2 * idx], py = (float)pat[2 * idx + 1];
            volatile float t0 = px * b, t1 = py * a, t2 = px * a, t3 = py * b;
            int dy = cv_round((float)(t0 + t1)), dx = cv_round((float)(t2 - t3));
            return center[dy * step + dx];
        };
        for (int i = 0; i < 32; ++i, pat += 32) {
            int val = 0;
            for (int j = 0; j < 8; ++j) {
                int t0 = get(2 * j), t1 = get(2 * j + 1);
                val |= (t0 < t1) << j;
            }
            desc[i] = (uint8_t)val;
        }
    }

    // operator(), src/ORBextractor.cc:1043-1105 (mask ignored, as in the reference).
    void extract(const Img& image, std::vector<KeyPoint>& keypoints, std::vector<uint8_t>& descriptors) {
        keypoints.clear();
        descriptors.clear();
        blurred.assign(nlevels, Img());
        if (image.w == 0 || image.h == 0) return;
        double t0 = now_s();
        compute_pyramid(image);
        stage_s[0] += now_s() - t0;
        std::vector<std::vector<KeyPoint>> all;
        compute_keypoints(all);
        size_t n = 0;
        for (auto& v : all) n += v.size();
        descriptors.assign(n * 32, 0);
        keypoints.reserve(n);
        size_t offset = 0;
        for (int level = 0; level < nlevels; ++level) {
            std::vector<KeyPoint>& kps = all[level];
            if (kps.empty()) continue;
            t0 = now_s();
            gaussian_blur7(pyramid[level], blurred[level], taps);
            stage_s[4] += now_s() - t0;
            t0 = now_s();
            for (size_t i = 0; i < kps.size(); ++i)
                orb_descriptor(kps[i], blurred[level], &descriptors[(offset + i) * 32]);
            stage_s[5] += now_s() - t0;
            offset += kps.size();
            if (level != 0) {
                float scale = mvScaleFactor[level];
                for (KeyPoint& kp : kps) { kp.x *= scale; kp.y *= scale; }
            }
            keypoints.insert(keypoints.end(), kps.begin(), kps.end());
        }
    }
};

// ------------------------------------------------------------------------------------------
// cv::BFMatcher(cv::NORM_HAMMING).match(query, train, matches) — one DMatch per query row in
// query order, nearest train row, strict '<' update so the LOWEST train index wins ties
// (call sites src/feature-tracker.cpp:161-163, :258, :303, :447, :494).
// ------------------------------------------------------------------------------------------
static inline int hamming256(const uint8_t* a, const uint8_t* b) {
    uint64_t x[4], y[4];
    std::memcpy(x, a, 32);
    std::memcpy(y, b, 32);
    return __builtin_popcountll(x[0] ^ y[0]) + __builtin_popcountll(x[1] ^ y[1]) +
           __builtin_popcountll(x[2] ^ y[2]) + __builtin_popcountll(x[3] ^ y[3]);
}
static void bf_match(const uint8_t* q, int nq, const uint8_t* t, int nt, int* trainIdx, int* dist) {
    for (int i = 0; i < nq; ++i) {
        int best = 0x7fffffff, bi = -1;
        for (int j = 0; j < nt; ++j) {
            int d = hamming256(q + (size_t)i * 32, t + (size_t)j * 32);
            if (d < best) { best = d; bi = j; }
        }
        trainIdx[i] = bi;
        dist[i] = best;
    }
}

// Global minimum distance + `distance < max(matchRatio*minDist, minMatchDist)` test of
// src/feature-tracker.cpp:164,211 / :259,293-296 / :304-307 (all fp32).  With an empty match list
// the reference dereferences end() (undefined); defined here as "nothing passes".
static float min_dist(const int* dist, int n) {
    float m = (float)dist[0];
    for (int i = 1; i < n; ++i) m = std::min(m, (float)dist[i]);
    return m;
}

// FeatureTracker::internalMatch, useRANSAC == false path (src/feature-tracker.cpp:161-164,208-218):
// keeps match i iff dist < max(ratio*minDist, minMatchDist) and |yL - yR| < maxVerticalPixelDist.
// Outputs the kept (queryIdx, trainIdx) pairs in query order; the caller gathers pixels and
// descriptor rows exactly as the reference's push_back loop does.
static int stereo_filter(const int* trainIdx, const int* dist, int nq, int nt, const KeyPoint* kpL,
                         const KeyPoint* kpR, float ratio, float minMatchDist, float maxDy, int* outL,
                         int* outR) {
    if (nq == 0 || nt == 0) return 0;
    const float thr = std::max(ratio * min_dist(dist, nq), minMatchDist);
    int m = 0;
    for (int i = 0; i < nq; ++i) {
        if ((float)dist[i] < thr && std::abs(kpL[i].y - kpR[trainIdx[i]].y) < maxDy) {
            outL[m] = i;
            outR[m] = trainIdx[i];
            ++m;
        }
    }
    return m;
}

// Distance-only selection of externalTrack (src/feature-tracker.cpp:258-259,293-296,303-307).
static int threshold_filter(const int* trainIdx, const int* dist, int nq, int nt, float ratio,
                            float minMatchDist, int* outQ, int* outT) {
    if (nq == 0 || nt == 0) return 0;
    const float thr = std::max(ratio * min_dist(dist, nq), minMatchDist);
    int m = 0;
    for (int i = 0; i < nq; ++i)
        if ((float)dist[i] < thr) { outQ[m] = i; outT[m] = trainIdx[i]; ++m; }
    return m;
}


// ------------------------------------------------------------------------------------------
// Next row f3: the feature pool of cfsd::FeatureTracker — the matching parts of externalTrack
// (src/feature-tracker.cpp:231-328, useRANSAC == false branch :293-296) against _histDescriptorsL/R, the
// _curFeatureMask / _matchedFeatureIDs bookkeeping (:311-321) and the pool rebuild of featurePoolUpdate
// (:330-412: age -= 1 for matched, += 2 for all, erase above maxFeatureAge, _hist* rebuilt by iterating _features,
// un-masked new features appended).  Triangulated positions are not kept (f4 computes them); the depth gate of :374 is
// an input (`accept`).
// _features is a std::unordered_map<size_t, Ptr<Feature>> (include/cfsd/feature-tracker.hpp:129), and the ORDER of
// _histFeatureIDs / _histDescriptors after a keyframe is that container's iteration order (:351-363) — it decides
// trainIdx and BFMatcher's lowest-index tie-break.  The restatement therefore uses the same container with the same
// key type, hash and sequence of operator[] / erase calls, so it follows libstdc++'s bucket order like the reference.
// ------------------------------------------------------------------------------------------
}  // namespace orc
#include <unordered_map>
namespace orc {

struct PoolFeature {
    uint8_t descL[32], descR[32];
    int age = 0;
};

struct FeaturePool {
    std::unordered_map<size_t, PoolFeature> features;          // _features
    std::vector<size_t> histIDs;                               // _histFeatureIDs
    std::vector<uint8_t> histL, histR;                         // _histDescriptorsL/R rows
    std::vector<size_t> matchedIDs;                            // _matchedFeatureIDs
    std::vector<uint8_t> curMask;                              // _curFeatureMask
    std::vector<uint8_t> curL, curR;                           // _curDescriptorsL/R of the last track call
    std::vector<int> circCur, circHist;                        // (queryIdx, trainIdx) of the circular matches
    size_t featureID = 0;                                      // _featureID
    int nLeft = 0, nRight = 0;                                 // verbose counters :297, :325
    float ratio = 2.f, minMatchDist = 30.f;
    int maxFeatureAge = 8;

    // processImage :102-103 then externalTrack(false) :231-328
    void track(const uint8_t* cl, const uint8_t* cr, int n) {
        curL.assign(cl, cl + (size_t)n * 32);
        curR.assign(cr, cr + (size_t)n * 32);
        curMask.assign(n, 1);
        circCur.clear(); circHist.clear();
        nLeft = nRight = 0;
        if (features.empty()) return;                          // :237-240 (matchedIDs keeps its old content)
        const int nh = (int)histIDs.size();
        std::vector<int> tl(n), dl(n), tr(n), dr(n), ql(n), hl(n), qr(n), hr(n);
        std::vector<int> mapCurHist(n, -1);
        if (n && nh) {
            bf_match(cl, n, histL.data(), nh, tl.data(), dl.data());
            nLeft = threshold_filter(tl.data(), dl.data(), n, nh, ratio, minMatchDist, ql.data(), hl.data());
            for (int i = 0; i < nLeft; ++i) mapCurHist[ql[i]] = hl[i];
        }
        matchedIDs.clear();                                    // :300
        if (n && nh) {
            bf_match(cr, n, histR.data(), nh, tr.data(), dr.data());
            nRight = threshold_filter(tr.data(), dr.data(), n, nh, ratio, minMatchDist, qr.data(), hr.data());
            for (int i = 0; i < nRight; ++i) {
                if (mapCurHist[qr[i]] == hr[i]) {              // :311-312
                    matchedIDs.push_back(histIDs[hr[i]]);
                    curMask[qr[i]] = 0;                        // :319
                    circCur.push_back(qr[i]);
                    circHist.push_back(hr[i]);
                }
            }
        }
    }

    // featurePoolUpdate :330-412.  accept[i] != 0 <=> !(depth > maxDepth) for cur feature i (:374); NULL = all.
    // Returns the number of inserted features; *erased = features dropped for age.
    int update(bool isKeyframe, const uint8_t* accept, int* erased) {
        int eraseCount = 0, insertCount = 0;
        if (features.empty() || isKeyframe) {
            for (size_t i = 0; i < matchedIDs.size(); ++i) features[matchedIDs[i]].age -= 1;   // :341-343
            histIDs.clear(); histL.clear(); histR.clear();
            auto f = features.begin();
            while (f != features.end()) {                      // :349-363
                f->second.age += 2;
                if (f->second.age > maxFeatureAge) { f = features.erase(f); ++eraseCount; }
                else {
                    histIDs.push_back(f->first);
                    histL.insert(histL.end(), f->second.descL, f->second.descL + 32);
                    histR.insert(histR.end(), f->second.descR, f->second.descR + 32);
                    ++f;
                }
            }
            for (size_t i = 0; i < curMask.size(); ++i) {      // :369-406
                if (!curMask[i] || (accept && !accept[i])) continue;
                PoolFeature pf;
                std::memcpy(pf.descL, &curL[i * 32], 32);
                std::memcpy(pf.descR, &curR[i * 32], 32);
                pf.age = 0;
                features[featureID] = pf;
                histIDs.push_back(featureID);
                histL.insert(histL.end(), pf.descL, pf.descL + 32);
                histR.insert(histR.end(), pf.descR, pf.descR + 32);
                ++featureID;
                ++insertCount;
            }
        }
        if (erased) *erased = eraseCount;
        return insertCount;
    }
};

}  // namespace orc

// ------------------------------------------------------------------------------------------
// C interface for ctypes (tests / bench cpu_baseline only).
// ------------------------------------------------------------------------------------------
using namespace orc;

static Img wrap(const uint8_t* p, int w, int h, int step) {
    Img im(w, h);
    for (int y = 0; y < h; ++y) std::memcpy(im.row(y), p + (size_t)y * step, w);
    return im;
}

// ------------------------------------------------------------------------------------------
// Next row f4: cv::triangulatePoints(P1, P2, pixelsL, pixelsR, points4D) and the depth gates of its two
// call sites (src/feature-tracker.cpp:365-379 featurePoolUpdate, :487-509 structFromMotion).
// OpenCV (modules/calib3d/src/triangulate.cpp, 4.x): per point the 4x4 system
//   A[2j+0][k] = x_j * P_j[2][k] - P_j[0][k],  A[2j+1][k] = y_j * P_j[2][k] - P_j[1][k]   (j = view 0/1)
// is handed to cv::SVD::compute and the point is the LAST row of Vt (smallest singular value).  A 4x4 double
// SVD is below OpenCV's LAPACK threshold, so it runs OpenCV's own one-sided Jacobi (modules/core/src/lapack.cpp,
// JacobiSVDImpl_<double>: eps = 10*DBL_EPSILON, at most max(m,30) sweeps, descending selection sort at the end)
// on At = A^T; restated here operation by operation (no FMA).  `hypot` inside lapack.cpp is OpenCV's own file-local
// template (scale by the larger magnitude, sqrt(1 + r*r)), not libm's — found by bisecting against cv2.SVDecomp.
// Pinned bit-exact against cv2.triangulatePoints by tests/test_oracle_vs_cv2.py.
// ------------------------------------------------------------------------------------------
static inline double cv_lapack_hypot(double a, double b) {
    a = std::fabs(a);
    b = std::fabs(b);
    if (a > b) { b /= a; return a * std::sqrt(1 + b * b); }
    if (b > 0) { a /= b; return b * std::sqrt(1 + a * a); }
    return 0;
}

static void jacobi_null_vector_4x4(double At[4][4], double out[4]) {
    const int m = 4, n = 4;
    const double eps = 2.220446049250313e-16 * 10;
    double W[4], Vt[4][4];
    for (int i = 0; i < n; ++i) {
        double sd = 0;
        for (int k = 0; k < m; ++k) { const double t = At[i][k]; sd += t * t; }
        W[i] = sd;
        for (int k = 0; k < n; ++k) Vt[i][k] = 0;
        Vt[i][i] = 1;
    }
    const int max_iter = 30;   // std::max(m, 30)
    for (int iter = 0; iter < max_iter; ++iter) {
        bool changed = false;
        for (int i = 0; i < n - 1; ++i)
            for (int j = i + 1; j < n; ++j) {
                double* Ai = At[i]; double* Aj = At[j];
                double a = W[i], p = 0, b = W[j];
                for (int k = 0; k < m; ++k) p += Ai[k] * Aj[k];
                if (std::fabs(p) <= eps * std::sqrt(a * b)) continue;
                p *= 2;
                const double beta = a - b, gamma = cv_lapack_hypot(p, beta);
                double c, s;
                if (beta < 0) {
                    const double delta = (gamma - beta) * 0.5;
                    s = std::sqrt(delta / gamma);
                    c = p / (gamma * s * 2);
                } else {
                    c = std::sqrt((gamma + beta) / (gamma * 2));
                    s = p / (gamma * c * 2);
                }
                a = b = 0;
                for (int k = 0; k < m; ++k) {
                    const double t0 = c * Ai[k] + s * Aj[k];
                    const double t1 = -s * Ai[k] + c * Aj[k];
                    Ai[k] = t0; Aj[k] = t1;
                    a += t0 * t0; b += t1 * t1;
                }
                W[i] = a; W[j] = b;
                changed = true;
                double* Vi = Vt[i]; double* Vj = Vt[j];
                for (int k = 0; k < n; ++k) {
                    const double t0 = c * Vi[k] + s * Vj[k];
                    const double t1 = -s * Vi[k] + c * Vj[k];
                    Vi[k] = t0; Vj[k] = t1;
                }
            }
        if (!changed) break;
    }
    for (int i = 0; i < n; ++i) {
        double sd = 0;
        for (int k = 0; k < m; ++k) { const double t = At[i][k]; sd += t * t; }
        W[i] = std::sqrt(sd);
    }
    int order[4] = {0, 1, 2, 3};   // the selection sort of the reference, tracked as a permutation of the Vt rows
    for (int i = 0; i < n - 1; ++i) {
        int j = i;
        for (int k = i + 1; k < n; ++k)
            if (W[j] < W[k]) j = k;
        if (i != j) { std::swap(W[i], W[j]); std::swap(order[i], order[j]); }
    }
    for (int k = 0; k < 4; ++k) out[k] = Vt[order[3]][k];
}

// points4d: 4 x n row-major (the cv::Mat CV_64F the reference indexes as points4D.at<double>(r, i)).
static void triangulate_points(const double* P1, const double* P2, const double* pixL, const double* pixR, int n,
                               double* points4d) {
    const double* P[2] = {P1, P2};
    for (int i = 0; i < n; ++i) {
        double At[4][4];
        for (int j = 0; j < 2; ++j) {
            const double* px = j ? pixR : pixL;
            const double x = px[2 * i], y = px[2 * i + 1];
            for (int k = 0; k < 4; ++k) {
                At[k][j * 2 + 0] = x * P[j][8 + k] - P[j][k];
                At[k][j * 2 + 1] = y * P[j][8 + k] - P[j][4 + k];
            }
        }
        double X[4];
        jacobi_null_vector_4x4(At, X);
        for (int k = 0; k < 4; ++k) points4d[(size_t)k * n + i] = X[k];
    }
}

// The consumers (src/feature-tracker.cpp:371-378 and :501-506): depth = Z/W, point_wrt_cam = (X/W, Y/W, depth);
// flag bit 0 = passes featurePoolUpdate's gate !(depth > maxDepth); bit 1 = passes structFromMotion's gate
// !(depth < 0.1 || depth > maxDepth).
static void depth_gate(const double* points4d, int n, double maxDepth, double* point_cam, uint8_t* flags) {
    for (int i = 0; i < n; ++i) {
        const double w = points4d[(size_t)3 * n + i];
        const double depth = points4d[(size_t)2 * n + i] / w;
        point_cam[3 * i + 0] = points4d[i] / w;
        point_cam[3 * i + 1] = points4d[(size_t)n + i] / w;
        point_cam[3 * i + 2] = depth;
        const bool pool_ok = !(depth > maxDepth);
        const bool sfm_ok = !(depth < 0.1 || depth > maxDepth);
        flags[i] = (uint8_t)((pool_ok ? 1 : 0) | (sfm_ok ? 2 : 0));
    }
}

extern "C" {

void* orc_create(int nf, float sf, int nl, int ini, int mn) { return new Extractor(nf, sf, nl, ini, mn); }
void orc_destroy(void* h) { delete (Extractor*)h; }
void orc_set_taps(void* h, const int* taps) { std::memcpy(((Extractor*)h)->taps, taps, 7 * sizeof(int)); }
void orc_set_tie_mode(void* h, int mode) { ((Extractor*)h)->tie_mode = mode < 0 ? -1 : 1; }
// ties[nlevels] of the last orc_extract: 0 = the reference's pointer comparison (src/ORBextractor.cc:684) could not
// have influenced that level, whatever the heap layout
void orc_get_ties(void* h, int* ties) {
    Extractor* e = (Extractor*)h;
    for (int i = 0; i < e->nlevels; ++i) ties[i] = i < (int)e->ties.size() ? e->ties[i] : 0;
}

// tables: scale[nl], inv_scale[nl], quota[nl], umax[16]
void orc_tables(void* h, float* scale, float* inv, int* quota, int* umax) {
    Extractor* e = (Extractor*)h;
    for (int i = 0; i < e->nlevels; ++i) {
        scale[i] = e->mvScaleFactor[i];
        inv[i] = e->mvInvScaleFactor[i];
        quota[i] = e->mnFeaturesPerLevel[i];
    }
    for (int i = 0; i < 16; ++i) umax[i] = e->umax[i];
}

int orc_extract(void* h, const uint8_t* img, int w, int hh, int step, KeyPoint* kps, uint8_t* desc, int cap) {
    Extractor* e = (Extractor*)h;
    std::vector<KeyPoint> k;
    std::vector<uint8_t> d;
    e->extract(wrap(img, w, hh, step), k, d);
    int n = (int)k.size();
    if (n > cap) return -n;
    if (n) {
        std::memcpy(kps, k.data(), (size_t)n * sizeof(KeyPoint));
        std::memcpy(desc, d.data(), (size_t)n * 32);
    }
    return n;
}

void orc_level_size(void* h, int level, int* w, int* hh) {
    Extractor* e = (Extractor*)h;
    *w = e->pyramid[level].w;
    *hh = e->pyramid[level].h;
}
// which: 0 = pyramid level, 1 = blurred level (empty if the level had no keypoints)
int orc_get_level(void* h, int level, int which, uint8_t* out) {
    Extractor* e = (Extractor*)h;
    const Img& im = which ? e->blurred[level] : e->pyramid[level];
    if (im.d.empty()) return 0;
    std::memcpy(out, im.d.data(), im.d.size());
    return 1;
}
// raw FAST corners of a level before the quadtree (x, y relative to the 16-px border, response)
int orc_get_raw(void* h, int level, float* xyr, int cap) {
    Extractor* e = (Extractor*)h;
    const auto& r = e->raw[level];
    int n = (int)r.size();
    for (int i = 0; i < n && i < cap; ++i) {
        xyr[3 * i] = r[i].x;
        xyr[3 * i + 1] = r[i].y;
        xyr[3 * i + 2] = r[i].response;
    }
    return n;
}

void orc_resize(const uint8_t* src, int sw, int sh, uint8_t* dst, int dw, int dh) {
    Img s = wrap(src, sw, sh, sw), d(dw, dh);
    resize_linear(s, d);
    std::memcpy(dst, d.d.data(), d.d.size());
}
// cv::remap(src, dst, map1, map2, INTER_LINEAR) with CV_16SC2 + CV_16UC1 fixed-point maps and the default
// BORDER_CONSTANT(0) — src/feature-tracker.cpp:75-76, maps from include/cfsd/camera-model.hpp:76-77.
// map1 = (sx, sy) per pixel, map2 = fx | fy << 5 in 1/32 px.  OpenCV's bilinear weight table is
// saturate_cast<short>((1-fy/32)(1-fx/32) * 32768) ...; all four products are exact multiples of 32 and sum to 32768, so
// out = (sum_w I + 16384) >> 15 == ((32-fy)(32-fx) I00 + (32-fy) fx I01 + fy (32-fx) I10 + fy fx I11 + 512) >> 10.
// Taps outside the source image read 0.
void orc_remap(const uint8_t* src, int sw, int sh, int sstep, const int16_t* map1, const uint16_t* map2, int dw, int dh,
               uint8_t* dst, int dstep) {
    for (int y = 0; y < dh; ++y)
        for (int x = 0; x < dw; ++x) {
            const int sx = map1[2 * ((size_t)y * dw + x)], sy = map1[2 * ((size_t)y * dw + x) + 1];
            const int m = map2[(size_t)y * dw + x];
            const int fx = m & 31, fy = (m >> 5) & 31;
            const auto tap = [&](int yy, int xx) -> int {
                return (xx >= 0 && xx < sw && yy >= 0 && yy < sh) ? src[(size_t)yy * sstep + xx] : 0;
            };
            const int s = (32 - fy) * (32 - fx) * tap(sy, sx) + (32 - fy) * fx * tap(sy, sx + 1) +
                          fy * (32 - fx) * tap(sy + 1, sx) + fy * fx * tap(sy + 1, sx + 1);
            dst[(size_t)y * dstep + x] = (uint8_t)((s + 512) >> 10);
        }
}
// cfsd live ingest, src/cfsd-state-estimation.cpp:117-127: the shared-memory frame (CV_8UC4, both eyes side by side) is
// cv::resize'd (INTER_LINEAR) to imgSize, cvtColor(BGR2GRAY) — channel 3 is ignored — and split into left / right halves.
// cv::resize on 4 channels is the 1-channel arithmetic per channel (checked against cv2 4.13); BGR2GRAY on 8-bit is
// (B*3735 + G*19235 + R*9798 + 16384) >> 15 in OpenCV 4.x (3.x used the 14-bit set 1868/9617/4899 — like the blur taps,
// the version on this box is the one that can be pinned).  dw = width of BOTH eyes together (even).
void orc_cfsd_ingest(const uint8_t* frame, int sw, int sh, int sstep, int dw, int dh, uint8_t* gray_l, uint8_t* gray_r,
                     int ostep) {
    Img ch[3];
    for (int c = 0; c < 3; ++c) {
        Img s(sw, sh);
        for (int y = 0; y < sh; ++y)
            for (int x = 0; x < sw; ++x) s.row(y)[x] = frame[(size_t)y * sstep + 4 * x + c];
        ch[c] = Img(dw, dh);
        resize_linear(s, ch[c]);
    }
    const int half = dw / 2;
    for (int y = 0; y < dh; ++y)
        for (int x = 0; x < dw; ++x) {
            const int g = (ch[0].row(y)[x] * 3735 + ch[1].row(y)[x] * 19235 + ch[2].row(y)[x] * 9798 + 16384) >> 15;
            (x < half ? gray_l : gray_r)[(size_t)y * ostep + (x < half ? x : x - half)] = (uint8_t)g;
        }
}
void orc_blur(const uint8_t* src, int w, int hh, uint8_t* dst, const int* taps) {
    Img s = wrap(src, w, hh, w), d;
    gaussian_blur7(s, d, taps ? taps : kTaps413);
    std::memcpy(dst, d.d.data(), d.d.size());
}
int orc_fast(const uint8_t* img, int w, int hh, int x0, int y0, int x1, int y1, int thr, float* xyr, int cap) {
    Img s = wrap(img, w, hh, w);
    std::vector<KeyPoint> k;
    fast9_16_roi(s, x0, y0, x1, y1, thr, k);
    int n = (int)k.size();
    for (int i = 0; i < n && i < cap; ++i) {
        xyr[3 * i] = k[i].x;
        xyr[3 * i + 1] = k[i].y;
        xyr[3 * i + 2] = k[i].response;
    }
    return n;
}
float orc_fast_atan2(float y, float x) { return fast_atan2_deg(y, x); }
float orc_cosf(float a) { return ref_cosf(a); }
float orc_sinf(float a) { return ref_sinf(a); }
// exhaustive check of the sincosf restatement against the libm of the running host over the
// float bit patterns [lo, hi]; returns the number of mismatching results.
long orc_sincos_check(uint32_t lo, uint32_t hi) {
    long bad = 0;
    for (uint64_t u = lo; u <= hi; ++u) {
        uint32_t b = (uint32_t)u;
        float a, c0, c1, s0, s1;
        std::memcpy(&a, &b, 4);
        c0 = cosf(a); c1 = ref_cosf(a);
        s0 = sinf(a); s1 = ref_sinf(a);
        if (std::memcmp(&c0, &c1, 4) || std::memcmp(&s0, &s1, 4)) ++bad;
    }
    return bad;
}
// quadtree alone: in = n x (x, y, response) relative coords; out = kept points in list order
int orc_octtree(const float* xyr, int n, int minX, int maxX, int minY, int maxY, int N, float* out, int cap) {
    std::vector<KeyPoint> in(n);
    for (int i = 0; i < n; ++i) in[i] = KeyPoint{xyr[3 * i], xyr[3 * i + 1], 7.f, -1.f, xyr[3 * i + 2], 0, -1};
    std::vector<KeyPoint> r = distribute_octtree(in, minX, maxX, minY, maxY, N);
    int m = (int)r.size();
    for (int i = 0; i < m && i < cap; ++i) {
        out[3 * i] = r[i].x;
        out[3 * i + 1] = r[i].y;
        out[3 * i + 2] = r[i].response;
    }
    return m;
}

void orc_bf_match(const uint8_t* q, int nq, const uint8_t* t, int nt, int* trainIdx, int* dist) {
    bf_match(q, nq, t, nt, trainIdx, dist);
}
int orc_stereo_filter(const int* trainIdx, const int* dist, int nq, int nt, const KeyPoint* kpL,
                      const KeyPoint* kpR, float ratio, float minMatchDist, float maxDy, int* outL, int* outR) {
    return stereo_filter(trainIdx, dist, nq, nt, kpL, kpR, ratio, minMatchDist, maxDy, outL, outR);
}
void orc_triangulate(const double* P1, const double* P2, const double* pixL, const double* pixR, int n,
                     double* points4d) {
    triangulate_points(P1, P2, pixL, pixR, n, points4d);
}
void orc_depth_gate(const double* points4d, int n, double maxDepth, double* point_cam, uint8_t* flags) {
    depth_gate(points4d, n, maxDepth, point_cam, flags);
}
int orc_threshold_filter(const int* trainIdx, const int* dist, int nq, int nt, float ratio,
                         float minMatchDist, int* outQ, int* outT) {
    return threshold_filter(trainIdx, dist, nq, nt, ratio, minMatchDist, outQ, outT);
}

// One stereo pair the way FeatureTracker::internalMatch runs it (src/feature-tracker.cpp:153-164,
// 208-218): left/right extraction in two threads (threads==2) or sequentially (threads==1),
// BF match L->R, min-dist/row filter.  hL/hR are two extractor instances.  Outputs may be NULL
// when only timing is wanted.  Returns the number of kept matches; *nL, *nR = keypoint counts.
int orc_stereo_pair(void* hL, void* hR, const uint8_t* imgL, const uint8_t* imgR, int w, int hh, int step,
                    float ratio, float minMatchDist, float maxDy, int threads, KeyPoint* kpLo, uint8_t* dLo,
                    KeyPoint* kpRo, uint8_t* dRo, int cap, int* nL, int* nR, int* trainIdx, int* dist,
                    int* outL, int* outR) {
    std::vector<KeyPoint> kL, kR;
    std::vector<uint8_t> dL, dR;
    Img iL = wrap(imgL, w, hh, step), iR = wrap(imgR, w, hh, step);
    if (threads >= 2) {
        std::thread tl([&] { ((Extractor*)hL)->extract(iL, kL, dL); });
        std::thread tr([&] { ((Extractor*)hR)->extract(iR, kR, dR); });
        tl.join();
        tr.join();
    } else {
        ((Extractor*)hL)->extract(iL, kL, dL);
        ((Extractor*)hR)->extract(iR, kR, dR);
    }
    const int a = (int)kL.size(), b = (int)kR.size();
    *nL = a;
    *nR = b;
    std::vector<int> ti(a), di(a), oL(a), oR(a);
    int m = 0;
    if (a && b) {
        bf_match(dL.data(), a, dR.data(), b, ti.data(), di.data());
        m = stereo_filter(ti.data(), di.data(), a, b, kL.data(), kR.data(), ratio, minMatchDist, maxDy,
                          oL.data(), oR.data());
    }
    if (kpLo && a <= cap && b <= cap) {
        std::memcpy(kpLo, kL.data(), (size_t)a * sizeof(KeyPoint));
        std::memcpy(dLo, dL.data(), (size_t)a * 32);
        std::memcpy(kpRo, kR.data(), (size_t)b * sizeof(KeyPoint));
        std::memcpy(dRo, dR.data(), (size_t)b * 32);
        std::memcpy(trainIdx, ti.data(), (size_t)a * 4);
        std::memcpy(dist, di.data(), (size_t)a * 4);
        std::memcpy(outL, oL.data(), (size_t)m * 4);
        std::memcpy(outR, oR.data(), (size_t)m * 4);
    }
    return m;
}

// CPU baseline driver: `npairs` stereo pairs (contiguous L then R images, each w*hh bytes) spread
// over `workers` host threads, each worker running its pairs sequentially with its own two extractor
// instances (per_pair_threads = 2 reproduces the reference's L/R std::threads inside a pair,
// src/feature-tracker.cpp:153-156).  temporal != 0 adds the matching parts of externalTrack
// (:258-259, 293-296, 303-321) between a worker's consecutive pairs.  Returns kept stereo matches +
// circular matches summed (so the work cannot be optimised away).
// stage seconds of the last orc_run_pairs, summed over workers: 6 extractor stages (both eyes), stereo match + filter,
// temporal match + circular check
static double g_stage_s[8];
void orc_last_stage_seconds(double* out) { std::memcpy(out, g_stage_s, sizeof g_stage_s); }

long orc_run_pairs(int nf, float sf, int nl, int ini, int mn, const uint8_t* imgsL, const uint8_t* imgsR,
                   int npairs, int w, int hh, float ratio, float minMatchDist, float maxDy, int workers,
                   int per_pair_threads, int temporal) {
    std::vector<long> tot(workers, 0);
    std::vector<std::vector<double>> st(workers, std::vector<double>(8, 0.0));
    std::vector<std::thread> th;
    for (int t = 0; t < workers; ++t) {
        th.emplace_back([&, t] {
            Extractor eL(nf, sf, nl, ini, mn), eR(nf, sf, nl, ini, mn);
            std::vector<uint8_t> histL, histR;
            for (int p = t; p < npairs; p += workers) {
                std::vector<KeyPoint> kL, kR;
                std::vector<uint8_t> dL, dR;
                Img iL = wrap(imgsL + (size_t)p * w * hh, w, hh, w), iR = wrap(imgsR + (size_t)p * w * hh, w, hh, w);
                if (per_pair_threads >= 2) {
                    std::thread tl([&] { eL.extract(iL, kL, dL); });
                    std::thread tr([&] { eR.extract(iR, kR, dR); });
                    tl.join();
                    tr.join();
                } else {
                    eL.extract(iL, kL, dL);
                    eR.extract(iR, kR, dR);
                }
                const int a = (int)kL.size(), b = (int)kR.size();
                std::vector<int> ti(a), di(a), oL(a), oR(a);
                int m = 0;
                double tm = Extractor::now_s();
                if (a && b) {
                    bf_match(dL.data(), a, dR.data(), b, ti.data(), di.data());
                    m = stereo_filter(ti.data(), di.data(), a, b, kL.data(), kR.data(), ratio, minMatchDist, maxDy,
                                      oL.data(), oR.data());
                }
                st[t][6] += Extractor::now_s() - tm;
                tot[t] += m;
                tm = Extractor::now_s();
                if (temporal) {
                    std::vector<uint8_t> curL((size_t)m * 32), curR((size_t)m * 32);
                    for (int i = 0; i < m; ++i) {
                        std::memcpy(&curL[(size_t)i * 32], &dL[(size_t)oL[i] * 32], 32);
                        std::memcpy(&curR[(size_t)i * 32], &dR[(size_t)oR[i] * 32], 32);
                    }
                    const int nh = (int)histL.size() / 32;
                    if (m && nh) {
                        std::vector<int> tl(m), dl(m), tr(m), dr(m), ql(m), hl(m), qr(m), hr(m);
                        bf_match(curL.data(), m, histL.data(), nh, tl.data(), dl.data());
                        bf_match(curR.data(), m, histR.data(), nh, tr.data(), dr.data());
                        const int cl = threshold_filter(tl.data(), dl.data(), m, nh, ratio, minMatchDist, ql.data(), hl.data());
                        const int cr = threshold_filter(tr.data(), dr.data(), m, nh, ratio, minMatchDist, qr.data(), hr.data());
                        std::vector<int> map(m, -1);
                        for (int i = 0; i < cl; ++i) map[ql[i]] = hl[i];
                        for (int i = 0; i < cr; ++i) tot[t] += (map[qr[i]] == hr[i]);
                    }
                    histL.swap(curL);
                    histR.swap(curR);
                }
                st[t][7] += Extractor::now_s() - tm;
            }
            for (int k = 0; k < 6; ++k) st[t][k] = eL.stage_s[k] + eR.stage_s[k];
        });
    }
    for (auto& x : th) x.join();
    long sum = 0;
    for (long v : tot) sum += v;
    for (int k = 0; k < 8; ++k) {
        g_stage_s[k] = 0;
        for (int t = 0; t < workers; ++t) g_stage_s[k] += st[t][k];
    }
    return sum;
}

// ---- f3: feature pool -------------------------------------------------------------------------
void* orc_pool_create(float ratio, float minMatchDist, int maxFeatureAge) {
    FeaturePool* p = new FeaturePool();
    p->ratio = ratio; p->minMatchDist = minMatchDist; p->maxFeatureAge = maxFeatureAge;
    return p;
}
void orc_pool_destroy(void* h) { delete (FeaturePool*)h; }
// externalTrack on n current (kept stereo) descriptor pairs.  Outputs: circ_cur / circ_hist [n], mask [n],
// matched_ids [n] (as uint64), counts[4] = {n_circ, n_left, n_right, n_hist}.
void orc_pool_track(void* h, const uint8_t* curL, const uint8_t* curR, int n, int* circ_cur, int* circ_hist,
                    uint8_t* mask, uint64_t* matched_ids, int* counts) {
    FeaturePool* p = (FeaturePool*)h;
    p->track(curL, curR, n);
    const int nc = (int)p->circCur.size();
    for (int i = 0; i < nc; ++i) { circ_cur[i] = p->circCur[i]; circ_hist[i] = p->circHist[i]; matched_ids[i] = p->matchedIDs[i]; }
    for (int i = 0; i < n; ++i) mask[i] = p->curMask[i];
    counts[0] = nc; counts[1] = p->nLeft; counts[2] = p->nRight; counts[3] = (int)p->histIDs.size();
}
int orc_pool_update(void* h, int isKeyframe, const uint8_t* accept, int* erased) {
    return ((FeaturePool*)h)->update(isKeyframe != 0, accept, erased);
}
int orc_pool_size(void* h) { return (int)((FeaturePool*)h)->histIDs.size(); }
// _histFeatureIDs, ages and descriptor rows in pool order
void orc_pool_get(void* h, uint64_t* ids, int* ages, uint8_t* histL, uint8_t* histR) {
    FeaturePool* p = (FeaturePool*)h;
    const size_t n = p->histIDs.size();
    for (size_t i = 0; i < n; ++i) {
        if (ids) ids[i] = p->histIDs[i];
        if (ages) ages[i] = p->features[p->histIDs[i]].age;
    }
    if (histL && n) std::memcpy(histL, p->histL.data(), n * 32);
    if (histR && n) std::memcpy(histR, p->histR.data(), n * 32);
}

}  // extern "C"
