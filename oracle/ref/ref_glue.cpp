// ref_glue.cpp — builds the REFERENCE'S OWN extractor into a checker (test infrastructure, NOT product code).
//
// oracle/ref/Makefile compiles /root/reference/src/ORBextractor.cc UNMODIFIED (from where it lies; it is never
// copied into this repository) together with this file into oracle/_ref/libref_orb*.so.  The reference file needs
// OpenCV only for cv::Mat / KeyPoint / Point / Size / Rect plumbing and six free functions
// (src/ORBextractor.cc:57-60); the plumbing is tests/mock_opencv/, and the functions are defined below by
// forwarding to the restatements in oracle/orb_oracle.cpp that tests/test_oracle_vs_cv2.py pins bit-exact to the
// real OpenCV 4.13 kernels of the cv2 wheel:
//     cv::FAST          -> orc_fast          (src/ORBextractor.cc:809-815)
//     cv::resize        -> orc_resize        (:1120)
//     cv::GaussianBlur  -> orc_blur          (:1086)
//     cv::fastAtan2     -> orc_fast_atan2    (:103)
//     cv::copyMakeBorder-> reflect-101 copy written here (:1122, :1127; never read afterwards)
//     cvRound/Floor/Ceil-> lrint / floor / ceil in the mock header
// cosf/sinf (:113) are the libm of the running host, as in the reference.
// So everything the reference's authors wrote — ctor tables, ComputePyramid, the per-cell FAST driver with the
// threshold fallback, DistributeOctTree / DivideNode, IC_Angle, computeOrbDescriptor, operator() — runs as THEIR
// code, and orb_oracle.cpp (and through it the CUDA path) is checked against it.
//
// Heap addresses.  DistributeOctTree sorts pair<int, ExtractorNode*> (src/ORBextractor.cc:684): ties in point
// count are broken by the ADDRESS of the std::list node, i.e. by the allocator.  To make that reproducible this
// library replaces operator new/delete *inside itself only* (the linker version script keeps every symbol but
// ref_* local) with a per-thread arena that is active during ref_extract:
//     REF_ARENA=+1  addresses grow with allocation order   -> "node created later is expanded first"  (the
//                   documented rule of SURVEY.md App. C, the oracle's default and the CUDA path's)
//     REF_ARENA=-1  addresses shrink with allocation order -> "node created earlier is expanded first"
//     REF_ARENA=0   the process's ordinary malloc (whatever order glibc's bins produce)
// Any of the three is a legitimate environment for the reference; the first two are exact models the oracle can
// follow (orc_set_tie_mode), the third is used by tests to show that results differ from them only on levels
// where the final partial pass really had equal-sized nodes.
#include <sys/mman.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <thread>
#include <vector>

#include <opencv2/core.hpp>

#include "ORBextractor.h"   // the reference's include/ORBextractor.h (via -I/root/reference/include)

#ifndef REF_ARENA
#define REF_ARENA 1
#endif

extern "C" {
// exported by oracle/liborb_oracle.so
void orc_resize(const uint8_t* src, int sw, int sh, uint8_t* dst, int dw, int dh);
void orc_blur(const uint8_t* src, int w, int hh, uint8_t* dst, const int* taps);
int orc_fast(const uint8_t* img, int w, int hh, int x0, int y0, int x1, int y1, int thr, float* xyr, int cap);
float orc_fast_atan2(float y, float x);
void orc_bf_match(const uint8_t* q, int nq, const uint8_t* t, int nt, int* trainIdx, int* dist);
int orc_stereo_filter(const int* trainIdx, const int* dist, int nq, int nt, const void* kpL, const void* kpR, float ratio,
                      float minMatchDist, float maxDy, int* outL, int* outR);
int orc_threshold_filter(const int* trainIdx, const int* dist, int nq, int nt, float ratio, float minMatchDist, int* outQ,
                         int* outT);
}

// ---------------------------------------------------------------------------------------------------------------
// arena operator new / delete (local to this shared object)
// ---------------------------------------------------------------------------------------------------------------
#if REF_ARENA != 0
namespace {
const size_t kArenaBytes = (size_t)48 << 30;   // virtual reservation; pages are touched on demand
struct Arena {
    char* base = nullptr;
    size_t used = 0, peak = 0;
    int depth = 0;   // > 0 while inside ref_extract
};
thread_local Arena g_arena;

inline bool in_arena(const void* p) {
    const Arena& a = g_arena;
    return a.base && (const char*)p >= a.base && (const char*)p < a.base + kArenaBytes;
}
void* arena_alloc(size_t n) {
    Arena& a = g_arena;
    if (a.depth == 0) return std::malloc(n ? n : 1);
    if (!a.base) {
        void* m = mmap(nullptr, kArenaBytes, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
        if (m == MAP_FAILED) { std::fprintf(stderr, "ref arena: mmap failed\n"); std::abort(); }
        a.base = (char*)m;
    }
    n = (n + 15) & ~(size_t)15;
    if (n == 0) n = 16;
    if (a.used + n > kArenaBytes) { std::fprintf(stderr, "ref arena: exhausted\n"); std::abort(); }
    a.used += n;
    if (a.used > a.peak) a.peak = a.used;
#if REF_ARENA > 0
    return a.base + (a.used - n);
#else
    return a.base + (kArenaBytes - a.used);
#endif
}
void arena_free(void* p) {
    if (!p || in_arena(p)) return;   // arena blocks are dropped wholesale when the call ends
    std::free(p);
}
void arena_enter() {
    Arena& a = g_arena;
    if (a.depth++ == 0) a.used = 0;
}
void arena_leave() {
    Arena& a = g_arena;
    if (--a.depth == 0 && a.base && a.peak > ((size_t)64 << 20)) {   // give the pages back after a large call
#if REF_ARENA > 0
        madvise(a.base, a.peak, MADV_DONTNEED);
#else
        madvise(a.base + (kArenaBytes - a.peak), a.peak, MADV_DONTNEED);
#endif
        a.peak = 0;
    }
}
}  // namespace

void* operator new(size_t n) { return arena_alloc(n); }
void* operator new[](size_t n) { return arena_alloc(n); }
void* operator new(size_t n, const std::nothrow_t&) noexcept { return arena_alloc(n); }
void* operator new[](size_t n, const std::nothrow_t&) noexcept { return arena_alloc(n); }
void operator delete(void* p) noexcept { arena_free(p); }
void operator delete[](void* p) noexcept { arena_free(p); }
void operator delete(void* p, size_t) noexcept { arena_free(p); }
void operator delete[](void* p, size_t) noexcept { arena_free(p); }
#else
namespace {
void arena_enter() {}
void arena_leave() {}
}  // namespace
#endif

// ---------------------------------------------------------------------------------------------------------------
// the OpenCV free functions src/ORBextractor.cc calls
// ---------------------------------------------------------------------------------------------------------------
namespace {
thread_local int g_taps[7] = {18, 34, 48, 56, 48, 34, 18};   // OpenCV 4.x bit-exact Gaussian kernel

// contiguous copy of a (possibly ROI) 8-bit Mat; malloc-backed so it never lands in the arena
struct Tight {
    uint8_t* p; int w, h;
    explicit Tight(const cv::Mat& m) : p((uint8_t*)std::malloc((size_t)m.rows * m.cols + 1)), w(m.cols), h(m.rows) {
        for (int y = 0; y < h; ++y) std::memcpy(p + (size_t)y * w, m.ptr(y), (size_t)w);
    }
    Tight(int w_, int h_) : p((uint8_t*)std::malloc((size_t)w_ * h_ + 1)), w(w_), h(h_) {}
    ~Tight() { std::free(p); }
    void store(cv::Mat& m) const { for (int y = 0; y < h; ++y) std::memcpy(m.ptr(y), p + (size_t)y * w, (size_t)w); }
};
inline int reflect101(int i, int n) {
    if (n == 1) return 0;
    while (i < 0 || i >= n) i = i < 0 ? -i : 2 * (n - 1) - i;
    return i;
}
}  // namespace

namespace cv {

void FAST(InputArray image, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression) {
    assert(nonmaxSuppression && image.type() == CV_8UC1);
    (void)nonmaxSuppression;
    keypoints.clear();
    if (image.empty()) return;
    Tight t(image);
    const int cap = (t.w * t.h) / 4 + 16;
    float* xyr = (float*)std::malloc((size_t)cap * 3 * sizeof(float));
    const int n = orc_fast(t.p, t.w, t.h, 0, 0, t.w, t.h, threshold, xyr, cap);
    assert(n <= cap);
    for (int i = 0; i < n; ++i) keypoints.push_back(KeyPoint(xyr[3 * i], xyr[3 * i + 1], 7.f, -1.f, xyr[3 * i + 2], 0, -1));
    std::free(xyr);
}

void resize(InputArray src, OutputArray dst, Size dsize, double, double, int interpolation) {
    assert(interpolation == INTER_LINEAR && src.type() == CV_8UC1);
    (void)interpolation;
    dst.create(dsize.height, dsize.width, src.type());   // no-op for the pre-sized ROI of ComputePyramid
    Tight s(src), d(dsize.width, dsize.height);
    orc_resize(s.p, s.w, s.h, d.p, d.w, d.h);
    d.store(dst);
}

void copyMakeBorder(InputArray src, OutputArray dst, int top, int bottom, int left, int right, int borderType) {
    assert((borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101);
    (void)borderType;
    Tight s(src);   // copy first: the source may be a view into dst (src/ORBextractor.cc:1122)
    dst.create(s.h + top + bottom, s.w + left + right, src.type());
    for (int y = 0; y < dst.rows; ++y) {
        const uint8_t* row = s.p + (size_t)reflect101(y - top, s.h) * s.w;
        uint8_t* out = dst.ptr(y);
        for (int x = 0; x < dst.cols; ++x) out[x] = row[reflect101(x - left, s.w)];
    }
}

void GaussianBlur(InputArray src, OutputArray dst, Size ksize, double sigmaX, double sigmaY, int borderType) {
    assert(ksize.width == 7 && ksize.height == 7 && sigmaX == 2 && sigmaY == 2 && borderType == BORDER_REFLECT_101);
    (void)ksize; (void)sigmaX; (void)sigmaY; (void)borderType;
    Tight s(src), d(src.cols, src.rows);
    orc_blur(s.p, s.w, s.h, d.p, g_taps);
    dst.create(src.rows, src.cols, src.type());
    d.store(dst);
}

float fastAtan2(float y, float x) { return orc_fast_atan2(y, x); }

// Only ComputeKeyPointsOld (dead code, call commented out at src/ORBextractor.cc:1057) uses this.
void KeyPointsFilter::retainBest(std::vector<KeyPoint>& keypoints, int npoints) {
    if (npoints < 0 || (int)keypoints.size() <= npoints) return;
    if (npoints == 0) { keypoints.clear(); return; }
    std::vector<float> r;
    for (const KeyPoint& k : keypoints) r.push_back(k.response);
    std::vector<float> sorted(r);
    for (size_t i = 1; i < sorted.size(); ++i)
        for (size_t j = i; j > 0 && sorted[j] > sorted[j - 1]; --j) std::swap(sorted[j], sorted[j - 1]);
    const float cut = sorted[npoints - 1];
    std::vector<KeyPoint> out;
    for (const KeyPoint& k : keypoints) if (k.response >= cut) out.push_back(k);
    keypoints.swap(out);
}

}  // namespace cv

// ---------------------------------------------------------------------------------------------------------------
// C entry points (ctypes: oracle/oracle.py RefExtractor)
// ---------------------------------------------------------------------------------------------------------------
namespace {
struct Probe : ORB_SLAM2::ORBextractor {   // reaches the protected tables without touching the reference source
    using ORB_SLAM2::ORBextractor::ORBextractor;
    const std::vector<int>& quota() const { return mnFeaturesPerLevel; }
    const std::vector<int>& umax_() const { return umax; }
};
}  // namespace

extern "C" {

int ref_arena_mode() { return REF_ARENA; }

void* ref_create(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST) {
    return new Probe(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST);
}
void ref_destroy(void* h) { delete (Probe*)h; }

// Gaussian taps used by cv::GaussianBlur on the calling thread (default: OpenCV 4.x; {18,34,49,55,49,34,18} = 3.2)
void ref_set_taps(const int* taps) { std::memcpy(g_taps, taps, sizeof(g_taps)); }

// scale[nl], inv_scale[nl], sigma2[nl], inv_sigma2[nl], quota[nl], umax[16] — src/ORBextractor.cc:415-469
void ref_tables(void* h, float* scale, float* inv, float* sigma2, float* inv_sigma2, int* quota, int* umax) {
    Probe* e = (Probe*)h;
    const int nl = e->GetLevels();
    std::vector<float> a = e->GetScaleFactors(), b = e->GetInverseScaleFactors(), c = e->GetScaleSigmaSquares(),
                       d = e->GetInverseScaleSigmaSquares();
    for (int i = 0; i < nl; ++i) {
        scale[i] = a[i]; inv[i] = b[i]; sigma2[i] = c[i]; inv_sigma2[i] = d[i];
        quota[i] = e->quota()[i];
    }
    for (int i = 0; i < 16; ++i) umax[i] = e->umax_()[i];
}

// ORBextractor::operator()(image, cv::Mat(), keypoints, descriptors).  kps = cv::KeyPoint[cap] (28 B each),
// desc = u8[cap][32].  Returns the keypoint count, or -count when cap is too small.
int ref_extract(void* h, const uint8_t* img, int w, int hh, int step, void* kps, uint8_t* desc, int cap) {
    Probe* e = (Probe*)h;
    int n;
    arena_enter();
    {
        cv::Mat image(hh, w, CV_8UC1, (void*)img, (size_t)step), mask, descriptors;
        std::vector<cv::KeyPoint> keypoints;
        (*e)(image, mask, keypoints, descriptors);
        n = (int)keypoints.size();
        if (n > cap) n = -n;
        else if (n) {
            std::memcpy(kps, keypoints.data(), (size_t)n * sizeof(cv::KeyPoint));
            for (int i = 0; i < n; ++i) std::memcpy(desc + (size_t)i * 32, descriptors.ptr(i), 32);
        }
    }
    arena_leave();
    return n;
}

// mvImagePyramid[level] after the last ref_extract (public member, include/ORBextractor.h:85)
void ref_level_size(void* h, int level, int* w, int* hh) {
    Probe* e = (Probe*)h;
    *w = e->mvImagePyramid[level].cols;
    *hh = e->mvImagePyramid[level].rows;
}
void ref_get_level(void* h, int level, uint8_t* out) {
    Probe* e = (Probe*)h;
    const cv::Mat& m = e->mvImagePyramid[level];
    for (int y = 0; y < m.rows; ++y) std::memcpy(out + (size_t)y * m.cols, m.ptr(y), (size_t)m.cols);
}
// mvImagePyramid[level] with its 19-px apron (the `temp` of ComputePyramid): (w+38) x (h+38) bytes
void ref_get_level_bordered(void* h, int level, uint8_t* out) {
    Probe* e = (Probe*)h;
    const cv::Mat& m = e->mvImagePyramid[level];
    const int W = m.cols + 38, H = m.rows + 38;
    const uint8_t* origin = m.data - 19 * m.step - 19;
    for (int y = 0; y < H; ++y) std::memcpy(out + (size_t)y * W, origin + (size_t)y * m.step, (size_t)W);
}

// CPU reference arm of bench.py (--impl reference): `npairs` stereo pairs spread over `workers` host threads, each
// worker owning two instances of the reference's ORBextractor (left / right, src/feature-tracker.cpp:26-27) and running
// its pairs one after the other: operator() x2 (on two std::threads when per_pair_threads >= 2, :153-156), BFMatcher +
// min-distance / row filter (:161-164, 208-218) and, with temporal != 0, the matching parts of externalTrack against the
// worker's previous pair (:258-259, 293-296, 303-321).  Extraction is the reference's own code; the matcher calls are
// OpenCV's in the reference and the cv2-pinned restatements here.  Returns kept + circular matches (keeps the work live).
long ref_run_pairs(int nf, float sf, int nl, int ini, int mn, const uint8_t* imgsL, const uint8_t* imgsR, int npairs, int w,
                   int hh, float ratio, float minMatchDist, float maxDy, int workers, int per_pair_threads, int temporal) {
    std::vector<long> tot(workers, 0);
    std::vector<std::thread> th;
    for (int t = 0; t < workers; ++t) {
        th.emplace_back([&, t] {
            Probe* eL = new Probe(nf, sf, nl, ini, mn);
            Probe* eR = new Probe(nf, sf, nl, ini, mn);
            std::vector<uint8_t> histL, histR;
            const int cap = nf + 4 * nl + 64;
            std::vector<cv::KeyPoint> kL(cap), kR(cap);
            std::vector<uint8_t> dL((size_t)cap * 32), dR((size_t)cap * 32);
            for (int p = t; p < npairs; p += workers) {
                int a = 0, b = 0;
                const uint8_t* iL = imgsL + (size_t)p * w * hh;
                const uint8_t* iR = imgsR + (size_t)p * w * hh;
                if (per_pair_threads >= 2) {
                    std::thread tl([&] { a = ref_extract(eL, iL, w, hh, w, kL.data(), dL.data(), cap); });
                    std::thread tr([&] { b = ref_extract(eR, iR, w, hh, w, kR.data(), dR.data(), cap); });
                    tl.join();
                    tr.join();
                } else {
                    a = ref_extract(eL, iL, w, hh, w, kL.data(), dL.data(), cap);
                    b = ref_extract(eR, iR, w, hh, w, kR.data(), dR.data(), cap);
                }
                if (a < 0 || b < 0) continue;
                std::vector<int> ti(a), di(a), oL(a), oR(a);
                int m = 0;
                if (a && b) {
                    orc_bf_match(dL.data(), a, dR.data(), b, ti.data(), di.data());
                    m = orc_stereo_filter(ti.data(), di.data(), a, b, kL.data(), kR.data(), ratio, minMatchDist, maxDy, oL.data(), oR.data());
                }
                tot[t] += m;
                if (temporal) {
                    std::vector<uint8_t> curL((size_t)m * 32), curR((size_t)m * 32);
                    for (int i = 0; i < m; ++i) {
                        std::memcpy(&curL[(size_t)i * 32], &dL[(size_t)oL[i] * 32], 32);
                        std::memcpy(&curR[(size_t)i * 32], &dR[(size_t)oR[i] * 32], 32);
                    }
                    const int nh = (int)histL.size() / 32;
                    if (m && nh) {
                        std::vector<int> tl(m), dl(m), tr(m), dr(m), ql(m), hl(m), qr(m), hr(m);
                        orc_bf_match(curL.data(), m, histL.data(), nh, tl.data(), dl.data());
                        orc_bf_match(curR.data(), m, histR.data(), nh, tr.data(), dr.data());
                        const int cl = orc_threshold_filter(tl.data(), dl.data(), m, nh, ratio, minMatchDist, ql.data(), hl.data());
                        const int cr = orc_threshold_filter(tr.data(), dr.data(), m, nh, ratio, minMatchDist, qr.data(), hr.data());
                        std::vector<int> map(m, -1);
                        for (int i = 0; i < cl; ++i) map[ql[i]] = hl[i];
                        for (int i = 0; i < cr; ++i) tot[t] += (map[qr[i]] == hr[i]);
                    }
                    histL.swap(curL);
                    histR.swap(curR);
                }
            }
            delete eL;
            delete eR;
        });
    }
    for (auto& x : th) x.join();
    long sum = 0;
    for (long v : tot) sum += v;
    return sum;
}

}  // extern "C"
