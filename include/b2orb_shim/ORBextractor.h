/* Drop-in replacement for the reference's include/ORBextractor.h (ORB_SLAM2::ORBextractor, lines 45-111) and for
 * the three cv::BFMatcher(cv::NORM_HAMMING).match call sites of src/feature-tracker.cpp (:161-163, :258, :303),
 * implemented over the C ABI in include/b2orb.h.  Same namespace, class name, constructor, operator(), getters and
 * public member; the arithmetic runs in libb2orb.so's sm_100a kernels.  There is no CPU fallback: construction
 * throws std::runtime_error when no CUDA device / library is available.
 *
 * Use: put this directory in front of the reference's include/ on the include path, drop src/ORBextractor.cc from
 * the build and link libb2orb.so.  Needs only OpenCV core + features2d headers.
 */
#ifndef ORBEXTRACTOR_H   /* same guard as the reference header so it cannot be included twice */
#define ORBEXTRACTOR_H

#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include <opencv2/core.hpp>
#include <opencv2/features2d.hpp>

#include "../b2orb.h"

namespace ORB_SLAM2 {

class ORBextractor {
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

    /* include/ORBextractor.h:51-52.  Workspace is sized lazily on the first image (and re-sized when a larger
     * one arrives), like the reference's mvImagePyramid. */
    ORBextractor(int nfeatures_, float scaleFactor_, int nlevels_, int iniThFAST_, int minThFAST_, int device = 0)
        : nfeatures(nfeatures_), scaleFactor(scaleFactor_), nlevels(nlevels_), iniThFAST(iniThFAST_),
          minThFAST(minThFAST_), device_(device) {
        ensure(64, 64);   /* also fills the scale tables (src/ORBextractor.cc:415-431) */
    }
    ~ORBextractor() { b2orb_destroy(h_); }
    ORBextractor(const ORBextractor&) = delete;
    ORBextractor& operator=(const ORBextractor&) = delete;

    /* include/ORBextractor.h:58-61, src/ORBextractor.cc:1043-1105.  Mask is ignored, as in the reference. */
    void operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint>& _keypoints,
                    cv::OutputArray _descriptors) {
        if (_image.empty()) return;                                   /* :1046-1047 */
        cv::Mat image = _image.getMat();
        CV_Assert(image.type() == CV_8UC1);                           /* :1050 */
        ensure(image.cols, image.rows);
        const int cap = b2orb_keypoint_capacity(h_);
        kp_.resize(cap);
        desc_.resize((size_t)cap * 32);
        int n = 0;
        check(b2orb_extract(h_, image.data, image.cols, image.rows, (int)image.step, kp_.data(), desc_.data(), cap, &n));
        _keypoints.clear();
        if (n == 0) { _descriptors.release(); return; }               /* :1064-1065 */
        _descriptors.create(n, 32, CV_8U);                            /* :1068 */
        cv::Mat descriptors = _descriptors.getMat();
        _keypoints.reserve(n);
        for (int i = 0; i < n; ++i) {
            const b2orb_keypoint& k = kp_[i];
            _keypoints.push_back(cv::KeyPoint(k.x, k.y, k.size, k.angle, k.response, k.octave, k.class_id));
            std::memcpy(descriptors.ptr(i), &desc_[(size_t)i * 32], 32);
        }
    }

    int inline GetLevels() { return nlevels; }
    float inline GetScaleFactor() { return (float)scaleFactor; }
    std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
    std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
    std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
    std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

    /* Public in the reference (:85) but never read outside the class; left empty (the pyramid lives in HBM). */
    std::vector<cv::Mat> mvImagePyramid;

    b2orb_handle* handle() { return h_; }

protected:
    void check(int rc) {
        if (rc != B2ORB_OK) throw std::runtime_error(std::string("b2orb: ") + b2orb_last_error(h_));
    }
    void ensure(int w, int h) {
        if (h_ && w <= max_w_ && h <= max_h_) return;
        if (h_) b2orb_destroy(h_);
        h_ = nullptr;
        max_w_ = w > max_w_ ? w : max_w_;
        max_h_ = h > max_h_ ? h : max_h_;
        b2orb_config cfg = {};
        cfg.nfeatures = nfeatures; cfg.scale_factor = (float)scaleFactor; cfg.nlevels = nlevels;
        cfg.ini_th_fast = iniThFAST; cfg.min_th_fast = minThFAST;
        cfg.max_width = max_w_; cfg.max_height = max_h_; cfg.max_images = 1;
        cfg.match_ratio = 2.f; cfg.min_match_dist = 30.f; cfg.max_vertical_pixel_dist = 0.1f;
        const int rc = b2orb_create(&h_, device_, &cfg);
        if (rc != B2ORB_OK) {
            std::string msg = h_ ? b2orb_last_error(h_) : "b2orb_create failed (no CUDA device?)";
            b2orb_destroy(h_);
            h_ = nullptr;
            throw std::runtime_error("b2orb: " + msg);
        }
        mvScaleFactor.resize(nlevels); mvInvScaleFactor.resize(nlevels);
        mvLevelSigma2.resize(nlevels); mvInvLevelSigma2.resize(nlevels); mnFeaturesPerLevel.resize(nlevels);
        b2orb_get_tables(h_, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(),
                         mvInvLevelSigma2.data(), mnFeaturesPerLevel.data());
    }

    int nfeatures;
    double scaleFactor;
    int nlevels;
    int iniThFAST;
    int minThFAST;
    std::vector<int> mnFeaturesPerLevel;
    std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;

    b2orb_handle* h_ = nullptr;
    int device_ = 0, max_w_ = 0, max_h_ = 0;
    std::vector<b2orb_keypoint> kp_;
    std::vector<uint8_t> desc_;
};

}  // namespace ORB_SLAM2

namespace b2orb {

/* Replacement for `cv::BFMatcher matcher(cv::NORM_HAMMING); matcher.match(q, t, matches);`
 * (src/feature-tracker.cpp:161-163, :258, :303, :447, :494): same result vector — one DMatch per query row in
 * query order, nearest train row with the lowest index on ties, distance as float. */
class BFMatcherHamming {
public:
    explicit BFMatcherHamming(b2orb_handle* h) : h_(h) {}
    void match(const cv::Mat& query, const cv::Mat& train, std::vector<cv::DMatch>& matches) {
        matches.clear();
        if (query.empty() || train.empty()) return;
        CV_Assert(query.type() == CV_8U && train.type() == CV_8U && query.cols == 32 && train.cols == 32);
        cv::Mat q = query.isContinuous() ? query : query.clone(), t = train.isContinuous() ? train : train.clone();
        idx_.resize(q.rows);
        dist_.resize(q.rows);
        if (b2orb_match(h_, q.data, q.rows, t.data, t.rows, idx_.data(), dist_.data()) != B2ORB_OK)
            throw std::runtime_error(std::string("b2orb: ") + b2orb_last_error(h_));
        matches.reserve(q.rows);
        for (int i = 0; i < q.rows; ++i) matches.push_back(cv::DMatch(i, idx_[i], 0, (float)dist_[i]));
    }

private:
    b2orb_handle* h_;
    std::vector<int32_t> idx_, dist_;
};

/* Replacement for `cv::triangulatePoints(_pCameraModel->_P1, _pCameraModel->_P2, pixelsL, pixelsR, points4D)`
 * (src/feature-tracker.cpp:367, :489; next row f4): P1, P2 = 3x4 CV_64F, pixels = std::vector<cv::Point2d>, points4D
 * becomes a 4 x N CV_64F matrix with the same values OpenCV's own routine produces (read by the reference as
 * points4D.at<double>(r, i)). */
inline void triangulatePoints(b2orb_handle* h, const cv::Mat& P1, const cv::Mat& P2, const std::vector<cv::Point2d>& pixelsL,
                              const std::vector<cv::Point2d>& pixelsR, cv::Mat& points4D) {
    CV_Assert(P1.type() == CV_64F && P2.type() == CV_64F && P1.rows == 3 && P1.cols == 4 && P2.rows == 3 && P2.cols == 4);
    CV_Assert(pixelsL.size() == pixelsR.size());
    static_assert(sizeof(cv::Point2d) == 2 * sizeof(double), "cv::Point2d is two packed doubles");
    cv::Mat p1 = P1.isContinuous() ? P1 : P1.clone(), p2 = P2.isContinuous() ? P2 : P2.clone();
    const int n = (int)pixelsL.size();
    points4D.create(4, n, CV_64F);
    if (b2orb_triangulate(h, (const double*)p1.data, (const double*)p2.data, n ? &pixelsL[0].x : nullptr,
                          n ? &pixelsR[0].x : nullptr, n, (double*)points4D.data) != B2ORB_OK)
        throw std::runtime_error(std::string("b2orb: ") + b2orb_last_error(h));
}

}  // namespace b2orb

#endif /* ORBEXTRACTOR_H */
