// Exercises the drop-in C++ shim exactly like the reference's call sites:
//   FeatureTracker::extractORB     (src/feature-tracker.cpp:133-138)  (*_ORBextractorLeft)(img, cv::Mat(), kps, desc)
//   FeatureTracker::internalMatch  (:161-164)                         matcher.match(descL, descR, matches)
//   FeatureTracker::featurePoolUpdate (:365-371)                      triangulatePoints(P1, P2, pixelsL, pixelsR, points4D)
// Input: raw u8 images (w h, then L and R pixel files); output: text dump compared by tests/test_cpp_shim.py.
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <functional>
#include "ORBextractor.h"

static std::vector<unsigned char> slurp(const char* path, size_t n) {
    std::vector<unsigned char> b(n);
    FILE* f = fopen(path, "rb");
    if (!f || fread(b.data(), 1, n, f) != n) { fprintf(stderr, "cannot read %s\n", path); exit(2); }
    fclose(f);
    return b;
}

int main(int argc, char** argv) {
    if (argc < 9) { fprintf(stderr, "usage: shim_demo w h nfeat ini min left.raw right.raw out.txt\n"); return 2; }
    const int w = atoi(argv[1]), h = atoi(argv[2]), nf = atoi(argv[3]), ini = atoi(argv[4]), mn = atoi(argv[5]);
    std::vector<unsigned char> bl = slurp(argv[6], (size_t)w * h), br = slurp(argv[7], (size_t)w * h);
    cv::Mat imgL(h, w, CV_8UC1, bl.data(), w), imgR(h, w, CV_8UC1, br.data(), w);
    ORB_SLAM2::ORBextractor left(nf, 1.2f, 8, ini, mn), right(nf, 1.2f, 8, ini, mn);
    std::vector<cv::KeyPoint> kl, kr;
    cv::Mat dl, dr;
    // two extractor instances driven from two threads, as in the reference (:153-156)
    std::thread tl([&] { left(imgL, cv::Mat(), kl, dl); });
    std::thread tr([&] { right(imgR, cv::Mat(), kr, dr); });
    tl.join(); tr.join();
    b2orb::BFMatcherHamming matcher(left.handle());
    std::vector<cv::DMatch> matches;
    matcher.match(dl, dr, matches);
    FILE* f = fopen(argv[8], "w");
    fprintf(f, "%zu %zu %zu %d %.9g\n", kl.size(), kr.size(), matches.size(), left.GetLevels(), left.GetScaleFactors()[7]);
    for (size_t i = 0; i < kl.size(); ++i) {
        fprintf(f, "%.9g %.9g %.9g %.9g %.9g %d", kl[i].pt.x, kl[i].pt.y, kl[i].size, kl[i].angle, kl[i].response, kl[i].octave);
        for (int b = 0; b < 32; ++b) fprintf(f, " %d", dl.ptr((int)i)[b]);
        fprintf(f, " %d %d\n", matches[i].trainIdx, (int)matches[i].distance);
    }
    // every match triangulated the way featurePoolUpdate does it (Point2f -> Point2d pixels, 3x4 CV_64F projections)
    cv::Mat P1(3, 4, CV_64F), P2(3, 4, CV_64F);
    const double fx = 435.2047, cx = 367.4517 * w / 752.0, cy = 252.2008 * h / 480.0, bfx = 47.9063;
    P1.at<double>(0, 0) = P2.at<double>(0, 0) = fx; P1.at<double>(1, 1) = P2.at<double>(1, 1) = fx;
    P1.at<double>(0, 2) = P2.at<double>(0, 2) = cx; P1.at<double>(1, 2) = P2.at<double>(1, 2) = cy;
    P1.at<double>(2, 2) = P2.at<double>(2, 2) = 1.0; P2.at<double>(0, 3) = bfx;
    std::vector<cv::Point2d> pixelsL, pixelsR;
    for (auto& m : matches) { pixelsL.push_back(kl[m.queryIdx].pt); pixelsR.push_back(kr[m.trainIdx].pt); }
    cv::Mat points4D;
    b2orb::triangulatePoints(left.handle(), P1, P2, pixelsL, pixelsR, points4D);
    fprintf(f, "P %.17g %.17g %.17g %.17g\n", fx, cx, cy, bfx);
    for (int i = 0; i < points4D.cols; ++i)
        fprintf(f, "T %.17g %.17g %.17g %.17g\n", points4D.at<double>(0, i), points4D.at<double>(1, i),
                points4D.at<double>(2, i), points4D.at<double>(3, i));
    fclose(f);
    return 0;
}
