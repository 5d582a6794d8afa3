// triangulate.cu — next row f4: cv::triangulatePoints + the depth gates of its two call sites.
//
// Restates (reference paths relative to /root/reference):
//   * featurePoolUpdate   src/feature-tracker.cpp:365-379   triangulatePoints(P1, P2, _curPixelsL, _curPixelsR, points4D),
//                         depth = Z/W, skip when depth > maxDepth, point_wrt_cam = (X/W, Y/W, depth)
//   * structFromMotion    src/feature-tracker.cpp:487-509   same call, gate depth < 0.1 || depth > maxDepth
// cv::triangulatePoints (OpenCV calib3d, un-vendored dependency): per point the 4x4 matrix
//   A[2j+0][k] = x_j * P_j[2][k] - P_j[0][k],  A[2j+1][k] = y_j * P_j[2][k] - P_j[1][k]
// goes through cv::SVD::compute and the point is the last row of Vt.  For a 4x4 double matrix that is OpenCV's own
// one-sided Jacobi (core/lapack.cpp JacobiSVDImpl_<double>) on At = A^T: eps = 10*DBL_EPSILON, at most 30 sweeps over
// the 6 column pairs, file-local hypot (scale by the larger magnitude), descending selection sort of the singular
// values.  One thread per match runs exactly that sequence of IEEE double operations (the file is compiled with
// --fmad=false; CUDA's double add / mul / div / sqrt are correctly rounded), so points4D is bit-identical to OpenCV's
// on the same inputs (tests/test_triangulate.py against the oracle, the oracle against cv2.triangulatePoints).
//
// Everything lives in registers: At (16) + Vt (16) + W (4) doubles; the pair loop is fully unrolled so no array is
// indexed dynamically.  fp64 work, ~2.5 kFLOP per point — a few hundred matches per stereo pair, so the kernel is
// latency-sized, not throughput-sized; it rides on the stereo filter's output without leaving the device.
#include "b2orb_internal.cuh"

namespace {

// lapack.cpp's file-local hypot: a > b ? a*sqrt(1+(b/a)^2) : b > 0 ? b*sqrt(1+(a/b)^2) : 0 on magnitudes — i.e. scale by
// the larger one (equal magnitudes give ratio 1 on either path), written without branches.
__device__ __forceinline__ double cv_lapack_hypot(double a, double b) {
    a = fabs(a);
    b = fabs(b);
    const double hi = a > b ? a : b, lo = a > b ? b : a;
    const double r = lo / hi;
    const double v = hi * sqrt(1 + r * r);
    return hi > 0 ? v : 0;
}

// One Jacobi rotation of columns (rows of At) i and j; returns whether it rotated.
__device__ __forceinline__ bool rotate_pair(double (&Ai)[4], double (&Aj)[4], double (&Vi)[4], double (&Vj)[4], double& Wi,
                                            double& Wj) {
    const double eps = 2.220446049250313e-16 * 10;
    double a = Wi, p = 0, b = Wj;
#pragma unroll
    for (int k = 0; k < 4; ++k) p += Ai[k] * Aj[k];
    if (fabs(p) <= eps * sqrt(a * b)) return false;
    p *= 2;
    const double beta = a - b, gamma = cv_lapack_hypot(p, beta);
    // The reference's two branches are the same operations on swapped operands; one code path keeps the warp converged:
    //   beta < 0:  delta = (gamma - beta)*0.5; s = sqrt(delta/gamma);       c = p/(gamma*s*2)
    //   else:      c = sqrt((gamma + beta)/(gamma*2));                      s = p/(gamma*c*2)
    const bool neg = beta < 0;
    const double num = neg ? (gamma - beta) * 0.5 : (gamma + beta);
    const double den = neg ? gamma : gamma * 2;
    const double r = sqrt(num / den);
    const double o = p / (gamma * r * 2);
    const double c = neg ? o : r, s = neg ? r : o;
    a = b = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const double t0 = c * Ai[k] + s * Aj[k];
        const double t1 = -s * Ai[k] + c * Aj[k];
        Ai[k] = t0;
        Aj[k] = t1;
        a += t0 * t0;
        b += t1 * t1;
    }
    Wi = a;
    Wj = b;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const double t0 = c * Vi[k] + s * Vj[k];
        const double t1 = -s * Vi[k] + c * Vj[k];
        Vi[k] = t0;
        Vj[k] = t1;
    }
    return true;
}

__device__ __forceinline__ void triangulate_one(const B2Projection& pr, double xl, double yl, double xr, double yr,
                                                double (&X)[4]) {
    double At[4][4], Vt[4][4], W[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        At[k][0] = xl * pr.P1[8 + k] - pr.P1[k];
        At[k][1] = yl * pr.P1[8 + k] - pr.P1[4 + k];
        At[k][2] = xr * pr.P2[8 + k] - pr.P2[k];
        At[k][3] = yr * pr.P2[8 + k] - pr.P2[4 + k];
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        double sd = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) sd += At[i][k] * At[i][k];
        W[i] = sd;
#pragma unroll
        for (int k = 0; k < 4; ++k) Vt[i][k] = (i == k) ? 1.0 : 0.0;
    }
#pragma unroll 1
    for (int iter = 0; iter < 30; ++iter) {
        bool changed = false;
        changed |= rotate_pair(At[0], At[1], Vt[0], Vt[1], W[0], W[1]);
        changed |= rotate_pair(At[0], At[2], Vt[0], Vt[2], W[0], W[2]);
        changed |= rotate_pair(At[0], At[3], Vt[0], Vt[3], W[0], W[3]);
        changed |= rotate_pair(At[1], At[2], Vt[1], Vt[2], W[1], W[2]);
        changed |= rotate_pair(At[1], At[3], Vt[1], Vt[3], W[1], W[3]);
        changed |= rotate_pair(At[2], At[3], Vt[2], Vt[3], W[2], W[3]);
        if (!changed) break;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        double sd = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) sd += At[i][k] * At[i][k];
        W[i] = sqrt(sd);
    }
    // descending selection sort (lapack.cpp): only the row that ends up last matters.  Track the permutation.
    int ord0 = 0, ord1 = 1, ord2 = 2, ord3 = 3;
    {   // i = 0
        int j = 0;
        double wj = W[0];
        if (wj < W[1]) { j = 1; wj = W[1]; }
        if (wj < W[2]) { j = 2; wj = W[2]; }
        if (wj < W[3]) { j = 3; wj = W[3]; }
        if (j == 1) { const double t = W[0]; W[0] = W[1]; W[1] = t; ord0 = 1; ord1 = 0; }
        else if (j == 2) { const double t = W[0]; W[0] = W[2]; W[2] = t; ord0 = 2; ord2 = 0; }
        else if (j == 3) { const double t = W[0]; W[0] = W[3]; W[3] = t; ord0 = 3; ord3 = 0; }
    }
    {   // i = 1
        int j = 1;
        double wj = W[1];
        if (wj < W[2]) { j = 2; wj = W[2]; }
        if (wj < W[3]) { j = 3; wj = W[3]; }
        if (j == 2) { const double t = W[1]; W[1] = W[2]; W[2] = t; const int o = ord1; ord1 = ord2; ord2 = o; }
        else if (j == 3) { const double t = W[1]; W[1] = W[3]; W[3] = t; const int o = ord1; ord1 = ord3; ord3 = o; }
    }
    if (W[2] < W[3]) { const int o = ord2; ord2 = ord3; ord3 = o; }   // i = 2
    (void)ord0; (void)ord1; (void)ord2;
#pragma unroll
    for (int k = 0; k < 4; ++k)
        X[k] = ord3 == 0 ? Vt[0][k] : ord3 == 1 ? Vt[1][k] : ord3 == 2 ? Vt[2][k] : Vt[3][k];
}

__device__ __forceinline__ void store_point(const double (&X)[4], double max_depth, double* p4, size_t row_stride,
                                            double* point_cam, uint8_t* flags, int i) {
    if (p4) {
#pragma unroll
        for (int k = 0; k < 4; ++k) p4[(size_t)k * row_stride + i] = X[k];
    }
    const double depth = X[2] / X[3];
    if (point_cam) {
        point_cam[3 * (size_t)i + 0] = X[0] / X[3];
        point_cam[3 * (size_t)i + 1] = X[1] / X[3];
        point_cam[3 * (size_t)i + 2] = depth;
    }
    if (flags) {
        const bool pool_ok = !(depth > max_depth);                    // src/feature-tracker.cpp:372
        const bool sfm_ok = !(depth < 0.1 || depth > max_depth);      // :503
        flags[i] = (uint8_t)((pool_ok ? 1 : 0) | (sfm_ok ? 2 : 0));
    }
}

// Generic form: n Point2d pairs -> 4 x n row-major points4D.
__global__ void __launch_bounds__(128) triangulate_kernel(B2Projection pr, const double2* __restrict__ pix_l,
                                                          const double2* __restrict__ pix_r, int n, double* points4d) {
    b2_pdl_wait();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double2 l = pix_l[i], r = pix_r[i];
    double X[4];
    triangulate_one(pr, l.x, l.y, r.x, r.y, X);
    store_point(X, pr.max_depth, points4d, (size_t)n, nullptr, nullptr, i);
}

// Fused form behind the stereo filter: pair p = blockIdx.y, kept match i -> keypoints match_l[i] (left image slot p)
// and match_r[i] (right image slot npairs + p); _curPixelsL/R are cv::Point2d built from the float keypoint
// coordinates (src/feature-tracker.cpp:214-215), i.e. an exact float -> double widening.
__global__ void __launch_bounds__(128) triangulate_matches_kernel(B2Projection pr, const b2orb_keypoint* __restrict__ kps,
                                                                  const int* __restrict__ match_l,
                                                                  const int* __restrict__ match_r,
                                                                  const int* __restrict__ n_match, int cap, int npairs,
                                                                  double* points4d, double* point_cam, uint8_t* flags) {
    b2_pdl_wait();
    const int p = blockIdx.y;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_match[p]) return;
    const size_t base = (size_t)p * cap;
    const b2orb_keypoint kl = kps[base + match_l[base + i]];
    const b2orb_keypoint kr = kps[((size_t)npairs + p) * cap + match_r[base + i]];
    double X[4];
    triangulate_one(pr, (double)kl.x, (double)kl.y, (double)kr.x, (double)kr.y, X);
    store_point(X, pr.max_depth, points4d + base * 4, (size_t)cap, point_cam + base * 3, flags + base, i);
}

}  // namespace

void b2_launch_triangulate(const B2Projection& pr, const double* pix_l, const double* pix_r, int n, double* points4d,
                           cudaStream_t s) {
    if (n <= 0) return;
    b2_launch_pdl(triangulate_kernel, dim3((n + 127) / 128), dim3(128), 0, s, pr, reinterpret_cast<const double2*>(pix_l),
                                                       reinterpret_cast<const double2*>(pix_r), n, points4d);
}

void b2_launch_triangulate_matches(const B2Projection& pr, const void* kps, const int* match_l, const int* match_r,
                                   const int* n_match, int cap, int npairs, double* points4d, double* point_cam,
                                   uint8_t* flags, cudaStream_t s) {
    if (npairs <= 0) return;
    dim3 grid((cap + 127) / 128, npairs);
    b2_launch_pdl(triangulate_matches_kernel, dim3(grid), dim3(128), 0, s, pr, static_cast<const b2orb_keypoint*>(kps), match_l, match_r, n_match,
                                                    cap, npairs, points4d, point_cam, flags);
}
