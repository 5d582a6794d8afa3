/* b2orb.h — C ABI of the B200-native ORB extraction + Hamming matching front-end.
 *
 * Drop-in boundary for the stereo visual front-end hot path of ruguangyou/pose-estimation
 * (paths below are relative to the reference tree):
 *   - ORB_SLAM2::ORBextractor            include/ORBextractor.h:45-111, src/ORBextractor.cc:410-470,1043-1105
 *   - cv::BFMatcher(NORM_HAMMING).match  src/feature-tracker.cpp:161-163, :258, :303, :447, :494
 *   - FeatureTracker::internalMatch      src/feature-tracker.cpp:140-229 (extract L/R, match, min-dist + row filter)
 *   - FeatureTracker::externalTrack      src/feature-tracker.cpp:231-328 (matching + distance filter + circular check)
 *
 * POD only: plain pointers and sizes, int status codes, no exceptions, no C++/torch/OpenCV types.
 * Every entry point runs hand-written sm_100a CUDA kernels; there is no CPU fallback — without a
 * CUDA device b2orb_create fails with B2ORB_ERR_CUDA.
 *
 * Threading (mirrors src/feature-tracker.cpp:153-156): one handle owns one device, its CUDA streams
 * and all workspace; calls on one handle must be serialised by the caller; different handles are
 * independent (8 handles on 8 GPUs may live in one process).
 */
#ifndef B2ORB_H_
#define B2ORB_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2ORB_ABI_VERSION 2

enum {
    B2ORB_OK = 0,
    B2ORB_ERR_INVALID = 1,  /* bad argument (NULL, non-positive size, image larger than configured, ...) */
    B2ORB_ERR_CAPACITY = 2, /* caller buffer or configured batch too small */
    B2ORB_ERR_CUDA = 3,     /* CUDA runtime error; text in b2orb_last_error() */
    B2ORB_ERR_NOMEM = 4
};

/* Same layout as cv::KeyPoint (28 bytes): what ORBextractor::operator() fills
 * (src/ORBextractor.cc:818-826, 837-847, 1095-1101).  class_id is always -1. */
typedef struct b2orb_keypoint {
    float x, y;     /* pt, level-0 pixel coordinates */
    float size;     /* (int)(31 * scale[octave]) */
    float angle;    /* degrees in [0,360), cv::fastAtan2 of the intensity centroid */
    float response; /* FAST score */
    int32_t octave;
    int32_t class_id;
} b2orb_keypoint;

/* Constructor arguments of ORBextractor (src/ORBextractor.cc:410-414) + the FeatureTracker matching
 * keys (src/feature-tracker.cpp:8-15; config/euroc.yml:106-112) + workspace sizing. */
typedef struct b2orb_config {
    int32_t nfeatures;    /* numberOfFeatures */
    float scale_factor;   /* scaleFactor, 1 < s <= 3 */
    int32_t nlevels;      /* levelPyramid, 1..12 */
    int32_t ini_th_fast;  /* iniThFAST */
    int32_t min_th_fast;  /* minThFAST */
    int32_t max_width;    /* largest image accepted (<= 4096) */
    int32_t max_height;   /* (<= 4096) */
    int32_t max_images;   /* images resident per launch set; a stereo pair needs 2 */
    int32_t blur_taps[7]; /* symmetric 7-tap integer Gaussian (Q8); all zero = OpenCV 4.x {18,34,48,56,48,34,18}; OpenCV 3.2 (the
                             version Dockerfile.amd64:20 names) rounds to {18,34,49,55,49,34,18} = 257, results saturate at 255 */
    float match_ratio;              /* matchRatio */
    float min_match_dist;           /* minMatchDist */
    float max_vertical_pixel_dist;  /* maxVerticalPixelDist */
} b2orb_config;

typedef struct b2orb_handle b2orb_handle;

int b2orb_abi_version(void);
int b2orb_create(b2orb_handle** out, int device, const b2orb_config* cfg);
void b2orb_destroy(b2orb_handle* h);
const char* b2orb_last_error(const b2orb_handle* h);

/* Upper bound on keypoints one image can return (nfeatures + 3 per level, DistributeOctTree can
 * overshoot its per-level quota by up to 3, src/ORBextractor.cc:673-738). */
int b2orb_keypoint_capacity(const b2orb_handle* h);

/* Getters of include/ORBextractor.h:63-83: out arrays hold nlevels entries. */
int b2orb_get_tables(const b2orb_handle* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2,
                     int32_t* features_per_level);

/* Pinned host memory helpers (cudaHostAlloc) so callers can stage images for full-rate copies. */
void* b2orb_host_alloc(size_t bytes);
void b2orb_host_free(void* p);

/* ORBextractor::operator()(image, mask, keypoints, descriptors), src/ORBextractor.cc:1043-1105.
 * image: host, 8-bit single channel, `step` bytes per row.  Writes *n keypoints (<= cap) and *n x 32
 * descriptor bytes.  Empty image (width or height 0) -> *n = 0 (reference :1046-1047). */
int b2orb_extract(b2orb_handle* h, const uint8_t* image, int width, int height, int step, b2orb_keypoint* kps,
                  uint8_t* desc, int cap, int* n);

/* Batched form: `count` images of identical size, image i at images + i*image_stride; outputs at
 * kps + i*cap, desc + i*cap*32, n[i].  count <= max_images. */
int b2orb_extract_batch(b2orb_handle* h, const uint8_t* images, int count, int width, int height, int step,
                        size_t image_stride, b2orb_keypoint* kps, uint8_t* desc, int cap, int32_t* n);

/* Rectification in front of the pyramid: cv::remap(gray, img, rmap[eye][0], rmap[eye][1], cv::INTER_LINEAR),
 * src/feature-tracker.cpp:75-76, 416-417, with the fixed-point maps initUndistortRectifyMap(..., CV_16SC2, ...)
 * produces (include/cfsd/camera-model.hpp:76-77): map1 = width*height (x, y) int16 pairs, map2 = width*height
 * uint16 (fx | fy << 5, 1/32 px); BORDER_CONSTANT 0.  After this call b2orb_stereo_frontend and
 * b2orb_stereo_frontend_device treat width x height images as UN-rectified camera frames and rectify them on the
 * device while staging them (no extra pass over the image); all four maps NULL switches it off again.  Host
 * pointers, copied. */
int b2orb_set_rectification(b2orb_handle* h, int width, int height, const int16_t* map1_l, const uint16_t* map2_l,
                            const int16_t* map1_r, const uint16_t* map2_r);

/* One cv::remap call with the maps above (eye 0 = left, 1 = right): host image in, host image out. */
int b2orb_remap(b2orb_handle* h, int eye, const uint8_t* image, int width, int height, int step, uint8_t* out,
                int out_step);

/* cfsd live ingest, src/cfsd-state-estimation.cpp:117-127: `frame` is the camera's CV_8UC4 image with both eyes side by
 * side (src_w x src_h, src_step bytes per row, multiple of 4); it is cv::resize'd (INTER_LINEAR) to 2*width x height,
 * converted with cvtColor(BGR2GRAY) (4th channel ignored) and split into the left and right grey eye (width x height
 * each) — one fused kernel.  b2orb_cfsd_ingest returns the two grey images (host in, host out; rectified as well when
 * rectification maps for width x height are set, as VisualInertialSLAM::process -> processImage would);
 * b2orb_stereo_frontend_packed continues on the device with the stereo front end of b2orb_stereo_frontend for
 * `nframes` frames (frame i at frames + i*frame_stride, 2*nframes <= max_images). */
int b2orb_cfsd_ingest(b2orb_handle* h, const uint8_t* frame, int src_w, int src_h, int src_step, int width, int height,
                      uint8_t* gray_l, uint8_t* gray_r, int out_step);

/* cv::BFMatcher(cv::NORM_HAMMING).match(query, train, matches): one result per query row in query
 * order; train_idx = nearest train row (lowest index on ties), dist = Hamming distance.
 * nt == 0 -> train_idx = -1, dist = INT32_MAX.  Host pointers, 32-byte descriptor rows, nt < 2^22. */
int b2orb_match(b2orb_handle* h, const uint8_t* query, int nq, const uint8_t* train, int nt, int32_t* train_idx,
                int32_t* dist);

/* Host-side views of one batch of stereo pairs (all arrays caller-owned, host memory; `cap` entries
 * per image, pair p's slots start at p*cap).  NULL members are skipped (not copied back). */
typedef struct b2orb_stereo_out {
    int32_t cap;              /* >= b2orb_keypoint_capacity() */
    b2orb_keypoint* kp_l;     /* [npairs*cap] */
    b2orb_keypoint* kp_r;
    uint8_t* desc_l;          /* [npairs*cap*32] */
    uint8_t* desc_r;
    int32_t* n_l;             /* [npairs] */
    int32_t* n_r;
    int32_t* train_idx;       /* [npairs*cap] BFMatcher result L->R */
    int32_t* dist;            /* [npairs*cap] */
    int32_t* match_l;         /* [npairs*cap] kept matches (queryIdx), in query order */
    int32_t* match_r;         /* [npairs*cap] (trainIdx) */
    int32_t* n_match;         /* [npairs] */
    /* temporal (externalTrack) part, filled when `temporal` != 0: circular matches between the kept
     * stereo matches of pair p (cur) and pair p-1 (hist); indices are positions in the kept lists. */
    int32_t* circ_cur;        /* [npairs*cap] */
    int32_t* circ_hist;       /* [npairs*cap] */
    int32_t* n_circ;          /* [npairs] (0 for the first pair after b2orb_reset_history) */
    /* triangulation of the kept matches, filled when b2orb_set_projection() is active (NULL = not copied back):
     * pair p, kept match i (same order as match_l / match_r). */
    double* points4d;         /* [npairs][4][cap]: pair p is a 4 x cap row-major matrix = cv::Mat(4, n, CV_64F, ptr, cap*8) */
    double* point_cam;        /* [npairs*cap][3]: (X/W, Y/W, Z/W), the reference's point_wrt_cam */
    uint8_t* depth_flags;     /* [npairs*cap]: bit 0 = !(depth > max_depth), bit 1 = !(depth < 0.1 || depth > max_depth) */
    /* What the reference itself keeps of a frame (src/feature-tracker.cpp:210-218), kept match i of pair p — a caller
     * that sets ONLY these (and the counts / match_* / circ_* lists) gets back a few KB per pair instead of every
     * keypoint and descriptor of both eyes; with pinned arrays the valid prefixes are written straight into them for
     * any batch size. */
    double* pix_l;            /* [npairs*cap][2]: _curPixelsL as cv::Point2d */
    double* pix_r;            /* [npairs*cap][2]: _curPixelsR */
    uint8_t* cur_desc_l;      /* [npairs*cap*32]: rows of _curDescriptorsL */
    uint8_t* cur_desc_r;      /* [npairs*cap*32]: rows of _curDescriptorsR */
    uint8_t* cur_mask;        /* [npairs*cap]: _curFeatureMask after externalTrack (temporal != 0): 0 = circularly matched */
} b2orb_stereo_out;

/* FeatureTracker::internalMatch (useRANSAC=false, src/feature-tracker.cpp:140-164,208-218) for `npairs`
 * independent stereo pairs: extract L and R, BF match L->R, keep matches with
 * dist < max(match_ratio*minDist, min_match_dist) and |yL - yR| < max_vertical_pixel_dist.
 * With temporal == 1 it also runs the matching parts of externalTrack (:258-259,293-296,303-321) between
 * consecutive pairs: cur = kept descriptors of pair p, hist = those of pair p-1 (the last pair of the
 * previous call for p == 0) — a previous-frame stand-in for the pool, NOT the reference's multi-frame feature pool;
 * temporal == 2 matches every pair against the device-resident feature pool instead (b2orb_pool_*, below), which is
 * what externalTrack computes.  2*npairs <= max_images.  Images are host buffers (pinned memory, see
 * b2orb_host_alloc, lets the copies run asynchronously).  Batches longer than 64 pairs are processed as a pipeline
 * of chunks on internal streams so that the H2D copy, the kernels and the D2H copy of different chunks overlap;
 * results do not depend on the chunking (b2orb_debug_set: "pipe_chunk" pairs per chunk, 0 = off; "pipe_lanes" 2..8
 * workspaces, default 3).
 * Calls of up to 4 pairs whose output arrays are all pinned host memory (b2orb_host_alloc, cudaHostAlloc,
 * cudaHostRegister) return their results through one kernel that writes entries [0, n) of every array directly
 * into the caller's buffers; otherwise whole arrays are copied back.  Either way only entries [0, n) are defined.
 * b2orb_debug_set(h, "export_max", 0) forces the copy path.  One-pair calls whose INPUT images are pinned host memory with
 * rows that need re-pitching (width not a multiple of 32) are read by the ingest kernel directly over the bus. */
int b2orb_stereo_frontend(b2orb_handle* h, const uint8_t* images_l, const uint8_t* images_r, int npairs, int width,
                          int height, int step, size_t image_stride, int temporal, const b2orb_stereo_out* out);

/* b2orb_cfsd_ingest followed by b2orb_stereo_frontend without leaving the device (see b2orb_cfsd_ingest). */
int b2orb_stereo_frontend_packed(b2orb_handle* h, const uint8_t* frames, int nframes, int src_w, int src_h, int src_step,
                                 size_t frame_stride, int width, int height, int temporal, const b2orb_stereo_out* out);

/* Same work with inputs ALREADY RESIDENT in device memory and results LEFT on the device (read them with
 * b2orb_stereo_download).  Asynchronous on the handle's stream; b2orb_sync() waits. */
int b2orb_stereo_frontend_device(b2orb_handle* h, const uint8_t* d_images_l, const uint8_t* d_images_r, int npairs,
                                 int width, int height, int step, size_t image_stride, int temporal);
int b2orb_stereo_download(b2orb_handle* h, int npairs, const b2orb_stereo_out* out);
int b2orb_sync(b2orb_handle* h);
void b2orb_reset_history(b2orb_handle* h);

/* cv::triangulatePoints(P1, P2, pixelsL, pixelsR, points4D) as the reference calls it (src/feature-tracker.cpp:367, :489):
 * P1, P2 = 3x4 row-major double projection matrices (CameraModel::_P1/_P2), pix_l / pix_r = n cv::Point2d (x, y
 * doubles), points4d = 4 x n row-major doubles (the CV_64F matrix the reference indexes as points4D.at<double>(r, i)).
 * Same arithmetic as OpenCV (4x4 one-sided Jacobi SVD per point, last row of Vt), one thread per point.  n == 0 is a
 * no-op.  Host pointers. */
int b2orb_triangulate(b2orb_handle* h, const double* P1, const double* P2, const double* pix_l, const double* pix_r,
                      int n, double* points4d);

/* Fused form: after this call b2orb_stereo_frontend* also triangulates the kept stereo matches of every pair on the
 * device (pixels = the float keypoint coordinates widened to double, as _curPixelsL/R are built at
 * src/feature-tracker.cpp:214-215) and fills b2orb_stereo_out.points4d / point_cam / depth_flags with the quantities
 * featurePoolUpdate (:371-378) and structFromMotion (:501-506) derive from points4D.  Both NULL switches it off. */
int b2orb_set_projection(b2orb_handle* h, const double* P1, const double* P2, double max_depth);

/* externalTrack matching parts as one call on host descriptor sets (src/feature-tracker.cpp:258-259,
 * 293-296, 303-321, useRANSAC=false): match(cur_l, hist_l) and match(cur_r, hist_r), each filtered by
 * dist < max(ratio*minDist, min_match_dist), then the circular test mapCurHist[q] == trainIdx.
 * Outputs (cur index, hist index) pairs in query order; n_left / n_right = matches passing the
 * distance filter on each side (the reference's verbose counters :297, :325). */
int b2orb_temporal_track(b2orb_handle* h, const uint8_t* cur_l, const uint8_t* cur_r, int ncur, const uint8_t* hist_l,
                         const uint8_t* hist_r, int nhist, int32_t* circ_cur, int32_t* circ_hist, int32_t* n_circ,
                         int32_t* n_left, int32_t* n_right);

/* ---- next row f3: the feature pool of FeatureTracker, resident on the device ------------------------------------
 * Reference: externalTrack matches _curDescriptorsL/R against _histDescriptorsL/R (src/feature-tracker.cpp:258, :303),
 * two cv::Mats that featurePoolUpdate (:330-412) rebuilds on every keyframe from the pool `_features`.  Here the pooled
 * descriptor rows never leave the GPU: per frame only the circular-match lists, a byte mask and counts come back, and a
 * keyframe update sends one small index list.  The host side mirrors the reference's bookkeeping (ages, feature ids,
 * std::unordered_map iteration order of :349-363, which decides trainIdx and tie-breaks), so ids, ages, pool order,
 * circular matches and mask equal the reference's (useRANSAC = false path; a caller that runs RANSAC between the two
 * matches takes the filtered lists of b2orb_temporal_track instead).  "cur" is always the kept stereo matches of pair
 * `pair` of the LAST b2orb_stereo_frontend* call on this handle (its _curDescriptorsL/R rows are still on the device).
 *
 *   b2orb_pool_reset        empty pool (`_features.clear()`, _featureID = 0), sets maxFeatureAge
 *   b2orb_pool_track        externalTrack(false): match + distance filter + circular test against the pool;
 *                           circ_cur / circ_hist = (queryIdx, trainIdx) in query order, cur_mask = _curFeatureMask,
 *                           matched_ids = _matchedFeatureIDs; arrays sized b2orb_keypoint_capacity(); any may be NULL
 *   b2orb_stereo_frontend(..., temporal = 2)  runs the same track fused behind the stereo filter (one launch chain,
 *                           every pair of the call against the same pool); b2orb_pool_last_track returns pair 0's lists
 *   b2orb_pool_update       featurePoolUpdate: on a keyframe (or while the pool is empty) ages -1 for matched, +2 for
 *                           all, erase above maxFeatureAge, rebuild in container order, append the un-masked current
 *                           features whose accept[i] != 0 (`!(depth > maxDepth)`, :374; NULL = bit 0 of the device depth
 *                           flags when b2orb_set_projection is active, else all).  new_ids = ids given to the appended
 *                           features in order
 *   b2orb_pool_get          _histFeatureIDs, ages and the descriptor rows in pool order (test / debug: copies rows back) */
int b2orb_pool_reset(b2orb_handle* h, int max_feature_age);
int b2orb_pool_size(b2orb_handle* h);
int b2orb_pool_track(b2orb_handle* h, int pair, int32_t* circ_cur, int32_t* circ_hist, int32_t* n_circ, uint8_t* cur_mask,
                     uint64_t* matched_ids, int32_t* n_cur, int32_t* n_left, int32_t* n_right);
int b2orb_pool_last_track(b2orb_handle* h, int32_t* circ_cur, int32_t* circ_hist, int32_t* n_circ, uint8_t* cur_mask,
                          uint64_t* matched_ids, int32_t* n_cur, int32_t* n_left, int32_t* n_right);
int b2orb_pool_update(b2orb_handle* h, int is_keyframe, const uint8_t* accept, uint64_t* new_ids, int32_t* n_inserted,
                      int32_t* n_erased);
int b2orb_pool_get(b2orb_handle* h, uint64_t* ids, int32_t* ages, uint8_t* hist_l, uint8_t* hist_r, int cap, int* n);

/* The host loops of internalMatch on their own (src/feature-tracker.cpp:164, 208-218, useRANSAC = false): global minDist,
 * keep match i iff dist < max(match_ratio*minDist, min_match_dist) and |yL - yR| < max_vertical_pixel_dist; outputs the
 * kept (queryIdx, trainIdx) in query order and, when asked for, _curPixelsL/R (cv::Point2d) and the gathered rows of
 * _curDescriptorsL/R (needs desc_l / desc_r).  Host arrays; nq, nt <= b2orb_keypoint_capacity(); max_images >= 2. */
int b2orb_stereo_filter(b2orb_handle* h, const int32_t* train_idx, const int32_t* dist, int nq, const b2orb_keypoint* kp_l,
                        const uint8_t* desc_l, const b2orb_keypoint* kp_r, const uint8_t* desc_r, int nt, int32_t* match_l,
                        int32_t* match_r, int32_t* n_match, double* pix_l, double* pix_r, uint8_t* cur_desc_l,
                        uint8_t* cur_desc_r);

/* Asynchronous form of b2orb_extract for the reference's two-instances pattern (src/feature-tracker.cpp:153-156): the
 * call enqueues upload + kernels on the handle's stream and returns; b2orb_extract_wait blocks and copies the results.
 * Two handles driven from ONE thread overlap exactly like the reference's two std::threads.  The image must stay valid
 * until the wait returns. */
int b2orb_extract_async(b2orb_handle* h, const uint8_t* image, int width, int height, int step);
int b2orb_extract_wait(b2orb_handle* h, b2orb_keypoint* kps, uint8_t* desc, int cap, int* n);

/* Test/debug access to intermediates of the LAST extract call for image slot `slot`:
 * which = 0 pyramid level, 1 blurred level (dst = w*h bytes, tightly packed). */
int b2orb_debug_level_size(const b2orb_handle* h, int level, int* w, int* hgt);
int b2orb_debug_get_level(b2orb_handle* h, int slot, int level, int which, uint8_t* dst);
/* raw per-cell FAST corners of a level before the quadtree: packed x | y<<12 | response<<24 with x,y
 * relative to the 16-px border (unordered).  Returns the count through *n. */
int b2orb_debug_get_candidates(b2orb_handle* h, int slot, int level, uint32_t* dst, int cap, int* n);

/* Launch-time choices the library normally makes by batch size, pinned per handle for tests and A/B runs (nothing is
 * read from the environment; two handles on two threads may differ).  Keys: "pdl" (-1 auto / 0 / 1 programmatic
 * dependent launch), "fork_max" (blur on a side stream for calls of at most this many images), "export_max" (direct
 * export into pinned outputs for calls of at most this many pairs), "fast_band" (-1 auto, 0 / 1 / 2 / 3 = 16 / 32 / 64 / 128-row
 * FAST bands), "blur_band" (-1 auto, 0 / 1 = 32 / 128-row blur bands), "octree_points" (1 = quadtree by point passes only),
 * "describe_tma" (1 default: descriptor windows fetched as TMA boxes; 0 = per-lane cp.async), "ingest_host" (1 default:
 * one-pair calls read pinned host images in the ingest kernel; 0 = DMA copy + re-pitching kernel), "pipe_chunk" /
 * "pipe_lanes" (pipelined host path, before its first use), "pipe_trace". */
int b2orb_debug_set(b2orb_handle* h, const char* key, int value);

/* Quadtree levels (since the library was loaded, on the handle's device) whose keypoints were so clustered that the
 * count-pyramid path had to be redone with point passes (csrc/octree.cu).  Test / tuning aid. */
uint64_t b2orb_debug_octree_fallbacks(b2orb_handle* h);

/* NO COUNTERPART IN THE REFERENCE.  An allocation-free form of the
 * quadtree keypoint culling (DistributeOctTree, csrc/octree.cu) for callers that cannot call cudaMalloc (CUDA graph
 * capture, tight serving loops, harnesses that forbid it): the call itself invokes no cudaMalloc / cudaFree / allocating
 * library routine; its workspace is a fixed file-scope device array (0.95 MB, held by every GPU that loads the library
 * until exit) instead of the handle's workspace.  Pyramid level `level` of the handle's CURRENT image size (the size of
 * its last call, or max_width x max_height) supplies the geometry: quota, quadtree roots, FAST cell grid.
 *   d_cand   DEVICE pointer: n_cand candidates x | y << 12 | response << 24 (x, y relative to the 16-px border, the
 *            format of b2orb_debug_get_candidates), any order
 *   d_sel    DEVICE pointer, room for quota(level) + 4 entries: the selected keypoints in the same packing, in
 *            DistributeOctTree's list order;  d_n_sel  DEVICE pointer to one int32: their count
 * Limits: quota(level) <= B2ORB_STATIC_MAX_FEATURES and n_cand <= B2ORB_STATIC_MAX_CANDIDATES (and the level's own
 * candidate capacity), else B2ORB_ERR_CAPACITY.  Asynchronous on the handle's stream (b2orb_sync waits).  ALL calls, on
 * every handle of a device, share the one workspace, which keeps the previous call's contents (the kernel initialises
 * what it reads): calls that overlap on the GPU race — serialise them. */
#define B2ORB_STATIC_MAX_FEATURES 4096
#define B2ORB_STATIC_MAX_CANDIDATES 262144
int b2orb_distribute_static(b2orb_handle* h, int level, const uint32_t* d_cand, int n_cand, uint32_t* d_sel, int32_t* d_n_sel);

/* The handle's CUDA stream (a cudaStream_t) so callers can record their own events on it. */
void* b2orb_stream(const b2orb_handle* h);

/* Per-kernel device timing: when enabled, CUDA events are recorded on the handle's stream between
 * launches (no extra synchronisation); b2orb_profile_read() synchronises, then returns the
 * accumulated milliseconds and launch counts per kernel id since b2orb_profile_enable(h, 1). */
enum {
    B2ORB_K_RESIZE = 0, B2ORB_K_FAST, B2ORB_K_OCTREE, B2ORB_K_BLUR, B2ORB_K_DESCRIBE, B2ORB_K_MATCH_STEREO,
    B2ORB_K_STEREO_FILTER, B2ORB_K_MATCH_TEMPORAL, B2ORB_K_CIRCULAR, B2ORB_K_FAST_CELLS, B2ORB_K_INGEST, B2ORB_K_TRIANGULATE,
    B2ORB_NUM_KERNELS
};
void b2orb_profile_enable(b2orb_handle* h, int on);
int b2orb_profile_read(b2orb_handle* h, double* ms /*[B2ORB_NUM_KERNELS]*/, uint64_t* launches /*[B2ORB_NUM_KERNELS]*/);

/* Kernel launches issued by this handle since creation (bench.py reports the per-step delta). */
uint64_t b2orb_launch_count(const b2orb_handle* h);
/* Device time (ms, CUDA events on the handle's stream) of the last *_device / frontend call. */
float b2orb_last_device_ms(const b2orb_handle* h);

#ifdef __cplusplus
}
#endif
#endif /* B2ORB_H_ */
