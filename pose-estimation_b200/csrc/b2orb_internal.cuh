// Internal declarations shared by the sm_100a kernels and the C-ABI layer (not installed).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/b2orb.h"

// ---- programmatic dependent launch ------------------------------------------------------------------------------
// Every kernel of the per-frame sequence is launched with programmatic stream serialisation and starts with
// b2_pdl_wait(): the next kernel's CTAs may be scheduled while the previous kernel drains (launch latency hidden —
// it matters for one stereo pair per call, where ~25 launches of 5-40 us each run back to back), but they touch no
// global memory before the previous grid has completed and flushed, so the result is the one of plain stream order.
// Early-scheduled CTAs hold SM slots while they wait, which costs throughput when another stream could have used
// them (measured: -5 % on 128-pair batches over two streams), so the C-ABI layer switches the attribute on only for
// small batches (b2_pdl_on, per host thread; b2orb_debug_set(h, "pdl", 0/1) overrides).  Without the attribute the wait
// is a no-op.
extern thread_local bool b2_pdl_on;
#ifdef __CUDACC__
__device__ __forceinline__ void b2_pdl_wait() {
    asm volatile("griddepcontrol.launch_dependents;\n\tgriddepcontrol.wait;" ::: "memory");
}
template <typename... KArgs, typename... Args>
inline cudaError_t b2_launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = b2_pdl_on ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

#define B2_MAX_LEVELS 12
#define B2_EDGE 19          // EDGE_THRESHOLD, src/ORBextractor.cc:74
#define B2_BORDER 16        // EDGE_THRESHOLD-3 = minBorderX/Y, src/ORBextractor.cc:773-776
#define B2_FAST_BAND 16     // corner-map rows per FAST score task (fast.cu): variant v uses B2_FAST_BAND << v rows
#define B2_FAST_VARIANTS 4  // 16 rows (latency-sized calls), 32, 64, 128 (large batches: less halo work)
#define B2_BLUR_BAND 32     // output rows per blur task (blur.cu): variant v uses B2_BLUR_BAND << (2 * v) rows
#define B2_BLUR_VARIANTS 2  // 32 rows, 128 rows (large batches: 6 halo rows per band are 19 % / 5 % extra work)
#define B2_HALF_PATCH 15    // HALF_PATCH_SIZE, src/ORBextractor.cc:73

// Per-level geometry, fixed for (config, image size).  Passed to kernels BY VALUE as one
// __grid_constant__ parameter, so several handles with different configs can coexist.
struct B2Level {
    int w, h, pitch;          // level image size, row pitch in bytes (multiple of 16)
    unsigned off;             // byte offset of the level inside one image's pyramid / blur slab
    int n_cols, n_rows;       // FAST cell grid (src/ORBextractor.cc:784-787)
    int w_cell, h_cell;
    int cell_base;            // first flattened cell index of this level
    int fast_task_base[B2_FAST_VARIANTS];    // first flattened FAST score task (band, strip) of this level, per band variant
    int fast_strips;          // 120-px strips per row band (fast.cu)
    int blur_task_base[B2_BLUR_VARIANTS];    // first flattened blur task (band, strip) of this level, per band variant
    int blur_strips;          // 128-px strips per row band
    int quota;                // mnFeaturesPerLevel[level]
    int sel_base;             // first slot of this level in the per-image selected-keypoint array
    int sel_cap;              // quota + 4
    unsigned cand_off;        // first slot of this level in the per-image candidate array
    int cand_cap;
    int node_cap;             // max(quota + 3, 4*n_ini) + 1: the list never grows beyond that (octree.cu)
    int ot_depth;             // depth D of the quadtree count pyramid (octree.cu); 0 = point passes only
    unsigned node_off;        // first node slot of this level in the per-image node workspace
    int n_ini;                // quadtree roots (src/ORBextractor.cc:543)
    float h_x;                // root width (src/ORBextractor.cc:545)
    float scale;              // mvScaleFactor[level]
    float kp_size;            // (float)(int)(31*scale)
};

struct B2Geom {
    int nlevels;
    int ini_th, min_th;
    int total_cells;          // over all levels
    int total_blur_tasks[B2_BLUR_VARIANTS];
    int total_fast_tasks[B2_FAST_VARIANTS];  // per band variant (short bands for small batches, long bands for less halo work)
    int sel_total;            // per-image selected-keypoint slots (sum of sel_cap)
    int kp_cap;               // per-image output capacity
    unsigned cand_total;      // per-image candidate slots
    unsigned node_total;      // per-image node slots
    unsigned slab;            // bytes of one image's pyramid (levels 1..) / blur (levels 0..) slab
    int umax[16];             // src/ORBextractor.cc:454-469
    int taps[7];
    int tap_sat;              // taps sum to more than 256: the blur saturates its result (OpenCV 3.2 kernel)
    unsigned tapv[10];        // blur.cu: u8x4 tap vectors for the 4 output pixels of a lane (host-built)
    B2Level lv[B2_MAX_LEVELS];
};

// Level-0 pixels are read where the caller (or the H2D copy) put them; levels >= 1 live in the
// pyramid slab.  All kernels use this one accessor.
struct B2Images {
    const uint8_t* lvl0[2];   // image b < split at lvl0[0] + b*stride0, else lvl0[1] + (b-split)*stride0
    int split;
    size_t stride0;
    int pitch0;
    uint8_t* pyr;             // image b's slab at pyr + b*slab (levels >= 1)
    uint8_t* blur;            // same layout, all levels
};

__device__ __forceinline__ const uint8_t* b2_level_ptr(const B2Geom& g, const B2Images& im, int b, int level,
                                                       int& pitch) {
    if (level == 0) {
        pitch = im.pitch0;
        return b < im.split ? im.lvl0[0] + (size_t)b * im.stride0 : im.lvl0[1] + (size_t)(b - im.split) * im.stride0;
    }
    pitch = g.lv[level].pitch;
    return im.pyr + (size_t)b * g.slab + g.lv[level].off;
}

// Quadtree node (list entry).  The node array IS the std::list of DistributeOctTree in list order.
struct B2Node {
    unsigned short x0, y0, x1, y1;  // UL.x, UL.y, BR.x, BR.y (relative to the 16-px border)
    unsigned cnt;                   // points inside
    unsigned code;                  // depth << 28 | path (root index, then 2 bits per split): slot in the count pyramid
};

// One Hamming matching problem (BFMatcher.match call): rows are 32-byte descriptors.
struct B2MatchProblem {
    const uint8_t* q;
    const uint8_t* t;
    const int* nq;       // device counts (read at kernel time, so upstream kernels can set them)
    const int* nt;
    int* train_idx;
    int* dist;
};

// ---- kernel launchers (each file documents the reference lines it restates) --------------------
void b2_launch_resize(const B2Geom& g, const B2Images& im, int level, const int* xofs, const short* xw,
                      const int* yofs, const short* yw, int nimg, cudaStream_t s);
// cv::remap with fixed-point maps (map1: short2 per pixel, map2: u16 per pixel), source and output of size w x h
void b2_launch_remap(const uint8_t* src, int step, size_t stride, const void* map1, const void* map2, uint8_t* dst,
                     int dpitch, size_t dstride, int w, int h, int count, cudaStream_t s);
// side-by-side 4-channel frame -> resize -> grey -> left / right eye (src/cfsd-state-estimation.cpp:117-127)
void b2_launch_ingest_bgra(const uint8_t* frames, int sstep, size_t sstride, int sw, int sh, const int* xofs, const short* xw,
                           const int* yofs, const short* yw, uint8_t* dst_l, uint8_t* dst_r, int dpitch, size_t dstride,
                           int half_w, int dh, int count, cudaStream_t s);
// src0 (and src1, optional) = device-visible addresses of pinned HOST images; false when the shape is out of its range
bool b2_launch_ingest_host(const uint8_t* src0, const uint8_t* src1, int step, uint8_t* dst, int dpitch, size_t dstride,
                           int w, int h, cudaStream_t s);
void b2_launch_repack(const uint8_t* src, int step, size_t stride, uint8_t* dst, int dpitch, size_t dstride, int w,
                      int h, int count, cudaStream_t s);
void b2_launch_blur(const B2Geom& g, const B2Images& im, int nimg, int forced_band, cudaStream_t s);
// fast_score appends the non-zero words of the NMS'd corner map (4 px of b + position, uint2) to a raw list and raises
// the per-cell "iniThFAST found something" flags; fast_vote applies the per-cell threshold vote and expands the words
// into candidates x | y<<12 | (b-1)<<24
void b2_launch_fast_score(const B2Geom& g, const B2Images& im, uint2* raw, int* raw_cnt, uint8_t* cell_ini, int nimg,
                          int forced_band, cudaStream_t s);
void b2_launch_fast_vote(const B2Geom& g, const uint2* raw, const int* raw_cnt, const uint8_t* cell_ini, uint32_t* cand,
                         int* cand_cnt, int nimg, cudaStream_t s);
void b2_launch_octree(const B2Geom& g, const uint32_t* cand, const int* cand_cnt, unsigned short* pnode,
                      unsigned* scratch, uint32_t* sel, int* sel_cnt, int nimg, int force_legacy, cudaStream_t s);
struct B2DescCache;   // describe.cu: tensor maps of a handle's slabs for the TMA-staged kernel (use_tma = 0: cp.async kernel)
B2DescCache* b2_desc_cache_create();
void b2_desc_cache_destroy(B2DescCache* c);
void b2_launch_describe(const B2Geom& g, const B2Images& im, const uint32_t* sel, const int* sel_cnt,
                        const unsigned* ic_table, void* kps, uint8_t* desc, int* n_out, int nimg, B2DescCache* cache,
                        int use_tma, cudaStream_t s);
#define B2_IC_WORDS (4 * 16 * 9 * 2)   // IC_Angle weight table, describe.cu
void b2_launch_match(const B2MatchProblem* probs, int nprob, int max_nq, cudaStream_t s);
// Problem table of a stereo batch, written on the device so that re-binding never synchronises the host.
struct B2BindBases {
    uint8_t *desc, *cur_l, *cur_r;
    int *nkp, *train, *dist, *nmatch, *tti_l, *tdi_l, *tti_r, *tdi_r;
    int cap;
    const uint8_t *pool_l, *pool_r;   // non-null: temporal problems match against the feature pool (pool.cu)
    const int* pool_n;
};
void b2_launch_bind(B2MatchProblem* probs, int np, const B2BindBases& b, cudaStream_t s);
void b2_launch_stereo_filter(const void* kps, const uint8_t* desc, const int* n_kp, const int* train_idx,
                             const int* dist, int kp_cap, float ratio, float min_match, float max_dy, int* match_l,
                             int* match_r, int* n_match, uint8_t* cur_l, uint8_t* cur_r, double* pix_l, double* pix_r,
                             int npairs, cudaStream_t s);   // left image slot = p, right image slot = npairs + p;
                                                            // pix_* (optional) = kept pixels as cv::Point2d
// circular_kernel's optional tail for one-pair calls with temporal = 1: copy the pair's kept rows into the history slot
struct B2HistHandover {
    uint8_t* dst_l; uint8_t* dst_r; const uint8_t* src_l; const uint8_t* src_r; int* n_dst;
};
void b2_launch_circular(const int* ti_l, const int* di_l, const int* ti_r, const int* di_r, const int* n_cur,
                        const int* n_hist, int hist_stride, int cap, float ratio, float min_match, int* circ_cur,
                        int* circ_hist, int* n_circ, int* n_left, int* n_right, uint8_t* mask, int nprob, cudaStream_t s,
                        const B2HistHandover* handover = nullptr);
// ---- f3: device-resident feature pool (pool.cu) ---------------------------------------------------------------
struct B2Pool;
B2Pool* b2_pool_create();
void b2_pool_destroy(B2Pool* p);
void b2_pool_reset(B2Pool* p);
int b2_pool_size(const B2Pool* p);                      // rows of _histDescriptorsL/R
bool b2_pool_empty(const B2Pool* p);                    // _features.empty()
const uint8_t* b2_pool_rows(const B2Pool* p, int eye);  // device rows in pool order (may change after an update)
const int* b2_pool_dcount(const B2Pool* p);             // device copy of the row count
const unsigned long long* b2_pool_ids(const B2Pool* p); // host: _histFeatureIDs
int b2_pool_age(const B2Pool* p, unsigned long long id);
// featurePoolUpdate bookkeeping; descriptors move device-to-device only.  Returns 0 or a cudaError_t.
int b2_pool_update(B2Pool* p, cudaStream_t s, bool keyframe, int max_age, const unsigned long long* matched_ids, int n_matched,
                   const uint8_t* mask, const uint8_t* accept, int n_cur, const uint8_t* d_cur_l, const uint8_t* d_cur_r,
                   unsigned long long* new_ids, int* n_inserted, int* n_erased);
// cv::triangulatePoints operands (triangulate.cu): 3x4 row-major projection matrices, passed by value
struct B2Projection {
    double P1[12], P2[12];
    double max_depth;
};
void b2_launch_triangulate(const B2Projection& pr, const double* pix_l, const double* pix_r, int n, double* points4d,
                           cudaStream_t s);
void b2_launch_triangulate_matches(const B2Projection& pr, const void* kps, const int* match_l, const int* match_r,
                                   const int* n_match, int cap, int npairs, double* points4d, double* point_cam,
                                   uint8_t* flags, cudaStream_t s);
#define B2_OT_WORDS 26   // quadtree workspace words per node slot (octree.cu)
// allocation-free quadtree (b2orb_distribute_static): limits of the file-scope device workspace in octree.cu
#define B2_STATIC_MAX_NODES (B2ORB_STATIC_MAX_FEATURES + 4)
#define B2_STATIC_MAX_CANDIDATES B2ORB_STATIC_MAX_CANDIDATES
int b2_launch_octree_static(const B2Geom& g1, const uint32_t* d_cand, int n_cand, uint32_t* d_sel, int* d_sel_cnt, cudaStream_t s);
int b2_octree_scratch_words(const B2Geom& g);
unsigned long long b2_octree_fallbacks();   // per device, since the library was loaded
