// C-ABI layer (include/b2orb.h): handle, host-side tables of ORBextractor's constructor
// (src/ORBextractor.cc:410-470), per-size geometry (ComputePyramid sizes :1111-1112, FAST cell grid
// :773-787, quadtree roots :543-545), workspace, and the launch sequence of one batch.
// Host C++ only orchestrates; all pixel/descriptor work is in the .cu kernels next to this file.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstdio>
#include <cstdint>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "../../include/b2orb.h"
#include "b2orb_internal.cuh"

thread_local bool b2_pdl_on = false;

namespace {

constexpr int B2_PDL_MAX_IMAGES = 8;   // programmatic dependent launch for latency-sized batches only (b2orb_internal.cuh)

inline int cv_round_f(float v) { return (int)lrintf(v); }   // cvRound: round-half-even
inline unsigned align_up(unsigned v, unsigned a) { return (v + a - 1) / a * a; }


struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
};

}  // namespace

struct B2ExportSeg {
    const uint8_t* src;
    uint8_t* dst;
    const int* count;          // device counter array (indexed by pair * count_stride + count_off) or nullptr = fixed
    long long src_pair, dst_pair;   // bytes between consecutive pairs
    int elem, fixed, count_off, count_max;
};
constexpr int B2_EXPORT_MAX_SEG = 30;
struct B2ExportPlan {
    B2ExportSeg seg[B2_EXPORT_MAX_SEG];
    int nseg;
};

// Launch-time choices the library normally makes by batch size; tests and A/B runs pin them per handle with
// b2orb_debug_set (no environment variables, nothing process-global: two handles on two threads can differ).
struct B2Tuning {
    int pdl = -1;            // programmatic dependent launch: -1 = by batch size (<= 8 images), 0 / 1 = forced
    int fork_max = 8;        // blur on the side stream for calls of at most this many images
    int export_max = 4;      // direct export into pinned outputs for calls of at most this many pairs
    int fast_band = -1;      // FAST band variant 0 / 1 / 2 / 3 (16 / 32 / 64 / 128 rows), -1 = by batch size
    int blur_band = -1;      // blur band variant 0 / 1 (32 / 128 rows), -1 = by batch size
    int octree_points = 0;   // 1 = quadtree by point passes only (the count-pyramid path's fallback)
    int describe_tma = 1;    // 0 = descriptor windows staged by per-lane cp.async instead of TMA boxes
    int ingest_host = 1;     // 0 = one-pair calls copy pinned host images by DMA + repack instead of reading them in the kernel
    int pipe_chunk = 64;     // pairs per chunk of the pipelined host path, 0 = off
    int pipe_lanes = 3;      // lanes (workspaces + streams) the chunks rotate over
    int pipe_trace = 0;      // per-chunk timeline on stderr
};

struct b2orb_handle {
    b2orb_config cfg{};
    B2Tuning tune{};
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    std::string err;
    uint64_t launches = 0;
    float last_ms = 0.f;
    // direct export (latency path): segment table of the current call
    B2ExportPlan exp_plan;
    bool exp_ok = false;

    // ctor tables
    std::vector<float> scale, inv_scale, sigma2, inv_sigma2;
    std::vector<int> quota;
    int umax[16]{};

    // geometry of the current image size
    int cur_w = 0, cur_h = 0;
    B2Geom geom{};
    std::vector<int*> d_xofs, d_yofs;
    std::vector<short*> d_xw, d_yw;

    // workspace (sized for cfg.max_images images of the current size)
    int nimg_cap = 0, npair_cap = 0;
    uint8_t *d_stage = nullptr, *d_raw = nullptr, *d_pyr = nullptr, *d_blur = nullptr;
    size_t stage_img_bytes = 0;
    int stage_pitch = 0;                                 // level-0 rows in the staging area: 32-byte aligned pitch
    uint32_t* d_cand = nullptr;                          // voted FAST candidates
    uint2* d_raw_cand = nullptr;                         // raw corner-map words (fast.cu)
    int* d_raw_cnt = nullptr;                            // [nimg*nlevels] ints, then the per-cell iniThFAST flags
    uint8_t* d_cell_ini = nullptr;
    size_t raw_state_bytes = 0;
    unsigned short* d_pnode = nullptr;
    int *d_cand_cnt = nullptr, *d_sel_cnt = nullptr, *d_nkp = nullptr;
    uint32_t* d_sel = nullptr;
    unsigned* d_ot_scratch = nullptr;
    unsigned* d_ic_table = nullptr;                      // IC_Angle dp4a weights (describe.cu)
    b2orb_keypoint* d_kps = nullptr;
    uint8_t* d_desc = nullptr;
    int *d_train = nullptr, *d_dist = nullptr, *d_match_l = nullptr, *d_match_r = nullptr, *d_nmatch = nullptr;
    uint8_t *d_cur_l = nullptr, *d_cur_r = nullptr;      // slot 0 = history, slot p+1 = pair p
    int *d_tti_l = nullptr, *d_tdi_l = nullptr, *d_tti_r = nullptr, *d_tdi_r = nullptr;
    int *d_circ_cur = nullptr, *d_circ_hist = nullptr, *d_ncirc = nullptr, *d_nleft = nullptr, *d_nright = nullptr;
    B2MatchProblem* d_probs = nullptr;                   // [npair_cap stereo][2*npair_cap temporal][3 generic]
    // generic matcher scratch (b2orb_match / b2orb_temporal_track)
    uint8_t *d_gq = nullptr, *d_gt = nullptr, *d_gq2 = nullptr, *d_gt2 = nullptr;
    int *d_gidx = nullptr, *d_gdist = nullptr, *d_gidx2 = nullptr, *d_gdist2 = nullptr, *d_gcnt = nullptr;
    int g_cap_q = 0, g_cap_t = 0;
    int bound_np = -1;                                   // batch size the problem table is bound to
    // optional per-kernel timing (b2orb_profile_*): events recorded between launches on the handle's stream
    bool profiling = false;
    // latency path: the blur only needs the pyramid, so it runs on a side stream next to FAST -> vote -> quadtree
    cudaStream_t s_aux = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    std::vector<cudaEvent_t> prof_ev;                    // pool
    std::vector<int> prof_tag;                           // tag of the launch that FOLLOWS event i (-1 = none)
    size_t prof_used = 0;
    double prof_ms[B2ORB_NUM_KERNELS] = {};
    uint64_t prof_cnt[B2ORB_NUM_KERNELS] = {};
    std::vector<void*> owned;
    // pipelined host path (b2orb_stereo_frontend with many pairs): the batch is cut into chunks that rotate over
    // `lanes` (child handles with their own stream + workspace) so the H2D copy of one chunk, the kernels of
    // another and the D2H copy of a third overlap without any host thread.
    std::vector<b2orb_handle*> lanes;
    b2orb_handle* root = nullptr;                        // owning handle (self for a user-created handle)
    b2orb_handle* hist_owner = nullptr;                  // root only: handle whose slot 0 holds the newest history
    cudaEvent_t ev_hist = nullptr, ev_done = nullptr;    // history slot written / kernels of the lane's last chunk done
    cudaEvent_t ev_in = nullptr, ev_out = nullptr;       // lane: chunk copied in / results copied out
    cudaStream_t s_in = nullptr, s_out = nullptr;        // root: copy-in and copy-out streams of the pipelined path
    // rectification maps of b2orb_set_rectification (root handle; lanes read the root's): [eye] map1 = short2, map2 = u16
    void* d_map1[2] = {nullptr, nullptr};
    void* d_map2[2] = {nullptr, nullptr};
    int rect_w = 0, rect_h = 0;                          // 0 = rectification off
    // packed colour frames of the cfsd ingest (b2orb_cfsd_ingest / b2orb_stereo_frontend_packed)
    uint8_t* d_packed = nullptr;
    size_t packed_bytes = 0;
    int ing_key[4] = {0, 0, 0, 0};                       // (src_w, src_h, dst_w, dst_h) the resize tables were built for
    int *d_ing_xofs = nullptr, *d_ing_yofs = nullptr;
    short *d_ing_xw = nullptr, *d_ing_yw = nullptr;
    // next row f4: cv::triangulatePoints of the kept stereo matches (b2orb_set_projection; root handle, lanes read it)
    B2Projection proj{};
    bool proj_on = false;
    double *d_pts4 = nullptr, *d_ptcam = nullptr;        // [npair_cap][4][cap] homogeneous points, [npair_cap][cap][3]
    uint8_t* d_dflag = nullptr;                          // [npair_cap][cap] depth-gate flags
    double *d_pix_l = nullptr, *d_pix_r = nullptr;       // [npair_cap][cap][2] kept pixels (_curPixelsL/R as cv::Point2d)
    uint8_t* d_mask = nullptr;                           // [npair_cap][cap] _curFeatureMask
    // next row f3: device-resident feature pool (root handle) + host copies of the last pool track
    B2Pool* pool = nullptr;
    B2DescCache* desc_cache = nullptr;   // tensor maps of this handle's slabs (describe.cu), created on first use
    int pool_max_age = 8;
    std::vector<int> trk_circ_hist;                      // trainIdx of the circular matches of the last pool track
    std::vector<unsigned long long> trk_ids;             // ... and the feature ids behind them (_matchedFeatureIDs)
    std::vector<uint8_t> trk_mask, trk_flags;            // _curFeatureMask / depth flags of the last pool track
    int trk_ncur = -1, trk_pair = 0;                     // -1 = lists of the last track not fetched yet
    b2orb_handle* trk_owner = nullptr;                   // handle (lane) whose workspace holds the tracked pair
    int bound_pool = 0;                                  // problem table bound for pool mode (and to which buffers)
    const uint8_t* bound_pool_rows = nullptr;
    // b2orb_extract_async: pinned count the device writes, state of the call in flight
    int32_t* h_async_n = nullptr;
    int async_count = 0;
    double *d_tri_in = nullptr, *d_tri_out = nullptr;    // b2orb_triangulate scratch (grow-only)
    int tri_cap = 0;
};

namespace {

bool cuda_ok(b2orb_handle* h, cudaError_t e, const char* what) {
    if (e == cudaSuccess) return true;
    char buf[256];
    snprintf(buf, sizeof buf, "%s: %s", what, cudaGetErrorString(e));
    h->err = buf;
    return false;
}
#define B2_CUDA(h, call)                                         \
    do {                                                         \
        if (!cuda_ok((h), (call), #call)) return B2ORB_ERR_CUDA; \
    } while (0)

inline bool pdl_for(const b2orb_handle* h, int nimg) {
    const int forced = h->root->tune.pdl;
    return forced >= 0 ? forced != 0 : nimg <= B2_PDL_MAX_IMAGES;
}

int fail(b2orb_handle* h, int code, const char* msg) {
    if (h) h->err = msg;
    return code;
}

// Profiling marker: an event before the launch tagged `tag`; prof_end() closes the last interval.
void prof_mark(b2orb_handle* h, int tag) {
    if (!h->profiling) return;
    if (h->prof_used == h->prof_ev.size()) {
        cudaEvent_t e;
        cudaEventCreate(&e);
        h->prof_ev.push_back(e);
        h->prof_tag.push_back(-1);
    }
    h->prof_tag[h->prof_used] = tag;
    cudaEventRecord(h->prof_ev[h->prof_used++], h->stream);
}

template <class T>
bool dev_alloc(b2orb_handle* h, T** p, size_t count) {
    void* q = nullptr;
    if (!cuda_ok(h, cudaMalloc(&q, count * sizeof(T) + 256), "cudaMalloc")) return false;
    cudaMemsetAsync(q, 0, count * sizeof(T) + 256, h->stream);
    h->owned.push_back(q);
    *p = static_cast<T*>(q);
    return true;
}

void free_workspace(b2orb_handle* h) {
    for (void* p : h->owned) cudaFree(p);
    h->owned.clear();
    h->d_xofs.clear(); h->d_yofs.clear(); h->d_xw.clear(); h->d_yw.clear();
    h->g_cap_q = h->g_cap_t = 0;
    h->bound_np = -1;
}

// ORBextractor::ORBextractor, src/ORBextractor.cc:410-470.
void build_ctor_tables(b2orb_handle* h) {
    const b2orb_config& c = h->cfg;
    const int nl = c.nlevels;
    const double sf = (double)c.scale_factor;   // member `scaleFactor` is a double initialised from the float arg
    h->scale.assign(nl, 1.f); h->sigma2.assign(nl, 1.f); h->inv_scale.assign(nl, 1.f); h->inv_sigma2.assign(nl, 1.f);
    for (int i = 1; i < nl; ++i) {
        h->scale[i] = (float)(h->scale[i - 1] * sf);
        h->sigma2[i] = h->scale[i] * h->scale[i];
    }
    for (int i = 0; i < nl; ++i) {
        h->inv_scale[i] = 1.0f / h->scale[i];
        h->inv_sigma2[i] = 1.0f / h->sigma2[i];
    }
    h->quota.assign(nl, 0);
    const float factor = (float)(1.0f / sf);
    float desired = c.nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nl));
    int sum = 0;
    for (int l = 0; l < nl - 1; ++l) {
        h->quota[l] = cv_round_f(desired);
        sum += h->quota[l];
        desired *= factor;
    }
    h->quota[nl - 1] = std::max(c.nfeatures - sum, 0);
    // umax, :454-469
    int v, v0;
    const int vmax = (int)std::floor(B2_HALF_PATCH * std::sqrt(2.f) / 2 + 1);
    const int vmin = (int)std::ceil(B2_HALF_PATCH * std::sqrt(2.f) / 2);
    const double hp2 = B2_HALF_PATCH * B2_HALF_PATCH;
    for (v = 0; v <= vmax; ++v) h->umax[v] = (int)lrint(std::sqrt(hp2 - v * v));
    for (v = B2_HALF_PATCH, v0 = 0; v >= vmin; --v) {
        while (h->umax[v0] == h->umax[v0 + 1]) ++v0;
        h->umax[v] = v0;
        ++v0;
    }
}

// cv::resize's per-axis tables (imgproc/resize.cpp): float source coordinate from a double scale,
// Q11 weights via saturate_cast<short>(w * 2048) (round-half-even).
void resize_axis_table(int src, int dst, std::vector<int>& ofs, std::vector<short>& w) {
    ofs.resize(dst);
    w.resize((size_t)2 * dst);
    const double scale = (double)src / dst;
    for (int d = 0; d < dst; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)std::floor(f);
        f -= s;
        if (s < 0) { s = 0; f = 0.f; }
        if (s >= src - 1) { s = src - 1; f = 0.f; }
        ofs[d] = s;
        w[2 * d] = (short)lrintf((1.f - f) * 2048.f);
        w[2 * d + 1] = (short)lrintf(f * 2048.f);
    }
}

// Geometry for one image size.  Returns an error string or nullptr.
const char* build_geometry(b2orb_handle* h, int w, int hgt, B2Geom& g) {
    const b2orb_config& c = h->cfg;
    memset(&g, 0, sizeof g);
    g.nlevels = c.nlevels;
    g.ini_th = c.ini_th_fast;
    g.min_th = c.min_th_fast;
    memcpy(g.umax, h->umax, sizeof g.umax);
    static const int def_taps[7] = {18, 34, 48, 56, 48, 34, 18};
    bool zero = true;
    for (int i = 0; i < 7; ++i) zero = zero && c.blur_taps[i] == 0;
    int tap_sum = 0;
    for (int i = 0; i < 7; ++i) { g.taps[i] = zero ? def_taps[i] : c.blur_taps[i]; tap_sum += g.taps[i]; }
    g.tap_sat = tap_sum > 256;
    {   // blur.cu: output pixel j of a lane sees columns x+j-3 .. x+j+3 of the words W0|W1|W2 = columns x-4 .. x+7
        const int k0 = g.taps[0], k1 = g.taps[1], k2 = g.taps[2], k3 = g.taps[3];
        auto tv = [](int a, int b, int c2, int d) { return (unsigned)a | ((unsigned)b << 8) | ((unsigned)c2 << 16) | ((unsigned)d << 24); };
        const unsigned v[10] = {tv(0, k0, k1, k2), tv(k3, k2, k1, k0), tv(0, 0, k0, k1), tv(k2, k3, k2, k1), tv(k0, 0, 0, 0),
                                tv(0, 0, 0, k0), tv(k1, k2, k3, k2), tv(k1, k0, 0, 0), tv(k0, k1, k2, k3), tv(k2, k1, k0, 0)};
        memcpy(g.tapv, v, sizeof v);
    }
    unsigned off = 0, cand_off = 0, node_off = 0;
    int cells = 0, tiles[B2_BLUR_VARIANTS] = {}, sel = 0, ftasks[B2_FAST_VARIANTS] = {};
    for (int l = 0; l < c.nlevels; ++l) {
        B2Level& L = g.lv[l];
        // ComputePyramid :1111-1112: Size(cvRound((float)cols*scale), cvRound((float)rows*scale))
        L.w = cv_round_f((float)w * h->inv_scale[l]);
        L.h = cv_round_f((float)hgt * h->inv_scale[l]);
        if (L.w < 1 || L.h < 1) return "pyramid level collapses to zero size";
        L.pitch = (int)align_up((unsigned)L.w, 32);
        L.off = off;
        off += align_up((unsigned)L.pitch * (unsigned)L.h, 256);
        L.blur_strips = (L.w + 127) / 128;
        for (int v = 0; v < B2_BLUR_VARIANTS; ++v) {
            const int band = B2_BLUR_BAND << (2 * v);
            L.blur_task_base[v] = tiles[v];
            tiles[v] += L.blur_strips * ((L.h + band - 1) / band);
        }
        L.quota = h->quota[l];
        L.scale = h->scale[l];
        L.kp_size = (float)(int)(31 * h->scale[l]);     // :836 `const int scaledPatchSize = PATCH_SIZE*mvScaleFactor[level]`
        // ComputeKeyPointsOctTree :773-787
        const int minB = B2_BORDER, maxBX = L.w - B2_BORDER, maxBY = L.h - B2_BORDER;
        const float width = (float)(maxBX - minB), height = (float)(maxBY - minB);
        const int n_cols = width > 0 ? (int)(width / 30.f) : 0, n_rows = height > 0 ? (int)(height / 30.f) : 0;
        L.cell_base = cells;
        for (int v = 0; v < B2_FAST_VARIANTS; ++v)
            L.fast_task_base[v] = ftasks[v];   // levels without cells own zero tasks but must keep the list monotonic
        L.n_ini = 1;
        L.h_x = 1.f;
        if (n_cols >= 1 && n_rows >= 1) {   // else: the reference divides by zero; the level yields no keypoints
            L.n_cols = n_cols;
            L.n_rows = n_rows;
            L.w_cell = (int)std::ceil(width / n_cols);
            L.h_cell = (int)std::ceil(height / n_rows);
            if (L.w_cell > 64 || L.h_cell > 64) return "FAST cell larger than 64 px";
            if ((long long)n_cols * n_rows * L.w_cell * L.h_cell >= (1 << 24)) return "level too large for 24-bit order keys";
            cells += n_cols * n_rows;
            L.fast_strips = (L.w - 2 * B2_BORDER + 119) / 120;
            for (int v = 0; v < B2_FAST_VARIANTS; ++v)
                ftasks[v] += L.fast_strips * ((L.h - 2 * B2_EDGE + (B2_FAST_BAND << v) - 1) / (B2_FAST_BAND << v));
            // DistributeOctTree :543-545
            L.n_ini = (int)std::round((float)(maxBX - minB) / (maxBY - minB));
            if (L.n_ini < 1) {   // taller than 2:1 -> the reference divides by zero; no keypoints
                L.n_ini = 1; L.n_cols = L.n_rows = 0; cells -= n_cols * n_rows;
                for (int v = 0; v < B2_FAST_VARIANTS; ++v) ftasks[v] = L.fast_task_base[v];
                L.fast_strips = 0;
            } else if (L.n_ini > 256) return "aspect ratio above 256:1";
            L.h_x = (float)(maxBX - minB) / L.n_ini;
        }
        L.cand_cap = L.n_cols * L.n_rows * ((L.w_cell + 1) / 2) * ((L.h_cell + 1) / 2);
        L.cand_off = cand_off;
        cand_off += align_up((unsigned)L.cand_cap, 64);
        L.node_cap = std::max(L.quota + 3, 4 * L.n_ini) + 1;
        if (L.node_cap > 60000) return "per-level feature quota too large";
        L.node_off = node_off;
        node_off += L.node_cap;
        L.sel_cap = L.node_cap;
        {   // count pyramid of the quadtree (octree.cu): deepest D <= 7 whose n_ini*(4^(D+1)-1)/3 counters fit into the
            // 13*node_cap workspace words the point-pass path uses for its child counts
            int D = 0;
            while (D < 7 && (long long)L.n_ini * (((1LL << (2 * (D + 2))) - 1) / 3) <= 13LL * L.node_cap) ++D;
            L.ot_depth = (long long)L.n_ini * (((1LL << (2 * (D + 1))) - 1) / 3) <= 13LL * L.node_cap ? D : 0;
        }
        L.sel_base = sel;
        sel += L.sel_cap;
    }
    if (w > 4096 + B2_BORDER || hgt > 4096 + B2_BORDER) return "image larger than 4096 px";
    g.total_cells = cells;
    for (int v = 0; v < B2_BLUR_VARIANTS; ++v) g.total_blur_tasks[v] = tiles[v];
    for (int v = 0; v < B2_FAST_VARIANTS; ++v) g.total_fast_tasks[v] = ftasks[v];
    g.sel_total = sel;
    g.kp_cap = sel;
    g.cand_total = cand_off;
    g.node_total = node_off;
    g.slab = align_up(off, 256);
    return nullptr;
}

}  // namespace

namespace {

// (Re)build geometry, resize tables and workspace for an image size.  Steady state: no-op.
int configure(b2orb_handle* h, int w, int hgt) {
    if (w == h->cur_w && hgt == h->cur_h) return B2ORB_OK;
    B2Geom g;
    if (const char* e = build_geometry(h, w, hgt, g)) return fail(h, B2ORB_ERR_INVALID, e);
    cudaStreamSynchronize(h->stream);
    free_workspace(h);
    h->cur_w = h->cur_h = 0;
    h->geom = g;
    const int nl = g.nlevels, nimg = h->cfg.max_images, np = std::max(nimg / 2, 1);
    h->nimg_cap = nimg;
    h->npair_cap = np;
    // resize tables per level (level l from level l-1)
    h->d_xofs.assign(nl, nullptr); h->d_yofs.assign(nl, nullptr); h->d_xw.assign(nl, nullptr); h->d_yw.assign(nl, nullptr);
    for (int l = 1; l < nl; ++l) {
        std::vector<int> xo, yo;
        std::vector<short> xw, yw;
        resize_axis_table(g.lv[l - 1].w, g.lv[l].w, xo, xw);
        resize_axis_table(g.lv[l - 1].h, g.lv[l].h, yo, yw);
        if (!dev_alloc(h, &h->d_xofs[l], xo.size()) || !dev_alloc(h, &h->d_yofs[l], yo.size()) ||
            !dev_alloc(h, &h->d_xw[l], xw.size()) || !dev_alloc(h, &h->d_yw[l], yw.size()))
            return B2ORB_ERR_NOMEM;
        B2_CUDA(h, cudaMemcpyAsync(h->d_xofs[l], xo.data(), xo.size() * 4, cudaMemcpyHostToDevice, h->stream));
        B2_CUDA(h, cudaMemcpyAsync(h->d_yofs[l], yo.data(), yo.size() * 4, cudaMemcpyHostToDevice, h->stream));
        B2_CUDA(h, cudaMemcpyAsync(h->d_xw[l], xw.data(), xw.size() * 2, cudaMemcpyHostToDevice, h->stream));
        B2_CUDA(h, cudaMemcpyAsync(h->d_yw[l], yw.data(), yw.size() * 2, cudaMemcpyHostToDevice, h->stream));
        B2_CUDA(h, cudaStreamSynchronize(h->stream));   // host vectors die at scope end
    }
    {   // IC_Angle weight words (describe.cu): row |v| of the radius-15 disc, read as 9 aligned words starting at
        // column (cx-15) & ~3; byte i of word k is u = 4k + i - 15 - a, a = (cx-15) & 3.  wu = u inside the disc
        // (|u| <= umax[|v|], src/ORBextractor.cc:83-100), wm = 1 inside the disc.
        std::vector<unsigned> tab(B2_IC_WORDS);
        for (int a = 0; a < 4; ++a)
            for (int dd = 0; dd < 16; ++dd)
                for (int k = 0; k < 9; ++k) {
                    unsigned wu = 0, wm = 0;
                    for (int i = 0; i < 4; ++i) {
                        const int u = 4 * k + i - 15 - a;
                        if (std::abs(u) <= h->umax[dd]) {
                            wu |= (unsigned)(uint8_t)(int8_t)u << (8 * i);
                            wm |= 1u << (8 * i);
                        }
                    }
                    tab[((a * 16 + dd) * 9 + k) * 2] = wu;
                    tab[((a * 16 + dd) * 9 + k) * 2 + 1] = wm;
                }
        if (!dev_alloc(h, &h->d_ic_table, tab.size())) return B2ORB_ERR_NOMEM;
        B2_CUDA(h, cudaMemcpyAsync(h->d_ic_table, tab.data(), tab.size() * 4, cudaMemcpyHostToDevice, h->stream));
        B2_CUDA(h, cudaStreamSynchronize(h->stream));
    }
    h->stage_pitch = (int)align_up((unsigned)w, 32);
    h->stage_img_bytes = (size_t)h->stage_pitch * hgt;
    const size_t cap = g.kp_cap;
    bool ok = dev_alloc(h, &h->d_stage, (size_t)nimg * h->stage_img_bytes) &&
              dev_alloc(h, &h->d_raw, (size_t)nimg * w * hgt) &&
              dev_alloc(h, &h->d_pyr, (size_t)nimg * g.slab) && dev_alloc(h, &h->d_blur, (size_t)nimg * g.slab) &&
              dev_alloc(h, &h->d_raw_cand, (size_t)nimg * g.cand_total) &&
              dev_alloc(h, &h->d_cand, (size_t)nimg * g.cand_total) &&
              dev_alloc(h, &h->d_pnode, (size_t)nimg * g.cand_total) &&
              dev_alloc(h, &h->d_sel_cnt, (size_t)nimg * nl) &&
              dev_alloc(h, &h->d_nkp, (size_t)nimg) && dev_alloc(h, &h->d_sel, (size_t)nimg * g.sel_total) &&
              dev_alloc(h, &h->d_kps, (size_t)nimg * cap) && dev_alloc(h, &h->d_desc, (size_t)nimg * cap * 32) &&
              dev_alloc(h, &h->d_train, (size_t)np * cap) && dev_alloc(h, &h->d_dist, (size_t)np * cap) &&
              dev_alloc(h, &h->d_match_l, (size_t)np * cap) && dev_alloc(h, &h->d_match_r, (size_t)np * cap) &&
              dev_alloc(h, &h->d_nmatch, (size_t)np + 1) && dev_alloc(h, &h->d_cur_l, (size_t)(np + 1) * cap * 32) &&
              dev_alloc(h, &h->d_cur_r, (size_t)(np + 1) * cap * 32) && dev_alloc(h, &h->d_tti_l, (size_t)np * cap) &&
              dev_alloc(h, &h->d_tdi_l, (size_t)np * cap) && dev_alloc(h, &h->d_tti_r, (size_t)np * cap) &&
              dev_alloc(h, &h->d_tdi_r, (size_t)np * cap) && dev_alloc(h, &h->d_circ_cur, (size_t)np * cap) &&
              dev_alloc(h, &h->d_circ_hist, (size_t)np * cap) && dev_alloc(h, &h->d_ncirc, (size_t)np) &&
              dev_alloc(h, &h->d_nleft, (size_t)np) && dev_alloc(h, &h->d_nright, (size_t)np) &&
              dev_alloc(h, &h->d_probs, (size_t)3 * np + 4) && dev_alloc(h, &h->d_gcnt, (size_t)8) &&
              dev_alloc(h, &h->d_pts4, (size_t)np * cap * 4) && dev_alloc(h, &h->d_ptcam, (size_t)np * cap * 3) &&
              dev_alloc(h, &h->d_dflag, (size_t)np * cap) && dev_alloc(h, &h->d_pix_l, (size_t)np * cap * 2) &&
              dev_alloc(h, &h->d_pix_r, (size_t)np * cap * 2) && dev_alloc(h, &h->d_mask, (size_t)np * cap);
    if (!ok) return B2ORB_ERR_NOMEM;
    {   // raw candidate counters and per-cell flags share one allocation: one memset per batch clears both
        const size_t cnt_bytes = (size_t)nimg * nl * sizeof(int);
        h->raw_state_bytes = 2 * cnt_bytes + (size_t)nimg * std::max(g.total_cells, 1);
        uint8_t* p = nullptr;
        if (!dev_alloc(h, &p, h->raw_state_bytes)) return B2ORB_ERR_NOMEM;
        h->d_raw_cnt = reinterpret_cast<int*>(p);
        h->d_cand_cnt = reinterpret_cast<int*>(p + cnt_bytes);
        h->d_cell_ini = p + 2 * cnt_bytes;
    }
    if ((size_t)b2_octree_scratch_words(g) * 4 > 200 * 1024) {   // quadtree state too big for shared memory
        if (!dev_alloc(h, &h->d_ot_scratch, (size_t)nimg * g.node_total * B2_OT_WORDS)) return B2ORB_ERR_NOMEM;
    } else h->d_ot_scratch = nullptr;
    // match problem table: [0,np) stereo pair p (left slot p vs right slot n+p is bound per call because the
    // right slot depends on the batch size) — filled by bind_problems().
    B2_CUDA(h, cudaStreamSynchronize(h->stream));
    h->cur_w = w;
    h->cur_h = hgt;
    return B2ORB_OK;
}

B2Pool* pool_of(b2orb_handle* h) {   // the feature pool lives on the root handle, created on first use
    b2orb_handle* r = h->root;
    if (!r->pool) r->pool = b2_pool_create();
    return r->pool;
}

// Stereo problems [0,np): desc(left p) vs desc(right np+p).  Temporal problems [np, 3np): cur slot p+1 vs
// hist slot p (or vs the feature pool), left then right.
int bind_problems(b2orb_handle* h, int np, int pool_mode = 0) {
    B2Pool* pool = pool_mode ? pool_of(h) : nullptr;
    if (pool_mode && !pool) return fail(h, B2ORB_ERR_NOMEM, "feature pool");
    const uint8_t* rows = pool ? b2_pool_rows(pool, 0) : nullptr;
    if (h->bound_np == np && h->bound_pool == pool_mode && h->bound_pool_rows == rows) return B2ORB_OK;
    B2BindBases b{h->d_desc, h->d_cur_l, h->d_cur_r, h->d_nkp, h->d_train, h->d_dist, h->d_nmatch,
                  h->d_tti_l, h->d_tdi_l, h->d_tti_r, h->d_tdi_r, h->geom.kp_cap,
                  rows, pool ? b2_pool_rows(pool, 1) : nullptr, pool ? b2_pool_dcount(pool) : nullptr};
    if (pool && !rows) {   // empty pool, nothing allocated yet: any valid pointer will do, the count is 0
        b.pool_l = h->d_cur_l; b.pool_r = h->d_cur_r;
    }
    h->bound_pool = pool_mode;
    h->bound_pool_rows = rows;
    b2_launch_bind(h->d_probs, np, b, h->stream);       // stream-ordered: no host synchronisation
    ++h->launches;
    h->bound_np = np;
    return B2ORB_OK;
}

// Launch sequence of ORBextractor::operator() for `nimg` resident images (src/ORBextractor.cc:1043-1105).
int launch_extract(b2orb_handle* h, const B2Images& im, int nimg) {
    const B2Geom& g = h->geom;
    cudaStream_t s = h->stream;
    const B2Tuning& tn = h->root->tune;
    b2_pdl_on = pdl_for(h, nimg);
    // counters first: nothing between the resize chain and the FAST kernel breaks the programmatic launch chain
    B2_CUDA(h, cudaMemsetAsync(h->d_raw_cnt, 0, h->raw_state_bytes, s));
    // (Two single-launch versions of the resize chain were measured and removed: one thread-block cluster per image with
    // barrier.cluster between levels took 38 us for a KITTI pair, a band-per-CTA version 35 us, against ~18 us for the
    // seven programmatically chained launches below.)
    for (int l = 1; l < g.nlevels; ++l) {
        prof_mark(h, B2ORB_K_RESIZE);
        b2_launch_resize(g, im, l, h->d_xofs[l], h->d_xw[l], h->d_yofs[l], h->d_yw[l], nimg, s);
        ++h->launches;
    }
    prof_mark(h, -1);
    // blur on the side stream for latency-sized calls (not while profiling per kernel)
    const bool fork = nimg <= tn.fork_max && !h->profiling && h->s_aux;
    if (fork) {
        B2_CUDA(h, cudaEventRecord(h->ev_fork, s));
        B2_CUDA(h, cudaStreamWaitEvent(h->s_aux, h->ev_fork, 0));
        b2_launch_blur(g, im, nimg, tn.blur_band, h->s_aux);
        ++h->launches;
        B2_CUDA(h, cudaEventRecord(h->ev_join, h->s_aux));
    }
    prof_mark(h, B2ORB_K_FAST);
    if (g.total_cells > 0) {
        b2_launch_fast_score(g, im, h->d_raw_cand, h->d_raw_cnt, h->d_cell_ini, nimg, tn.fast_band, s);
        prof_mark(h, B2ORB_K_FAST_CELLS);
        b2_launch_fast_vote(g, h->d_raw_cand, h->d_raw_cnt, h->d_cell_ini, h->d_cand, h->d_cand_cnt, nimg, s);
        h->launches += 2;
    }
    prof_mark(h, B2ORB_K_OCTREE);
    b2_launch_octree(g, h->d_cand, h->d_cand_cnt, h->d_pnode, h->d_ot_scratch, h->d_sel, h->d_sel_cnt, nimg, tn.octree_points, s);
    ++h->launches;
    if (fork) {
        B2_CUDA(h, cudaStreamWaitEvent(s, h->ev_join, 0));
    } else {
        prof_mark(h, B2ORB_K_BLUR);
        b2_launch_blur(g, im, nimg, tn.blur_band, s);
        ++h->launches;
    }
    prof_mark(h, B2ORB_K_DESCRIBE);
    if (!h->desc_cache) h->desc_cache = b2_desc_cache_create();
    b2_launch_describe(g, im, h->d_sel, h->d_sel_cnt, h->d_ic_table, h->d_kps, h->d_desc, h->d_nkp, nimg, h->desc_cache, tn.describe_tma, s);
    ++h->launches;
    prof_mark(h, -1);
    B2_CUDA(h, cudaGetLastError());
    return B2ORB_OK;
}

B2Images make_images(b2orb_handle* h, const uint8_t* a, const uint8_t* b, int split, size_t stride0, int pitch0) {
    B2Images im;
    im.lvl0[0] = a;
    im.lvl0[1] = b;
    im.split = split;
    im.stride0 = stride0;
    im.pitch0 = pitch0;
    im.pyr = h->d_pyr;
    im.blur = h->d_blur;
    return im;
}

// Host (or device) images -> staging area with an aligned row pitch (kernels read level 0 with 32-bit / 128-bit
// accesses).  Host images whose rows are already aligned go straight in; tight rows of odd width (1241 ...) are
// copied linearly (a pitched 2-D DMA of odd-width rows runs at half the PCIe rate) and re-pitched by a kernel.
enum UploadPath { UP_DEVICE, UP_PITCHED, UP_TIGHT, UP_ROWS };
UploadPath upload_path(const b2orb_handle* h, int w, int hgt, int step, size_t image_stride, bool on_device) {
    if (on_device) return UP_DEVICE;
    if (step == h->stage_pitch && image_stride == (size_t)step * hgt) return UP_PITCHED;
    if (step == w && image_stride == (size_t)w * hgt) return UP_TIGHT;
    return UP_ROWS;
}
// Rectification (b2orb_set_rectification) turns the ingest step into cv::remap: raw images are staged as tight rows
// in d_raw and remap_kernel writes the rectified image straight into the pitched staging area.  eye: -1 = images are
// already rectified, 0 / 1 = left / right maps.
bool rectifying(const b2orb_handle* h, int w, int hgt, int eye) {
    return eye >= 0 && h->root->rect_w == w && h->root->rect_h == hgt;
}
// the host->device part, on stream `cs` (the handle's own stream, or the pipelined path's copy-in stream)
int upload_copy(b2orb_handle* h, const uint8_t* src, int count, int w, int hgt, int step, size_t image_stride, int slot0,
                cudaStream_t cs, int eye = -1) {
    uint8_t* dst = h->d_stage + (size_t)slot0 * h->stage_img_bytes;
    uint8_t* raw = h->d_raw + (size_t)slot0 * w * hgt;
    const UploadPath path = upload_path(h, w, hgt, step, image_stride, false);
    if (path == UP_TIGHT) {
        B2_CUDA(h, cudaMemcpyAsync(raw, src, (size_t)count * w * hgt, cudaMemcpyHostToDevice, cs));
    } else if (rectifying(h, w, hgt, eye)) {
        for (int i = 0; i < count; ++i)
            B2_CUDA(h, cudaMemcpy2DAsync(raw + (size_t)i * w * hgt, w, src + i * image_stride, step, w, hgt,
                                         cudaMemcpyHostToDevice, cs));
    } else if (path == UP_PITCHED) {
        B2_CUDA(h, cudaMemcpyAsync(dst, src, (size_t)count * step * hgt, cudaMemcpyHostToDevice, cs));
    } else {
        for (int i = 0; i < count; ++i)
            B2_CUDA(h, cudaMemcpy2DAsync(dst + (size_t)i * h->stage_img_bytes, h->stage_pitch, src + i * image_stride, step, w,
                                         hgt, cudaMemcpyHostToDevice, cs));
    }
    return B2ORB_OK;
}
// the kernel part of the ingest (re-pitching, or rectification), on the handle's stream
int upload_repack(b2orb_handle* h, const uint8_t* src, int count, int w, int hgt, int step, size_t image_stride, int slot0,
                  bool on_device, int eye = -1) {
    const UploadPath path = upload_path(h, w, hgt, step, image_stride, on_device);
    const bool rect = rectifying(h, w, hgt, eye);
    if (!rect && path != UP_DEVICE && path != UP_TIGHT) return B2ORB_OK;
    b2_pdl_on = pdl_for(h, 2 * count);
    uint8_t* dst = h->d_stage + (size_t)slot0 * h->stage_img_bytes;
    const uint8_t* from = path == UP_DEVICE ? src : h->d_raw + (size_t)slot0 * w * hgt;
    const int fstep = path == UP_DEVICE ? step : w;
    const size_t fstride = path == UP_DEVICE ? image_stride : (size_t)w * hgt;
    prof_mark(h, B2ORB_K_INGEST);
    if (rect)
        b2_launch_remap(from, fstep, fstride, h->root->d_map1[eye], h->root->d_map2[eye], dst, h->stage_pitch,
                        h->stage_img_bytes, w, hgt, count, h->stream);
    else
        b2_launch_repack(from, fstep, fstride, dst, h->stage_pitch, h->stage_img_bytes, w, hgt, count, h->stream);
    ++h->launches;
    prof_mark(h, -1);
    return B2ORB_OK;
}
// Device-visible address of a pinned host buffer (cudaHostAlloc / cudaHostRegister / b2orb_host_alloc), else nullptr.
const uint8_t* pinned_devptr(const void* host) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, host) != cudaSuccess) { (void)cudaGetLastError(); return nullptr; }
    if (a.type != cudaMemoryTypeHost || !a.devicePointer) return nullptr;
    return static_cast<const uint8_t*>(a.devicePointer);
}
// Latency path: one image per eye, pinned host memory, no rectification -> ONE kernel reads the image(s) over the bus and
// writes the pitched staging slots slot0 (and slot0 + 1 for src_r).  false = not applicable, use upload().
bool upload_zero_copy(b2orb_handle* h, const uint8_t* src_l, const uint8_t* src_r, int w, int hgt, int step, int slot0) {
    if (!h->root->tune.ingest_host || rectifying(h, w, hgt, 0)) return false;
    // rows that already have the staging pitch go straight in by DMA: nothing to re-pitch, nothing to gain
    if (upload_path(h, w, hgt, step, (size_t)step * hgt, false) == UP_PITCHED) return false;
    const uint8_t* dl = pinned_devptr(src_l);
    const uint8_t* dr = src_r ? pinned_devptr(src_r) : nullptr;
    if (!dl || (src_r && !dr)) return false;
    b2_pdl_on = pdl_for(h, 2);
    prof_mark(h, B2ORB_K_INGEST);
    const bool ok = b2_launch_ingest_host(dl, dr, step, h->d_stage + (size_t)slot0 * h->stage_img_bytes, h->stage_pitch,
                                          h->stage_img_bytes, w, hgt, h->stream);
    if (ok) ++h->launches;
    prof_mark(h, -1);
    return ok;
}
int upload(b2orb_handle* h, const uint8_t* src, int count, int w, int hgt, int step, size_t image_stride, int slot0,
           bool on_device = false, int eye = -1) {
    if (!on_device && count == 1 && eye < 0 && upload_zero_copy(h, src, nullptr, w, hgt, step, slot0)) return B2ORB_OK;
    if (!on_device)
        if (int rc = upload_copy(h, src, count, w, hgt, step, image_stride, slot0, h->stream, eye)) return rc;
    return upload_repack(h, src, count, w, hgt, step, image_stride, slot0, on_device, eye);
}

int launch_stereo(b2orb_handle* h, const B2Images& im, int np, int temporal) {
    const B2Geom& g = h->geom;
    cudaStream_t s = h->stream;
    int rc = launch_extract(h, im, 2 * np);
    if (rc) return rc;
    rc = bind_problems(h, np, temporal == 2);   // stream-ordered kernel; the table depends on np and the history source
    if (rc) return rc;
    prof_mark(h, B2ORB_K_MATCH_STEREO);
    b2_launch_match(h->d_probs, np, g.kp_cap, s);
    ++h->launches;
    prof_mark(h, B2ORB_K_STEREO_FILTER);
    b2_launch_stereo_filter(h->d_kps, h->d_desc, h->d_nkp, h->d_train, h->d_dist, g.kp_cap, h->cfg.match_ratio,
                            h->cfg.min_match_dist, h->cfg.max_vertical_pixel_dist, h->d_match_l, h->d_match_r,
                            h->d_nmatch + 1, h->d_cur_l + (size_t)g.kp_cap * 32, h->d_cur_r + (size_t)g.kp_cap * 32,
                            h->d_pix_l, h->d_pix_r, np, s);
    ++h->launches;
    if (h->root->proj_on) {   // f4: triangulate the kept matches where they lie (src/feature-tracker.cpp:365-379)
        prof_mark(h, B2ORB_K_TRIANGULATE);
        b2_launch_triangulate_matches(h->root->proj, h->d_kps, h->d_match_l, h->d_match_r, h->d_nmatch + 1, g.kp_cap, np,
                                      h->d_pts4, h->d_ptcam, h->d_dflag, s);
        ++h->launches;
    }
    prof_mark(h, -1);
    if (temporal == 2) {   // f3: every pair against the device-resident feature pool (externalTrack :258, :303)
        B2Pool* pool = h->root->pool;
        h->root->trk_ncur = -1;          // the lists of this call are fetched on demand (b2orb_pool_last_track / _update)
        h->root->trk_owner = h;
        h->root->trk_pair = 0;
        prof_mark(h, B2ORB_K_MATCH_TEMPORAL);
        b2_launch_match(h->d_probs + np, 2 * np, g.kp_cap, s);
        ++h->launches;
        prof_mark(h, B2ORB_K_CIRCULAR);
        b2_launch_circular(h->d_tti_l, h->d_tdi_l, h->d_tti_r, h->d_tdi_r, h->d_nmatch + 1, b2_pool_dcount(pool), 0, g.kp_cap,
                           h->cfg.match_ratio, h->cfg.min_match_dist, h->d_circ_cur, h->d_circ_hist, h->d_ncirc,
                           h->d_nleft, h->d_nright, h->d_mask, np, s);
        ++h->launches;
        prof_mark(h, -1);
    } else if (temporal) {
        const size_t row = (size_t)g.kp_cap * 32;
        b2orb_handle* src = h->root->hist_owner;
        if (src && src != h) {   // the previous pair of the sequence was processed on another lane: fetch its slot 0
            B2_CUDA(h, cudaStreamWaitEvent(s, src->ev_hist, 0));
            B2_CUDA(h, cudaMemcpyAsync(h->d_cur_l, src->d_cur_l, row, cudaMemcpyDeviceToDevice, s));
            B2_CUDA(h, cudaMemcpyAsync(h->d_cur_r, src->d_cur_r, row, cudaMemcpyDeviceToDevice, s));
            B2_CUDA(h, cudaMemcpyAsync(h->d_nmatch, src->d_nmatch, sizeof(int), cudaMemcpyDeviceToDevice, s));
        }
        prof_mark(h, B2ORB_K_MATCH_TEMPORAL);
        b2_launch_match(h->d_probs + np, 2 * np, g.kp_cap, s);
        ++h->launches;
        prof_mark(h, B2ORB_K_CIRCULAR);
        // the last pair of this call becomes the history of the next call (slot 0): for one pair the circular kernel does it
        // in its tail (its only CTA has read the old count by then), else three device-to-device copies
        const B2HistHandover ho = {h->d_cur_l, h->d_cur_r, h->d_cur_l + row, h->d_cur_r + row, h->d_nmatch};
        b2_launch_circular(h->d_tti_l, h->d_tdi_l, h->d_tti_r, h->d_tdi_r, h->d_nmatch + 1, h->d_nmatch, 1, g.kp_cap,
                           h->cfg.match_ratio, h->cfg.min_match_dist, h->d_circ_cur, h->d_circ_hist, h->d_ncirc,
                           h->d_nleft, h->d_nright, h->d_mask, np, s, np == 1 ? &ho : nullptr);
        ++h->launches;
        prof_mark(h, -1);
        if (np > 1) {
            B2_CUDA(h, cudaMemcpyAsync(h->d_cur_l, h->d_cur_l + (size_t)np * row, row, cudaMemcpyDeviceToDevice, s));
            B2_CUDA(h, cudaMemcpyAsync(h->d_cur_r, h->d_cur_r + (size_t)np * row, row, cudaMemcpyDeviceToDevice, s));
            B2_CUDA(h, cudaMemcpyAsync(h->d_nmatch, h->d_nmatch + np, sizeof(int), cudaMemcpyDeviceToDevice, s));
        }
        B2_CUDA(h, cudaEventRecord(h->ev_hist, s));
        h->root->hist_owner = h;
    }
    B2_CUDA(h, cudaGetLastError());
    return B2ORB_OK;
}

// ---- direct export (latency path) -----------------------------------------------------------------
// One stereo pair per call goes back to the host as ~14 device-to-host copies of 4-130 KB; each costs a few
// microseconds of DMA set-up and they run one after the other on the stream (measured: 33 us for the large arrays,
// ~70 us between the last kernel and the caller for a KITTI pair).  When every output array is pinned host memory
// the device can address (b2orb_host_alloc / cudaHostAlloc / cudaHostRegister), one kernel stores the VALID prefix of
// every array straight into the caller's buffers over the bus instead: coalesced 4-byte words, posted writes.
__global__ void __launch_bounds__(256)
export_kernel(const __grid_constant__ B2ExportPlan plan) {
    b2_pdl_wait();
    const B2ExportSeg& sg = plan.seg[blockIdx.x];
    const int p = blockIdx.y;
    int n = sg.fixed;
    if (sg.count) n = min(max(sg.count[p + sg.count_off], 0), sg.count_max);
    if (!sg.count && p != 0) return;                      // fixed-size segments (the count arrays) go once
    const long long bytes = (long long)n * sg.elem;
    const uint8_t* src = sg.src + (sg.count ? p * sg.src_pair : 0);
    uint8_t* dst = sg.dst + (sg.count ? p * sg.dst_pair : 0);
    if ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 3) {
        // 1-byte segments (depth flags) of pair p start at p * cap, which need not be a multiple of 4: byte copies
        for (long long i = blockIdx.z * 256 + threadIdx.x; i < bytes; i += 256 * gridDim.z) dst[i] = src[i];
        return;
    }
    const int words = (int)(bytes >> 2);
    const unsigned* s4 = reinterpret_cast<const unsigned*>(src);
    unsigned* d4 = reinterpret_cast<unsigned*>(dst);
    int w0 = 0;
    if (!((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15)) {
        // the large arrays (keypoints, descriptors: 50-130 KB) as 16-byte vectors, four loads in flight per thread: the copy
        // is a chain of L2 round trips per thread, not a bandwidth problem
        const int quads = words >> 2;
        const uint4* s16 = reinterpret_cast<const uint4*>(src);
        uint4* d16 = reinterpret_cast<uint4*>(dst);
        const int stride = 256 * gridDim.z;
        for (int i = blockIdx.z * 256 + threadIdx.x; i < quads; i += 4 * stride) {
            uint4 v[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) if (i + k * stride < quads) v[k] = s16[i + k * stride];
#pragma unroll
            for (int k = 0; k < 4; ++k) if (i + k * stride < quads) d16[i + k * stride] = v[k];
        }
        w0 = quads << 2;
    }
    for (int i = w0 + blockIdx.z * 256 + threadIdx.x; i < words; i += 256 * gridDim.z) d4[i] = s4[i];
    if (blockIdx.z == 0 && threadIdx.x < (int)(bytes & 3)) dst[(words << 2) + threadIdx.x] = src[(words << 2) + threadIdx.x];
}

// Builds the plan for `o`; false when some array is not device-visible pinned memory (caller uses the copy path).
bool export_plan(b2orb_handle* h, int np, const b2orb_stereo_out* o, B2ExportPlan& plan) {
    const int kcap = h->geom.kp_cap;
    const long long cap = o->cap;
    plan.nseg = 0;
    bool ok = true;
    auto devptr = [&](void* host) -> uint8_t* {
        cudaPointerAttributes a;
        if (cudaPointerGetAttributes(&a, host) != cudaSuccess) { cudaGetLastError(); ok = false; return nullptr; }
        if (a.type != cudaMemoryTypeHost || !a.devicePointer) { ok = false; return nullptr; }
        return static_cast<uint8_t*>(a.devicePointer);
    };
    auto add = [&](void* host, const void* src, int elem, long long src_pair_elems, long long dst_pair_elems,
                   const int* count, int count_off, int fixed) {
        if (!host || !ok) return;
        if (plan.nseg >= B2_EXPORT_MAX_SEG) { ok = false; return; }
        B2ExportSeg& sg = plan.seg[plan.nseg];
        sg.dst = devptr(host);
        if (!ok) return;
        sg.src = static_cast<const uint8_t*>(src);
        sg.elem = elem;
        sg.src_pair = src_pair_elems * elem;
        sg.dst_pair = dst_pair_elems * elem;
        sg.count = count;
        sg.count_off = count_off;
        sg.fixed = fixed;
        sg.count_max = kcap;
        ++plan.nseg;
    };
    const int* nl = h->d_nkp;
    const int* nm = h->d_nmatch + 1;
    add(o->kp_l, h->d_kps, sizeof(b2orb_keypoint), kcap, cap, nl, 0, 0);
    add(o->kp_r, h->d_kps + (size_t)np * kcap, sizeof(b2orb_keypoint), kcap, cap, nl, np, 0);
    add(o->desc_l, h->d_desc, 32, kcap, cap, nl, 0, 0);
    add(o->desc_r, h->d_desc + (size_t)np * kcap * 32, 32, kcap, cap, nl, np, 0);
    add(o->train_idx, h->d_train, 4, kcap, cap, nl, 0, 0);
    add(o->dist, h->d_dist, 4, kcap, cap, nl, 0, 0);
    add(o->match_l, h->d_match_l, 4, kcap, cap, nm, 0, 0);
    add(o->match_r, h->d_match_r, 4, kcap, cap, nm, 0, 0);
    add(o->circ_cur, h->d_circ_cur, 4, kcap, cap, h->d_ncirc, 0, 0);
    add(o->circ_hist, h->d_circ_hist, 4, kcap, cap, h->d_ncirc, 0, 0);
    if (h->root->proj_on) {
        for (int r = 0; r < 4 && o->points4d; ++r)
            add(o->points4d + (size_t)r * cap, h->d_pts4 + (size_t)r * kcap, 8, 4LL * kcap, 4LL * cap, nm, 0, 0);
        add(o->point_cam, h->d_ptcam, 24, kcap, cap, nm, 0, 0);
        add(o->depth_flags, h->d_dflag, 1, kcap, cap, nm, 0, 0);
    }
    add(o->pix_l, h->d_pix_l, 16, kcap, cap, nm, 0, 0);
    add(o->pix_r, h->d_pix_r, 16, kcap, cap, nm, 0, 0);
    add(o->cur_desc_l, h->d_cur_l + (size_t)kcap * 32, 32, kcap, cap, nm, 0, 0);
    add(o->cur_desc_r, h->d_cur_r + (size_t)kcap * 32, 32, kcap, cap, nm, 0, 0);
    add(o->cur_mask, h->d_mask, 1, kcap, cap, nm, 0, 0);
    add(o->n_l, h->d_nkp, 4, 0, 0, nullptr, 0, np);
    add(o->n_r, h->d_nkp + np, 4, 0, 0, nullptr, 0, np);
    add(o->n_match, h->d_nmatch + 1, 4, 0, 0, nullptr, 0, np);
    add(o->n_circ, h->d_ncirc, 4, 0, 0, nullptr, 0, np);
    return ok && plan.nseg > 0;
}

// Results of `np` pairs to the caller and wait: direct export for latency-sized calls with pinned outputs, else the
// copy path (b2orb_stereo_download).
// A caller that asks only for what the reference keeps of a frame (kept pixels / descriptor rows / lists / counts,
// src/feature-tracker.cpp:210-218) gets a few KB per pair: valid prefixes by export_kernel beat whole-slab copies at
// any batch size then.
bool kept_only(const b2orb_stereo_out* o) {
    return !o->kp_l && !o->kp_r && !o->desc_l && !o->desc_r && !o->train_idx && !o->dist;
}

int finish_outputs(b2orb_handle* h, int np, const b2orb_stereo_out* o) {
    if (o && (np <= h->root->tune.export_max || (kept_only(o) && h->root->tune.export_max > 0)) && o->cap >= h->geom.kp_cap) {
        static_assert(sizeof(B2ExportPlan) < 3500, "plan must fit the kernel parameter space");
        // The pointer attributes are queried on every call (a few hundred ns each): a cached verdict would go stale when
        // the caller frees a pinned array and a pageable one lands on the same address, or toggles b2orb_set_projection.
        h->exp_ok = export_plan(h, np, o, h->exp_plan);
        if (h->exp_ok) {
            b2_launch_pdl(export_kernel, dim3(h->exp_plan.nseg, np, 4), dim3(256), 0, h->stream, h->exp_plan);
            ++h->launches;
            B2_CUDA(h, cudaGetLastError());
            B2_CUDA(h, cudaStreamSynchronize(h->stream));
            return B2ORB_OK;
        }
    }
    return b2orb_stereo_download(h, np, o);
}

}  // namespace

// =================================================================================================
// exported C ABI
// =================================================================================================
extern "C" {

int b2orb_abi_version(void) { return B2ORB_ABI_VERSION; }

int b2orb_create(b2orb_handle** out, int device, const b2orb_config* cfg) {
    if (!out || !cfg) return B2ORB_ERR_INVALID;
    *out = nullptr;
    if (cfg->nfeatures < 1 || cfg->nlevels < 1 || cfg->nlevels > B2_MAX_LEVELS || !(cfg->scale_factor > 1.f) || cfg->scale_factor > 3.f ||
        cfg->ini_th_fast < 1 || cfg->min_th_fast < 1 || cfg->min_th_fast > cfg->ini_th_fast ||
        cfg->ini_th_fast > 254 || cfg->max_width < 1 || cfg->max_height < 1 || cfg->max_images < 1)
        return B2ORB_ERR_INVALID;
    {   // blur taps: all zero (default = OpenCV 4.x {18,34,48,56,48,34,18}) or a symmetric u8 kernel whose Q8 x Q8
        // accumulation fits 32 bits; sums above 256 (OpenCV 3.2's {18,34,49,55,49,34,18} = 257) saturate (blur.cu)
        int sum = 0; bool zero = true, ok = true;
        for (int i = 0; i < 7; ++i) {
            zero = zero && cfg->blur_taps[i] == 0;
            ok = ok && cfg->blur_taps[i] >= 0 && cfg->blur_taps[i] <= 255 && cfg->blur_taps[i] == cfg->blur_taps[6 - i];
            sum += cfg->blur_taps[i];
        }
        if (!zero && (!ok || sum < 1 || sum > 1024)) return B2ORB_ERR_INVALID;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return B2ORB_ERR_CUDA;
    b2orb_handle* h = new (std::nothrow) b2orb_handle();
    if (!h) return B2ORB_ERR_NOMEM;
    h->cfg = *cfg;
    h->device = device;
    *out = h;   // returned even on failure so the caller can read b2orb_last_error, then destroy
    if (!cuda_ok(h, cudaSetDevice(device), "cudaSetDevice")) return B2ORB_ERR_CUDA;
    if (!cuda_ok(h, cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking), "cudaStreamCreate")) return B2ORB_ERR_CUDA;
    cudaEventCreate(&h->ev0);
    cudaEventCreate(&h->ev1);
    cudaEventCreateWithFlags(&h->ev_hist, cudaEventDisableTiming);
    cudaStreamCreateWithFlags(&h->s_aux, cudaStreamNonBlocking);
    cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&h->ev_done, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&h->ev_in, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&h->ev_out, cudaEventDisableTiming);
    h->root = h;
    build_ctor_tables(h);
    return configure(h, cfg->max_width, cfg->max_height);
}

void b2orb_destroy(b2orb_handle* h) {
    if (!h) return;
    cudaSetDevice(h->device);
    for (b2orb_handle* l : h->lanes) b2orb_destroy(l);
    h->lanes.clear();
    if (h->stream) cudaStreamSynchronize(h->stream);
    free_workspace(h);
    for (int e = 0; e < 2; ++e) {
        if (h->d_map1[e]) cudaFree(h->d_map1[e]);
        if (h->d_map2[e]) cudaFree(h->d_map2[e]);
    }
    if (h->pool) b2_pool_destroy(h->pool);
    if (h->desc_cache) b2_desc_cache_destroy(h->desc_cache);
    if (h->h_async_n) cudaFreeHost(h->h_async_n);
    if (h->d_packed) cudaFree(h->d_packed);
    if (h->d_tri_in) cudaFree(h->d_tri_in);
    if (h->d_tri_out) cudaFree(h->d_tri_out);
    for (void* p : {(void*)h->d_ing_xofs, (void*)h->d_ing_yofs, (void*)h->d_ing_xw, (void*)h->d_ing_yw})
        if (p) cudaFree(p);
    if (h->ev_hist) cudaEventDestroy(h->ev_hist);
    if (h->ev_fork) cudaEventDestroy(h->ev_fork);
    if (h->ev_join) cudaEventDestroy(h->ev_join);
    if (h->s_aux) cudaStreamDestroy(h->s_aux);
    if (h->ev_done) cudaEventDestroy(h->ev_done);
    if (h->ev_in) cudaEventDestroy(h->ev_in);
    if (h->ev_out) cudaEventDestroy(h->ev_out);
    if (h->s_in) cudaStreamDestroy(h->s_in);
    if (h->s_out) cudaStreamDestroy(h->s_out);
    for (cudaEvent_t e : h->prof_ev) cudaEventDestroy(e);
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

const char* b2orb_last_error(const b2orb_handle* h) { return h ? h->err.c_str() : "null handle"; }
int b2orb_keypoint_capacity(const b2orb_handle* h) { return h ? h->geom.kp_cap : 0; }
uint64_t b2orb_launch_count(const b2orb_handle* h) {
    if (!h) return 0;
    uint64_t n = h->launches;
    for (const b2orb_handle* l : h->lanes) n += l->launches;
    return n;
}
float b2orb_last_device_ms(const b2orb_handle* h) { return h ? h->last_ms : 0.f; }

int b2orb_get_tables(const b2orb_handle* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2,
                     int32_t* features_per_level) {
    if (!h) return B2ORB_ERR_INVALID;
    for (int i = 0; i < h->cfg.nlevels; ++i) {
        if (scale) scale[i] = h->scale[i];
        if (inv_scale) inv_scale[i] = h->inv_scale[i];
        if (sigma2) sigma2[i] = h->sigma2[i];
        if (inv_sigma2) inv_sigma2[i] = h->inv_sigma2[i];
        if (features_per_level) features_per_level[i] = h->quota[i];
    }
    return B2ORB_OK;
}

void* b2orb_host_alloc(size_t bytes) {
    void* p = nullptr;
    return cudaHostAlloc(&p, bytes, cudaHostAllocDefault) == cudaSuccess ? p : nullptr;
}
void b2orb_host_free(void* p) {
    if (p) cudaFreeHost(p);
}

int b2orb_sync(b2orb_handle* h) {
    if (!h) return B2ORB_ERR_INVALID;
    B2_CUDA(h, cudaSetDevice(h->device));
    B2_CUDA(h, cudaStreamSynchronize(h->stream));
    cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1);
    return B2ORB_OK;
}

void b2orb_reset_history(b2orb_handle* h) {
    if (!h || !h->d_nmatch) return;
    cudaSetDevice(h->device);
    cudaMemsetAsync(h->d_nmatch, 0, sizeof(int), h->stream);
    for (b2orb_handle* l : h->lanes)
        if (l->d_nmatch) cudaMemsetAsync(l->d_nmatch, 0, sizeof(int), l->stream);
    h->hist_owner = nullptr;
}

int b2orb_extract_batch(b2orb_handle* h, const uint8_t* images, int count, int width, int height, int step,
                        size_t image_stride, b2orb_keypoint* kps, uint8_t* desc, int cap, int32_t* n) {
    if (!h || !n || count < 0) return B2ORB_ERR_INVALID;
    if (count == 0) return B2ORB_OK;
    if (width == 0 || height == 0) {   // `if(_image.empty()) return;` src/ORBextractor.cc:1046-1047
        for (int i = 0; i < count; ++i) n[i] = 0;
        return B2ORB_OK;
    }
    if (!images || !kps || !desc || width < 0 || height < 0 || step < width) return fail(h, B2ORB_ERR_INVALID, "bad image arguments");
    if (count > h->cfg.max_images) return fail(h, B2ORB_ERR_CAPACITY, "count exceeds max_images");
    if (width > h->cfg.max_width || height > h->cfg.max_height) return fail(h, B2ORB_ERR_INVALID, "image larger than configured maximum");
    B2_CUDA(h, cudaSetDevice(h->device));
    int rc = configure(h, width, height);
    if (rc) return rc;
    const int kcap = h->geom.kp_cap;
    cudaEventRecord(h->ev0, h->stream);
    if ((rc = upload(h, images, count, width, height, step, image_stride, 0))) return rc;
    B2Images im = make_images(h, h->d_stage, h->d_stage, count, h->stage_img_bytes, h->stage_pitch);
    if ((rc = launch_extract(h, im, count))) return rc;
    B2_CUDA(h, cudaMemcpyAsync(n, h->d_nkp, (size_t)count * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    cudaEventRecord(h->ev1, h->stream);
    B2_CUDA(h, cudaStreamSynchronize(h->stream));
    cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1);
    for (int i = 0; i < count; ++i) {
        if (n[i] > cap) return fail(h, B2ORB_ERR_CAPACITY, "keypoint buffer too small");
        if (n[i] == 0) continue;
        B2_CUDA(h, cudaMemcpyAsync(kps + (size_t)i * cap, h->d_kps + (size_t)i * kcap, (size_t)n[i] * sizeof(b2orb_keypoint),
                                   cudaMemcpyDeviceToHost, h->stream));
        B2_CUDA(h, cudaMemcpyAsync(desc + (size_t)i * cap * 32, h->d_desc + (size_t)i * kcap * 32, (size_t)n[i] * 32,
                                   cudaMemcpyDeviceToHost, h->stream));
    }
    B2_CUDA(h, cudaStreamSynchronize(h->stream));
    return B2ORB_OK;
}

int b2orb_extract(b2orb_handle* h, const uint8_t* image, int width, int height, int step, b2orb_keypoint* kps,
                  uint8_t* desc, int cap, int* n) {
    if (!n) return B2ORB_ERR_INVALID;
    int32_t cnt = 0;
    const int rc = b2orb_extract_batch(h, image, 1, width, height, step, (size_t)step * (size_t)(height > 0 ? height : 0),
                                       kps, desc, cap, &cnt);
    *n = cnt;
    return rc;
}

int b2orb_stereo_frontend_device(b2orb_handle* h, const uint8_t* d_images_l, const uint8_t* d_images_r, int npairs,
                                 int width, int height, int step, size_t image_stride, int temporal) {
    if (!h || !d_images_l || !d_images_r || npairs < 1 || width < 1 || height < 1 || step < width)
        return fail(h, B2ORB_ERR_INVALID, "bad stereo arguments");
    if (2 * npairs > h->cfg.max_images) return fail(h, B2ORB_ERR_CAPACITY, "2*npairs exceeds max_images");
    if (width > h->cfg.max_width || height > h->cfg.max_height) return fail(h, B2ORB_ERR_INVALID, "image larger than configured maximum");
    if (npairs == 1) image_stride = (size_t)step * height;
    B2_CUDA(h, cudaSetDevice(h->device));
    int rc = configure(h, width, height);
    if (rc) return rc;
    if ((rc = bind_problems(h, npairs))) return rc;
    cudaEventRecord(h->ev0, h->stream);
    B2Images im;
    const auto aligned = [](const void* p, size_t a, size_t b) { return ((uintptr_t)p | a | b) % 16 == 0; };
    if (!rectifying(h, width, height, 0) && aligned(d_images_l, (size_t)step, image_stride) &&
        aligned(d_images_r, (size_t)step, image_stride)) {
        im = make_images(h, d_images_l, d_images_r, npairs, image_stride, step);   // read in place
    } else {   // kernels need 16-byte aligned rows: repack into the staging area (device-to-device)
        if ((rc = upload(h, d_images_l, npairs, width, height, step, image_stride, 0, true, 0))) return rc;
        if ((rc = upload(h, d_images_r, npairs, width, height, step, image_stride, npairs, true, 1))) return rc;
        im = make_images(h, h->d_stage, h->d_stage + (size_t)npairs * h->stage_img_bytes, npairs, h->stage_img_bytes,
                         h->stage_pitch);
    }
    rc = launch_stereo(h, im, npairs, temporal);
    cudaEventRecord(h->ev1, h->stream);
    return rc;
}

static int enqueue_download(b2orb_handle* h, int npairs, const b2orb_stereo_out* o, cudaStream_t s) {
    if (!h || !o || npairs < 1 || 2 * npairs > h->nimg_cap) return B2ORB_ERR_INVALID;
    const int kcap = h->geom.kp_cap;
    if (o->cap < kcap) return fail(h, B2ORB_ERR_CAPACITY, "b2orb_stereo_out.cap below b2orb_keypoint_capacity()");
    B2_CUDA(h, cudaSetDevice(h->device));
    const size_t np = npairs;
    // cap == kcap lets every slab go back in one copy; otherwise row-wise 2D copies
    auto copy2d = [&](void* dst, const void* src, size_t elem, size_t rows) -> cudaError_t {
        if (!dst) return cudaSuccess;
        if (o->cap == kcap) return cudaMemcpyAsync(dst, src, (size_t)kcap * elem * rows, cudaMemcpyDeviceToHost, s);
        return cudaMemcpy2DAsync(dst, (size_t)o->cap * elem, src, (size_t)kcap * elem, (size_t)kcap * elem, rows,
                                 cudaMemcpyDeviceToHost, s);
    };
    B2_CUDA(h, copy2d(o->kp_l, h->d_kps, sizeof(b2orb_keypoint), np));
    B2_CUDA(h, copy2d(o->kp_r, h->d_kps + np * kcap, sizeof(b2orb_keypoint), np));
    B2_CUDA(h, copy2d(o->desc_l, h->d_desc, 32, np));
    B2_CUDA(h, copy2d(o->desc_r, h->d_desc + np * kcap * 32, 32, np));
    B2_CUDA(h, copy2d(o->train_idx, h->d_train, 4, np));
    B2_CUDA(h, copy2d(o->dist, h->d_dist, 4, np));
    B2_CUDA(h, copy2d(o->match_l, h->d_match_l, 4, np));
    B2_CUDA(h, copy2d(o->match_r, h->d_match_r, 4, np));
    B2_CUDA(h, copy2d(o->circ_cur, h->d_circ_cur, 4, np));
    B2_CUDA(h, copy2d(o->circ_hist, h->d_circ_hist, 4, np));
    if (h->root->proj_on) {
        B2_CUDA(h, copy2d(o->points4d, h->d_pts4, 8, 4 * np));
        B2_CUDA(h, copy2d(o->point_cam, h->d_ptcam, 24, np));
        B2_CUDA(h, copy2d(o->depth_flags, h->d_dflag, 1, np));
    }
    B2_CUDA(h, copy2d(o->pix_l, h->d_pix_l, 16, np));
    B2_CUDA(h, copy2d(o->pix_r, h->d_pix_r, 16, np));
    B2_CUDA(h, copy2d(o->cur_desc_l, h->d_cur_l + (size_t)kcap * 32, 32, np));
    B2_CUDA(h, copy2d(o->cur_desc_r, h->d_cur_r + (size_t)kcap * 32, 32, np));
    B2_CUDA(h, copy2d(o->cur_mask, h->d_mask, 1, np));
    if (o->n_l) B2_CUDA(h, cudaMemcpyAsync(o->n_l, h->d_nkp, np * 4, cudaMemcpyDeviceToHost, s));
    if (o->n_r) B2_CUDA(h, cudaMemcpyAsync(o->n_r, h->d_nkp + np, np * 4, cudaMemcpyDeviceToHost, s));
    if (o->n_match) B2_CUDA(h, cudaMemcpyAsync(o->n_match, h->d_nmatch + 1, np * 4, cudaMemcpyDeviceToHost, s));
    if (o->n_circ) B2_CUDA(h, cudaMemcpyAsync(o->n_circ, h->d_ncirc, np * 4, cudaMemcpyDeviceToHost, s));
    return B2ORB_OK;
}

int b2orb_stereo_download(b2orb_handle* h, int npairs, const b2orb_stereo_out* o) {
    const int rc = enqueue_download(h, npairs, o, h->stream);
    if (rc) return rc;
    B2_CUDA(h, cudaStreamSynchronize(h->stream));
    return B2ORB_OK;
}

// b2orb_stereo_frontend for a long batch: chunks of `pipe_chunk` pairs rotate over the lanes (workspaces).  Three
// stages run on three kinds of stream and are tied together by events only:
//   copy-in stream  : H2D of chunk k into lane k%L's staging area, as soon as that lane's previous kernels are done
//   lane stream     : re-pitch + the whole launch sequence, once the chunk is in and the lane's previous results are out
//   copy-out stream : D2H of the chunk's results
// so the H2D of chunk k+1 and the D2H of chunk k-1 overlap the kernels of chunk k with no host thread involved.
// Results are identical to one big call: images are independent and the temporal history is handed from lane to
// lane on the device (launch_stereo).
static int stereo_frontend_pipelined(b2orb_handle* h, const uint8_t* images_l, const uint8_t* images_r, int npairs,
                                     int width, int height, int step, size_t image_stride, int temporal,
                                     const b2orb_stereo_out* out) {
    const int chunk = h->tune.pipe_chunk;
    while ((int)h->lanes.size() < h->tune.pipe_lanes) {
        b2orb_config c = h->cfg;
        c.max_images = 2 * chunk;
        b2orb_handle* l = nullptr;
        const int rc = b2orb_create(&l, h->device, &c);
        if (rc) {
            if (l) { h->err = l->err; b2orb_destroy(l); }
            return rc;
        }
        l->root = h;
        h->lanes.push_back(l);
    }
    if (!h->s_in) {
        B2_CUDA(h, cudaStreamCreateWithFlags(&h->s_in, cudaStreamNonBlocking));
        B2_CUDA(h, cudaStreamCreateWithFlags(&h->s_out, cudaStreamNonBlocking));
    }
    const int nl = (int)h->lanes.size();
    cudaEventRecord(h->ev0, h->stream);
    B2_CUDA(h, cudaStreamWaitEvent(h->s_in, h->ev0, 0));   // ordered after earlier work on this handle
    const size_t cap = out->cap;
    int k = 0, rc = B2ORB_OK;
    const bool trace = h->tune.pipe_trace != 0;   // debugging aid: per-chunk timeline on stderr
    struct Tr { cudaEvent_t e[4]; double h0, h1; };
    std::vector<Tr> tr;
    const auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t_start = now();
    b2orb_handle* l = nullptr;
    for (int p0 = 0; p0 < npairs; p0 += chunk, ++k) {
        l = h->lanes[k % nl];
        const int n = std::min(chunk, npairs - p0);
        if (trace) {
            Tr t{};
            for (auto& e : t.e) cudaEventCreate(&e);
            t.h0 = now() - t_start;
            tr.push_back(t);
        }
        if ((rc = configure(l, width, height))) break;
        // ---- copy-in ----
        if (k >= nl) cudaStreamWaitEvent(h->s_in, l->ev_done, 0);             // the lane's staging area is free again
        if (trace) cudaEventRecord(tr.back().e[0], h->s_in);
        if ((rc = upload_copy(l, images_l + (size_t)p0 * image_stride, n, width, height, step, image_stride, 0, h->s_in, 0))) break;
        if ((rc = upload_copy(l, images_r + (size_t)p0 * image_stride, n, width, height, step, image_stride, n, h->s_in, 1))) break;
        cudaEventRecord(l->ev_in, h->s_in);
        if (trace) cudaEventRecord(tr.back().e[1], h->s_in);
        // ---- kernels ----
        cudaStreamWaitEvent(l->stream, l->ev_in, 0);
        if (k >= nl) cudaStreamWaitEvent(l->stream, l->ev_out, 0);            // the lane's previous results are on the host
        else cudaStreamWaitEvent(l->stream, h->ev0, 0);
        if ((rc = bind_problems(l, n))) break;
        if ((rc = upload_repack(l, nullptr, n, width, height, step, image_stride, 0, false, 0))) break;
        if ((rc = upload_repack(l, nullptr, n, width, height, step, image_stride, n, false, 1))) break;
        B2Images im = make_images(l, l->d_stage, l->d_stage + (size_t)n * l->stage_img_bytes, n, l->stage_img_bytes,
                                  l->stage_pitch);
        if ((rc = launch_stereo(l, im, n, temporal))) break;
        cudaEventRecord(l->ev_done, l->stream);
        if (trace) cudaEventRecord(tr.back().e[2], l->stream);
        // ---- copy-out ----
        cudaStreamWaitEvent(h->s_out, l->ev_done, 0);
        b2orb_stereo_out o = *out;
        const auto shift = [&](auto*& ptr, size_t per_pair) { if (ptr) ptr += (size_t)p0 * per_pair; };
        shift(o.kp_l, cap); shift(o.kp_r, cap); shift(o.desc_l, cap * 32); shift(o.desc_r, cap * 32);
        shift(o.train_idx, cap); shift(o.dist, cap); shift(o.match_l, cap); shift(o.match_r, cap);
        shift(o.circ_cur, cap); shift(o.circ_hist, cap);
        shift(o.points4d, cap * 4); shift(o.point_cam, cap * 3); shift(o.depth_flags, cap);
        shift(o.n_l, 1); shift(o.n_r, 1); shift(o.n_match, 1); shift(o.n_circ, 1);
        shift(o.pix_l, cap * 2); shift(o.pix_r, cap * 2); shift(o.cur_desc_l, cap * 32); shift(o.cur_desc_r, cap * 32);
        shift(o.cur_mask, cap);
        if (kept_only(&o) && h->tune.export_max > 0 && o.cap >= l->geom.kp_cap && export_plan(l, n, &o, l->exp_plan)) {
            // kept-only outputs in pinned memory: the lane writes the valid prefixes itself, nothing for the copy-out stream
            b2_pdl_on = false;
            b2_launch_pdl(export_kernel, dim3(l->exp_plan.nseg, n, 4), dim3(256), 0, l->stream, l->exp_plan);
            ++l->launches;
            cudaEventRecord(l->ev_done, l->stream);
            cudaStreamWaitEvent(h->s_out, l->ev_done, 0);
        } else if ((rc = enqueue_download(l, n, &o, h->s_out))) break;
        cudaEventRecord(l->ev_out, h->s_out);
        if (trace) { cudaEventRecord(tr.back().e[3], h->s_out); tr.back().h1 = now() - t_start; }
    }
    if (rc && l) h->err = l->err;
    cudaStreamSynchronize(h->s_in);
    for (b2orb_handle* q : h->lanes) cudaStreamSynchronize(q->stream);
    cudaEventRecord(h->ev1, h->s_out);
    B2_CUDA(h, cudaStreamSynchronize(h->s_out));
    cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1);
    if (trace && !rc) {
        fprintf(stderr, "[b2orb pipe] chunk: host enqueue [h0,h1] | device: h2d start, h2d done, kernels done, d2h done (ms); total %.3f host %.3f\n", h->last_ms, now() - t_start);
        for (size_t i = 0; i < tr.size(); ++i) {
            float t[4];
            for (int j = 0; j < 4; ++j) cudaEventElapsedTime(&t[j], h->ev0, tr[i].e[j]);
            fprintf(stderr, "[b2orb pipe] %2zu: [%.3f, %.3f] | %.3f %.3f %.3f %.3f\n", i, tr[i].h0, tr[i].h1, t[0], t[1], t[2], t[3]);
        }
    }
    for (Tr& t : tr) for (auto& e : t.e) if (e) cudaEventDestroy(e);
    return rc;
}

int b2orb_stereo_frontend(b2orb_handle* h, const uint8_t* images_l, const uint8_t* images_r, int npairs, int width,
                          int height, int step, size_t image_stride, int temporal, const b2orb_stereo_out* out) {
    if (!h || !images_l || !images_r || !out || npairs < 1 || width < 1 || height < 1 || step < width)
        return fail(h, B2ORB_ERR_INVALID, "bad stereo arguments");
    if (2 * npairs > h->cfg.max_images) return fail(h, B2ORB_ERR_CAPACITY, "2*npairs exceeds max_images");
    if (width > h->cfg.max_width || height > h->cfg.max_height) return fail(h, B2ORB_ERR_INVALID, "image larger than configured maximum");
    if (npairs == 1) image_stride = (size_t)step * height;   // unused for one pair: keep the one-copy upload paths
    B2_CUDA(h, cudaSetDevice(h->device));
    if (h->tune.pipe_chunk > 0 && npairs > h->tune.pipe_chunk && h->root == h)
        return stereo_frontend_pipelined(h, images_l, images_r, npairs, width, height, step, image_stride, temporal, out);
    int rc = configure(h, width, height);
    if (rc) return rc;
    if ((rc = bind_problems(h, npairs))) return rc;
    cudaEventRecord(h->ev0, h->stream);
    if (!(npairs == 1 && upload_zero_copy(h, images_l, images_r, width, height, step, 0))) {
        if ((rc = upload(h, images_l, npairs, width, height, step, image_stride, 0, false, 0))) return rc;
        if ((rc = upload(h, images_r, npairs, width, height, step, image_stride, npairs, false, 1))) return rc;
    }
    B2Images im = make_images(h, h->d_stage, h->d_stage + (size_t)npairs * h->stage_img_bytes, npairs,
                              h->stage_img_bytes, h->stage_pitch);
    if ((rc = launch_stereo(h, im, npairs, temporal))) return rc;
    cudaEventRecord(h->ev1, h->stream);
    rc = finish_outputs(h, npairs, out);
    cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1);
    return rc;
}

// ---- triangulation (next row f4: cv::triangulatePoints + depth gates) -------------------------------
int b2orb_set_projection(b2orb_handle* h, const double* P1, const double* P2, double max_depth) {
    if (!h || h->root != h) return B2ORB_ERR_INVALID;
    if (!P1 != !P2) return fail(h, B2ORB_ERR_INVALID, "b2orb_set_projection needs both matrices (or both NULL)");
    B2_CUDA(h, cudaSetDevice(h->device));
    if (!P1) { h->proj_on = false; return B2ORB_OK; }
    memcpy(h->proj.P1, P1, sizeof h->proj.P1);
    memcpy(h->proj.P2, P2, sizeof h->proj.P2);
    h->proj.max_depth = max_depth;
    h->proj_on = true;   // passed by value with every launch: nothing in flight reads it
    return B2ORB_OK;
}

int b2orb_triangulate(b2orb_handle* h, const double* P1, const double* P2, const double* pix_l, const double* pix_r,
                      int n, double* points4d) {
    if (!h || !P1 || !P2 || n < 0 || (n > 0 && (!pix_l || !pix_r || !points4d)))
        return fail(h, B2ORB_ERR_INVALID, "bad triangulation arguments");
    if (n == 0) return B2ORB_OK;   // cv::triangulatePoints on empty vectors yields a 4 x 0 matrix
    B2_CUDA(h, cudaSetDevice(h->device));
    if (n > h->tri_cap) {
        B2_CUDA(h, cudaStreamSynchronize(h->stream));
        if (h->d_tri_in) cudaFree(h->d_tri_in);
        if (h->d_tri_out) cudaFree(h->d_tri_out);
        h->d_tri_in = h->d_tri_out = nullptr;
        h->tri_cap = 0;
        const int want = std::max(n, 4096);
        if (cudaMalloc(&h->d_tri_in, (size_t)want * 4 * sizeof(double)) != cudaSuccess ||
            cudaMalloc(&h->d_tri_out, (size_t)want * 4 * sizeof(double)) != cudaSuccess)
            return fail(h, B2ORB_ERR_NOMEM, "triangulation scratch");
        h->tri_cap = want;
    }
    B2Projection pr;
    memcpy(pr.P1, P1, sizeof pr.P1);
    memcpy(pr.P2, P2, sizeof pr.P2);
    pr.max_depth = 0;
    cudaStream_t s = h->stream;
    double* d_l = h->d_tri_in;
    double* d_r = h->d_tri_in + (size_t)2 * n;
    B2_CUDA(h, cudaMemcpyAsync(d_l, pix_l, (size_t)n * 16, cudaMemcpyHostToDevice, s));
    B2_CUDA(h, cudaMemcpyAsync(d_r, pix_r, (size_t)n * 16, cudaMemcpyHostToDevice, s));
    prof_mark(h, B2ORB_K_TRIANGULATE);
    b2_launch_triangulate(pr, d_l, d_r, n, h->d_tri_out, s);
    ++h->launches;
    prof_mark(h, -1);
    B2_CUDA(h, cudaGetLastError());
    B2_CUDA(h, cudaMemcpyAsync(points4d, h->d_tri_out, (size_t)n * 32, cudaMemcpyDeviceToHost, s));
    B2_CUDA(h, cudaStreamSynchronize(s));
    return B2ORB_OK;
}

// ---- rectification (next row f1: cv::remap in front of the pyramid) --------------------------------
int b2orb_set_rectification(b2orb_handle* h, int width, int height, const int16_t* map1_l, const uint16_t* map2_l,
                            const int16_t* map1_r, const uint16_t* map2_r) {
    if (!h || h->root != h) return B2ORB_ERR_INVALID;
    B2_CUDA(h, cudaSetDevice(h->device));
    const bool off = !map1_l && !map2_l && !map1_r && !map2_r;
    if (!off && (!map1_l || !map2_l || !map1_r || !map2_r || width < 1 || height < 1 || width > h->cfg.max_width ||
                 height > h->cfg.max_height))
        return fail(h, B2ORB_ERR_INVALID, "bad rectification maps");
    // no launch may still be reading the old maps
    B2_CUDA(h, cudaStreamSynchronize(h->stream));
    for (b2orb_handle* l : h->lanes) B2_CUDA(h, cudaStreamSynchronize(l->stream));
    for (int e = 0; e < 2; ++e) {
        if (h->d_map1[e]) cudaFree(h->d_map1[e]);
        if (h->d_map2[e]) cudaFree(h->d_map2[e]);
        h->d_map1[e] = h->d_map2[e] = nullptr;
    }
    h->rect_w = h->rect_h = 0;
    if (off) return B2ORB_OK;
    const size_t n = (size_t)width * height;
    const int16_t* m1[2] = {map1_l, map1_r};
    const uint16_t* m2[2] = {map2_l, map2_r};
    for (int e = 0; e < 2; ++e) {
        B2_CUDA(h, cudaMalloc(&h->d_map1[e], n * 4));
        B2_CUDA(h, cudaMalloc(&h->d_map2[e], n * 2));
        B2_CUDA(h, cudaMemcpy(h->d_map1[e], m1[e], n * 4, cudaMemcpyHostToDevice));
        B2_CUDA(h, cudaMemcpy(h->d_map2[e], m2[e], n * 2, cudaMemcpyHostToDevice));
    }
    h->rect_w = width;
    h->rect_h = height;
    return B2ORB_OK;
}

int b2orb_remap(b2orb_handle* h, int eye, const uint8_t* image, int width, int height, int step, uint8_t* out,
                int out_step) {
    if (!h || !image || !out || eye < 0 || eye > 1 || step < width || out_step < width)
        return fail(h, B2ORB_ERR_INVALID, "bad remap arguments");
    if (h->rect_w != width || h->rect_h != height || width < 1)
        return fail(h, B2ORB_ERR_INVALID, "no rectification maps for this image size");
    B2_CUDA(h, cudaSetDevice(h->device));
    int rc = configure(h, width, height);
    if (rc) return rc;
    cudaEventRecord(h->ev0, h->stream);
    if ((rc = upload(h, image, 1, width, height, step, (size_t)step * height, 0, false, eye))) return rc;
    cudaEventRecord(h->ev1, h->stream);
    B2_CUDA(h, cudaMemcpy2DAsync(out, out_step, h->d_stage, h->stage_pitch, width, height, cudaMemcpyDeviceToHost, h->stream));
    B2_CUDA(h, cudaStreamSynchronize(h->stream));
    cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1);
    return B2ORB_OK;
}

// ---- cfsd live ingest (next row f2: colour frame -> resize -> grey -> split) -----------------------
// Uploads `count` packed frames and leaves the grey eyes of frame i in staging slots i (left) and count + i (right);
// with rectification on, the grey eyes go through remap_kernel on the way (process() -> processImage()).
static int ingest_packed(b2orb_handle* h, const uint8_t* frames, int count, int src_w, int src_h, int src_step,
                         size_t frame_stride, int width, int height) {
    const int key[4] = {src_w, src_h, 2 * width, height};
    if (std::memcmp(key, h->ing_key, sizeof key) != 0) {   // cv::resize tables for this geometry
        std::vector<int> xo, yo;
        std::vector<short> xw, yw;
        resize_axis_table(src_w, 2 * width, xo, xw);
        resize_axis_table(src_h, height, yo, yw);
        B2_CUDA(h, cudaStreamSynchronize(h->stream));
        for (void* p : {(void*)h->d_ing_xofs, (void*)h->d_ing_yofs, (void*)h->d_ing_xw, (void*)h->d_ing_yw})
            if (p) cudaFree(p);
        B2_CUDA(h, cudaMalloc(&h->d_ing_xofs, xo.size() * 4));
        B2_CUDA(h, cudaMalloc(&h->d_ing_yofs, yo.size() * 4));
        B2_CUDA(h, cudaMalloc(&h->d_ing_xw, xw.size() * 2));
        B2_CUDA(h, cudaMalloc(&h->d_ing_yw, yw.size() * 2));
        B2_CUDA(h, cudaMemcpy(h->d_ing_xofs, xo.data(), xo.size() * 4, cudaMemcpyHostToDevice));
        B2_CUDA(h, cudaMemcpy(h->d_ing_yofs, yo.data(), yo.size() * 4, cudaMemcpyHostToDevice));
        B2_CUDA(h, cudaMemcpy(h->d_ing_xw, xw.data(), xw.size() * 2, cudaMemcpyHostToDevice));
        B2_CUDA(h, cudaMemcpy(h->d_ing_yw, yw.data(), yw.size() * 2, cudaMemcpyHostToDevice));
        std::memcpy(h->ing_key, key, sizeof key);
    }
    const size_t fbytes = (size_t)src_step * src_h;
    if (h->packed_bytes < fbytes * count) {
        B2_CUDA(h, cudaStreamSynchronize(h->stream));
        if (h->d_packed) cudaFree(h->d_packed);
        h->d_packed = nullptr;
        h->packed_bytes = 0;
        B2_CUDA(h, cudaMalloc(&h->d_packed, fbytes * count));
        h->packed_bytes = fbytes * count;
    }
    if (frame_stride == fbytes) {
        B2_CUDA(h, cudaMemcpyAsync(h->d_packed, frames, fbytes * count, cudaMemcpyHostToDevice, h->stream));
    } else {
        for (int i = 0; i < count; ++i)
            B2_CUDA(h, cudaMemcpyAsync(h->d_packed + i * fbytes, frames + i * frame_stride, fbytes, cudaMemcpyHostToDevice, h->stream));
    }
    const bool rect = rectifying(h, width, height, 0);
    // with rectification the grey eyes are an intermediate: park them in the (not yet used) blur slab, pitched
    uint8_t* gl = rect ? h->d_blur : h->d_stage;
    const size_t gstride = rect ? (size_t)h->geom.slab : h->stage_img_bytes;
    uint8_t* gr = gl + (size_t)count * gstride;
    prof_mark(h, B2ORB_K_INGEST);
    b2_launch_ingest_bgra(h->d_packed, src_step, fbytes, src_w, src_h, h->d_ing_xofs, h->d_ing_xw, h->d_ing_yofs, h->d_ing_yw,
                          gl, gr, h->stage_pitch, gstride, width, height, count, h->stream);
    ++h->launches;
    if (rect) {
        for (int e = 0; e < 2; ++e) {
            b2_launch_remap(e ? gr : gl, h->stage_pitch, gstride, h->root->d_map1[e], h->root->d_map2[e],
                            h->d_stage + (size_t)e * count * h->stage_img_bytes, h->stage_pitch, h->stage_img_bytes, width, height,
                            count, h->stream);
            ++h->launches;
        }
    }
    prof_mark(h, -1);
    return B2ORB_OK;
}

static int check_packed_args(b2orb_handle* h, const void* frames, int count, int src_w, int src_h, int src_step, int width,
                             int height) {
    if (!h || !frames || count < 1 || src_w < 2 || src_h < 1 || src_step < 4 * src_w || (src_step & 3) || width < 1 || height < 1)
        return fail(h, B2ORB_ERR_INVALID, "bad packed-frame arguments");
    if (2 * count > h->cfg.max_images) return fail(h, B2ORB_ERR_CAPACITY, "2*frames exceeds max_images");
    if (width > h->cfg.max_width || height > h->cfg.max_height) return fail(h, B2ORB_ERR_INVALID, "image larger than configured maximum");
    return B2ORB_OK;
}

int b2orb_cfsd_ingest(b2orb_handle* h, const uint8_t* frame, int src_w, int src_h, int src_step, int width, int height,
                      uint8_t* gray_l, uint8_t* gray_r, int out_step) {
    int rc = check_packed_args(h, frame, 1, src_w, src_h, src_step, width, height);
    if (rc) return rc;
    if (!gray_l || !gray_r || out_step < width) return fail(h, B2ORB_ERR_INVALID, "bad packed-frame arguments");
    B2_CUDA(h, cudaSetDevice(h->device));
    if ((rc = configure(h, width, height))) return rc;
    cudaEventRecord(h->ev0, h->stream);
    if ((rc = ingest_packed(h, frame, 1, src_w, src_h, src_step, (size_t)src_step * src_h, width, height))) return rc;
    cudaEventRecord(h->ev1, h->stream);
    B2_CUDA(h, cudaMemcpy2DAsync(gray_l, out_step, h->d_stage, h->stage_pitch, width, height, cudaMemcpyDeviceToHost, h->stream));
    B2_CUDA(h, cudaMemcpy2DAsync(gray_r, out_step, h->d_stage + h->stage_img_bytes, h->stage_pitch, width, height,
                                 cudaMemcpyDeviceToHost, h->stream));
    B2_CUDA(h, cudaStreamSynchronize(h->stream));
    cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1);
    return B2ORB_OK;
}

int b2orb_stereo_frontend_packed(b2orb_handle* h, const uint8_t* frames, int nframes, int src_w, int src_h, int src_step,
                                 size_t frame_stride, int width, int height, int temporal, const b2orb_stereo_out* out) {
    int rc = check_packed_args(h, frames, nframes, src_w, src_h, src_step, width, height);
    if (rc) return rc;
    if (!out) return fail(h, B2ORB_ERR_INVALID, "bad packed-frame arguments");
    B2_CUDA(h, cudaSetDevice(h->device));
    if ((rc = configure(h, width, height))) return rc;
    if ((rc = bind_problems(h, nframes))) return rc;
    cudaEventRecord(h->ev0, h->stream);
    if ((rc = ingest_packed(h, frames, nframes, src_w, src_h, src_step, frame_stride, width, height))) return rc;
    B2Images im = make_images(h, h->d_stage, h->d_stage + (size_t)nframes * h->stage_img_bytes, nframes,
                              h->stage_img_bytes, h->stage_pitch);
    if ((rc = launch_stereo(h, im, nframes, temporal))) return rc;
    cudaEventRecord(h->ev1, h->stream);
    rc = finish_outputs(h, nframes, out);
    cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1);
    return rc;
}

// ---- generic matcher entry points (host descriptor sets) ------------------------------------------
static int ensure_generic(b2orb_handle* h, int nq, int nt) {
    if (nq > h->g_cap_q) {
        const int c = std::max(nq, 2 * h->g_cap_q);
        if (!dev_alloc(h, &h->d_gq, (size_t)c * 32) || !dev_alloc(h, &h->d_gq2, (size_t)c * 32) ||
            !dev_alloc(h, &h->d_gidx, (size_t)c) || !dev_alloc(h, &h->d_gdist, (size_t)c) ||
            !dev_alloc(h, &h->d_gidx2, (size_t)c) || !dev_alloc(h, &h->d_gdist2, (size_t)c))
            return B2ORB_ERR_NOMEM;
        h->g_cap_q = c;
    }
    if (nt > h->g_cap_t) {
        const int c = std::max(nt, 2 * h->g_cap_t);
        if (!dev_alloc(h, &h->d_gt, (size_t)c * 32) || !dev_alloc(h, &h->d_gt2, (size_t)c * 32)) return B2ORB_ERR_NOMEM;
        h->g_cap_t = c;
    }
    return B2ORB_OK;
}

int b2orb_match(b2orb_handle* h, const uint8_t* query, int nq, const uint8_t* train, int nt, int32_t* train_idx,
                int32_t* dist) {
    if (!h || nq < 0 || nt < 0) return B2ORB_ERR_INVALID;
    if (nq == 0) return B2ORB_OK;
    if (!query || !train_idx || !dist || (nt > 0 && !train)) return fail(h, B2ORB_ERR_INVALID, "bad match arguments");
    if (nt >= (1 << 22)) return fail(h, B2ORB_ERR_INVALID, "train set larger than 2^22 rows");
    if (nt == 0) {   // BFMatcher on an empty train set returns no matches
        for (int i = 0; i < nq; ++i) { train_idx[i] = -1; dist[i] = 0x7fffffff; }
        return B2ORB_OK;
    }
    B2_CUDA(h, cudaSetDevice(h->device));
    int rc = ensure_generic(h, nq, nt);
    if (rc) return rc;
    cudaStream_t s = h->stream;
    const int cnt[2] = {nq, nt};
    B2MatchProblem pr = {h->d_gq, h->d_gt, h->d_gcnt, h->d_gcnt + 1, h->d_gidx, h->d_gdist};
    B2MatchProblem* slot = h->d_probs + (size_t)3 * h->npair_cap;
    cudaEventRecord(h->ev0, s);
    B2_CUDA(h, cudaMemcpyAsync(h->d_gq, query, (size_t)nq * 32, cudaMemcpyHostToDevice, s));
    B2_CUDA(h, cudaMemcpyAsync(h->d_gt, train, (size_t)nt * 32, cudaMemcpyHostToDevice, s));
    B2_CUDA(h, cudaMemcpyAsync(h->d_gcnt, cnt, sizeof cnt, cudaMemcpyHostToDevice, s));
    B2_CUDA(h, cudaMemcpyAsync(slot, &pr, sizeof pr, cudaMemcpyHostToDevice, s));
    b2_launch_match(slot, 1, nq, s);
    ++h->launches;
    B2_CUDA(h, cudaGetLastError());
    B2_CUDA(h, cudaMemcpyAsync(train_idx, h->d_gidx, (size_t)nq * 4, cudaMemcpyDeviceToHost, s));
    B2_CUDA(h, cudaMemcpyAsync(dist, h->d_gdist, (size_t)nq * 4, cudaMemcpyDeviceToHost, s));
    cudaEventRecord(h->ev1, s);
    B2_CUDA(h, cudaStreamSynchronize(s));
    cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1);
    return B2ORB_OK;
}

int b2orb_temporal_track(b2orb_handle* h, const uint8_t* cur_l, const uint8_t* cur_r, int ncur, const uint8_t* hist_l,
                         const uint8_t* hist_r, int nhist, int32_t* circ_cur, int32_t* circ_hist, int32_t* n_circ,
                         int32_t* n_left, int32_t* n_right) {
    if (!h || !n_circ || ncur < 0 || nhist < 0) return B2ORB_ERR_INVALID;
    *n_circ = 0;
    if (n_left) *n_left = 0;
    if (n_right) *n_right = 0;
    if (ncur == 0 || nhist == 0) return B2ORB_OK;   // externalTrack early-out, src/feature-tracker.cpp:237-240
    if (!cur_l || !cur_r || !hist_l || !hist_r || !circ_cur || !circ_hist) return fail(h, B2ORB_ERR_INVALID, "bad temporal arguments");
    if (nhist >= (1 << 22)) return fail(h, B2ORB_ERR_INVALID, "history larger than 2^22 rows");
    B2_CUDA(h, cudaSetDevice(h->device));
    int rc = ensure_generic(h, ncur, nhist);
    if (rc) return rc;
    cudaStream_t s = h->stream;
    const int cnt[2] = {ncur, nhist};
    B2MatchProblem pr[2] = {{h->d_gq, h->d_gt, h->d_gcnt, h->d_gcnt + 1, h->d_gidx, h->d_gdist},
                            {h->d_gq2, h->d_gt2, h->d_gcnt, h->d_gcnt + 1, h->d_gidx2, h->d_gdist2}};
    B2MatchProblem* slot = h->d_probs + (size_t)3 * h->npair_cap + 1;
    int* d_out = h->d_gcnt + 2;   // n_circ, n_left, n_right
    cudaEventRecord(h->ev0, s);
    B2_CUDA(h, cudaMemcpyAsync(h->d_gq, cur_l, (size_t)ncur * 32, cudaMemcpyHostToDevice, s));
    B2_CUDA(h, cudaMemcpyAsync(h->d_gq2, cur_r, (size_t)ncur * 32, cudaMemcpyHostToDevice, s));
    B2_CUDA(h, cudaMemcpyAsync(h->d_gt, hist_l, (size_t)nhist * 32, cudaMemcpyHostToDevice, s));
    B2_CUDA(h, cudaMemcpyAsync(h->d_gt2, hist_r, (size_t)nhist * 32, cudaMemcpyHostToDevice, s));
    B2_CUDA(h, cudaMemcpyAsync(h->d_gcnt, cnt, sizeof cnt, cudaMemcpyHostToDevice, s));
    B2_CUDA(h, cudaMemcpyAsync(slot, pr, sizeof pr, cudaMemcpyHostToDevice, s));
    b2_launch_match(slot, 2, ncur, s);
    ++h->launches;
    // circular check reuses the cur-side index buffers as outputs' staging: circ lists go to d_gdist/d_gdist2
    // only after the kernel has read them, so use separate regions: the q buffers are free after matching.
    int* d_cc = reinterpret_cast<int*>(h->d_gq);
    int* d_ch = reinterpret_cast<int*>(h->d_gq2);
    b2_launch_circular(h->d_gidx, h->d_gdist, h->d_gidx2, h->d_gdist2, h->d_gcnt, h->d_gcnt + 1, 1, 0, h->cfg.match_ratio,
                       h->cfg.min_match_dist, d_cc, d_ch, d_out, d_out + 1, d_out + 2, nullptr, 1, s);
    ++h->launches;
    B2_CUDA(h, cudaGetLastError());
    int out[3] = {0, 0, 0};
    B2_CUDA(h, cudaMemcpyAsync(out, d_out, sizeof out, cudaMemcpyDeviceToHost, s));
    cudaEventRecord(h->ev1, s);
    B2_CUDA(h, cudaStreamSynchronize(s));
    cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1);
    *n_circ = out[0];
    if (n_left) *n_left = out[1];
    if (n_right) *n_right = out[2];
    if (out[0] > 0) {
        B2_CUDA(h, cudaMemcpyAsync(circ_cur, d_cc, (size_t)out[0] * 4, cudaMemcpyDeviceToHost, s));
        B2_CUDA(h, cudaMemcpyAsync(circ_hist, d_ch, (size_t)out[0] * 4, cudaMemcpyDeviceToHost, s));
        B2_CUDA(h, cudaStreamSynchronize(s));
    }
    return B2ORB_OK;
}

// ---- allocation-free quadtree (no counterpart in the reference) ------
int b2orb_distribute_static(b2orb_handle* h, int level, const uint32_t* d_cand, int n_cand, uint32_t* d_sel, int32_t* d_n_sel) {
    if (!h) return B2ORB_ERR_INVALID;
    if (level < 0 || level >= h->geom.nlevels || !d_cand || !d_sel || !d_n_sel || n_cand < 0)
        return fail(h, B2ORB_ERR_INVALID, "bad b2orb_distribute_static arguments");
    const B2Level& L = h->geom.lv[level];
    if (L.quota > B2ORB_STATIC_MAX_FEATURES || L.node_cap > B2_STATIC_MAX_NODES)
        return fail(h, B2ORB_ERR_CAPACITY, "b2orb_distribute_static: level quota exceeds B2ORB_STATIC_MAX_FEATURES (4096)");
    if (n_cand > B2ORB_STATIC_MAX_CANDIDATES)
        return fail(h, B2ORB_ERR_CAPACITY, "b2orb_distribute_static: n_cand exceeds B2ORB_STATIC_MAX_CANDIDATES (262144)");
    if (n_cand > L.cand_cap) return fail(h, B2ORB_ERR_CAPACITY, "b2orb_distribute_static: more candidates than the level can hold");
    B2_CUDA(h, cudaSetDevice(h->device));
    B2Geom g1 = h->geom;                    // one-level geometry: the level's slots start at 0
    g1.nlevels = 1;
    g1.lv[0] = L;
    g1.lv[0].cand_off = 0; g1.lv[0].node_off = 0; g1.lv[0].sel_base = 0;
    g1.cand_total = (unsigned)L.cand_cap;
    g1.node_total = (unsigned)L.node_cap;
    g1.sel_total = L.sel_cap;
    const int e = b2_launch_octree_static(g1, d_cand, n_cand, d_sel, d_n_sel, h->stream);
    h->launches += 2;
    if (e) { cuda_ok(h, (cudaError_t)e, "b2orb_distribute_static"); return B2ORB_ERR_CUDA; }
    return B2ORB_OK;
}

// ---- next row f3: device-resident feature pool ------------------------------------------------------
// featurePoolUpdate / externalTrack bookkeeping (src/feature-tracker.cpp:231-328, 330-412): descriptors of pooled
// features stay on the device (pool.cu); per frame only index lists, a byte mask and counts cross the bus.
int b2orb_pool_reset(b2orb_handle* h, int max_feature_age) {
    if (!h || h->root != h || max_feature_age < 0) return B2ORB_ERR_INVALID;
    B2_CUDA(h, cudaSetDevice(h->device));
    B2_CUDA(h, cudaStreamSynchronize(h->stream));
    B2Pool* p = pool_of(h);
    if (!p) return fail(h, B2ORB_ERR_NOMEM, "feature pool");
    b2_pool_reset(p);
    h->pool_max_age = max_feature_age;
    h->trk_ncur = -1;
    h->trk_owner = nullptr;
    h->trk_circ_hist.clear(); h->trk_mask.clear(); h->trk_flags.clear(); h->trk_ids.clear();
    return B2ORB_OK;
}

int b2orb_pool_size(b2orb_handle* h) { return h && h->root->pool ? b2_pool_size(h->root->pool) : 0; }

int b2orb_pool_get(b2orb_handle* h, uint64_t* ids, int32_t* ages, uint8_t* hist_l, uint8_t* hist_r, int cap, int* n) {
    if (!h || !n) return B2ORB_ERR_INVALID;
    B2Pool* p = pool_of(h);
    if (!p) return fail(h, B2ORB_ERR_NOMEM, "feature pool");
    const int m = b2_pool_size(p);
    *n = m;
    if (m > cap) return (ids || ages || hist_l || hist_r) ? fail(h, B2ORB_ERR_CAPACITY, "pool larger than the caller's arrays") : B2ORB_OK;
    const unsigned long long* pid = b2_pool_ids(p);
    for (int i = 0; i < m; ++i) {
        if (ids) ids[i] = pid[i];
        if (ages) ages[i] = b2_pool_age(p, pid[i]);
    }
    B2_CUDA(h, cudaSetDevice(h->device));
    if (m && hist_l) B2_CUDA(h, cudaMemcpyAsync(hist_l, b2_pool_rows(p, 0), (size_t)m * 32, cudaMemcpyDeviceToHost, h->stream));
    if (m && hist_r) B2_CUDA(h, cudaMemcpyAsync(hist_r, b2_pool_rows(p, 1), (size_t)m * 32, cudaMemcpyDeviceToHost, h->stream));
    B2_CUDA(h, cudaStreamSynchronize(h->stream));
    return B2ORB_OK;
}

// Host copies of what featurePoolUpdate needs from the tracked pair (a few KB): circular trainIdx list, mask, depth
// flags.  `o` = the handle (root or lane) whose workspace holds the pair.
static int pool_fetch_track(b2orb_handle* o, int pair) {
    b2orb_handle* r = o->root;
    const int kcap = o->geom.kp_cap;
    int cnt[2] = {0, 0};
    B2_CUDA(o, cudaMemcpyAsync(&cnt[0], o->d_nmatch + 1 + pair, 4, cudaMemcpyDeviceToHost, o->stream));
    B2_CUDA(o, cudaMemcpyAsync(&cnt[1], o->d_ncirc + pair, 4, cudaMemcpyDeviceToHost, o->stream));
    B2_CUDA(o, cudaStreamSynchronize(o->stream));
    const int n = std::min(std::max(cnt[0], 0), kcap), nc = std::min(std::max(cnt[1], 0), kcap);
    r->trk_circ_hist.resize(nc);
    r->trk_mask.resize(n);
    r->trk_flags.assign(n, 1);
    if (nc) B2_CUDA(o, cudaMemcpyAsync(r->trk_circ_hist.data(), o->d_circ_hist + (size_t)pair * kcap, (size_t)nc * 4, cudaMemcpyDeviceToHost, o->stream));
    if (n) B2_CUDA(o, cudaMemcpyAsync(r->trk_mask.data(), o->d_mask + (size_t)pair * kcap, n, cudaMemcpyDeviceToHost, o->stream));
    if (n && r->proj_on) B2_CUDA(o, cudaMemcpyAsync(r->trk_flags.data(), o->d_dflag + (size_t)pair * kcap, n, cudaMemcpyDeviceToHost, o->stream));
    B2_CUDA(o, cudaStreamSynchronize(o->stream));
    r->trk_ids.resize(nc);                 // _matchedFeatureIDs.push_back(_histFeatureIDs[m.trainIdx]), :315-317
    const unsigned long long* pid = b2_pool_ids(pool_of(o));
    const int psz = b2_pool_size(pool_of(o));
    for (int i = 0; i < nc; ++i) r->trk_ids[i] = r->trk_circ_hist[i] >= 0 && r->trk_circ_hist[i] < psz ? pid[r->trk_circ_hist[i]] : ~0ull;
    r->trk_ncur = n;
    r->trk_pair = pair;
    r->trk_owner = o;
    return B2ORB_OK;
}

// Results of the last pool track (b2orb_pool_track, or b2orb_stereo_frontend with temporal = 2).
int b2orb_pool_last_track(b2orb_handle* h, int32_t* circ_cur, int32_t* circ_hist, int32_t* n_circ, uint8_t* cur_mask,
                          uint64_t* matched_ids, int32_t* n_cur, int32_t* n_left, int32_t* n_right) {
    if (!h || h->root != h || !n_circ) return B2ORB_ERR_INVALID;
    B2_CUDA(h, cudaSetDevice(h->device));
    b2orb_handle* o = h->trk_owner ? h->trk_owner : h;
    if (h->trk_ncur < 0) {   // a fused call (temporal = 2) left the lists on the device: fetch pair 0
        const int rc = pool_fetch_track(o, 0);
        if (rc) { h->err = o->err; return rc; }
    }
    B2Pool* p = pool_of(h);
    if (!p) return fail(h, B2ORB_ERR_NOMEM, "feature pool");
    const int kcap = o->geom.kp_cap, pair = h->trk_pair;
    const int nc = (int)h->trk_circ_hist.size(), n = h->trk_ncur;
    *n_circ = nc;
    if (n_cur) *n_cur = n;
    for (int i = 0; i < nc; ++i) {
        if (circ_hist) circ_hist[i] = h->trk_circ_hist[i];
        if (matched_ids) matched_ids[i] = h->trk_ids[i];
    }
    if (cur_mask && n) memcpy(cur_mask, h->trk_mask.data(), n);
    int lr[2] = {0, 0};
    if (nc && circ_cur) B2_CUDA(h, cudaMemcpyAsync(circ_cur, o->d_circ_cur + (size_t)pair * kcap, (size_t)nc * 4, cudaMemcpyDeviceToHost, o->stream));
    if (n_left) B2_CUDA(h, cudaMemcpyAsync(&lr[0], o->d_nleft + pair, 4, cudaMemcpyDeviceToHost, o->stream));
    if (n_right) B2_CUDA(h, cudaMemcpyAsync(&lr[1], o->d_nright + pair, 4, cudaMemcpyDeviceToHost, o->stream));
    B2_CUDA(h, cudaStreamSynchronize(o->stream));
    if (n_left) *n_left = lr[0];
    if (n_right) *n_right = lr[1];
    return B2ORB_OK;
}

int b2orb_pool_track(b2orb_handle* h, int pair, int32_t* circ_cur, int32_t* circ_hist, int32_t* n_circ, uint8_t* cur_mask,
                     uint64_t* matched_ids, int32_t* n_cur, int32_t* n_left, int32_t* n_right) {
    if (!h || h->root != h || pair < 0 || pair >= h->npair_cap || !n_circ) return B2ORB_ERR_INVALID;
    B2_CUDA(h, cudaSetDevice(h->device));
    B2Pool* p = pool_of(h);
    if (!p) return fail(h, B2ORB_ERR_NOMEM, "feature pool");
    const int kcap = h->geom.kp_cap;
    const size_t row = (size_t)kcap * 32;
    cudaStream_t s = h->stream;
    // two problems in the generic slots: cur (kept rows of `pair`) against the pool, left and right
    const uint8_t* pl = b2_pool_rows(p, 0);
    const uint8_t* pr_ = b2_pool_rows(p, 1);
    B2MatchProblem pr[2] = {
        {h->d_cur_l + (pair + 1) * row, pl ? pl : h->d_cur_l, h->d_nmatch + 1 + pair, b2_pool_dcount(p), h->d_tti_l + (size_t)pair * kcap, h->d_tdi_l + (size_t)pair * kcap},
        {h->d_cur_r + (pair + 1) * row, pr_ ? pr_ : h->d_cur_r, h->d_nmatch + 1 + pair, b2_pool_dcount(p), h->d_tti_r + (size_t)pair * kcap, h->d_tdi_r + (size_t)pair * kcap}};
    B2MatchProblem* slot = h->d_probs + (size_t)3 * h->npair_cap + 1;
    b2_pdl_on = false;
    cudaEventRecord(h->ev0, s);
    B2_CUDA(h, cudaMemcpyAsync(slot, pr, sizeof pr, cudaMemcpyHostToDevice, s));
    prof_mark(h, B2ORB_K_MATCH_TEMPORAL);
    b2_launch_match(slot, 2, kcap, s);
    prof_mark(h, B2ORB_K_CIRCULAR);
    b2_launch_circular(h->d_tti_l + (size_t)pair * kcap, h->d_tdi_l + (size_t)pair * kcap, h->d_tti_r + (size_t)pair * kcap,
                       h->d_tdi_r + (size_t)pair * kcap, h->d_nmatch + 1 + pair, b2_pool_dcount(p), 0, 0, h->cfg.match_ratio,
                       h->cfg.min_match_dist, h->d_circ_cur + (size_t)pair * kcap, h->d_circ_hist + (size_t)pair * kcap,
                       h->d_ncirc + pair, h->d_nleft + pair, h->d_nright + pair, h->d_mask + (size_t)pair * kcap, 1, s);
    h->launches += 2;
    prof_mark(h, -1);
    cudaEventRecord(h->ev1, s);
    B2_CUDA(h, cudaGetLastError());
    B2_CUDA(h, cudaStreamSynchronize(s));   // `pr` is a stack array: its copy must have left the host
    const int rc = pool_fetch_track(h, pair);
    if (rc) return rc;
    return b2orb_pool_last_track(h, circ_cur, circ_hist, n_circ, cur_mask, matched_ids, n_cur, n_left, n_right);
}

int b2orb_pool_update(b2orb_handle* h, int is_keyframe, const uint8_t* accept, uint64_t* new_ids, int32_t* n_inserted,
                      int32_t* n_erased) {
    if (!h || h->root != h) return B2ORB_ERR_INVALID;
    B2_CUDA(h, cudaSetDevice(h->device));
    B2Pool* p = pool_of(h);
    if (!p) return fail(h, B2ORB_ERR_NOMEM, "feature pool");
    b2orb_handle* o = h->trk_owner ? h->trk_owner : h;
    if (h->trk_ncur < 0) {
        const int rc = pool_fetch_track(o, 0);
        if (rc) { h->err = o->err; return rc; }
    }
    const size_t row = (size_t)o->geom.kp_cap * 32;
    const int n = h->trk_ncur;
    // :374 `depth > _maxDepth` -> skip: the caller's flags, else bit 0 of the device depth flags when a projection is set
    std::vector<uint8_t> acc;
    if (!accept && h->proj_on) {
        acc.resize(n);
        for (int i = 0; i < n; ++i) acc[i] = h->trk_flags[i] & 1;
        accept = acc.data();
    }
    int ins = 0, er = 0;
    static_assert(sizeof(unsigned long long) == sizeof(uint64_t), "id type");
    const int e = b2_pool_update(p, o->stream, is_keyframe != 0, h->pool_max_age, h->trk_ids.data(), (int)h->trk_ids.size(),
                                 h->trk_mask.data(), accept, n, o->d_cur_l + (h->trk_pair + 1) * row, o->d_cur_r + (h->trk_pair + 1) * row,
                                 reinterpret_cast<unsigned long long*>(new_ids), &ins, &er);
    if (e) { cuda_ok(h, (cudaError_t)e, "feature pool update"); return B2ORB_ERR_CUDA; }
    if (n_inserted) *n_inserted = ins;
    if (n_erased) *n_erased = er;
    if (b2_pool_size(p) > 0 && (ins || er || is_keyframe)) ++h->launches;
    return B2ORB_OK;
}

// ---- stand-alone stereo filter (src/feature-tracker.cpp:164, 208-218 on host arrays) -----------------
int b2orb_stereo_filter(b2orb_handle* h, const int32_t* train_idx, const int32_t* dist, int nq, const b2orb_keypoint* kp_l,
                        const uint8_t* desc_l, const b2orb_keypoint* kp_r, const uint8_t* desc_r, int nt, int32_t* match_l,
                        int32_t* match_r, int32_t* n_match, double* pix_l, double* pix_r, uint8_t* cur_desc_l,
                        uint8_t* cur_desc_r) {
    if (!h || !n_match || nq < 0 || nt < 0) return B2ORB_ERR_INVALID;
    *n_match = 0;
    if (nq == 0 || nt == 0) return B2ORB_OK;   // min_element on an empty vector (:164): defined as "no matches"
    if (!train_idx || !dist || !kp_l || !kp_r || !match_l || !match_r || ((cur_desc_l || cur_desc_r) && (!desc_l || !desc_r)))
        return fail(h, B2ORB_ERR_INVALID, "bad stereo filter arguments");
    const int kcap = h->geom.kp_cap;
    if (nq > kcap || nt > kcap) return fail(h, B2ORB_ERR_CAPACITY, "more rows than b2orb_keypoint_capacity()");
    if (h->nimg_cap < 2) return fail(h, B2ORB_ERR_CAPACITY, "handle needs max_images >= 2");
    for (int i = 0; i < nq; ++i)
        if (train_idx[i] < 0 || train_idx[i] >= nt) return fail(h, B2ORB_ERR_INVALID, "train_idx out of range");
    B2_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t s = h->stream;
    const int cnt[2] = {nq, nt};
    // image slots 0 (left) and 1 (right) of a one-pair batch
    B2_CUDA(h, cudaMemcpyAsync(h->d_kps, kp_l, (size_t)nq * sizeof(b2orb_keypoint), cudaMemcpyHostToDevice, s));
    B2_CUDA(h, cudaMemcpyAsync(h->d_kps + kcap, kp_r, (size_t)nt * sizeof(b2orb_keypoint), cudaMemcpyHostToDevice, s));
    if (desc_l) B2_CUDA(h, cudaMemcpyAsync(h->d_desc, desc_l, (size_t)nq * 32, cudaMemcpyHostToDevice, s));
    if (desc_r) B2_CUDA(h, cudaMemcpyAsync(h->d_desc + (size_t)kcap * 32, desc_r, (size_t)nt * 32, cudaMemcpyHostToDevice, s));
    B2_CUDA(h, cudaMemcpyAsync(h->d_train, train_idx, (size_t)nq * 4, cudaMemcpyHostToDevice, s));
    B2_CUDA(h, cudaMemcpyAsync(h->d_dist, dist, (size_t)nq * 4, cudaMemcpyHostToDevice, s));
    B2_CUDA(h, cudaMemcpyAsync(h->d_nkp, cnt, sizeof cnt, cudaMemcpyHostToDevice, s));
    b2_pdl_on = false;
    prof_mark(h, B2ORB_K_STEREO_FILTER);
    b2_launch_stereo_filter(h->d_kps, h->d_desc, h->d_nkp, h->d_train, h->d_dist, kcap, h->cfg.match_ratio, h->cfg.min_match_dist,
                            h->cfg.max_vertical_pixel_dist, h->d_match_l, h->d_match_r, h->d_nmatch + 1,
                            h->d_cur_l + (size_t)kcap * 32, h->d_cur_r + (size_t)kcap * 32, h->d_pix_l, h->d_pix_r, 1, s);
    ++h->launches;
    prof_mark(h, -1);
    B2_CUDA(h, cudaGetLastError());
    int m = 0;
    B2_CUDA(h, cudaMemcpyAsync(&m, h->d_nmatch + 1, 4, cudaMemcpyDeviceToHost, s));
    B2_CUDA(h, cudaStreamSynchronize(s));
    *n_match = m;
    if (m > 0) {
        B2_CUDA(h, cudaMemcpyAsync(match_l, h->d_match_l, (size_t)m * 4, cudaMemcpyDeviceToHost, s));
        B2_CUDA(h, cudaMemcpyAsync(match_r, h->d_match_r, (size_t)m * 4, cudaMemcpyDeviceToHost, s));
        if (pix_l) B2_CUDA(h, cudaMemcpyAsync(pix_l, h->d_pix_l, (size_t)m * 16, cudaMemcpyDeviceToHost, s));
        if (pix_r) B2_CUDA(h, cudaMemcpyAsync(pix_r, h->d_pix_r, (size_t)m * 16, cudaMemcpyDeviceToHost, s));
        if (cur_desc_l) B2_CUDA(h, cudaMemcpyAsync(cur_desc_l, h->d_cur_l + (size_t)kcap * 32, (size_t)m * 32, cudaMemcpyDeviceToHost, s));
        if (cur_desc_r) B2_CUDA(h, cudaMemcpyAsync(cur_desc_r, h->d_cur_r + (size_t)kcap * 32, (size_t)m * 32, cudaMemcpyDeviceToHost, s));
        B2_CUDA(h, cudaStreamSynchronize(s));
    }
    return B2ORB_OK;
}

// ---- asynchronous extraction (two instances driven from one or two threads, src/feature-tracker.cpp:153-156) -----
int b2orb_extract_async(b2orb_handle* h, const uint8_t* image, int width, int height, int step) {
    if (!h) return B2ORB_ERR_INVALID;
    h->async_count = 0;
    if (width == 0 || height == 0) { h->async_count = -1; return B2ORB_OK; }   // empty image: wait reports 0 keypoints
    if (!image || width < 0 || height < 0 || step < width) return fail(h, B2ORB_ERR_INVALID, "bad image arguments");
    if (width > h->cfg.max_width || height > h->cfg.max_height) return fail(h, B2ORB_ERR_INVALID, "image larger than configured maximum");
    B2_CUDA(h, cudaSetDevice(h->device));
    int rc = configure(h, width, height);
    if (rc) return rc;
    if (!h->h_async_n) B2_CUDA(h, cudaHostAlloc(&h->h_async_n, sizeof(int32_t), cudaHostAllocDefault));
    cudaEventRecord(h->ev0, h->stream);
    if ((rc = upload(h, image, 1, width, height, step, (size_t)step * height, 0))) return rc;
    B2Images im = make_images(h, h->d_stage, h->d_stage, 1, h->stage_img_bytes, h->stage_pitch);
    if ((rc = launch_extract(h, im, 1))) return rc;
    B2_CUDA(h, cudaMemcpyAsync(h->h_async_n, h->d_nkp, sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
    cudaEventRecord(h->ev1, h->stream);
    h->async_count = 1;
    return B2ORB_OK;
}

int b2orb_extract_wait(b2orb_handle* h, b2orb_keypoint* kps, uint8_t* desc, int cap, int* n) {
    if (!h || !n) return B2ORB_ERR_INVALID;
    *n = 0;
    if (h->async_count == -1) { h->async_count = 0; return B2ORB_OK; }
    if (h->async_count != 1) return fail(h, B2ORB_ERR_INVALID, "b2orb_extract_wait without b2orb_extract_async");
    h->async_count = 0;
    B2_CUDA(h, cudaSetDevice(h->device));
    B2_CUDA(h, cudaStreamSynchronize(h->stream));
    cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1);
    const int m = *h->h_async_n;
    if (m > cap) return fail(h, B2ORB_ERR_CAPACITY, "keypoint buffer too small");
    if (m > 0) {
        if (!kps || !desc) return fail(h, B2ORB_ERR_INVALID, "bad output arguments");
        B2_CUDA(h, cudaMemcpyAsync(kps, h->d_kps, (size_t)m * sizeof(b2orb_keypoint), cudaMemcpyDeviceToHost, h->stream));
        B2_CUDA(h, cudaMemcpyAsync(desc, h->d_desc, (size_t)m * 32, cudaMemcpyDeviceToHost, h->stream));
        B2_CUDA(h, cudaStreamSynchronize(h->stream));
    }
    *n = m;
    return B2ORB_OK;
}

// ---- per-handle launch-time options (tests, A/B runs) ---------------------------------------------
int b2orb_debug_set(b2orb_handle* h, const char* key, int value) {
    if (!h || !key || h->root != h) return B2ORB_ERR_INVALID;
    B2Tuning& t = h->tune;
    const std::string k(key);
    if (k == "pdl") t.pdl = value < 0 ? -1 : (value ? 1 : 0);
    else if (k == "fork_max") t.fork_max = value;
    else if (k == "export_max") t.export_max = value;
    else if (k == "fast_band") t.fast_band = value;
    else if (k == "blur_band") t.blur_band = value;
    else if (k == "octree_points") t.octree_points = value ? 1 : 0;
    else if (k == "describe_tma") t.describe_tma = value ? 1 : 0;
    else if (k == "ingest_host") t.ingest_host = value ? 1 : 0;
    else if (k == "pipe_trace") t.pipe_trace = value;
    else if (k == "pipe_chunk" || k == "pipe_lanes") {
        if (!h->lanes.empty()) return fail(h, B2ORB_ERR_INVALID, "pipeline options must be set before the first pipelined call");
        if (k == "pipe_chunk") t.pipe_chunk = std::max(0, value);
        else t.pipe_lanes = std::min(std::max(2, value), 8);
    } else return fail(h, B2ORB_ERR_INVALID, "unknown option");
    return B2ORB_OK;
}

// ---- stream + per-kernel timing ------------------------------------------------------------------
uint64_t b2orb_debug_octree_fallbacks(b2orb_handle* h) {
    if (!h || cudaSetDevice(h->device) != cudaSuccess) return 0;
    cudaStreamSynchronize(h->stream);
    return b2_octree_fallbacks();
}

void* b2orb_stream(const b2orb_handle* h) { return h ? (void*)h->stream : nullptr; }

void b2orb_profile_enable(b2orb_handle* h, int on) {
    if (!h) return;
    h->profiling = on != 0;
    h->prof_used = 0;
    for (int k = 0; k < B2ORB_NUM_KERNELS; ++k) { h->prof_ms[k] = 0; h->prof_cnt[k] = 0; }
}

int b2orb_profile_read(b2orb_handle* h, double* ms, uint64_t* launches) {
    if (!h || !ms || !launches) return B2ORB_ERR_INVALID;
    B2_CUDA(h, cudaSetDevice(h->device));
    B2_CUDA(h, cudaStreamSynchronize(h->stream));
    for (size_t i = 0; i + 1 < h->prof_used; ++i) {
        const int tag = h->prof_tag[i];
        if (tag < 0) continue;
        float t = 0.f;
        if (cudaEventElapsedTime(&t, h->prof_ev[i], h->prof_ev[i + 1]) == cudaSuccess) {
            h->prof_ms[tag] += t;
            h->prof_cnt[tag] += 1;
        }
    }
    h->prof_used = 0;
    for (int k = 0; k < B2ORB_NUM_KERNELS; ++k) { ms[k] = h->prof_ms[k]; launches[k] = h->prof_cnt[k]; }
    return B2ORB_OK;
}

// ---- debug / test access -----------------------------------------------------------------------
int b2orb_debug_level_size(const b2orb_handle* h, int level, int* w, int* hgt) {
    if (!h || level < 0 || level >= h->geom.nlevels || !w || !hgt) return B2ORB_ERR_INVALID;
    *w = h->geom.lv[level].w;
    *hgt = h->geom.lv[level].h;
    return B2ORB_OK;
}

int b2orb_debug_get_level(b2orb_handle* h, int slot, int level, int which, uint8_t* dst) {
    if (!h || !dst || slot < 0 || slot >= h->nimg_cap || level < 0 || level >= h->geom.nlevels) return B2ORB_ERR_INVALID;
    B2_CUDA(h, cudaSetDevice(h->device));
    const B2Level& L = h->geom.lv[level];
    const uint8_t* src;
    size_t pitch;
    if (which == 0 && level == 0) {   // level 0 of the last host-input call lives in the staging area
        src = h->d_stage + (size_t)slot * h->stage_img_bytes;
        pitch = h->stage_pitch;
    } else {
        src = (which ? h->d_blur : h->d_pyr) + (size_t)slot * h->geom.slab + L.off;
        pitch = L.pitch;
    }
    B2_CUDA(h, cudaMemcpy2DAsync(dst, L.w, src, pitch, L.w, L.h, cudaMemcpyDeviceToHost, h->stream));
    B2_CUDA(h, cudaStreamSynchronize(h->stream));
    return B2ORB_OK;
}

int b2orb_debug_get_candidates(b2orb_handle* h, int slot, int level, uint32_t* dst, int cap, int* n) {
    if (!h || !n || slot < 0 || slot >= h->nimg_cap || level < 0 || level >= h->geom.nlevels) return B2ORB_ERR_INVALID;
    B2_CUDA(h, cudaSetDevice(h->device));
    int cnt = 0;
    B2_CUDA(h, cudaMemcpyAsync(&cnt, h->d_cand_cnt + slot * h->geom.nlevels + level, 4, cudaMemcpyDeviceToHost, h->stream));
    B2_CUDA(h, cudaStreamSynchronize(h->stream));
    *n = cnt;
    const int m = std::min(std::min(cnt, cap), h->geom.lv[level].cand_cap);
    if (m > 0 && dst) {
        B2_CUDA(h, cudaMemcpyAsync(dst, h->d_cand + (size_t)slot * h->geom.cand_total + h->geom.lv[level].cand_off,
                                   (size_t)m * 4, cudaMemcpyDeviceToHost, h->stream));
        B2_CUDA(h, cudaStreamSynchronize(h->stream));
    }
    return B2ORB_OK;
}

}  // extern "C"
