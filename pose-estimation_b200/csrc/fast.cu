// Grid-cell FAST-9/16 with threshold fallback — ORBextractor::ComputeKeyPointsOctTree cell loop,
// src/ORBextractor.cc:778-829 (K3).  The reference calls cv::FAST(cellROI, iniThFAST, nms=true) on every
// ~30x30 cell (ROI = cell + 3-px inset on each side) and re-runs it with minThFAST when the cell returned
// nothing.  Facts used (checked against cv2 in tests/):
//   * corner at threshold t  <=>  b(p) > t, where
//       b(p) = max over the 16 arcs of 9 contiguous circle pixels of max(min_k d_k, min_k(-d_k)),
//       d_k = I(p) - I(p + o_k);  cv::FAST's score is b - 1, independent of t;
//   * with nms a corner survives iff its score is strictly larger than its 8 neighbours', where
//     neighbours that are not corners or lie outside the ROI's detect area count as 0 — equivalently
//     b(p) > b(q) for the neighbours q inside the cell's own detect area;
//   * detect areas of neighbouring cells tile the bordered rectangle without overlap.
// One CTA owns one cell: it stages the cell + 3-px halo in shared memory as u16x2 pixel pairs (pixel x
// and pixel x + half-cell-width share a 32-bit word, so consecutive lanes hit consecutive banks), computes
// b for two pixels per instruction with the native VIMNMX/VIMNMX3 .U16x2 min/max, does the cell-local
// non-max suppression, votes on the threshold (any survivor with b > ini ? ini : min) and appends the
// survivors to the level's candidate list.  Candidate order is irrelevant downstream: the quadtree breaks
// response ties with a key that reproduces the reference's cell-row-major order.
#include "b2orb_internal.cuh"

namespace {

constexpr int FT_THREADS = 128;
constexpr int FT_MAX_CELL = 64;                         // cells are < 60 px on a side (ceil(W'/floor(W'/30)))
constexpr int FT_MAX_HALF = FT_MAX_CELL / 2;
constexpr int FT_TILE_WORDS = (FT_MAX_CELL + 6) * (FT_MAX_HALF + 6);
constexpr int FT_ITEMS_PER_THREAD = 32;                 // >= ceil(64*64 / 128)

__device__ __forceinline__ uint32_t umin2(uint32_t a, uint32_t b) { return __vminu2(a, b); }
__device__ __forceinline__ uint32_t umax2(uint32_t a, uint32_t b) { return __vmaxu2(a, b); }
__device__ __forceinline__ uint32_t umin3(uint32_t a, uint32_t b, uint32_t c) { return __vimin3_u16x2(a, b, c); }
__device__ __forceinline__ uint32_t umax3(uint32_t a, uint32_t b, uint32_t c) { return __vimax3_u16x2(a, b, c); }

// b for the two pixels packed in `c` (u16 lanes), ring pixels in r[0..15] (same packing).
// Biased differences d' = 256 + I(p) - I(q) keep both lanes positive so one 32-bit subtract serves
// both; max over arcs of min over 9 uses sliding min3-of-min3.
__device__ __forceinline__ uint32_t fast_b_pair(uint32_t c, const uint32_t (&r)[16]) {
    uint32_t d[16], m3[16], M3[16];
    const uint32_t cb = c + 0x01000100u;
#pragma unroll
    for (int k = 0; k < 16; ++k) d[k] = cb - r[k];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        m3[k] = umin3(d[k], d[(k + 1) & 15], d[(k + 2) & 15]);
        M3[k] = umax3(d[k], d[(k + 1) & 15], d[(k + 2) & 15]);
    }
    uint32_t A = 0, B = 0xFFFFFFFFu;  // A = max_arcs min9(d'), B = min_arcs max9(d')
#pragma unroll
    for (int k = 0; k < 16; k += 2) {
        uint32_t a0 = umin3(m3[k], m3[(k + 3) & 15], m3[(k + 6) & 15]);
        uint32_t a1 = umin3(m3[(k + 1) & 15], m3[(k + 4) & 15], m3[(k + 7) & 15]);
        A = umax3(A, a0, a1);
        uint32_t b0 = umax3(M3[k], M3[(k + 3) & 15], M3[(k + 6) & 15]);
        uint32_t b1 = umax3(M3[(k + 1) & 15], M3[(k + 4) & 15], M3[(k + 7) & 15]);
        B = umin3(B, b0, b1);
    }
    // b = max(A' - 256, 256 - B', 0) per lane
    const uint32_t bright = __vsubus2(A, 0x01000100u);
    const uint32_t dark = __vsubus2(0x01000100u, B);
    return umax2(bright, dark);
}

__global__ void __launch_bounds__(FT_THREADS)
fast_cells_kernel(const __grid_constant__ B2Geom g, const B2Images im, uint32_t* __restrict__ cand,
                  int* __restrict__ cand_cnt) {
    __shared__ uint32_t s_tile[FT_TILE_WORDS];                    // packed pixel pairs, cell + halo
    __shared__ uint8_t s_score[(FT_MAX_CELL + 2) * (FT_MAX_CELL + 4)];  // b per pixel, 1-px zero ring
    __shared__ int s_base;

    const int b = blockIdx.y;
    int level = 0;
    while (level + 1 < g.nlevels && (int)blockIdx.x >= g.lv[level + 1].cell_base) ++level;
    const B2Level& L = g.lv[level];
    const int cell = blockIdx.x - L.cell_base;
    const int ci = cell / L.n_cols, cj = cell - ci * L.n_cols;
    // detect area of this cell in level coordinates (ROI inset by 3, clipped to maxBorder-3 = dim-19)
    const int x0 = B2_EDGE + cj * L.w_cell, y0 = B2_EDGE + ci * L.h_cell;
    const int x1 = min(x0 + L.w_cell, L.w - B2_EDGE), y1 = min(y0 + L.h_cell, L.h - B2_EDGE);
    const int cw = x1 - x0, ch = y1 - y0;
    if (cw <= 0 || ch <= 0) return;
    const int half = (cw + 1) >> 1;
    const int tw = half + 6;            // packed tile width in words
    const int sp = cw + 2;              // score tile pitch
    int pitch;
    const uint8_t* img = b2_level_ptr(g, im, b, level, pitch);
    const int tid = threadIdx.x;

    // stage: word (r,c) = I[y0-3+r][x0-3+c] | I[y0-3+r][x0-3+c+half] << 16
    const uint8_t* org = img + (size_t)(y0 - 3) * pitch + (x0 - 3);
    const int xlim = L.w - (x0 - 3);    // columns beyond the image are never used for a valid pixel
    for (int i = tid; i < (ch + 6) * tw; i += FT_THREADS) {
        const int r = i / tw, c = i - r * tw;
        const uint8_t* p = org + (size_t)r * pitch + c;
        const uint32_t lo = p[0];
        const uint32_t hi = (c + half < xlim) ? p[half] : 0u;
        s_tile[i] = lo | (hi << 16);
    }
    for (int i = tid; i < (ch + 2) * sp; i += FT_THREADS) s_score[i] = 0;
    __syncthreads();

    // b for every pixel of the detect area, two at a time
    for (int i = tid; i < ch * half; i += FT_THREADS) {
        const int r = i / half, c = i - r * half;
        const uint32_t* t = &s_tile[(r + 3) * tw + (c + 3)];
        uint32_t ring[16];
        ring[0] = t[3 * tw];       ring[1] = t[3 * tw + 1];   ring[2] = t[2 * tw + 2];   ring[3] = t[tw + 3];
        ring[4] = t[3];            ring[5] = t[-tw + 3];      ring[6] = t[-2 * tw + 2];  ring[7] = t[-3 * tw + 1];
        ring[8] = t[-3 * tw];      ring[9] = t[-3 * tw - 1];  ring[10] = t[-2 * tw - 2]; ring[11] = t[-tw - 3];
        ring[12] = t[-3];          ring[13] = t[tw - 3];      ring[14] = t[2 * tw - 2];  ring[15] = t[3 * tw - 1];
        const uint32_t bb = fast_b_pair(t[0], ring);
        s_score[(r + 1) * sp + (c + 1)] = (uint8_t)(bb & 0xFFFFu);
        if (c + half < cw) s_score[(r + 1) * sp + (c + half + 1)] = (uint8_t)(bb >> 16);
    }
    __syncthreads();

    // cell-local strict 3x3 non-max suppression; survivors above each threshold as per-thread bit masks
    uint32_t m_ini = 0, m_min = 0;
    const int npx = ch * cw;
#pragma unroll 1
    for (int k = 0; k < FT_ITEMS_PER_THREAD; ++k) {
        const int i = tid + k * FT_THREADS;
        if (i >= npx) break;
        const int r = i / cw, c = i - r * cw;
        const uint8_t* s = &s_score[(r + 1) * sp + (c + 1)];
        const int v = s[0];
        if (v > g.min_th) {
            const bool mx = v > s[-1] && v > s[1] && v > s[-sp - 1] && v > s[-sp] && v > s[-sp + 1] &&
                            v > s[sp - 1] && v > s[sp] && v > s[sp + 1];
            if (mx) {
                m_min |= 1u << k;
                if (v > g.ini_th) m_ini |= 1u << k;
            }
        }
    }
    const int any_ini = __syncthreads_or(m_ini != 0);
    uint32_t m = any_ini ? m_ini : m_min;   // src/ORBextractor.cc:809-816: min threshold only if ini found nothing

    // append survivors: block-aggregated reservation, then one write per survivor
    const int mine = __popc(m);
    // warp scan + cross-warp offsets
    __shared__ int s_warp[FT_THREADS / 32];
    const int lane = tid & 31, wid = tid >> 5;
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int n = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= o) incl += n;
    }
    if (lane == 31) s_warp[wid] = incl;
    __syncthreads();
    if (tid == 0) {
        int tot = 0;
        for (int wdx = 0; wdx < FT_THREADS / 32; ++wdx) { int v = s_warp[wdx]; s_warp[wdx] = tot; tot += v; }
        s_base = tot ? atomicAdd(&cand_cnt[b * g.nlevels + level], tot) : 0;
    }
    __syncthreads();
    int pos = s_base + s_warp[wid] + incl - mine;
    uint32_t* out = cand + (size_t)b * g.cand_total + L.cand_off;
    while (m) {
        const int k = __ffs(m) - 1;
        m &= m - 1;
        const int i = tid + k * FT_THREADS;
        const int r = i / cw, c = i - r * cw;
        const uint32_t v = s_score[(r + 1) * sp + (c + 1)];
        const uint32_t xr = x0 + c - B2_BORDER, yr = y0 + r - B2_BORDER;   // relative to minBorderX/Y
        if (pos < L.cand_cap) out[pos] = xr | (yr << 12) | ((v - 1u) << 24);
        ++pos;
    }
}

}  // namespace

void b2_launch_fast(const B2Geom& g, const B2Images& im, uint32_t* cand, int* cand_cnt, int nimg, cudaStream_t s) {
    dim3 grid(g.total_cells, nimg);
    fast_cells_kernel<<<grid, FT_THREADS, 0, s>>>(g, im, cand, cand_cnt);
}
