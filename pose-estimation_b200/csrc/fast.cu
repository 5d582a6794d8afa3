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
//
// Two kernels, both organised around instructions per pixel (the stage is ALU-pipe bound, not HBM bound):
//
// fast_score_kernel — one WARP owns a 112-px strip (lanes 2..29 produce 4 px each, lanes 0/1/30/31 only feed
//   the halo of the score and of the fused NMS) and marches down a band of rows.  Per image row a lane loads one 32-bit word, splits it into two
//   u16x2 "pixel pairs" (x,x+2) and (x+1,x+3) and gets the six neighbouring pairs by shuffle, so the 16 ring
//   pixels of TWO output pixels are always one register each; the last 7 rows of pairs live in registers
//   (7-phase unrolled rotation).  b is evaluated on packed pairs with VIMNMX/VIMNMX3.U16x2:
//       Rmax = max_arcs min_arc r,  Rmin = min_arcs max_arc r,  b = max(Rmax, c) - min(Rmin, c)
//   (at most one side can be positive), arcs taken two at a time: max(min9[k], min9[k+1]) =
//   min(min8[k+1..k+8], max(r[k], r[k+9])) — 36 min/max per side per pixel pair.  The 3x3 non-max suppression
//   with the cell-boundary rule is fused in (3-row window of packed scores, column masks per lane, row masks
//   per step) and the kernel writes a sparse corner map M: M = b where the pixel survives NMS inside its
//   cell and b > minThFAST, else 0.
// fast_cells_kernel — one warp per cell scans its part of M, votes on the threshold (any survivor with
//   b > ini ? ini : min — src/ORBextractor.cc:809-816) and appends the survivors to the level's candidate
//   list.  Candidate order is irrelevant downstream: the quadtree breaks response ties with a key that
//   reproduces the reference's cell-row-major order.
#include "b2orb_internal.cuh"

namespace {

constexpr int FS_WARPS = 4;
constexpr int FS_PF = 4;        // rows in flight per warp

__device__ __forceinline__ unsigned vmin2(unsigned a, unsigned b) { return __vminu2(a, b); }
__device__ __forceinline__ unsigned vmax2(unsigned a, unsigned b) { return __vmaxu2(a, b); }
__device__ __forceinline__ unsigned vmin3(unsigned a, unsigned b, unsigned c) { return __vimin3_u16x2(a, b, c); }
__device__ __forceinline__ unsigned vmax3(unsigned a, unsigned b, unsigned c) { return __vimax3_u16x2(a, b, c); }

// b for the two pixels packed in c (u16 lanes), ring pixels v[0..15] in circle order (same packing).
__device__ __forceinline__ unsigned fast_b_pair(const unsigned (&v)[16], unsigned c) {
    unsigned qn[8], qx[8];     // min / max of v[j], v[j+1] for odd j = 2i+1
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        qn[i] = vmin2(v[2 * i + 1], v[(2 * i + 2) & 15]);
        qx[i] = vmax2(v[2 * i + 1], v[(2 * i + 2) & 15]);
    }
    unsigned rn[8], rx[8];     // min / max of v[j .. j+3]
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        rn[i] = vmin2(qn[i], qn[(i + 1) & 7]);
        rx[i] = vmax2(qx[i], qx[(i + 1) & 7]);
    }
    unsigned pn[8], px[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        // arcs k = 2i and k+1 share v[k+1 .. k+8] = rn[i], rn[i+2]; the ends are v[k] and v[k+9]
        const unsigned a = v[2 * i], e = v[(2 * i + 9) & 15];
        pn[i] = vmin3(rn[i], rn[(i + 2) & 7], vmax2(a, e));   // max(min9[k], min9[k+1])
        px[i] = vmax3(rx[i], rx[(i + 2) & 7], vmin2(a, e));   // min(max9[k], max9[k+1])
    }
    const unsigned rmax = vmax3(vmax3(pn[0], pn[1], pn[2]), vmax3(pn[3], pn[4], pn[5]), vmax2(pn[6], pn[7]));
    const unsigned rmin = vmin3(vmin3(px[0], px[1], px[2]), vmin3(px[3], px[4], px[5]), vmin2(px[6], px[7]));
    return vmax2(rmax, c) - vmin2(rmin, c);   // per-lane non-negative: no borrow between the halves
}

// out = (b > nb) ? b : 0 per u16 lane, all values < 256
__device__ __forceinline__ unsigned keep_if_greater(unsigned b, unsigned nb) {
    const unsigned d = b + 0x00FF00FFu - nb;                 // bit 8 of each lane: b > nb
    const unsigned one = __byte_perm(d, 0u, 0x4341);         // (d.b1, 0, d.b3, 0): 0 or 1 per lane
    return b & (one * 255u);
}

__global__ void __launch_bounds__(FS_WARPS * 32)
fast_score_kernel(const __grid_constant__ B2Geom g, const B2Images im, uint8_t* __restrict__ score) {
    const int lane = threadIdx.x & 31;
    const int task = blockIdx.x * FS_WARPS + (threadIdx.x >> 5);
    if (task >= g.total_fast_tasks) return;
    const int b = blockIdx.y;
    int level = 0;
    while (level + 1 < g.nlevels && task >= g.lv[level + 1].fast_task_base) ++level;
    const B2Level& L = g.lv[level];
    const int t = task - L.fast_task_base;
    const int band = t / L.fast_strips, strip = t - band * L.fast_strips;
    int pitch;
    const uint8_t* img = b2_level_ptr(g, im, b, level, pitch);
    uint8_t* M = score + (size_t)b * g.slab + L.off;
    const int w = L.w, h = L.h;
    const int x = B2_BORDER - 8 + strip * 112 + lane * 4;      // lanes 2..29 own columns x .. x+3
    const int ya = B2_EDGE + band * B2_FAST_BAND;              // M rows [ya, yb)
    const int yb = min(ya + B2_FAST_BAND, h - B2_EDGE);
    const int nsteps = yb - ya + 8;                            // pixel rows ya-4 .. yb+3
    const int xo = min(x, pitch - 4);
    const uint8_t* col = img + xo;

    // column masks of this lane's pixels (x, x+2 | x+1, x+3): inside [19, w-19), left / right neighbour in the same cell
    unsigned inE = 0, inO = 0, mlE = 0, mlO = 0, mrE = 0, mrO = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int X = x + j;
        const bool in = X >= B2_EDGE && X < w - B2_EDGE;
        const int cx = (X - B2_EDGE) % L.w_cell;
        const bool left = in && cx != 0;
        const bool right = in && cx != L.w_cell - 1 && X + 1 < w - B2_EDGE;
        const unsigned lanebits = (j & 2) ? 0xFFFF0000u : 0x0000FFFFu;
        if (j & 1) { if (in) inO |= lanebits; if (left) mlO |= lanebits; if (right) mrO |= lanebits; }
        else       { if (in) inE |= lanebits; if (left) mlE |= lanebits; if (right) mrE |= lanebits; }
    }
    const unsigned thr2 = (unsigned)g.min_th * 0x00010001u;
    const bool store_lane = lane >= 2 && lane <= 29 && x < w - B2_BORDER;   // b valid on lanes 1..30, NMS on 2..29

    unsigned raw[7];
    unsigned P[7][8];            // pairs (p[a], p[a+2]) for a = x-3 .. x+4, last 7 rows
    // Rows are addressed by a 32-bit offset that is clamped to the last row of the band, so the prefetch needs no
    // branch; pixel row r of the band sits at col0 + r * pitch.
    const uint8_t* col0 = col + (size_t)(ya - 4) * pitch;
    const unsigned offmax = (unsigned)(nsteps - 1) * (unsigned)pitch;
    unsigned off = 0;
#pragma unroll
    for (int i = 0; i < FS_PF; ++i) {
        raw[i] = *reinterpret_cast<const unsigned*>(col0 + min(off, offmax));
        off += pitch;
    }
#define B2_FAST_PAIRS(p)                                                                                        \
    {                                                                                                           \
        const unsigned R1 = raw[p];                                                                             \
        raw[((p) + FS_PF) % 7] = *reinterpret_cast<const unsigned*>(col0 + min(off, offmax));                   \
        off += pitch;                                                                                           \
        const unsigned E1 = __byte_perm(R1, 0u, 0x4240), O1 = __byte_perm(R1, 0u, 0x4341);                      \
        const unsigned E2 = __shfl_down_sync(0xFFFFFFFFu, E1, 1), O2 = __shfl_down_sync(0xFFFFFFFFu, O1, 1);    \
        const unsigned c2 = __byte_perm(E1, E2, 0x5432), c3 = __byte_perm(O1, O2, 0x5432);                      \
        P[p][0] = __shfl_up_sync(0xFFFFFFFFu, O1, 1);                                                           \
        P[p][1] = __shfl_up_sync(0xFFFFFFFFu, c2, 1);                                                           \
        P[p][2] = __shfl_up_sync(0xFFFFFFFFu, c3, 1);                                                           \
        P[p][3] = E1; P[p][4] = O1; P[p][5] = c2; P[p][6] = c3; P[p][7] = E2;                                   \
    }
    // prologue: pixel rows 0..5 only fill the window
#pragma unroll
    for (int p = 0; p < 6; ++p) B2_FAST_PAIRS(p)

    // NMS state: rows s-2 (rm only), s-1 (pending centre), s (newest)
    unsigned rmE_a = 0, rmO_a = 0, rmE_b = 0, rmO_b = 0, bE_b = 0, bO_b = 0, sdE_b = 0, sdO_b = 0;
    uint8_t* mrow = M + (size_t)ya * L.pitch + x;             // output row of step r = 8
    int y = ya;
    int cy = (band * B2_FAST_BAND) % L.h_cell;                 // row of y inside its cell
    const int hcell1 = L.h_cell - 1, ylast = h - B2_EDGE - 1;

#define B2_FAST_STEP(p, OUT)                                                                                   \
    {                                                                                                           \
        B2_FAST_PAIRS(p)                                                                                     \
        const int s3 = p, s2 = (p + 6) % 7, s1 = (p + 5) % 7, s0 = (p + 4) % 7, t1 = (p + 3) % 7,            \
                  t2 = (p + 2) % 7, t3 = (p + 1) % 7;                                                        \
        unsigned bE, bO;                                                                                     \
        {                                                                                                    \
            const unsigned v[16] = {P[s3][3], P[s3][4], P[s2][5], P[s1][6], P[s0][6], P[t1][6], P[t2][5], P[t3][4], \
                                    P[t3][3], P[t3][2], P[t2][1], P[t1][0], P[s0][0], P[s1][0], P[s2][1], P[s3][2]}; \
            bE = fast_b_pair(v, P[s0][3]);                                                                   \
        }                                                                                                    \
        {                                                                                                    \
            const unsigned v[16] = {P[s3][4], P[s3][5], P[s2][6], P[s1][7], P[s0][7], P[t1][7], P[t2][6], P[t3][5], \
                                    P[t3][4], P[t3][3], P[t2][2], P[t1][1], P[s0][1], P[s1][1], P[s2][2], P[s3][3]}; \
            bO = fast_b_pair(v, P[s0][4]);                                                                   \
        }                                                                                                    \
        const unsigned bOp = __shfl_up_sync(0xFFFFFFFFu, bO, 1), bEn = __shfl_down_sync(0xFFFFFFFFu, bE, 1); \
        const unsigned lE = __byte_perm(bOp, bO, 0x5432) & mlE;                                              \
        const unsigned rE = bO & mrE;                                                                        \
        const unsigned lO = bE & mlO;                                                                        \
        const unsigned rO = __byte_perm(bE, bEn, 0x5432) & mrO;                                              \
        const unsigned rmE = vmax3(lE, bE, rE), rmO = vmax3(lO, bO, rO);                                     \
        if (OUT) {                                                                                           \
            const unsigned up = cy != 0 ? 0xFFFFFFFFu : 0u;                                                  \
            const unsigned dn = (cy != hcell1 && y != ylast) ? 0xFFFFFFFFu : 0u;                             \
            const unsigned nbE = vmax3(rmE_a & up, rmE & dn, vmax2(sdE_b, thr2));                            \
            const unsigned nbO = vmax3(rmO_a & up, rmO & dn, vmax2(sdO_b, thr2));                            \
            const unsigned oE = keep_if_greater(bE_b, nbE) & inE, oO = keep_if_greater(bO_b, nbO) & inO;     \
            if (store_lane) *reinterpret_cast<unsigned*>(mrow) = __byte_perm(oE, oO, 0x6240);                \
            mrow += L.pitch;                                                                                 \
            ++y;                                                                                             \
            cy = cy == hcell1 ? 0 : cy + 1;                                                                  \
        }                                                                                                    \
        rmE_a = rmE_b; rmO_a = rmO_b;                                                                        \
        rmE_b = rmE; rmO_b = rmO; bE_b = bE; bO_b = bO;                                                      \
        sdE_b = vmax2(lE, rE); sdO_b = vmax2(lO, rO);                                                        \
    }
    // score rows ya-1 and ya only prime the NMS window (pixel rows r = 6, 7); outputs start at r = 8
    if (6 < nsteps) B2_FAST_STEP(6, false)
    if (7 < nsteps) B2_FAST_STEP(0, false)
    for (int base = 8; base < nsteps; base += 7) {
#pragma unroll
        for (int q = 0; q < 7; ++q) {
            if (base + q < nsteps) B2_FAST_STEP((1 + q) % 7, true)
        }
    }
#undef B2_FAST_STEP
#undef B2_FAST_PAIRS
}

// ---------------------------------------------------------------------------------------------------
constexpr int FC_WARPS = 4;
constexpr int FC_CHUNK = 8;
constexpr int FC_BUF = 1024;     // >= ceil(64/2) * ceil(64/2) survivors of one cell (NMS keeps <= 1 of 4)

__global__ void __launch_bounds__(FC_WARPS * 32)
fast_cells_kernel(const __grid_constant__ B2Geom g, const uint8_t* __restrict__ score, uint32_t* __restrict__ cand,
                  int* __restrict__ cand_cnt) {
    __shared__ uint32_t s_buf[FC_WARPS][FC_BUF];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int cellid = blockIdx.x * FC_WARPS + wid;
    if (cellid >= g.total_cells) return;
    const int b = blockIdx.y;
    int level = 0;
    while (level + 1 < g.nlevels && cellid >= g.lv[level + 1].cell_base) ++level;
    const B2Level& L = g.lv[level];
    const int cell = cellid - L.cell_base;
    const int ci = cell / L.n_cols, cj = cell - ci * L.n_cols;
    // detect area of this cell in level coordinates (ROI inset by 3, clipped to maxBorder-3 = dim-19)
    const int x0 = B2_EDGE + cj * L.w_cell, y0 = B2_EDGE + ci * L.h_cell;
    const int x1 = min(x0 + L.w_cell, L.w - B2_EDGE), y1 = min(y0 + L.h_cell, L.h - B2_EDGE);
    if (x1 <= x0 || y1 <= y0) return;
    const uint8_t* M = score + (size_t)b * g.slab + L.off;
    uint32_t* buf = s_buf[wid];
    const int ini = g.ini_th;
    __shared__ int s_cnt[FC_WARPS];
    if (lane == 0) s_cnt[wid] = 0;
    __syncwarp();
    // Word-wise scan: a lane owns one aligned 32-bit column of the cell (4 px) and, when the cell is at most 16
    // words wide, every other row.  M is sparse (NMS keeps <= 1 px of 4), so the append path is rarely taken;
    // buffer order is irrelevant (the quadtree re-derives the reference order from coordinates).
    const int xa = x0 & ~3;
    const int wpr = (((x1 + 3) & ~3) - xa) >> 2;               // words per cell row, <= 17
    const int rpi = wpr <= 16 ? 2 : 1;                          // rows per iteration
    const int wc = rpi == 2 ? (lane & 15) : lane;
    const int rsub = rpi == 2 ? (lane >> 4) : 0;
    unsigned bytemask = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int X = xa + 4 * wc + j;
        if (wc < wpr && X >= x0 && X < x1) bytemask |= 0xFFu << (8 * j);
    }
    bool mine_ini = false;
    const uint8_t* colp = M + xa + 4 * wc;
    // loads are issued FC_CHUNK rows at a time so their latencies overlap; the (rare) append path runs afterwards
    for (int yc = y0 + rsub; yc < y1; yc += rpi * FC_CHUNK) {
        unsigned vv[FC_CHUNK];
#pragma unroll
        for (int k = 0; k < FC_CHUNK; ++k) {
            const int y = yc + k * rpi;
            vv[k] = (bytemask && y < y1) ? (*reinterpret_cast<const unsigned*>(colp + (size_t)y * L.pitch) & bytemask) : 0u;
        }
#pragma unroll
        for (int k = 0; k < FC_CHUNK; ++k) {
            unsigned v = vv[k];
            const int y = yc + k * rpi;
            while (v) {
                const int j = (__ffs(v) - 1) >> 3;
                const unsigned bv = (v >> (8 * j)) & 0xFFu;
                v &= ~(0xFFu << (8 * j));
                mine_ini |= (int)bv > ini;
                const int pos = atomicAdd(&s_cnt[wid], 1);
                if (pos < FC_BUF)
                    buf[pos] = (unsigned)(xa + 4 * wc + j - B2_BORDER) | ((unsigned)(y - B2_BORDER) << 12) | ((bv - 1u) << 24);
            }
        }
    }
    const bool any_ini = __any_sync(0xFFFFFFFFu, mine_ini);
    __syncwarp();
    int n = s_cnt[wid];
    __syncwarp();
    n = min(n, FC_BUF);
    // src/ORBextractor.cc:809-816: the min threshold applies only if the ini threshold found nothing in this cell
    const unsigned thr = any_ini ? (unsigned)ini : 0u;     // entries hold b-1; keep b > thr  <=>  b-1 >= thr
    int kept = 0;
    for (int i0 = 0; i0 < n; i0 += 32) {
        const int i = i0 + lane;
        const bool k = i < n && (buf[i] >> 24) >= thr;
        kept += __popc(__ballot_sync(0xFFFFFFFFu, k));
    }
    if (kept == 0) return;
    int basepos = 0;
    if (lane == 0) basepos = atomicAdd(&cand_cnt[b * g.nlevels + level], kept);
    basepos = __shfl_sync(0xFFFFFFFFu, basepos, 0);
    uint32_t* out = cand + (size_t)b * g.cand_total + L.cand_off;
    for (int i0 = 0; i0 < n; i0 += 32) {
        const int i = i0 + lane;
        const uint32_t e = i < n ? buf[i] : 0u;
        const bool k = i < n && (e >> 24) >= thr;
        const unsigned bal = __ballot_sync(0xFFFFFFFFu, k);
        if (k) {
            const int pos = basepos + __popc(bal & ((1u << lane) - 1));
            if (pos < L.cand_cap) out[pos] = e;
        }
        basepos += __popc(bal);
    }
}

}  // namespace

void b2_launch_fast_score(const B2Geom& g, const B2Images& im, uint8_t* score, int nimg, cudaStream_t s) {
    if (g.total_fast_tasks <= 0) return;
    dim3 grid((g.total_fast_tasks + FS_WARPS - 1) / FS_WARPS, nimg);
    fast_score_kernel<<<grid, FS_WARPS * 32, 0, s>>>(g, im, score);
}

void b2_launch_fast_cells(const B2Geom& g, const uint8_t* score, uint32_t* cand, int* cand_cnt, int nimg, cudaStream_t s) {
    if (g.total_cells <= 0) return;
    dim3 grid((g.total_cells + FC_WARPS - 1) / FC_WARPS, nimg);
    fast_cells_kernel<<<grid, FC_WARPS * 32, 0, s>>>(g, score, cand, cand_cnt);
}
