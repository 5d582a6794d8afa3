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
// Organised around instructions per pixel (the stage is ALU-pipe bound, not HBM bound):
//
// fast_score_kernel — one WARP owns a 120-px strip (lanes 1..30 produce 4 px each; lanes 0 and 31 only feed
//   the halo of the score and of the fused NMS) and marches down a band of rows.  Per image row a lane loads one 32-bit word, splits it into two
//   u16x2 "pixel pairs" (x,x+2) and (x+1,x+3) and gets the six neighbouring pairs by shuffle, so the 16 ring
//   pixels of TWO output pixels are always one register each; the last 7 rows of pairs live in registers
//   (7-phase unrolled rotation).  b is evaluated on packed pairs with VIMNMX/VIMNMX3.U16x2:
//       Rmax = max_arcs min_arc r,  Rmin = min_arcs max_arc r,  b = max(Rmax, c) - min(Rmin, c)
//   (at most one side can be positive), arcs taken two at a time: max(min9[k], min9[k+1]) =
//   min(min8[k+1..k+8], max(r[k], r[k+9])) — 36 min/max per side per pixel pair.  The 3x3 non-max suppression
//   with the cell-boundary rule is fused in (3-row window of packed scores, column masks per lane, row masks
//   per step) and the kernel writes a sparse corner map M: M = b where the pixel survives NMS inside its
//   cell and b > minThFAST, else 0.  M itself is never stored: its non-zero words (few: NMS keeps <= 3 px of 4)
//   are compacted by ballot into a per-warp shared-memory queue and copied to the level's RAW list (uint2 = the
//   4 scores + position) every FS_FLUSH rows with one global atomic; "b > iniThFAST seen" is OR-ed into two
//   registers per lane and written to the per-cell flag array whenever the band crosses a cell-row boundary.
// fast_vote_kernel — the threshold vote of src/ORBextractor.cc:809-816 (a cell falls back to minThFAST only when
//   iniThFAST found nothing in it) needs every strip of a cell, i.e. a kernel boundary: one thread per raw word
//   keeps a corner iff b > ini or its cell's flag is clear and appends x | y<<12 | (b-1)<<24 to the candidate list
//   (warp-aggregated atomics).  Candidate order is irrelevant downstream: the quadtree breaks response ties with
//   a key that reproduces the reference's cell-row-major order.
#include <cuda_fp16.h>
#include "b2orb_internal.cuh"

namespace {

constexpr int FS_WARPS = 4;
constexpr int FS_PF = 5;        // rows in flight per warp (ring of 7 slots, slot = row % 7 = unroll phase)
constexpr int FS_FLUSH = 7;     // output rows between queue flushes (one pass of the 7-phase unrolled loop)
constexpr int FS_QUEUE = 2 * 30 * FS_FLUSH;   // at most one non-zero word per storing lane and row; flushed when half full

__device__ __forceinline__ unsigned vmin2(unsigned a, unsigned b) { return __vminu2(a, b); }
__device__ __forceinline__ unsigned vmax2(unsigned a, unsigned b) { return __vmaxu2(a, b); }
__device__ __forceinline__ unsigned vmin3(unsigned a, unsigned b, unsigned c) { return __vimin3_u16x2(a, b, c); }
__device__ __forceinline__ unsigned vmax3(unsigned a, unsigned b, unsigned c) { return __vimax3_u16x2(a, b, c); }

// Pixels travel as 0x6400 | v per u16 lane: as an unsigned short that orders like v (VIMNMX.U16x2 on the ALU pipe),
// and as a binary16 it IS the number 1024 + v, exactly — so the same registers can also be compared on the FMA pipe,
// which this ALU-bound kernel leaves idle:  d = relu(a - b) (HFMA2.RELU),  max = b + d,  min = a - d (HADD2).
// Every intermediate is an integer below 2048, hence exact in binary16, and the results carry the same 0x6400 tag.
__device__ __forceinline__ void hminmax2(unsigned a, unsigned b, unsigned& mn, unsigned& mx) {
    const __half2 ha = *reinterpret_cast<const __half2*>(&a), hb = *reinterpret_cast<const __half2*>(&b);
    const __half2 d = __hfma2_relu(hb, __floats2half2_rn(-1.f, -1.f), ha);
    const __half2 hx = __hadd2(hb, d), hn = __hsub2(ha, d);
    mx = *reinterpret_cast<const unsigned*>(&hx);
    mn = *reinterpret_cast<const unsigned*>(&hn);
}

// b for the two pixels packed in c (u16 lanes), ring pixels v[0..15] in circle order (same packing).
// NQ of the 8 first-layer (min, max) pairs and NE of the 8 arc-end (min, max) pairs go to the FMA pipe.
template <int NQ, int NE>
__device__ __forceinline__ unsigned fast_b_pair(const unsigned (&v)[16], unsigned c) {
    unsigned qn[8], qx[8];     // min / max of v[j], v[j+1] for odd j = 2i+1
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        if (i < NQ) {
            hminmax2(v[2 * i + 1], v[(2 * i + 2) & 15], qn[i], qx[i]);
        } else {
            qn[i] = vmin2(v[2 * i + 1], v[(2 * i + 2) & 15]);
            qx[i] = vmax2(v[2 * i + 1], v[(2 * i + 2) & 15]);
        }
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
        unsigned en, ex;
        if (i < NE) {
            hminmax2(a, e, en, ex);
        } else {
            en = vmin2(a, e);
            ex = vmax2(a, e);
        }
        pn[i] = vmin3(rn[i], rn[(i + 2) & 7], ex);   // max(min9[k], min9[k+1])
        px[i] = vmax3(rx[i], rx[(i + 2) & 7], en);   // min(max9[k], max9[k+1])
    }
    const unsigned rmax = vmax3(vmax3(pn[0], pn[1], pn[2]), vmax3(pn[3], pn[4], pn[5]), vmax2(pn[6], pn[7]));
    const unsigned rmin = vmin3(vmin3(px[0], px[1], px[2]), vmin3(px[3], px[4], px[5]), vmin2(px[6], px[7]));
    return vmax2(rmax, c) - vmin2(rmin, c);   // per-lane non-negative (the 0x6400 tags cancel): no borrow between the halves
}

// one 4-byte async copy + its own commit group
__device__ __forceinline__ void fs_cp_async4(unsigned smem_addr, const void* gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n\tcp.async.commit_group;" ::"r"(smem_addr), "l"(gmem) : "memory");
}
// wait until at most FS_PF groups are pending (the oldest needed row has landed), then read this lane's word
__device__ __forceinline__ unsigned fs_wait_lds(unsigned smem_addr) {
    unsigned v;
    asm volatile("cp.async.wait_group %1;\n\tld.shared.u32 %0, [%2];" : "=r"(v) : "n"(FS_PF), "r"(smem_addr) : "memory");
    return v;
}

// out = (b > nb) ? b : 0 per u16 lane, all values < 256
__device__ __forceinline__ unsigned keep_if_greater(unsigned b, unsigned nb) {
    const unsigned d = b + 0x00FF00FFu - nb;                 // bit 8 of each lane: b > nb
    const unsigned one = __byte_perm(d, 0u, 0x4341);         // (d.b1, 0, d.b3, 0): 0 or 1 per lane
    return b & (one * 255u);
}

// out = (b > nb) ? (b & m) : 0 per u16 lane, all values < 256: bit 15 of (b - nb + 0x7FFF) is "b > nb"; PRMT replicates it
// over the low byte of each lane (selector nibbles 9 / B = sign of byte 1 / 3, 4 = zero byte), one LOP3 applies both masks
__device__ __forceinline__ unsigned keep_if_greater_and(unsigned b, unsigned nb, unsigned m) {
    const unsigned d = b + 0x7FFF7FFFu - nb;
    unsigned keep;                                           // (__byte_perm masks the selector's replicate bit: PTX it is)
    asm("prmt.b32 %0, %1, %2, 0x4B49;" : "=r"(keep) : "r"(d), "r"(0u));
    return b & keep & m;
}

// ---------------------------------------------------------------------------------------------------
// fast_score_win_kernel — the 7-row window of pixel pairs lives in SHARED memory, not registers.  Why: a register
// window forces a 7-phase unrolled loop (static register indices), 36-40 KB of SASS for the loop alone — more than
// the 32 KB instruction cache level, and ncu showed "no instruction" as the top stall of that first version (removed;
// git history).  Here the loop body exists once (~5 KB): each lane keeps its 8 pairs of the six older rows in a private
// 6-slot ring (two 16-byte units per row: Q = pairs 2,5,3,4, W = pairs 1,6,0,7; lane stride 16 B, conflict-free
// LDS.128 / STS.128), re-reads the 9 units the ring pixels of this step need, and overwrites the oldest slot with the
// new row.  Slot addresses rotate through six registers (moves on the FMA pipe).  Only the lane itself reads what it
// stored, so no warp synchronisation is involved.  Half the registers -> more resident warps.
constexpr int FW_SLOTS = 6;
constexpr int FW_RING = 8;      // cp.async staging slots per lane (power of two: slot = row & 7)
#ifndef FW_MINB
#define FW_MINB 4
#endif

__device__ __forceinline__ uint4 fw_lds128(unsigned a) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ uint2 fw_lds64(unsigned a) {
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void fw_sts128(unsigned a, unsigned x, unsigned y, unsigned z, unsigned w) {
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(x), "r"(y), "r"(z), "r"(w) : "memory");
}

template <int NQ, int NE>
__global__ void __launch_bounds__(FS_WARPS * 32, FW_MINB)
fast_score_win_kernel(const __grid_constant__ B2Geom g, const B2Images im, uint2* __restrict__ raw_list,
                      int* __restrict__ raw_cnt, uint8_t* __restrict__ cell_ini, int variant) {
    b2_pdl_wait();
    __shared__ uint2 s_queue[FS_WARPS][FS_QUEUE];
    __shared__ unsigned s_ring[FS_WARPS][FW_RING][32];
    __shared__ uint4 s_win[FS_WARPS][FW_SLOTS][64];
    const int lane = threadIdx.x & 31;
    // broadcast from lane 0: tells the compiler the warp index (and everything derived: task, level, band, row counters,
    // "special") is warp-uniform -> uniform registers and branches, no convergence bookkeeping around the shuffles
    const int wid = __shfl_sync(0xFFFFFFFFu, (int)(threadIdx.x >> 5), 0);
    const int task = blockIdx.x * FS_WARPS + wid;
    if (task >= g.total_fast_tasks[variant]) return;
    const int band_rows = B2_FAST_BAND << variant;
    const int b = blockIdx.y;
    int level = 0;
    while (level + 1 < g.nlevels && task >= g.lv[level + 1].fast_task_base[variant]) ++level;
    const B2Level& L = g.lv[level];
    const int t = task - L.fast_task_base[variant];
    const int band = t / L.fast_strips, strip = t - band * L.fast_strips;
    int pitch;
    const uint8_t* img = b2_level_ptr(g, im, b, level, pitch);
    uint2* queue = s_queue[wid];
    int qn = 0;                                                // queue fill, warp-uniform
    uint2* out = raw_list + (size_t)b * g.cand_total + L.cand_off;
    int* out_cnt = raw_cnt + b * g.nlevels + level;
    uint8_t* flags = cell_ini + (size_t)b * g.total_cells + L.cell_base;
    const int w = L.w, h = L.h;
    const int x = B2_BORDER - 4 + strip * 120 + lane * 4;      // lanes 1..30 own columns x .. x+3
    const int ya = B2_EDGE + band * band_rows;                 // output rows [ya, yb)
    const int yb = min(ya + band_rows, h - B2_EDGE);
    const int nsteps = yb - ya + 8;                            // pixel rows ya-4 .. yb+3
    const int xo = min(x, pitch - 4);

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
    const int cjA = max(x - B2_EDGE, 0) / L.w_cell;
    unsigned sameA = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j)
        if (max(x + j - B2_EDGE, 0) / L.w_cell == cjA) sameA |= 1u << j;
    const unsigned kini = 0x00FF00FFu - (unsigned)g.ini_th * 0x00010001u;
    const unsigned lt_mask = (1u << lane) - 1;
    unsigned accE = 0, accO = 0;
    const unsigned thr2 = (unsigned)g.min_th * 0x00010001u;
    const bool store_lane = lane >= 1 && lane <= 30 && x < w - B2_BORDER;

    const uint8_t* col0 = img + xo + (size_t)(ya - 4) * pitch;
    const unsigned offmax = (unsigned)(nsteps - 1) * (unsigned)pitch;
    const unsigned ring_s = (unsigned)__cvta_generic_to_shared(&s_ring[wid][0][lane]);
    unsigned off = 0;
#pragma unroll
    for (int i = 0; i < FS_PF; ++i) {
        fs_cp_async4(ring_s + i * 128, col0 + min(off, offmax));
        off += pitch;
    }
    // slot addresses by age: a1 = previous row (dy = +2 of the output row) ... a6 = oldest (dy = -3)
    unsigned a1 = (unsigned)__cvta_generic_to_shared(&s_win[wid][0][lane]);
    unsigned a2 = a1 + 1024, a3 = a1 + 2048, a4 = a1 + 3072, a5 = a1 + 4096, a6 = a1 + 5120;

    unsigned rmE_a = 0, rmO_a = 0, rmE_b = 0, rmO_b = 0, bE_b = 0, bO_b = 0, sdE_b = 0, sdO_b = 0;
    int ci = (band * band_rows) / L.h_cell;                     // cell row of the output row
    unsigned meta = (unsigned)(x - B2_BORDER) | ((unsigned)(ya - B2_BORDER) << 12) | ((unsigned)ci << 24);
    const int hcell = L.h_cell, ylast = h - B2_EDGE - 1;
    // loop invariants the compiler would otherwise rebuild from the constant bank / special registers every row
    unsigned thr2_r = thr2, queue_s = (unsigned)__cvta_generic_to_shared(queue);
    asm volatile("" : "+r"(thr2_r), "+r"(queue_s));

    // "b > ini seen" per pixel: acc holds the running max of the emitted scores (one VIMNMX per row instead of add + or)
#define B2_FW_FLAGS()                                                                                          \
    {                                                                                                           \
        const unsigned tE = accE + kini, tO = accO + kini;      /* bit 8 of each u16 lane: score > ini */       \
        const unsigned hit = ((tE >> 8) & 1u) | ((tO >> 7) & 2u) | ((tE >> 22) & 4u) | ((tO >> 21) & 8u);       \
        if (store_lane && (hit & sameA)) flags[ci * L.n_cols + cjA] = 1;                                       \
        if (store_lane && (hit & ~sameA)) flags[ci * L.n_cols + cjA + 1] = 1;                                  \
        accE = accO = 0;                                                                                        \
    }
#define B2_FW_FLUSH()                                                                                          \
    {                                                                                                           \
        __syncwarp();                                                                                           \
        if (qn) {                                                                                               \
            int base = 0;                                                                                       \
            if (lane == 0) base = atomicAdd(out_cnt, qn);                                                       \
            base = __shfl_sync(0xFFFFFFFFu, base, 0);                                                           \
            for (int i = lane; i < qn; i += 32)                                                                 \
                if (base + i < L.cand_cap) out[base + i] = queue[i];                                            \
            __syncwarp();                                                                                       \
            qn = 0;                                                                                             \
        }                                                                                                       \
    }
    // ---- one row step.  MODE 0: rows 0..5 only fill the window.  MODE 1: rows 6, 7 also compute scores (they prime the
    // 3-row NMS state).  MODE 2: steady state, emits output row y = ya + r - 8.  The steady state has a COMMON form (the
    // row is neither the first nor the last row of its FAST cell: no masks, no cell bookkeeping) and a SPECIAL form taken
    // when the countdown `ev` hits zero: cell-boundary rows (neighbour rows of another cell do not suppress), the cell's
    // iniThFAST flags, the queue flush, and the distance to the next special row.
    int ev = 0;                                                 // common rows left before the next special row
#ifdef FWX_PLAIN_STORE
#define FWX_STORE(a, m, t, has) if (has) asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(a), "r"(m), "r"(t) : "memory");
#else
#define FWX_STORE(a, m, t, has)                                                                                \
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %3, 0;\n\t@p st.shared.v2.u32 [%0], {%1, %2};\n\t}" \
                 ::"r"(a), "r"(m), "r"(t), "r"(has) : "memory");
#endif
#define B2_FW_ROW(MODE)                                                                                        \
    {                                                                                                           \
        fs_cp_async4(ring_s + (((r + FS_PF) & (FW_RING - 1)) << 7), col0 + min(off, offmax));                  \
        off += pitch;                                                                                           \
        const unsigned R1 = fs_wait_lds(ring_s + ((r & (FW_RING - 1)) << 7));                                  \
        const unsigned E1 = __byte_perm(R1, 0x64646464u, 0x4240), O1 = __byte_perm(R1, 0x64646464u, 0x4341);  \
        const unsigned E2 = __shfl_down_sync(0xFFFFFFFFu, E1, 1), O2 = __shfl_down_sync(0xFFFFFFFFu, O1, 1);  \
        const unsigned c2 = __byte_perm(E1, E2, 0x5432), c3 = __byte_perm(O1, O2, 0x5432);                    \
        const unsigned Ep = __shfl_up_sync(0xFFFFFFFFu, E1, 1), n0 = __shfl_up_sync(0xFFFFFFFFu, O1, 1);      \
        const unsigned n1 = __byte_perm(Ep, E1, 0x5432);       /* (x-2, x): the high half is this lane's own pixel, */ \
        const unsigned n2 = __byte_perm(n0, O1, 0x5432);       /* (x-1, x+1)   so lane 0's last pixel gets a valid score */ \
        if (MODE >= 1) {                                                                                        \
            /* unit Q = pairs (2, 5, 3, 4), unit W = pairs (1, 6, 0, 7): every load asks for exactly what it uses */ \
            const uint2 q1 = fw_lds64(a1), w1 = fw_lds64(a1 + 512);       /* dy +2: pairs 2,5 | 1,6 */         \
            const uint4 w2 = fw_lds128(a2 + 512);                          /* dy +1: pairs 1,6,0,7 */           \
            const uint2 q3 = fw_lds64(a3 + 8);                             /* dy  0: pairs 3,4 */               \
            const uint4 w3 = fw_lds128(a3 + 512);                          /*        pairs 1,6,0,7 */           \
            const uint4 w4 = fw_lds128(a4 + 512);                          /* dy -1 */                          \
            const uint2 q5 = fw_lds64(a5), w5 = fw_lds64(a5 + 512);       /* dy -2 */                          \
            const uint4 q6 = fw_lds128(a6);                                /* dy -3: pairs 2,5,3,4 */           \
            unsigned bE, bO;                                                                                    \
            {                                                                                                   \
                const unsigned v[16] = {E1, O1, q1.y, w2.y, w3.y, w4.y, q5.y, q6.w,                            \
                                        q6.z, q6.x, w5.x, w4.z, w3.z, w2.z, w1.x, n2};                         \
                bE = fast_b_pair<NQ, NE>(v, q3.x);                                                              \
            }                                                                                                   \
            {                                                                                                   \
                const unsigned v[16] = {O1, c2, w1.y, w2.w, w3.w, w4.w, w5.y, q6.y,                            \
                                        q6.w, q6.z, q5.x, w4.x, w3.x, w2.x, q1.x, E1};                         \
                bO = fast_b_pair<NQ, NE>(v, q3.y);                                                              \
            }                                                                                                   \
            const unsigned bOp = __shfl_up_sync(0xFFFFFFFFu, bO, 1), bEn = __shfl_down_sync(0xFFFFFFFFu, bE, 1); \
            const unsigned lE = __byte_perm(bOp, bO, 0x5432) & mlE;                                             \
            const unsigned rE = bO & mrE;                                                                       \
            const unsigned lO = bE & mlO;                                                                       \
            const unsigned rO = __byte_perm(bE, bEn, 0x5432) & mrO;                                             \
            const unsigned rmE = vmax3(lE, bE, rE), rmO = vmax3(lO, bO, rO);                                    \
            if (MODE == 2) {                                                                                    \
                unsigned nbE, nbO;                                                                              \
                const bool special = ev == 0;                                                                   \
                int cy = 0, yy = 0;                                                                             \
                if (!special) {                                                                                 \
                    --ev;                                                                                       \
                    nbE = vmax3(rmE_a, rmE, vmax2(sdE_b, thr2_r));                                              \
                    nbO = vmax3(rmO_a, rmO, vmax2(sdO_b, thr2_r));                                              \
                } else {                                                                                        \
                    yy = ya + r - 8;                            /* (not from meta: lane 0's x field is negative) */ \
                    cy = yy - B2_EDGE - ci * hcell;                                                             \
                    const unsigned up = cy != 0 ? 0xFFFFFFFFu : 0u;                                             \
                    const unsigned dn = (cy != hcell - 1 && yy != ylast) ? 0xFFFFFFFFu : 0u;                    \
                    nbE = vmax3(rmE_a & up, rmE & dn, vmax2(sdE_b, thr2_r));                                    \
                    nbO = vmax3(rmO_a & up, rmO & dn, vmax2(sdO_b, thr2_r));                                    \
                }                                                                                               \
                const unsigned oE = keep_if_greater_and(bE_b, nbE, inE), oO = keep_if_greater_and(bO_b, nbO, inO); \
                const unsigned m4 = __byte_perm(oE, oO, 0x6240);                                                \
                accE = vmax2(accE, oE); accO = vmax2(accO, oO);                                                 \
                const int has = store_lane && m4 != 0;                                                          \
                const unsigned bal = __ballot_sync(0xFFFFFFFFu, has);                                           \
                FWX_STORE(queue_s + 8u * (unsigned)(qn + __popc(bal & lt_mask)), m4, meta, has)                 \
                qn += __popc(bal);                                                                              \
                meta += 0x1000u;                                                                                \
                if (special) {                                                                                  \
                    int cyn = cy + 1;                           /* row of the NEXT output row inside its cell */ \
                    if (cy == hcell - 1) { B2_FW_FLAGS() ++ci; meta += 0x1000000u; cyn = 0; }                   \
                    if (qn > FS_QUEUE - 30 * FS_FLUSH) B2_FW_FLUSH()                                            \
                    /* common rows before the next special one: the next row is special when it is the first row of a \
                       cell; otherwise run until the last row of the cell / of the level, at most FS_FLUSH - 1 rows */ \
                    ev = cyn == 0 ? 0 : min(min(hcell - 1 - cyn, ylast - (yy + 1)), FS_FLUSH - 1);              \
                    ev = max(ev, 0);                                                                            \
                }                                                                                               \
            }                                                                                                   \
            rmE_a = rmE_b; rmO_a = rmO_b;                                                                       \
            rmE_b = rmE; rmO_b = rmO; bE_b = bE; bO_b = bO;                                                     \
            sdE_b = vmax2(lE, rE); sdO_b = vmax2(lO, rO);                                                       \
        }                                                                                                       \
        fw_sts128(a6, n2, c2, E1, O1);          /* Q = pairs 2, 5, 3, 4 */                                      \
        fw_sts128(a6 + 512, n1, c3, n0, E2);    /* W = pairs 1, 6, 0, 7 */                                      \
        const unsigned tmp = a6;                                                                                \
        a6 = a5; a5 = a4; a4 = a3; a3 = a2; a2 = a1; a1 = tmp;                                                  \
    }

    int r = 0;
#pragma unroll 1
    for (; r < 6; ++r) B2_FW_ROW(0)
#pragma unroll 1
    for (; r < 8; ++r) B2_FW_ROW(1)
    // one row per trip.  (Round 1 unrolled by two to give the scheduler two independent trees per warp; since the warp
    // index is compiler-visibly uniform the body is 25 % shorter and the smaller loop wins: 8.35 -> 8.22 us/pair; a fifth
    // CTA per SM (93 registers) measured 8.39.)
#pragma unroll 1
    for (; r < nsteps; ++r) B2_FW_ROW(2)
    B2_FW_FLUSH()
    B2_FW_FLAGS()
#undef B2_FW_ROW
#undef B2_FW_FLAGS
#undef B2_FW_FLUSH
}

// ---------------------------------------------------------------------------------------------------
constexpr int FV_THREADS = 256;
constexpr int FV_MAX_CHUNKS = 32;  // CTAs per (image, level): 4 for large batches, more when few images are in flight

constexpr int FV_Q = 256;          // per-warp staging queue (entries); one iteration adds at most 96

__global__ void __launch_bounds__(FV_THREADS)
fast_vote_kernel(const __grid_constant__ B2Geom g, const uint2* __restrict__ raw_list, const int* __restrict__ raw_cnt,
                 const uint8_t* __restrict__ cell_ini, uint32_t* __restrict__ cand, int* __restrict__ cand_cnt,
                 int chunks) {
    b2_pdl_wait();
    __shared__ uint32_t s_q[FV_THREADS / 32][FV_Q];
    const int level = blockIdx.x / chunks, chunk = blockIdx.x - level * chunks, b = blockIdx.y;
    const B2Level& L = g.lv[level];
    const int n = min(raw_cnt[b * g.nlevels + level], L.cand_cap);
    const int per = ((n + chunks * FV_THREADS - 1) / (chunks * FV_THREADS)) * FV_THREADS;
    const int lo = chunk * per, hi = min(lo + per, n);
    if (lo >= hi) return;
    const uint2* raw = raw_list + (size_t)b * g.cand_total + L.cand_off;
    const uint8_t* flags = cell_ini + (size_t)b * g.total_cells + L.cell_base;
    uint32_t* out = cand + (size_t)b * g.cand_total + L.cand_off;
    int* out_cnt = cand_cnt + b * g.nlevels + level;
    const int lane = threadIdx.x & 31;
    uint32_t* q = s_q[threadIdx.x >> 5];
    const unsigned lt_mask = (1u << lane) - 1;
    const unsigned ini = (unsigned)g.ini_th;
    const int wc = L.w_cell, cap = L.cand_cap;
    // cell column by multiplication: exact for n < 2^20 / w_cell (n < 4096, w_cell < 60)
    const unsigned magic = ((1u << 20) + (unsigned)wc - 1u) / (unsigned)wc;
    int qn = 0;                                                 // queue fill, warp-uniform
    const auto flush = [&]() {
        __syncwarp();
        if (qn) {
            int base = 0;
            if (lane == 0) base = atomicAdd(out_cnt, qn);
            base = __shfl_sync(0xFFFFFFFFu, base, 0);
            for (int i = lane; i < qn; i += 32)
                if (base + i < cap) out[base + i] = q[i];
            __syncwarp();
            qn = 0;
        }
    };
    uint2 nxt = make_uint2(0u, 0u);
    if (lo + (int)threadIdx.x < hi) nxt = raw[lo + threadIdx.x];
    for (int i0 = lo; i0 < lo + per; i0 += FV_THREADS) {       // whole warps iterate together (ballots below)
        const uint2 e = nxt;
        const int in = i0 + FV_THREADS + (int)threadIdx.x;      // next word in flight while this one is voted on
        nxt = make_uint2(0u, 0u);
        if (in < hi) nxt = raw[in];
        // NMS leaves at most three corners in a word of 4 horizontally adjacent pixels (neighbours in different
        // cells do not suppress each other, so two may be adjacent across a cell boundary).  The word spans at most two
        // cells (w_cell >= 30): pixels [0, jb) lie in cell column colA, the rest in colA + 1.
        uint32_t keep[3];
        int nk = 0;
        const unsigned m4 = e.x;
        if (m4) {
            const unsigned xb = e.y & 0xFFFu;                   // relative to the 16-px border
            const unsigned rowbase = (e.y >> 24) * (unsigned)L.n_cols;
            // (the first word of a row starts left of the first cell: its leading pixels are always 0)
            const unsigned colA = ((max(xb, (unsigned)(B2_EDGE - B2_BORDER)) - (B2_EDGE - B2_BORDER)) * magic) >> 20;
            const int jb = (int)((colA + 1u) * (unsigned)wc + (B2_EDGE - B2_BORDER)) - (int)xb;
            // a flag is only needed for corners that pass minThFAST but not iniThFAST
            const unsigned le = __vcmpleu4(m4, ini * 0x01010101u) & __vcmpne4(m4, 0u);
            unsigned fA = 1, fB = 1;                            // "iniThFAST found something in the cell"
            if (le) {
                fA = flags[rowbase + colA];
                if (jb < 4) fB = flags[rowbase + colA + 1];
            }
            const unsigned ypart = e.y & 0x00FFF000u;
            unsigned m = m4;                                    // usually a single corner per word: visit set bytes only
            while (m) {
                const int j = (__ffs(m) - 1) >> 3;
                const unsigned bv = (m4 >> (8 * j)) & 0xFFu;
                m &= ~(0xFFu << (8 * j));
                const unsigned f = j < jb ? fA : fB;
                if ((bv > ini || f == 0) && nk < 3) keep[nk++] = (xb + j) | ypart | ((bv - 1u) << 24);
            }
        }
        const unsigned b0 = __ballot_sync(0xFFFFFFFFu, nk & 1), b1 = __ballot_sync(0xFFFFFFFFu, nk & 2);
        const int total = __popc(b0) + 2 * __popc(b1);
        int pos = qn + __popc(b0 & lt_mask) + 2 * __popc(b1 & lt_mask);
        if (nk > 0) q[pos] = keep[0];
        if (nk > 1) q[pos + 1] = keep[1];
        if (nk > 2) q[pos + 2] = keep[2];
        qn += total;
        if (qn > FV_Q - 96) flush();
    }
    flush();
}

}  // namespace

void b2_launch_fast_score(const B2Geom& g, const B2Images& im, uint2* raw, int* raw_cnt, uint8_t* cell_ini, int nimg,
                          int forced_band, cudaStream_t s) {
    if (g.total_fast_tasks[0] <= 0) return;
    // long bands (less halo work per output row) once the batch alone fills the machine several times over
    // and short ones while a 32-row launch would not even give every SM two CTAs (per-CTA latency is what counts then)
    int variant = (long long)g.total_fast_tasks[3] * nimg >= 148LL * 16 * 4 ? 3
                  : (long long)g.total_fast_tasks[2] * nimg >= 148LL * 16 * 4 ? 2
                  : (long long)g.total_fast_tasks[1] * nimg >= 148LL * 2 * 4 ? 1 : 0;
    if (forced_band >= 0 && forced_band < B2_FAST_VARIANTS) variant = forced_band;   // b2orb_debug_set("fast_band")
    dim3 grid((g.total_fast_tasks[variant] + FS_WARPS - 1) / FS_WARPS, nimg);
    // the 8 first-layer (min, max) pairs on the FMA pipe, everything else on the ALU pipe: every other mix that was
    // measured (0 / 8 / 16 pairs) landed within 1 %, the kernel runs at the mixed-issue limit
    b2_launch_pdl(fast_score_win_kernel<8, 0>, dim3(grid), dim3(FS_WARPS * 32), 0, s, g, im, raw, raw_cnt, cell_ini, variant);
}

void b2_launch_fast_vote(const B2Geom& g, const uint2* raw, const int* raw_cnt, const uint8_t* cell_ini, uint32_t* cand,
                         int* cand_cnt, int nimg, cudaStream_t s) {
    int chunks = 4;
    while (chunks < FV_MAX_CHUNKS && g.nlevels * chunks * nimg < 148 * 8) chunks *= 2;
    dim3 grid(g.nlevels * chunks, nimg);
    b2_launch_pdl(fast_vote_kernel, dim3(grid), dim3(FV_THREADS), 0, s, g, raw, raw_cnt, cell_ini, cand, cand_cnt, chunks);
}
