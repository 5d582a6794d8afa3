// 7x7 Gaussian blur of every pyramid level — GaussianBlur(workingMat, workingMat, Size(7,7), 2, 2,
// BORDER_REFLECT_101), src/ORBextractor.cc:1085-1086 (K6).  OpenCV's 8-bit fixed-point path: separable
// integer taps (sum 256 in OpenCV 4.x; 257 for the kernel OpenCV 3.2 rounds, which then needs the saturating cast);
// horizontal pass exact in Q8, vertical pass accumulates in Q16 and rounds with (acc + 32768) >> 16.  Both passes are exact integer sums with ONE rounding at the end, so any
// evaluation order is bit-identical.
//
// The kernel is instruction-issue bound, not HBM bound, so it is organised around instructions per pixel:
// one WARP owns a 128-px wide strip (4 px per lane) and marches down a band of rows.  Per input row a lane
// reads its own word and its two neighbours from a shared-memory ring the warp fills with cp.async (10 lanes
// copy one 16-byte chunk each, BL_PF rows in flight), and computes the 4 horizontal sums with 10 IDP.4A (u8x4
// dot products against shifted tap vectors — no byte unpacking at all).  The vertical pass keeps the last rows as
// 16-bit PAIRS in registers (6-phase unrolled rotation): 1 PRMT + 1 IMAD + 3 IDP.2A per pixel.  Results are packed with 3 PRMT and stored as one 32-bit word.
// Image borders (reflect-101) are folded into per-lane PRMT selectors computed once per strip; interior strips
// skip that code entirely.
// One launch covers all levels of all images: blockIdx.x walks a flattened (level, band, strip) task list.
#include "b2orb_internal.cuh"
#ifdef B2ORB_VARIANTS
#include <cstring>
#include <cuda.h>   // CUtensorMap (types only; the encoder is fetched through cudaGetDriverEntryPoint)
#endif

namespace {

constexpr int BL_WARPS = 4;                 // independent warp tasks per CTA
constexpr int BL_PF = 6;                    // rows in flight per warp
constexpr int BL_SLOTS = 8;                 // shared-memory ring (>= BL_PF + 2: a slot is rewritten two steps after its last read)
constexpr int BL_ROWB = 160;                // staged bytes per row: columns [x0-16, x0+144)

// Rows are staged with cp.async (LDGSTS): unlike plain loads, whose completion is tracked by 6 shared scoreboard
// slots (ncu showed the 5-deep register prefetch waiting on the NEWEST load), async-copy groups complete in
// order and wait_group N leaves exactly N rows in flight.
__device__ __forceinline__ void cp_async16(unsigned smem_addr, const void* gmem) {   // smem_addr: shared-window address
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
// PRMT with a run-time selector (__byte_perm first masks the selector with 0x7777: one more instruction per use)
__device__ __forceinline__ unsigned bl_prmt(unsigned a, unsigned b, unsigned sel) {
    unsigned d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel));
    return d;
}

__device__ __forceinline__ int reflect101_once(int i, int n) {
    // BORDER_REFLECT_101: -i -> i, (n-1)+i -> (n-1)-i.  One bounce is exact for n >= 4 (halo is 3); smaller
    // levels never carry keypoints, there the clamp only keeps the access in bounds.
    if (i < 0) i = -i;
    if (i >= n) i = 2 * (n - 1) - i;
    return min(max(i, 0), n - 1);
}

template <bool EDGE, bool SAT>
__device__ __forceinline__ void blur_task(const uint8_t* __restrict__ src, int spitch, uint8_t* __restrict__ dst,
                                          int dpitch, int w, int h, int x, int y0, int rows, int lane, uint8_t* ring,
                                          const B2Geom& g) {
    // ---- per-lane load plan: lane c < 10 copies the 16-byte chunk c of every row (skipped outside the row) ----
    const int x0 = x - lane * 4;                               // strip origin (multiple of 128)
    const int cb = x0 - 16 + lane * 16;                        // first column of this lane's chunk
    const bool copier = lane < BL_ROWB / 16 && cb >= 0 && cb + 16 <= spitch;
    const uint8_t* csrc = src + (copier ? cb : 0);
    const unsigned cdst = (unsigned)__cvta_generic_to_shared(ring) + lane * 16;   // converted once, not per row
    const unsigned* myw = reinterpret_cast<const unsigned*>(ring + 16 + lane * 4);   // own word inside a ring row

    // ---- border selectors (EDGE strips only) --------------------------------------------------
    // Column c = x-4+j (j = 0..11) must read reflect101(c).  Needed columns always resolve inside the lane's
    // own three raw words (R0 | R1 | R2 = columns x-4 .. x+7); W0 and W1 pick from (R0,R1) or (R1,R2).
    unsigned s0 = 0x3210, s1 = 0x7654, s2 = 0x7654;
    bool w0_hi = false, w2_lo = false;                         // W0 from (R1,R2) / W2 from (R0,R1) instead
    if (EDGE) {
        int idx[12];
#pragma unroll
        for (int j = 0; j < 12; ++j) {
            const int c = x - 4 + j;
            int rc = c;
            if (c < 0) rc = -c;
            if (c >= w) rc = 2 * (w - 1) - c;
            rc = min(max(rc, 0), w - 1);
            int k = rc - (x - 4);                              // position inside R0|R1|R2
            idx[j] = min(max(k, 0), 11);                       // out of reach == never needed for a valid output
        }
        // W0: columns j = 0..3
        w0_hi = idx[1] >= 4 || idx[2] >= 4 || idx[3] >= 4;     // left image edge: sources are in R1 (and R2)
        s0 = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int k = w0_hi ? idx[j] - 4 : idx[j];
            s0 |= (unsigned)min(max(k, 0), 7) << (4 * j);
        }
        // W1: columns j = 4..7 from (R0,R1)
        s1 = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) s1 |= (unsigned)min(idx[4 + j], 7) << (4 * j);
        // W2: columns j = 8..11 from (R1,R2) unless a needed source lies in R0
        w2_lo = idx[8] < 4 || idx[9] < 4 || idx[10] < 4;
        s2 = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int k = w2_lo ? idx[8 + j] : idx[8 + j] - 4;
            s2 |= (unsigned)min(max(k, 0), 7) << (4 * j);
        }
    }

    // ---- row pipeline -------------------------------------------------------------------------
    // Vertical pass: the horizontal sums are 16-bit (<= 255 * 257), so consecutive rows travel in PAIRS P_j = (H_j, H_j+1)
    // and a pair meets two taps in one IDP.2A: out(r) = P_{r-6}.(t0,t1) + P_{r-4}.(t2,t3) + P_{r-2}.(t2,t1) + H_r * t0 + 2^15
    // — 1 PRMT + 1 IMAD + 3 IDP.2A per pixel instead of 3 IADD + 4 IMAD.  Six pair slots, 6-phase unrolled rotation.
    const int nsteps = rows + 6;
    const unsigned T0123 = (unsigned)g.taps[0] | ((unsigned)g.taps[1] << 8) | ((unsigned)g.taps[2] << 16) | ((unsigned)g.taps[3] << 24);
    const unsigned T21 = (unsigned)g.taps[2] | ((unsigned)g.taps[1] << 8);
    const unsigned t0 = (unsigned)g.taps[0];
    unsigned P[6][4];
    unsigned Hp[4] = {0, 0, 0, 0};
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) P[i][j] = 0;
#pragma unroll
    for (int i = 0; i < BL_PF; ++i) {
        if (i < nsteps && copier) cp_async16(cdst + i * BL_ROWB, csrc + (size_t)reflect101_once(y0 - 3 + i, h) * spitch);
        cp_async_commit();
    }
    for (int base = 0; base < nsteps; base += 6) {
#pragma unroll
        for (int p = 0; p < 6; ++p) {
            const int r = base + p;
            if (r < nsteps) {
                // prefetch row r + BL_PF (always commit, possibly an empty group, so the wait count stays fixed)
                if (r + BL_PF < nsteps && copier)
                    cp_async16(cdst + ((r + BL_PF) & (BL_SLOTS - 1)) * BL_ROWB,
                               csrc + (size_t)reflect101_once(y0 - 3 + r + BL_PF, h) * spitch);
                cp_async_commit();
                cp_async_wait<BL_PF>();                        // row r has landed (for this lane's chunk)
                __syncwarp();                                  // ... and for every other lane's
                const unsigned* rw = myw + (r & (BL_SLOTS - 1)) * (BL_ROWB / 4);
                const unsigned R0 = rw[-1], R1 = rw[0], R2 = rw[1];
                unsigned W0 = R0, W1 = R1, W2 = R2;
                if (EDGE) {
                    W0 = bl_prmt(w0_hi ? R1 : R0, w0_hi ? R2 : R1, s0);
                    W1 = bl_prmt(R0, R1, s1);
                    W2 = bl_prmt(w2_lo ? R0 : R1, w2_lo ? R1 : R2, s2);
                }
                // horizontal pass: 4 pixels, 10 dot products
                unsigned H[4];
                H[0] = __dp4a(W1, g.tapv[1], __dp4a(W0, g.tapv[0], 0u));
                H[1] = __dp4a(W2, g.tapv[4], __dp4a(W1, g.tapv[3], __dp4a(W0, g.tapv[2], 0u)));
                H[2] = __dp4a(W2, g.tapv[7], __dp4a(W1, g.tapv[6], __dp4a(W0, g.tapv[5], 0u)));
                H[3] = __dp4a(W2, g.tapv[9], __dp4a(W1, g.tapv[8], 0u));
                if (r >= 6) {
                    // vertical pass for output row y0 + r - 6 (window rows r-6 .. r, centre r-3)
                    unsigned o[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        unsigned acc = H[j] * t0 + 32768u;
                        acc = __dp2a_lo(P[p][j], T0123, acc);                  // rows r-6, r-5
                        acc = __dp2a_hi(P[(p + 2) % 6][j], T0123, acc);        // rows r-4, r-3
                        acc = __dp2a_lo(P[(p + 4) % 6][j], T21, acc);          // rows r-2, r-1
                        // taps summing to more than 256 (OpenCV 3.2 rounds the sigma-2 kernel to a sum of 257) can push
                        // the rounded value past 255: saturate_cast<uchar>, as FixedPtCastEx does
                        o[j] = SAT ? min(acc, 0x00FFFFFFu) : acc;
                    }
                    // bits 16..23 of each accumulator are the result (never above 255 when the taps sum to <= 256)
                    const unsigned lo = __byte_perm(o[0], o[1], 0x0062), hi = __byte_perm(o[2], o[3], 0x0062);
                    const unsigned packed = __byte_perm(lo, hi, 0x5410);
                    if (x < w) *reinterpret_cast<unsigned*>(dst + (size_t)(y0 + r - 6) * dpitch + x) = packed;
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    P[(p + 5) % 6][j] = __byte_perm(Hp[j], H[j], 0x5410);      // P_{r-1} = (H_{r-1}, H_r); its slot held P_{r-7}
                    Hp[j] = H[j];
                }
            }
        }
    }
}

__global__ void __launch_bounds__(BL_WARPS * 32)
blur_kernel(const __grid_constant__ B2Geom g, const B2Images im, int variant) {
    b2_pdl_wait();
    __shared__ __align__(16) uint8_t s_ring[BL_WARPS][BL_SLOTS * BL_ROWB];
    const int lane = threadIdx.x & 31;
    const int wid = __shfl_sync(0xFFFFFFFFu, (int)(threadIdx.x >> 5), 0);   // broadcast: the compiler then knows it is warp-uniform
    const int task = blockIdx.x * BL_WARPS + wid;
    if (task >= g.total_blur_tasks[variant]) return;
    const int band_rows = B2_BLUR_BAND << (2 * variant);
    const int b = blockIdx.y;
    int level = 0;
    while (level + 1 < g.nlevels && task >= g.lv[level + 1].blur_task_base[variant]) ++level;
    const B2Level& L = g.lv[level];
    const int t = task - L.blur_task_base[variant];
    const int band = t / L.blur_strips, strip = t - band * L.blur_strips;
    int spitch;
    const uint8_t* src = b2_level_ptr(g, im, b, level, spitch);
    uint8_t* dst = im.blur + (size_t)b * g.slab + L.off;
    const int x = strip * 128 + lane * 4;
    const int y0 = band * band_rows;
    const int rows = min(band_rows, L.h - y0);

    // a strip needs the border code if it touches column 0 or its last lane's halo reaches beyond w-1
    const bool edge = strip == 0 || strip * 128 + 128 + 3 > L.w - 1;
    uint8_t* ring = s_ring[wid];
    if (g.tap_sat) {   // warp-uniform (whole launch)
        if (edge) blur_task<true, true>(src, spitch, dst, L.pitch, L.w, L.h, x, y0, rows, lane, ring, g);
        else blur_task<false, true>(src, spitch, dst, L.pitch, L.w, L.h, x, y0, rows, lane, ring, g);
    } else {
        if (edge) blur_task<true, false>(src, spitch, dst, L.pitch, L.w, L.h, x, y0, rows, lane, ring, g);
        else blur_task<false, false>(src, spitch, dst, L.pitch, L.w, L.h, x, y0, rows, lane, ring, g);
    }
}

#ifdef B2ORB_VARIANTS
// ---------------------------------------------------------------------------------------------------------------
// TMA variant of the row staging (measured A/B, not in the default build: see DESIGN.md "TMA staging").  The cp.async
// ring above costs one LDGSTS (10 lanes x 16 B) + a reflected row address + commit + wait_group per row and warp.  Here
// one elected lane issues ONE cp.async.bulk.tensor.3d per EIGHT rows (box = 160 B x 8 rows x 1 image, completion on an
// mbarrier, out-of-bounds columns zero-filled by the copy engine) into a two-box ring.  Only bands whose 6 halo rows lie
// inside the image can use it (the box cannot reflect rows); edge bands take the cp.async path.
constexpr int BT_BOX = 8;                   // rows per TMA box
constexpr int BT_RING = 2 * BT_BOX;         // two boxes in flight per warp

struct B2BlurMaps {
    CUtensorMap m[B2_MAX_LEVELS + 1];       // [0] level 0 left block, [1] level 0 right block, [l + 1] level l >= 1
};

__device__ __forceinline__ void bt_wait(unsigned mbar, unsigned parity) {
    asm volatile("{\n\t.reg .pred P1;\n\tBT_WAIT:\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
                 "@P1 bra BT_DONE;\n\tbra BT_WAIT;\n\tBT_DONE:\n\t}" ::"r"(mbar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bt_issue(unsigned dst, const CUtensorMap* map, unsigned mbar, int x, int y, int z) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(BT_BOX * BL_ROWB) : "memory");
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(dst), "l"(map), "r"(mbar), "r"(x), "r"(y), "r"(z) : "memory");
}

template <bool EDGE, bool SAT>
__device__ __forceinline__ void blur_task_tma(const CUtensorMap* map, int z, uint8_t* __restrict__ dst, int dpitch, int w,
                                              int x, int y0, int rows, int lane, uint8_t* ring, unsigned mbar0,
                                              const B2Geom& g) {
    const int x0 = x - lane * 4;
    const unsigned ring_s = (unsigned)__cvta_generic_to_shared(ring);
    const unsigned* myw = reinterpret_cast<const unsigned*>(ring + 16 + lane * 4);
    unsigned s0 = 0x3210, s1 = 0x7654, s2 = 0x7654;
    bool w0_hi = false, w2_lo = false;
    if (EDGE) {   // same column selectors as blur_task
        int idx[12];
#pragma unroll
        for (int j = 0; j < 12; ++j) {
            const int c = x - 4 + j;
            int rc = c;
            if (c < 0) rc = -c;
            if (c >= w) rc = 2 * (w - 1) - c;
            rc = min(max(rc, 0), w - 1);
            idx[j] = min(max(rc - (x - 4), 0), 11);
        }
        w0_hi = idx[1] >= 4 || idx[2] >= 4 || idx[3] >= 4;
        s0 = s1 = s2 = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            s0 |= (unsigned)min(max(w0_hi ? idx[j] - 4 : idx[j], 0), 7) << (4 * j);
            s1 |= (unsigned)min(idx[4 + j], 7) << (4 * j);
        }
        w2_lo = idx[8] < 4 || idx[9] < 4 || idx[10] < 4;
#pragma unroll
        for (int j = 0; j < 4; ++j) s2 |= (unsigned)min(max(w2_lo ? idx[8 + j] : idx[8 + j] - 4, 0), 7) << (4 * j);
    }
    const int nsteps = rows + 6;
    const int nboxes = (nsteps + BT_BOX - 1) / BT_BOX;
    // the last box may reach below the rows this band needs; rows past the image are zero-filled and never used
    if (lane == 0) bt_issue(ring_s, map, mbar0, x0 - 16, y0 - 3, z);
    int hw[7][4];
    for (int base = 0; base < nsteps; base += 7) {
#pragma unroll
        for (int p = 0; p < 7; ++p) {
            const int r = base + p;
            if (r < nsteps) {
                if ((r & (BT_BOX - 1)) == 0) {             // warp-uniform: a new box starts
                    const int bx = r / BT_BOX;
                    __syncwarp();                          // every lane is done with box bx - 1 (its slot is reused now)
                    if (lane == 0 && bx + 1 < nboxes)
                        bt_issue(ring_s + ((bx + 1) & 1) * BT_BOX * BL_ROWB, map, mbar0 + 8 * ((bx + 1) & 1), x0 - 16,
                                 y0 - 3 + (bx + 1) * BT_BOX, z);
                    bt_wait(mbar0 + 8 * (bx & 1), (bx >> 1) & 1);
                }
                const unsigned* rw = myw + (r & (BT_RING - 1)) * (BL_ROWB / 4);
                const unsigned R0 = rw[-1], R1 = rw[0], R2 = rw[1];
                unsigned W0 = R0, W1 = R1, W2 = R2;
                if (EDGE) {
                    W0 = bl_prmt(w0_hi ? R1 : R0, w0_hi ? R2 : R1, s0);
                    W1 = bl_prmt(R0, R1, s1);
                    W2 = bl_prmt(w2_lo ? R0 : R1, w2_lo ? R1 : R2, s2);
                }
                hw[p][0] = __dp4a(W1, g.tapv[1], __dp4a(W0, g.tapv[0], 0u));
                hw[p][1] = __dp4a(W2, g.tapv[4], __dp4a(W1, g.tapv[3], __dp4a(W0, g.tapv[2], 0u)));
                hw[p][2] = __dp4a(W2, g.tapv[7], __dp4a(W1, g.tapv[6], __dp4a(W0, g.tapv[5], 0u)));
                hw[p][3] = __dp4a(W2, g.tapv[9], __dp4a(W1, g.tapv[8], 0u));
                if (r >= 6) {
                    unsigned o[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int e0 = hw[p][j] + hw[(p + 1) % 7][j];
                        const int e1 = hw[(p + 6) % 7][j] + hw[(p + 2) % 7][j];
                        const int e2 = hw[(p + 5) % 7][j] + hw[(p + 3) % 7][j];
                        o[j] = (unsigned)(g.taps[0] * e0 + g.taps[1] * e1 + g.taps[2] * e2 + g.taps[3] * hw[(p + 4) % 7][j] + 32768);
                        if (SAT) o[j] = min(o[j], 0x00FFFFFFu);
                    }
                    const unsigned lo = __byte_perm(o[0], o[1], 0x0062), hi = __byte_perm(o[2], o[3], 0x0062);
                    if (x < w) *reinterpret_cast<unsigned*>(dst + (size_t)(y0 + r - 6) * dpitch + x) = __byte_perm(lo, hi, 0x5410);
                }
            }
        }
    }
}

__global__ void __launch_bounds__(BL_WARPS * 32)
blur_tma_kernel(const __grid_constant__ B2Geom g, const B2Images im, int variant, const __grid_constant__ B2BlurMaps maps) {
    b2_pdl_wait();
    __shared__ __align__(128) uint8_t s_ring[BL_WARPS][BT_RING * BL_ROWB];
    __shared__ __align__(8) unsigned long long s_mbar[BL_WARPS][2];
    const int lane = threadIdx.x & 31, wid = __shfl_sync(0xFFFFFFFFu, (int)(threadIdx.x >> 5), 0);
    const int task = blockIdx.x * BL_WARPS + wid;
    if (task >= g.total_blur_tasks[variant]) return;
    const int band_rows = B2_BLUR_BAND << (2 * variant);
    const int b = blockIdx.y;
    int level = 0;
    while (level + 1 < g.nlevels && task >= g.lv[level + 1].blur_task_base[variant]) ++level;
    const B2Level& L = g.lv[level];
    const int t = task - L.blur_task_base[variant];
    const int band = t / L.blur_strips, strip = t - band * L.blur_strips;
    int spitch;
    const uint8_t* src = b2_level_ptr(g, im, b, level, spitch);
    uint8_t* dst = im.blur + (size_t)b * g.slab + L.off;
    const int x = strip * 128 + lane * 4;
    const int y0 = band * band_rows;
    const int rows = min(band_rows, L.h - y0);
    const bool edge = strip == 0 || strip * 128 + 128 + 3 > L.w - 1;
    uint8_t* ring = s_ring[wid];
    const bool interior = y0 >= 3 && y0 + rows + 3 <= L.h;      // all halo rows exist: no row reflection needed
    if (!interior) {   // top / bottom bands: the cp.async path (rows are reflected per row)
        if (g.tap_sat) {
            if (edge) blur_task<true, true>(src, spitch, dst, L.pitch, L.w, L.h, x, y0, rows, lane, ring, g);
            else blur_task<false, true>(src, spitch, dst, L.pitch, L.w, L.h, x, y0, rows, lane, ring, g);
        } else {
            if (edge) blur_task<true, false>(src, spitch, dst, L.pitch, L.w, L.h, x, y0, rows, lane, ring, g);
            else blur_task<false, false>(src, spitch, dst, L.pitch, L.w, L.h, x, y0, rows, lane, ring, g);
        }
        return;
    }
    const unsigned mbar0 = (unsigned)__cvta_generic_to_shared(&s_mbar[wid][0]);
    if (lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mbar0));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mbar0 + 8));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    const CUtensorMap* map = level == 0 ? &maps.m[b < im.split ? 0 : 1] : &maps.m[level + 1];
    const int z = level == 0 ? (b < im.split ? b : b - im.split) : b;
    if (g.tap_sat) {
        if (edge) blur_task_tma<true, true>(map, z, dst, L.pitch, L.w, x, y0, rows, lane, ring, mbar0, g);
        else blur_task_tma<false, true>(map, z, dst, L.pitch, L.w, x, y0, rows, lane, ring, mbar0, g);
    } else {
        if (edge) blur_task_tma<true, false>(map, z, dst, L.pitch, L.w, x, y0, rows, lane, ring, mbar0, g);
        else blur_task_tma<false, false>(map, z, dst, L.pitch, L.w, x, y0, rows, lane, ring, mbar0, g);
    }
}
#endif  // B2ORB_VARIANTS

}  // namespace

void b2_launch_blur(const B2Geom& g, const B2Images& im, int nimg, int forced_band, cudaStream_t s) {
    // long bands once the batch alone fills the machine several times over (6 halo rows per band)
    int variant = (long long)g.total_blur_tasks[1] * nimg >= 148LL * 16 * 4 ? 1 : 0;
    if (forced_band >= 0 && forced_band < B2_BLUR_VARIANTS) variant = forced_band;   // b2orb_debug_set("blur_band")
    dim3 grid((g.total_blur_tasks[variant] + BL_WARPS - 1) / BL_WARPS, nimg);
#ifdef B2ORB_VARIANTS
    if (forced_band >= 16) {   // b2orb_debug_set("blur_band", 16 + v): the TMA-staged variant with band variant v
        variant = min(forced_band - 16, B2_BLUR_VARIANTS - 1);
        grid = dim3((g.total_blur_tasks[variant] + BL_WARPS - 1) / BL_WARPS, nimg);
        typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                     const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                     CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
        static EncodeFn encode = nullptr;
        if (!encode) {
            void* fn = nullptr;
            cudaDriverEntryPointQueryResult qr;
            if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr) != cudaSuccess || !fn) return;
            encode = (EncodeFn)fn;
        }
        B2BlurMaps maps;
        memset(&maps, 0, sizeof maps);
        const cuuint32_t box[3] = {BL_ROWB, BT_BOX, 1}, es[3] = {1, 1, 1};
        auto mk = [&](CUtensorMap* m, const void* base, int pitch, int h, size_t img_stride, int n) {
            const cuuint64_t dims[3] = {(cuuint64_t)pitch, (cuuint64_t)h, (cuuint64_t)max(n, 1)};
            const cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)img_stride};
            return encode(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(base), dims, strides, box, es,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
        };
        bool ok = mk(&maps.m[0], im.lvl0[0], im.pitch0, g.lv[0].h, im.stride0, im.split) &&
                  mk(&maps.m[1], im.lvl0[1], im.pitch0, g.lv[0].h, im.stride0, nimg - im.split);
        for (int l = 1; l < g.nlevels && ok; ++l)
            ok = mk(&maps.m[l + 1], im.pyr + g.lv[l].off, g.lv[l].pitch, g.lv[l].h, g.slab, nimg);
        if (ok) {
            b2_launch_pdl(blur_tma_kernel, dim3(grid), dim3(BL_WARPS * 32), 0, s, g, im, variant, maps);
            return;
        }
    }
#endif
    b2_launch_pdl(blur_kernel, dim3(grid), dim3(BL_WARPS * 32), 0, s, g, im, variant);
}
