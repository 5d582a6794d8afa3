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
// dot products against shifted tap vectors — no byte unpacking at all).  The vertical pass keeps the last 7
// horizontal rows in registers (7-phase unrolled rotation) and uses tap symmetry: 3 IADD + 4 IMAD per pixel,
// which also balances the ALU and FMA pipes.  Results are packed with 3 PRMT and stored as one 32-bit word.
// Image borders (reflect-101) are folded into per-lane PRMT selectors computed once per strip; interior strips
// skip that code entirely.
// One launch covers all levels of all images: blockIdx.x walks a flattened (level, band, strip) task list.
#include "b2orb_internal.cuh"

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
    const int nsteps = rows + 6;
    int hw[7][4];
#pragma unroll
    for (int i = 0; i < BL_PF; ++i) {
        if (i < nsteps && copier) cp_async16(cdst + i * BL_ROWB, csrc + (size_t)reflect101_once(y0 - 3 + i, h) * spitch);
        cp_async_commit();
    }
    for (int base = 0; base < nsteps; base += 7) {
#pragma unroll
        for (int p = 0; p < 7; ++p) {
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
                    W0 = __byte_perm(w0_hi ? R1 : R0, w0_hi ? R2 : R1, s0);
                    W1 = __byte_perm(R0, R1, s1);
                    W2 = __byte_perm(w2_lo ? R0 : R1, w2_lo ? R1 : R2, s2);
                }
                // horizontal pass: 4 pixels, 10 dot products
                hw[p][0] = __dp4a(W1, g.tapv[1], __dp4a(W0, g.tapv[0], 0u));
                hw[p][1] = __dp4a(W2, g.tapv[4], __dp4a(W1, g.tapv[3], __dp4a(W0, g.tapv[2], 0u)));
                hw[p][2] = __dp4a(W2, g.tapv[7], __dp4a(W1, g.tapv[6], __dp4a(W0, g.tapv[5], 0u)));
                hw[p][3] = __dp4a(W2, g.tapv[9], __dp4a(W1, g.tapv[8], 0u));
                if (r >= 6) {
                    // vertical pass for output row y0 + r - 6 (window rows r-6 .. r, centre r-3)
                    unsigned o[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int e0 = hw[p][j] + hw[(p + 1) % 7][j];
                        const int e1 = hw[(p + 6) % 7][j] + hw[(p + 2) % 7][j];
                        const int e2 = hw[(p + 5) % 7][j] + hw[(p + 3) % 7][j];
                        o[j] = (unsigned)(g.taps[0] * e0 + g.taps[1] * e1 + g.taps[2] * e2 + g.taps[3] * hw[(p + 4) % 7][j] + 32768);
                        // taps summing to more than 256 (OpenCV 3.2 rounds the sigma-2 kernel to a sum of 257) can push
                        // the rounded value past 255: saturate_cast<uchar>, as FixedPtCastEx does
                        if (SAT) o[j] = min(o[j], 0x00FFFFFFu);
                    }
                    // bits 16..23 of each accumulator are the result (never above 255 when the taps sum to <= 256)
                    const unsigned lo = __byte_perm(o[0], o[1], 0x0062), hi = __byte_perm(o[2], o[3], 0x0062);
                    const unsigned packed = __byte_perm(lo, hi, 0x5410);
                    if (x < w) *reinterpret_cast<unsigned*>(dst + (size_t)(y0 + r - 6) * dpitch + x) = packed;
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
    const int task = blockIdx.x * BL_WARPS + (threadIdx.x >> 5);
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
    uint8_t* ring = s_ring[threadIdx.x >> 5];
    if (g.tap_sat) {   // warp-uniform (whole launch)
        if (edge) blur_task<true, true>(src, spitch, dst, L.pitch, L.w, L.h, x, y0, rows, lane, ring, g);
        else blur_task<false, true>(src, spitch, dst, L.pitch, L.w, L.h, x, y0, rows, lane, ring, g);
    } else {
        if (edge) blur_task<true, false>(src, spitch, dst, L.pitch, L.w, L.h, x, y0, rows, lane, ring, g);
        else blur_task<false, false>(src, spitch, dst, L.pitch, L.w, L.h, x, y0, rows, lane, ring, g);
    }
}

}  // namespace

void b2_launch_blur(const B2Geom& g, const B2Images& im, int nimg, int forced_band, cudaStream_t s) {
    // long bands once the batch alone fills the machine several times over (6 halo rows per band)
    int variant = (long long)g.total_blur_tasks[1] * nimg >= 148LL * 16 * 4 ? 1 : 0;
    if (forced_band >= 0 && forced_band < B2_BLUR_VARIANTS) variant = forced_band;   // b2orb_debug_set("blur_band")
    dim3 grid((g.total_blur_tasks[variant] + BL_WARPS - 1) / BL_WARPS, nimg);
    b2_launch_pdl(blur_kernel, dim3(grid), dim3(BL_WARPS * 32), 0, s, g, im, variant);
}
