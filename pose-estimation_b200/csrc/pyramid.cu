// Image pyramid: level l = cv::resize(level l-1, INTER_LINEAR) — ORBextractor::ComputePyramid,
// src/ORBextractor.cc:1107-1132 (K1 in SURVEY.md §2.3).  Fixed-point arithmetic of OpenCV's 8-bit
// bilinear path: Q11 weights per axis (host-computed tables, exactly as cv::resize builds them),
// horizontal pass in int32, vertical pass ((b0*(S0>>4))>>16) + ((b1*(S1>>4))>>16) + 2) >> 2.
// The 19-px reflect apron of the reference (copyMakeBorder, :1122,1127) is never read by later
// stages, so levels are stored un-padded with a 16-byte-aligned pitch.
#include "b2orb_internal.cuh"

namespace {

constexpr int RS_WARPS = 4;     // independent warp tasks per CTA
constexpr int RS_PF = 4;        // source rows in flight per warp
constexpr int RS_SLOTS = 8;     // shared-memory ring (>= RS_PF + 2)
constexpr int RS_ROWB = 512;    // staged bytes per source row: one 16-byte chunk per lane (scale <= 3 -> <= 26 chunks)

__device__ __forceinline__ void rs_cp_async16(unsigned smem_addr, const void* gmem) {   // shared-window address
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr), "l"(gmem) : "memory");
}
__device__ __forceinline__ unsigned rs_lds(unsigned smem_addr) {
    unsigned v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(smem_addr));
    return v;
}
__device__ __forceinline__ void rs_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void rs_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
// PRMT with a run-time selector (__byte_perm first masks the selector with 0x7777: one more instruction per use)
__device__ __forceinline__ unsigned rs_prmt(unsigned a, unsigned b, unsigned sel) {
    unsigned d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel));
    return d;
}

// One WARP owns a 128-px wide strip of the destination level (4 px per lane) and walks down RS_BAND output rows.
// The source rows it needs are streamed through a cp.async shared-memory ring (deep, in-order prefetch).
// Horizontal pass of one source row for the lane's 4 pixels: the two source bytes of pixel j are cut out of two
// aligned words with one PRMT (selector fixed per lane) and combined with one IDP.2A against the packed Q11
// weights (a0 | a1 << 16).  Every source row is interpolated once and used by the 1-2 output rows that need it.
// Vertical pass: two IMAD.HI with the weights pre-shifted by 16, one add; the rounding shift is folded into a
// multiply by 64 so the result byte can be picked with PRMT.
// One task = (band, strip) of `level` of image b, run by the calling warp with its private ring.
__device__ __forceinline__ void resize_task(const B2Geom& g, const B2Images& im, int b, int level, int band_rows, int task,
                                            const int* __restrict__ xofs, const short* __restrict__ xw,
                                            const int* __restrict__ yofs, const short* __restrict__ yw, uint8_t* ring, uint4* vtab) {
    const B2Level& L = g.lv[level];
    const int lane = threadIdx.x & 31;
    const int strips = (L.w + 127) >> 7;
    const int band = task / strips, strip = task - band * strips;
    const int dy0 = band * band_rows;
    if (dy0 >= L.h) return;
    const int dy1 = min(dy0 + band_rows, L.h);
    int spitch;
    const uint8_t* src = b2_level_ptr(g, im, b, level - 1, spitch);
    const int sh = g.lv[level - 1].h;
    uint8_t* dst = im.pyr + (size_t)b * g.slab + L.off;
    const int dx0 = strip * 128 + lane * 4;

    // source span of the strip, 16-byte aligned; lane c copies chunk c of every source row
    const int gx0 = xofs[strip * 128] & ~15;
    const int gx1 = xofs[min(strip * 128 + 127, L.w - 1)] + 2;           // one past the last byte used
    const int cb = gx0 + lane * 16;
    const bool copier = cb < gx1 && cb + 16 <= spitch;
    const unsigned ring_s = (unsigned)__cvta_generic_to_shared(ring);   // all ring addressing in 32-bit shared space
    const unsigned cdst = ring_s + lane * 16;

    // per-lane horizontal plan
    unsigned A[4], sel[4];
    unsigned o01, o23;                                                    // ring-row byte offsets of the two word pairs
    {
        int sx[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int dx = min(dx0 + j, L.w - 1);
            sx[j] = xofs[dx] - gx0;
            A[j] = (unsigned)(unsigned short)xw[2 * dx] | ((unsigned)(unsigned short)xw[2 * dx + 1] << 16);
        }
        const int b01 = sx[0] & ~3, b23 = sx[2] & ~3;
        // scale <= 3: sx[1] - b01 <= 6 and sx[3] - b23 <= 6, so both bytes of a pixel sit in the 8 bytes at its base
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int o = min(sx[j] - (j < 2 ? b01 : b23), 6);
            sel[j] = (unsigned)o | ((unsigned)(o + 1) << 4);
        }
        o01 = ring_s + (unsigned)min(b01, RS_ROWB - 8);
        o23 = ring_s + (unsigned)min(b23, RS_ROWB - 8);
    }

    const int s_first = yofs[dy0];
    const int s_last = min(yofs[dy1 - 1] + 1, sh - 1);
    const int nrows = s_last - s_first + 1;
    const uint8_t* gsrc = src + (size_t)s_first * spitch + (copier ? cb : 0);   // next source row to issue
    // Loop invariants the compiler would otherwise rebuild every row (shared-window base via S2R, constant-bank
    // reloads of the level geometry): pinned in registers.
    unsigned cdst_r = cdst;
    int sh1 = sh - 1, dpitch = L.pitch, sp = spitch;
    asm volatile("" : "+r"(cdst_r), "+r"(sh1), "+r"(dpitch), "+r"(sp));
    int issued = 0;
#pragma unroll
    for (int i = 0; i < RS_PF; ++i) {
        if (issued < nrows && copier) rs_cp_async16(cdst_r + i * RS_ROWB, gsrc);
        rs_commit();
        gsrc += sp;
        ++issued;
    }
    unsigned HA[4] = {0, 0, 0, 0}, HB[4] = {0, 0, 0, 0};
    uint8_t* dptr = dst + (size_t)dy0 * dpitch + dx0;                     // next output row of this lane
    int store = dx0 < L.w;
    asm volatile("" : "+r"(store));
    // vertical table of the band in shared memory: entry e = (source row, b0 << 16, b1 << 16) of output row dy0 + e, one
    // LDS.128 per output row (lane e fills entry e; bands have at most 32 rows)
    const unsigned vt_s = (unsigned)__cvta_generic_to_shared(vtab);
    {
        const int dy = min(dy0 + lane, dy1 - 1);
        const unsigned w2 = *reinterpret_cast<const unsigned*>(yw + 2 * dy);
        vtab[lane] = make_uint4((unsigned)yofs[dy], w2 << 16, w2 & 0xFFFF0000u, 0u);
    }
    __syncwarp();
    int left = dy1 - dy0;                                                 // output rows still to emit
    unsigned vt_next = vt_s;                                              // address of the next output row's entry
    int sy; unsigned b0, b1;
    const auto next_row = [&]() {
        unsigned x, y, z, w_;
        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(x), "=r"(y), "=r"(z), "=r"(w_) : "r"(vt_next));
        sy = (int)x; b0 = y; b1 = z;
        vt_next += 16;
    };
    next_row();
    unsigned slot = 0;                                                    // ring byte offset of source row s
    // one source row: stage the next one, interpolate this one horizontally into Hc, emit the output rows whose source
    // pair is (s - 1, s) from (Hp, Hc).  Called with the two register sets swapped on alternate rows: no copies.
    const auto step = [&](int s, unsigned (&Hp)[4], unsigned (&Hc)[4]) {
        if (issued < nrows && copier) rs_cp_async16(cdst_r + ((slot + RS_PF * RS_ROWB) & (RS_SLOTS * RS_ROWB - 1)), gsrc);
        rs_commit();
        gsrc += sp;
        ++issued;
        rs_wait<RS_PF>();
        __syncwarp();
        const unsigned wa0 = rs_lds(o01 + slot), wa1 = rs_lds(o01 + slot + 4);
        const unsigned wb0 = rs_lds(o23 + slot), wb1 = rs_lds(o23 + slot + 4);
        slot = (slot + RS_ROWB) & (RS_SLOTS * RS_ROWB - 1);
        Hc[0] = __dp2a_lo(A[0], rs_prmt(wa0, wa1, sel[0]), 0u) >> 4;
        Hc[1] = __dp2a_lo(A[1], rs_prmt(wa0, wa1, sel[1]), 0u) >> 4;
        Hc[2] = __dp2a_lo(A[2], rs_prmt(wb0, wb1, sel[2]), 0u) >> 4;
        Hc[3] = __dp2a_lo(A[3], rs_prmt(wb0, wb1, sel[3]), 0u) >> 4;
        // rows come in ascending source order, so once sy == s the pair (s-1, s) is no longer needed
        const auto emit = [&](const unsigned (&H0)[4], const unsigned (&H1)[4]) {
            unsigned u[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) u[j] = (__umulhi(b0, H0[j]) + __umulhi(b1, H1[j]) + 2u) * 64u;   // byte 1 = (..) >> 2
            const unsigned lo = __byte_perm(u[0], u[1], 0x0051), hi = __byte_perm(u[2], u[3], 0x0051);
            if (store) *reinterpret_cast<unsigned*>(dptr) = __byte_perm(lo, hi, 0x5410);
            dptr += dpitch;
            --left;
            next_row();                                                   // (one entry past the band is read, never used)
        };
        while (left > 0 && sy == s - 1) emit(Hp, Hc);
        if (s == sh1)                                                     // last source row of the level: rows clamped to
            while (left > 0 && sy == s) emit(Hc, Hc);                     // (s, s), their second weight is 0
    };
    int s = s_first;
    for (; s + 1 <= s_last; s += 2) {
        step(s, HA, HB);
        step(s + 1, HB, HA);
    }
    if (s <= s_last) step(s, HA, HB);
}

// Level-0 ingest: images with arbitrary row step (e.g. tight 1241-byte rows) -> staging area whose rows start on
// 32-byte boundaries, which is what every other kernel's 32/128-bit accesses assume.  One thread per output word.
__global__ void __launch_bounds__(256)
repack_kernel(const uint8_t* __restrict__ src, int step, size_t stride, uint8_t* __restrict__ dst, int dpitch,
              size_t dstride, int w, int h) {
    b2_pdl_wait();
    const int words = dpitch >> 2;
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= words * h) return;
    const int y = i / words, xw = (i - y * words) * 4;
    const uint8_t* s = src + blockIdx.y * stride + (size_t)y * step + xw;
    uint32_t v = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k)
        if (xw + k < w) v |= (uint32_t)s[k] << (8 * k);
    *reinterpret_cast<uint32_t*>(dst + blockIdx.y * dstride + (size_t)y * dpitch + xw) = v;
}

// Latency-path ingest: the image(s) of a one-pair call are read by the kernel straight out of the caller's PINNED host
// memory and written to the pitched staging area — one pass over the bus instead of a DMA copy into a raw buffer followed
// by the re-pitching kernel (one KITTI pair: 21 us of copies + 16 us of repack launches -> one launch bound by the link).
// A CTA owns IH_ROWS rows: their source bytes are one contiguous range (rows are `step` apart), fetched with aligned
// 16-byte non-coherent loads into shared memory (every source byte crosses the bus once; a 16-byte chunk never crosses a
// page, so the aligned over-read at both ends stays inside mapped memory), then re-pitched from shared memory.
constexpr int IH_ROWS = 8;
constexpr int IH_MAX_STEP = 4096;
__global__ void __launch_bounds__(256)
ingest_host_kernel(const uint8_t* __restrict__ src0, const uint8_t* __restrict__ src1, int step, uint8_t* __restrict__ dst,
                   int dpitch, size_t dstride, int w, int h) {
    b2_pdl_wait();
    __shared__ __align__(16) uint8_t s_rows[IH_ROWS * IH_MAX_STEP + 32];
    const uint8_t* src = blockIdx.y ? src1 : src0;
    const int y0 = blockIdx.x * IH_ROWS, ny = min(IH_ROWS, h - y0);
    const uintptr_t first = (uintptr_t)src + (size_t)y0 * step;
    const uintptr_t a0 = first & ~(uintptr_t)15;
    const int skew = (int)(first - a0);
    const int nbytes = (ny - 1) * step + w;
    const int nchunk = (skew + nbytes + 15) >> 4;
    for (int c = threadIdx.x; c < nchunk; c += 256) {
        uint4 v;
        asm volatile("ld.global.nc.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                     : "l"(a0 + (size_t)c * 16));
        reinterpret_cast<uint4*>(s_rows)[c] = v;
    }
    __syncthreads();
    const int words = dpitch >> 2;
    uint8_t* d = dst + blockIdx.y * dstride + (size_t)y0 * dpitch;
    for (int i = threadIdx.x; i < words * ny; i += 256) {
        const int r = i / words, xw = (i - r * words) * 4;
        const uint8_t* s = s_rows + skew + r * step + xw;
        uint32_t v = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k)
            if (xw + k < w) v |= (uint32_t)s[k] << (8 * k);
        *reinterpret_cast<uint32_t*>(d + (size_t)r * dpitch + xw) = v;
    }
}

// Rectification in front of the pyramid — cv::remap(gray, img, rmap[eye][0], rmap[eye][1], INTER_LINEAR),
// src/feature-tracker.cpp:75-76 (maps: include/cfsd/camera-model.hpp:76-77, CV_16SC2 + CV_16UC1).  map1 = (sx, sy),
// map2 = fx | fy << 5 in 1/32 px; OpenCV's fixed-point weights are exact multiples of 32 that sum to 32768, so
//   out = ((32-fy)(32-fx) I00 + (32-fy) fx I01 + fy (32-fx) I10 + fy fx I11 + 512) >> 10,   taps outside the image = 0
// (BORDER_CONSTANT).  One thread per output word; the result goes straight into the 32-byte pitched staging area, so
// it replaces the re-pitching pass (no extra image round trip).  A pure gather: 6 map bytes + 4 taps per pixel.
__global__ void __launch_bounds__(256)
remap_kernel(const uint8_t* __restrict__ src, int step, size_t stride, const short2* __restrict__ map1,
             const unsigned short* __restrict__ map2, uint8_t* __restrict__ dst, int dpitch, size_t dstride, int w, int h) {
    b2_pdl_wait();
    const int words = dpitch >> 2;
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= words * h) return;
    const int y = i / words, xw = (i - y * words) * 4;
    const uint8_t* s = src + blockIdx.y * stride;
    const size_t mrow = (size_t)y * w;
    uint32_t v = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (xw + k >= w) break;
        const short2 p = map1[mrow + xw + k];
        const int m = map2[mrow + xw + k];
        const int sx = p.x, sy = p.y, fx = m & 31, fy = (m >> 5) & 31;
        const bool x0 = sx >= 0 && sx < w, x1 = sx + 1 >= 0 && sx + 1 < w;
        const bool y0 = sy >= 0 && sy < h, y1 = sy + 1 >= 0 && sy + 1 < h;
        const uint8_t* r0 = s + (size_t)sy * step + sx;
        const uint8_t* r1 = r0 + step;
        const int i00 = (x0 && y0) ? r0[0] : 0, i01 = (x1 && y0) ? r0[1] : 0;
        const int i10 = (x0 && y1) ? r1[0] : 0, i11 = (x1 && y1) ? r1[1] : 0;
        const int acc = (32 - fy) * ((32 - fx) * i00 + fx * i01) + fy * ((32 - fx) * i10 + fx * i11);
        v |= (uint32_t)((acc + 512) >> 10) << (8 * k);
    }
    *reinterpret_cast<uint32_t*>(dst + blockIdx.y * dstride + (size_t)y * dpitch + xw) = v;
}

// cfsd live ingest (next row f2), src/cfsd-state-estimation.cpp:117-127: CV_8UC4 frame with both eyes side by side ->
__global__ void __launch_bounds__(RS_WARPS * 32)
resize_kernel(const __grid_constant__ B2Geom g, const B2Images im, int level, int band_rows,
              const int* __restrict__ xofs, const short* __restrict__ xw, const int* __restrict__ yofs,
              const short* __restrict__ yw) {
    b2_pdl_wait();
    __shared__ __align__(16) uint8_t s_ring[RS_WARPS][RS_SLOTS * RS_ROWB];
    __shared__ uint4 s_vtab[RS_WARPS][33];                                  // (one entry past a 32-row band is read)
    const int wid = __shfl_sync(0xFFFFFFFFu, (int)(threadIdx.x >> 5), 0);   // broadcast: warp-uniform for the compiler
    resize_task(g, im, blockIdx.y, level, band_rows, blockIdx.x * RS_WARPS + wid, xofs, xw, yofs, yw, s_ring[wid], s_vtab[wid]);
}

// cv::resize(INTER_LINEAR) to 2*half_w x dh -> cvtColor(BGR2GRAY) -> left / right halves, fused: one thread per output
// word of one eye; four 32-bit pixel loads per output pixel, the Q11 resize arithmetic of resize_kernel per channel
// (H = S0*wx0 + S1*wx1;  v = ((wy0*(H0>>4))>>16) + ((wy1*(H1>>4))>>16) + 2) >> 2), then OpenCV 4.x's 15-bit grey
// weights (B*3735 + G*19235 + R*9798 + 16384) >> 15.  The 7.4 MB colour frame is read once; nothing but the two grey
// eyes is written.
__global__ void __launch_bounds__(256)
ingest_bgra_kernel(const uint8_t* __restrict__ frames, int sstep, size_t sstride, int sw, int sh,
                   const int* __restrict__ xofs, const short* __restrict__ xw, const int* __restrict__ yofs,
                   const short* __restrict__ yw, uint8_t* __restrict__ dst_l, uint8_t* __restrict__ dst_r, int dpitch,
                   size_t dstride, int half_w, int dh) {
    b2_pdl_wait();
    const int words = dpitch >> 2;
    int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= 2 * words * dh) return;
    const int eye = i >= words * dh;
    i -= eye * words * dh;
    const int y = i / words, x4 = (i - y * words) * 4;
    const uint8_t* fr = frames + blockIdx.y * sstride;
    const int sy0 = yofs[y], sy1 = min(sy0 + 1, sh - 1);
    const int wy0 = yw[2 * y], wy1 = yw[2 * y + 1];
    const uint8_t* r0 = fr + (size_t)sy0 * sstep;
    const uint8_t* r1 = fr + (size_t)sy1 * sstep;
    uint32_t out = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int x = x4 + k;
        if (x >= half_w) break;
        const int X = x + eye * half_w;
        const int sx0 = xofs[X], sx1 = min(sx0 + 1, sw - 1);
        const int wx0 = xw[2 * X], wx1 = xw[2 * X + 1];
        const uint32_t p00 = *reinterpret_cast<const uint32_t*>(r0 + 4 * sx0), p01 = *reinterpret_cast<const uint32_t*>(r0 + 4 * sx1);
        const uint32_t p10 = *reinterpret_cast<const uint32_t*>(r1 + 4 * sx0), p11 = *reinterpret_cast<const uint32_t*>(r1 + 4 * sx1);
        int ch[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const int h0 = (int)((p00 >> (8 * c)) & 255u) * wx0 + (int)((p01 >> (8 * c)) & 255u) * wx1;
            const int h1 = (int)((p10 >> (8 * c)) & 255u) * wx0 + (int)((p11 >> (8 * c)) & 255u) * wx1;
            const int v = (((wy0 * (h0 >> 4)) >> 16) + ((wy1 * (h1 >> 4)) >> 16) + 2) >> 2;
            ch[c] = min(max(v, 0), 255);
        }
        out |= (uint32_t)((ch[0] * 3735 + ch[1] * 19235 + ch[2] * 9798 + 16384) >> 15) << (8 * k);
    }
    uint8_t* d = (eye ? dst_r : dst_l) + blockIdx.y * dstride + (size_t)y * dpitch + x4;
    *reinterpret_cast<uint32_t*>(d) = out;
}

}  // namespace

void b2_launch_ingest_bgra(const uint8_t* frames, int sstep, size_t sstride, int sw, int sh, const int* xofs, const short* xw,
                           const int* yofs, const short* yw, uint8_t* dst_l, uint8_t* dst_r, int dpitch, size_t dstride,
                           int half_w, int dh, int count, cudaStream_t s) {
    dim3 grid((2 * (dpitch >> 2) * dh + 255) / 256, count);
    b2_launch_pdl(ingest_bgra_kernel, dim3(grid), dim3(256), 0, s, frames, sstep, sstride, sw, sh, xofs, xw, yofs, yw, dst_l, dst_r, dpitch, dstride,
                                            half_w, dh);
}

void b2_launch_remap(const uint8_t* src, int step, size_t stride, const void* map1, const void* map2, uint8_t* dst,
                     int dpitch, size_t dstride, int w, int h, int count, cudaStream_t s) {
    dim3 grid(((dpitch >> 2) * h + 255) / 256, count);
    b2_launch_pdl(remap_kernel, dim3(grid), dim3(256), 0, s, src, step, stride, static_cast<const short2*>(map1),
                                      static_cast<const unsigned short*>(map2), dst, dpitch, dstride, w, h);
}

void b2_launch_resize(const B2Geom& g, const B2Images& im, int level, const int* xofs, const short* xw,
                      const int* yofs, const short* yw, int nimg, cudaStream_t s) {
    const B2Level& L = g.lv[level];
    // A task is one latency chain (a row every few hundred cycles), so the band shrinks until the launch has
    // enough warps to fill the machine (148 SMs x 48 warps); 32 rows amortise the per-task setup best.
    const int strips = (L.w + 127) / 128;
    int band = 32;
    while (band > 4 && (long long)strips * ((L.h + band - 1) / band) * nimg < 148 * 48) band >>= 1;
    const int tasks = strips * ((L.h + band - 1) / band);
    dim3 grid((tasks + RS_WARPS - 1) / RS_WARPS, nimg);
    b2_launch_pdl(resize_kernel, dim3(grid), dim3(RS_WARPS * 32), 0, s, g, im, level, band, xofs, xw, yofs, yw);
}

bool b2_launch_ingest_host(const uint8_t* src0, const uint8_t* src1, int step, uint8_t* dst, int dpitch, size_t dstride,
                           int w, int h, cudaStream_t s) {
    if (step > IH_MAX_STEP || w < 1 || h < 1) return false;
    dim3 grid((h + IH_ROWS - 1) / IH_ROWS, src1 ? 2 : 1);
    b2_launch_pdl(ingest_host_kernel, dim3(grid), dim3(256), 0, s, src0, src1, step, dst, dpitch, dstride, w, h);
    return true;
}

void b2_launch_repack(const uint8_t* src, int step, size_t stride, uint8_t* dst, int dpitch, size_t dstride, int w,
                      int h, int count, cudaStream_t s) {
    dim3 grid(((dpitch >> 2) * h + 255) / 256, count);
    b2_launch_pdl(repack_kernel, dim3(grid), dim3(256), 0, s, src, step, stride, dst, dpitch, dstride, w, h);
}
