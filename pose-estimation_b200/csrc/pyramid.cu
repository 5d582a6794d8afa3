// Image pyramid: level l = cv::resize(level l-1, INTER_LINEAR) — ORBextractor::ComputePyramid,
// src/ORBextractor.cc:1107-1132 (K1 in SURVEY.md §2.3).  Fixed-point arithmetic of OpenCV's 8-bit
// bilinear path: Q11 weights per axis (host-computed tables, exactly as cv::resize builds them),
// horizontal pass in int32, vertical pass ((b0*(S0>>4))>>16) + ((b1*(S1>>4))>>16) + 2) >> 2.
// The 19-px reflect apron of the reference (copyMakeBorder, :1122,1127) is never read by later
// stages, so levels are stored un-padded with a 16-byte-aligned pitch.
#include "b2orb_internal.cuh"

namespace {

constexpr int RS_WARPS = 4;     // independent warp tasks per CTA
constexpr int RS_BAND = 32;     // output rows per task

// One WARP owns a 128-px wide strip of the destination level (4 px per lane) and walks down RS_BAND output rows.
// Horizontal pass of one source row for the lane's 4 pixels: the two source bytes of pixel j are cut out of two
// aligned words with one PRMT (selector fixed per lane) and combined with one IDP.2A against the packed Q11
// weights (a0 | a1 << 16).  A source row is interpolated once and reused by the 1-2 output rows that need it
// (H0/H1 registers).  Vertical pass: two IMAD.HI with the weights pre-shifted by 16, one add, and the rounding
// shift is folded into a multiply by 64 so the result byte can be picked with PRMT.
__global__ void __launch_bounds__(RS_WARPS * 32)
resize_kernel(const __grid_constant__ B2Geom g, const B2Images im, int level, const int* __restrict__ xofs,
              const short* __restrict__ xw, const int* __restrict__ yofs, const short* __restrict__ yw) {
    const B2Level& L = g.lv[level];
    const int lane = threadIdx.x & 31;
    const int strips = (L.w + 127) >> 7;
    const int task = blockIdx.x * RS_WARPS + (threadIdx.x >> 5);
    const int band = task / strips, strip = task - band * strips;
    const int dy0 = band * RS_BAND;
    if (dy0 >= L.h) return;
    const int dy1 = min(dy0 + RS_BAND, L.h);
    const int b = blockIdx.y;
    int spitch;
    const uint8_t* src = b2_level_ptr(g, im, b, level - 1, spitch);
    const int sh = g.lv[level - 1].h;
    uint8_t* dst = im.pyr + (size_t)b * g.slab + L.off;
    const int dx0 = strip * 128 + lane * 4;

    // per-lane horizontal plan
    unsigned A[4], sel[4];
    int base01, base23;
    {
        int sx[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int dx = min(dx0 + j, L.w - 1);
            sx[j] = xofs[dx];
            A[j] = (unsigned)(unsigned short)xw[2 * dx] | ((unsigned)(unsigned short)xw[2 * dx + 1] << 16);
        }
        base01 = sx[0] & ~3;
        base23 = sx[2] & ~3;
        // downscaling by <= 2 per level: sx[1] - base01 <= 5 and sx[3] - base23 <= 5, so both bytes of a pixel
        // sit in the 8 bytes starting at its base (a larger step would need more words: rejected at create)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int o = min(sx[j] - (j < 2 ? base01 : base23), 6);
            sel[j] = (unsigned)o | ((unsigned)(o + 1) << 4);
        }
    }
    const int o0a = base01, o0b = min(base01 + 4, spitch - 4), o1a = base23, o1b = min(base23 + 4, spitch - 4);

    int cur0 = -1, cur1 = -1;
    unsigned H0[4], H1[4];
    auto hrow = [&](int s, unsigned (&H)[4]) {
        const uint8_t* r = src + (size_t)s * spitch;
        const unsigned w0 = *reinterpret_cast<const unsigned*>(r + o0a), w1 = *reinterpret_cast<const unsigned*>(r + o0b);
        const unsigned v0 = *reinterpret_cast<const unsigned*>(r + o1a), v1 = *reinterpret_cast<const unsigned*>(r + o1b);
        H[0] = __dp2a_lo(A[0], __byte_perm(w0, w1, sel[0]), 0u) >> 4;
        H[1] = __dp2a_lo(A[1], __byte_perm(w0, w1, sel[1]), 0u) >> 4;
        H[2] = __dp2a_lo(A[2], __byte_perm(v0, v1, sel[2]), 0u) >> 4;
        H[3] = __dp2a_lo(A[3], __byte_perm(v0, v1, sel[3]), 0u) >> 4;
    };
    for (int dy = dy0; dy < dy1; ++dy) {
        const int sy = yofs[dy];
        const int s1 = min(sy + 1, sh - 1);
        const unsigned b0 = (unsigned)yw[2 * dy] << 16, b1 = (unsigned)yw[2 * dy + 1] << 16;
        if (sy != cur0) {
            if (sy == cur1) {
#pragma unroll
                for (int j = 0; j < 4; ++j) H0[j] = H1[j];
            } else hrow(sy, H0);
            cur0 = sy;
        }
        if (s1 != cur1) {
            if (s1 == cur0) {
#pragma unroll
                for (int j = 0; j < 4; ++j) H1[j] = H0[j];
            } else hrow(s1, H1);
            cur1 = s1;
        }
        unsigned u[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) u[j] = (__umulhi(b0, H0[j]) + __umulhi(b1, H1[j]) + 2u) * 64u;   // byte 1 = (..)>>2
        const unsigned lo = __byte_perm(u[0], u[1], 0x0051), hi = __byte_perm(u[2], u[3], 0x0051);
        if (dx0 < L.w) *reinterpret_cast<unsigned*>(dst + (size_t)dy * L.pitch + dx0) = __byte_perm(lo, hi, 0x5410);
    }
}

// Level-0 ingest: images with arbitrary row step (e.g. tight 1241-byte rows) -> staging area whose rows start on
// 32-byte boundaries, which is what every other kernel's 32/128-bit accesses assume.  One thread per output word.
__global__ void __launch_bounds__(256)
repack_kernel(const uint8_t* __restrict__ src, int step, size_t stride, uint8_t* __restrict__ dst, int dpitch,
              size_t dstride, int w, int h) {
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

}  // namespace

void b2_launch_resize(const B2Geom& g, const B2Images& im, int level, const int* xofs, const short* xw,
                      const int* yofs, const short* yw, int nimg, cudaStream_t s) {
    const B2Level& L = g.lv[level];
    const int tasks = ((L.w + 127) / 128) * ((L.h + RS_BAND - 1) / RS_BAND);
    dim3 grid((tasks + RS_WARPS - 1) / RS_WARPS, nimg);
    resize_kernel<<<grid, RS_WARPS * 32, 0, s>>>(g, im, level, xofs, xw, yofs, yw);
}

void b2_launch_repack(const uint8_t* src, int step, size_t stride, uint8_t* dst, int dpitch, size_t dstride, int w,
                      int h, int count, cudaStream_t s) {
    dim3 grid(((dpitch >> 2) * h + 255) / 256, count);
    repack_kernel<<<grid, 256, 0, s>>>(src, step, stride, dst, dpitch, dstride, w, h);
}
