// Image pyramid: level l = cv::resize(level l-1, INTER_LINEAR) — ORBextractor::ComputePyramid,
// src/ORBextractor.cc:1107-1132 (K1 in SURVEY.md §2.3).  Fixed-point arithmetic of OpenCV's 8-bit
// bilinear path: Q11 weights per axis (host-computed tables, exactly as cv::resize builds them),
// horizontal pass in int32, vertical pass ((b0*(S0>>4))>>16) + ((b1*(S1>>4))>>16) + 2) >> 2.
// The 19-px reflect apron of the reference (copyMakeBorder, :1122,1127) is never read by later
// stages, so levels are stored un-padded with a 16-byte-aligned pitch.
#include "b2orb_internal.cuh"

namespace {

constexpr int RS_TX = 32;  // threads in x, each produces 4 consecutive pixels (one 32-bit store)
constexpr int RS_TY = 8;

__global__ void __launch_bounds__(RS_TX* RS_TY)
resize_kernel(const __grid_constant__ B2Geom g, const B2Images im, int level, const int* __restrict__ xofs,
              const short* __restrict__ xw, const int* __restrict__ yofs, const short* __restrict__ yw) {
    const int b = blockIdx.z;
    const B2Level& L = g.lv[level];
    const int dx0 = (blockIdx.x * RS_TX + threadIdx.x) * 4;
    const int dy = blockIdx.y * RS_TY + threadIdx.y;
    if (dx0 >= L.w || dy >= L.h) return;
    int spitch;
    const uint8_t* src = b2_level_ptr(g, im, b, level - 1, spitch);
    const int sw = g.lv[level - 1].w, sh = g.lv[level - 1].h;
    uint8_t* dst = im.pyr + (size_t)b * g.slab + L.off + (size_t)dy * L.pitch;

    const int sy = yofs[dy];
    const int sy1 = min(sy + 1, sh - 1);
    const int b0 = yw[2 * dy], b1 = yw[2 * dy + 1];
    const uint8_t* r0 = src + (size_t)sy * spitch;
    const uint8_t* r1 = src + (size_t)sy1 * spitch;
    uint32_t packed = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int dx = dx0 + k;
        if (dx < L.w) {
            const int sx = xofs[dx];
            const int sx1 = min(sx + 1, sw - 1);
            const int a0 = xw[2 * dx], a1 = xw[2 * dx + 1];
            const int h0 = r0[sx] * a0 + r0[sx1] * a1;
            const int h1 = r1[sx] * a0 + r1[sx1] * a1;
            int v = (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
            v = min(max(v, 0), 255);
            packed |= (uint32_t)v << (8 * k);
        }
    }
    *reinterpret_cast<uint32_t*>(dst + dx0) = packed;  // pitch is a multiple of 16: padding bytes are ours
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
    dim3 block(RS_TX, RS_TY);
    dim3 grid((L.w + RS_TX * 4 - 1) / (RS_TX * 4), (L.h + RS_TY - 1) / RS_TY, nimg);
    resize_kernel<<<grid, block, 0, s>>>(g, im, level, xofs, xw, yofs, yw);
}

void b2_launch_repack(const uint8_t* src, int step, size_t stride, uint8_t* dst, int dpitch, size_t dstride, int w,
                      int h, int count, cudaStream_t s) {
    dim3 grid(((dpitch >> 2) * h + 255) / 256, count);
    repack_kernel<<<grid, 256, 0, s>>>(src, step, stride, dst, dpitch, dstride, w, h);
}
