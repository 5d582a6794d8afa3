// 7x7 Gaussian blur of every pyramid level — GaussianBlur(workingMat, workingMat, Size(7,7), 2, 2,
// BORDER_REFLECT_101), src/ORBextractor.cc:1085-1086 (K6).  OpenCV's 8-bit fixed-point path: separable
// integer taps (sum 256); horizontal pass exact in Q8 (<= 65280, u16), vertical pass accumulates in Q16 and
// rounds with (acc + 32768) >> 16.  One launch covers all levels of all images: blockIdx.x walks a
// flattened (level, tile) list, blockIdx.y is the image.
#include "b2orb_internal.cuh"

namespace {

constexpr int BT_W = 64, BT_H = 16, BT_THREADS = 256;
constexpr int IN_W = BT_W + 6, IN_H = BT_H + 6, IN_P = 72;

__device__ __forceinline__ int reflect101(int i, int n) {
    // BORDER_REFLECT_101: -i -> i, (n-1)+i -> (n-1)-i; loops only for images narrower than the halo
    while (i < 0 || i >= n) i = (i < 0) ? -i : 2 * (n - 1) - i;
    return i;
}

__global__ void __launch_bounds__(BT_THREADS)
blur_kernel(const __grid_constant__ B2Geom g, const B2Images im) {
    __shared__ uint8_t s_in[IN_H][IN_P];
    __shared__ uint16_t s_h[IN_H][BT_W];
    const int b = blockIdx.y;
    int level = 0;
    while (level + 1 < g.nlevels && (int)blockIdx.x >= g.lv[level + 1].blur_tile_base) ++level;
    const B2Level& L = g.lv[level];
    const int t = blockIdx.x - L.blur_tile_base;
    const int x0 = (t % L.blur_tiles_x) * BT_W, y0 = (t / L.blur_tiles_x) * BT_H;
    int spitch;
    const uint8_t* src = b2_level_ptr(g, im, b, level, spitch);
    const int w = L.w, h = L.h;
    const int tid = threadIdx.x;

    for (int i = tid; i < IN_H * IN_W; i += BT_THREADS) {
        const int r = i / IN_W, c = i - r * IN_W;
        const int yy = (h == 1) ? 0 : reflect101(y0 + r - 3, h);
        const int xx = (w == 1) ? 0 : reflect101(x0 + c - 3, w);
        s_in[r][c] = src[(size_t)yy * spitch + xx];
    }
    __syncthreads();
    const int k0 = g.taps[0], k1 = g.taps[1], k2 = g.taps[2], k3 = g.taps[3], k4 = g.taps[4], k5 = g.taps[5],
              k6 = g.taps[6];
    for (int i = tid; i < IN_H * BT_W; i += BT_THREADS) {
        const int r = i / BT_W, c = i - r * BT_W;
        const uint8_t* p = &s_in[r][c];
        s_h[r][c] = (uint16_t)(k0 * p[0] + k1 * p[1] + k2 * p[2] + k3 * p[3] + k4 * p[4] + k5 * p[5] + k6 * p[6]);
    }
    __syncthreads();
    const int r = tid / (BT_W / 4), c0 = (tid % (BT_W / 4)) * 4;
    const int y = y0 + r, x = x0 + c0;
    if (y < h && x < w) {
        uint32_t packed = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int c = c0 + k;
            unsigned acc = k0 * s_h[r][c] + k1 * s_h[r + 1][c] + k2 * s_h[r + 2][c] + k3 * s_h[r + 3][c] +
                           k4 * s_h[r + 4][c] + k5 * s_h[r + 5][c] + k6 * s_h[r + 6][c];
            unsigned v = min((acc + 32768u) >> 16, 255u);
            packed |= v << (8 * k);
        }
        uint8_t* dst = im.blur + (size_t)b * g.slab + L.off + (size_t)y * L.pitch + x;
        *reinterpret_cast<uint32_t*>(dst) = packed;  // pitch padding absorbs the tail
    }
}

}  // namespace

void b2_launch_blur(const B2Geom& g, const B2Images& im, int nimg, cudaStream_t s) {
    dim3 grid(g.total_blur_tiles, nimg);
    blur_kernel<<<grid, BT_THREADS, 0, s>>>(g, im);
}
