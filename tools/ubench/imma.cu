// Micro-benchmark: throughput of the legacy warp-level integer MMA (mma.sync m16n8k32 u8 -> IMMA.16832) on sm_100a,
// alone and with ALU work interleaved.  Decides whether the Hamming matcher's 256-bit dot products are cheaper on
// this path than on POPC (XU pipe, 0.5 warp-op/clk/SM).
#include <cstdio>
#include <cuda_runtime.h>

#define ITERS 4096
__device__ __forceinline__ void imma(int (&d)[4], const unsigned (&a)[4], const unsigned (&b)[2]) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
template <int NACC, int ALU>
__global__ void __launch_bounds__(1024) k(int* out, unsigned seed) {
    unsigned a[4] = {seed + threadIdx.x, seed * 3 + 1, seed * 5 + 2, seed * 7};
    unsigned b[2] = {seed * 11, seed * 13 + threadIdx.x};
    int d[NACC][4];
#pragma unroll
    for (int i = 0; i < NACC; ++i) d[i][0] = d[i][1] = d[i][2] = d[i][3] = i;
    unsigned x0 = seed, x1 = seed + 1, x2 = seed + 2, x3 = seed + 3;
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < NACC; ++i) {
            imma(d[i], a, b);
            if (ALU) {   // ALU ops per IMMA, independent chains
#pragma unroll
                for (int u = 0; u < ALU; u += 4) { x0 = min(x0 ^ x1, x2); x1 = max(x1 ^ x2, x3); x2 = min(x2 ^ x3, x0); x3 = max(x3 ^ x0, x1); }
            }
        }
    }
    int s = x0 + x1 + x2 + x3;
#pragma unroll
    for (int i = 0; i < NACC; ++i) s += d[i][0] + d[i][1] + d[i][2] + d[i][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int NACC, int ALU>
void run(const char* name, int* d_out, int threads) {
    int dev; cudaGetDevice(&dev); cudaDeviceProp p; cudaGetDeviceProperties(&p, dev);
    int clk_khz; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, dev);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<NACC, ALU><<<p.multiProcessorCount, threads>>>(d_out, 1);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<NACC, ALU><<<p.multiProcessorCount, threads>>>(d_out, 2);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double immas = (double)ITERS * NACC * (threads / 32);            // warp-level MMAs per SM
    double clk = ms * 1e-3 * clk_khz * 1e3;                           // at the nominal max clock
    printf("%-28s threads=%4d  %.3f ms  IMMA/clk/SM=%.3f (nominal clk)  => %.0f TOPS chip\n", name, threads, ms, immas / clk,
           immas * 16 * 8 * 32 * 2 * p.multiProcessorCount / (ms * 1e-3) / 1e12);
}
int main() {
    int* d_out; cudaMalloc(&d_out, 1024 * 256 * 4);
    run<4, 0>("imma x4acc", d_out, 1024);
    run<8, 0>("imma x8acc", d_out, 1024);
    run<8, 0>("imma x8acc", d_out, 512);
    run<8, 0>("imma x8acc", d_out, 256);
    run<8, 4>("imma x8acc + 4 alu", d_out, 512);
    run<8, 8>("imma x8acc + 8 alu", d_out, 512);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
