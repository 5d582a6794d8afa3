// Micro-benchmark: which issue pipes serve the packed min/max, dp4a and popc instructions on sm_100a, and
// whether mixing integer (ALU pipe) and half2 (FMA pipe?) min/max raises throughput.  One CTA of 1024 threads per SM.
#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#define ITERS 2048
__device__ __forceinline__ unsigned HMAX(unsigned a, unsigned b){ __half2 r=__hmax2(*(__half2*)&a,*(__half2*)&b); return *(unsigned*)&r; }
__device__ __forceinline__ unsigned HMIN(unsigned a, unsigned b){ __half2 r=__hmin2(*(__half2*)&a,*(__half2*)&b); return *(unsigned*)&r; }
__device__ __forceinline__ unsigned HFMA(unsigned a, unsigned b, unsigned c){ __half2 r=__hfma2(*(__half2*)&a,*(__half2*)&b,*(__half2*)&c); return *(unsigned*)&r; }
__device__ __forceinline__ unsigned FMAXU(unsigned a, unsigned b){ return __float_as_uint(fmaxf(__uint_as_float(a),__uint_as_float(b))); }
__device__ __forceinline__ unsigned FMINU(unsigned a, unsigned b){ return __float_as_uint(fminf(__uint_as_float(a),__uint_as_float(b))); }
template <int MODE>
__global__ void __launch_bounds__(1024) k(unsigned* out, unsigned seed) {
    unsigned a0 = seed + threadIdx.x, a1 = a0 * 3 + 1, a2 = a0 * 5 + 2, a3 = a0 * 7 + 3;
    unsigned b0 = a0 | 0x64006400u, b1 = a1 | 0x64006400u, b2 = a2 | 0x64006400u, b3 = a3 | 0x64006400u;
    const unsigned c = seed * 77 + 5, d = seed * 13 + 9;
#pragma unroll 1
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            unsigned t;
            if (MODE == 0) {   // VIMNMX.U16x2 only (8 ops)
                a0 = __vmaxu2(a0, a1); a1 = __vminu2(a1, a2); a2 = __vmaxu2(a2, a3); a3 = __vminu2(a3, b0);
                b0 = __vmaxu2(b0, b1); b1 = __vminu2(b1, b2); b2 = __vmaxu2(b2, b3); b3 = __vminu2(b3, a0);
            } else if (MODE == 1) {   // half2 min/max only (8 ops)
                a0 = HMAX(a0, a1); a1 = HMIN(a1, a2); a2 = HMAX(a2, a3); a3 = HMIN(a3, b0);
                b0 = HMAX(b0, b1); b1 = HMIN(b1, b2); b2 = HMAX(b2, b3); b3 = HMIN(b3, a0);
            } else if (MODE == 2) {   // 4 VIMNMX + 4 half2
                a0 = __vmaxu2(a0, a1); a1 = __vminu2(a1, a2); a2 = __vmaxu2(a2, a3); a3 = __vminu2(a3, a0);
                b0 = HMAX(b0, b1); b1 = HMIN(b1, b2); b2 = HMAX(b2, b3); b3 = HMIN(b3, b0);
            } else if (MODE == 3) {   // VIMNMX3.U16x2 only (8 ops)
                a0 = __vimax3_u16x2(a0, a1, a2); a1 = __vimin3_u16x2(a1, a2, a3); a2 = __vimax3_u16x2(a2, a3, b0); a3 = __vimin3_u16x2(a3, b0, b1);
                b0 = __vimax3_u16x2(b0, b1, b2); b1 = __vimin3_u16x2(b1, b2, b3); b2 = __vimax3_u16x2(b2, b3, a0); b3 = __vimin3_u16x2(b3, a0, a1);
            } else if (MODE == 4) {   // IMAD only (8 ops)
                a0 = a0 * c + a1; a1 = a1 * c + a2; a2 = a2 * c + a3; a3 = a3 * c + b0;
                b0 = b0 * c + b1; b1 = b1 * c + b2; b2 = b2 * c + b3; b3 = b3 * c + a0;
            } else if (MODE == 5) {   // 4 IMAD + 4 VIMNMX
                a0 = a0 * c + a1; a1 = a1 * c + a2; a2 = a2 * c + a3; a3 = a3 * c + a0;
                b0 = __vmaxu2(b0, b1); b1 = __vminu2(b1, b2); b2 = __vmaxu2(b2, b3); b3 = __vminu2(b3, b0);
            } else if (MODE == 6) {   // IDP4A only (8)
                a0 = __dp4a(a0, c, a1); a1 = __dp4a(a1, c, a2); a2 = __dp4a(a2, c, a3); a3 = __dp4a(a3, c, b0);
                b0 = __dp4a(b0, c, b1); b1 = __dp4a(b1, c, b2); b2 = __dp4a(b2, c, b3); b3 = __dp4a(b3, c, a0);
            } else if (MODE == 7) {   // POPC only (8)
                a0 ^= __popc(a1); a1 ^= __popc(a2); a2 ^= __popc(a3); a3 ^= __popc(b0);
                b0 ^= __popc(b1); b1 ^= __popc(b2); b2 ^= __popc(b3); b3 ^= __popc(a0);
            } else if (MODE == 8) {   // LOP3 only (8)
                a0 = (a0 ^ c) & a1; a1 = (a1 ^ c) | a2; a2 = (a2 & c) ^ a3; a3 = (a3 | c) ^ b0;
                b0 = (b0 ^ c) & b1; b1 = (b1 ^ c) | b2; b2 = (b2 & c) ^ b3; b3 = (b3 | c) ^ a0;
            } else if (MODE == 9) {   // 4 IDP4A + 4 LOP3
                a0 = __dp4a(a0, c, a1); a1 = __dp4a(a1, c, a2); a2 = __dp4a(a2, c, a3); a3 = __dp4a(a3, c, a0);
                b0 = (b0 ^ c) & b1; b1 = (b1 ^ c) | b2; b2 = (b2 & c) ^ b3; b3 = (b3 | c) ^ b0;
            } else if (MODE == 10) {  // HFMA2 only
                a0 = HFMA(a0, c, a1); a1 = HFMA(a1, c, a2); a2 = HFMA(a2, c, a3); a3 = HFMA(a3, c, b0);
                b0 = HFMA(b0, c, b1); b1 = HFMA(b1, c, b2); b2 = HFMA(b2, c, b3); b3 = HFMA(b3, c, a0);
            } else if (MODE == 11) {  // 4 VIMNMX3 + 4 half2 minmax
                a0 = __vimax3_u16x2(a0, a1, a2); a1 = __vimin3_u16x2(a1, a2, a3); a2 = __vimax3_u16x2(a2, a3, a0); a3 = __vimin3_u16x2(a3, a0, a1);
                b0 = HMAX(b0, b1); b1 = HMIN(b1, b2); b2 = HMAX(b2, b3); b3 = HMIN(b3, b0);
            } else if (MODE == 12) {  // 4 POPC + 4 LOP3
                a0 ^= __popc(a1); a1 ^= __popc(a2); a2 ^= __popc(a3); a3 ^= __popc(a0);
                b0 = (b0 ^ c) & b1; b1 = (b1 ^ c) | b2; b2 = (b2 & c) ^ b3; b3 = (b3 | c) ^ b0;
            } else if (MODE == 13) {  // 4 VIMNMX + 4 HFMA2
                a0 = __vmaxu2(a0, a1); a1 = __vminu2(a1, a2); a2 = __vmaxu2(a2, a3); a3 = __vminu2(a3, a0);
                b0 = HFMA(b0, c, b1); b1 = HFMA(b1, c, b2); b2 = HFMA(b2, c, b3); b3 = HFMA(b3, c, b0);
            } else if (MODE == 14) {  // 4 VIMNMX + 4 FMNMX(f32)
                a0 = __vmaxu2(a0, a1); a1 = __vminu2(a1, a2); a2 = __vmaxu2(a2, a3); a3 = __vminu2(a3, a0);
                b0 = FMAXU(b0, b1); b1 = FMINU(b1, b2); b2 = FMAXU(b2, b3); b3 = FMINU(b3, b0);
            } else if (MODE == 15) {  // PRMT only
                a0 = __byte_perm(a0, a1, 0x5410); a1 = __byte_perm(a1, a2, 0x7632); a2 = __byte_perm(a2, a3, 0x6420); a3 = __byte_perm(a3, b0, 0x7531);
                b0 = __byte_perm(b0, b1, 0x5410); b1 = __byte_perm(b1, b2, 0x7632); b2 = __byte_perm(b2, b3, 0x6420); b3 = __byte_perm(b3, a0, 0x7531);
            } else if (MODE == 16) {  // 4 VIMNMX3 + 4 IMAD
                a0 = __vimax3_u16x2(a0, a1, a2); a1 = __vimin3_u16x2(a1, a2, a3); a2 = __vimax3_u16x2(a2, a3, a0); a3 = __vimin3_u16x2(a3, a0, a1);
                b0 = b0 * c + b1; b1 = b1 * c + b2; b2 = b2 * c + b3; b3 = b3 * c + b0;
            }
            (void)t;
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + b0 + b1 + b2 + b3;
}

template <int MODE>
void run(const char* name, unsigned* d_out) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int grid = 148 * 2;
    k<MODE><<<grid, 1024>>>(d_out, 1u);
    cudaEventRecord(e0);
    k<MODE><<<grid, 1024>>>(d_out, 3u);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double warp_instr = (double)grid * 32 * ITERS * 32.0;   // 32 "ops" per iteration (8 x unroll 4) per warp
    // per SM: 2 CTAs x 32 warps; cycles at 1.965 GHz
    const double cyc = ms * 1e-3 * 1.965e9;
    printf("%-28s %8.3f ms  %6.2f warp-ops/clk/SM (if 1965 MHz)\n", name, ms, warp_instr / 148.0 / cyc);
}

int main() {
    unsigned* d_out;
    cudaMalloc(&d_out, 148 * 2 * 1024 * 4);
    run<0>("VIMNMX.U16x2", d_out);
    run<1>("HMNMX2", d_out);
    run<2>("VIMNMX + HMNMX2 (1:1)", d_out);
    run<3>("VIMNMX3.U16x2", d_out);
    run<11>("VIMNMX3 + HMNMX2 (1:1)", d_out);
    run<4>("IMAD", d_out);
    run<5>("IMAD + VIMNMX (1:1)", d_out);
    run<6>("IDP4A", d_out);
    run<9>("IDP4A + LOP3 (1:1)", d_out);
    run<7>("POPC(+IADD)", d_out);
    run<12>("POPC(+IADD) + LOP3 (1:1)", d_out);
    run<8>("LOP3", d_out);
    run<10>("HFMA2", d_out);
    run<13>("VIMNMX + HFMA2 (1:1)", d_out);
    run<14>("VIMNMX + FMNMX (1:1)", d_out);
    run<15>("PRMT", d_out);
    run<16>("VIMNMX3 + IMAD (1:1)", d_out);
    cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
