// Probe for the tcgen05.mma.kind::i8 operand layout / descriptor fields used by match.cu: one 128x128x256 tile with
// small signed test matrices, accumulator dumped and compared with the CPU product.
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>

constexpr int R = 128, K = 256, TILE = R * K;
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46);
}
__global__ void __launch_bounds__(128) probe(const int8_t* A, const int8_t* B, int* D, uint32_t idesc, int lbo, int sbo, int kstep_bytes) {
    extern __shared__ __align__(1024) uint8_t sm[];
    __shared__ __align__(8) uint64_t mbar;
    __shared__ uint32_t s_tmem;
    const int tid = threadIdx.x, wid = tid >> 5;
    uint8_t* sA = sm; uint8_t* sB = sm + TILE;
    // logical (row r, k) -> (k/16)*2048 + r*16 + k%16
    for (int i = tid; i < TILE; i += 128) {
        const int r = i / K, k = i % K;
        sA[(k / 16) * 2048 + r * 16 + k % 16] = (uint8_t)A[i];
        sB[(k / 16) * 2048 + r * 16 + k % 16] = (uint8_t)B[i];
    }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&mbar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (wid == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(128) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = s_tmem;
    if (tid == 0) {
        for (int s = 0; s < 8; ++s) {
            const uint64_t da = desc(smem_u32(sA) + s * kstep_bytes, lbo, sbo), db = desc(smem_u32(sB) + s * kstep_bytes, lbo, sbo);
            const uint32_t acc = s > 0, zero = 0;
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                         "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}"
                         ::"r"(tmem), "l"(da), "l"(db), "r"(idesc), "r"(acc), "r"(zero) : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&mbar)) : "memory");
    }
    asm volatile("{\n\t.reg .pred P1;\n\tLAB_WAIT:\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
                 "@P1 bra DONE;\n\tbra LAB_WAIT;\n\tDONE:\n\t}" ::"r"(smem_u32(&mbar)), "r"(0) : "memory");
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    for (int c = 0; c < 4; ++c) {
        int v[32];
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                     "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                     "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                       "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                       "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                       "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                     : "r"(tmem + c * 32 + ((uint32_t)(wid * 32) << 16)));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int k = 0; k < 32; ++k) D[tid * 128 + c * 32 + k] = v[k];
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (wid == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(128) : "memory");
}

int main(int argc, char** argv) {
    static int8_t hA[TILE], hB[TILE];
    static int hD[R * R], ref[R * R];
    int8_t *dA, *dB; int* dD;
    cudaMalloc(&dA, TILE); cudaMalloc(&dB, TILE); cudaMalloc(&dD, R * R * 4);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * TILE);
    struct Cfg { const char* name; uint32_t afmt, bfmt; int lbo, sbo, kstep; };
    const Cfg cfgs[] = {{"fmt=1 lbo=2048 sbo=128", 1, 1, 2048, 128, 4096}, {"fmt=0 lbo=2048 sbo=128", 0, 0, 2048, 128, 4096},
                        {"fmt=1 lbo=128 sbo=2048", 1, 1, 128, 2048, 4096}};
    for (int mode = 0; mode < 2; ++mode) {
        for (int i = 0; i < TILE; ++i) {
            const int r = i / K, k = i % K;
            hA[i] = mode == 0 ? (int8_t)(((r * 7 + k * 3) % 5) - 2) : (int8_t)((r * 31 + k * 17) % 3 == 0 ? -1 : 1);
            hB[i] = mode == 0 ? (int8_t)(((r * 5 + k) % 7) - 3) : (int8_t)((r * 13 + k * 7) % 5 < 2 ? -1 : 1);
        }
        for (int r = 0; r < R; ++r)
            for (int n = 0; n < R; ++n) {
                int s = 0;
                for (int k = 0; k < K; ++k) s += (int)hA[r * K + k] * (int)hB[n * K + k];
                ref[r * R + n] = s;
            }
        cudaMemcpy(dA, hA, TILE, cudaMemcpyHostToDevice); cudaMemcpy(dB, hB, TILE, cudaMemcpyHostToDevice);
        for (const Cfg& c : cfgs) {
            const uint32_t idesc = (2u << 4) | (c.afmt << 7) | (c.bfmt << 10) | ((128u >> 3) << 17) | ((128u >> 4) << 24);
            cudaMemset(dD, 0xEE, R * R * 4);
            probe<<<1, 128, 2 * TILE>>>(dA, dB, dD, idesc, c.lbo, c.sbo, c.kstep);
            cudaError_t e = cudaDeviceSynchronize();
            cudaMemcpy(hD, dD, R * R * 4, cudaMemcpyDeviceToHost);
            int bad = 0;
            for (int i = 0; i < R * R; ++i) bad += hD[i] != ref[i];
            printf("mode %d  %-26s err=%s mismatches=%d  D[0][0..3]=%d %d %d %d ref=%d %d %d %d  D[1][0]=%d ref=%d D[9][17]=%d ref=%d\n", mode, c.name,
                   cudaGetErrorString(e), bad, hD[0], hD[1], hD[2], hD[3], ref[0], ref[1], ref[2], ref[3], hD[128], ref[128], hD[9 * 128 + 17], ref[9 * 128 + 17]);
        }
    }
    return 0;
}
