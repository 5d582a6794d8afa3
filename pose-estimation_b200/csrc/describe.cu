// Orientation + rotated-BRIEF descriptors + keypoint assembly, one warp per selected keypoint.
//   * IC_Angle, src/ORBextractor.cc:77-104 (K5): m10, m01 over the radius-15 disc (row extents umax) of the
//     UN-blurred level, angle = cv::fastAtan2((float)m01, (float)m10) — OpenCV's fp32 7th-order polynomial,
//     evaluated without FMA (__fmul_rn/__fadd_rn; an FMA changes 2 % of results).
//   * computeOrbDescriptor, src/ORBextractor.cc:108-147 (K7): angle*factorPI, a = cosf, b = sinf (glibc
//     2.39's algorithm restated in double, verified bit-identical over every float in [0, 2*pi]),
//     offsets cvRound(x*b + y*a), cvRound(x*a - y*b) with separately rounded products (no FMA) and
//     round-half-even, 256 comparisons on the BLURRED level, LSB-first bytes.
//   * keypoint fields, :818-826, 837-847, 1095-1101 (K8): level offset +16, octave, size, pt *= scale.
// Levels are concatenated in ascending order (:1076-1104); slot offsets come from the per-level counts.
#include <algorithm>
#include <cstring>
#include <cuda.h>   // CUtensorMap (types only; the encoder is fetched through cudaGetDriverEntryPoint, no libcuda link)

#include "b2orb_internal.cuh"

namespace {

__device__ __align__(16) const signed char kPattern[1024] = {
#include "orb_pattern.inc"
};

struct KeyPointOut {   // == cv::KeyPoint == b2orb_keypoint
    float x, y, size, angle, response;
    int octave, class_id;
};

// cv::fastAtan2 (degrees), fp32, no contraction.
__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
    const float scale = (float)(180.0 / 3.14159265358979323846);
    const float p1 = 0.9997878412794807f * scale, p3 = -0.3258083974640975f * scale,
                p5 = 0.1555786518463281f * scale, p7 = -0.04432655554792128f * scale;
    const float eps = (float)2.2204460492503131e-16;
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, eps));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, eps));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

// glibc 2.39 sincosf kernel (sysdeps/ieee754/flt-32/sincosf.h): n even -> sine poly, odd -> cosine poly.
__device__ __forceinline__ float sincos_poly(double x, double x2, int n, bool neg) {
    const double C0 = 0x1p0, C1 = -0x1.ffffffd0c621cp-2, C2 = 0x1.55553e1068f19p-5, C3 = -0x1.6c087e89a359dp-10,
                 C4 = 0x1.99343027bf8c3p-16;
    const double S1 = -0x1.555545995a603p-3, S2 = 0x1.1107605230bc4p-7, S3 = -0x1.994eb3774cf24p-13;
    if ((n & 1) == 0) {
        const double x3 = __dmul_rn(x, x2);
        const double s1 = __dadd_rn(S2, __dmul_rn(x2, S3));
        const double x7 = __dmul_rn(x3, x2);
        const double s = __dadd_rn(x, __dmul_rn(x3, S1));
        return __double2float_rn(__dadd_rn(s, __dmul_rn(x7, s1)));
    }
    const double sg = neg ? -1.0 : 1.0;
    const double x4 = __dmul_rn(x2, x2);
    const double c2 = __dadd_rn(sg * C3, __dmul_rn(x2, sg * C4));
    const double c1 = __dadd_rn(sg * C0, __dmul_rn(x2, sg * C1));
    const double x6 = __dmul_rn(x4, x2);
    const double c = __dadd_rn(c1, __dmul_rn(x4, sg * C2));
    return __double2float_rn(__dadd_rn(c, __dmul_rn(x6, c2)));
}

// cosf and sinf of y in [0, 2*pi] exactly as glibc computes them.
__device__ __forceinline__ void glibc_sincosf(float y, float& sn, float& cs) {
    const double x = (double)y;
    const unsigned top = (__float_as_uint(y) >> 20) & 0x7ff;
    if (top < ((__float_as_uint(0x1.921FB6p-1f) >> 20) & 0x7ff)) {   // |y| < pi/4
        if (top < ((__float_as_uint(0x1p-12f) >> 20) & 0x7ff)) { sn = y; cs = 1.0f; return; }
        const double x2 = __dmul_rn(x, x);
        sn = sincos_poly(x, x2, 0, false);
        cs = sincos_poly(x, x2, 1, false);
        return;
    }
    const double r = __dmul_rn(x, 0x1.45F306DC9C883p+23);
    const int n = ((int)r + 0x800000) >> 24;                       // (int32_t)r truncates; r >= 0 here
    const double xr = __dsub_rn(x, __dmul_rn((double)n, 0x1.921FB54442D18p0));
    const double sgn = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
    const bool neg = (n & 2) != 0;
    const double xs = __dmul_rn(xr, sgn), x2 = __dmul_rn(xr, xr);
    sn = sincos_poly(xs, x2, n, neg);
    cs = sincos_poly(xs, x2, n ^ 1, neg);
}

constexpr int DS_WARPS = 8;
#ifndef DS_MINB
#define DS_MINB 3
#endif
constexpr int IC_WORDS = 4 * 16 * 9 * 2;   // [alignment][|v|][word][wu, wm]

__device__ __forceinline__ void ds_cp_async4(unsigned smem_addr, const void* gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_addr), "l"(gmem) : "memory");
}

__device__ __forceinline__ int dp4a_us(unsigned a, int b, int c) {   // unsigned bytes x signed bytes
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}


#ifndef DS_KPW_N
#define DS_KPW_N 16
#endif
constexpr int DS_KPW = DS_KPW_N;          // keypoints per warp in large batches (amortises the weight-table copy); 4 when latency counts
constexpr int WA_ROWS = 31, WA_WORDS = 9;    // un-blurred window: rows cy-15..cy+15, 36 bytes from (cx-15) & ~3
constexpr int WB_ROWS = 37, WB_WORDS = 10;   // blurred window:    rows cy-18..cy+18, 40 bytes from (cx-18) & ~3

// The descriptor's 512 single-byte gathers and the orientation's 31 short rows are L1-wavefront bound when done
// straight from global memory (ncu: lg/mio throttle), so each warp first copies the two windows of its keypoint into
// shared memory with coalesced 32-bit loads and gathers from there.
__global__ void __launch_bounds__(DS_WARPS * 32, DS_MINB)
describe_kernel(const __grid_constant__ B2Geom g, const B2Images im, const uint32_t* __restrict__ sel,
                const int* __restrict__ sel_cnt, const unsigned* __restrict__ ic_table, KeyPointOut* __restrict__ kps,
                uint8_t* __restrict__ desc, int* __restrict__ n_out, int kpw) {
    b2_pdl_wait();
    __shared__ unsigned s_ic[IC_WORDS];
    __shared__ unsigned s_wa[DS_WARPS][2][WA_ROWS * WA_WORDS];      // double-buffered: keypoint k+1 lands while k is worked on
    __shared__ unsigned s_wb[DS_WARPS][2][WB_ROWS * WB_WORDS + 2];
    for (int i = threadIdx.x; i < IC_WORDS; i += DS_WARPS * 32) s_ic[i] = ic_table[i];
    __syncthreads();
    const int b = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int wid = __shfl_sync(0xFFFFFFFFu, (int)(threadIdx.x >> 5), 0);   // broadcast: warp-uniform for the compiler
    // per-level counts -> prefix (same for every keypoint of this image)
    int cnt_l = 0;
    if (lane < g.nlevels) cnt_l = max(sel_cnt[b * g.nlevels + lane], 0);
    int pre = cnt_l;                                   // inclusive prefix over levels
#pragma unroll
    for (int o = 1; o < 16; o <<= 1) {
        const int v = __shfl_up_sync(0xFFFFFFFFu, pre, o);
        if (lane >= o) pre += v;
    }
    const int total = __shfl_sync(0xFFFFFFFFu, pre, 15);   // nlevels <= 12: lane 15 holds the sum
    if (blockIdx.x == 0 && threadIdx.x == 0) n_out[b] = min(total, g.kp_cap);
    const unsigned wa_s = (unsigned)__cvta_generic_to_shared(&s_wa[wid][0][lane]);   // lane's staging slot, buffer 0
    const unsigned wb_s = (unsigned)__cvta_generic_to_shared(&s_wb[wid][0][lane]);
    // Rotated-BRIEF pattern: output byte `lane` uses the 32 coordinates kPattern[32 * lane ..], the same for every
    // keypoint, so they stay in 8 registers as offset-128 bytes.  A coordinate becomes a float without the XU pipe and
    // without a shared-memory read (the kernel is bound by LSU wavefronts): PRMT drops the byte into the mantissa of
    // 1.5 * 2^23 and one FADD removes the bias, exactly.
    unsigned pk[8];
    {
        const uint4* pp = reinterpret_cast<const uint4*>(kPattern + lane * 32);
        const uint4 p0 = pp[0], p1 = pp[1];
        pk[0] = p0.x ^ 0x80808080u; pk[1] = p0.y ^ 0x80808080u; pk[2] = p0.z ^ 0x80808080u; pk[3] = p0.w ^ 0x80808080u;
        pk[4] = p1.x ^ 0x80808080u; pk[5] = p1.y ^ 0x80808080u; pk[6] = p1.z ^ 0x80808080u; pk[7] = p1.w ^ 0x80808080u;
    }
    const int ra_row = lane / 9, ra_col = lane - ra_row * 9;      // staging roles (fixed per lane)
    const int rb_row = lane / 10, rb_col = lane - rb_row * 10;

    // Software pipeline over the warp's keypoints: the windows of keypoint k+1 are copied global -> shared with 4-byte
    // cp.async into the other buffer while k's orientation + descriptor math runs (no staging registers, no STS).
    struct Kp { int slot, level, cx, cy; float response; bool valid; };
    const auto fetch = [&](int it) {
        const unsigned bufa = (it & 1) * (unsigned)sizeof(s_wa[0][0]), bufb = (it & 1) * (unsigned)sizeof(s_wb[0][0]);
        Kp k;
        k.slot = (blockIdx.x * DS_WARPS + wid) * kpw + it;      // output slot in this image
        k.valid = it < kpw && k.slot < total && k.slot < g.kp_cap;
        k.level = 0; k.cx = k.cy = 0; k.response = 0.f;
        if (!k.valid) { asm volatile("cp.async.commit_group;" ::: "memory"); return k; }
        // level = number of levels whose inclusive prefix is <= slot
        k.level = __popc(__ballot_sync(0xFFFFFFFFu, lane < g.nlevels && pre <= k.slot));
        const int base = k.level ? __shfl_sync(0xFFFFFFFFu, pre, k.level - 1) : 0;
        const B2Level& L = g.lv[k.level];
        const uint32_t v = sel[(size_t)b * g.sel_total + L.sel_base + (k.slot - base)];
        k.cx = (int)(v & 0xFFF) + B2_BORDER;
        k.cy = (int)((v >> 12) & 0xFFF) + B2_BORDER;
        k.response = (float)(v >> 24);
        int pitch;
        const uint8_t* img = b2_level_ptr(g, im, b, k.level, pitch);
        const uint8_t* bl = im.blur + (size_t)b * g.slab + L.off;
        const int bp = L.pitch;
        // lane -> (row within a group of 3, word column): 27 lanes x 11 groups cover the 31 x 9 window A,
        // 30 lanes x 13 groups the 37 x 10 window B; no per-element index arithmetic
        const int ax = (k.cx - B2_HALF_PATCH) & ~3, a = (k.cx - B2_HALF_PATCH) & 3;
        const int bx = (k.cx - 18) & ~3;
        const uint8_t* pa = img + (size_t)(k.cy - B2_HALF_PATCH + ra_row) * pitch + ax + 4 * ra_col;
        const bool oka = lane < 27 && (ra_col < 8 || a >= 2);
        // (words that are skipped keep stale but finite bytes; their IC_Angle weights are zero)
#pragma unroll
        for (int i = 0; i < 11; ++i) {
            if (oka && 3 * i + ra_row < WA_ROWS) ds_cp_async4(wa_s + bufa + i * (27 * 4), pa);
            pa += 3 * pitch;
        }
        const uint8_t* pb = bl + (size_t)(k.cy - 18 + rb_row) * bp + bx + 4 * rb_col;
        const bool okb = lane < 30;
#pragma unroll
        for (int i = 0; i < 13; ++i) {
            if (okb && 3 * i + rb_row < WB_ROWS) ds_cp_async4(wb_s + bufb + i * (30 * 4), pb);
            pb += 3 * bp;
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        return k;
    };

    Kp nxt = fetch(0);
    for (int it = 0; it < kpw; ++it) {
        const Kp cur = nxt;
        if (!cur.valid) break;
        const int slot = cur.slot, level = cur.level, cx = cur.cx, cy = cur.cy;
        const float response = cur.response;
        const B2Level& L = g.lv[level];
        const int a = (cx - B2_HALF_PATCH) & 3, ab = (cx - 18) & 3;
        asm volatile("cp.async.wait_group 0;" ::: "memory");   // this keypoint's windows have landed (this lane's words)
        __syncwarp();                                   // ... every lane's; and the previous keypoint's gathers are done
        nxt = fetch(it + 1);                            // copies in flight during the math below
        const unsigned* wa = s_wa[wid][it & 1];
        const unsigned* wb = s_wb[wid][it & 1];
        const uint8_t* wb8 = reinterpret_cast<const uint8_t*>(wb);

        // ---- IC_Angle (src/ORBextractor.cc:77-104): lane = patch row v = lane-15, IDP.4A against per-(alignment, |v|)
        // weights: wu = u inside the disc |u| <= umax[|v|] (-> m10), wm = 1 inside the disc (-> row sum -> m01)
        int m01 = 0, m10 = 0;
        if (lane < 31) {
            const int vv = lane - B2_HALF_PATCH;
            const unsigned* tw = s_ic + ((a * 16 + abs(vv)) * 9) * 2;
            const unsigned* row = wa + lane * WA_WORDS;
            int rs = 0;
#pragma unroll
            for (int k = 0; k < 9; ++k) {
                const unsigned px = row[k];
                m10 = dp4a_us(px, (int)tw[2 * k], m10);
                rs = (int)__dp4a(px, tw[2 * k + 1], (unsigned)rs);
            }
            m01 = vv * rs;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            m10 += __shfl_xor_sync(0xFFFFFFFFu, m10, o);
            m01 += __shfl_xor_sync(0xFFFFFFFFu, m01, o);
        }
        const float angle = fast_atan2_deg((float)m01, (float)m10);

        // ---- descriptor on the blurred window: lane = output byte (src/ORBextractor.cc:108-147)
        const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
        float ca, sn;
        glibc_sincosf(__fmul_rn(angle, factorPI), sn, ca);
        const uint8_t* centre = wb8 + 18 * (WB_WORDS * 4) + 18 + ab;
        unsigned val = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            int t[2];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int c = 4 * j + 2 * e;                  // compile-time after unrolling
                const float px = __fadd_rn(__uint_as_float(__byte_perm(pk[c >> 2], 0x4B400000u, 0x7650 + (c & 3))), -12583040.f);
                const float py = __fadd_rn(__uint_as_float(__byte_perm(pk[c >> 2], 0x4B400000u, 0x7651 + (c & 3))), -12583040.f);
                // cvRound == round-half-even; |value| <= 18.4, so adding 1.5 * 2^23 rounds exactly like cvtss2si
                // and leaves the integer in the low mantissa bits (FADD + IADD instead of an XU-pipe F2I)
                const float fy = __fadd_rn(__fadd_rn(__fmul_rn(px, sn), __fmul_rn(py, ca)), 12582912.f);
                const float fx = __fadd_rn(__fsub_rn(__fmul_rn(px, ca), __fmul_rn(py, sn)), 12582912.f);
                const int dy = __float_as_int(fy) - 0x4B400000, dx = __float_as_int(fx) - 0x4B400000;
                t[e] = centre[dy * (WB_WORDS * 4) + dx];
            }
            val |= (unsigned)(t[0] < t[1]) << j;
        }
        desc[((size_t)b * g.kp_cap + slot) * 32 + lane] = (uint8_t)val;
        if (lane == 0) {
            KeyPointOut k;
            k.x = (float)cx;
            k.y = (float)cy;
            if (level != 0) { k.x = __fmul_rn(k.x, L.scale); k.y = __fmul_rn(k.y, L.scale); }
            k.size = L.kp_size;
            k.angle = angle;
            k.response = response;
            k.octave = level;
            k.class_id = -1;
            kps[(size_t)b * g.kp_cap + slot] = k;
        }
    }
}


// ---------------------------------------------------------------------------------------------------------------
// describe_tma_kernel — the same computation with the two windows of a keypoint fetched by the TMA unit.
//
// In describe_kernel above, staging costs every lane 24 four-byte cp.async with a 64-bit row address each (a fifth of the
// kernel's instructions and 85 L1 tag requests per keypoint in an LSU-bound kernel).  Here one elected lane asks the copy
// engine for two boxes per keypoint — 48 x 31 bytes of the un-blurred level from column (cx - 15) & ~15, 64 x 37 bytes of
// the blurred level from column (cx - 18) & ~15 — with cp.async.bulk.tensor.3d, completion on an mbarrier, double-buffered
// per warp: no per-lane address arithmetic, no LSU traffic for the staging.  (The innermost box coordinate has to be a
// multiple of 16 bytes: an exact start at cx - 15 faults with "illegal instruction" on sm_100 — measured.  So the
// alignment residue stays, now 0..15 instead of 0..3.)  The patch rows are read with three conflict-free LDS.128 per lane
// (row stride 48 B: a quarter-warp covers all 32 banks) and the 4-byte part of the residue picks, warp-uniformly, which
// nine of the twelve words meet the IC_Angle weights; the descriptor's byte gathers take one LEA for the address (the
// float-to-int rounding bias is folded into the base).
constexpr int WTA_ROWB = 48, WTB_ROWB = 64;          // box widths in bytes (multiples of 16; 31 + 15 and 37 + 15 are needed)
constexpr int WTA_BYTES = 1536;                      // 31 rows x 48 B = 1488, padded to the 128-B alignment TMA wants
constexpr int WTB_BYTES = 2432;                      // 37 rows x 64 B = 2368
constexpr int WT_BUF = WTB_BYTES;                    // one staging buffer holds either window; two per warp
constexpr int WT_SMEM = DS_WARPS * 2 * WT_BUF + DS_WARPS * 2 * 8;

struct DescMaps {
    CUtensorMap raw0[2];                 // level 0: left block (images < split), right block
    CUtensorMap raw[B2_MAX_LEVELS];      // levels >= 1 of the pyramid slab ([0] unused)
    CUtensorMap blur[B2_MAX_LEVELS];     // blurred levels
};

__device__ __forceinline__ void dt_wait(unsigned mbar, unsigned parity) {
    asm volatile("{\n\t.reg .pred P1;\n\tDT_WAIT:\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
                 "@P1 bra DT_DONE;\n\tbra DT_WAIT;\n\tDT_DONE:\n\t}" ::"r"(mbar), "r"(parity) : "memory");
}
__device__ __forceinline__ void dt_load(unsigned dst, const CUtensorMap* map, unsigned mbar, int x, int y, int z) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(dst), "l"(map), "r"(mbar), "r"(x), "r"(y), "r"(z) : "memory");
}
__device__ __forceinline__ uint4 dt_lds128(unsigned a) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ int dt_lds8(unsigned a) {
    unsigned v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return (int)v;
}
// row words W[SW .. SW + 8] against the (wu, wm) weight pairs of the lane's row
template <int SW>
__device__ __forceinline__ void dt_ic_row(const unsigned (&W)[12], const uint2* tw, int& m10, int& rs) {
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        const uint2 t = tw[k];
        m10 = dp4a_us(W[SW + k], (int)t.x, m10);
        rs = (int)__dp4a(W[SW + k], t.y, (unsigned)rs);
    }
}

__global__ void __launch_bounds__(DS_WARPS * 32, DS_MINB)
describe_tma_kernel(const __grid_constant__ B2Geom g, const B2Images im, const DescMaps* __restrict__ maps,
                    const uint32_t* __restrict__ sel, const int* __restrict__ sel_cnt, const unsigned* __restrict__ ic_table,
                    KeyPointOut* __restrict__ kps, uint8_t* __restrict__ desc, int* __restrict__ n_out, int kpw) {
    b2_pdl_wait();
    extern __shared__ __align__(128) uint8_t s_dyn[];
    __shared__ __align__(8) unsigned s_ic[IC_WORDS];
    for (int i = threadIdx.x; i < IC_WORDS; i += DS_WARPS * 32) s_ic[i] = ic_table[i];
    const int b = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int wid = __shfl_sync(0xFFFFFFFFu, (int)(threadIdx.x >> 5), 0);   // broadcast: warp-uniform for the compiler
    const unsigned win_s = (unsigned)__cvta_generic_to_shared(s_dyn) + wid * 2 * WT_BUF;
    const unsigned bar_s = (unsigned)__cvta_generic_to_shared(s_dyn) + DS_WARPS * 2 * WT_BUF + wid * 16;
    if (lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_s));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_s + 8));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();                                    // weight table + this warp's barriers
    // per-level counts -> inclusive prefix in lanes 0..nlevels-1 (same for every keypoint of this image)
    int cnt_l = 0;
    if (lane < g.nlevels) cnt_l = max(sel_cnt[b * g.nlevels + lane], 0);
    int pre = cnt_l;
#pragma unroll
    for (int o = 1; o < 16; o <<= 1) {
        const int v = __shfl_up_sync(0xFFFFFFFFu, pre, o);
        if (lane >= o) pre += v;
    }
    const int total = __shfl_sync(0xFFFFFFFFu, pre, 15);   // nlevels <= 12: lane 15 holds the sum
    if (blockIdx.x == 0 && threadIdx.x == 0) n_out[b] = min(total, g.kp_cap);

    // ---- lane i owns keypoint i of this warp's kpw consecutive output slots: level, position and response are looked up
    // once, in parallel; the per-keypoint passes below fetch them by shuffle
    const int slot0 = (blockIdx.x * DS_WARPS + wid) * kpw;
    const int my_slot = slot0 + lane;
    const bool my_valid = lane < kpw && my_slot < total && my_slot < g.kp_cap;
    int my_level = 0, my_base = 0;
    for (int l = 0; l < g.nlevels; ++l) {              // level = number of levels whose inclusive prefix is <= slot
        const int p = __shfl_sync(0xFFFFFFFFu, pre, l);
        if (p <= my_slot) { my_level = l + 1; my_base = p; }
    }
    my_level = min(my_level, g.nlevels - 1);
    unsigned my_v = 0;
    if (my_valid) my_v = sel[(size_t)b * g.sel_total + g.lv[my_level].sel_base + (my_slot - my_base)];
    const int nv = __popc(__ballot_sync(0xFFFFFFFFu, my_valid));   // valid keypoints are a prefix of the lanes
    if (nv == 0) return;

    // box n of the warp: n < nv the un-blurred window of keypoint n, else the blurred window of keypoint n - nv; buffer n & 1
    const auto issue = [&](int n) {
        const int kp = n < nv ? n : n - nv;
        const unsigned v = __shfl_sync(0xFFFFFFFFu, my_v, kp);
        const int level = __shfl_sync(0xFFFFFFFFu, my_level, kp);
        if (lane == 0) {
            const int cx = (int)(v & 0xFFF) + B2_BORDER, cy = (int)((v >> 12) & 0xFFF) + B2_BORDER;
            const unsigned buf = win_s + (n & 1) * WT_BUF, bar = bar_s + (n & 1) * 8;
            if (n < nv) {
                const CUtensorMap* ma = level ? &maps->raw[level] : &maps->raw0[b < im.split ? 0 : 1];
                const int za = level ? b : (b < im.split ? b : b - im.split);
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(WA_ROWS * WTA_ROWB) : "memory");
                dt_load(buf, ma, bar, (cx - B2_HALF_PATCH) & ~15, cy - B2_HALF_PATCH, za);
            } else {
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(WB_ROWS * WTB_ROWB) : "memory");
                dt_load(buf, &maps->blur[level], bar, (cx - 18) & ~15, cy - 18, b);
            }
        }
    };

    // ---- pass 1: IC_Angle moments (src/ORBextractor.cc:77-104) of every keypoint; lane = patch row v = lane - 15
    const int vv = lane - B2_HALF_PATCH;
    const unsigned row_off = (unsigned)min(lane, 30) * WTA_ROWB;
    const unsigned* tw_row = s_ic + (min(abs(vv), 15) * 9) * 2;   // + residue * 16 * 9 * 2 per keypoint
    int my_m10 = 0, my_m01 = 0;
    issue(0);
    for (int n = 0; n < nv; ++n) {
        __syncwarp();                                   // every lane is done with the buffer the next box lands in
        issue(n + 1);                                   // (n + 1 == nv: the first blurred window) in flight during the math
        dt_wait(bar_s + (n & 1) * 8, (n >> 1) & 1);
        const unsigned wa_s = win_s + (n & 1) * WT_BUF;
        const int cx = (int)(__shfl_sync(0xFFFFFFFFu, my_v, n) & 0xFFF) + B2_BORDER;
        const int ra = (cx - B2_HALF_PATCH) & 15;       // the row's 31 bytes start at byte ra of the 48
        int m10 = 0, rs = 0;
        if (lane < 31) {
            const uint4 r0 = dt_lds128(wa_s + row_off), r1 = dt_lds128(wa_s + row_off + 16), r2 = dt_lds128(wa_s + row_off + 32);
            const unsigned W[12] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, r2.x, r2.y, r2.z, r2.w};
            const uint2* tw = reinterpret_cast<const uint2*>(tw_row + (ra & 3) * (16 * 9 * 2));
            switch (ra >> 2) {                          // warp-uniform
                case 0: dt_ic_row<0>(W, tw, m10, rs); break;
                case 1: dt_ic_row<1>(W, tw, m10, rs); break;
                case 2: dt_ic_row<2>(W, tw, m10, rs); break;
                default: dt_ic_row<3>(W, tw, m10, rs); break;
            }
        }
        int m01 = vv * rs;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            m10 += __shfl_xor_sync(0xFFFFFFFFu, m10, o);
            m01 += __shfl_xor_sync(0xFFFFFFFFu, m01, o);
        }
        if (lane == n) { my_m10 = m10; my_m01 = m01; }
    }

    // ---- between the passes, one lane per keypoint: angle, its sine / cosine, and the keypoint record
    // (:837-847, 1095-1101) — the scalar math runs once per keypoint instead of on all 32 lanes
    float my_sn = 0.f, my_ca = 1.f;
    if (my_valid) {
        const float angle = fast_atan2_deg((float)my_m01, (float)my_m10);
        const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
        glibc_sincosf(__fmul_rn(angle, factorPI), my_sn, my_ca);
        const B2Level& L = g.lv[my_level];
        KeyPointOut k;
        k.x = (float)((int)(my_v & 0xFFF) + B2_BORDER);
        k.y = (float)((int)((my_v >> 12) & 0xFFF) + B2_BORDER);
        if (my_level != 0) { k.x = __fmul_rn(k.x, L.scale); k.y = __fmul_rn(k.y, L.scale); }
        k.size = L.kp_size;
        k.angle = angle;
        k.response = (float)(my_v >> 24);
        k.octave = my_level;
        k.class_id = -1;
        kps[(size_t)b * g.kp_cap + my_slot] = k;
    }

    // ---- pass 2: descriptors on the blurred windows; lane = output byte (src/ORBextractor.cc:108-147)
    // Rotated-BRIEF pattern: output byte `lane` uses the 32 coordinates kPattern[32 * lane ..], the same for every keypoint,
    // kept in 8 registers as offset-128 bytes (PRMT into the mantissa of 1.5 * 2^23 + one FADD makes them floats, exactly)
    unsigned pk[8];
    {
        const uint4* pp = reinterpret_cast<const uint4*>(kPattern + lane * 32);
        const uint4 p0 = pp[0], p1 = pp[1];
        pk[0] = p0.x ^ 0x80808080u; pk[1] = p0.y ^ 0x80808080u; pk[2] = p0.z ^ 0x80808080u; pk[3] = p0.w ^ 0x80808080u;
        pk[4] = p1.x ^ 0x80808080u; pk[5] = p1.y ^ 0x80808080u; pk[6] = p1.z ^ 0x80808080u; pk[7] = p1.w ^ 0x80808080u;
    }
    for (int it = 0; it < nv; ++it) {
        const int n = nv + it;
        __syncwarp();
        if (it + 1 < nv) issue(n + 1);
        dt_wait(bar_s + (n & 1) * 8, (n >> 1) & 1);
        const unsigned wb_s = win_s + (n & 1) * WT_BUF;
        const float sn = __shfl_sync(0xFFFFFFFFu, my_sn, it), ca = __shfl_sync(0xFFFFFFFFu, my_ca, it);
        const int cx = (int)(__shfl_sync(0xFFFFFFFFu, my_v, it) & 0xFFF) + B2_BORDER;
        const int rb = (cx - 18) & 15;
        // address of window byte (18 + dy, rb + 18 + dx) with dy, dx still carrying the 1.5 * 2^23 rounding bias (0x4B400000)
        const unsigned centre = wb_s + 18 * WTB_ROWB + 18 + rb - 0x4B400000u * (unsigned)WTB_ROWB - 0x4B400000u;
        unsigned val = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            int t[2];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int c = 4 * j + 2 * e;
                const float px = __fadd_rn(__uint_as_float(__byte_perm(pk[c >> 2], 0x4B400000u, 0x7650 + (c & 3))), -12583040.f);
                const float py = __fadd_rn(__uint_as_float(__byte_perm(pk[c >> 2], 0x4B400000u, 0x7651 + (c & 3))), -12583040.f);
                // cvRound == round-half-even; |value| <= 18.4, so adding 1.5 * 2^23 rounds exactly like cvtss2si and leaves
                // the integer in the low mantissa bits
                const float fy = __fadd_rn(__fadd_rn(__fmul_rn(px, sn), __fmul_rn(py, ca)), 12582912.f);
                const float fx = __fadd_rn(__fsub_rn(__fmul_rn(px, ca), __fmul_rn(py, sn)), 12582912.f);
                t[e] = dt_lds8(centre + __float_as_uint(fy) * (unsigned)WTB_ROWB + __float_as_uint(fx));
            }
            val |= (unsigned)(t[0] < t[1]) << j;
        }
        desc[((size_t)b * g.kp_cap + slot0 + it) * 32 + lane] = (uint8_t)val;
    }
}

}  // namespace

// Tensor maps of one handle's slabs, re-encoded only when what they describe changes (the slabs are allocated once per
// handle; the level-0 maps follow the caller's buffer when it is read in place).
struct B2DescCache {
    DescMaps maps;             // host copy; the kernel reads the device copy (with 26 maps the kernel parameters would pass 4 KB)
    DescMaps* d_maps = nullptr;
    struct Key {
        const void* lvl0[2]; const void* pyr; const void* blur;
        size_t stride0; int pitch0, split, nimg, nlevels; unsigned slab;
        int geo[B2_MAX_LEVELS][4];
    } key;
    bool valid = false;        // maps match key
    bool unavailable = false;  // no encoder in this driver / an encode failed for this geometry: use the cp.async kernel
    bool attr_set = false;
};

B2DescCache* b2_desc_cache_create() {
    B2DescCache* c = new B2DescCache();
    memset(&c->maps, 0, sizeof c->maps);
    if (cudaMalloc(&c->d_maps, sizeof(DescMaps)) != cudaSuccess) { c->d_maps = nullptr; c->unavailable = true; (void)cudaGetLastError(); }
    return c;
}
void b2_desc_cache_destroy(B2DescCache* c) {
    if (!c) return;
    if (c->d_maps) cudaFree(c->d_maps);
    delete c;
}

namespace {

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

bool desc_maps_update(B2DescCache* c, const B2Geom& g, const B2Images& im, int nimg, cudaStream_t s) {
    if (c->unavailable) return false;
    B2DescCache::Key k;
    memset(&k, 0, sizeof k);
    k.lvl0[0] = im.lvl0[0]; k.lvl0[1] = im.lvl0[1]; k.pyr = im.pyr; k.blur = im.blur;
    k.stride0 = im.stride0; k.pitch0 = im.pitch0; k.split = im.split; k.nimg = nimg; k.nlevels = g.nlevels; k.slab = g.slab;
    for (int l = 0; l < g.nlevels; ++l) { k.geo[l][0] = g.lv[l].w; k.geo[l][1] = g.lv[l].h; k.geo[l][2] = g.lv[l].pitch; k.geo[l][3] = (int)g.lv[l].off; }
    if (c->valid && memcmp(&k, &c->key, sizeof k) == 0) return true;
    static EncodeFn encode = nullptr;
    if (!encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr) != cudaSuccess || !fn) {
            c->unavailable = true;
            return false;
        }
        encode = (EncodeFn)fn;
    }
    const cuuint32_t es[3] = {1, 1, 1};
    auto mk = [&](CUtensorMap* m, const void* base, int pitch, int h, size_t img_stride, int n, int box_w, int box_rows) {
        if (!base || ((uintptr_t)base & 15) || (pitch & 15) || (img_stride & 15)) return false;
        const cuuint32_t box[3] = {(cuuint32_t)box_w, (cuuint32_t)box_rows, 1};
        const cuuint64_t dims[3] = {(cuuint64_t)pitch, (cuuint64_t)h, (cuuint64_t)std::max(n, 1)};
        const cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)std::max(img_stride, (size_t)16)};
        return encode(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(base), dims, strides, box, es,
                      CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
    };
    c->valid = false;
    const int n_left = std::min(im.split, nimg), n_right = nimg - n_left;
    bool ok = true;
    if (n_left > 0) ok = mk(&c->maps.raw0[0], im.lvl0[0], im.pitch0, g.lv[0].h, im.stride0, n_left, WTA_ROWB, WA_ROWS);
    if (ok && n_right > 0) ok = mk(&c->maps.raw0[1], im.lvl0[1], im.pitch0, g.lv[0].h, im.stride0, n_right, WTA_ROWB, WA_ROWS);
    for (int l = 1; l < g.nlevels && ok; ++l)
        ok = mk(&c->maps.raw[l], im.pyr + g.lv[l].off, g.lv[l].pitch, g.lv[l].h, g.slab, nimg, WTA_ROWB, WA_ROWS);
    for (int l = 0; l < g.nlevels && ok; ++l)
        ok = mk(&c->maps.blur[l], im.blur + g.lv[l].off, g.lv[l].pitch, g.lv[l].h, g.slab, nimg, WTB_ROWB, WB_ROWS);
    if (!ok) return false;      // e.g. a caller buffer that is not 16-byte aligned: this launch takes the cp.async kernel
    // stream-ordered: launches already queued keep reading the old maps until they are done
    if (cudaMemcpyAsync(c->d_maps, &c->maps, sizeof(DescMaps), cudaMemcpyHostToDevice, s) != cudaSuccess) return false;
    c->key = k;
    c->valid = true;
    return true;
}

}  // namespace

void b2_launch_describe(const B2Geom& g, const B2Images& im, const uint32_t* sel, const int* sel_cnt,
                        const unsigned* ic_table, void* kps, uint8_t* desc, int* n_out, int nimg, B2DescCache* cache,
                        int use_tma, cudaStream_t s) {
    // few images: short warps on many SMs (latency); large batches: long warps (the per-CTA table copy amortises)
    const int kpw = (long long)g.kp_cap * nimg <= 148LL * DS_WARPS * 16 ? 4 : DS_KPW;
    const int per_cta = DS_WARPS * kpw;
    dim3 grid((g.kp_cap + per_cta - 1) / per_cta, nimg);
    if (use_tma && cache && desc_maps_update(cache, g, im, nimg, s)) {
        if (!cache->attr_set) {
            if (cudaFuncSetAttribute(describe_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WT_SMEM) != cudaSuccess) {
                cache->unavailable = true;
                (void)cudaGetLastError();
            }
            cache->attr_set = true;
        }
        if (!cache->unavailable) {
            b2_launch_pdl(describe_tma_kernel, dim3(grid), dim3(DS_WARPS * 32), (size_t)WT_SMEM, s, g, im, static_cast<const DescMaps*>(cache->d_maps), sel, sel_cnt,
                          ic_table, static_cast<KeyPointOut*>(kps), desc, n_out, kpw);
            return;
        }
    }
    b2_launch_pdl(describe_kernel, dim3(grid), dim3(DS_WARPS * 32), 0, s, g, im, sel, sel_cnt, ic_table, static_cast<KeyPointOut*>(kps), desc, n_out, kpw);
}
