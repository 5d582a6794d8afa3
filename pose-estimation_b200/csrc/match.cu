// Brute-force Hamming matching and the FeatureTracker filters.
//   * match_tc_kernel — cv::BFMatcher(cv::NORM_HAMMING).match(query, train), src/feature-tracker.cpp:161-163,
//     :258, :303 (K9): for every query row the nearest train row by popcount(q xor t) over 32-byte
//     descriptors; the lowest train index wins ties exactly like OpenCV's batch-distance loop (strict '<' in
//     ascending train order).  The distance is a dense contraction and runs on tcgen05 (see below).  Two earlier
//     versions were measured and removed (git history): POPC + carry-save tree on the XU pipe 5.0 us per
//     KITTI pair, warp-level mma.sync.m16n8k32.u8 2.5 us; this one 1.3 us.
//   * stereo_filter_kernel — internalMatch, :164 and :208-218 (K10): global minDist, keep
//     dist < max(matchRatio*minDist, minMatchDist) && |yL - yR| < maxVerticalPixelDist (fp32), ordered
//     compaction in query order, gather of the kept descriptor rows (_curDescriptorsL/R).
//   * circular_kernel — externalTrack, :258-259, 293-296, 303-321 (useRANSAC=false): distance filter on both
//     sides and the circular test mapCurHist[queryIdx] == trainIdx.
#include "b2orb_internal.cuh"

namespace {

constexpr int SF_THREADS = 1024;   // one CTA per pair: the passes over ~2000 matches are latency chains, so use the widest CTA

__device__ __forceinline__ int block_min(int v, int* s_red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(0xFFFFFFFFu, v, o));
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = v;
    __syncthreads();
    int r = s_red[0];
    for (int w = 1; w < (int)blockDim.x / 32; ++w) r = min(r, s_red[w]);
    __syncthreads();
    return r;
}

// exclusive scan of one 0/1 flag per thread over the block; returns the offset, total via *tot
__device__ __forceinline__ int block_scan_flag(int f, int* s_red, int* tot) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const unsigned bal = __ballot_sync(0xFFFFFFFFu, f);
    const int within = __popc(bal & ((1u << lane) - 1));
    if (lane == 0) s_red[wid] = __popc(bal);
    __syncthreads();
    int off = 0, t = 0;
    for (int w = 0; w < (int)blockDim.x / 32; ++w) {
        const int c = s_red[w];
        if (w < wid) off += c;
        t += c;
    }
    __syncthreads();
    *tot = t;
    return off + within;
}

struct KeyPointIn {
    float x, y, size, angle, response;
    int octave, class_id;
};

// One CTA per stereo pair.  Image slots: left = p, right = npairs + p.
constexpr int SF_B = 4;   // rows per thread whose loads are in flight together
__global__ void __launch_bounds__(SF_THREADS)
stereo_filter_kernel(const KeyPointIn* __restrict__ kps, const uint8_t* __restrict__ desc, const int* __restrict__ n_kp,
                     const int* __restrict__ train_idx, const int* __restrict__ dist, int cap, float ratio,
                     float min_match, float max_dy, int* __restrict__ match_l, int* __restrict__ match_r,
                     int* __restrict__ n_match, uint8_t* __restrict__ cur_l, uint8_t* __restrict__ cur_r,
                     double2* __restrict__ pix_l, double2* __restrict__ pix_r) {
    b2_pdl_wait();
    __shared__ int s_red[SF_THREADS / 32];
    const int p = blockIdx.x;
    const int np = gridDim.x;
    const int nl = n_kp[p], nr = n_kp[np + p];
    const KeyPointIn* kl = kps + (size_t)p * cap;
    const KeyPointIn* kr = kps + (size_t)(np + p) * cap;
    const int* ti = train_idx + (size_t)p * cap;
    const int* di = dist + (size_t)p * cap;
    if (nl == 0 || nr == 0) {   // reference: min_element on an empty vector (undefined) -> defined as no matches
        if (threadIdx.x == 0) n_match[p] = 0;
        return;
    }
    // A latency chain, not a throughput problem (one CTA, ~2000 rows): every thread first issues ALL its independent loads
    // (distance, train index, left row coordinate of up to SF_B rows per batch), then the dependent right-row loads, and
    // only then meets the block-wide min — three memory round trips in total instead of one per row and step.
    int mn = 0x7fffffff;
    unsigned long long dyok = 0;                     // bit k: row k * SF_THREADS + tid passes the row test (:208-209)
    const bool bits_ok = nl <= 64 * SF_THREADS;      // (larger sets: the test is redone in pass 2)
    for (int k0 = 0; k0 * SF_THREADS < nl; k0 += SF_B) {
        int d[SF_B], t[SF_B];
        float yl[SF_B];
#pragma unroll
        for (int k = 0; k < SF_B; ++k) {
            const int i = (k0 + k) * SF_THREADS + threadIdx.x;
            const bool in = i < nl;
            d[k] = in ? di[i] : 0x7fffffff;
            t[k] = in ? ti[i] : 0;
            yl[k] = in ? kl[i].y : 0.f;
        }
#pragma unroll
        for (int k = 0; k < SF_B; ++k) {
            const int i = (k0 + k) * SF_THREADS + threadIdx.x;
            mn = min(mn, d[k]);
            if (i < nl && bits_ok && fabsf(__fsub_rn(yl[k], kr[t[k]].y)) < max_dy) dyok |= 1ull << (k0 + k);
        }
    }
    mn = block_min(mn, s_red);
    const float thr = fmaxf(__fmul_rn(ratio, (float)mn), min_match);
    int base = 0;
    for (int i0 = 0, k = 0; i0 < nl; i0 += SF_THREADS, ++k) {
        const int i = i0 + threadIdx.x;
        int keep = 0, t = 0;
        if (i < nl) {
            t = ti[i];
            const bool row_ok = bits_ok ? ((dyok >> k) & 1ull) != 0 : fabsf(__fsub_rn(kl[i].y, kr[t].y)) < max_dy;
            keep = ((float)di[i] < thr) && row_ok;
        }
        int tot;
        const int off = block_scan_flag(keep, s_red, &tot);
        if (keep) {
            const int m = base + off;
            match_l[(size_t)p * cap + m] = i;
            match_r[(size_t)p * cap + m] = t;
            const uint4* sl = reinterpret_cast<const uint4*>(desc + ((size_t)p * cap + i) * 32);
            const uint4* sr = reinterpret_cast<const uint4*>(desc + ((size_t)(np + p) * cap + t) * 32);
            uint4* dl = reinterpret_cast<uint4*>(cur_l + ((size_t)p * cap + m) * 32);
            uint4* dr = reinterpret_cast<uint4*>(cur_r + ((size_t)p * cap + m) * 32);
            const uint4 l0 = sl[0], l1 = sl[1], r0 = sr[0], r1 = sr[1];
            dl[0] = l0; dl[1] = l1;
            dr[0] = r0; dr[1] = r1;
            if (pix_l) {   // _curPixelsL/R.push_back(keypoint.pt): Point2f widened to Point2d (:214-215)
                pix_l[(size_t)p * cap + m] = make_double2((double)kl[i].x, (double)kl[i].y);
                pix_r[(size_t)p * cap + m] = make_double2((double)kr[t].x, (double)kr[t].y);
            }
        }
        base += tot;
    }
    if (threadIdx.x == 0) n_match[p] = base;
}

// One CTA per temporal problem: left and right match lists of the same cur set against the same hist set.
__global__ void __launch_bounds__(SF_THREADS)
circular_kernel(const int* __restrict__ ti_l, const int* __restrict__ di_l, const int* __restrict__ ti_r,
                const int* __restrict__ di_r, const int* __restrict__ n_cur, const int* __restrict__ n_hist,
                int hist_stride, int cap, float ratio, float min_match, int* __restrict__ circ_cur,
                int* __restrict__ circ_hist, int* __restrict__ n_circ, int* __restrict__ n_left, int* __restrict__ n_right,
                uint8_t* __restrict__ mask, B2HistHandover ho) {
    b2_pdl_wait();
    __shared__ int s_red[SF_THREADS / 32];
    const int p = blockIdx.x;
    const int nc = n_cur[p], nh = n_hist[p * hist_stride];   // hist_stride 0: every problem against the same set (the pool)
    if (nc == 0 || nh == 0) {   // :237-240 early-out on an empty pool; empty cur set -> nothing to match
        if (threadIdx.x == 0) { n_circ[p] = 0; if (n_left) n_left[p] = 0; if (n_right) n_right[p] = 0; }
        if (mask) for (int i = threadIdx.x; i < nc; i += SF_THREADS) mask[(size_t)p * cap + i] = 1;   // _curFeatureMask all true
    } else {
        const int* tl = ti_l + (size_t)p * cap; const int* dl = di_l + (size_t)p * cap;
        const int* tr = ti_r + (size_t)p * cap; const int* dr = di_r + (size_t)p * cap;
        int mnl = 0x7fffffff, mnr = 0x7fffffff;
        for (int i = threadIdx.x; i < nc; i += SF_THREADS) { mnl = min(mnl, dl[i]); mnr = min(mnr, dr[i]); }
        mnl = block_min(mnl, s_red);
        mnr = block_min(mnr, s_red);
        const float thl = fmaxf(__fmul_rn(ratio, (float)mnl), min_match);
        const float thr = fmaxf(__fmul_rn(ratio, (float)mnr), min_match);
        int base = 0, cl = 0, cr = 0;
        for (int i0 = 0; i0 < nc; i0 += SF_THREADS) {
            const int i = i0 + threadIdx.x;
            int okl = 0, okr = 0, t = -1;
            if (i < nc) { okl = (float)dl[i] < thl; okr = (float)dr[i] < thr; t = tl[i]; }
            const int keep = okl && okr && t == (i < nc ? tr[i] : -2);
            int tot;
            const int off = block_scan_flag(keep, s_red, &tot);
            if (keep) { circ_cur[(size_t)p * cap + base + off] = i; circ_hist[(size_t)p * cap + base + off] = t; }
            if (mask && i < nc) mask[(size_t)p * cap + i] = keep ? 0 : 1;   // _curFeatureMask[m.queryIdx] = false (:319)
            base += tot;
            cl += __syncthreads_count(okl);                                  // sizes of the two filtered lists (:259, :304)
            cr += __syncthreads_count(okr);
        }
        if (threadIdx.x == 0) { n_circ[p] = base; if (n_left) n_left[p] = cl; if (n_right) n_right[p] = cr; }
    }
    // One-pair calls of the consecutive-frame mode: this pair's kept descriptor rows become the history of the next call
    // (three device-to-device copies on the stream otherwise).  Nothing reads the history slot after the temporal matcher,
    // which has completed before this kernel started; the count was read above by every thread.
    if (ho.dst_l && p == gridDim.x - 1) {
        __syncthreads();
        const uint4* sl = reinterpret_cast<const uint4*>(ho.src_l); const uint4* sr = reinterpret_cast<const uint4*>(ho.src_r);
        uint4* dl4 = reinterpret_cast<uint4*>(ho.dst_l); uint4* dr4 = reinterpret_cast<uint4*>(ho.dst_r);
        for (int i = threadIdx.x; i < 2 * nc; i += SF_THREADS) { dl4[i] = sl[i]; dr4[i] = sr[i]; }
        if (threadIdx.x == 0) *ho.n_dst = nc;
    }
}

}  // namespace

// =================================================================================================
// tcgen05 matcher: the same contraction on the 5th-generation tensor cores.
//   popcount(q xor t) = |q| + sum_k (2 q_k - 1)(-t_k)      (signed 8-bit operands: q' in {+1,-1}, t' in {0,-128}),
// so the accumulator alone orders the candidates of a query row (|q| is added once at the end).  The train side carries
// -128 instead of -1 and a ninth K slice holds (1, 0, ...) on the query side and (column, 0, ...) on the train side, so the
// accumulator IS the key 128 * (distance - |q|) + column: the epilogue is a plain running min over TMEM values (BFMatcher's
// strict '<' in ascending train order == lexicographic min), no per-column arithmetic.
// One CTA = 128 query rows (TMEM lane = query row = thread).  Bits are expanded to bytes in registers (shift, PRMT
// sign-replicate: 0x00 / 0xFF = 0 / -1 for the train side, one more LOP3 makes +-1 on the query side) and stored to shared memory in the canonical K-major no-swizzle UMMA layout:
//   byte (row r, k) of a 128 x 256 operand tile at  (k / 16) * 2048 + r * 16 + k % 16
// (8-row x 16-byte core matrices; LBO = 2048 B between 16-byte K chunks, SBO = 128 B between 8-row groups).
// Per train tile of 128 rows one elected thread issues 8 x tcgen05.mma.kind::i8 (M128 N128 K32) into one of two
// 128-column TMEM accumulator stages and commits to an mbarrier; the expansion of the next tile and the epilogue of
// the previous one overlap the tensor pipe.  k order inside a row is free (both operands use the same one):
// byte 32 s + 4 j + i is bit 8 i + j of 32-bit word s.
constexpr int TC_ROWS = 128;                       // M and N of one MMA
constexpr int TC_KS = 9;                           // K slices of 32 bytes: 8 of descriptor bits + 1 that carries the column index
constexpr int TC_TILE = TC_ROWS * 32 * TC_KS;      // bytes of one expanded operand tile
constexpr uint32_t TC_IDESC = (2u << 4) | (1u << 7) | (1u << 10) | ((TC_ROWS >> 3) << 17) | ((TC_ROWS >> 4) << 24);
//                             D = s32     A = s8      B = s8       N >> 3                  M >> 4        (K-major A, B)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t tc_smem_desc(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(2048 >> 4) << 16) | ((uint64_t)(128 >> 4) << 32) | (1ull << 46);
}

// msb of every byte replicated over the byte (prmt with the sign-replicate bit of each selector nibble set;
// __byte_perm only honours the low three selector bits, hence the PTX)
__device__ __forceinline__ unsigned tc_sign_bytes(unsigned x) {
    unsigned d;
    asm("prmt.b32 %0, %1, %1, 0xBA98;" : "=r"(d) : "r"(x));
    return d;
}

// one descriptor row (8 words) -> 256 bytes of +-1 at row r of the tile at generic address `tile`
// QUERY: bit 1 -> +1, bit 0 -> -1;  train: bit 1 -> -128, bit 0 -> 0.  Rows beyond the set are all-zero words.
// One thread expands HALF a row: words 4*half .. 4*half+3 of row r.
template <bool QUERY>
__device__ __forceinline__ void tc_expand_half(uint8_t* tile, int r, int half, const uint4& q) {
    const unsigned w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int s = 0; s < 4; ++s) {
        unsigned e[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {   // bit 8i + j -> msb of byte i
            const unsigned m = w[s] << (7 - j);
            // query: msb replicated (0x00 / 0xFF), then 0xFF -> 0x01, 0x00 -> 0xFF (+1 / -1); train: the msb alone (0 / -128)
            e[j] = QUERY ? (~tc_sign_bytes(m) | 0x01010101u) : (m & 0x80808080u);
        }
        uint8_t* dst = tile + (size_t)(2 * (4 * half + s)) * 2048 + r * 16;
        *reinterpret_cast<uint4*>(dst) = make_uint4(e[0], e[1], e[2], e[3]);
        *reinterpret_cast<uint4*>(dst + 2048) = make_uint4(e[4], e[5], e[6], e[7]);
    }
}

__device__ __forceinline__ void tc_mbar_wait(uint32_t mbar, uint32_t parity) {
    asm volatile("{\n\t.reg .pred P1;\n\tLAB_WAIT:\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
                 "@P1 bra DONE;\n\tbra LAB_WAIT;\n\tDONE:\n\t}" ::"r"(mbar), "r"(parity) : "memory");
}

// TC_KS MMAs (K = 288) of one 128 x 128 tile into TMEM columns [tmem_d, tmem_d + 128), then commit -> mbarrier
__device__ __forceinline__ void tc_issue_tile(uint32_t tmem_d, uint32_t sa, uint32_t sb, uint32_t mbar) {
#pragma unroll
    for (int s = 0; s < TC_KS; ++s) {
        const uint64_t da = tc_smem_desc(sa + s * 4096), db = tc_smem_desc(sb + s * 4096);
        const uint32_t acc = s > 0, zero = 0;
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}"
                     ::"r"(tmem_d), "l"(da), "l"(db), "r"(TC_IDESC), "r"(acc), "r"(zero) : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(mbar) : "memory");
}

constexpr int TC_THREADS = 2 * TC_ROWS;   // two thread groups: each expands half of every row and scans half of the columns

__global__ void __launch_bounds__(TC_THREADS, 2)
match_tc_kernel(const B2MatchProblem* __restrict__ probs) {
    b2_pdl_wait();
    extern __shared__ __align__(1024) uint8_t tc_smem[];       // A tile, then two B tiles
    __shared__ __align__(8) uint64_t s_mbar[2];
    __shared__ uint32_t s_tmem;
    __shared__ int s_pop[TC_ROWS], s_key[TC_ROWS], s_t0[TC_ROWS];
    const B2MatchProblem P = probs[blockIdx.y];
    const int nq = *P.nq, nt = *P.nt;
    const int q0 = blockIdx.x * TC_ROWS;
    if (q0 >= nq) return;
    const int tid = threadIdx.x, wid = tid >> 5;
    const int r = tid & (TC_ROWS - 1), grp = tid >> 7;          // row inside the tile, thread group
    const int row = q0 + r;
    if (nt <= 0) {
        if (grp == 0 && row < nq) { P.train_idx[row] = -1; P.dist[row] = 0x7fffffff; }
        return;
    }
    uint8_t* sA = tc_smem;
    uint8_t* sB = tc_smem + TC_TILE;
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_mbar[0])));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_mbar[1])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (wid == 0) {   // two accumulator stages of 128 columns
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(256) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    {   // query tile
        uint4 q = make_uint4(0, 0, 0, 0);
        if (row < nq) q = reinterpret_cast<const uint4*>(P.q + (size_t)row * 32)[grp];
        tc_expand_half<true>(sA, r, grp, q);
        if (grp == 1) s_pop[r] = __popc(q.x) + __popc(q.y) + __popc(q.z) + __popc(q.w);
        else s_key[r] = __popc(q.x) + __popc(q.y) + __popc(q.z) + __popc(q.w);   // read back by the same thread below
    }
    {   // ninth K slice (constant): query byte 0 = 1, train byte 0 = row (= accumulator column) of the tile, rest 0
        const uint4 z = make_uint4(0u, 0u, 0u, 0u);
        const size_t o = (size_t)(16 + grp) * 2048 + r * 16;
        *reinterpret_cast<uint4*>(sA + o) = grp ? z : make_uint4(1u, 0u, 0u, 0u);
        *reinterpret_cast<uint4*>(sB + o) = grp ? z : make_uint4((unsigned)r, 0u, 0u, 0u);
        *reinterpret_cast<uint4*>(sB + TC_TILE + o) = grp ? z : make_uint4((unsigned)r, 0u, 0u, 0u);
    }
    const uint4* tp = reinterpret_cast<const uint4*>(P.t);
    const int ntiles = (nt + TC_ROWS - 1) / TC_ROWS;
    // This thread's half row of the next two train tiles lives in registers (two sets, the loop below is unrolled by
    // two so that no register copy — which would wait for the load — sits between a load and its use a tile later).
    const auto fetch_b = [&](int tile) {
        const int tr = tile * TC_ROWS + r;
        return tr < nt ? tp[2 * (size_t)tr + grp] : make_uint4(0, 0, 0, 0);
    };
    uint4 rb0 = fetch_b(0), rb1 = fetch_b(1);
    tc_expand_half<false>(sB, r, grp, rb0);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy stores -> visible to the tensor core
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = s_tmem;
    const uint32_t sa = smem_u32(sA), sb = smem_u32(sB);
    const uint32_t mb0 = smem_u32(&s_mbar[0]);
    if (tid == 0) tc_issue_tile(tmem, sa, sb, mb0);
    const int qpop_half = s_key[r];                              // (group 0: own value; group 1: unused)

    // running min over 128 * accumulator + column: strict-'<' ascending scan of BFMatcher == lexicographic min
    int best_key = 0x7FFFFFFF, best_t0 = 0;
    // one tile step: cur = registers holding tile i+1's half row, nxt = registers to refill with tile i+2's
    const auto step = [&](int i, const uint4& cur, uint4& nxt) {
        if (i + 1 < ntiles) tc_expand_half<false>(sB + ((i + 1) & 1) * TC_TILE, r, grp, cur);   // buffer last read by MMA i-1 (completed)
        nxt = fetch_b(i + 2);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();                                        // also: epilogue i-1 drained accumulator stage (i+1)&1
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (i + 1 < ntiles && tid == 0)
            tc_issue_tile(tmem + ((i + 1) & 1) * TC_ROWS, sa, sb + ((i + 1) & 1) * TC_TILE, mb0 + ((i + 1) & 1) * 8);
        tc_mbar_wait(mb0 + (i & 1) * 8, (i >> 1) & 1);          // MMA i done
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        // ---- epilogue: this thread's row (TMEM lane), its group's 64 columns in two chunks of 32 ----
        const int t0 = i * TC_ROWS;
        const int nvalid = nt - t0;                             // >= 1
        const uint32_t taddr = tmem + (i & 1) * TC_ROWS + ((uint32_t)((wid & 3) * 32) << 16) + grp * 64;
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
            const int col0 = grp * 64 + cc * 32;
            int v[32];
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                         "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                         "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                         : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                           "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                           "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                           "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                         : "r"(taddr + cc * 32));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            // v = 128 * acc + column-in-tile (acc in [-256, 256]) straight from the tensor core; padded train rows (last
            // tile) never win
            int (&key)[32] = v;
            if (nvalid < TC_ROWS) {
#pragma unroll
                for (int k = 0; k < 32; ++k)
                    if (col0 + k >= nvalid) key[k] = 0x7FFFFFFF;
            }
            int cm = key[0];
#pragma unroll
            for (int k = 1; k < 31; k += 2) cm = __vimin3_s32(cm, key[k], key[k + 1]);
            cm = min(cm, key[31]);
            // compare accumulators across chunks and tiles (a later one wins only with a strictly smaller distance)
            if ((cm >> 7) < (best_key >> 7)) { best_key = cm; best_t0 = t0; }
        }
    };
    for (int i = 0; i < ntiles; i += 2) {
        step(i, rb1, rb0);
        if (i + 1 < ntiles) step(i + 1, rb0, rb1);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    if (grp == 1) { s_key[r] = best_key; s_t0[r] = best_t0; }
    __syncthreads();
    if (wid == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256) : "memory");
    if (grp == 0 && row < nq) {
        // merge with the other group's columns: smaller distance, then smaller train index
        const int ok = s_key[r], ot = s_t0[r];
        const int da = best_key >> 7, db = ok >> 7;
        const int ia = best_t0 + (best_key & 127), ib = ot + (ok & 127);
        const bool other = db < da || (db == da && ib < ia);
        P.train_idx[row] = other ? ib : ia;
        P.dist[row] = (other ? db : da) + qpop_half + s_pop[r];
    }
}

// Stereo problems [0,np): desc(left p) vs desc(right np+p).  Temporal problems [np, 3np): cur slot p+1 vs hist
// slot p, left then right (src/feature-tracker.cpp:161-163, 258, 303).
__global__ void bind_kernel(B2MatchProblem* pr, int np, B2BindBases b) {
    b2_pdl_wait();
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= np) return;
    const size_t cap = b.cap, row = cap * 32;
    pr[p] = {b.desc + p * row, b.desc + (np + p) * row, b.nkp + p, b.nkp + np + p, b.train + p * cap, b.dist + p * cap};
    if (b.pool_l) {   // temporal problems against the device-resident feature pool (f3): the same train set for every pair
        pr[np + p] = {b.cur_l + (p + 1) * row, b.pool_l, b.nmatch + p + 1, b.pool_n, b.tti_l + p * cap, b.tdi_l + p * cap};
        pr[2 * np + p] = {b.cur_r + (p + 1) * row, b.pool_r, b.nmatch + p + 1, b.pool_n, b.tti_r + p * cap, b.tdi_r + p * cap};
        return;
    }
    pr[np + p] = {b.cur_l + (p + 1) * row, b.cur_l + p * row, b.nmatch + p + 1, b.nmatch + p, b.tti_l + p * cap, b.tdi_l + p * cap};
    pr[2 * np + p] = {b.cur_r + (p + 1) * row, b.cur_r + p * row, b.nmatch + p + 1, b.nmatch + p, b.tti_r + p * cap, b.tdi_r + p * cap};
}

void b2_launch_bind(B2MatchProblem* probs, int np, const B2BindBases& b, cudaStream_t s) {
    b2_launch_pdl(bind_kernel, dim3((np + 127) / 128), dim3(128), 0, s, probs, np, b);
}

void b2_launch_match(const B2MatchProblem* probs, int nprob, int max_nq, cudaStream_t s) {
    if (nprob <= 0 || max_nq <= 0) return;
    const int smem = 3 * TC_TILE;
    cudaFuncSetAttribute(match_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);   // per device; cheap
    dim3 grid((max_nq + TC_ROWS - 1) / TC_ROWS, nprob);
    b2_launch_pdl(match_tc_kernel, dim3(grid), dim3(TC_THREADS), smem, s, probs);
}

void b2_launch_stereo_filter(const void* kps, const uint8_t* desc, const int* n_kp, const int* train_idx,
                             const int* dist, int kp_cap, float ratio, float min_match, float max_dy, int* match_l,
                             int* match_r, int* n_match, uint8_t* cur_l, uint8_t* cur_r, double* pix_l, double* pix_r,
                             int npairs, cudaStream_t s) {
    b2_launch_pdl(stereo_filter_kernel, dim3(npairs), dim3(SF_THREADS), 0, s, static_cast<const KeyPointIn*>(kps), desc, n_kp, train_idx, dist,
                  kp_cap, ratio, min_match, max_dy, match_l, match_r, n_match, cur_l, cur_r,
                  reinterpret_cast<double2*>(pix_l), reinterpret_cast<double2*>(pix_r));
}

void b2_launch_circular(const int* ti_l, const int* di_l, const int* ti_r, const int* di_r, const int* n_cur,
                        const int* n_hist, int hist_stride, int cap, float ratio, float min_match, int* circ_cur,
                        int* circ_hist, int* n_circ, int* n_left, int* n_right, uint8_t* mask, int nprob, cudaStream_t s,
                        const B2HistHandover* handover) {
    B2HistHandover ho = {};
    if (handover) ho = *handover;
    b2_launch_pdl(circular_kernel, dim3(nprob), dim3(SF_THREADS), 0, s, ti_l, di_l, ti_r, di_r, n_cur, n_hist, hist_stride, cap, ratio,
                  min_match, circ_cur, circ_hist, n_circ, n_left, n_right, mask, ho);
}
