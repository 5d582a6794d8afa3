// Brute-force Hamming matching and the FeatureTracker filters.
//   * match_kernel — cv::BFMatcher(cv::NORM_HAMMING).match(query, train), src/feature-tracker.cpp:161-163,
//     :258, :303 (K9): for every query row the nearest train row by popcount(q xor t) over 32-byte
//     descriptors; train rows are scanned in ascending order with a strict '<' update, so the lowest
//     train index wins ties exactly like OpenCV's batch-distance loop.  Integer pipe only (LOP3 + POPC).
//   * stereo_filter_kernel — internalMatch, :164 and :208-218 (K10): global minDist, keep
//     dist < max(matchRatio*minDist, minMatchDist) && |yL - yR| < maxVerticalPixelDist (fp32), ordered
//     compaction in query order, gather of the kept descriptor rows (_curDescriptorsL/R).
//   * circular_kernel — externalTrack, :258-259, 293-296, 303-321 (useRANSAC=false): distance filter on both
//     sides and the circular test mapCurHist[queryIdx] == trainIdx.
#include "b2orb_internal.cuh"

namespace {

constexpr int MT_THREADS = 128;   // queries per CTA (one per thread)
constexpr int MT_TILE = 128;      // train descriptors staged per iteration

// Carry-save adder on 32 bit columns: a + b + c = s + 2*cy (two LOP3).
__device__ __forceinline__ void csa(unsigned a, unsigned b, unsigned c, unsigned& s, unsigned& cy) {
    s = a ^ b ^ c;
    cy = (a & b) | (a & c) | (b & c);
}

// Hamming distance of two 256-bit rows.  POPC issues on the quarter-rate XU pipe (ncu: 92 % busy in the plain
// 8-POPC loop), so three carry-save adders (6 LOP3 on the ALU pipe) first compress 7 of the 8 xor words into one
// "ones" word and three "twos" words: 5 POPC instead of 8, the two pipes end up about equally loaded.
__device__ __forceinline__ int hamming256(const uint4& qa, const uint4& qb, const uint4& ta, const uint4& tb) {
    unsigned o1, t1, o2, t2, o3, t3;
    csa(qa.x ^ ta.x, qa.y ^ ta.y, qa.z ^ ta.z, o1, t1);
    csa(qa.w ^ ta.w, qb.x ^ tb.x, qb.y ^ tb.y, o2, t2);
    csa(o1, o2, qb.z ^ tb.z, o3, t3);
    return __popc(o3) + __popc(qb.w ^ tb.w) + 2 * (__popc(t1) + __popc(t2) + __popc(t3));
}

__global__ void __launch_bounds__(MT_THREADS)
match_kernel(const B2MatchProblem* __restrict__ probs) {
    __shared__ uint4 s_t[MT_TILE * 2];
    const B2MatchProblem P = probs[blockIdx.y];
    const int nq = *P.nq, nt = *P.nt;
    const int q0 = blockIdx.x * MT_THREADS;
    if (q0 >= nq) return;
    const int qi = q0 + threadIdx.x;
    uint4 qa = make_uint4(0, 0, 0, 0), qb = qa;
    if (qi < nq) {
        const uint4* qp = reinterpret_cast<const uint4*>(P.q + (size_t)qi * 32);
        qa = qp[0];
        qb = qp[1];
    }
    // best = min over (distance << 23 | train index)  (distance <= 256, index < 2^23): ascending scan + strict '<' of BFMatcher == lexicographic min
    unsigned best = 0xFFFFFFFFu;
    const uint4* tp = reinterpret_cast<const uint4*>(P.t);
    for (int t0 = 0; t0 < nt; t0 += MT_TILE) {
        const int n = min(MT_TILE, nt - t0);
        __syncthreads();
        for (int i = threadIdx.x; i < 2 * n; i += MT_THREADS) s_t[i] = tp[(size_t)2 * t0 + i];
        __syncthreads();
#pragma unroll 4
        for (int j = 0; j < n; ++j) {
            const int d = hamming256(qa, qb, s_t[2 * j], s_t[2 * j + 1]);
            best = min(best, ((unsigned)d << 23) + (unsigned)(t0 + j));
        }
    }
    if (qi < nq) {
        P.train_idx[qi] = nt > 0 ? (int)(best & 0x7FFFFFu) : -1;
        P.dist[qi] = nt > 0 ? (int)(best >> 23) : 0x7fffffff;
    }
}

constexpr int SF_THREADS = 256;

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
__global__ void __launch_bounds__(SF_THREADS)
stereo_filter_kernel(const KeyPointIn* __restrict__ kps, const uint8_t* __restrict__ desc, const int* __restrict__ n_kp,
                     const int* __restrict__ train_idx, const int* __restrict__ dist, int cap, float ratio,
                     float min_match, float max_dy, int* __restrict__ match_l, int* __restrict__ match_r,
                     int* __restrict__ n_match, uint8_t* __restrict__ cur_l, uint8_t* __restrict__ cur_r) {
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
    int mn = 0x7fffffff;
    for (int i = threadIdx.x; i < nl; i += SF_THREADS) mn = min(mn, di[i]);
    mn = block_min(mn, s_red);
    const float thr = fmaxf(__fmul_rn(ratio, (float)mn), min_match);
    int base = 0;
    for (int i0 = 0; i0 < nl; i0 += SF_THREADS) {
        const int i = i0 + threadIdx.x;
        int keep = 0, t = 0;
        if (i < nl) {
            t = ti[i];
            keep = ((float)di[i] < thr) && (fabsf(__fsub_rn(kl[i].y, kr[t].y)) < max_dy);
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
            dl[0] = sl[0]; dl[1] = sl[1];
            dr[0] = sr[0]; dr[1] = sr[1];
        }
        base += tot;
    }
    if (threadIdx.x == 0) n_match[p] = base;
}

// One CTA per temporal problem: left and right match lists of the same cur set against the same hist set.
__global__ void __launch_bounds__(SF_THREADS)
circular_kernel(const int* __restrict__ ti_l, const int* __restrict__ di_l, const int* __restrict__ ti_r,
                const int* __restrict__ di_r, const int* __restrict__ n_cur, const int* __restrict__ n_hist, int cap,
                float ratio, float min_match, int* __restrict__ circ_cur, int* __restrict__ circ_hist,
                int* __restrict__ n_circ, int* __restrict__ n_left, int* __restrict__ n_right) {
    __shared__ int s_red[SF_THREADS / 32];
    const int p = blockIdx.x;
    const int nc = n_cur[p], nh = n_hist[p];
    if (nc == 0 || nh == 0) {   // :237-240 early-out on an empty pool; empty cur set -> nothing to match
        if (threadIdx.x == 0) { n_circ[p] = 0; if (n_left) n_left[p] = 0; if (n_right) n_right[p] = 0; }
        return;
    }
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
        int okl = 0, okr = 0;
        if (i < nc) { okl = (float)dl[i] < thl; okr = (float)dr[i] < thr; }
        const int keep = okl && okr && tl[i] == tr[i];
        int tot, t2;
        const int off = block_scan_flag(keep, s_red, &tot);
        if (keep) { circ_cur[(size_t)p * cap + base + off] = i; circ_hist[(size_t)p * cap + base + off] = tl[i]; }
        base += tot;
        block_scan_flag(okl, s_red, &t2); cl += t2;
        block_scan_flag(okr, s_red, &t2); cr += t2;
    }
    if (threadIdx.x == 0) { n_circ[p] = base; if (n_left) n_left[p] = cl; if (n_right) n_right[p] = cr; }
}

}  // namespace

// Stereo problems [0,np): desc(left p) vs desc(right np+p).  Temporal problems [np, 3np): cur slot p+1 vs hist
// slot p, left then right (src/feature-tracker.cpp:161-163, 258, 303).
__global__ void bind_kernel(B2MatchProblem* pr, int np, B2BindBases b) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= np) return;
    const size_t cap = b.cap, row = cap * 32;
    pr[p] = {b.desc + p * row, b.desc + (np + p) * row, b.nkp + p, b.nkp + np + p, b.train + p * cap, b.dist + p * cap};
    pr[np + p] = {b.cur_l + (p + 1) * row, b.cur_l + p * row, b.nmatch + p + 1, b.nmatch + p, b.tti_l + p * cap, b.tdi_l + p * cap};
    pr[2 * np + p] = {b.cur_r + (p + 1) * row, b.cur_r + p * row, b.nmatch + p + 1, b.nmatch + p, b.tti_r + p * cap, b.tdi_r + p * cap};
}

void b2_launch_bind(B2MatchProblem* probs, int np, const B2BindBases& b, cudaStream_t s) {
    bind_kernel<<<(np + 127) / 128, 128, 0, s>>>(probs, np, b);
}

void b2_launch_match(const B2MatchProblem* probs, int nprob, int max_nq, cudaStream_t s) {
    if (nprob <= 0 || max_nq <= 0) return;
    dim3 grid((max_nq + MT_THREADS - 1) / MT_THREADS, nprob);
    match_kernel<<<grid, MT_THREADS, 0, s>>>(probs);
}

void b2_launch_stereo_filter(const void* kps, const uint8_t* desc, const int* n_kp, const int* train_idx,
                             const int* dist, int kp_cap, float ratio, float min_match, float max_dy, int* match_l,
                             int* match_r, int* n_match, uint8_t* cur_l, uint8_t* cur_r, int npairs,
                             cudaStream_t s) {
    stereo_filter_kernel<<<npairs, SF_THREADS, 0, s>>>(static_cast<const KeyPointIn*>(kps), desc, n_kp, train_idx, dist,
                                                       kp_cap, ratio, min_match, max_dy, match_l, match_r, n_match,
                                                       cur_l, cur_r);
}

void b2_launch_circular(const int* ti_l, const int* di_l, const int* ti_r, const int* di_r, const int* n_cur,
                        const int* n_hist, int cap, float ratio, float min_match, int* circ_cur, int* circ_hist,
                        int* n_circ, int* n_left, int* n_right, int nprob, cudaStream_t s) {
    circular_kernel<<<nprob, SF_THREADS, 0, s>>>(ti_l, di_l, ti_r, di_r, n_cur, n_hist, cap, ratio, min_match,
                                                 circ_cur, circ_hist, n_circ, n_left, n_right);
}
