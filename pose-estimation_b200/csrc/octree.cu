// Quadtree keypoint culling — ORBextractor::DistributeOctTree + ExtractorNode::DivideNode,
// src/ORBextractor.cc:481-763 (K4).  Exact semantics, including list order (it is the output order) and
// the partial last pass; the per-cell top-k idea of the north star survives only as the final
// "best response per node" segmented arg-max, because keypoints must be bit-exact.
//
// Parallel restatement.  The node ARRAY is the reference's std::list in list order; a "round" replaces a
// set of multi-point nodes E_1..E_m by their non-empty children.  Because children are pushed to the list
// front one expansion after another (:618-658, :695-735), the list after a round is
//     rev(children(E_m)) ++ ... ++ rev(children(E_1)) ++ (old list without E_1..E_m, order kept),
// so every new position is a prefix sum.  Outer pass (:594-672): E = all multi-point nodes in list order,
// no early stop.  Partial pass (:673-738): E sorted by (point count desc, creation order desc) and m = the
// first prefix that brings the list to >= N nodes.  The reference breaks count ties by heap address
// (:684 sorts pair<int, ExtractorNode*>); the documented deterministic rule shared with the oracle is
// "later created first", and since list order among freshly created nodes is reverse creation order that
// is simply "smaller list position first".  The final pick per node is the largest response, first in
// the reference's candidate order on ties (:741-760); candidates arrive unordered here, so the tie-break
// key re-derives the cell-row-major / pixel-row-major order of the cell loop (:789-826) from coordinates.
// One CTA per (image, level); all list state lives in shared memory.
//
// Two ways of knowing how many points fall into each child of a node:
//  * COUNT PYRAMID (the normal path).  The splits of DivideNode depend only on the node rectangle, never on the
//    points, so the node a point ends up in at depth d is a function of (x, y) alone: root index, then d quadrant
//    choices = a 2-bit-per-level path code.  ONE pass over the points histograms them at depth D (L.ot_depth), a
//    4-ary reduction gives the counts of every possible node of depths D-1..0, and the rounds then run on nodes only
//    (child counts = pyramid look-ups).  A second pass finds each point's final node by walking its code up a
//    pyramid of node ids and takes the per-node best.  The points are read twice in total instead of once per round.
//  * POINT PASSES (fallback, the first version): every round moves every point to its new node and counts it into
//    that node's children.  Used when the level has few points, and when a node at depth D still has to be split
//    (sparse, clustered points), in which case the level is simply redone this way.
#include "b2orb_internal.cuh"

namespace {

#ifdef OT_TRACE
__device__ long long g_ot_trace[64];
#define OT_MARK(k) do { __syncthreads(); if (threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0) g_ot_trace[k] = clock64(); } while (0)
#else
#define OT_MARK(k) do { } while (0)
#endif
__device__ unsigned long long g_ot_fallbacks;   // levels redone with point passes because the pyramid was too shallow
constexpr int OT_THREADS = 512;
constexpr int OT_MAX_ROOTS = 256;
constexpr int OT_ILP = 4;
constexpr int OT_ILP_PYR = 8;   // loads in flight per thread in the two passes of the count-pyramid path

struct OtShared {
    unsigned warp_sums[OT_THREADS / 32];
    unsigned total;
    int lsz, m_exp, t_children, n_multi, n_cand, too_deep;
    unsigned root_cnt[OT_MAX_ROOTS];
    int root_map[OT_MAX_ROOTS];
};

// In-place exclusive scan of a[0..n) by the whole block; returns the total.
__device__ unsigned block_scan_excl(unsigned* a, int n, OtShared& sh) {
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int ipt = (n + OT_THREADS - 1) / OT_THREADS;
    const int lo = min(tid * ipt, n), hi = min(lo + ipt, n);
    unsigned sum = 0;
    for (int i = lo; i < hi; ++i) sum += a[i];
    unsigned incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        unsigned v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) sh.warp_sums[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        unsigned v = lane < OT_THREADS / 32 ? sh.warp_sums[lane] : 0;
        unsigned inc2 = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            unsigned u = __shfl_up_sync(0xFFFFFFFFu, inc2, o);
            if (lane >= o) inc2 += u;
        }
        if (lane < OT_THREADS / 32) sh.warp_sums[lane] = inc2 - v;
        if (lane == OT_THREADS / 32 - 1) sh.total = inc2;
    }
    __syncthreads();
    unsigned run = sh.warp_sums[wid] + incl - sum;
    for (int i = lo; i < hi; ++i) {
        unsigned v = a[i];
        a[i] = run;
        run += v;
    }
    __syncthreads();
    return sh.total;
}

__device__ __forceinline__ int quadrant(const B2Node& nd, int x, int y) {
    // DivideNode :483-484: halfX = ceil((UR.x-UL.x)/2); points go left iff x < UL.x+halfX, up iff y < UL.y+halfY
    const int mx = nd.x0 + ((nd.x1 - nd.x0 + 1) >> 1);
    const int my = nd.y0 + ((nd.y1 - nd.y0 + 1) >> 1);
    return (x >= mx ? 1 : 0) + (y >= my ? 2 : 0);
}

__device__ __forceinline__ B2Node child_rect(const B2Node& nd, int q, unsigned cnt) {
    const int mx = nd.x0 + ((nd.x1 - nd.x0 + 1) >> 1);
    const int my = nd.y0 + ((nd.y1 - nd.y0 + 1) >> 1);
    B2Node c;
    c.x0 = (q & 1) ? mx : nd.x0;
    c.x1 = (q & 1) ? nd.x1 : mx;
    c.y0 = (q & 2) ? my : nd.y0;
    c.y1 = (q & 2) ? nd.y1 : my;
    c.cnt = cnt;
    c.code = ((nd.code & 0x0FFFFFFFu) << 2 | (unsigned)q) | ((nd.code >> 28) + 1u) << 28;   // depth in the top 4 bits
    return c;
}

// Count pyramid: level d holds n_ini * 4^d counters, level d starts at n_ini * (4^d - 1) / 3.
__device__ __forceinline__ unsigned pyr_off(int n_ini, int d) { return (unsigned)n_ini * (((1u << (2 * d)) - 1u) / 3u); }
__device__ __forceinline__ unsigned node_depth(const B2Node& nd) { return nd.code >> 28; }
__device__ __forceinline__ unsigned node_path(const B2Node& nd) { return nd.code & 0x0FFFFFFFu; }

// Path code of a point at depth D: root by the float division of :566-570, then D midpoint decisions (DivideNode).
// DivideNode halves x and y independently, so the path is the bit-interleave of an x path and a y path, each a function
// of one coordinate: the CTA tabulates them once (s_xc[x] = root << 7 | x path bits, s_yc[y] = y path bits) and a
// point's code costs four shared-memory loads.
constexpr int OT_MAX_DIM = 4096 + 32;
struct OtTables {
    unsigned short xc[OT_MAX_DIM], yc[OT_MAX_DIM], spread[128];
};

__device__ void build_code_tables(OtTables& t, int w, int h, int n_ini, float h_x, int hgt, int D) {
    for (int x = threadIdx.x; x < w; x += OT_THREADS) {
        int r = (int)((float)x / h_x);
        r = min(r, n_ini - 1);
        int x0 = (int)(h_x * (float)r), x1 = (int)(h_x * (float)(r + 1));
        unsigned bits = 0;
        for (int d = 0; d < D; ++d) {
            const int mx = x0 + ((x1 - x0 + 1) >> 1);
            const bool right = x >= mx;
            x0 = right ? mx : x0; x1 = right ? x1 : mx;
            bits = bits << 1 | (right ? 1u : 0u);
        }
        t.xc[x] = (unsigned short)((unsigned)r << 7 | bits);
    }
    for (int y = threadIdx.x; y < h; y += OT_THREADS) {
        int y0 = 0, y1 = hgt;
        unsigned bits = 0;
        for (int d = 0; d < D; ++d) {
            const int my = y0 + ((y1 - y0 + 1) >> 1);
            const bool down = y >= my;
            y0 = down ? my : y0; y1 = down ? y1 : my;
            bits = bits << 1 | (down ? 1u : 0u);
        }
        t.yc[y] = (unsigned short)bits;
    }
    if (threadIdx.x < 128) {
        unsigned v = threadIdx.x, sp = 0;
        for (int i = 0; i < 7; ++i) sp |= ((v >> i) & 1u) << (2 * i);
        t.spread[threadIdx.x] = (unsigned short)sp;
    }
}

__device__ __forceinline__ unsigned point_code(const OtTables& t, uint32_t v, int D) {
    const unsigned xc = t.xc[v & 0xFFF], yc = t.yc[(v >> 12) & 0xFFF];
    return (xc >> 7) << (2 * D) | t.spread[xc & 127] | (unsigned)t.spread[yc] << 1;
}

// One expansion round.  partial == false: outer pass; true: sorted partial pass with early stop at N.
// cc holds the child point counts of every multi-point node of `cur` (filled by the previous round's point pass, or
// by the root pass); ccn receives the same for the new list while the points are moved: ONE pass over the points
// per round.
template <bool PYR>
__device__ void ot_round(bool partial, int N, const uint32_t* __restrict__ pts, unsigned short* __restrict__ pnode,
                         int M, B2Node* cur, B2Node* nxt, const unsigned* cc, unsigned* ccn, int* cpos, int* newpos,
                         unsigned* rank, unsigned* order, unsigned* epre, unsigned* flag, int node_cap, OtShared& sh,
                         const unsigned* pyr, int n_ini, int D) {
    const int tid = threadIdx.x;
    const int lsz = sh.lsz;
    // points in child q of node i: pyramid look-up (PYR) or the counts left by the previous point pass
    auto child_cnt = [&](int i, int q) -> unsigned {
        if (PYR) {
            const B2Node nd = cur[i];
            return pyr[pyr_off(n_ini, (int)node_depth(nd) + 1) + (node_path(nd) << 2) + q];
        }
        return cc[4 * i + q];
    };
    // 1. candidates = multi-point nodes
    for (int i = tid; i < lsz; i += OT_THREADS) {
        const B2Node nd = cur[i];
        flag[i] = nd.cnt > 1 ? 1u : 0u;
        if (PYR && nd.cnt > 1 && (int)node_depth(nd) >= D) sh.too_deep = 1;   // its children are below the pyramid
    }
    __syncthreads();
    if (PYR && sh.too_deep) return;
    // 3. expansion order: rank r of candidate i, order[r] = i
    //    flag -> compact candidate index (list order)
    const int n_cand = (int)block_scan_excl(flag, lsz, sh);
    for (int i = tid; i < lsz; i += OT_THREADS)
        if (cur[i].cnt > 1) epre[flag[i]] = i;          // epre temporarily = candidate list
    __syncthreads();
    if (!partial) {
        for (int c = tid; c < n_cand; c += OT_THREADS) { const int i = epre[c]; rank[i] = c; order[c] = i; }
    } else {
        // rank = number of candidates that come first: larger node, or equal size and earlier in the list (the candidate
        // list is in list order, so "earlier" is the smaller candidate index).  O(n_cand^2) broadcast reads: the sizes are
        // staged densely in `flag` (free until step 5) so the inner loop is one independent shared-memory load per step.
        for (int c = tid; c < n_cand; c += OT_THREADS) flag[c] = cur[epre[c]].cnt;
        __syncthreads();
        for (int c = tid; c < n_cand; c += OT_THREADS) {
            const unsigned ci = flag[c];
            unsigned r = 0;
            int d = 0;
#pragma unroll 4
            for (; d < c; ++d) r += flag[d] >= ci ? 1u : 0u;            // earlier in the list: ties count
#pragma unroll 4
            for (d = c + 1; d < n_cand; ++d) r += flag[d] > ci ? 1u : 0u;
            const int i = epre[c];
            rank[i] = r;
            order[r] = i;
        }
    }
    __syncthreads();
    // 4. e_r = non-empty children of the r-th expansion; inclusive prefix; early-stop index m
    for (int r = tid; r < n_cand; r += OT_THREADS) {
        const int i = order[r];
        epre[r] = (child_cnt(i, 0) != 0) + (child_cnt(i, 1) != 0) + (child_cnt(i, 2) != 0) + (child_cnt(i, 3) != 0);
    }
    __syncthreads();
    block_scan_excl(epre, n_cand, sh);   // epre[r] = sum_{t<r} e_t
    if (tid == 0) { sh.m_exp = n_cand; sh.n_multi = 0; }
    __syncthreads();
    if (partial) {
        // size after expanding ranks 0..r  =  lsz + sum_{t<=r} (e_t - 1); take the first r reaching N
        for (int r = tid; r < n_cand; r += OT_THREADS) {
            const int i = order[r];
            const int e_r = (child_cnt(i, 0) != 0) + (child_cnt(i, 1) != 0) + (child_cnt(i, 2) != 0) + (child_cnt(i, 3) != 0);
            const int after = lsz + (int)epre[r] + e_r - (r + 1);
            const int before = lsz + (int)epre[r] - r;
            if (after >= N && before < N) sh.m_exp = r + 1;
        }
        __syncthreads();
    }
    const int m_exp = sh.m_exp;
    if (tid == 0) {
        int tch = 0;
        if (m_exp > 0) {
            const int i = order[m_exp - 1];
            tch = (int)epre[m_exp - 1] + (child_cnt(i, 0) != 0) + (child_cnt(i, 1) != 0) + (child_cnt(i, 2) != 0) +
                  (child_cnt(i, 3) != 0);
        }
        sh.t_children = tch;
    }
    // 5. positions of surviving (unexpanded) nodes
    for (int i = tid; i < lsz; i += OT_THREADS) flag[i] = (cur[i].cnt > 1 && (int)rank[i] < m_exp) ? 0u : 1u;
    __syncthreads();
    const int t_children = sh.t_children;
    const int n_keep = (int)block_scan_excl(flag, lsz, sh);
    const int new_lsz = t_children + n_keep;
    if (new_lsz > node_cap) {           // cannot happen (see b2orb_capi.cu sizing); never corrupt memory
        if (tid == 0) sh.lsz = -1;
        __syncthreads();
        return;
    }
    // 6. write the new list
    int multi = 0;
    for (int i = tid; i < lsz; i += OT_THREADS) {
        const B2Node nd = cur[i];
        if (nd.cnt > 1 && (int)rank[i] < m_exp) {
            const int r = rank[i];
            const unsigned c0 = child_cnt(i, 0), c1 = child_cnt(i, 1), c2 = child_cnt(i, 2), c3 = child_cnt(i, 3);
            const int e_r = (c0 != 0) + (c1 != 0) + (c2 != 0) + (c3 != 0);
            // children pushed n1..n4 to the front -> n4 first; block of this expansion starts at
            // t_children - (epre[r] + e_r)
            int pos = t_children - ((int)epre[r] + e_r);
            const unsigned cs[4] = {c0, c1, c2, c3};
#pragma unroll
            for (int q = 3; q >= 0; --q) {
                if (cs[q]) {
                    nxt[pos] = child_rect(nd, q, cs[q]);
                    if (!PYR) {
                        ccn[4 * pos] = ccn[4 * pos + 1] = ccn[4 * pos + 2] = ccn[4 * pos + 3] = 0;
                        cpos[4 * i + q] = pos;
                    }
                    multi += cs[q] > 1;
                    ++pos;
                } else if (!PYR) cpos[4 * i + q] = -1;
            }
            if (!PYR) newpos[i] = -1;
        } else {
            const int pos = t_children + (int)flag[i];
            nxt[pos] = nd;
            if (!PYR) {
                ccn[4 * pos] = ccn[4 * pos + 1] = ccn[4 * pos + 2] = ccn[4 * pos + 3] = 0;
                newpos[i] = pos;
            }
            multi += nd.cnt > 1;
        }
    }
    if (multi) atomicAdd(&sh.n_multi, multi);
    __syncthreads();
    // 7. move the points and count them into the children of their new node (next round's cc)
    // (point loops are batched OT_ILP deep: the pass is a chain of dependent L2 loads, so the loads of several
    //  points are issued before any of them is used)
    for (int p0 = tid; !PYR && p0 < M; p0 += OT_ILP * OT_THREADS) {
        int nn[OT_ILP];
        uint32_t vv[OT_ILP];
#pragma unroll
        for (int k = 0; k < OT_ILP; ++k) {
            const int p = p0 + k * OT_THREADS;
            nn[k] = p < M ? (int)pnode[p] : -1;
            vv[k] = p < M ? pts[p] : 0u;
        }
#pragma unroll
        for (int k = 0; k < OT_ILP; ++k) {
            if (nn[k] < 0) continue;
            const int n = nn[k];
            const int x = vv[k] & 0xFFF, y = (vv[k] >> 12) & 0xFFF;
            int np = newpos[n];
            if (np < 0) {
                np = cpos[4 * n + quadrant(cur[n], x, y)];
                pnode[p0 + k * OT_THREADS] = (unsigned short)np;
            } else if (np != n) pnode[p0 + k * OT_THREADS] = (unsigned short)np;
            const B2Node nd = nxt[np];
            if (nd.cnt > 1) atomicAdd(&ccn[4 * np + quadrant(nd, x, y)], 1u);
        }
    }
    __syncthreads();
    if (tid == 0) sh.lsz = new_lsz;
    __syncthreads();
}

// A FULL expansion round on the count pyramid (every multi-point node splits, :594-672): the same new list as
// ot_round<true>(false, ...), with ONE block scan instead of three.  Candidate r (in list order) contributes e_r non-empty
// children, a surviving node contributes itself; both running sums travel in one word (children in the low half, survivors
// in the high half), so a node knows its new position from its own prefix: children of candidate r at
// t_children - (sum_{t<r} e_t + e_r) .. (pushed to the front, last child first), survivors behind all children in list order.
__device__ void ot_round_full_pyr(B2Node* cur, B2Node* nxt, unsigned* flag, int node_cap, OtShared& sh, const unsigned* pyr,
                                  int n_ini, int D) {
    const int tid = threadIdx.x;
    const int lsz = sh.lsz;
    if (tid == 0) sh.n_multi = 0;
    for (int i = tid; i < lsz; i += OT_THREADS) {
        const B2Node nd = cur[i];
        unsigned v = 1u << 16;
        if (nd.cnt > 1) {
            if ((int)node_depth(nd) >= D) sh.too_deep = 1;   // its children are below the pyramid
            else {
                const unsigned* c = pyr + pyr_off(n_ini, (int)node_depth(nd) + 1) + (node_path(nd) << 2);
                v = (c[0] != 0) + (c[1] != 0) + (c[2] != 0) + (c[3] != 0);
            }
        }
        flag[i] = v;
    }
    __syncthreads();
    if (sh.too_deep) return;
    const unsigned total = block_scan_excl(flag, lsz, sh);
    const int t_children = (int)(total & 0xFFFFu), n_keep = (int)(total >> 16);
    const int new_lsz = t_children + n_keep;
    if (new_lsz > node_cap) {           // cannot happen (see b2orb_capi.cu sizing); never corrupt memory
        if (tid == 0) sh.lsz = -1;
        __syncthreads();
        return;
    }
    int multi = 0;
    for (int i = tid; i < lsz; i += OT_THREADS) {
        const B2Node nd = cur[i];
        const unsigned pre = flag[i];
        if (nd.cnt > 1) {
            const unsigned* c = pyr + pyr_off(n_ini, (int)node_depth(nd) + 1) + (node_path(nd) << 2);
            const unsigned cs[4] = {c[0], c[1], c[2], c[3]};
            const int e_r = (cs[0] != 0) + (cs[1] != 0) + (cs[2] != 0) + (cs[3] != 0);
            int pos = t_children - ((int)(pre & 0xFFFFu) + e_r);
#pragma unroll
            for (int q = 3; q >= 0; --q)
                if (cs[q]) {
                    nxt[pos++] = child_rect(nd, q, cs[q]);
                    multi += cs[q] > 1;
                }
        } else {
            nxt[t_children + (int)(pre >> 16)] = nd;
        }
    }
    if (multi) atomicAdd(&sh.n_multi, multi);
    __syncthreads();
    if (tid == 0) sh.lsz = new_lsz;
    __syncthreads();
}

// The main loop of DistributeOctTree (:592-739) on whichever source of child counts; leaves the final list in `cur`
// (returned through the reference) and its size in sh.lsz (-1 = workspace guard tripped).  With PYR it may stop early
// with sh.too_deep set.
template <bool PYR>
__device__ void ot_main_loop(int N, const uint32_t* pts, unsigned short* pnode, int M, B2Node*& cur, B2Node*& nxt,
                             unsigned*& cc, unsigned*& ccn, int* cpos, int* newpos, unsigned* rank, unsigned* order,
                             unsigned* epre, unsigned* flag, int Q, OtShared& sh, const unsigned* pyr, int n_ini, int D) {
    bool finish = false;
#ifdef OT_TRACE
    int tr = 20;
#define OT_RMARK() do { if (tr < 60) { OT_MARK(tr); ++tr; } } while (0)
#else
#define OT_RMARK() do { } while (0)
#endif
    while (!finish) {
        int prev = sh.lsz;
        OT_RMARK();
        if (PYR && Q <= 16000) ot_round_full_pyr(cur, nxt, flag, Q, sh, pyr, n_ini, D);   // (both packed sums stay below 2^16)
        else ot_round<PYR>(false, N, pts, pnode, M, cur, nxt, cc, ccn, cpos, newpos, rank, order, epre, flag, Q, sh, pyr, n_ini, D);
        if (PYR && sh.too_deep) return;
        { B2Node* t = cur; cur = nxt; nxt = t; unsigned* u = cc; cc = ccn; ccn = u; }
        int lsz = sh.lsz;
        if (lsz < 0) break;
        if (lsz >= N || lsz == prev) finish = true;
        else if (lsz + 3 * sh.n_multi > N) {
            while (!finish) {
                prev = sh.lsz;
                OT_RMARK();
                ot_round<PYR>(true, N, pts, pnode, M, cur, nxt, cc, ccn, cpos, newpos, rank, order, epre, flag, Q, sh, pyr, n_ini, D);
                if (PYR && sh.too_deep) return;
                { B2Node* t = cur; cur = nxt; nxt = t; unsigned* u = cc; cc = ccn; ccn = u; }
                lsz = sh.lsz;
                if (lsz < 0 || lsz >= N || lsz == prev) finish = true;
            }
        }
    }
}

// Response + tie-break key of a candidate for the per-node arg-max (:741-760): largest response, then the earliest
// position in the reference's cell-row-major / pixel-row-major candidate order.  x / w_cell and y / h_cell by
// multiplication with ceil(2^22 / d): exact for x < 4096, d >= 3 (x * d < 2^22), no 32-bit overflow (x * m < 2^32 / 3).
__device__ __forceinline__ unsigned best_key(uint32_t v, int wc, int hc, int ncol, unsigned mwc, unsigned mhc) {
    const int x = (int)(v & 0xFFF) - 3, y = (int)((v >> 12) & 0xFFF) - 3;
    const int cj = (int)(((unsigned)x * mwc) >> 22), ci = (int)(((unsigned)y * mhc) >> 22);
    const unsigned key = (unsigned)(((ci * ncol + cj) * hc + (y - ci * hc)) * wc + (x - cj * wc));
    return (v & 0xFF000000u) | (0x00FFFFFFu - key);
}

__global__ void __launch_bounds__(OT_THREADS)
octree_kernel(const __grid_constant__ B2Geom g, const uint32_t* __restrict__ cand, const int* __restrict__ cand_cnt,
              unsigned short* __restrict__ pnode_all, unsigned* __restrict__ gscratch, int use_global,
              uint32_t* __restrict__ sel, int* __restrict__ sel_cnt, int force_legacy) {
    b2_pdl_wait();
    extern __shared__ __align__(16) unsigned dyn[];
    __shared__ OtShared sh;
    __shared__ OtTables tab;
    const int level = blockIdx.x, b = blockIdx.y;
    const B2Level& L = g.lv[level];
    const int tid = threadIdx.x;
    const int N = L.quota;
    const int Q = L.node_cap;
    int M = cand_cnt[b * g.nlevels + level];
    M = min(M, L.cand_cap);
    const uint32_t* pts = cand + (size_t)b * g.cand_total + L.cand_off;
    unsigned short* pnode = pnode_all + (size_t)b * g.cand_total + L.cand_off;
    uint32_t* out = sel + (size_t)b * g.sel_total + L.sel_base;
    if (M == 0) {
        if (tid == 0) sel_cnt[b * g.nlevels + level] = 0;
        return;
    }
    // workspace carve-up (B2_OT_WORDS*Q words): two node arrays (4 words each), two child-count arrays, scratch.
    // The count pyramid (at most 13*Q words, see build_geometry) lives where the point-pass path keeps cc .. newpos.
    unsigned* ws = use_global ? gscratch + ((size_t)b * g.node_total + L.node_off) * B2_OT_WORDS : dyn;
    B2Node* cur = reinterpret_cast<B2Node*>(ws);
    B2Node* nxt = reinterpret_cast<B2Node*>(ws + 4 * Q);
    unsigned* cc = ws + 8 * Q;
    unsigned* ccn = ws + 12 * Q;
    int* cpos = reinterpret_cast<int*>(ws + 16 * Q);
    int* newpos = reinterpret_cast<int*>(ws + 20 * Q);
    unsigned* rank = ws + 21 * Q;
    unsigned* order = ws + 22 * Q;
    unsigned* epre = ws + 23 * Q;
    unsigned* flag = ws + 24 * Q;
    unsigned* best = ws + 25 * Q;
    unsigned* pyr = ws + 8 * Q;

    const int n_ini = L.n_ini;
    const float h_x = L.h_x;
    const int hgt = L.h - 2 * B2_BORDER;
    const int wc = L.w_cell, hc = L.h_cell, ncol = L.n_cols;
    const int D = L.ot_depth;
    const unsigned mwc = ((1u << 22) + wc - 1) / wc, mhc = ((1u << 22) + hc - 1) / hc;
    // few points: the point passes are cheap and deep, sparse trees are likely
    bool use_pyr = D >= 1 && !force_legacy && M >= 2 * N && M >= 1024;
    if (tid == 0) sh.too_deep = 0;

    OT_MARK(0);
    if (use_pyr) {
        // ---- count pyramid: histogram at depth D, then 4-ary sums upwards
        const int pyr_words = (int)pyr_off(n_ini, D + 1);
        for (int i = tid; i < pyr_words; i += OT_THREADS) pyr[i] = 0;
        build_code_tables(tab, min(L.w, OT_MAX_DIM), min(L.h, OT_MAX_DIM), n_ini, h_x, hgt, D);
        __syncthreads();
        unsigned* leaf = pyr + pyr_off(n_ini, D);
        for (int p0 = tid; p0 < M; p0 += OT_ILP_PYR * OT_THREADS) {
            uint32_t vv[OT_ILP_PYR];
#pragma unroll
            for (int k = 0; k < OT_ILP_PYR; ++k) vv[k] = p0 + k * OT_THREADS < M ? pts[p0 + k * OT_THREADS] : 0xFFFFFFFFu;
#pragma unroll
            for (int k = 0; k < OT_ILP_PYR; ++k)
                if (p0 + k * OT_THREADS < M)
                    atomicAdd(&leaf[point_code(tab, vv[k], D)], 1u);
        }
        __syncthreads();
        OT_MARK(1);
        for (int d = D - 1; d >= 0; --d) {
            unsigned* up = pyr + pyr_off(n_ini, d);
            const unsigned* dn = pyr + pyr_off(n_ini, d + 1);
            const int cnt = n_ini << (2 * d);
            for (int i = tid; i < cnt; i += OT_THREADS) up[i] = dn[4 * i] + dn[4 * i + 1] + dn[4 * i + 2] + dn[4 * i + 3];
            __syncthreads();
        }
        // roots (:543-590): nIni = round(W'/H'), hX = W'/nIni; point -> root (int)(x / hX); empty roots dropped
        if (tid == 0) {
            int k = 0;
            for (int i = 0; i < n_ini; ++i) {
                if (pyr[i]) {
                    B2Node nd;
                    nd.x0 = (unsigned short)(int)(h_x * (float)i);
                    nd.x1 = (unsigned short)(int)(h_x * (float)(i + 1));
                    nd.y0 = 0;
                    nd.y1 = (unsigned short)hgt;
                    nd.cnt = pyr[i];
                    nd.code = (unsigned)i;          // depth 0
                    cur[k++] = nd;
                }
            }
            sh.lsz = k;
        }
        __syncthreads();
        OT_MARK(2);
        ot_main_loop<true>(N, pts, pnode, M, cur, nxt, cc, ccn, cpos, newpos, rank, order, epre, flag, Q, sh, pyr, n_ini, D);
        __syncthreads();
        if (sh.too_deep) {          // a node below the pyramid must be split: redo the level with point passes
            use_pyr = false;
            if (tid == 0) atomicAdd(&g_ot_fallbacks, 1ull);
            cur = reinterpret_cast<B2Node*>(ws);
            nxt = reinterpret_cast<B2Node*>(ws + 4 * Q);
            cc = ws + 8 * Q;
            ccn = ws + 12 * Q;
        }
    }
    if (!use_pyr) {
        for (int i = tid; i < n_ini; i += OT_THREADS) sh.root_cnt[i] = 0;
        __syncthreads();
        for (int p = tid; p < M; p += OT_THREADS) {
            int r = (int)((float)(pts[p] & 0xFFF) / h_x);
            r = min(r, n_ini - 1);
            atomicAdd(&sh.root_cnt[r], 1u);
        }
        __syncthreads();
        if (tid == 0) {
            int k = 0;
            for (int i = 0; i < n_ini; ++i) {
                if (sh.root_cnt[i]) {
                    B2Node nd;
                    nd.x0 = (unsigned short)(int)(h_x * (float)i);
                    nd.x1 = (unsigned short)(int)(h_x * (float)(i + 1));
                    nd.y0 = 0;
                    nd.y1 = (unsigned short)hgt;
                    nd.cnt = sh.root_cnt[i];
                    nd.code = (unsigned)i;
                    cur[k] = nd;
                    sh.root_map[i] = k++;
                } else sh.root_map[i] = -1;
            }
            sh.lsz = k;
        }
        __syncthreads();
        for (int i = tid; i < 4 * sh.lsz; i += OT_THREADS) cc[i] = 0;
        __syncthreads();
        for (int p = tid; p < M; p += OT_THREADS) {
            const uint32_t v = pts[p];
            int r = (int)((float)(v & 0xFFF) / h_x);
            r = min(r, n_ini - 1);
            const int n = sh.root_map[r];
            pnode[p] = (unsigned short)n;
            const B2Node nd = cur[n];
            if (nd.cnt > 1) atomicAdd(&cc[4 * n + quadrant(nd, v & 0xFFF, (v >> 12) & 0xFFF)], 1u);
        }
        __syncthreads();
        ot_main_loop<false>(N, pts, pnode, M, cur, nxt, cc, ccn, cpos, newpos, rank, order, epre, flag, Q, sh, pyr, n_ini, D);
        __syncthreads();
    }
    OT_MARK(3);
    int lsz = sh.lsz;
    if (lsz < 0) {   // workspace guard tripped: report nothing rather than garbage
        if (tid == 0) sel_cnt[b * g.nlevels + level] = -1;
        return;
    }
    lsz = min(lsz, L.sel_cap);

    // best point per node (:741-760): max response, earliest in the reference's candidate order
    for (int i = tid; i < sh.lsz; i += OT_THREADS) best[i] = 0;
    if (use_pyr) {
        // pyramid of node ids: final node (depth d, path c) -> id + 1 at its own slot; every depth-D cell then walks its
        // code upwards until it meets one (the final nodes partition the area, so at most one lies on every path) and
        // keeps the id, so a point finds its node with one look-up
        const int pyr_words = (int)pyr_off(n_ini, D + 1);
        for (int i = tid; i < pyr_words; i += OT_THREADS) pyr[i] = 0;
        __syncthreads();
        for (int i = tid; i < sh.lsz; i += OT_THREADS) {
            const B2Node nd = cur[i];
            pyr[pyr_off(n_ini, (int)node_depth(nd)) + node_path(nd)] = (unsigned)i + 1u;
        }
        __syncthreads();
        unsigned* leaf = pyr + pyr_off(n_ini, D);
        for (int c = tid; c < (n_ini << (2 * D)); c += OT_THREADS) {
            if (leaf[c]) continue;                 // a depth-D node: only this thread touches the slot
            unsigned id = 0, code = (unsigned)c >> 2;
            for (int d = D - 1; d >= 0 && !id; --d, code >>= 2) id = pyr[pyr_off(n_ini, d) + code];   // read-only levels
            leaf[c] = id;
        }
        __syncthreads();
        for (int p0 = tid; p0 < M; p0 += OT_ILP_PYR * OT_THREADS) {
            uint32_t vv[OT_ILP_PYR];
#pragma unroll
            for (int k = 0; k < OT_ILP_PYR; ++k) vv[k] = p0 + k * OT_THREADS < M ? pts[p0 + k * OT_THREADS] : 0xFFFFFFFFu;
#pragma unroll
            for (int k = 0; k < OT_ILP_PYR; ++k) {
                if (p0 + k * OT_THREADS >= M) continue;
                const unsigned id = leaf[point_code(tab, vv[k], D)];
                if (id) atomicMax(&best[id - 1], best_key(vv[k], wc, hc, ncol, mwc, mhc));
            }
        }
    } else {
        __syncthreads();
        for (int p0 = tid; p0 < M; p0 += OT_ILP * OT_THREADS) {
            int nn[OT_ILP];
            uint32_t vv[OT_ILP];
#pragma unroll
            for (int k = 0; k < OT_ILP; ++k) {
                const int p = p0 + k * OT_THREADS;
                nn[k] = p < M ? (int)pnode[p] : -1;
                vv[k] = p < M ? pts[p] : 0u;
            }
#pragma unroll
            for (int k = 0; k < OT_ILP; ++k) {
                if (nn[k] < 0) continue;
                atomicMax(&best[nn[k]], best_key(vv[k], wc, hc, ncol, mwc, mhc));
            }
        }
    }
    __syncthreads();
    OT_MARK(4);
    for (int i = tid; i < lsz; i += OT_THREADS) {
        const unsigned v = best[i];
        unsigned key = 0x00FFFFFFu - (v & 0x00FFFFFFu);
        const int lx = key % wc; key /= wc;
        const int ly = key % hc; key /= hc;
        const int cj = key % ncol, ci = key / ncol;
        const unsigned x = cj * wc + lx + 3, y = ci * hc + ly + 3;
        out[i] = x | (y << 12) | (v & 0xFF000000u);
    }
    if (tid == 0) sel_cnt[b * g.nlevels + level] = lsz;
    OT_MARK(5);
#ifdef OT_TRACE
    if (tid == 0 && blockIdx.x == 0 && blockIdx.y == 0) { g_ot_trace[10] = M; g_ot_trace[11] = sh.lsz; g_ot_trace[12] = use_pyr; g_ot_trace[13] = D; }
#endif
}

}  // namespace

#ifdef OT_TRACE
extern "C" void b2orb_debug_octree_trace(long long* out) { cudaMemcpyFromSymbol(out, g_ot_trace, sizeof(long long) * 64); }
#endif

unsigned long long b2_octree_fallbacks() {
    unsigned long long v = 0;
    cudaMemcpyFromSymbol(&v, g_ot_fallbacks, sizeof v);
    return v;
}

int b2_octree_scratch_words(const B2Geom& g) {
    int q = 0;
    for (int l = 0; l < g.nlevels; ++l) q = q > g.lv[l].node_cap ? q : g.lv[l].node_cap;
    return B2_OT_WORDS * q;
}

void b2_launch_octree(const B2Geom& g, const uint32_t* cand, const int* cand_cnt, unsigned short* pnode,
                      unsigned* scratch, uint32_t* sel, int* sel_cnt, int nimg, int force_legacy, cudaStream_t s) {
    const size_t smem = (size_t)b2_octree_scratch_words(g) * sizeof(unsigned);
    const int use_global = scratch != nullptr;
    if (!use_global && smem > 24 * 1024)   // static (tables) + dynamic above 48 KB needs the opt-in; per-device
                                           // attribute, cheap enough to set on every launch
        cudaFuncSetAttribute(octree_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    dim3 grid(g.nlevels, nimg);
    b2_launch_pdl(octree_kernel, dim3(grid), dim3(OT_THREADS), use_global ? 0 : smem, s, g, cand, cand_cnt, pnode, scratch, use_global, sel,
                                                              sel_cnt, force_legacy);
}

// ---------------------------------------------------------------------------------------------------------------
// Allocation-free variant (b2orb_distribute_static, include/b2orb.h).  No counterpart in the reference: added
// for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops,
// harnesses that forbid it).  The quadtree's workspace — node lists, child counts / count pyramid, per-point node ids,
// and the candidate count the kernel reads — lives in ONE file-scope device array of fixed size instead of the
// handle's cudaMalloc'ed (or dynamic shared) workspace.  Every GPU that loads the library holds the array until exit.
// All calls share it and it keeps the last call's contents: the kernel initialises every word it reads (it has to,
// its normal workspace is uninitialised shared memory) and the count word is rewritten by every call; two calls that
// overlap on the GPU (different streams / handles) race on it.
constexpr int OT_STATIC_WS_WORDS = B2_OT_WORDS * B2_STATIC_MAX_NODES;                 // 26 words per node slot
constexpr int OT_STATIC_PNODE_WORDS = B2_STATIC_MAX_CANDIDATES / 2;                   // one u16 per candidate
constexpr int OT_STATIC_HDR_WORDS = 64;                                               // [0] = candidate count
constexpr int OT_STATIC_WORDS = OT_STATIC_HDR_WORDS + OT_STATIC_WS_WORDS + OT_STATIC_PNODE_WORDS;
static __device__ __align__(256) unsigned g_ot_static[OT_STATIC_WORDS];              // 0.95 MB per GPU

namespace {
// device code may name the array directly; host code needs cudaGetSymbolAddress (below)
__global__ void octree_static_prep_kernel(int n_cand) { g_ot_static[0] = (unsigned)n_cand; }
}  // namespace

// One level of one image: candidates (x | y << 12 | response << 24, relative to the 16-px border) -> selected keypoints
// in DistributeOctTree's list order.  g1 = one-level geometry.  Returns a cudaError_t (0 = ok).  Allocates nothing.
int b2_launch_octree_static(const B2Geom& g1, const uint32_t* d_cand, int n_cand, uint32_t* d_sel, int* d_sel_cnt, cudaStream_t s) {
    // In host code `g_ot_static` is a host-side stand-in, not a device address: resolve it once per device, with that
    // device current (the caller has made the handle's device current).
    static unsigned* s_base[64] = {};
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= 64) return (int)cudaErrorInvalidDevice;
    if (!s_base[dev]) {
        unsigned* ptr = nullptr;
        e = cudaGetSymbolAddress((void**)&ptr, g_ot_static);
        if (e != cudaSuccess) return (int)e;
        s_base[dev] = ptr;
    }
    unsigned* base = s_base[dev];
    octree_static_prep_kernel<<<1, 1, 0, s>>>(n_cand);
    b2_pdl_on = false;
    b2_launch_pdl(octree_kernel, dim3(1, 1), dim3(OT_THREADS), 0, s, g1, d_cand, reinterpret_cast<const int*>(base),
                  reinterpret_cast<unsigned short*>(base + OT_STATIC_HDR_WORDS + OT_STATIC_WS_WORDS),
                  base + OT_STATIC_HDR_WORDS, 1, d_sel, d_sel_cnt, 0);
    return (int)cudaGetLastError();
}
