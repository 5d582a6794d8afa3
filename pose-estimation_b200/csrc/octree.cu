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
#include "b2orb_internal.cuh"

namespace {

constexpr int OT_THREADS = 512;
constexpr int OT_MAX_ROOTS = 256;
constexpr int OT_ILP = 4;

struct OtShared {
    unsigned warp_sums[OT_THREADS / 32];
    unsigned total;
    int lsz, m_exp, t_children, n_multi, n_cand;
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
    return c;
}

// One expansion round.  partial == false: outer pass; true: sorted partial pass with early stop at N.
// cc holds the child point counts of every multi-point node of `cur` (filled by the previous round's point pass, or
// by the root pass); ccn receives the same for the new list while the points are moved: ONE pass over the points
// per round.
__device__ void ot_round(bool partial, int N, const uint32_t* __restrict__ pts, unsigned short* __restrict__ pnode,
                         int M, B2Node* cur, B2Node* nxt, const unsigned* cc, unsigned* ccn, int* cpos, int* newpos,
                         unsigned* rank, unsigned* order, unsigned* epre, unsigned* flag, int node_cap, OtShared& sh) {
    const int tid = threadIdx.x;
    const int lsz = sh.lsz;
    // 1. candidates = multi-point nodes
    for (int i = tid; i < lsz; i += OT_THREADS) flag[i] = cur[i].cnt > 1 ? 1u : 0u;
    __syncthreads();
    // 3. expansion order: rank r of candidate i, order[r] = i
    //    flag -> compact candidate index (list order)
    const int n_cand = (int)block_scan_excl(flag, lsz, sh);
    for (int i = tid; i < lsz; i += OT_THREADS)
        if (cur[i].cnt > 1) epre[flag[i]] = i;          // epre temporarily = candidate list
    __syncthreads();
    if (!partial) {
        for (int c = tid; c < n_cand; c += OT_THREADS) { const int i = epre[c]; rank[i] = c; order[c] = i; }
    } else {
        for (int c = tid; c < n_cand; c += OT_THREADS) {
            const int i = epre[c];
            const unsigned ci = cur[i].cnt;
            unsigned r = 0;
            for (int d = 0; d < n_cand; ++d) {
                const int j = epre[d];
                const unsigned cj = cur[j].cnt;
                r += (cj > ci || (cj == ci && j < i)) ? 1u : 0u;
            }
            rank[i] = r;
            order[r] = i;
        }
    }
    __syncthreads();
    // 4. e_r = non-empty children of the r-th expansion; inclusive prefix; early-stop index m
    for (int r = tid; r < n_cand; r += OT_THREADS) {
        const int i = order[r];
        epre[r] = (cc[4 * i] != 0) + (cc[4 * i + 1] != 0) + (cc[4 * i + 2] != 0) + (cc[4 * i + 3] != 0);
    }
    __syncthreads();
    block_scan_excl(epre, n_cand, sh);   // epre[r] = sum_{t<r} e_t
    if (tid == 0) { sh.m_exp = n_cand; sh.n_multi = 0; }
    __syncthreads();
    if (partial) {
        // size after expanding ranks 0..r  =  lsz + sum_{t<=r} (e_t - 1); take the first r reaching N
        for (int r = tid; r < n_cand; r += OT_THREADS) {
            const int i = order[r];
            const int e_r = (cc[4 * i] != 0) + (cc[4 * i + 1] != 0) + (cc[4 * i + 2] != 0) + (cc[4 * i + 3] != 0);
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
            tch = (int)epre[m_exp - 1] + (cc[4 * i] != 0) + (cc[4 * i + 1] != 0) + (cc[4 * i + 2] != 0) +
                  (cc[4 * i + 3] != 0);
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
            const unsigned c0 = cc[4 * i], c1 = cc[4 * i + 1], c2 = cc[4 * i + 2], c3 = cc[4 * i + 3];
            const int e_r = (c0 != 0) + (c1 != 0) + (c2 != 0) + (c3 != 0);
            // children pushed n1..n4 to the front -> n4 first; block of this expansion starts at
            // t_children - (epre[r] + e_r)
            int pos = t_children - ((int)epre[r] + e_r);
            const unsigned cs[4] = {c0, c1, c2, c3};
#pragma unroll
            for (int q = 3; q >= 0; --q) {
                if (cs[q]) {
                    nxt[pos] = child_rect(nd, q, cs[q]);
                    ccn[4 * pos] = ccn[4 * pos + 1] = ccn[4 * pos + 2] = ccn[4 * pos + 3] = 0;
                    cpos[4 * i + q] = pos;
                    multi += cs[q] > 1;
                    ++pos;
                } else cpos[4 * i + q] = -1;
            }
            newpos[i] = -1;
        } else {
            const int pos = t_children + (int)flag[i];
            nxt[pos] = nd;
            ccn[4 * pos] = ccn[4 * pos + 1] = ccn[4 * pos + 2] = ccn[4 * pos + 3] = 0;
            newpos[i] = pos;
            multi += nd.cnt > 1;
        }
    }
    if (multi) atomicAdd(&sh.n_multi, multi);
    __syncthreads();
    // 7. move the points and count them into the children of their new node (next round's cc)
    // (point loops are batched OT_ILP deep: the pass is a chain of dependent L2 loads, so the loads of several
    //  points are issued before any of them is used)
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

__global__ void __launch_bounds__(OT_THREADS)
octree_kernel(const __grid_constant__ B2Geom g, const uint32_t* __restrict__ cand, const int* __restrict__ cand_cnt,
              unsigned short* __restrict__ pnode_all, unsigned* __restrict__ gscratch, int use_global,
              uint32_t* __restrict__ sel, int* __restrict__ sel_cnt) {
    extern __shared__ __align__(16) unsigned dyn[];
    __shared__ OtShared sh;
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
    // workspace carve-up (B2_OT_WORDS*Q words): two node arrays (3 words each), two child-count arrays, scratch
    unsigned* ws = use_global ? gscratch + ((size_t)b * g.node_total + L.node_off) * B2_OT_WORDS : dyn;
    B2Node* cur = reinterpret_cast<B2Node*>(ws);
    B2Node* nxt = reinterpret_cast<B2Node*>(ws + 3 * Q);
    unsigned* cc = ws + 6 * Q;
    unsigned* ccn = ws + 10 * Q;
    int* cpos = reinterpret_cast<int*>(ws + 14 * Q);
    int* newpos = reinterpret_cast<int*>(ws + 18 * Q);
    unsigned* rank = ws + 19 * Q;
    unsigned* order = ws + 20 * Q;
    unsigned* epre = ws + 21 * Q;
    unsigned* flag = ws + 22 * Q;
    unsigned* best = ws + 23 * Q;

    // roots (:543-590): nIni = round(W'/H'), hX = W'/nIni; point -> root (int)(x / hX); empty roots dropped
    const int n_ini = L.n_ini;
    const float h_x = L.h_x;
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
        const int hgt = L.h - 2 * B2_BORDER;
        for (int i = 0; i < n_ini; ++i) {
            if (sh.root_cnt[i]) {
                B2Node nd;
                nd.x0 = (unsigned short)(int)(h_x * (float)i);
                nd.x1 = (unsigned short)(int)(h_x * (float)(i + 1));
                nd.y0 = 0;
                nd.y1 = (unsigned short)hgt;
                nd.cnt = sh.root_cnt[i];
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

    // main loop (:592-739)
    bool finish = false;
    while (!finish) {
        int prev = sh.lsz;
        ot_round(false, N, pts, pnode, M, cur, nxt, cc, ccn, cpos, newpos, rank, order, epre, flag, Q, sh);
        { B2Node* t = cur; cur = nxt; nxt = t; unsigned* u = cc; cc = ccn; ccn = u; }
        int lsz = sh.lsz;
        if (lsz < 0) break;
        if (lsz >= N || lsz == prev) finish = true;
        else if (lsz + 3 * sh.n_multi > N) {
            while (!finish) {
                prev = sh.lsz;
                ot_round(true, N, pts, pnode, M, cur, nxt, cc, ccn, cpos, newpos, rank, order, epre, flag, Q, sh);
                { B2Node* t = cur; cur = nxt; nxt = t; unsigned* u = cc; cc = ccn; ccn = u; }
                lsz = sh.lsz;
                if (lsz < 0 || lsz >= N || lsz == prev) finish = true;
            }
        }
    }
    int lsz = sh.lsz;
    if (lsz < 0) {   // workspace guard tripped: report nothing rather than garbage
        if (tid == 0) sel_cnt[b * g.nlevels + level] = -1;
        return;
    }
    lsz = min(lsz, L.sel_cap);

    // best point per node (:741-760): max response, earliest in the reference's candidate order
    for (int i = tid; i < sh.lsz; i += OT_THREADS) best[i] = 0;
    __syncthreads();
    const int wc = L.w_cell, hc = L.h_cell, ncol = L.n_cols;
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
            const uint32_t v = vv[k];
            const int x = (int)(v & 0xFFF) - 3, y = (int)((v >> 12) & 0xFFF) - 3;
            const int cj = x / wc, ci = y / hc;
            const unsigned key = (unsigned)(((ci * ncol + cj) * hc + (y - ci * hc)) * wc + (x - cj * wc));
            atomicMax(&best[nn[k]], (v & 0xFF000000u) | (0x00FFFFFFu - key));
        }
    }
    __syncthreads();
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
}

}  // namespace

int b2_octree_scratch_words(const B2Geom& g) {
    int q = 0;
    for (int l = 0; l < g.nlevels; ++l) q = q > g.lv[l].node_cap ? q : g.lv[l].node_cap;
    return B2_OT_WORDS * q;
}

void b2_launch_octree(const B2Geom& g, const uint32_t* cand, const int* cand_cnt, unsigned short* pnode,
                      B2Node* nodes, unsigned* scratch, uint32_t* sel, int* sel_cnt, int nimg, cudaStream_t s) {
    (void)nodes;
    const size_t smem = (size_t)b2_octree_scratch_words(g) * sizeof(unsigned);
    const int use_global = scratch != nullptr;
    if (!use_global && smem > 48 * 1024)   // per-device attribute; cheap enough to set on every launch
        cudaFuncSetAttribute(octree_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    dim3 grid(g.nlevels, nimg);
    octree_kernel<<<grid, OT_THREADS, use_global ? 0 : smem, s>>>(g, cand, cand_cnt, pnode, scratch, use_global, sel,
                                                              sel_cnt);
}
