// Next row f3 — the feature pool of cfsd::FeatureTracker kept RESIDENT ON THE DEVICE.
//
// Reference: externalTrack matches the kept stereo descriptors of the current frame against _histDescriptorsL/R
// (src/feature-tracker.cpp:258, :303), cv::Mats that featurePoolUpdate rebuilds on every keyframe from the pool
// `_features` (:345-363 ageing + erase + rebuild, :393-395 append of the un-masked new features).  On the host path
// those two matrices (64 B per pooled feature) would have to be uploaded for every frame; here they never leave the
// GPU: the pool's descriptor rows live in two device arrays (double-buffered), a keyframe update is ONE gather kernel
// driven by a small index list, and the matcher reads them in place.
//
// What stays on the host is the bookkeeping the reference also does on the host — and it has to be done with the same
// container: `_features` is a std::unordered_map<size_t, Ptr<Feature>> (include/cfsd/feature-tracker.hpp:129) and the
// row ORDER of _histDescriptors after a keyframe is that map's iteration order (:349-363).  The order is observable
// (trainIdx, BFMatcher's lowest-index tie-break, _histFeatureIDs), so this file mirrors the map with the same key type
// and the same sequence of operator[] / erase / insert calls, and derives the gather list from its iteration.
#include <unordered_map>
#include <vector>

#include "b2orb_internal.cuh"

namespace {

struct PoolEntry {
    int age = 0;
    int row = -1;   // row of this feature in the CURRENT device arrays
};

// new row i <- old row idx[i] (i < n_keep)  |  current-frame row idx[i] (otherwise); two threads per row, 16 B each per eye
__global__ void __launch_bounds__(256)
pool_gather_kernel(const uint4* __restrict__ old_l, const uint4* __restrict__ old_r, const uint4* __restrict__ cur_l,
                   const uint4* __restrict__ cur_r, const int* __restrict__ idx, int n_keep, int n_total,
                   uint4* __restrict__ new_l, uint4* __restrict__ new_r, int* __restrict__ d_count) {
    const int t = blockIdx.x * 256 + threadIdx.x;
    if (t == 0) *d_count = n_total;
    const int row = t >> 1, half = t & 1;
    if (row >= n_total) return;
    const int src = idx[row];
    const uint4* sl = row < n_keep ? old_l : cur_l;
    const uint4* sr = row < n_keep ? old_r : cur_r;
    new_l[2 * row + half] = sl[2 * src + half];
    new_r[2 * row + half] = sr[2 * src + half];
}

}  // namespace

struct B2Pool {
    std::unordered_map<size_t, PoolEntry> features;   // _features (ages; descriptors are on the device)
    std::vector<unsigned long long> hist_ids;         // _histFeatureIDs
    size_t feature_id = 0;                            // _featureID
    uint8_t* d_rows[2][2] = {{nullptr, nullptr}, {nullptr, nullptr}};   // [buffer][eye]
    int cur = 0, cap = 0;
    int* d_count = nullptr;
    int* d_idx = nullptr;                             // gather list (device)
    int* h_idx = nullptr;                             // gather list (pinned host)
    int idx_cap = 0;
};

B2Pool* b2_pool_create() {
    B2Pool* p = new B2Pool();
    if (cudaMalloc(&p->d_count, sizeof(int)) != cudaSuccess) { delete p; return nullptr; }
    cudaMemset(p->d_count, 0, sizeof(int));
    return p;
}

void b2_pool_destroy(B2Pool* p) {
    if (!p) return;
    for (int b = 0; b < 2; ++b)
        for (int e = 0; e < 2; ++e)
            if (p->d_rows[b][e]) cudaFree(p->d_rows[b][e]);
    if (p->d_count) cudaFree(p->d_count);
    if (p->d_idx) cudaFree(p->d_idx);
    if (p->h_idx) cudaFreeHost(p->h_idx);
    delete p;
}

void b2_pool_reset(B2Pool* p) {
    std::unordered_map<size_t, PoolEntry>().swap(p->features);   // a FRESH container: the bucket history restarts too
    p->hist_ids.clear();
    p->feature_id = 0;
    cudaMemset(p->d_count, 0, sizeof(int));
}

int b2_pool_size(const B2Pool* p) { return (int)p->hist_ids.size(); }
bool b2_pool_empty(const B2Pool* p) { return p->features.empty(); }
const uint8_t* b2_pool_rows(const B2Pool* p, int eye) { return p->d_rows[p->cur][eye]; }
const int* b2_pool_dcount(const B2Pool* p) { return p->d_count; }
const unsigned long long* b2_pool_ids(const B2Pool* p) { return p->hist_ids.data(); }
int b2_pool_age(const B2Pool* p, unsigned long long id) {
    auto it = p->features.find((size_t)id);
    return it == p->features.end() ? -1 : it->second.age;
}

// FeatureTracker::featurePoolUpdate, src/feature-tracker.cpp:330-412 (without the triangulated positions).
//   matched_ids        _matchedFeatureIDs of the last externalTrack (:315-317)
//   mask[n_cur]        _curFeatureMask; accept[n_cur] (optional) = !(depth > maxDepth) of :374
//   d_cur_l / d_cur_r  device rows of _curDescriptorsL/R
// Returns 0 or a cudaError_t.
int b2_pool_update(B2Pool* p, cudaStream_t s, bool keyframe, int max_age, const unsigned long long* matched_ids, int n_matched,
                   const uint8_t* mask, const uint8_t* accept, int n_cur, const uint8_t* d_cur_l, const uint8_t* d_cur_r,
                   unsigned long long* new_ids, int* n_inserted, int* n_erased) {
    *n_inserted = *n_erased = 0;
    if (!(p->features.empty() || keyframe)) return 0;                      // :337
    for (int i = 0; i < n_matched; ++i) {                                  // :341-343
        auto it = p->features.find((size_t)matched_ids[i]);
        if (it != p->features.end()) it->second.age -= 1;                  // (the reference would dereference a null Ptr)
    }
    const int need = (int)p->features.size() + n_cur;
    if (need > p->idx_cap) {
        const int c = need * 2 + 1024;
        cudaStreamSynchronize(s);
        if (p->d_idx) cudaFree(p->d_idx);
        if (p->h_idx) cudaFreeHost(p->h_idx);
        p->d_idx = p->h_idx = nullptr;
        p->idx_cap = 0;
        cudaError_t e = cudaMalloc(&p->d_idx, (size_t)c * sizeof(int));
        if (e == cudaSuccess) e = cudaHostAlloc(&p->h_idx, (size_t)c * sizeof(int), cudaHostAllocDefault);
        if (e != cudaSuccess) return (int)e;
        p->idx_cap = c;
    }
    // ageing, erase, rebuild in the container's iteration order (:349-363)
    std::vector<unsigned long long> ids;
    ids.reserve(need);
    int n_keep = 0;
    auto f = p->features.begin();
    while (f != p->features.end()) {
        f->second.age += 2;
        if (f->second.age > max_age) {
            f = p->features.erase(f);
            ++*n_erased;
        } else {
            ids.push_back(f->first);
            p->h_idx[n_keep] = f->second.row;
            f->second.row = n_keep++;
            ++f;
        }
    }
    // new features (:369-406)
    int n_total = n_keep;
    for (int i = 0; i < n_cur; ++i) {
        if (!mask[i] || (accept && !accept[i])) continue;
        PoolEntry e;
        e.age = 0;
        e.row = n_total;
        p->features[p->feature_id] = e;
        ids.push_back(p->feature_id);
        if (new_ids) new_ids[*n_inserted] = p->feature_id;
        p->h_idx[n_total++] = i;
        ++p->feature_id;
        ++*n_inserted;
    }
    p->hist_ids.swap(ids);
    // device side: gather into the other buffer pair
    if (n_total > p->cap) {   // grow both pairs; the current one is the gather source, its content moves along
        const int c = n_total * 2 + 4096;
        cudaError_t err = cudaDeviceSynchronize();   // rare: nothing may still be reading the old arrays
        if (err != cudaSuccess) return (int)err;
        for (int b = 0; b < 2; ++b)
            for (int e = 0; e < 2; ++e) {
                uint8_t* q = nullptr;
                err = cudaMalloc(&q, (size_t)c * 32);
                if (err != cudaSuccess) return (int)err;
                if (p->d_rows[b][e]) {
                    if (b == p->cur) cudaMemcpy(q, p->d_rows[b][e], (size_t)p->cap * 32, cudaMemcpyDeviceToDevice);
                    cudaFree(p->d_rows[b][e]);
                }
                p->d_rows[b][e] = q;
            }
        p->cap = c;
    }
    const int nb = p->cur ^ 1;
    if (n_total > 0) {
        cudaError_t err = cudaMemcpyAsync(p->d_idx, p->h_idx, (size_t)n_total * sizeof(int), cudaMemcpyHostToDevice, s);
        if (err != cudaSuccess) return (int)err;
        pool_gather_kernel<<<(2 * n_total + 255) / 256, 256, 0, s>>>(
            reinterpret_cast<const uint4*>(p->d_rows[p->cur][0]), reinterpret_cast<const uint4*>(p->d_rows[p->cur][1]),
            reinterpret_cast<const uint4*>(d_cur_l), reinterpret_cast<const uint4*>(d_cur_r), p->d_idx, n_keep, n_total,
            reinterpret_cast<uint4*>(p->d_rows[nb][0]), reinterpret_cast<uint4*>(p->d_rows[nb][1]), p->d_count);
        err = cudaGetLastError();
        if (err != cudaSuccess) return (int)err;
        err = cudaStreamSynchronize(s);   // h_idx is reused by the next update
        if (err != cudaSuccess) return (int)err;
    } else {
        cudaMemsetAsync(p->d_count, 0, sizeof(int), s);
    }
    p->cur = nb;
    return 0;
}
