// Gradient-boosted-tree training kernels for the classical-ML library (hetuml GBDT, device='gpu').
//
// reference: third_party/HetuML/hetuml/cc/ml/include/model/gbdt/{hist/histogram.h,helper/hist_manager.h,
// helper/node_indexer.h} - host code that builds one (gradient, hessian) histogram per tree node
// and re-indexes instances after every split.  Here a whole tree LEVEL is one launch: every
// instance carries the id of the node it currently sits in, one warp walks the stored entries of
// one instance and adds its (g, h) into the histogram of that node with red.global.add.f32.  Only
// stored (non-zero) entries are visited; the mass of the zero-valued rows of a feature is
// recovered afterwards as "node total minus visited mass" (done by the caller on the device).
#include "common.cuh"

// hist: [num_slots, total_bins, 2]; slot_of_node[node] = histogram slot of an active node, -1 otherwise
__global__ void gbdt_hist_kernel(const long long* __restrict__ indptr, const int* __restrict__ indices,
                                 const unsigned char* __restrict__ bins, const int* __restrict__ node_of,
                                 const int* __restrict__ slot_of_node, const float* __restrict__ g,
                                 const float* __restrict__ h, const long long* __restrict__ bin_off,
                                 const unsigned char* __restrict__ feat_on, float* __restrict__ hist,
                                 long long total_bins, long long n) {
    long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    int lane = threadIdx.x & 31;
    for (long long i = warp; i < n; i += nwarps) {
        int nd = node_of[i];
        if (nd < 0) continue;
        int slot = slot_of_node[nd];
        if (slot < 0) continue;
        float gi = g[i], hi = h[i];
        if (gi == 0.f && hi == 0.f) continue;                 // instance not sampled for this tree
        float* base = hist + (size_t)slot * (size_t)total_bins * 2;
        long long end = indptr[i + 1];
        for (long long q = indptr[i] + lane; q < end; q += 32) {
            int f = indices[q];
            if (!feat_on[f]) continue;
            float* p = base + (bin_off[f] + (long long)bins[q]) * 2;
            atomicAdd(p, gi);
            atomicAdd(p + 1, hi);
        }
    }
}

// After a level's splits are chosen: move every instance of a split node to its left / right child.
// split_feat[node] < 0 = node was not split.  Rows are sorted by column, so the feature is found by
// binary search; an absent entry means the value 0 (bin zero_bin[f]).
__global__ void gbdt_route_kernel(const long long* __restrict__ indptr, const int* __restrict__ indices,
                                  const unsigned char* __restrict__ bins, int* __restrict__ node_of,
                                  const int* __restrict__ split_feat, const int* __restrict__ split_bin,
                                  const int* __restrict__ zero_bin, const int* __restrict__ left_child, long long n) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int nd = node_of[i];
    if (nd < 0) return;
    int f = split_feat[nd];
    if (f < 0) return;
    long long lo = indptr[i], hi = indptr[i + 1];
    int b = zero_bin[f];
    while (lo < hi) {
        long long mid = (lo + hi) >> 1;
        int c = indices[mid];
        if (c == f) { b = bins[mid]; break; }
        if (c < f) lo = mid + 1; else hi = mid;
    }
    node_of[i] = b <= split_bin[nd] ? left_child[nd] : left_child[nd] + 1;
}

HETU_API int hetu_gbdt_hist(const long long* indptr, const int* indices, const unsigned char* bins, const int* node_of,
                            const int* slot_of_node, const float* g, const float* h, const long long* bin_off,
                            const unsigned char* feat_on, float* hist, long long total_bins, long long n, cudaStream_t s) {
    if (n <= 0) return 0;
    int threads = 256;
    long long warps_needed = n;
    long long blocks = (warps_needed * 32 + threads - 1) / threads;
    long long cap = (long long)kNumSMs * 8;
    gbdt_hist_kernel<<<(unsigned)(blocks < cap ? blocks : cap), threads, 0, s>>>(indptr, indices, bins, node_of, slot_of_node, g, h,
                                                                                 bin_off, feat_on, hist, total_bins, n);
    HETU_CHECK_LAUNCH("gbdt_hist");
    return 0;
}

HETU_API int hetu_gbdt_route(const long long* indptr, const int* indices, const unsigned char* bins, int* node_of,
                             const int* split_feat, const int* split_bin, const int* zero_bin, const int* left_child,
                             long long n, cudaStream_t s) {
    if (n <= 0) return 0;
    int threads = 256;
    gbdt_route_kernel<<<(unsigned)((n + threads - 1) / threads), threads, 0, s>>>(indptr, indices, bins, node_of, split_feat, split_bin,
                                                                                  zero_bin, left_child, n);
    HETU_CHECK_LAUNCH("gbdt_route");
    return 0;
}
