// HET embedding cache resident in HBM (GPU tier in front of the host-DRAM parameter server).
//
// reference: src/hetu_cache/include/embedding.h:18-149 (Line: data, lazily kept gradient, version, update count),
// src/hetu_cache/src/cache.cc:59-106,131-196,217-283 (lookup with pull_bound, update with push_bound, eviction),
// src/hetu_cache/src/lfuopt_cache.cc:5-63 (LFUOpt: lines used often enough become permanent).  The reference keeps the
// cache in host memory: every lookup walks std::unordered_map on the CPU and copies the rows to the GPU afterwards.
//
// Here the lines live in HBM in an 8-way set-associative table: a warp probes one set with one 8-lane compare, rows are
// served by 128-bit copies at HBM speed, local updates are applied to the cached row AND kept in its gradient row until
// the push bound is passed.  The host only sees what has to cross to the parameter server: ids + cached versions go up
// (staleness check against pull_bound), missing / stale rows come down, dirty evictions and rows over push_bound go up
// (pinned staging, cudaMemcpyAsync, one per copy - parallel/gpu_cache.py).
#include "common.cuh"

namespace {
constexpr int kWays = 8;
constexpr unsigned kPermanent = 16;          // LFUOpt: use count from which a line is never evicted (kUseCntMax upstream)
enum { POLICY_LRU = 0, POLICY_LFU = 1, POLICY_LFUOPT = 2 };

struct HetCache {
    long long* tags;      // [nsets * 8], -1 = empty
    float* data;          // [nsets * 8, width]
    float* grad;          // [nsets * 8, width]: local updates not yet pushed (already scaled by -lr)
    long long* version;   // server version of the row when it was fetched (+ own pushed updates)
    int* updates;         // local updates since the last push
    unsigned* score;      // LRU: tick of the last use; LFU / LFUOpt: use count
    int* queued;          // 1 while the line sits on the push list
    int* lock;            // per set
    int nsets, width, policy;
};

__device__ __forceinline__ int set_of(long long id, int nsets) {
    unsigned long long h = (unsigned long long)id * 0x9E3779B97F4A7C15ull;
    return (int)((h >> 24) % (unsigned long long)nsets);
}
// way holding `id` in `set` (all lanes get the answer), -1 if absent
__device__ __forceinline__ int find_way(const HetCache& c, int set, long long id, int lane) {
    const long long t = lane < kWays ? *reinterpret_cast<volatile long long*>(c.tags + (size_t)set * kWays + lane) : -2;
    const unsigned m = __ballot_sync(0xffffffffu, t == id);
    return m ? __ffs(m) - 1 : -1;
}

// slot[i] = line holding ids[i] or -1; ver[i] = its version or -1 (what the staleness check sends to the server)
__global__ void __launch_bounds__(256) het_probe_kernel(HetCache c, const long long* __restrict__ ids, size_t n, int* __restrict__ slot,
                                                        long long* __restrict__ ver, unsigned tick, int touch,
                                                        unsigned long long* __restrict__ stats) {
    size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t i = warp; i < n; i += nwarp) {
        const long long id = ids[i];
        int s = -1;
        if (id >= 0) {
            const int set = set_of(id, c.nsets);
            const int way = find_way(c, set, id, lane);
            if (way >= 0) s = set * kWays + way;
        }
        if (lane == 0) {
            slot[i] = s;
            if (ver != nullptr) ver[i] = s >= 0 ? c.version[s] : -1;
            if (touch && s >= 0) {
                if (c.policy == POLICY_LRU) c.score[s] = tick;
                else if (c.score[s] < 0x7fffffffu) atomicAdd(c.score + s, 1u);
            }
            if (touch && stats != nullptr) { atomicAdd(stats, 1ull); if (s < 0) atomicAdd(stats + 1, 1ull); }
        }
    }
}

// Install (miss) or refresh (stale hit) rows that came down from the server.  One warp per row, the set is locked while
// its victim is chosen.  A dirty victim leaves through the eviction list (id, gradient row, update count).
__global__ void __launch_bounds__(256) het_install_kernel(HetCache c, const long long* __restrict__ ids, const long long* __restrict__ vers,
                                                          const float* __restrict__ rows, size_t m, unsigned tick,
                                                          long long* __restrict__ ev_ids, float* __restrict__ ev_grads,
                                                          long long* __restrict__ ev_upd, unsigned* __restrict__ ev_count, int ev_cap,
                                                          unsigned long long* __restrict__ stats) {
    size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    const int w = c.width;
    for (size_t k = warp; k < m; k += nwarp) {
        const long long id = ids[k];
        if (id < 0) continue;
        const int set = set_of(id, c.nsets);
        if (lane == 0) { while (atomicCAS(c.lock + set, 0, 1) != 0) __nanosleep(20); }
        __syncwarp();
        __threadfence();
        int way = find_way(c, set, id, lane);
        const float* src = rows + k * (size_t)w;
        if (way >= 0) {                                   // refresh: server row + the local gradient that is still here
            const size_t s = (size_t)set * kWays + way;
            for (int j = lane; j < w; j += 32) c.data[s * w + j] = src[j] + c.grad[s * w + j];
            if (lane == 0) c.version[s] = vers[k];
        } else {
            // victim: an empty way, else the smallest score; LFUOpt never evicts permanent lines
            unsigned best = 0xffffffffu;
            int bway = -1;
            if (lane < kWays) {
                const size_t s = (size_t)set * kWays + lane;
                const long long t = *reinterpret_cast<volatile long long*>(c.tags + s);
                const unsigned cur = *reinterpret_cast<volatile unsigned*>(c.score + s);
                unsigned sc = t < 0 ? 0u : (cur < 0xfffffffeu ? cur + 1u : cur);
                if (c.policy == POLICY_LFUOPT && t >= 0 && cur >= kPermanent) sc = 0xffffffffu;
                best = sc;
                bway = lane;
            }
#pragma unroll
            for (int o = 4; o > 0; o >>= 1) {
                const unsigned ob = __shfl_xor_sync(0xffffffffu, best, o);
                const int ow = __shfl_xor_sync(0xffffffffu, bway, o);
                if (ob < best || (ob == best && ow >= 0 && (bway < 0 || ow < bway))) { best = ob; bway = ow; }
            }
            best = __shfl_sync(0xffffffffu, best, 0);
            way = __shfl_sync(0xffffffffu, bway, 0);
            if (best != 0xffffffffu && way >= 0) {
                const size_t s = (size_t)set * kWays + way;
                const long long old = *reinterpret_cast<volatile long long*>(c.tags + s);
                int slot_out = -1;
                if (old >= 0 && *reinterpret_cast<volatile int*>(c.updates + s) > 0) {       // dirty eviction
                    if (lane == 0) slot_out = (int)atomicAdd(ev_count, 1u);
                    slot_out = __shfl_sync(0xffffffffu, slot_out, 0);
                    if (slot_out < ev_cap) {
                        for (int j = lane; j < w; j += 32) ev_grads[(size_t)slot_out * w + j] = c.grad[s * w + j];
                        if (lane == 0) { ev_ids[slot_out] = old; ev_upd[slot_out] = c.updates[s]; }
                    }
                }
                for (int j = lane; j < w; j += 32) { c.data[s * w + j] = src[j]; c.grad[s * w + j] = 0.f; }
                if (lane == 0) {
                    if (old >= 0 && stats != nullptr) atomicAdd(stats + 2, 1ull);
                    c.tags[s] = id; c.version[s] = vers[k]; c.updates[s] = 0; c.queued[s] = 0;
                    c.score[s] = c.policy == POLICY_LRU ? tick : 1u;
                }
            }                                             // else: every way is permanent - the row is served uncached
        }
        __threadfence();
        __syncwarp();
        if (lane == 0) atomicExch(c.lock + set, 0);
    }
}

// out[i] = cached row, or (line absent: set full of permanent lines / more misses than ways in one batch) the freshly
// fetched row fallback[fb_index[i]], or zeros
__global__ void __launch_bounds__(256) het_gather_kernel(HetCache c, const int* __restrict__ slot, size_t n, float* __restrict__ out,
                                                         const float* __restrict__ fallback, const int* __restrict__ fb_index) {
    size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    const int w = c.width;
    for (size_t i = warp; i < n; i += nwarp) {
        const int s = slot[i];
        const float* src = s >= 0 ? c.data + (size_t)s * w
                                  : (fb_index != nullptr && fb_index[i] >= 0 ? fallback + (size_t)fb_index[i] * w : nullptr);
        for (int j = lane; j < w; j += 32) out[i * w + j] = src != nullptr ? src[j] : 0.f;
    }
}

// delta rows (already scaled by -lr) are added to the cached row and to its kept gradient; lines that pass push_bound go
// on the push list once; rows that are not cached go on the direct list (position in the batch)
__global__ void __launch_bounds__(256) het_update_kernel(HetCache c, const int* __restrict__ slot, const float* __restrict__ delta, size_t n,
                                                         int push_bound, int* __restrict__ push_slots, unsigned* __restrict__ push_count,
                                                         int* __restrict__ direct, unsigned* __restrict__ direct_count) {
    size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    const int w = c.width;
    for (size_t i = warp; i < n; i += nwarp) {
        const int s = slot[i];
        if (s < 0) {
            if (lane == 0) direct[atomicAdd(direct_count, 1u)] = (int)i;
            continue;
        }
        for (int j = lane; j < w; j += 32) {
            const float d = delta[i * w + j];
            atomicAdd(c.data + (size_t)s * w + j, d);
            atomicAdd(c.grad + (size_t)s * w + j, d);
        }
        if (lane == 0) {
            const int u = atomicAdd(c.updates + s, 1) + 1;
            if (u > push_bound && atomicExch(c.queued + s, 1) == 0) push_slots[atomicAdd(push_count, 1u)] = s;
        }
    }
}

// hand the listed lines' gradients over (ids, rows, update counts) and reset them; all_dirty != 0: every dirty line
__global__ void __launch_bounds__(256) het_drain_kernel(HetCache c, const int* __restrict__ push_slots, const unsigned* __restrict__ push_count,
                                                        int all_dirty, long long* __restrict__ out_ids, float* __restrict__ out_grads,
                                                        long long* __restrict__ out_upd, unsigned* __restrict__ out_count, int out_cap) {
    size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    const int w = c.width;
    const size_t total = all_dirty ? (size_t)c.nsets * kWays : (size_t)*push_count;
    for (size_t k = warp; k < total; k += nwarp) {
        const size_t s = all_dirty ? k : (size_t)push_slots[k];
        if (c.tags[s] < 0 || c.updates[s] <= 0) continue;
        int pos = 0;
        if (lane == 0) pos = (int)atomicAdd(out_count, 1u);
        pos = __shfl_sync(0xffffffffu, pos, 0);
        if (pos >= out_cap) continue;
        for (int j = lane; j < w; j += 32) { out_grads[(size_t)pos * w + j] = c.grad[s * w + j]; c.grad[s * w + j] = 0.f; }
        if (lane == 0) {
            out_ids[pos] = c.tags[s];
            out_upd[pos] = c.updates[s];
            if (c.version[s] >= 0) c.version[s] += c.updates[s];
            c.updates[s] = 0;
            c.queued[s] = 0;
        }
    }
}

HetCache make(void* tags, float* data, float* grad, void* version, void* updates, void* score, void* queued, void* lock, int nsets,
              int width, int policy) {
    HetCache c;
    c.tags = (long long*)tags; c.data = data; c.grad = grad; c.version = (long long*)version; c.updates = (int*)updates;
    c.score = (unsigned*)score; c.queued = (int*)queued; c.lock = (int*)lock; c.nsets = nsets; c.width = width; c.policy = policy;
    return c;
}
}  // namespace

#define HET_ARGS void* tags, float* data, float* grad, void* version, void* updates, void* score, void* queued, void* lock, int nsets, int width, int policy
#define HET_MAKE make(tags, data, grad, version, updates, score, queued, lock, nsets, width, policy)

HETU_API int hetu_het_probe(HET_ARGS, const void* ids, size_t n, int* slot, void* ver, unsigned tick, int touch, void* stats,
                            cudaStream_t s) {
    if (n == 0) return 0;
    het_probe_kernel<<<ew_grid(n * 32, 256, 16), 256, 0, s>>>(HET_MAKE, (const long long*)ids, n, slot, (long long*)ver, tick, touch,
                                                             (unsigned long long*)stats);
    HETU_CHECK_LAUNCH("hetu_het_probe");
    return 0;
}
HETU_API int hetu_het_install(HET_ARGS, const void* ids, const void* vers, const float* rows, size_t m, unsigned tick, void* ev_ids,
                              float* ev_grads, void* ev_upd, void* ev_count, int ev_cap, void* stats, cudaStream_t s) {
    if (m == 0) return 0;
    het_install_kernel<<<ew_grid(m * 32, 256, 16), 256, 0, s>>>(HET_MAKE, (const long long*)ids, (const long long*)vers, rows, m, tick,
                                                               (long long*)ev_ids, ev_grads, (long long*)ev_upd, (unsigned*)ev_count,
                                                               ev_cap, (unsigned long long*)stats);
    HETU_CHECK_LAUNCH("hetu_het_install");
    return 0;
}
HETU_API int hetu_het_gather(HET_ARGS, const int* slot, size_t n, float* out, const float* fallback, const int* fb_index,
                             cudaStream_t s) {
    if (n == 0) return 0;
    het_gather_kernel<<<ew_grid(n * 32, 256, 16), 256, 0, s>>>(HET_MAKE, slot, n, out, fallback, fb_index);
    HETU_CHECK_LAUNCH("hetu_het_gather");
    return 0;
}
HETU_API int hetu_het_update(HET_ARGS, const int* slot, const float* delta, size_t n, int push_bound, int* push_slots,
                             void* push_count, int* direct, void* direct_count, cudaStream_t s) {
    if (n == 0) return 0;
    het_update_kernel<<<ew_grid(n * 32, 256, 16), 256, 0, s>>>(HET_MAKE, slot, delta, n, push_bound, push_slots, (unsigned*)push_count,
                                                              direct, (unsigned*)direct_count);
    HETU_CHECK_LAUNCH("hetu_het_update");
    return 0;
}
HETU_API int hetu_het_drain(HET_ARGS, const int* push_slots, const void* push_count, int all_dirty, void* out_ids, float* out_grads,
                            void* out_upd, void* out_count, int out_cap, cudaStream_t s) {
    het_drain_kernel<<<kNumSMs * 4, 256, 0, s>>>(HET_MAKE, push_slots, (const unsigned*)push_count, all_dirty, (long long*)out_ids,
                                                 out_grads, (long long*)out_upd, (unsigned*)out_count, out_cap);
    HETU_CHECK_LAUNCH("hetu_het_drain");
    return 0;
}
