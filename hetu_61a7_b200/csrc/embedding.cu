// Embedding gather, sparse-row optimizers, on-device de-duplication - local and
// NVLink-sharded variants.
// reference: src/ops/EmbeddingLookup.cu:3-19 (one THREAD per id copies the whole row
// serially), :67-79 (zero the whole table, then per-element atomicAdd),
// src/ops/OptimizersSparse.cu:54-67,287-330 (dedup index computed on the HOST with numpy,
// python/hetu/ndarray.py:535-552); PS path: SparsePull / SparsePush RPCs through host DRAM.
//
// Here one warp moves one row with 128-bit accesses.  The sharded variants take a table of
// peer base pointers (CUDA-IPC mapped, parallel/symm.py): row r lives on GPU r % nshard at
// local row r / nshard, so a lookup reads the owner's HBM directly over NVLink and the SGD
// update is a remote red.add - the PS pull/push and the optimizer are one kernel each.
#include "common.cuh"

constexpr int kMaxShards = 16;
struct ShardTable { float* base[kMaxShards]; int nshard; };

__device__ __forceinline__ void copy_row(const float* __restrict__ src, float* __restrict__ dst, int width, int lane) {
    if ((width & 3) == 0 && ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0) {
        const float4* s4 = reinterpret_cast<const float4*>(src);
        float4* d4 = reinterpret_cast<float4*>(dst);
        for (int j = lane; j < (width >> 2); j += 32) d4[j] = s4[j];
    } else {
        for (int j = lane; j < width; j += 32) dst[j] = src[j];
    }
}

__global__ void __launch_bounds__(256) gather_rows_kernel(const float* __restrict__ table, const void* __restrict__ idx,
                                                          int idx_code, float* __restrict__ out, size_t n, int width,
                                                          size_t vocab) {
    size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t i = warp; i < n; i += nwarp) {
        long long r = load_index(idx, idx_code, i);
        float* dst = out + i * width;
        if (r < 0 || (size_t)r >= vocab) {
            for (int j = lane; j < width; j += 32) dst[j] = 0.f;
        } else {
            copy_row(table + (size_t)r * width, dst, width, lane);
        }
    }
}
HETU_API int hetu_gather_rows(const float* table, const void* idx, int idx_code, float* out, size_t n, int width,
                              size_t vocab, cudaStream_t s) {
    if (n == 0) return 0;
    gather_rows_kernel<<<ew_grid(n * 32, 256, 16), 256, 0, s>>>(table, idx, idx_code, out, n, width, vocab);
    HETU_CHECK_LAUNCH("hetu_gather_rows");
    return 0;
}

__global__ void __launch_bounds__(256) gather_rows_sharded_kernel(ShardTable tab, const void* __restrict__ idx, int idx_code,
                                                                  float* __restrict__ out, size_t n, int width, size_t vocab) {
    size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t i = warp; i < n; i += nwarp) {
        long long r = load_index(idx, idx_code, i);
        float* dst = out + i * width;
        if (r < 0 || (size_t)r >= vocab) {
            for (int j = lane; j < width; j += 32) dst[j] = 0.f;
        } else {
            int owner = (int)(r % tab.nshard);
            size_t local = (size_t)(r / tab.nshard);
            copy_row(tab.base[owner] + local * width, dst, width, lane);   // peer ld.global over NVLink
        }
    }
}
HETU_API int hetu_gather_rows_sharded(void* const* bases, int nshard, const void* idx, int idx_code, float* out,
                                      size_t n, int width, size_t vocab, cudaStream_t s) {
    if (n == 0) return 0;
    if (nshard > kMaxShards) { hetu_set_error("gather_rows_sharded: nshard %d > %d", nshard, kMaxShards); return -1; }
    ShardTable tab;
    tab.nshard = nshard;
    for (int i = 0; i < kMaxShards; ++i) tab.base[i] = i < nshard ? (float*)bases[i] : nullptr;
    gather_rows_sharded_kernel<<<ew_grid(n * 32, 256, 16), 256, 0, s>>>(tab, idx, idx_code, out, n, width, vocab);
    HETU_CHECK_LAUNCH("hetu_gather_rows_sharded");
    return 0;
}

// Sorted ids -> per-run sums.  Warp i handles position i of the sorted order.  The run [lo, hi) of equal
// ids is found by binary search.  Short runs (<= 16 rows) are summed by the run head alone, in a fixed
// order; long runs (BERT position / token-type tables: thousands of duplicates of a handful of ids) are
// accumulated by all their members with atomic adds into the head's zero-initialised row, so no warp
// ever walks a long run serially.  Non-head positions get id -1 and are skipped by the optimizer
// kernel.  Static shapes: the step stays CUDA-graph capturable (torch.unique is data-dependent).
__global__ void __launch_bounds__(256) dedup_sorted_rows_kernel(const long long* __restrict__ sidx,
                                                                const long long* __restrict__ perm,
                                                                const float* __restrict__ vals, long long* __restrict__ didx,
                                                                float* __restrict__ dval, size_t n, int width) {
    size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t i = warp; i < n; i += nwarp) {
        const long long id = sidx[i];
        size_t lo = 0, hi = i;                                  // first position holding id
        while (lo < hi) { size_t mid = (lo + hi) >> 1; if (sidx[mid] < id) lo = mid + 1; else hi = mid; }
        const size_t head = lo;
        size_t l2 = i + 1, h2 = n;                              // one past the last position holding id
        while (l2 < h2) { size_t mid = (l2 + h2) >> 1; if (sidx[mid] <= id) l2 = mid + 1; else h2 = mid; }
        const size_t end = l2;
        if (lane == 0) didx[i] = (i == head) ? id : -1;
        if (end - head <= 16) {
            if (i != head) continue;
            for (int j = lane; j < width; j += 32) {
                float acc = 0.f;
                for (size_t k = head; k < end; ++k) acc += vals[(size_t)perm[k] * width + j];
                dval[head * width + j] = acc;
            }
        } else {
            const float* src = vals + (size_t)perm[i] * width;
            float* dst = dval + head * width;
            for (int j = lane; j < width; j += 32) atomicAdd(dst + j, src[j]);
        }
    }
}
HETU_API int hetu_dedup_sorted_rows(const void* sidx, const void* perm, const float* vals, void* didx, float* dval,
                                    size_t n, int width, cudaStream_t s) {
    if (n == 0) return 0;
    HETU_CUDA(cudaMemsetAsync(dval, 0, sizeof(float) * n * (size_t)width, s));
    dedup_sorted_rows_kernel<<<ew_grid(n * 32, 256, 16), 256, 0, s>>>((const long long*)sidx, (const long long*)perm, vals,
                                                                      (long long*)didx, dval, n, width);
    HETU_CHECK_LAUNCH("hetu_dedup_sorted_rows");
    return 0;
}

enum { OPT_SGD = 0, OPT_MOMENTUM, OPT_NESTEROV, OPT_ADAGRAD, OPT_ADAM, OPT_ADAMW, OPT_LAMB };

// one warp per (deduplicated) gradient row; ids < 0 are skipped.
template <int KIND>
__global__ void __launch_bounds__(256) sparse_opt_kernel(float* __restrict__ param, const long long* __restrict__ idx,
                                                         const float* __restrict__ grad, float* __restrict__ s1,
                                                         float* __restrict__ s2, size_t n, int width,
                                                         const float* __restrict__ lr_p, const long long* __restrict__ step_p,
                                                         float h0, float h1, float h2, float h3, float l2reg, int use_atomic) {
    size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    float lr = *lr_p;
    float b1t = 0.f, b2t = 0.f;
    if (KIND >= OPT_ADAM) {
        float t = (float)(*step_p);
        b1t = 1.f - powf(h0, t);
        b2t = 1.f - powf(h1, t);
    }
    for (size_t i = warp; i < n; i += nwarp) {
        long long r = idx[i];
        if (r < 0) continue;
        float* p = param + (size_t)r * width;
        const float* g = grad + i * width;
        for (int j = lane; j < width; j += 32) {
            float gv = g[j];
            if (KIND == OPT_SGD) {
                if (use_atomic) { atomicAdd(p + j, -lr * (gv + l2reg * p[j])); }
                else p[j] -= lr * (gv + l2reg * p[j]);
                continue;
            }
            float pv = p[j];
            gv += l2reg * pv;
            size_t so = (size_t)r * width + j;
            if (KIND == OPT_MOMENTUM) {
                float v = h0 * s1[so] - lr * gv;
                s1[so] = v;
                p[j] = pv + v;
            } else if (KIND == OPT_NESTEROV) {
                float tmp = lr * gv;
                float v = h0 * (s1[so] - tmp);
                s1[so] = v;
                p[j] = pv + v - tmp;
            } else if (KIND == OPT_ADAGRAD) {
                float a = s1[so] + gv * gv;
                s1[so] = a;
                p[j] = pv - lr * gv / (sqrtf(a) + h0);
            } else {
                float m = h0 * s1[so] + (1.f - h0) * gv;
                float v = h1 * s2[so] + (1.f - h1) * gv * gv;
                s1[so] = m;
                s2[so] = v;
                float upd = (m / b1t) / (sqrtf(v / b2t) + h2);
                if (KIND == OPT_ADAM) p[j] = pv - lr * upd;
                else p[j] = pv - lr * (upd + h3 * pv);      // AdamW; sparse Lamb uses the AdamW form per row
            }
        }
    }
}
HETU_API int hetu_sparse_opt(int kind, float* param, const void* idx, const float* grad, float* s1, float* s2, size_t n,
                             int width, const float* lr_p, const void* step_p, float h0, float h1, float h2, float h3,
                             float l2reg, int use_atomic, cudaStream_t s) {
    if (n == 0) return 0;
    int grid = ew_grid(n * 32, 256, 16);
    const long long* ix = (const long long*)idx;
    const long long* st = (const long long*)step_p;
#define SP(K) sparse_opt_kernel<K><<<grid, 256, 0, s>>>(param, ix, grad, s1, s2, n, width, lr_p, st, h0, h1, h2, h3, l2reg, use_atomic)
    switch (kind) {
        case OPT_SGD: SP(OPT_SGD); break;
        case OPT_MOMENTUM: SP(OPT_MOMENTUM); break;
        case OPT_NESTEROV: SP(OPT_NESTEROV); break;
        case OPT_ADAGRAD: SP(OPT_ADAGRAD); break;
        case OPT_ADAM: SP(OPT_ADAM); break;
        case OPT_ADAMW: SP(OPT_ADAMW); break;
        case OPT_LAMB: SP(OPT_LAMB); break;
        default: hetu_set_error("hetu_sparse_opt: bad kind %d", kind); return -1;
    }
#undef SP
    HETU_CHECK_LAUNCH("hetu_sparse_opt");
    return 0;
}

// Sharded SGD / AdaGrad push: every worker adds -lr*g into the owner's rows with remote
// atomics (exactly the PS "server does +=" semantics, ParameterServerCommunicate.py:21-25,
// without the host round trip).  AdaGrad keeps its accumulator shard next to the table.
__global__ void __launch_bounds__(256) sparse_sgd_sharded_kernel(ShardTable tab, const void* __restrict__ idx, int idx_code,
                                                                 const float* __restrict__ grad, size_t n, int width,
                                                                 size_t vocab, const float* __restrict__ lr_p, float scale) {
    size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    float k = -(*lr_p) * scale;
    for (size_t i = warp; i < n; i += nwarp) {
        long long r = load_index(idx, idx_code, i);
        if (r < 0 || (size_t)r >= vocab) continue;
        float* p = tab.base[(int)(r % tab.nshard)] + (size_t)(r / tab.nshard) * width;
        const float* g = grad + i * width;
        for (int j = lane; j < width; j += 32) atomicAdd(p + j, k * g[j]);      // red.global.add.f32 over NVLink
    }
}
HETU_API int hetu_sparse_sgd_sharded(void* const* bases, int nshard, const void* idx, int idx_code, const float* grad,
                                     size_t n, int width, size_t vocab, const float* lr_p, float scale, cudaStream_t s) {
    if (n == 0) return 0;
    if (nshard > kMaxShards) { hetu_set_error("sparse_sgd_sharded: nshard %d > %d", nshard, kMaxShards); return -1; }
    ShardTable tab;
    tab.nshard = nshard;
    for (int i = 0; i < kMaxShards; ++i) tab.base[i] = i < nshard ? (float*)bases[i] : nullptr;
    sparse_sgd_sharded_kernel<<<ew_grid(n * 32, 256, 16), 256, 0, s>>>(tab, idx, idx_code, grad, n, width, vocab, lr_p, scale);
    HETU_CHECK_LAUNCH("hetu_sparse_sgd_sharded");
    return 0;
}

// dense += sparse rows (IndexedSlices -> dense accumulate); reference IndexedSlices.cu
__global__ void __launch_bounds__(256) scatter_add_rows_kernel(float* __restrict__ dense, const void* __restrict__ idx,
                                                               int idx_code, const float* __restrict__ vals, size_t n,
                                                               int width, size_t vocab) {
    size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t i = warp; i < n; i += nwarp) {
        long long r = load_index(idx, idx_code, i);
        if (r < 0 || (size_t)r >= vocab) continue;
        for (int j = lane; j < width; j += 32) atomicAdd(dense + (size_t)r * width + j, vals[i * width + j]);
    }
}
HETU_API int hetu_scatter_add_rows(float* dense, const void* idx, int idx_code, const float* vals, size_t n, int width,
                                   size_t vocab, cudaStream_t s) {
    if (n == 0) return 0;
    scatter_add_rows_kernel<<<ew_grid(n * 32, 256, 16), 256, 0, s>>>(dense, idx, idx_code, vals, n, width, vocab);
    HETU_CHECK_LAUNCH("hetu_scatter_add_rows");
    return 0;
}

// ============================================================================================
// Sharded table, optimizers with state (Momentum / Nesterov / AdaGrad / Adam / AdamW) applied AT THE OWNER.
// reference: the PS server applies them per pushed row (ps-lite/include/ps/server/optimizer.h:36-357) after the
// worker has de-duplicated its ids on the host (src/ops/OptimizersSparse.cu:54-67,287-330).
//
// Here: phase 1 (every rank, hetu_sparse_push_sharded): each gradient row is ADDED into the owner's accumulation
// shard with remote vector atomics - duplicates, within a rank and across ranks, meet there, which is the
// de-duplication - and the first toucher of a row (atomicExch on the owner's per-row flag) appends the row to the
// owner's work list.  A cross-GPU barrier follows (host side).  Phase 2 (owner only, hetu_sparse_apply_owner):
// one warp per listed row reads the summed gradient once, applies the optimizer to the row and its state rows,
// clears the accumulator and the flag; the last CTA resets the list counter.
struct OwnerPush {
    float* acc[kMaxShards];          // [local_rows, width] accumulation shard of every rank
    unsigned* flag[kMaxShards];      // [local_rows]
    int* list[kMaxShards];           // work list of every rank
    unsigned* count[kMaxShards];     // its length
    int nshard, cap;
};

__global__ void __launch_bounds__(256) sparse_push_sharded_kernel(OwnerPush op, const void* __restrict__ idx, int idx_code,
                                                                  const float* __restrict__ grad, size_t n, int width,
                                                                  size_t vocab, float scale) {
    size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t i = warp; i < n; i += nwarp) {
        const long long r = load_index(idx, idx_code, i);
        if (r < 0 || (size_t)r >= vocab) continue;
        const int owner = (int)(r % op.nshard);
        const size_t local = (size_t)(r / op.nshard);
        float* dst = op.acc[owner] + local * width;
        const float* g = grad + i * width;
        if ((width & 3) == 0) {
            for (int j = lane; j < (width >> 2); j += 32) {
                const float4 v = reinterpret_cast<const float4*>(g)[j];
                atomicAdd(reinterpret_cast<float4*>(dst) + j, make_float4(scale * v.x, scale * v.y, scale * v.z, scale * v.w));
            }
        } else {
            for (int j = lane; j < width; j += 32) atomicAdd(dst + j, scale * g[j]);
        }
        if (lane == 0 && atomicExch(op.flag[owner] + local, 1u) == 0u) {
            const unsigned pos = atomicAdd(op.count[owner], 1u);
            if ((int)pos < op.cap) op.list[owner][pos] = (int)local;
        }
    }
}

template <int KIND>
__global__ void __launch_bounds__(256) sparse_apply_owner_kernel(float* __restrict__ table, float* __restrict__ acc,
                                                                 unsigned* __restrict__ flag, const int* __restrict__ list,
                                                                 unsigned* __restrict__ count, unsigned* __restrict__ done,
                                                                 float* __restrict__ s1, float* __restrict__ s2, int width, int cap,
                                                                 const float* __restrict__ lr_p, const long long* __restrict__ step_p,
                                                                 float h0, float h1, float h2, float h3, float l2reg) {
    const unsigned n = min(*reinterpret_cast<volatile unsigned*>(count), (unsigned)cap);
    size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    const float lr = *lr_p;
    float b1t = 1.f, b2t = 1.f;
    if (KIND >= OPT_ADAM) {
        const float t = (float)(*step_p);
        b1t = 1.f - powf(h0, t);
        b2t = 1.f - powf(h1, t);
    }
    for (size_t i = warp; i < n; i += nwarp) {
        const size_t row = (size_t)list[i];
        float* p = table + row * width;
        float* a = acc + row * width;
        for (int j = lane; j < width; j += 32) {
            const float pv = p[j];
            const float gv = a[j] + l2reg * pv;
            a[j] = 0.f;
            const size_t so = row * width + j;
            if (KIND == OPT_MOMENTUM) {
                const float v = h0 * s1[so] - lr * gv;
                s1[so] = v;
                p[j] = pv + v;
            } else if (KIND == OPT_NESTEROV) {
                const float tmp = lr * gv;
                const float v = h0 * (s1[so] - tmp);
                s1[so] = v;
                p[j] = pv + v - tmp;
            } else if (KIND == OPT_ADAGRAD) {
                const float q = s1[so] + gv * gv;
                s1[so] = q;
                p[j] = pv - lr * gv / (sqrtf(q) + h0);
            } else {
                const float m = h0 * s1[so] + (1.f - h0) * gv;
                const float v = h1 * s2[so] + (1.f - h1) * gv * gv;
                s1[so] = m;
                s2[so] = v;
                const float upd = (m / b1t) / (sqrtf(v / b2t) + h2);
                p[j] = (KIND == OPT_ADAM) ? pv - lr * upd : pv - lr * (upd + h3 * pv);
            }
        }
        if (lane == 0) flag[row] = 0u;
    }
    // the last CTA to finish empties the work list for the next step
    __shared__ bool last;
    __syncthreads();
    if (threadIdx.x == 0) { __threadfence(); last = atomicAdd(done, 1u) == gridDim.x - 1; }
    __syncthreads();
    if (last && threadIdx.x == 0) { *count = 0u; *done = 0u; }
}

HETU_API int hetu_sparse_push_sharded(void* const* acc, void* const* flag, void* const* list, void* const* count, int nshard,
                                      int cap, const void* idx, int idx_code, const float* grad, size_t n, int width,
                                      size_t vocab, float scale, cudaStream_t s) {
    if (n == 0) return 0;
    if (nshard < 1 || nshard > kMaxShards) { hetu_set_error("sparse_push_sharded: nshard %d outside [1, %d]", nshard, kMaxShards); return -1; }
    OwnerPush op;
    op.nshard = nshard; op.cap = cap;
    for (int i = 0; i < kMaxShards; ++i) {
        op.acc[i] = i < nshard ? (float*)acc[i] : nullptr;
        op.flag[i] = i < nshard ? (unsigned*)flag[i] : nullptr;
        op.list[i] = i < nshard ? (int*)list[i] : nullptr;
        op.count[i] = i < nshard ? (unsigned*)count[i] : nullptr;
    }
    sparse_push_sharded_kernel<<<ew_grid(n * 32, 256, 16), 256, 0, s>>>(op, idx, idx_code, grad, n, width, vocab, scale);
    HETU_CHECK_LAUNCH("hetu_sparse_push_sharded");
    return 0;
}

HETU_API int hetu_sparse_apply_owner(int kind, float* table, float* acc, void* flag, const void* list, void* count, void* done,
                                     float* s1, float* s2, int width, int cap, const float* lr_p, const void* step_p, float h0,
                                     float h1, float h2, float h3, float l2reg, cudaStream_t s) {
    const int grid = kNumSMs * 4;
    const long long* st = (const long long*)step_p;
#define AP(K) sparse_apply_owner_kernel<K><<<grid, 256, 0, s>>>(table, acc, (unsigned*)flag, (const int*)list, (unsigned*)count, \
                                                                 (unsigned*)done, s1, s2, width, cap, lr_p, st, h0, h1, h2, h3, l2reg)
    switch (kind) {
        case OPT_MOMENTUM: AP(OPT_MOMENTUM); break;
        case OPT_NESTEROV: AP(OPT_NESTEROV); break;
        case OPT_ADAGRAD: AP(OPT_ADAGRAD); break;
        case OPT_ADAM: AP(OPT_ADAM); break;
        case OPT_ADAMW: case OPT_LAMB: AP(OPT_ADAMW); break;
        default: hetu_set_error("hetu_sparse_apply_owner: kind %d has no state (SGD uses hetu_sparse_sgd_sharded)", kind); return -1;
    }
#undef AP
    HETU_CHECK_LAUNCH("hetu_sparse_apply_owner");
    return 0;
}
