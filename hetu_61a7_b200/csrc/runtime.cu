// Native runtime under the kernels: array / stream / event C ABI and the workspace chunk cache.
//
// reference: src/common/dlarray.h:22-65 (DLContext, DLArray, DLStream, DLEvent),
// src/common/c_runtime_api.cc:93-142 (DLArrayAlloc / DLArrayFree / DLArrayCopyFromTo),
// src/common/{cpu,cuda}_device_api.cc (pinned host allocation; raw cudaMalloc per array; peer
// copies), src/cuda_common/gpu_runtime.cc:61-118 (stream / event API) and gpu_chunk.cc:18-86
// (workspace chunks cached in a multiset keyed by exact size).
//
// Same entry points for callers that do not go through torch.  Differences: device arrays come
// from the stream-ordered CUDA memory pool of the device (cudaMallocAsync, release threshold
// raised so freed blocks stay cached - with 180 GB of HBM3e there is no reason to give memory back
// between steps) instead of one cudaMalloc per array, and the chunk cache hands out the smallest
// cached chunk that fits (best fit) rather than exact-size matches only.
#include "common.cuh"

#include <map>
#include <mutex>
#include <vector>

#include "../../include/hetu_b200.h"

namespace {

std::mutex g_mu;
bool g_pool_ready[64] = {};

int ensure_pool(int dev) {
    if (dev < 0 || dev >= 64) { hetu_set_error("device id %d out of range", dev); return -1; }
    if (g_pool_ready[dev]) return 0;
    cudaMemPool_t pool;
    HETU_CUDA(cudaDeviceGetDefaultMemPool(&pool, dev));
    unsigned long long keep = ~0ull;                      // never trim: freed blocks stay in the pool
    HETU_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
    g_pool_ready[dev] = true;
    return 0;
}

size_t numel(const DLArray* a) {
    size_t n = 1;
    for (int i = 0; i < a->ndim; ++i) n *= (size_t)a->shape[i];
    return n;
}

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (dev != prev) cudaSetDevice(dev); else prev = -1; }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

// per-device cache of free workspace chunks: size -> pointers
std::multimap<size_t, void*> g_free_chunks[64];
std::map<void*, size_t> g_chunk_size[64];

}  // namespace

HETU_API int DLArrayAlloc(const long long* shape, int ndim, DLContext ctx, DLArrayHandle* out) {
    if (ndim < 0 || ndim > 16 || !out) { hetu_set_error("DLArrayAlloc: bad arguments"); return -1; }
    DLArray* a = new DLArray();
    a->ctx = ctx; a->ndim = ndim;
    a->shape = new long long[ndim > 0 ? ndim : 1];
    a->stride = new long long[ndim > 0 ? ndim : 1];
    long long st = 1;
    for (int i = ndim - 1; i >= 0; --i) { a->shape[i] = shape[i]; a->stride[i] = st; st *= shape[i]; }
    size_t bytes = numel(a) * sizeof(float);
    if (bytes == 0) bytes = sizeof(float);
    cudaError_t e;
    if (ctx.device_type == kDLGPU) {
        std::lock_guard<std::mutex> l(g_mu);
        if (ensure_pool(ctx.device_id) != 0) { delete[] a->shape; delete[] a->stride; delete a; return -1; }
        DeviceGuard g(ctx.device_id);
        e = cudaMallocAsync(&a->data, bytes, (cudaStream_t)0);
    } else {
        e = cudaMallocHost(&a->data, bytes);                 // pinned, so H2D / D2H copies are asynchronous
    }
    if (e != cudaSuccess) {
        hetu_set_error("DLArrayAlloc(%zu bytes): %s", bytes, cudaGetErrorString(e));
        delete[] a->shape; delete[] a->stride; delete a;
        return (int)e;
    }
    *out = a;
    return 0;
}

HETU_API int DLArrayFree(DLArrayHandle a) {
    if (!a) return 0;
    cudaError_t e = cudaSuccess;
    if (a->data) {
        if (a->ctx.device_type == kDLGPU) { DeviceGuard g(a->ctx.device_id); e = cudaFreeAsync(a->data, (cudaStream_t)0); }
        else e = cudaFreeHost(a->data);
    }
    delete[] a->shape; delete[] a->stride; delete a;
    if (e != cudaSuccess) { hetu_set_error("DLArrayFree: %s", cudaGetErrorString(e)); return (int)e; }
    return 0;
}

HETU_API int DLArrayCopyFromTo(DLArrayHandle from, DLArrayHandle to, DLStreamHandle stream) {
    if (!from || !to) { hetu_set_error("DLArrayCopyFromTo: null array"); return -1; }
    size_t n = numel(from);
    if (n != numel(to)) { hetu_set_error("DLArrayCopyFromTo: %zu elements into %zu", n, numel(to)); return -1; }
    cudaStream_t s = stream ? (cudaStream_t)stream->handle : (cudaStream_t)0;
    size_t bytes = n * sizeof(float);
    bool fg = from->ctx.device_type == kDLGPU, tg = to->ctx.device_type == kDLGPU;
    if (fg && tg && from->ctx.device_id != to->ctx.device_id)
        HETU_CUDA(cudaMemcpyPeerAsync(to->data, to->ctx.device_id, from->data, from->ctx.device_id, bytes, s));
    else if (!fg && !tg)
        memcpy(to->data, from->data, bytes);
    else
        HETU_CUDA(cudaMemcpyAsync(to->data, from->data, bytes, fg ? (tg ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost)
                                                                   : cudaMemcpyHostToDevice, s));
    if (!stream && (fg || tg)) HETU_CUDA(cudaStreamSynchronize(s));     // no stream given: blocking copy, like the reference
    return 0;
}

HETU_API int DLStreamCreate(int device_id, DLStreamHandle* out) {
    DeviceGuard g(device_id);
    cudaStream_t s;
    HETU_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    *out = new DLStream{device_id, (void*)s};
    return 0;
}
HETU_API int DLStreamDestroy(DLStreamHandle h) {
    if (!h) return 0;
    cudaError_t e = cudaStreamDestroy((cudaStream_t)h->handle);
    delete h;
    if (e != cudaSuccess) { hetu_set_error("DLStreamDestroy: %s", cudaGetErrorString(e)); return (int)e; }
    return 0;
}
HETU_API int DLStreamSync(DLStreamHandle h) {
    HETU_CUDA(cudaStreamSynchronize(h ? (cudaStream_t)h->handle : (cudaStream_t)0));
    return 0;
}
HETU_API int DLEventCreate(int device_id, DLEventHandle* out) {
    DeviceGuard g(device_id);
    cudaEvent_t e;
    HETU_CUDA(cudaEventCreate(&e));
    *out = new DLEvent{device_id, (void*)e};
    return 0;
}
HETU_API int DLEventDestroy(DLEventHandle h) {
    if (!h) return 0;
    cudaError_t e = cudaEventDestroy((cudaEvent_t)h->handle);
    delete h;
    if (e != cudaSuccess) { hetu_set_error("DLEventDestroy: %s", cudaGetErrorString(e)); return (int)e; }
    return 0;
}
HETU_API int DLEventRecord(DLStreamHandle s, DLEventHandle e) {
    HETU_CUDA(cudaEventRecord((cudaEvent_t)e->handle, s ? (cudaStream_t)s->handle : (cudaStream_t)0));
    return 0;
}
HETU_API int DLEventSync(DLEventHandle e) {
    HETU_CUDA(cudaEventSynchronize((cudaEvent_t)e->handle));
    return 0;
}
/* the consumer stream waits on the device; the host does not block (the reference synced on the host) */
HETU_API int DLStreamWaitEvent(DLStreamHandle s, DLEventHandle e) {
    HETU_CUDA(cudaStreamWaitEvent(s ? (cudaStream_t)s->handle : (cudaStream_t)0, (cudaEvent_t)e->handle, 0));
    return 0;
}
HETU_API int DLEventElapsedTime(DLEventHandle start, DLEventHandle end, float* ms) {
    HETU_CUDA(cudaEventElapsedTime(ms, (cudaEvent_t)start->handle, (cudaEvent_t)end->handle));
    return 0;
}

/* ---- workspace chunk cache --------------------------------------------------------------- */
HETU_API void* hetu_find_chunk(size_t bytes, int device_id) {
    if (device_id < 0 || device_id >= 64) { hetu_set_error("device id %d out of range", device_id); return nullptr; }
    std::lock_guard<std::mutex> l(g_mu);
    auto& free_ = g_free_chunks[device_id];
    auto it = free_.lower_bound(bytes);                  // smallest cached chunk that fits
    if (it != free_.end() && it->first <= bytes * 2 + 4096) {
        void* p = it->second;
        free_.erase(it);
        return p;
    }
    if (ensure_pool(device_id) != 0) return nullptr;
    DeviceGuard g(device_id);
    void* p = nullptr;
    cudaError_t e = cudaMallocAsync(&p, bytes ? bytes : 4, (cudaStream_t)0);
    if (e != cudaSuccess) { hetu_set_error("hetu_find_chunk(%zu): %s", bytes, cudaGetErrorString(e)); return nullptr; }
    g_chunk_size[device_id][p] = bytes;
    return p;
}
HETU_API int hetu_del_chunk(void* ptr, int device_id) {
    if (!ptr) return 0;
    std::lock_guard<std::mutex> l(g_mu);
    auto it = g_chunk_size[device_id].find(ptr);
    if (it == g_chunk_size[device_id].end()) { hetu_set_error("hetu_del_chunk: pointer was not handed out by hetu_find_chunk"); return -1; }
    g_free_chunks[device_id].emplace(it->second, ptr);   // back to the cache, not to the driver
    return 0;
}
HETU_API size_t hetu_chunk_cache_bytes(int device_id) {
    std::lock_guard<std::mutex> l(g_mu);
    size_t b = 0;
    for (auto& kv : g_free_chunks[device_id]) b += kv.first;
    return b;
}
HETU_API int hetu_chunk_cache_release(int device_id) {
    std::lock_guard<std::mutex> l(g_mu);
    DeviceGuard g(device_id);
    for (auto& kv : g_free_chunks[device_id]) { cudaFreeAsync(kv.second, (cudaStream_t)0); g_chunk_size[device_id].erase(kv.second); }
    g_free_chunks[device_id].clear();
    return 0;
}
