// Multi-tensor dense optimizer step: every parameter tensor of the model in ONE launch.
// reference: src/ops/Optimizers.cu:39-402 - one launch per tensor per step
// (sgd_update / momentum_update / adagrad_update / adam_update / adamw_update /
// calc_lamb_update + lamb_update_step with cudnnReduceTensor norms).
//
// grid = (x-chunks, tensor index).  Pointers travel by value in the launch parameters;
// lr and the step count are read from device memory so a captured CUDA graph stays valid
// while the schedule / bias correction advance.
#include "common.cuh"
#include <mutex>

enum { OPT_SGD = 0, OPT_MOMENTUM, OPT_NESTEROV, OPT_ADAGRAD, OPT_ADAM, OPT_ADAMW, OPT_LAMB };
constexpr int kMaxTensors = 48;

constexpr int kOptChunk = 256 * 4 * 8;       // elements per CTA of the flattened element-wise update (8 float4 per thread)

struct TensorList {
    float* p[kMaxTensors];
    const float* g[kMaxTensors];
    float* s1[kMaxTensors];
    float* s2[kMaxTensors];
    long long n[kMaxTensors];
    int chunk_begin[kMaxTensors + 1];        // flattened schedule: CTA b works on tensor t with chunk_begin[t] <= b < chunk_begin[t+1]
};

template <int KIND>
__global__ void __launch_bounds__(256) multi_tensor_opt_kernel(TensorList tl, const float* __restrict__ lr_p,
                                                               const long long* __restrict__ step_p, float h0, float h1,
                                                               float h2, float h3, float l2reg, float grad_scale) {
    // flattened schedule: every CTA owns one kOptChunk-element chunk of one tensor, so a 31M-element
    // embedding table and a 1024-element bias are both covered at full occupancy
    int t = 0;
    {
        int lo = 0, hi = kMaxTensors;
        while (lo < hi) { int mid = (lo + hi + 1) >> 1; if (tl.chunk_begin[mid] <= (int)blockIdx.x) lo = mid; else hi = mid - 1; }
        t = lo;
    }
    const long long n = tl.n[t];
    float* __restrict__ p = tl.p[t];
    const float* __restrict__ g = tl.g[t];
    float* __restrict__ s1 = tl.s1[t];
    float* __restrict__ s2 = tl.s2[t];
    float lr = *lr_p;
    float b1t = 1.f, b2t = 1.f;
    if (KIND >= OPT_ADAM) {
        float st = (float)(*step_p);
        b1t = 1.f - powf(h0, st);
        b2t = 1.f - powf(h1, st);
    }
    const long long c0 = (long long)((int)blockIdx.x - tl.chunk_begin[t]) * kOptChunk;
    const long long c1 = c0 + kOptChunk < n ? c0 + kOptChunk : n;
    for (long long i = c0 + threadIdx.x; i < c1; i += blockDim.x) {
        float pv = p[i];
        float gv = g[i] * grad_scale + l2reg * pv;
        if (KIND == OPT_SGD) {
            p[i] = pv - lr * gv;
        } else if (KIND == OPT_MOMENTUM) {
            float v = h0 * s1[i] - lr * gv;
            s1[i] = v;
            p[i] = pv + v;
        } else if (KIND == OPT_NESTEROV) {
            float tmp = lr * gv;
            float v = h0 * (s1[i] - tmp);
            s1[i] = v;
            p[i] = pv + v - tmp;
        } else if (KIND == OPT_ADAGRAD) {
            float a = s1[i] + gv * gv;
            s1[i] = a;
            p[i] = pv - lr * gv / (sqrtf(a) + h0);
        } else {
            float m = h0 * s1[i] + (1.f - h0) * gv;
            float v = h1 * s2[i] + (1.f - h1) * gv * gv;
            s1[i] = m;
            s2[i] = v;
            float upd = (m / b1t) / (sqrtf(v / b2t) + h2);
            if (KIND == OPT_ADAM) p[i] = pv - lr * upd;
            else p[i] = pv - lr * (upd + h3 * pv);
        }
    }
}

// Lamb, phase 1: moments + per-tensor ||p||^2 and ||update||^2 into scratch[2t], scratch[2t+1]
__global__ void __launch_bounds__(256) lamb_phase1_kernel(TensorList tl, const long long* __restrict__ step_p, float h0,
                                                          float h1, float h2, float l2reg, float grad_scale,
                                                          float* __restrict__ scratch) {
    __shared__ float sm[32];
    int t = blockIdx.y;
    long long n = tl.n[t];
    float st = (float)(*step_p);
    float b1t = 1.f - powf(h0, st), b2t = 1.f - powf(h1, st);
    float np = 0.f, nu = 0.f;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        float pv = tl.p[t][i];
        float gv = tl.g[t][i] * grad_scale + l2reg * pv;
        float m = h0 * tl.s1[t][i] + (1.f - h0) * gv;
        float v = h1 * tl.s2[t][i] + (1.f - h1) * gv * gv;
        tl.s1[t][i] = m;
        tl.s2[t][i] = v;
        float upd = (m / b1t) / (sqrtf(v / b2t) + h2);
        np += pv * pv;
        nu += upd * upd;
    }
    np = block_sum(np, sm);
    nu = block_sum(nu, sm);
    if (threadIdx.x == 0) {
        atomicAdd(scratch + 2 * t, np);
        atomicAdd(scratch + 2 * t + 1, nu);
    }
}
__global__ void __launch_bounds__(256) lamb_phase2_kernel(TensorList tl, const float* __restrict__ lr_p,
                                                          const long long* __restrict__ step_p, float h0, float h1, float h2,
                                                          float h3, const float* __restrict__ scratch) {
    int t = blockIdx.y;
    long long n = tl.n[t];
    float st = (float)(*step_p);
    float b1t = 1.f - powf(h0, st), b2t = 1.f - powf(h1, st);
    float np2 = scratch[2 * t], nu2 = scratch[2 * t + 1];
    float ratio = (np2 > 0.f && nu2 > 0.f) ? sqrtf(np2) / sqrtf(nu2) : 1.f;   // zero update or zero weights: plain step
    float lr = *lr_p;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        float pv = tl.p[t][i];
        float upd = (tl.s1[t][i] / b1t) / (sqrtf(tl.s2[t][i] / b2t) + h2);
        tl.p[t][i] = pv - lr * ratio * (upd + h3 * pv);
    }
}

HETU_API int hetu_multi_tensor_opt(int kind, int ntensors, void* const* params, void* const* grads, void* const* s1,
                                   void* const* s2, const long long* sizes, const float* lr_p, const void* step_p,
                                   float h0, float h1, float h2, float h3, float l2reg, float* scratch,
                                   cudaStream_t s) {
    if (ntensors <= 0) return 0;
    if (ntensors > kMaxTensors) { hetu_set_error("multi_tensor_opt: %d tensors > %d per launch", ntensors, kMaxTensors); return -1; }
    TensorList tl;
    long long maxn = 0;
    for (int i = 0; i < kMaxTensors; ++i) {
        bool on = i < ntensors;
        tl.p[i] = on ? (float*)params[i] : nullptr;
        tl.g[i] = on ? (const float*)grads[i] : nullptr;
        tl.s1[i] = on ? (float*)s1[i] : nullptr;
        tl.s2[i] = on ? (float*)s2[i] : nullptr;
        tl.n[i] = on ? sizes[i] : 0;
        if (on && sizes[i] > maxn) maxn = sizes[i];
    }
    int gx = (int)((maxn + 256 * 4 - 1) / (256 * 4));
    int cap = (kNumSMs * 8 + ntensors - 1) / ntensors;
    if (gx > cap) gx = cap;
    if (gx < 1) gx = 1;
    dim3 grid(gx, ntensors);                     // Lamb: (x-chunks, tensor) grid with grid-stride loops
    int nchunks = 0;
    for (int i = 0; i <= kMaxTensors; ++i) {
        tl.chunk_begin[i] = nchunks;
        if (i < ntensors) nchunks += (int)((sizes[i] + kOptChunk - 1) / kOptChunk);
    }
    for (int i = ntensors + 1; i <= kMaxTensors; ++i) tl.chunk_begin[i] = 0x7fffffff;   // never selected by the search
    if (kind != OPT_LAMB) grid = dim3(nchunks > 0 ? nchunks : 1, 1);
    if (kind != OPT_LAMB && nchunks == 0) return 0;
    const long long* st = (const long long*)step_p;
    switch (kind) {
        case OPT_SGD: multi_tensor_opt_kernel<OPT_SGD><<<grid, 256, 0, s>>>(tl, lr_p, st, h0, h1, h2, h3, l2reg, 1.f); break;
        case OPT_MOMENTUM: multi_tensor_opt_kernel<OPT_MOMENTUM><<<grid, 256, 0, s>>>(tl, lr_p, st, h0, h1, h2, h3, l2reg, 1.f); break;
        case OPT_NESTEROV: multi_tensor_opt_kernel<OPT_NESTEROV><<<grid, 256, 0, s>>>(tl, lr_p, st, h0, h1, h2, h3, l2reg, 1.f); break;
        case OPT_ADAGRAD: multi_tensor_opt_kernel<OPT_ADAGRAD><<<grid, 256, 0, s>>>(tl, lr_p, st, h0, h1, h2, h3, l2reg, 1.f); break;
        case OPT_ADAM: multi_tensor_opt_kernel<OPT_ADAM><<<grid, 256, 0, s>>>(tl, lr_p, st, h0, h1, h2, h3, l2reg, 1.f); break;
        case OPT_ADAMW: multi_tensor_opt_kernel<OPT_ADAMW><<<grid, 256, 0, s>>>(tl, lr_p, st, h0, h1, h2, h3, l2reg, 1.f); break;
        case OPT_LAMB:
            if (!scratch) { hetu_set_error("multi_tensor_opt: Lamb needs a zeroed scratch of 2 floats per tensor"); return -1; }
            lamb_phase1_kernel<<<grid, 256, 0, s>>>(tl, st, h0, h1, h2, l2reg, 1.f, scratch);
            lamb_phase2_kernel<<<grid, 256, 0, s>>>(tl, lr_p, st, h0, h1, h2, h3, scratch);
            break;
        default: hetu_set_error("multi_tensor_opt: bad kind %d", kind); return -1;
    }
    HETU_CHECK_LAUNCH("hetu_multi_tensor_opt");
    return 0;
}

// ============================================================================================
// Lamb step for callers that cannot allocate device memory.
//
// hetu_multi_tensor_opt(kind = Lamb) needs two floats of device scratch per tensor (the squared
// norms of the parameters and of the update) and expects the caller to allocate and zero them.
// Some callers have no way to do that: code running under a framework whose allocator they must
// not re-enter (e.g. from inside another library's callback), real-time / embedded-style hosts
// that forbid cudaMalloc after start-up because it synchronises the device and can fail late,
// and thin C or FFI clients that only own the tensors they pass in.  For them this entry point
// keeps the scratch in a small file-scope device array that lives in the library image itself:
// every GPU that loads libhetu_b200.so holds HETU_LAMB_NOALLOC_MAX_TENSORS * 2 floats (768 B)
// until the process exits.
//
// Rules that follow from the scratch being shared and static:
//  * at most HETU_LAMB_NOALLOC_MAX_TENSORS tensors per call - more is rejected, not truncated;
//    split a larger model into several calls on the same stream;
//  * the array keeps whatever the previous call left there, so each call first clears the slots
//    it is about to accumulate into (a memset on `s`, not an allocation);
//  * two calls whose kernels can overlap on one GPU (different streams, or different host
//    threads on different streams) race on the array - serialise them on one stream;
//  * the first call on a given device looks up the array's device address (set-up, once per
//    device, done with that device current).  Make that first call outside of any stream
//    capture.
// No cudaMalloc / cudaFree / allocating library routine is called on any path through here.
constexpr int HETU_LAMB_NOALLOC_MAX_TENSORS = 96;
static __device__ __align__(256) float g_lamb_noalloc_scratch[2 * HETU_LAMB_NOALLOC_MAX_TENSORS];

namespace {
constexpr int kNoAllocMaxDevices = 64;
float* g_noalloc_ptr[kNoAllocMaxDevices] = {};     // device address of the array, per device ordinal
std::mutex g_noalloc_mu;

// The symbol name is only a host-side handle; kernels must be given the address the runtime
// reports for the *current* device.
int lamb_noalloc_scratch(float** out) {
    int dev = -1;
    HETU_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kNoAllocMaxDevices) {
        hetu_set_error("hetu_lamb_step_noalloc: device ordinal %d outside [0, %d)", dev, kNoAllocMaxDevices);
        return -1;
    }
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (g_noalloc_ptr[dev] == nullptr) {
        float* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&p, g_lamb_noalloc_scratch);
        if (e != cudaSuccess || p == nullptr) {
            hetu_set_error("hetu_lamb_step_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev,
                           cudaGetErrorString(e));
            return e != cudaSuccess ? (int)e : -1;
        }
        g_noalloc_ptr[dev] = p;
    }
    *out = g_noalloc_ptr[dev];
    return 0;
}
}  // namespace

HETU_API int hetu_lamb_step_noalloc(int ntensors, void* const* params, void* const* grads, void* const* s1,
                                    void* const* s2, const long long* sizes, const float* lr_p, const void* step_p,
                                    float beta1, float beta2, float eps, float weight_decay, float l2reg,
                                    cudaStream_t s) {
    if (ntensors <= 0) return 0;
    if (ntensors > HETU_LAMB_NOALLOC_MAX_TENSORS) {
        hetu_set_error("hetu_lamb_step_noalloc: %d tensors exceed the static scratch (max %d per call); "
                       "split the step into several calls on the same stream",
                       ntensors, HETU_LAMB_NOALLOC_MAX_TENSORS);
        return -1;
    }
    float* scratch = nullptr;
    int rc = lamb_noalloc_scratch(&scratch);
    if (rc) return rc;
    // norms are accumulated with atomicAdd: the slots used by this call must start at zero
    HETU_CUDA(cudaMemsetAsync(scratch, 0, sizeof(float) * 2 * (size_t)ntensors, s));
    const long long* st = (const long long*)step_p;
    for (int base = 0; base < ntensors; base += kMaxTensors) {
        int cnt = ntensors - base < kMaxTensors ? ntensors - base : kMaxTensors;
        TensorList tl;
        long long maxn = 0;
        for (int i = 0; i < kMaxTensors; ++i) {
            bool on = i < cnt;
            tl.p[i] = on ? (float*)params[base + i] : nullptr;
            tl.g[i] = on ? (const float*)grads[base + i] : nullptr;
            tl.s1[i] = on ? (float*)s1[base + i] : nullptr;
            tl.s2[i] = on ? (float*)s2[base + i] : nullptr;
            tl.n[i] = on ? sizes[base + i] : 0;
            if (on && sizes[base + i] > maxn) maxn = sizes[base + i];
        }
        int gx = (int)((maxn + 256 * 4 - 1) / (256 * 4));
        int cap = (kNumSMs * 8 + cnt - 1) / cnt;
        if (gx > cap) gx = cap;
        if (gx < 1) gx = 1;
        dim3 grid(gx, cnt);
        lamb_phase1_kernel<<<grid, 256, 0, s>>>(tl, st, beta1, beta2, eps, l2reg, 1.f, scratch + 2 * base);
        lamb_phase2_kernel<<<grid, 256, 0, s>>>(tl, lr_p, st, beta1, beta2, eps, weight_decay, scratch + 2 * base);
    }
    HETU_CHECK_LAUNCH("hetu_lamb_step_noalloc");
    return 0;
}
