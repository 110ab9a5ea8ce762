// Shared helpers for the sm_100a kernel library (libhetu_b200.so).
// Every exported function is `extern "C" int hetu_<name>(..., cudaStream_t)`,
// returns 0 on success, records a message readable through hetu_last_error().
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstddef>

#define HETU_API extern "C" __attribute__((visibility("default")))

void hetu_set_error(const char* fmt, ...);

#define HETU_CHECK_LAUNCH(name)                                                   \
    do {                                                                          \
        cudaError_t _e = cudaGetLastError();                                      \
        if (_e != cudaSuccess) {                                                  \
            hetu_set_error("%s: %s", name, cudaGetErrorString(_e));               \
            return (int)_e;                                                       \
        }                                                                         \
    } while (0)

#define HETU_CUDA(call)                                                           \
    do {                                                                          \
        cudaError_t _e = (call);                                                  \
        if (_e != cudaSuccess) {                                                  \
            hetu_set_error("%s failed: %s", #call, cudaGetErrorString(_e));       \
            return (int)_e;                                                       \
        }                                                                         \
    } while (0)

constexpr int kNumSMs = 148;   // B200: 2 dies x 74 SMs

// grid for a grid-stride element-wise kernel: a few waves of 148 SMs at most
static inline int ew_grid(size_t work_items, int threads, int max_ctas_per_sm = 8) {
    size_t need = (work_items + threads - 1) / threads;
    size_t cap = (size_t)kNumSMs * max_ctas_per_sm;
    if (need < 1) need = 1;
    return (int)(need < cap ? need : cap);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// block-wide sum for blockDim.x <= 1024 (result valid in every thread)
__device__ __forceinline__ float block_sum(float v, float* smem32) {
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) smem32[warp] = v;
    __syncthreads();
    int nw = (blockDim.x + 31) >> 5;
    v = (threadIdx.x < nw) ? smem32[threadIdx.x] : 0.f;
    if (warp == 0) v = warp_sum(v);
    if (threadIdx.x == 0) smem32[0] = v;
    __syncthreads();
    return smem32[0];
}
__device__ __forceinline__ float block_max(float v, float* smem32) {
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) smem32[warp] = v;
    __syncthreads();
    int nw = (blockDim.x + 31) >> 5;
    v = (threadIdx.x < nw) ? smem32[threadIdx.x] : -INFINITY;
    if (warp == 0) v = warp_max(v);
    if (threadIdx.x == 0) smem32[0] = v;
    __syncthreads();
    return smem32[0];
}

// float / int32 / int64 encoded row ids (reference convention is float32)
__device__ __forceinline__ long long load_index(const void* p, int code, size_t i) {
    if (code == 0) return (long long)((const float*)p)[i];
    if (code == 1) return (long long)((const int*)p)[i];
    return ((const long long*)p)[i];
}

// Philox4x32-10 counter-based generator (random.cu, dropout)
struct Philox {
    uint32_t key[2];
    __device__ Philox(unsigned long long seed) {
        key[0] = (uint32_t)seed;
        key[1] = (uint32_t)(seed >> 32);
    }
    __device__ __forceinline__ uint4 operator()(unsigned long long ctr, unsigned long long sub = 0) const {
        uint32_t c0 = (uint32_t)ctr, c1 = (uint32_t)(ctr >> 32), c2 = (uint32_t)sub, c3 = (uint32_t)(sub >> 32);
        uint32_t k0 = key[0], k1 = key[1];
#pragma unroll
        for (int r = 0; r < 10; ++r) {
            uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
            uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
            uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
            c0 = n0; c1 = n1; c2 = n2; c3 = n3;
            k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
        }
        return make_uint4(c0, c1, c2, c3);
    }
};
__device__ __forceinline__ float u32_to_unit(uint32_t x) {   // (0, 1]
    return ((x >> 8) + 1) * (1.0f / 16777216.0f);
}
