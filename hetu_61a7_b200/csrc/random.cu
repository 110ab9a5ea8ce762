// Parameter initialisation with counter-based Philox (constant / uniform / normal /
// truncated normal).  reference: src/ops/Initializers.cu (curand Philox per thread).
// Element i always gets the same draw for a given seed, independent of the launch
// geometry, so sharded and unsharded initialisation agree.
#include "common.cuh"

enum { K_CONST = 0, K_UNIFORM, K_NORMAL, K_TRUNC };

__device__ __forceinline__ void box_muller(uint32_t a, uint32_t b, float& z0, float& z1) {
    float u1 = u32_to_unit(a), u2 = u32_to_unit(b);
    float r = sqrtf(-2.f * __logf(u1));
    float s, c;
    __sincosf(6.283185307179586f * u2, &s, &c);
    z0 = r * c;
    z1 = r * s;
}

__global__ void __launch_bounds__(256) init_random_kernel(float* __restrict__ out, size_t n, int kind, float a, float b,
                                                          unsigned long long seed) {
    Philox rng(seed);
    size_t n4 = (n + 3) >> 2;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        float v[4];
        if (kind == K_CONST) {
            v[0] = v[1] = v[2] = v[3] = a;
        } else if (kind == K_UNIFORM) {
            uint4 r = rng(i);
            v[0] = a + (b - a) * u32_to_unit(r.x); v[1] = a + (b - a) * u32_to_unit(r.y);
            v[2] = a + (b - a) * u32_to_unit(r.z); v[3] = a + (b - a) * u32_to_unit(r.w);
        } else {
            uint4 r = rng(i);
            box_muller(r.x, r.y, v[0], v[1]);
            box_muller(r.z, r.w, v[2], v[3]);
            if (kind == K_TRUNC) {
                // resample values outside [-2, 2] from further Philox sub-streams
                for (int k = 0; k < 4; ++k) {
                    unsigned long long sub = 1;
                    while (fabsf(v[k]) > 2.f && sub < 64) {
                        uint4 q = rng(i, sub * 4 + k);
                        float z0, z1;
                        box_muller(q.x, q.y, z0, z1);
                        v[k] = fabsf(z0) <= 2.f ? z0 : z1;
                        ++sub;
                    }
                    v[k] = fminf(fmaxf(v[k], -2.f), 2.f);
                }
            }
            for (int k = 0; k < 4; ++k) v[k] = a + b * v[k];
        }
        size_t j = i * 4;
        if (j + 3 < n && (reinterpret_cast<uintptr_t>(out) & 15) == 0) {
            *reinterpret_cast<float4*>(out + j) = make_float4(v[0], v[1], v[2], v[3]);
        } else {
            for (int k = 0; k < 4 && j + k < n; ++k) out[j + k] = v[k];
        }
    }
}

HETU_API int hetu_init_random(float* out, size_t n, int kind, float a, float b, unsigned long long seed, cudaStream_t s) {
    if (n == 0) return 0;
    init_random_kernel<<<ew_grid((n + 3) / 4, 256), 256, 0, s>>>(out, n, kind, a, b, seed);
    HETU_CHECK_LAUNCH("hetu_init_random");
    return 0;
}
