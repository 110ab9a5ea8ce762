// Loss kernels: fused softmax-cross-entropy (dense / sparse labels) and BCE.
// reference: src/ops/SoftmaxCrossEntropy.cu:5-30 (one thread per row, serial over the
// classes), CudnnSoftmaxEntropy.cu (cudnnSoftmax + kernel + cudnnReduceTensor),
// SoftmaxCrossEntropySparse.cu, BinaryCrossEntropy.cu.
// One warp per row, online max/sum in a single pass over the logits.
#include "common.cuh"

__device__ __forceinline__ void online_lse(const float* __restrict__ p, int cols, int lane, float& m, float& s) {
    m = -INFINITY; s = 0.f;
    for (int j = lane; j < cols; j += 32) {
        float v = p[j];
        float nm = fmaxf(m, v);
        s = s * __expf(m - nm) + __expf(v - nm);
        m = nm;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        float om = __shfl_xor_sync(0xffffffffu, m, o);
        float os = __shfl_xor_sync(0xffffffffu, s, o);
        float nm = fmaxf(m, om);
        s = (m == -INFINITY ? 0.f : s * __expf(m - nm)) + (om == -INFINITY ? 0.f : os * __expf(om - nm));
        m = nm;
    }
}

// loss[r] = sum_c y[r,c] * (lse - x[r,c])
__global__ void __launch_bounds__(256) softmax_xent_kernel(const float* __restrict__ x, const float* __restrict__ y,
                                                           float* __restrict__ loss, int rows, int cols) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= rows) return;
    const float* p = x + (size_t)warp * cols;
    const float* q = y + (size_t)warp * cols;
    float m, s;
    online_lse(p, cols, lane, m, s);
    float lse = m + __logf(s);
    float acc = 0.f;
    for (int j = lane; j < cols; j += 32) acc += q[j] * (lse - p[j]);
    acc = warp_sum(acc);
    if (lane == 0) loss[warp] = acc;
}
HETU_API int hetu_softmax_xent(const float* x, const float* y, float* loss, int rows, int cols, cudaStream_t s) {
    if (rows == 0) return 0;
    softmax_xent_kernel<<<(rows * 32 + 255) / 256, 256, 0, s>>>(x, y, loss, rows, cols);
    HETU_CHECK_LAUNCH("hetu_softmax_xent");
    return 0;
}
// dx[r,c] = (softmax(x)[r,c] * sum_c y[r,:] - y[r,c]) * g[r]
__global__ void __launch_bounds__(256) softmax_xent_grad_kernel(const float* __restrict__ x, const float* __restrict__ y,
                                                                const float* __restrict__ g, float* __restrict__ dx,
                                                                int rows, int cols) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= rows) return;
    const float* p = x + (size_t)warp * cols;
    const float* q = y + (size_t)warp * cols;
    float m, s;
    online_lse(p, cols, lane, m, s);
    float ysum = 0.f;
    for (int j = lane; j < cols; j += 32) ysum += q[j];
    ysum = warp_sum(ysum);
    float inv = 1.f / s, gr = g[warp];
    float* o = dx + (size_t)warp * cols;
    for (int j = lane; j < cols; j += 32) o[j] = (__expf(p[j] - m) * inv * ysum - q[j]) * gr;
}
HETU_API int hetu_softmax_xent_grad(const float* x, const float* y, const float* g, float* dx, int rows, int cols,
                                    cudaStream_t s) {
    if (rows == 0) return 0;
    softmax_xent_grad_kernel<<<(rows * 32 + 255) / 256, 256, 0, s>>>(x, y, g, dx, rows, cols);
    HETU_CHECK_LAUNCH("hetu_softmax_xent_grad");
    return 0;
}

__global__ void __launch_bounds__(256) softmax_xent_sparse_kernel(const float* __restrict__ x, const float* __restrict__ lab,
                                                                  float* __restrict__ loss, int rows, int cols, int ignored) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= rows) return;
    int t = (int)lab[warp];
    if (t == ignored || t < 0) { if (lane == 0) loss[warp] = 0.f; return; }
    const float* p = x + (size_t)warp * cols;
    float m, s;
    online_lse(p, cols, lane, m, s);
    if (lane == 0) loss[warp] = m + __logf(s) - p[t];
}
HETU_API int hetu_softmax_xent_sparse(const float* x, const float* lab, float* loss, int rows, int cols, int ignored,
                                      cudaStream_t s) {
    if (rows == 0) return 0;
    softmax_xent_sparse_kernel<<<(rows * 32 + 255) / 256, 256, 0, s>>>(x, lab, loss, rows, cols, ignored);
    HETU_CHECK_LAUNCH("hetu_softmax_xent_sparse");
    return 0;
}
__global__ void __launch_bounds__(256) softmax_xent_sparse_grad_kernel(const float* __restrict__ x, const float* __restrict__ lab,
                                                                       const float* __restrict__ g, float* __restrict__ dx,
                                                                       int rows, int cols, int ignored) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= rows) return;
    int t = (int)lab[warp];
    float* o = dx + (size_t)warp * cols;
    if (t == ignored || t < 0) { for (int j = lane; j < cols; j += 32) o[j] = 0.f; return; }
    const float* p = x + (size_t)warp * cols;
    float m, s;
    online_lse(p, cols, lane, m, s);
    float inv = 1.f / s, gr = g[warp];
    for (int j = lane; j < cols; j += 32) o[j] = (__expf(p[j] - m) * inv - (j == t ? 1.f : 0.f)) * gr;
}
HETU_API int hetu_softmax_xent_sparse_grad(const float* x, const float* lab, const float* g, float* dx, int rows,
                                           int cols, int ignored, cudaStream_t s) {
    if (rows == 0) return 0;
    softmax_xent_sparse_grad_kernel<<<(rows * 32 + 255) / 256, 256, 0, s>>>(x, lab, g, dx, rows, cols, ignored);
    HETU_CHECK_LAUNCH("hetu_softmax_xent_sparse_grad");
    return 0;
}

__global__ void __launch_bounds__(256) bce_kernel(const float* __restrict__ p, const float* __restrict__ y,
                                                  float* __restrict__ o, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        o[i] = -y[i] * logf(p[i]) - (1.f - y[i]) * logf(1.f - p[i]);
}
HETU_API int hetu_bce(const float* p, const float* y, float* o, size_t n, cudaStream_t s) {
    if (n == 0) return 0;
    bce_kernel<<<ew_grid(n, 256), 256, 0, s>>>(p, y, o, n);
    HETU_CHECK_LAUNCH("hetu_bce");
    return 0;
}
__global__ void __launch_bounds__(256) bce_grad_kernel(const float* __restrict__ p, const float* __restrict__ y,
                                                       const float* __restrict__ g, float* __restrict__ o, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        o[i] = g[i] * (-y[i] / p[i] + (1.f - y[i]) / (1.f - p[i]));
}
HETU_API int hetu_bce_grad(const float* p, const float* y, const float* g, float* o, size_t n, cudaStream_t s) {
    if (n == 0) return 0;
    bce_grad_kernel<<<ew_grid(n, 256), 256, 0, s>>>(p, y, g, o, n);
    HETU_CHECK_LAUNCH("hetu_bce_grad");
    return 0;
}
