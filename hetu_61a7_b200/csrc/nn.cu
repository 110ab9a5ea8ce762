// NN kernels: softmax, layer/batch norm (fwd+bwd), dropout, im2col/col2im, pooling.
// reference: src/ops/Softmax.cu (1 thread per row), CudnnSoftmax.cu, LayerNorm.cu
// (fwd 1 block/row; bwd = 3 kernels + 2 cudnnReduceTensor over 3 full-size
// workspaces), CudnnBn.cu, Dropout.cu (curand_init per thread), Conv2d.cu im2col,
// MaxPool.cu / AvgPool.cu.
#include "common.cuh"

// ------------------------------------------------------------------ softmax
// one warp per row when cols <= 1024 (row held in registers), else one CTA per row.
template <int PER_LANE>
__global__ void __launch_bounds__(256) softmax_rows_warp(const float* __restrict__ x, float* __restrict__ y,
                                                         int rows, int cols) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= rows) return;
    const float* p = x + (size_t)warp * cols;
    float v[PER_LANE];
    float m = -INFINITY;
#pragma unroll
    for (int i = 0; i < PER_LANE; ++i) {
        int j = lane + i * 32;
        v[i] = j < cols ? p[j] : -INFINITY;
        m = fmaxf(m, v[i]);
    }
    m = warp_max(m);
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < PER_LANE; ++i) {
        v[i] = (lane + i * 32 < cols) ? __expf(v[i] - m) : 0.f;
        s += v[i];
    }
    s = warp_sum(s);
    float inv = 1.f / s;
    float* q = y + (size_t)warp * cols;
#pragma unroll
    for (int i = 0; i < PER_LANE; ++i) {
        int j = lane + i * 32;
        if (j < cols) q[j] = v[i] * inv;
    }
}
__global__ void __launch_bounds__(512) softmax_rows_block(const float* __restrict__ x, float* __restrict__ y,
                                                          int rows, int cols) {
    __shared__ float sm[32];
    for (int row = blockIdx.x; row < rows; row += gridDim.x) {
        const float* p = x + (size_t)row * cols;
        float m = -INFINITY;
        for (int j = threadIdx.x; j < cols; j += blockDim.x) m = fmaxf(m, p[j]);
        m = block_max(m, sm);
        float s = 0.f;
        for (int j = threadIdx.x; j < cols; j += blockDim.x) s += __expf(p[j] - m);
        s = block_sum(s, sm);
        float inv = 1.f / s;
        float* q = y + (size_t)row * cols;
        for (int j = threadIdx.x; j < cols; j += blockDim.x) q[j] = __expf(p[j] - m) * inv;
    }
}
HETU_API int hetu_softmax_rows(const float* x, float* y, int rows, int cols, cudaStream_t s) {
    if (rows == 0) return 0;
    int grid = (rows * 32 + 255) / 256;
    if (cols <= 32) softmax_rows_warp<1><<<grid, 256, 0, s>>>(x, y, rows, cols);
    else if (cols <= 128) softmax_rows_warp<4><<<grid, 256, 0, s>>>(x, y, rows, cols);
    else if (cols <= 512) softmax_rows_warp<16><<<grid, 256, 0, s>>>(x, y, rows, cols);
    else if (cols <= 1024) softmax_rows_warp<32><<<grid, 256, 0, s>>>(x, y, rows, cols);
    else softmax_rows_block<<<rows < kNumSMs * 4 ? rows : kNumSMs * 4, 512, 0, s>>>(x, y, rows, cols);
    HETU_CHECK_LAUNCH("hetu_softmax_rows");
    return 0;
}

// dx = y * (dy - sum(dy*y))
__global__ void __launch_bounds__(256) softmax_grad_rows_kernel(const float* __restrict__ y, const float* __restrict__ dy,
                                                                float* __restrict__ dx, int rows, int cols) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= rows) return;
    const float* py = y + (size_t)warp * cols;
    const float* pd = dy + (size_t)warp * cols;
    float dot = 0.f;
    for (int j = lane; j < cols; j += 32) dot += py[j] * pd[j];
    dot = warp_sum(dot);
    float* q = dx + (size_t)warp * cols;
    for (int j = lane; j < cols; j += 32) q[j] = py[j] * (pd[j] - dot);
}
HETU_API int hetu_softmax_grad_rows(const float* y, const float* dy, float* dx, int rows, int cols, cudaStream_t s) {
    if (rows == 0) return 0;
    softmax_grad_rows_kernel<<<(rows * 32 + 255) / 256, 256, 0, s>>>(y, dy, dx, rows, cols);
    HETU_CHECK_LAUNCH("hetu_softmax_grad_rows");
    return 0;
}

// ---------------------------------------------------------------- layer norm
// One warp per row (any cols; row re-read from L1 for the second moment).  gamma/beta may
// be null (instance norm uses the same kernel with rows = N*C, cols = H*W).
__global__ void __launch_bounds__(256) layernorm_fwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                            const float* __restrict__ beta, float* __restrict__ y,
                                                            float* __restrict__ mean_out, float* __restrict__ rstd_out,
                                                            int rows, int cols, float eps) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= rows) return;
    const float* p = x + (size_t)warp * cols;
    float s = 0.f;
    for (int j = lane; j < cols; j += 32) s += p[j];
    float mean = warp_sum(s) / cols;
    float v = 0.f;
    for (int j = lane; j < cols; j += 32) { float d = p[j] - mean; v += d * d; }
    float rstd = rsqrtf(warp_sum(v) / cols + eps);
    float* q = y + (size_t)warp * cols;
    for (int j = lane; j < cols; j += 32) {
        float h = (p[j] - mean) * rstd;
        q[j] = gamma ? h * gamma[j] + (beta ? beta[j] : 0.f) : h;
    }
    if (lane == 0) { mean_out[warp] = mean; rstd_out[warp] = rstd; }
}
HETU_API int hetu_layernorm_fwd(const float* x, const float* gamma, const float* beta, float* y, float* mean,
                                float* rstd, int rows, int cols, float eps, cudaStream_t s) {
    if (rows == 0) return 0;
    layernorm_fwd_kernel<<<(rows * 32 + 255) / 256, 256, 0, s>>>(x, gamma, beta, y, mean, rstd, rows, cols, eps);
    HETU_CHECK_LAUNCH("hetu_layernorm_fwd");
    return 0;
}

// dX: one warp per row.  dgamma/dbeta: each CTA owns a slab of rows, accumulates per-column
// partials in registers (thread = column, coalesced) and adds them once with atomics.
__global__ void __launch_bounds__(256) layernorm_bwd_dx_kernel(const float* __restrict__ dy, const float* __restrict__ x,
                                                               const float* __restrict__ gamma, const float* __restrict__ mean,
                                                               const float* __restrict__ rstd, float* __restrict__ dx,
                                                               int rows, int cols) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= rows) return;
    const float* pd = dy + (size_t)warp * cols;
    const float* px = x + (size_t)warp * cols;
    float mu = mean[warp], rs = rstd[warp];
    float s1 = 0.f, s2 = 0.f;
    for (int j = lane; j < cols; j += 32) {
        float g = pd[j] * (gamma ? gamma[j] : 1.f);
        float h = (px[j] - mu) * rs;
        s1 += g;
        s2 += g * h;
    }
    s1 = warp_sum(s1) / cols;
    s2 = warp_sum(s2) / cols;
    float* q = dx + (size_t)warp * cols;
    for (int j = lane; j < cols; j += 32) {
        float g = pd[j] * (gamma ? gamma[j] : 1.f);
        float h = (px[j] - mu) * rs;
        q[j] = rs * (g - s1 - h * s2);
    }
}
__global__ void __launch_bounds__(256) layernorm_bwd_param_kernel(const float* __restrict__ dy, const float* __restrict__ x,
                                                                  const float* __restrict__ mean, const float* __restrict__ rstd,
                                                                  float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                                  int rows, int cols, int rows_per_cta) {
    int col = blockIdx.x * blockDim.x + threadIdx.x;
    if (col >= cols) return;
    int r0 = blockIdx.y * rows_per_cta, r1 = min(r0 + rows_per_cta, rows);
    float g = 0.f, b = 0.f;
    for (int r = r0; r < r1; ++r) {
        float d = dy[(size_t)r * cols + col];
        g += d * (x[(size_t)r * cols + col] - mean[r]) * rstd[r];
        b += d;
    }
    atomicAdd(dgamma + col, g);
    atomicAdd(dbeta + col, b);
}
HETU_API int hetu_layernorm_bwd(const float* dy, const float* x, const float* gamma, const float* mean,
                                const float* rstd, float* dx, float* dgamma, float* dbeta, int rows, int cols,
                                cudaStream_t s) {
    if (rows == 0) return 0;
    layernorm_bwd_dx_kernel<<<(rows * 32 + 255) / 256, 256, 0, s>>>(dy, x, gamma, mean, rstd, dx, rows, cols);
    if (dgamma) {
        cudaMemsetAsync(dgamma, 0, cols * sizeof(float), s);
        cudaMemsetAsync(dbeta, 0, cols * sizeof(float), s);
        int gx = (cols + 255) / 256;
        int gy = (kNumSMs * 2 + gx - 1) / gx;
        if (gy > rows) gy = rows;
        int rpc = (rows + gy - 1) / gy;
        gy = (rows + rpc - 1) / rpc;
        layernorm_bwd_param_kernel<<<dim3(gx, gy), 256, 0, s>>>(dy, x, mean, rstd, dgamma, dbeta, rows, cols, rpc);
    }
    HETU_CHECK_LAUNCH("hetu_layernorm_bwd");
    return 0;
}

// ---------------------------------------------------------------- batch norm (NCHW)
// one CTA per channel; stats over N*HW, then normalise.  CIFAR-size maps fit L2.
__global__ void __launch_bounds__(512) batchnorm_fwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                            const float* __restrict__ beta, float* __restrict__ y,
                                                            float* __restrict__ run_mean, float* __restrict__ run_var,
                                                            float* __restrict__ save_mean, float* __restrict__ save_rstd,
                                                            int N, int C, int HW, float momentum, float eps, int inference) {
    __shared__ float sm[32];
    int c = blockIdx.x;
    size_t cnt = (size_t)N * HW;
    float mean, rstd;
    if (inference) {
        mean = run_mean[c];
        rstd = rsqrtf(run_var[c] + eps);
    } else {
        float s = 0.f;
        for (size_t i = threadIdx.x; i < cnt; i += blockDim.x) {
            size_t n = i / HW, hw = i % HW;
            s += x[(n * C + c) * HW + hw];
        }
        mean = block_sum(s, sm) / cnt;
        float v = 0.f;
        for (size_t i = threadIdx.x; i < cnt; i += blockDim.x) {
            size_t n = i / HW, hw = i % HW;
            float d = x[(n * C + c) * HW + hw] - mean;
            v += d * d;
        }
        float var = block_sum(v, sm) / cnt;
        rstd = rsqrtf(var + eps);
        if (threadIdx.x == 0) {
            save_mean[c] = mean;
            save_rstd[c] = rstd;
            float unbiased = cnt > 1 ? var * cnt / (cnt - 1) : var;
            run_mean[c] = (1.f - momentum) * run_mean[c] + momentum * mean;
            run_var[c] = (1.f - momentum) * run_var[c] + momentum * unbiased;
        }
    }
    float g = gamma[c], b = beta[c];
    for (size_t i = threadIdx.x; i < cnt; i += blockDim.x) {
        size_t n = i / HW, hw = i % HW;
        size_t o = (n * C + c) * HW + hw;
        y[o] = (x[o] - mean) * rstd * g + b;
    }
}
// Cluster variant (training): a thread-block cluster of 8 CTAs owns one channel.  CTA r reduces images
// n = r, r+8, ... (contiguous HW runs: coalesced, no integer division per element), publishes its partial
// (sum, sum of squares) in shared memory, and after one cluster barrier every CTA reads the 8 partials
// through distributed shared memory and normalises the images it has just touched (still in L1/L2).
// One launch, C*8 CTAs instead of C, no workspace, no atomics.
constexpr int kBnCluster = 8;

__device__ __forceinline__ float2 block_sum2(float a, float b, float* sm /*64*/) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { a += __shfl_xor_sync(0xffffffffu, a, o); b += __shfl_xor_sync(0xffffffffu, b, o); }
    if (lane == 0) { sm[warp] = a; sm[32 + warp] = b; }
    __syncthreads();
    const int nw = blockDim.x >> 5;
    a = lane < nw ? sm[lane] : 0.f;
    b = lane < nw ? sm[32 + lane] : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { a += __shfl_xor_sync(0xffffffffu, a, o); b += __shfl_xor_sync(0xffffffffu, b, o); }
    __syncthreads();
    return make_float2(a, b);
}

__device__ __forceinline__ float2 cluster_sum2(float2 mine, float2* slot /*shared*/) {
    if (threadIdx.x == 0) *slot = mine;
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    float2 tot = make_float2(0.f, 0.f);
    const uint32_t local = (uint32_t)__cvta_generic_to_shared(slot);
#pragma unroll
    for (int r = 0; r < kBnCluster; ++r) {
        uint32_t remote;
        float a, b;
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local), "r"(r));
        asm volatile("ld.shared::cluster.v2.f32 {%0, %1}, [%2];" : "=f"(a), "=f"(b) : "r"(remote) : "memory");
        tot.x += a; tot.y += b;
    }
    // nobody may leave (and free its shared memory) while peers still read it
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    return tot;
}

__global__ void __cluster_dims__(kBnCluster, 1, 1) __launch_bounds__(512)
batchnorm_fwd_cluster_kernel(const float* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta,
                             float* __restrict__ y, float* __restrict__ run_mean, float* __restrict__ run_var,
                             float* __restrict__ save_mean, float* __restrict__ save_rstd, int N, int C, int HW,
                             float momentum, float eps) {
    __shared__ float sm[64];
    __shared__ float2 slot;
    const int c = blockIdx.x / kBnCluster, r = blockIdx.x % kBnCluster;
    float s = 0.f, ss = 0.f;
    for (int n = r; n < N; n += kBnCluster) {
        const float* xp = x + ((size_t)n * C + c) * HW;
        for (int i = threadIdx.x; i < HW; i += blockDim.x) { const float v = xp[i]; s += v; ss += v * v; }
    }
    const float2 tot = cluster_sum2(block_sum2(s, ss, sm), &slot);
    const float cnt = (float)N * HW;
    const float mean = tot.x / cnt;
    const float var = fmaxf(tot.y / cnt - mean * mean, 0.f);
    const float rstd = rsqrtf(var + eps);
    if (r == 0 && threadIdx.x == 0) {
        save_mean[c] = mean;
        save_rstd[c] = rstd;
        const float unbiased = cnt > 1.f ? var * cnt / (cnt - 1.f) : var;
        run_mean[c] = (1.f - momentum) * run_mean[c] + momentum * mean;
        run_var[c] = (1.f - momentum) * run_var[c] + momentum * unbiased;
    }
    const float g = gamma[c] * rstd, b = beta[c] - mean * rstd * gamma[c];
    for (int n = r; n < N; n += kBnCluster) {
        const size_t o = ((size_t)n * C + c) * HW;
        for (int i = threadIdx.x; i < HW; i += blockDim.x) y[o + i] = x[o + i] * g + b;
    }
}

__global__ void __cluster_dims__(kBnCluster, 1, 1) __launch_bounds__(512)
batchnorm_bwd_cluster_kernel(const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ gamma,
                             const float* __restrict__ save_mean, const float* __restrict__ save_rstd, float* __restrict__ dx,
                             float* __restrict__ dgamma, float* __restrict__ dbeta, int N, int C, int HW) {
    __shared__ float sm[64];
    __shared__ float2 slot;
    const int c = blockIdx.x / kBnCluster, r = blockIdx.x % kBnCluster;
    const float mean = save_mean[c], rstd = save_rstd[c];
    float sb = 0.f, sg = 0.f;
    for (int n = r; n < N; n += kBnCluster) {
        const size_t o = ((size_t)n * C + c) * HW;
        for (int i = threadIdx.x; i < HW; i += blockDim.x) {
            const float d = dy[o + i];
            sb += d;
            sg += d * (x[o + i] - mean) * rstd;
        }
    }
    const float2 tot = cluster_sum2(block_sum2(sb, sg, sm), &slot);
    if (r == 0 && threadIdx.x == 0) { dbeta[c] = tot.x; dgamma[c] = tot.y; }
    const float cnt = (float)N * HW;
    const float k = gamma[c] * rstd / cnt;
    for (int n = r; n < N; n += kBnCluster) {
        const size_t o = ((size_t)n * C + c) * HW;
        for (int i = threadIdx.x; i < HW; i += blockDim.x) {
            const float h = (x[o + i] - mean) * rstd;
            dx[o + i] = k * (cnt * dy[o + i] - tot.x - h * tot.y);
        }
    }
}

HETU_API int hetu_batchnorm_fwd(const float* x, const float* gamma, const float* beta, float* y, float* run_mean,
                                float* run_var, float* save_mean, float* save_rstd, int N, int C, int HW,
                                float momentum, float eps, int inference, cudaStream_t s) {
    if (C == 0) return 0;
    if (!inference && N >= kBnCluster) {
        const int threads = HW >= 512 ? 512 : (HW >= 128 ? 256 : 128);
        batchnorm_fwd_cluster_kernel<<<C * kBnCluster, threads, 0, s>>>(x, gamma, beta, y, run_mean, run_var, save_mean,
                                                                       save_rstd, N, C, HW, momentum, eps);
        HETU_CHECK_LAUNCH("hetu_batchnorm_fwd(cluster)");
        return 0;
    }
    batchnorm_fwd_kernel<<<C, 512, 0, s>>>(x, gamma, beta, y, run_mean, run_var, save_mean, save_rstd, N, C, HW,
                                           momentum, eps, inference);
    HETU_CHECK_LAUNCH("hetu_batchnorm_fwd");
    return 0;
}
__global__ void __launch_bounds__(512) batchnorm_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ x,
                                                            const float* __restrict__ gamma, const float* __restrict__ save_mean,
                                                            const float* __restrict__ save_rstd, float* __restrict__ dx,
                                                            float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                            int N, int C, int HW) {
    __shared__ float sm[32];
    int c = blockIdx.x;
    size_t cnt = (size_t)N * HW;
    float mean = save_mean[c], rstd = save_rstd[c];
    float sb = 0.f, sg = 0.f;
    for (size_t i = threadIdx.x; i < cnt; i += blockDim.x) {
        size_t n = i / HW, hw = i % HW;
        size_t o = (n * C + c) * HW + hw;
        float d = dy[o];
        sb += d;
        sg += d * (x[o] - mean) * rstd;
    }
    sb = block_sum(sb, sm);
    sg = block_sum(sg, sm);
    if (threadIdx.x == 0) { dgamma[c] = sg; dbeta[c] = sb; }
    float k = gamma[c] * rstd / cnt;
    for (size_t i = threadIdx.x; i < cnt; i += blockDim.x) {
        size_t n = i / HW, hw = i % HW;
        size_t o = (n * C + c) * HW + hw;
        float h = (x[o] - mean) * rstd;
        dx[o] = k * (cnt * dy[o] - sb - h * sg);
    }
}
HETU_API int hetu_batchnorm_bwd(const float* dy, const float* x, const float* gamma, const float* save_mean,
                                const float* save_rstd, float* dx, float* dgamma, float* dbeta, int N, int C, int HW,
                                cudaStream_t s) {
    if (C == 0) return 0;
    if (N >= kBnCluster) {
        const int threads = HW >= 512 ? 512 : (HW >= 128 ? 256 : 128);
        batchnorm_bwd_cluster_kernel<<<C * kBnCluster, threads, 0, s>>>(dy, x, gamma, save_mean, save_rstd, dx, dgamma, dbeta,
                                                                       N, C, HW);
        HETU_CHECK_LAUNCH("hetu_batchnorm_bwd(cluster)");
        return 0;
    }
    batchnorm_bwd_kernel<<<C, 512, 0, s>>>(dy, x, gamma, save_mean, save_rstd, dx, dgamma, dbeta, N, C, HW);
    HETU_CHECK_LAUNCH("hetu_batchnorm_bwd");
    return 0;
}

// ------------------------------------------------------------------- dropout
// y = x * mask / (1-p); mask from Philox(seed ^ step)(element/4).  The same call with the
// same (seed, step) regenerates the mask in backward - nothing is stored.
__global__ void __launch_bounds__(256) dropout_kernel(const float* __restrict__ x, float* __restrict__ y, size_t n,
                                                      float p, unsigned long long seed, const long long* __restrict__ step) {
    Philox rng(seed ^ ((unsigned long long)(*step) * 0x9E3779B97F4A7C15ull));
    float scale = 1.f / (1.f - p);
    size_t n4 = (n + 3) >> 2;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        uint4 r = rng(i);
        uint32_t rr[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            size_t j = i * 4 + k;
            if (j < n) y[j] = (u32_to_unit(rr[k]) > p) ? x[j] * scale : 0.f;
        }
    }
}
HETU_API int hetu_dropout(const float* x, float* y, size_t n, float p, unsigned long long seed, const void* step,
                          cudaStream_t s) {
    if (n == 0) return 0;
    dropout_kernel<<<ew_grid((n + 3) / 4, 256), 256, 0, s>>>(x, y, n, p, seed, (const long long*)step);
    HETU_CHECK_LAUNCH("hetu_dropout");
    return 0;
}

// ------------------------------------------------------------ im2col / col2im
// col[(n*OH+oh)*OW+ow][(c*kh+ky)*kw+kx] = x[n][c][oh*sh-ph+ky][ow*sw-pw+kx]
__global__ void __launch_bounds__(256) im2col_kernel(const float* __restrict__ x, float* __restrict__ col, int N, int C,
                                                     int H, int W, int kh, int kw, int ph, int pw, int sh, int sw,
                                                     int OH, int OW) {
    size_t K = (size_t)C * kh * kw;
    size_t total = (size_t)N * OH * OW * K;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        size_t row = i / K;
        int kk = (int)(i % K);
        int kx = kk % kw, ky = (kk / kw) % kh, c = kk / (kw * kh);
        int ow = (int)(row % OW), oh = (int)((row / OW) % OH), n = (int)(row / ((size_t)OW * OH));
        int h = oh * sh - ph + ky, w = ow * sw - pw + kx;
        col[i] = (h >= 0 && h < H && w >= 0 && w < W) ? x[(((size_t)n * C + c) * H + h) * W + w] : 0.f;
    }
}
// Tiled variant: one CTA per (image, output row).  The kh input rows of every channel that this output
// row touches are staged in shared memory with coalesced row reads, then the OW x K patch rows are
// written out coalesced (K fastest).  All index math is 32-bit.
__global__ void __launch_bounds__(256) im2col_tiled_kernel(const float* __restrict__ x, float* __restrict__ col, int C, int H,
                                                           int W, int kh, int kw, int ph, int pw, int sh, int sw, int OH,
                                                           int OW) {
    extern __shared__ float im2col_smem[];               // [C][kh][W]
    const int n = blockIdx.x / OH, oh = blockIdx.x - n * OH;
    const int K = C * kh * kw, rows = C * kh;
    const float* xn = x + (size_t)n * C * H * W;
    for (int i = threadIdx.x; i < rows * W; i += 256) {
        const int r = i / W, w = i - r * W;
        const int c = r / kh, ky = r - c * kh;
        const int h = oh * sh - ph + ky;
        im2col_smem[i] = (h >= 0 && h < H) ? __ldg(xn + ((size_t)c * H + h) * W + w) : 0.f;
    }
    __syncthreads();
    float* out = col + (size_t)blockIdx.x * OW * K;
    const int khw = kh * kw;
    for (int i = threadIdx.x; i < OW * K; i += 256) {
        const int ow = i / K, kk = i - ow * K;
        const int c = kk / khw, rem = kk - c * khw;
        const int ky = rem / kw, kx = rem - ky * kw;
        const int w = ow * sw - pw + kx;
        out[i] = (w >= 0 && w < W) ? im2col_smem[(c * kh + ky) * W + w] : 0.f;
    }
}

HETU_API int hetu_im2col(const float* x, float* col, int N, int C, int H, int W, int kh, int kw, int ph, int pw,
                         int sh, int sw, int OH, int OW, cudaStream_t s) {
    size_t total = (size_t)N * OH * OW * C * kh * kw;
    if (total == 0) return 0;
    const size_t smem = (size_t)C * kh * W * sizeof(float);
    if (smem <= 48 * 1024 && (size_t)OW * C * kh * kw < (1u << 30)) {
        im2col_tiled_kernel<<<N * OH, 256, smem, s>>>(x, col, C, H, W, kh, kw, ph, pw, sh, sw, OH, OW);
        HETU_CHECK_LAUNCH("hetu_im2col(tiled)");
        return 0;
    }
    im2col_kernel<<<ew_grid(total, 256, 16), 256, 0, s>>>(x, col, N, C, H, W, kh, kw, ph, pw, sh, sw, OH, OW);
    HETU_CHECK_LAUNCH("hetu_im2col");
    return 0;
}
// gather form (no atomics): each input pixel sums the col entries that were copied from it
__global__ void __launch_bounds__(256) col2im_kernel(const float* __restrict__ col, float* __restrict__ x, int N, int C,
                                                     int H, int W, int kh, int kw, int ph, int pw, int sh, int sw,
                                                     int OH, int OW) {
    size_t K = (size_t)C * kh * kw;
    size_t total = (size_t)N * C * H * W;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        int w = (int)(i % W), h = (int)((i / W) % H), c = (int)((i / ((size_t)W * H)) % C), n = (int)(i / ((size_t)W * H * C));
        float acc = 0.f;
        for (int ky = 0; ky < kh; ++ky) {
            int t = h + ph - ky;
            if (t < 0 || t % sh) continue;
            int oh = t / sh;
            if (oh >= OH) continue;
            for (int kx = 0; kx < kw; ++kx) {
                int u = w + pw - kx;
                if (u < 0 || u % sw) continue;
                int ow = u / sw;
                if (ow >= OW) continue;
                acc += col[(((size_t)n * OH + oh) * OW + ow) * K + ((size_t)c * kh + ky) * kw + kx];
            }
        }
        x[i] = acc;
    }
}
HETU_API int hetu_col2im(const float* col, float* x, int N, int C, int H, int W, int kh, int kw, int ph, int pw,
                         int sh, int sw, int OH, int OW, cudaStream_t s) {
    size_t total = (size_t)N * C * H * W;
    if (total == 0) return 0;
    col2im_kernel<<<ew_grid(total, 256, 16), 256, 0, s>>>(col, x, N, C, H, W, kh, kw, ph, pw, sh, sw, OH, OW);
    HETU_CHECK_LAUNCH("hetu_col2im");
    return 0;
}

// same gather from the transposed layout colT [N, C*kh*kw, OH*OW] (what the batched GEMM W^T dY_n writes when dY stays
// NCHW): neighbouring threads (w, w+1) read neighbouring ow of the same row, so the reads coalesce - the [N*OH*OW, C*kh*kw]
// layout above makes every thread of a warp touch its own 32-byte sector
// grid: x = pixel blocks of one (image, channel) plane, y = plane.  All index arithmetic is 32-bit and the stride is a
// template parameter: the generic kernel above spends ~400 instructions per element on 64-bit div / mod and `% stride`,
// which - not memory - was its limit (235 us for a 302 MB buffer).
template <int S>
__global__ void __launch_bounds__(256) col2im_t_kernel(const float* __restrict__ colT, float* __restrict__ x, int H, int W, int kh,
                                                       int kw, int ph, int pw, int OH, int OW) {
    const int P = OH * OW, HW = H * W;
    const size_t nc = blockIdx.y;
    const float* base = colT + nc * (size_t)(kh * kw) * P;            // rows (c, ky, kx) of image n
    float* out = x + nc * HW;
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < HW; p += gridDim.x * blockDim.x) {
        const int h = p / W, w = p - h * W;
        float acc = 0.f;
        for (int ky = 0; ky < kh; ++ky) {
            const int t = h + ph - ky;
            if (t < 0 || (S > 1 && (t % S))) continue;
            const int oh = t / S;
            if (oh >= OH) continue;
            const float* row = base + (size_t)(ky * kw) * P + oh * OW;
            for (int kx = 0; kx < kw; ++kx) {
                const int u = w + pw - kx;
                if (u < 0 || (S > 1 && (u % S))) continue;
                const int ow = u / S;
                if (ow >= OW) continue;
                acc += row[(size_t)kx * P + ow];
            }
        }
        out[p] = acc;
    }
}
HETU_API int hetu_col2im_t(const float* colT, float* x, int N, int C, int H, int W, int kh, int kw, int ph, int pw,
                           int sh, int sw, int OH, int OW, cudaStream_t s) {
    if ((size_t)N * C * H * W == 0) return 0;
    if (sh != sw || sh < 1 || sh > 2 || (long long)N * C > 2147483647LL / 4) {
        hetu_set_error("col2im_t: stride (%d, %d) not supported", sh, sw);
        return -1;
    }
    const int HW = H * W;
    dim3 grid((HW + 255) / 256 > 64 ? 64 : (HW + 255) / 256, N * C);
    if (grid.y > 65535) { hetu_set_error("col2im_t: %d planes > 65535", N * C); return -1; }
    if (sh == 1) col2im_t_kernel<1><<<grid, 256, 0, s>>>(colT, x, H, W, kh, kw, ph, pw, OH, OW);
    else col2im_t_kernel<2><<<grid, 256, 0, s>>>(colT, x, H, W, kh, kw, ph, pw, OH, OW);
    HETU_CHECK_LAUNCH("hetu_col2im_t");
    return 0;
}

// ------------------------------------------------------------------- pooling
__global__ void __launch_bounds__(256) pool2d_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int NC, int H,
                                                         int W, int kh, int kw, int pad, int stride, int OH, int OW, int is_max) {
    size_t total = (size_t)NC * OH * OW;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        int ow = (int)(i % OW), oh = (int)((i / OW) % OH);
        size_t nc = i / ((size_t)OW * OH);
        const float* p = x + nc * H * W;
        float acc = is_max ? -INFINITY : 0.f;
        for (int ky = 0; ky < kh; ++ky) {
            int h = oh * stride - pad + ky;
            if (h < 0 || h >= H) continue;
            for (int kx = 0; kx < kw; ++kx) {
                int w = ow * stride - pad + kx;
                if (w < 0 || w >= W) continue;
                float v = p[h * W + w];
                acc = is_max ? fmaxf(acc, v) : acc + v;
            }
        }
        y[i] = is_max ? acc : acc / (kh * kw);
    }
}
HETU_API int hetu_pool2d_fwd(const float* x, float* y, int NC, int H, int W, int kh, int kw, int pad, int stride,
                             int OH, int OW, int is_max, cudaStream_t s) {
    size_t total = (size_t)NC * OH * OW;
    if (total == 0) return 0;
    pool2d_fwd_kernel<<<ew_grid(total, 256), 256, 0, s>>>(x, y, NC, H, W, kh, kw, pad, stride, OH, OW, is_max);
    HETU_CHECK_LAUNCH("hetu_pool2d_fwd");
    return 0;
}
// gather form: each input pixel collects from the windows that contain it
__global__ void __launch_bounds__(256) pool2d_bwd_kernel(const float* __restrict__ x, const float* __restrict__ y,
                                                         const float* __restrict__ dy, float* __restrict__ dx, int NC, int H,
                                                         int W, int kh, int kw, int pad, int stride, int OH, int OW, int is_max) {
    size_t total = (size_t)NC * H * W;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        int w = (int)(i % W), h = (int)((i / W) % H);
        size_t nc = i / ((size_t)W * H);
        float xv = x[i], acc = 0.f;
        for (int ky = 0; ky < kh; ++ky) {
            int t = h + pad - ky;
            if (t < 0 || t % stride) continue;
            int oh = t / stride;
            if (oh >= OH) continue;
            for (int kx = 0; kx < kw; ++kx) {
                int u = w + pad - kx;
                if (u < 0 || u % stride) continue;
                int ow = u / stride;
                if (ow >= OW) continue;
                size_t o = (nc * OH + oh) * OW + ow;
                if (is_max) { if (y[o] == xv) acc += dy[o]; }
                else acc += dy[o] / (kh * kw);
            }
        }
        dx[i] = acc;
    }
}
HETU_API int hetu_pool2d_bwd(const float* x, const float* y, const float* dy, float* dx, int NC, int H, int W, int kh,
                             int kw, int pad, int stride, int OH, int OW, int is_max, cudaStream_t s) {
    size_t total = (size_t)NC * H * W;
    if (total == 0) return 0;
    pool2d_bwd_kernel<<<ew_grid(total, 256), 256, 0, s>>>(x, y, dy, dx, NC, H, W, kh, kw, pad, stride, OH, OW, is_max);
    HETU_CHECK_LAUNCH("hetu_pool2d_bwd");
    return 0;
}
