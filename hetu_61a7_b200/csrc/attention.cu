// Attention softmax kernels (scores stay in one buffer; scale, additive key mask, softmax and the
// log-sum-exp needed by the backward pass are one pass over each row).
//
// reference: examples/nlp/bert/hetu_bert.py:249-264 builds attention from five graph ops
// (batch_matmul, broadcast mask add, softmax, dropout, batch_matmul).  Here the two GEMMs run on the
// tcgen05 batched GEMM and everything between them is hetu_attn_softmax / hetu_attn_softmax_bwd.
#include "common.cuh"

// one warp per row (rows = B*H*Sq, cols = Sk); in place: s <- softmax(s * scale + mask[b, :])
__global__ void __launch_bounds__(256) attn_softmax_kernel(float* __restrict__ s, const float* __restrict__ mask,
                                                           float* __restrict__ lse, long long rows, int cols,
                                                           long long rows_per_batch, float scale) {
    const long long row = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    float* r = s + row * cols;
    const float* m = mask ? mask + (row / rows_per_batch) * cols : nullptr;
    float mx = -INFINITY;
    for (int c = lane; c < cols; c += 32) {
        float v = r[c] * scale + (m ? __ldg(m + c) : 0.f);
        r[c] = v;
        mx = fmaxf(mx, v);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float sum = 0.f;
    for (int c = lane; c < cols; c += 32) {
        float e = __expf(r[c] - mx);
        r[c] = e;
        sum += e;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float inv = 1.f / sum;
    for (int c = lane; c < cols; c += 32) r[c] *= inv;
    if (lane == 0 && lse) lse[row] = mx + __logf(sum);
}

// in place on dp: ds = p * (dp - sum_c(dp * p)) * scale
__global__ void __launch_bounds__(256) attn_softmax_bwd_kernel(const float* __restrict__ p, float* __restrict__ dp,
                                                               long long rows, int cols, float scale) {
    const long long row = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    const float* pr = p + row * cols;
    float* dr = dp + row * cols;
    float dot = 0.f;
    for (int c = lane; c < cols; c += 32) dot += pr[c] * dr[c];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
    for (int c = lane; c < cols; c += 32) dr[c] = pr[c] * (dr[c] - dot) * scale;
}

// backward after the flash forward (csrc/attention_sm100.cu keeps no probabilities): s holds raw scores q k^T on entry,
// s <- exp(s * scale + mask[b, :] - lse[row]); one warp per row
__global__ void __launch_bounds__(256) attn_probs_from_lse_kernel(float* __restrict__ s, const float* __restrict__ mask,
                                                                  const float* __restrict__ lse, long long rows, int cols,
                                                                  long long rows_per_batch, float scale) {
    const long long row = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    float* r = s + row * cols;
    const float* m = mask ? mask + (row / rows_per_batch) * cols : nullptr;
    const float l = lse[row];
    for (int c = lane; c < cols; c += 32) {
        const float v = r[c] * scale + (m ? __ldg(m + c) : 0.f);
        r[c] = (l == -INFINITY) ? 0.f : __expf(v - l);
    }
}

HETU_API int hetu_attn_probs_from_lse(float* s, const float* mask, const float* lse, long long rows, int cols,
                                      long long rows_per_batch, float scale, cudaStream_t st) {
    if (rows <= 0 || cols <= 0) return 0;
    attn_probs_from_lse_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(s, mask, lse, rows, cols, rows_per_batch, scale);
    HETU_CHECK_LAUNCH("hetu_attn_probs_from_lse");
    return 0;
}

HETU_API int hetu_attn_softmax(float* s, const float* mask, float* lse, long long rows, int cols, long long rows_per_batch,
                               float scale, cudaStream_t st) {
    if (rows <= 0 || cols <= 0) return 0;
    attn_softmax_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(s, mask, lse, rows, cols, rows_per_batch, scale);
    HETU_CHECK_LAUNCH("hetu_attn_softmax");
    return 0;
}

HETU_API int hetu_attn_softmax_bwd(const float* p, float* dp, long long rows, int cols, float scale, cudaStream_t st) {
    if (rows <= 0 || cols <= 0) return 0;
    attn_softmax_bwd_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(p, dp, rows, cols, scale);
    HETU_CHECK_LAUNCH("hetu_attn_softmax_bwd");
    return 0;
}
