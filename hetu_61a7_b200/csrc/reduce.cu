// Reductions over a contiguous run of axes, arg-max, row top-k, column scan.
// reference: src/ops/ReduceSum.cu, ReduceSumAxisZero.cu (tree reduce + workspace),
// Max.cu/Min.cu, Argmax.cu, TopKIdx.cu:3-26 (1 thread per row, k serial scans over
// a 2048-float local array), CumSum.cu:3-18 (1 thread per column, serial over T).
#include "common.cuh"

enum { R_SUM = 0, R_MEAN, R_MAX, R_MIN, R_SQSUM };

template <int MODE> __device__ __forceinline__ float r_init() {
    return MODE == R_MAX ? -INFINITY : (MODE == R_MIN ? INFINITY : 0.f);
}
template <int MODE> __device__ __forceinline__ float r_op(float acc, float v) {
    if (MODE == R_MAX) return fmaxf(acc, v);
    if (MODE == R_MIN) return fminf(acc, v);
    if (MODE == R_SQSUM) return acc + v * v;
    return acc + v;
}
template <int MODE> __device__ __forceinline__ float r_merge(float a, float b) {
    if (MODE == R_MAX) return fmaxf(a, b);
    if (MODE == R_MIN) return fminf(a, b);
    return a + b;
}
template <int MODE> __device__ __forceinline__ float r_warp(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = r_merge<MODE>(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// B == 1: x viewed [A, R]; one warp per row (R <= 2048) or one CTA per row.
template <int MODE>
__global__ void __launch_bounds__(256) reduce_rows_warp(const float* __restrict__ x, float* __restrict__ out,
                                                        size_t A, size_t R) {
    size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t row = warp; row < A; row += nwarp) {
        const float* p = x + row * R;
        float acc = r_init<MODE>();
        for (size_t j = lane; j < R; j += 32) acc = r_op<MODE>(acc, p[j]);
        acc = r_warp<MODE>(acc);
        if (lane == 0) out[row] = (MODE == R_MEAN) ? acc / (float)R : acc;
    }
}
template <int MODE>
__global__ void __launch_bounds__(512) reduce_rows_block(const float* __restrict__ x, float* __restrict__ out,
                                                         size_t A, size_t R) {
    __shared__ float sm[32];
    for (size_t row = blockIdx.x; row < A; row += gridDim.x) {
        const float* p = x + row * R;
        float acc = r_init<MODE>();
        for (size_t j = threadIdx.x; j < R; j += blockDim.x) acc = r_op<MODE>(acc, p[j]);
        acc = r_warp<MODE>(acc);
        __syncthreads();
        if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = acc;
        __syncthreads();
        if (threadIdx.x < 32) {
            float v = threadIdx.x < (blockDim.x >> 5) ? sm[threadIdx.x] : r_init<MODE>();
            v = r_warp<MODE>(v);
            if (threadIdx.x == 0) out[row] = (MODE == R_MEAN) ? v / (float)R : v;
        }
    }
}
// B > 1: x viewed [A, R, B]; thread owns column b (coalesced across threads), loops over R.
// For tall-skinny cases (R >> A*B) the R loop is split across blockIdx.y with atomics for sums.
template <int MODE>
__global__ void __launch_bounds__(256) reduce_mid_kernel(const float* __restrict__ x, float* __restrict__ out,
                                                         size_t A, size_t R, size_t B, int rsplit) {
    size_t ab = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (ab >= A * B) return;
    size_t a = ab / B, b = ab % B;
    size_t chunk = (R + rsplit - 1) / rsplit;
    size_t r0 = (size_t)blockIdx.y * chunk, r1 = r0 + chunk < R ? r0 + chunk : R;
    const float* p = x + a * R * B + b;
    float acc = r_init<MODE>();
    for (size_t r = r0; r < r1; ++r) acc = r_op<MODE>(acc, p[r * B]);
    if (rsplit == 1) out[ab] = (MODE == R_MEAN) ? acc / (float)R : acc;
    else atomicAdd(out + ab, (MODE == R_MEAN) ? acc / (float)R : acc);
}

template <int MODE>
static int launch_reduce(const float* x, float* out, size_t A, size_t R, size_t B, cudaStream_t s) {
    if (B == 1) {
        if (R <= 2048) {
            size_t threads = A * 32;
            int grid = ew_grid(threads, 256);
            reduce_rows_warp<MODE><<<grid, 256, 0, s>>>(x, out, A, R);
        } else {
            int grid = (int)(A < (size_t)kNumSMs * 4 ? A : (size_t)kNumSMs * 4);
            reduce_rows_block<MODE><<<grid, 512, 0, s>>>(x, out, A, R);
        }
    } else {
        size_t cols = A * B;
        int gx = (int)((cols + 255) / 256);
        int rsplit = 1;
        if ((MODE == R_SUM || MODE == R_MEAN || MODE == R_SQSUM) && gx < kNumSMs && R >= 256) {
            rsplit = (int)((kNumSMs * 2 + gx - 1) / gx);
            if ((size_t)rsplit > R / 64) rsplit = (int)(R / 64);
            if (rsplit < 1) rsplit = 1;
        }
        if (rsplit > 1) cudaMemsetAsync(out, 0, cols * sizeof(float), s);
        reduce_mid_kernel<MODE><<<dim3(gx, rsplit), 256, 0, s>>>(x, out, A, R, B, rsplit);
    }
    return 0;
}

HETU_API int hetu_reduce_mid(const float* x, float* out, size_t A, size_t R, size_t B, int mode, cudaStream_t s) {
    if (A * B == 0) return 0;
    switch (mode) {
        case R_SUM: launch_reduce<R_SUM>(x, out, A, R, B, s); break;
        case R_MEAN: launch_reduce<R_MEAN>(x, out, A, R, B, s); break;
        case R_MAX: launch_reduce<R_MAX>(x, out, A, R, B, s); break;
        case R_MIN: launch_reduce<R_MIN>(x, out, A, R, B, s); break;
        case R_SQSUM: launch_reduce<R_SQSUM>(x, out, A, R, B, s); break;
        default: hetu_set_error("hetu_reduce_mid: bad mode %d", mode); return -1;
    }
    HETU_CHECK_LAUNCH("hetu_reduce_mid");
    return 0;
}

__global__ void __launch_bounds__(256) argmax_rows_kernel(const float* __restrict__ x, float* __restrict__ out,
                                                          size_t rows, size_t cols) {
    size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t row = warp; row < rows; row += nwarp) {
        const float* p = x + row * cols;
        float best = -INFINITY;
        int bi = 0x7fffffff;
        for (size_t j = lane; j < cols; j += 32) {
            float v = p[j];
            if (v > best) { best = v; bi = (int)j; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            float ov = __shfl_xor_sync(0xffffffffu, best, o);
            int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
        }
        if (lane == 0) out[row] = (float)bi;
    }
}
HETU_API int hetu_argmax_rows(const float* x, float* out, size_t rows, size_t cols, cudaStream_t s) {
    if (rows == 0) return 0;
    argmax_rows_kernel<<<ew_grid(rows * 32, 256), 256, 0, s>>>(x, out, rows, cols);
    HETU_CHECK_LAUNCH("hetu_argmax_rows");
    return 0;
}

// Row top-k (k <= 8): one warp per row, k rounds of warp arg-max with already-taken
// columns masked (every lane keeps the list of taken columns in registers).
__global__ void __launch_bounds__(256) topk_rows_kernel(const float* __restrict__ x, float* __restrict__ idx_out,
                                                        float* __restrict__ val_out, size_t rows, size_t cols, int k) {
    size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t row = warp; row < rows; row += nwarp) {
        const float* p = x + row * cols;
        int taken[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) taken[t] = -1;
#pragma unroll
        for (int t = 0; t < 8; ++t) {
            if (t >= k) break;
            float best = -INFINITY;
            int bi = 0x7fffffff;
            for (size_t j = lane; j < cols; j += 32) {
                bool skip = false;
#pragma unroll
                for (int u = 0; u < 8; ++u) skip |= (taken[u] == (int)j);
                float v = p[j];
                if (!skip && (v > best)) { best = v; bi = (int)j; }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                float ov = __shfl_xor_sync(0xffffffffu, best, o);
                int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
            }
            taken[t] = bi;
            if (lane == 0) {
                idx_out[row * k + t] = (float)bi;
                if (val_out) val_out[row * k + t] = best;
            }
        }
    }
}
HETU_API int hetu_topk_rows(const float* x, float* idx, float* val, size_t rows, size_t cols, int k, cudaStream_t s) {
    if (rows == 0) return 0;
    if (k > 8) { hetu_set_error("hetu_topk_rows: k=%d > 8", k); return -1; }
    topk_rows_kernel<<<ew_grid(rows * 32, 256), 256, 0, s>>>(x, idx, val, rows, cols, k);
    HETU_CHECK_LAUNCH("hetu_topk_rows");
    return 0;
}

// Inclusive column scan of x[T, E] (+bias).  T is large (16384 tokens), E small.
// A CTA of 32 warps owns 32 columns: warp w sums rows [w*chunk, (w+1)*chunk) of those
// columns (lane = column, coalesced 128 B rows), a 32-entry exclusive scan across warps in
// shared memory gives every warp its offset, then each warp rescans its chunk and writes.
__global__ void __launch_bounds__(1024) cumsum_cols_kernel(const float* __restrict__ x, float* __restrict__ out,
                                                           size_t T, size_t E, float bias) {
    __shared__ float part[32][33];
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    size_t col = (size_t)blockIdx.x * 32 + lane;
    size_t chunk = (T + 31) / 32;
    size_t r0 = (size_t)warp * chunk, r1 = r0 + chunk < T ? r0 + chunk : T;
    float acc = 0.f;
    if (col < E)
        for (size_t r = r0; r < r1; ++r) acc += x[r * E + col];
    part[warp][lane] = acc;
    __syncthreads();
    if (warp == 0) {
        float run = 0.f;
        for (int w = 0; w < 32; ++w) {
            float v = part[w][lane];
            part[w][lane] = run;
            run += v;
        }
    }
    __syncthreads();
    float run = part[warp][lane] + bias;
    if (col < E)
        for (size_t r = r0; r < r1; ++r) {
            run += x[r * E + col];
            out[r * E + col] = run;
        }
}
// Few columns (MoE gates: E = number of experts): one CTA per column, 1024 threads each own a run of
// consecutive rows; thread sums -> block-wide exclusive scan (shuffle + one smem hop) -> rescan.
__global__ void __launch_bounds__(1024) cumsum_fewcols_kernel(const float* __restrict__ x, float* __restrict__ out,
                                                              size_t T, size_t E, float bias) {
    __shared__ float wsum[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t col = blockIdx.x;
    const size_t chunk = (T + 1023) / 1024;
    const size_t r0 = (size_t)threadIdx.x * chunk, r1 = r0 + chunk < T ? r0 + chunk : T;
    float acc = 0.f;
    for (size_t r = r0; r < r1; ++r) acc += __ldg(x + r * E + col);
    float incl = acc;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        float v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        float w = wsum[lane], wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            float v = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += v;
        }
        wsum[lane] = wi - w;                       // exclusive prefix of the warp totals
    }
    __syncthreads();
    float run = wsum[warp] + (incl - acc) + bias;
    for (size_t r = r0; r < r1; ++r) {
        run += __ldg(x + r * E + col);
        out[r * E + col] = run;
    }
}

HETU_API int hetu_cumsum_cols(const float* x, float* out, size_t T, size_t E, float bias, cudaStream_t s) {
    if (T == 0 || E == 0) return 0;
    if (E <= 64 && T >= 2048) {
        cumsum_fewcols_kernel<<<(int)E, 1024, 0, s>>>(x, out, T, E, bias);
        HETU_CHECK_LAUNCH("hetu_cumsum_cols(few)");
        return 0;
    }
    cumsum_cols_kernel<<<(int)((E + 31) / 32), 1024, 0, s>>>(x, out, T, E, bias);
    HETU_CHECK_LAUNCH("hetu_cumsum_cols");
    return 0;
}
