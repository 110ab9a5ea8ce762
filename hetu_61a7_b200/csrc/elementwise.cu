// Element-wise kernels: unary / binary functor switch, row/column broadcast
// arithmetic, where, generic strided copy, one-hot.
// reference kernels: src/ops/{Abs,AddConst,...,Where}.cu (one file, one launch
// config "threads=min(size,1024), one element per thread" each - SURVEY 2.5).
// Here: float4 main loop + scalar tail, grid-stride over <= 148*8 CTAs.
#include "common.cuh"

enum { U_ABS = 0, U_ADD_CONST, U_MUL_CONST, U_EXP, U_FLOOR, U_GELU, U_LEAKY_RELU, U_LOG,
       U_MINUS_BY_CONST, U_OPPOSITE, U_POW, U_CONST_POW, U_RELU, U_SIGMOID, U_SIN, U_COS, U_SQRT,
       U_RSQRT, U_TANH, U_DIV_CONST, U_CLAMP, U_BOOL, U_BOOL_EQ, U_BOOL_LT, U_BOOL_GT, U_BOOL_LE,
       U_BOOL_GE, U_FILL, U_COPY, U_SIGN, U_CLAMP_MIN, U_CLAMP_MAX };
enum { B_ADD = 0, B_SUB, B_MUL, B_DIV, B_RELU_GRAD, B_GELU_GRAD, B_LEAKY_RELU_GRAD, B_TANH_GRAD,
       B_ABS_GRAD, B_POW_GRAD, B_CONST_POW_GRAD, B_LOG_GRAD, B_EQ, B_LT, B_GT, B_LE, B_GE,
       B_MASKED_FILL, B_WHERE_CONST, B_MAX, B_MIN, B_SIGMOID_GRAD };

template <int OP>
__device__ __forceinline__ float unary_fn(float x, float a, float b) {
    if (OP == U_ABS) return fabsf(x);
    if (OP == U_ADD_CONST) return x + a;
    if (OP == U_MUL_CONST) return x * a;
    if (OP == U_EXP) return expf(x);
    if (OP == U_FLOOR) return floorf(x);
    if (OP == U_GELU) return x * 0.5f * (1.0f + erff(x * 0.70710678118654752f));
    if (OP == U_LEAKY_RELU) return x > 0.f ? x : a * x;
    if (OP == U_LOG) return logf(x);
    if (OP == U_MINUS_BY_CONST) return a - x;
    if (OP == U_OPPOSITE) return -x;
    if (OP == U_POW) return powf(x, a);
    if (OP == U_CONST_POW) return powf(a, x);
    if (OP == U_RELU) return fmaxf(x, 0.f);
    if (OP == U_SIGMOID) return 1.0f / (1.0f + expf(-x));
    if (OP == U_SIN) return sinf(x);
    if (OP == U_COS) return cosf(x);
    if (OP == U_SQRT) return sqrtf(x);
    if (OP == U_RSQRT) return rsqrtf(x);
    if (OP == U_TANH) return tanhf(x);
    if (OP == U_DIV_CONST) return a / x;
    if (OP == U_CLAMP) return fminf(fmaxf(x, a), b);
    if (OP == U_CLAMP_MIN) return fmaxf(x, a);
    if (OP == U_CLAMP_MAX) return fminf(x, a);
    if (OP == U_BOOL) return x > 0.f ? 1.f : 0.f;
    if (OP == U_BOOL_EQ) return fabsf(x - a) < 1e-6f ? 1.f : 0.f;
    if (OP == U_BOOL_LT) return (x - a) < -1e-6f ? 1.f : 0.f;
    if (OP == U_BOOL_GT) return (x - a) > 1e-6f ? 1.f : 0.f;
    if (OP == U_BOOL_LE) return (x - a) < 1e-6f ? 1.f : 0.f;
    if (OP == U_BOOL_GE) return (x - a) > -1e-6f ? 1.f : 0.f;
    if (OP == U_FILL) return a;
    if (OP == U_SIGN) return (x > 0.f) - (x < 0.f);
    return x;  // U_COPY
}

template <int OP>
__device__ __forceinline__ float binary_fn(float x, float y, float a) {
    if (OP == B_ADD) return x + y;
    if (OP == B_SUB) return x - y;
    if (OP == B_MUL) return x * y;
    if (OP == B_DIV) return x / y;
    if (OP == B_RELU_GRAD) return x > 0.f ? y : 0.f;
    if (OP == B_GELU_GRAD)
        return y * (0.5f + 0.5f * erff(x * 0.70710678118654752f) +
                    x * 0.3989422804014327f * expf(-0.5f * x * x));
    if (OP == B_LEAKY_RELU_GRAD) return x > 0.f ? y : a * y;
    if (OP == B_TANH_GRAD) return y * (1.f - x * x);
    if (OP == B_SIGMOID_GRAD) return y * x * (1.f - x);
    if (OP == B_ABS_GRAD) return y * ((x > 0.f) - (x < 0.f));
    if (OP == B_POW_GRAD) return a * powf(x, a - 1.f) * y;
    if (OP == B_CONST_POW_GRAD) return x * y * logf(a);
    if (OP == B_LOG_GRAD) return x / y;
    if (OP == B_EQ) return fabsf(x - y) < 1e-6f ? 1.f : 0.f;
    if (OP == B_LT) return (x - y) < -1e-6f ? 1.f : 0.f;
    if (OP == B_GT) return (x - y) > 1e-6f ? 1.f : 0.f;
    if (OP == B_LE) return (x - y) < 1e-6f ? 1.f : 0.f;
    if (OP == B_GE) return (x - y) > -1e-6f ? 1.f : 0.f;
    if (OP == B_MASKED_FILL) return y != 0.f ? a : x;
    if (OP == B_WHERE_CONST) return x != 0.f ? y : a;
    if (OP == B_MAX) return fmaxf(x, y);
    if (OP == B_MIN) return fminf(x, y);
    return x;
}

template <int OP>
__global__ void __launch_bounds__(256) unary_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                    size_t n, float a, float b) {
    size_t n4 = n >> 2;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const float4* x4 = reinterpret_cast<const float4*>(x);
    float4* y4 = reinterpret_cast<float4*>(y);
    for (size_t i = tid; i < n4; i += stride) {
        float4 v = (OP == U_FILL) ? make_float4(0, 0, 0, 0) : x4[i];
        v.x = unary_fn<OP>(v.x, a, b); v.y = unary_fn<OP>(v.y, a, b);
        v.z = unary_fn<OP>(v.z, a, b); v.w = unary_fn<OP>(v.w, a, b);
        y4[i] = v;
    }
    for (size_t i = (n4 << 2) + tid; i < n; i += stride)
        y[i] = unary_fn<OP>((OP == U_FILL) ? 0.f : x[i], a, b);
}

template <int OP>
__global__ void __launch_bounds__(256) binary_kernel(const float* __restrict__ x, const float* __restrict__ y,
                                                     float* __restrict__ o, size_t n, float a) {
    size_t n4 = n >> 2;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const float4* x4 = reinterpret_cast<const float4*>(x);
    const float4* y4 = reinterpret_cast<const float4*>(y);
    float4* o4 = reinterpret_cast<float4*>(o);
    for (size_t i = tid; i < n4; i += stride) {
        float4 u = x4[i], v = y4[i], r;
        r.x = binary_fn<OP>(u.x, v.x, a); r.y = binary_fn<OP>(u.y, v.y, a);
        r.z = binary_fn<OP>(u.z, v.z, a); r.w = binary_fn<OP>(u.w, v.w, a);
        o4[i] = r;
    }
    for (size_t i = (n4 << 2) + tid; i < n; i += stride) o[i] = binary_fn<OP>(x[i], y[i], a);
}

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// scalar fallbacks for mis-aligned views
template <int OP>
__global__ void unary_kernel_s(const float* x, float* y, size_t n, float a, float b) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        y[i] = unary_fn<OP>((OP == U_FILL) ? 0.f : x[i], a, b);
}
template <int OP>
__global__ void binary_kernel_s(const float* x, const float* y, float* o, size_t n, float a) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        o[i] = binary_fn<OP>(x[i], y[i], a);
}

#define UCASE(OP)                                                                          \
    case OP:                                                                               \
        if (vec) unary_kernel<OP><<<grid, 256, 0, s>>>(x, y, n, a, b);                     \
        else unary_kernel_s<OP><<<grid, 256, 0, s>>>(x, y, n, a, b);                       \
        break;
#define BCASE(OP)                                                                          \
    case OP:                                                                               \
        if (vec) binary_kernel<OP><<<grid, 256, 0, s>>>(x, y, o, n, a);                    \
        else binary_kernel_s<OP><<<grid, 256, 0, s>>>(x, y, o, n, a);                      \
        break;

HETU_API int hetu_unary(int op, const float* x, float* y, size_t n, float a, float b, cudaStream_t s) {
    if (n == 0) return 0;
    bool vec = aligned16(x) && aligned16(y);
    int grid = ew_grid(vec ? (n + 3) / 4 : n, 256);
    switch (op) {
        UCASE(U_ABS) UCASE(U_ADD_CONST) UCASE(U_MUL_CONST) UCASE(U_EXP) UCASE(U_FLOOR) UCASE(U_GELU)
        UCASE(U_LEAKY_RELU) UCASE(U_LOG) UCASE(U_MINUS_BY_CONST) UCASE(U_OPPOSITE) UCASE(U_POW)
        UCASE(U_CONST_POW) UCASE(U_RELU) UCASE(U_SIGMOID) UCASE(U_SIN) UCASE(U_COS) UCASE(U_SQRT)
        UCASE(U_RSQRT) UCASE(U_TANH) UCASE(U_DIV_CONST) UCASE(U_CLAMP) UCASE(U_BOOL) UCASE(U_BOOL_EQ)
        UCASE(U_BOOL_LT) UCASE(U_BOOL_GT) UCASE(U_BOOL_LE) UCASE(U_BOOL_GE) UCASE(U_FILL) UCASE(U_COPY)
        UCASE(U_SIGN) UCASE(U_CLAMP_MIN) UCASE(U_CLAMP_MAX)
        default: hetu_set_error("hetu_unary: bad op %d", op); return -1;
    }
    HETU_CHECK_LAUNCH("hetu_unary");
    return 0;
}

HETU_API int hetu_binary(int op, const float* x, const float* y, float* o, size_t n, float a, cudaStream_t s) {
    if (n == 0) return 0;
    bool vec = aligned16(x) && aligned16(y) && aligned16(o);
    int grid = ew_grid(vec ? (n + 3) / 4 : n, 256);
    switch (op) {
        BCASE(B_ADD) BCASE(B_SUB) BCASE(B_MUL) BCASE(B_DIV) BCASE(B_RELU_GRAD) BCASE(B_GELU_GRAD)
        BCASE(B_LEAKY_RELU_GRAD) BCASE(B_TANH_GRAD) BCASE(B_ABS_GRAD) BCASE(B_POW_GRAD)
        BCASE(B_CONST_POW_GRAD) BCASE(B_LOG_GRAD) BCASE(B_EQ) BCASE(B_LT) BCASE(B_GT) BCASE(B_LE)
        BCASE(B_GE) BCASE(B_MASKED_FILL) BCASE(B_WHERE_CONST) BCASE(B_MAX) BCASE(B_MIN)
        BCASE(B_SIGMOID_GRAD)
        default: hetu_set_error("hetu_binary: bad op %d", op); return -1;
    }
    HETU_CHECK_LAUNCH("hetu_binary");
    return 0;
}

// x[outer, inner] (op) y: mode 0 -> y[inner] broadcast over rows (bias add);
//                         mode 1 -> y[outer] broadcast along each row (`inner` = row length)
template <int OP>
__global__ void __launch_bounds__(256) bcast_kernel(const float* __restrict__ x, const float* __restrict__ y,
                                                    float* __restrict__ o, size_t n, size_t inner, int mode) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        float yv = mode == 0 ? __ldg(y + (i % inner)) : __ldg(y + (i / inner));
        o[i] = binary_fn<OP>(x[i], yv, 0.f);
    }
}

HETU_API int hetu_binary_bcast(int op, const float* x, const float* y, float* o, size_t n, size_t inner,
                               int mode, cudaStream_t s) {
    if (n == 0) return 0;
    int grid = ew_grid(n, 256);
    switch (op) {
        case B_ADD: bcast_kernel<B_ADD><<<grid, 256, 0, s>>>(x, y, o, n, inner, mode); break;
        case B_SUB: bcast_kernel<B_SUB><<<grid, 256, 0, s>>>(x, y, o, n, inner, mode); break;
        case B_MUL: bcast_kernel<B_MUL><<<grid, 256, 0, s>>>(x, y, o, n, inner, mode); break;
        case B_DIV: bcast_kernel<B_DIV><<<grid, 256, 0, s>>>(x, y, o, n, inner, mode); break;
        default: hetu_set_error("hetu_binary_bcast: bad op %d", op); return -1;
    }
    HETU_CHECK_LAUNCH("hetu_binary_bcast");
    return 0;
}

__global__ void __launch_bounds__(256) where_kernel(const float* __restrict__ c, const float* __restrict__ a,
                                                    const float* __restrict__ b, float* __restrict__ o, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        o[i] = c[i] != 0.f ? a[i] : b[i];
}
HETU_API int hetu_where(const float* c, const float* a, const float* b, float* o, size_t n, cudaStream_t s) {
    if (n == 0) return 0;
    where_kernel<<<ew_grid(n, 256), 256, 0, s>>>(c, a, b, o, n);
    HETU_CHECK_LAUNCH("hetu_where");
    return 0;
}

// ---------------------------------------------------------------- strided copy
struct StridedDesc { long long shape[8], ss[8], ds[8]; int nd; };

__global__ void __launch_bounds__(256) strided_copy_kernel(const float* __restrict__ src, float* __restrict__ dst,
                                                           StridedDesc d, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        size_t rem = i;
        long long so = 0, dof = 0;
#pragma unroll
        for (int k = 7; k >= 0; --k) {
            if (k < d.nd) {
                long long idx = (long long)(rem % (size_t)d.shape[k]);
                rem /= (size_t)d.shape[k];
                so += idx * d.ss[k];
                dof += idx * d.ds[k];
            }
        }
        dst[dof] = src[so];
    }
}

// fast path: innermost dim a multiple of 4, unit-stride in dst and unit-stride (or broadcast, BCAST) in src -> float4 rows
template <bool BCAST>
__global__ void __launch_bounds__(256) strided_copy_vec4_kernel(const float* __restrict__ src, float* __restrict__ dst,
                                                                StridedDesc d, size_t n4) {
    long long inner4 = d.shape[d.nd - 1] >> 2;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        size_t rem = i / (size_t)inner4;
        long long col = (long long)(i % (size_t)inner4) << 2;
        long long so = BCAST ? 0 : col, dof = col;
#pragma unroll
        for (int k = 6; k >= 0; --k) {
            if (k < d.nd - 1) {
                long long idx = (long long)(rem % (size_t)d.shape[k]);
                rem /= (size_t)d.shape[k];
                so += idx * d.ss[k];
                dof += idx * d.ds[k];
            }
        }
        if (BCAST) {
            const float v = __ldg(src + so);
            *reinterpret_cast<float4*>(dst + dof) = make_float4(v, v, v, v);
        } else {
            *reinterpret_cast<float4*>(dst + dof) = *reinterpret_cast<const float4*>(src + so);
        }
    }
}

HETU_API int hetu_strided_copy(const float* src, float* dst, int nd, const long long* shape,
                               const long long* ss, const long long* ds, size_t n, cudaStream_t s) {
    if (n == 0) return 0;
    StridedDesc d;
    d.nd = nd;
    const bool bcast = nd >= 1 && ss[nd - 1] == 0;           // e.g. the gradient of a row reduction, expanded
    bool vec = nd >= 1 && (ss[nd - 1] == 1 || bcast) && ds[nd - 1] == 1 && (shape[nd - 1] & 3) == 0 &&
               (bcast || aligned16(src)) && aligned16(dst);
    for (int i = 0; i < 8; ++i) {
        d.shape[i] = i < nd ? shape[i] : 1;
        d.ss[i] = i < nd ? ss[i] : 0;
        d.ds[i] = i < nd ? ds[i] : 0;
        if (i < nd - 1 && ((!bcast && (ss[i] & 3)) || (ds[i] & 3))) vec = false;
    }
    if (vec) {
        size_t n4 = n >> 2;
        if (bcast) strided_copy_vec4_kernel<true><<<ew_grid(n4, 256), 256, 0, s>>>(src, dst, d, n4);
        else strided_copy_vec4_kernel<false><<<ew_grid(n4, 256), 256, 0, s>>>(src, dst, d, n4);
    } else {
        strided_copy_kernel<<<ew_grid(n, 256), 256, 0, s>>>(src, dst, d, n);
    }
    HETU_CHECK_LAUNCH("hetu_strided_copy");
    return 0;
}

__global__ void __launch_bounds__(256) one_hot_kernel(const float* __restrict__ idx, float* __restrict__ out,
                                                      size_t n, int classes) {
    size_t total = n * (size_t)classes;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        size_t r = i / classes;
        int c = (int)(i % classes);
        out[i] = ((int)idx[r] == c) ? 1.f : 0.f;
    }
}
HETU_API int hetu_one_hot(const float* idx, float* out, size_t n, int classes, cudaStream_t s) {
    if (n == 0) return 0;
    one_hot_kernel<<<ew_grid(n * classes, 256), 256, 0, s>>>(idx, out, n, classes);
    HETU_CHECK_LAUNCH("hetu_one_hot");
    return 0;
}

// ---------------------------------------------------------------- 3xTF32 operand split
// x = hi + lo with hi the nearest tf32 value (10 explicit mantissa bits) and lo = x - hi (exact in fp32).
// ops/linalg.py runs A B as A_lo B_hi + A_hi B_lo + A_hi B_hi on the tensor cores: fp32-faithful products at a
// third of the tf32 rate instead of the CUDA-core rate (reference precision: cublasSgemm, src/ops/MatrixMult.cu:22-26).
__global__ void __launch_bounds__(256) split_tf32_kernel(const float* __restrict__ x, float* __restrict__ hi, float* __restrict__ lo,
                                                         size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float v = x[i];
        const uint32_t b = __float_as_uint(v);
        const float h = ((b & 0x7f800000u) == 0x7f800000u) ? v : __uint_as_float((b + 0x1000u) & 0xffffe000u);
        hi[i] = h;
        lo[i] = v - h;
    }
}
HETU_API int hetu_split_tf32(const float* x, float* hi, float* lo, size_t n, cudaStream_t s) {
    if (n == 0) return 0;
    split_tf32_kernel<<<ew_grid(n, 256), 256, 0, s>>>(x, hi, lo, n);
    HETU_CHECK_LAUNCH("hetu_split_tf32");
    return 0;
}

// ---------------------------------------------------------------- batched 2-D transpose
// src [B, R, C] -> dst [B, C, R] through 32 x 32 shared-memory tiles (both sides coalesced).  NCHW <-> NHWC changes
// around the convolution GEMMs are this pattern; the generic strided copy does a 64-bit div / mod chain per element and
// reads with a stride.
__global__ void __launch_bounds__(256) transpose_batched_kernel(const float* __restrict__ src, float* __restrict__ dst, int R, int C) {
    __shared__ float tile[32][33];
    const size_t boff = (size_t)blockIdx.z * R * C;
    const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;          // 32 x 8
#pragma unroll
    for (int j = 0; j < 32; j += 8) {
        const int r = r0 + ty + j, c = c0 + tx;
        if (r < R && c < C) tile[ty + j][tx] = src[boff + (size_t)r * C + c];
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 32; j += 8) {
        const int c = c0 + ty + j, r = r0 + tx;
        if (r < R && c < C) dst[boff + (size_t)c * R + r] = tile[tx][ty + j];
    }
}
HETU_API int hetu_transpose_batched(const float* src, float* dst, int B, int R, int C, cudaStream_t s) {
    if (B <= 0 || R <= 0 || C <= 0) return 0;
    if (B > 65535) { hetu_set_error("transpose_batched: batch %d > 65535", B); return -1; }
    dim3 grid((C + 31) / 32, (R + 31) / 32, B);
    transpose_batched_kernel<<<grid, 256, 0, s>>>(src, dst, R, C);
    HETU_CHECK_LAUNCH("hetu_transpose_batched");
    return 0;
}
