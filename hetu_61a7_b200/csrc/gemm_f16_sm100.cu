// 16-bit operand GEMM for sm_100a: C(fp32) = act(alpha * op(A) op(B) + beta * C + bias) with A, B rounded to bf16 or
// fp16 and multiplied on the tensor cores with fp32 accumulation (tcgen05.mma kind::f16, accumulator in TMEM).
//
// reference: src/ops/MatrixMult.cu:22-26, BatchMatrixMult.cu:31-36 - cublasSgemm in fp32; the reference's Python
// layer has float16 plumbing (python/hetu/gpu_ops/executor.py dtype handling) but no half GEMM kernel.
//
// Storage stays fp32 (master weights, activations, gradients); only the two operands of a product are converted, in
// ONE pass each that also puts them into the layout the tensor core likes best - K-major rows of 128 bytes
// (SWIZZLE_128B, 64 elements): op(A) as [M, K], op(B) as [N, K].  A transposed operand is transposed by the convert
// kernel (32 x 32 shared-memory tiles), so the GEMM kernel has exactly one operand layout.  The convert pass reads
// 4 and writes 2 bytes per element - O(MK + NK) against O(MNK) MMA work.
//
// GEMM kernel = the warp-specialised structure of gemm_sm100.cu: warp 0 TMA producer (4-stage ring, mbarrier
// expect-tx), warp 1 TMEM allocator + one elected MMA thread (128 x BLOCK_N x 16 per instruction), warps 2..5 epilogue
// (tcgen05.ld 32x32b.x32 -> warp-private transpose tile -> coalesced 128-bit stores with the fused epilogue).
#include "tcgen05.cuh"
#include <cuda_bf16.h>
#include <cuda_fp16.h>

namespace {
constexpr int FBM = 128;            // rows per CTA tile = UMMA M
constexpr int FBK = 64;             // 64 x 2 bytes = one 128-byte swizzle row
constexpr int FUK = 16;             // K per tcgen05.mma kind::f16
constexpr int kF16Threads = 192;

__device__ __forceinline__ float f16_act(float v, int act) {
    if (act == 1) return fmaxf(v, 0.f);
    if (act == 2) return v * 0.5f * (1.0f + erff(v * 0.70710678118654752f));
    if (act == 3) return 1.f / (1.f + __expf(-v));
    return v;
}

__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accum) : "memory");
}

template <int BLOCK_N>
struct F16Cfg {
    static constexpr int kStages = 4;
    static constexpr int kABytes = FBM * FBK * 2;          // 16 KB
    static constexpr int kBBytes = BLOCK_N * FBK * 2;
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kSmem = kStages * kStageBytes + 1024 + 256;
    static constexpr int kTmemCols = BLOCK_N < 32 ? 32 : BLOCK_N;
};

struct F16Params {
    float* C;
    const float* bias;
    int M, N, K;
    long long sC;
    float alpha, beta;
    int act, is_bf16;
};

template <int BLOCK_N>
__global__ void __launch_bounds__(kF16Threads, 1)
gemm_f16_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, F16Params p) {
    using Cfg = F16Cfg<BLOCK_N>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + Cfg::kStages * Cfg::kABytes;
    float* smem_epi = reinterpret_cast<float*>(smem);       // reused once the last MMA has retired
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
    uint64_t* empty_bar = full_bar + Cfg::kStages;
    uint64_t* tmem_full_bar = empty_bar + Cfg::kStages;
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.y * FBM, n0 = blockIdx.x * BLOCK_N, bz = blockIdx.z;
    const int num_kb = (p.K + FBK - 1) / FBK;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmA);
        prefetch_tmap(&tmB);
        for (int i = 0; i < Cfg::kStages; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
        mbar_init(tmem_full_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_ptr_smem, Cfg::kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    if (warp == 0) {
        if (lane == 0) {
            for (int kb = 0; kb < num_kb; ++kb) {
                const int st = kb % Cfg::kStages;
                if (kb >= Cfg::kStages) mbar_wait(&empty_bar[st], ((kb / Cfg::kStages) - 1) & 1);
                mbar_expect_tx(&full_bar[st], Cfg::kStageBytes);
                tma_load_3d(smem_a + st * Cfg::kABytes, &tmA, kb * FBK, m0, bz, &full_bar[st]);     // box {64 k, 128 rows}
                tma_load_3d(smem_b + st * Cfg::kBBytes, &tmB, kb * FBK, n0, bz, &full_bar[st]);     // box {64 k, BLOCK_N rows}
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // instruction descriptor (kind::f16): D = f32 (bit 4), A / B format @7 / @10 (0 = f16, 1 = bf16), both K-major,
            // N >> 3 @17, M >> 4 @24
            const uint32_t fmt = p.is_bf16 ? 1u : 0u;
            const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(BLOCK_N >> 3) << 17) | ((uint32_t)(FBM >> 4) << 24);
            for (int kb = 0; kb < num_kb; ++kb) {
                const int st = kb % Cfg::kStages;
                mbar_wait(&full_bar[st], (kb / Cfg::kStages) & 1);
                tc_fence_after();
                const uint32_t a_base = smem_u32(smem_a + st * Cfg::kABytes);
                const uint32_t b_base = smem_u32(smem_b + st * Cfg::kBBytes);
#pragma unroll
                for (int ki = 0; ki < FBK / FUK; ++ki) {
                    // K-major, 128-byte rows, 8-row groups 1024 B apart; one MMA consumes 16 elements = 32 bytes of a row
                    const uint64_t a_desc = make_smem_desc(a_base + ki * (FUK * 2), 16, 1024, 2);
                    const uint64_t b_desc = make_smem_desc(b_base + ki * (FUK * 2), 16, 1024, 2);
                    umma_f16(tmem_base, a_desc, b_desc, idesc, (kb > 0 || ki > 0) ? 1u : 0u);
                }
                umma_commit(&empty_bar[st]);
            }
            umma_commit(tmem_full_bar);
        }
    } else {
        const int q = warp & 3;
        float* tile = smem_epi + (warp - 2) * (32 * 36);
        mbar_wait(tmem_full_bar, 0);
        tc_fence_after();
        float* Cb = p.C + (size_t)bz * p.sC;
        const int row_base = m0 + q * 32;
        const bool vec_ok = (p.N & 3) == 0 && ((reinterpret_cast<uintptr_t>(Cb) & 15) == 0);
        const int sub_row = lane >> 3, c4 = (lane & 7) * 4;
#pragma unroll 1
        for (int c0 = 0; c0 < BLOCK_N; c0 += 32) {
            if (n0 + c0 >= p.N) break;
            float v[32];
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
#pragma unroll
            for (int j = 0; j < 8; ++j)
                *reinterpret_cast<float4*>(&tile[lane * 36 + j * 4]) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            __syncwarp();
            const int col = n0 + c0 + c4;
            if (vec_ok && col + 3 < p.N) {
                float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
                if (p.bias != nullptr) bv = __ldg(reinterpret_cast<const float4*>(p.bias + col));
#pragma unroll
                for (int r = 0; r < 32; r += 4) {
                    const int row = row_base + r + sub_row;
                    if (row < p.M) {
                        const float4 a = *reinterpret_cast<const float4*>(&tile[(r + sub_row) * 36 + c4]);
                        float4 o = make_float4(p.alpha * a.x + bv.x, p.alpha * a.y + bv.y, p.alpha * a.z + bv.z, p.alpha * a.w + bv.w);
                        float4* dst = reinterpret_cast<float4*>(Cb + (size_t)row * p.N + col);
                        if (p.beta != 0.f) {
                            const float4 c = *dst;
                            o.x += p.beta * c.x; o.y += p.beta * c.y; o.z += p.beta * c.z; o.w += p.beta * c.w;
                        }
                        if (p.act != 0) { o.x = f16_act(o.x, p.act); o.y = f16_act(o.y, p.act); o.z = f16_act(o.z, p.act); o.w = f16_act(o.w, p.act); }
                        *dst = o;
                    }
                }
            } else {
                for (int r = 0; r < 32; r += 4) {
                    const int row = row_base + r + sub_row;
                    if (row >= p.M) continue;
                    for (int u = 0; u < 4; ++u) {
                        const int cc = col + u;
                        if (cc >= p.N) break;
                        float o = p.alpha * tile[(r + sub_row) * 36 + c4 + u] + (p.bias ? __ldg(p.bias + cc) : 0.f);
                        float* dst = Cb + (size_t)row * p.N + cc;
                        if (p.beta != 0.f) o += p.beta * (*dst);
                        *dst = f16_act(o, p.act);
                    }
                }
            }
            __syncwarp();
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, Cfg::kTmemCols);
}

// ---- operand conversion: fp32 -> bf16 / fp16, optionally transposed, K padded with zeros to a multiple of 8 ------------
template <typename T> __device__ __forceinline__ T cvt16(float v);
template <> __device__ __forceinline__ __nv_bfloat16 cvt16<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ __half cvt16<__half>(float v) { return __float2half_rn(v); }

// src [rows, cols] (pitch cols) -> dst [rows, cols_pad]
template <typename T>
__global__ void __launch_bounds__(256) cvt_rows_kernel(const float* __restrict__ src, T* __restrict__ dst, long long rows, int cols,
                                                       int cols_pad, long long src_batch, long long dst_batch) {
    const float* s = src + (size_t)blockIdx.y * src_batch;
    T* d = dst + (size_t)blockIdx.y * dst_batch;
    const long long total = rows * cols_pad;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long r = i / cols_pad;
        const int c = (int)(i - r * cols_pad);
        d[i] = cvt16<T>(c < cols ? s[r * cols + c] : 0.f);
    }
}
// src [rows_s, cols_s] -> dst [cols_s, rows_pad] (transpose; rows_s becomes the contiguous K)
template <typename T>
__global__ void __launch_bounds__(256) cvt_transpose_kernel(const float* __restrict__ src, T* __restrict__ dst, int rows_s, int cols_s,
                                                            int rows_pad, long long src_batch, long long dst_batch) {
    __shared__ float tile[32][33];
    const float* s = src + (size_t)blockIdx.z * src_batch;
    T* d = dst + (size_t)blockIdx.z * dst_batch;
    const int r0 = blockIdx.y * 32, c0 = blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;       // 32 x 8
    for (int j = ty; j < 32; j += 8) {
        const int r = r0 + j, c = c0 + tx;
        tile[j][tx] = (r < rows_s && c < cols_s) ? s[(size_t)r * cols_s + c] : 0.f;
    }
    __syncthreads();
    for (int j = ty; j < 32; j += 8) {
        const int c = c0 + j, r = r0 + tx;                        // output row = source column
        if (c < cols_s && r < rows_pad) d[(size_t)c * rows_pad + r] = cvt16<T>(tile[tx][j]);
    }
}

int make_tmap16(CUtensorMap* map, const void* base, bool bf16, long long inner, long long rows, long long batch, int box_rows) {
    auto fn = get_encode_fn();
    if (!fn) { hetu_set_error("cuTensorMapEncodeTiled entry point not found"); return -1; }
    cuuint64_t dims[3] = {(cuuint64_t)inner, (cuuint64_t)rows, (cuuint64_t)(batch > 0 ? batch : 1)};
    cuuint64_t strides[2] = {(cuuint64_t)inner * 2, (cuuint64_t)inner * rows * 2};
    cuuint32_t box[3] = {(cuuint32_t)FBK, (cuuint32_t)box_rows, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = fn(map, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void*>(base), dims,
                    strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { hetu_set_error("cuTensorMapEncodeTiled(16-bit) failed with %d (inner=%lld rows=%lld)", (int)r, inner, rows); return -1; }
    return 0;
}

template <int BLOCK_N>
int launch_f16(const void* Ah, const void* Bh, F16Params& p, int Kp, int batch, cudaStream_t s) {
    using Cfg = F16Cfg<BLOCK_N>;
    CUtensorMap tmA, tmB;
    if (make_tmap16(&tmA, Ah, p.is_bf16, Kp, p.M, batch, FBM)) return -1;
    if (make_tmap16(&tmB, Bh, p.is_bf16, Kp, p.N, batch, BLOCK_N)) return -1;
    static bool attr_set = false;
    if (!attr_set) {
        HETU_CUDA(cudaFuncSetAttribute(gemm_f16_tcgen05_kernel<BLOCK_N>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmem));
        attr_set = true;
    }
    dim3 grid((p.N + BLOCK_N - 1) / BLOCK_N, (p.M + FBM - 1) / FBM, batch);
    gemm_f16_tcgen05_kernel<BLOCK_N><<<grid, kF16Threads, Cfg::kSmem, s>>>(tmA, tmB, p);
    return 0;
}
}  // namespace

// bytes of 16-bit scratch the caller must provide for op(A) and op(B) (each: batch x rows x K rounded up to 8)
HETU_API long long hetu_gemm_f16_scratch_elems(int rows, int K, int batch) { return (long long)batch * rows * ((K + 7) / 8 * 8); }

// A: [M, K] (transA == 0) or [K, M]; B: [K, N] (transB == 0) or [N, K]; fp32, dense, batch strides sA / sB (0 = shared).
// a16 / b16: scratch from hetu_gemm_f16_scratch_elems (2 bytes per element).  half_kind: 0 = fp16, 1 = bf16.
HETU_API int hetu_gemm_f16(const float* A, const float* B, float* C, const float* bias, void* a16, void* b16, int M, int N, int K,
                           int transA, int transB, int batch, long long sA, long long sB, long long sC, float alpha, float beta,
                           int act, int half_kind, cudaStream_t s) {
    if (M <= 0 || N <= 0 || K <= 0 || batch <= 0) return 0;
    const int Kp = (K + 7) / 8 * 8;                      // 16-byte row pitch for TMA
    const int ba = (batch > 1 && sA != 0) ? batch : 1, bb = (batch > 1 && sB != 0) ? batch : 1;
    if ((ba != batch && ba != 1) || (bb != batch && bb != 1)) return -1;
    auto convert = [&](const float* src, void* dst, int rows, bool transposed, int nb, long long sb) {
        // rows = M or N; source is [rows, K] or (transposed) [K, rows]
        if (!transposed) {
            dim3 g(ew_grid((size_t)rows * Kp, 256), nb);
            if (half_kind) cvt_rows_kernel<__nv_bfloat16><<<g, 256, 0, s>>>(src, (__nv_bfloat16*)dst, rows, K, Kp, sb, (long long)rows * Kp);
            else cvt_rows_kernel<__half><<<g, 256, 0, s>>>(src, (__half*)dst, rows, K, Kp, sb, (long long)rows * Kp);
        } else {
            dim3 g((rows + 31) / 32, (Kp + 31) / 32, nb);
            if (half_kind) cvt_transpose_kernel<__nv_bfloat16><<<g, 256, 0, s>>>(src, (__nv_bfloat16*)dst, K, rows, Kp, sb, (long long)rows * Kp);
            else cvt_transpose_kernel<__half><<<g, 256, 0, s>>>(src, (__half*)dst, K, rows, Kp, sb, (long long)rows * Kp);
        }
    };
    convert(A, a16, M, transA != 0, ba, sA);
    convert(B, b16, N, transB == 0, bb, sB);
    HETU_CHECK_LAUNCH("hetu_gemm_f16(convert)");
    if (ba != batch || bb != batch) {
        if (batch > 1) { hetu_set_error("hetu_gemm_f16: broadcast batch operands are not supported"); return -1; }
    }
    F16Params p;
    p.C = C; p.bias = bias; p.M = M; p.N = N; p.K = Kp; p.sC = sC; p.alpha = alpha; p.beta = beta; p.act = act; p.is_bf16 = half_kind ? 1 : 0;
    int rc;
    if (N > 128) rc = launch_f16<256>(a16, b16, p, Kp, batch, s);
    else if (N > 64) rc = launch_f16<128>(a16, b16, p, Kp, batch, s);
    else if (N > 32) rc = launch_f16<64>(a16, b16, p, Kp, batch, s);
    else rc = launch_f16<32>(a16, b16, p, Kp, batch, s);
    if (rc) return rc;
    HETU_CHECK_LAUNCH("hetu_gemm_f16");
    return 0;
}
