// CPU op library (OpenMP) behind the CPU executor and the numpy-oracle tests.
//
// reference: src/dnnl_ops/*.cpp (oneDNN primitives + OpenMP loops, 47 exported functions listed in
// python/hetu/_base.py:16-58).  oneDNN is not part of this image; the primitives are written out as
// cache-blocked OpenMP loops the compiler vectorises (AVX2/FMA); the GEMM has hand-written AVX-512 / AVX2 micro-kernels.  All tensors are dense row-major
// fp32; every function returns 0 on success.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>
#include <omp.h>

#define API extern "C" __attribute__((visibility("default")))

// ------------------------------------------------------------------------------------------------
// GEMM: C[M,N] = alpha * op(A) op(B) + beta * C          (reference MatrixMult.cpp:18-56)
//
// C is cut into tiles that are handed to the OpenMP threads; a thread walks K in chunks, packs the
// op(A) / op(B) sub-panels of its tile into thread-local buffers (so the micro-kernel streams
// unit-stride data whatever the transposes are) and runs a register-blocked micro-kernel:
// 12 x 32 on AVX-512 (24 zmm accumulators + 2 B vectors + 1 broadcast), 6 x 16 on AVX2.
// The instruction set is picked once at run time.
#include <immintrin.h>

namespace {
// K chunk walked per micro-kernel call.  384 measured best on the MLP shapes of BASELINE config 1 (a sweep of
// 96..384 on a shared 8-vCPU box stayed within noise); HETU_CPU_GEMM_KC overrides it (tuning knob).
int kc_setting() {
    static const int v = [] {
        const char* e = getenv("HETU_CPU_GEMM_KC");
        int k = e ? atoi(e) : 0;
        return k >= 16 ? k : 384;
    }();
    return v;
}

struct Kern { int MR, NR; void (*fn)(int, const float*, const float*, int, float*, int, int, int, float, float); };

// a: packed [kc][MR], b: packed [kc][NR]; C tile (mr x nr valid) = alpha * a.b + beta * C
__attribute__((target("avx512f"), noinline))
void kern_avx512(int kc, const float* __restrict a, const float* __restrict b, int ldbk, float* __restrict c, int ldc, int mr,
                 int nr, float alpha, float beta) {
    constexpr int MR = 12;
    __m512 acc0[MR], acc1[MR];
    for (int i = 0; i < MR; ++i) { acc0[i] = _mm512_setzero_ps(); acc1[i] = _mm512_setzero_ps(); }
    for (int p = 0; p < kc; ++p) {
        __m512 b0 = _mm512_loadu_ps(b + (size_t)p * ldbk), b1 = _mm512_loadu_ps(b + (size_t)p * ldbk + 16);
        const float* ap = a + p * MR;
#pragma GCC unroll 12
        for (int i = 0; i < MR; ++i) {
            __m512 av = _mm512_set1_ps(ap[i]);
            acc0[i] = _mm512_fmadd_ps(av, b0, acc0[i]);
            acc1[i] = _mm512_fmadd_ps(av, b1, acc1[i]);
        }
    }
    __m512 va = _mm512_set1_ps(alpha), vb = _mm512_set1_ps(beta);
    __mmask16 m0 = nr >= 16 ? (__mmask16)0xffff : (__mmask16)((1u << nr) - 1);
    __mmask16 m1 = nr >= 32 ? (__mmask16)0xffff : (nr > 16 ? (__mmask16)((1u << (nr - 16)) - 1) : (__mmask16)0);
    for (int i = 0; i < mr; ++i) {
        float* cr = c + (size_t)i * ldc;
        __m512 r0 = _mm512_mul_ps(va, acc0[i]), r1 = _mm512_mul_ps(va, acc1[i]);
        if (beta != 0.f) {
            r0 = _mm512_fmadd_ps(vb, _mm512_maskz_loadu_ps(m0, cr), r0);
            if (m1) r1 = _mm512_fmadd_ps(vb, _mm512_maskz_loadu_ps(m1, cr + 16), r1);
        }
        _mm512_mask_storeu_ps(cr, m0, r0);
        if (m1) _mm512_mask_storeu_ps(cr + 16, m1, r1);
    }
}

// same panel layout, at most 16 valid columns: half the FMAs (the 10-class output layer and its gradients)
__attribute__((target("avx512f"), noinline))
void kern_avx512_n16(int kc, const float* __restrict a, const float* __restrict b, int ldbk, float* __restrict c, int ldc, int mr,
                     int nr, float alpha, float beta) {
    constexpr int MR = 12;
    __m512 acc[MR];
    for (int i = 0; i < MR; ++i) acc[i] = _mm512_setzero_ps();
    for (int p = 0; p < kc; ++p) {
        __m512 b0 = _mm512_loadu_ps(b + (size_t)p * ldbk);
        const float* ap = a + p * MR;
#pragma GCC unroll 12
        for (int i = 0; i < MR; ++i) acc[i] = _mm512_fmadd_ps(_mm512_set1_ps(ap[i]), b0, acc[i]);
    }
    __m512 va = _mm512_set1_ps(alpha), vb = _mm512_set1_ps(beta);
    __mmask16 m0 = nr >= 16 ? (__mmask16)0xffff : (__mmask16)((1u << nr) - 1);
    for (int i = 0; i < mr; ++i) {
        float* cr = c + (size_t)i * ldc;
        __m512 r0 = _mm512_mul_ps(va, acc[i]);
        if (beta != 0.f) r0 = _mm512_fmadd_ps(vb, _mm512_maskz_loadu_ps(m0, cr), r0);
        _mm512_mask_storeu_ps(cr, m0, r0);
    }
}

__attribute__((target("avx2,fma"), noinline))
void kern_avx2(int kc, const float* __restrict a, const float* __restrict b, int ldbk, float* __restrict c, int ldc, int mr,
               int nr, float alpha, float beta) {
    constexpr int MR = 6;
    __m256 acc0[MR], acc1[MR];
    for (int i = 0; i < MR; ++i) { acc0[i] = _mm256_setzero_ps(); acc1[i] = _mm256_setzero_ps(); }
    for (int p = 0; p < kc; ++p) {
        __m256 b0 = _mm256_loadu_ps(b + (size_t)p * ldbk), b1 = _mm256_loadu_ps(b + (size_t)p * ldbk + 8);
        const float* ap = a + p * MR;
#pragma GCC unroll 6
        for (int i = 0; i < MR; ++i) {
            __m256 av = _mm256_broadcast_ss(ap + i);
            acc0[i] = _mm256_fmadd_ps(av, b0, acc0[i]);
            acc1[i] = _mm256_fmadd_ps(av, b1, acc1[i]);
        }
    }
    alignas(32) float tmp[16];
    for (int i = 0; i < mr; ++i) {
        _mm256_store_ps(tmp, acc0[i]);
        _mm256_store_ps(tmp + 8, acc1[i]);
        float* cr = c + (size_t)i * ldc;
        for (int j = 0; j < nr; ++j) cr[j] = alpha * tmp[j] + (beta != 0.f ? beta * cr[j] : 0.f);
    }
}

void kern_scalar(int kc, const float* a, const float* b, int ldbk, float* c, int ldc, int mr, int nr, float alpha, float beta) {
    constexpr int MR = 4, NR = 8;
    float acc[MR][NR] = {};
    for (int p = 0; p < kc; ++p)
        for (int i = 0; i < MR; ++i)
            for (int j = 0; j < NR; ++j) acc[i][j] += a[p * MR + i] * b[(size_t)p * ldbk + j];
    for (int i = 0; i < mr; ++i)
        for (int j = 0; j < nr; ++j) c[(size_t)i * ldc + j] = alpha * acc[i][j] + (beta != 0.f ? beta * c[(size_t)i * ldc + j] : 0.f);
}

// Packs up to 16 rows x columns [p0, p1) of a row-major matrix into a panel of width W: dst[(p - p0) * W + off + i]
// (rows beyond hgt read as zero; the first popcount(sm) lanes of a column are stored).  16 x 16 blocks are transposed in registers - unpack 32-bit, unpack 64-bit, two
// rounds of 128-bit lane shuffles - so both the reads and the writes are whole vectors instead of a 4-byte scatter.
__attribute__((target("avx512f"), noinline))
void pack_rows_avx512(const float* src, int ld, int hgt, int p0, int p1, float* dst, int W, int off, unsigned sm) {
    for (int p = p0; p < p1; p += 16) {
        const int cols = std::min(16, p1 - p);
        const __mmask16 lm = cols >= 16 ? (__mmask16)0xffff : (__mmask16)((1u << cols) - 1);
        __m512 r[16], t[16], u[16];
        for (int i = 0; i < 16; ++i)
            r[i] = i < hgt ? _mm512_maskz_loadu_ps(lm, src + (size_t)i * ld + p) : _mm512_setzero_ps();
        for (int i = 0; i < 8; ++i) {
            t[2 * i] = _mm512_unpacklo_ps(r[2 * i], r[2 * i + 1]);
            t[2 * i + 1] = _mm512_unpackhi_ps(r[2 * i], r[2 * i + 1]);
        }
        for (int g = 0; g < 4; ++g) {                    // u[4g + j], lane L: rows 4g..4g+3 of column 4L + j
            __m512d a0 = _mm512_castps_pd(t[4 * g]), a1 = _mm512_castps_pd(t[4 * g + 1]);
            __m512d a2 = _mm512_castps_pd(t[4 * g + 2]), a3 = _mm512_castps_pd(t[4 * g + 3]);
            u[4 * g] = _mm512_castpd_ps(_mm512_unpacklo_pd(a0, a2));
            u[4 * g + 1] = _mm512_castpd_ps(_mm512_unpackhi_pd(a0, a2));
            u[4 * g + 2] = _mm512_castpd_ps(_mm512_unpacklo_pd(a1, a3));
            u[4 * g + 3] = _mm512_castpd_ps(_mm512_unpackhi_pd(a1, a3));
        }
        __m512 out[16];
        for (int j = 0; j < 4; ++j) {
            __m512 a = _mm512_shuffle_f32x4(u[j], u[4 + j], 0x88), b = _mm512_shuffle_f32x4(u[8 + j], u[12 + j], 0x88);
            __m512 c = _mm512_shuffle_f32x4(u[j], u[4 + j], 0xdd), d = _mm512_shuffle_f32x4(u[8 + j], u[12 + j], 0xdd);
            out[j] = _mm512_shuffle_f32x4(a, b, 0x88);
            out[8 + j] = _mm512_shuffle_f32x4(a, b, 0xdd);
            out[4 + j] = _mm512_shuffle_f32x4(c, d, 0x88);
            out[12 + j] = _mm512_shuffle_f32x4(c, d, 0xdd);
        }
        for (int q = 0; q < cols; ++q) _mm512_mask_storeu_ps(dst + (size_t)(p - p0 + q) * W + off, (__mmask16)sm, out[q]);
    }
}

// op(B)[p0:p1, col0:col0 + w] -> dst[(p - p0) * NR + j], columns beyond w zero-filled
void pack_b_chunk(const float* B, int ldb, int transB, int col0, int w, int p0, int p1, int NR, float* dst) {
    if (!transB) {
        for (int p = p0; p < p1; ++p) {
            float* d = dst + (size_t)(p - p0) * NR;
            const float* src = B + (size_t)p * ldb + col0;
            if (w == NR) { std::memcpy(d, src, sizeof(float) * NR); continue; }
            for (int j = 0; j < w; ++j) d[j] = src[j];
            for (int j = w; j < NR; ++j) d[j] = 0.f;
        }
    } else if (NR == 32) {                                           // AVX-512 kernel: two 16-row register transposes
        for (int o = 0; o < 32; o += 16)
            pack_rows_avx512(B + (size_t)(col0 + o) * ldb, ldb, std::max(0, std::min(16, w - o)), p0, p1, dst, 32, o, 0xffffu);
    } else {
        for (int j = 0; j < NR; ++j) {                               // contiguous reads along K
            if (j < w) {
                const float* src = B + (size_t)(col0 + j) * ldb;
                for (int p = p0; p < p1; ++p) dst[(size_t)(p - p0) * NR + j] = src[p];
            } else {
                for (int p = p0; p < p1; ++p) dst[(size_t)(p - p0) * NR + j] = 0.f;
            }
        }
    }
}

const Kern& pick_kernel() {
    static const Kern k = [] {
        __builtin_cpu_init();
        if (__builtin_cpu_supports("avx512f")) return Kern{12, 32, kern_avx512};
        if (__builtin_cpu_supports("avx2") && __builtin_cpu_supports("fma")) return Kern{6, 16, kern_avx2};
        return Kern{4, 8, kern_scalar};
    }();
    return k;
}

inline int round_up(int x, int m) { return (x + m - 1) / m * m; }
}  // namespace

API int hetu_cpu_matmul(const float* A, const float* B, float* C, int M, int N, int K, int transA, int transB,
                        float alpha, float beta) {
    if (M <= 0 || N <= 0) return 0;
    const Kern& kr = pick_kernel();
    const int MR = kr.MR, NR = kr.NR, KC = kc_setting();
    if (K <= 0) {
#pragma omp parallel for
        for (int64_t i = 0; i < (int64_t)M * N; ++i) C[i] = beta == 0.f ? 0.f : beta * C[i];
        return 0;
    }
    const int lda = transA ? M : K, ldb = transB ? K : N;
    const int mpanels = (M + MR - 1) / MR, npanels = (N + NR - 1) / NR;
    // op(A) is packed once into row panels [panel][K][MR] (each element is copied once, in parallel).
    // op(B) is read in place when it is row-major K x N with a short K (a panel row is 32 consecutive floats).
    // Otherwise - transposed layout, long K (rows ldb floats apart defeat the L1 prefetcher and collide in its
    // sets), ragged last panel - the thread that is about to use a KC x NR chunk packs it into its own buffer
    // right before the micro-kernel calls: the chunk stays in that core's L1/L2 and there is no B stage or barrier.
    static thread_local std::vector<float> apack_store, bchunk_store;
    std::vector<float>& apack = apack_store;
    apack.resize((size_t)mpanels * K * MR);
    const bool pack_all_b = transB != 0 || (K >= 512 && getenv("HETU_CPU_GEMM_INPLACE_B") == nullptr);
    float* ap_base = apack.data();
    const bool go_parallel = (int64_t)M * N * K >= (1 << 18);          // tiny products: the fork/join costs more than it saves
#pragma omp parallel if (go_parallel)
    {
        const int kchunks = (K + KC - 1) / KC;                       // packing work items = panel x K chunk, so few panels still fill all threads
#pragma omp for schedule(static) collapse(2) nowait
        for (int ip = 0; ip < mpanels; ++ip)
            for (int kb = 0; kb < kchunks; ++kb) {
                const int p0 = kb * KC, p1 = std::min(K, p0 + KC);
                float* dst = ap_base + (size_t)ip * K * MR;
                const int row0 = ip * MR, hgt = std::min(MR, M - row0);
                if (!transA && MR == 12) {
                    pack_rows_avx512(A + (size_t)row0 * lda, lda, hgt, p0, p1, dst + (size_t)p0 * MR, 12, 0, 0x0fffu);
                } else if (!transA) {
                    for (int i = 0; i < MR; ++i) {                   // contiguous reads along K, strided writes
                        if (i < hgt) {
                            const float* src = A + (size_t)(row0 + i) * lda;
                            for (int p = p0; p < p1; ++p) dst[(size_t)p * MR + i] = src[p];
                        } else {
                            for (int p = p0; p < p1; ++p) dst[(size_t)p * MR + i] = 0.f;
                        }
                    }
                } else {
                    for (int p = p0; p < p1; ++p) {
                        const float* src = A + (size_t)p * lda + row0;
                        float* d = dst + (size_t)p * MR;
                        for (int i = 0; i < hgt; ++i) d[i] = src[i];
                        for (int i = hgt; i < MR; ++i) d[i] = 0.f;
                    }
                }
            }
#pragma omp barrier
        bchunk_store.resize((size_t)KC * NR);                        // this thread's B chunk
        float* bchunk = bchunk_store.data();
        // tiles of C: a few row panels x a few column panels each, about two tiles per thread at least
        const int nthr = omp_get_num_threads();
        int tp_m = std::max(1, std::min(mpanels, 96 / MR)), tp_n = std::max(1, std::min(npanels, 128 / NR));
        auto ntiles = [&](int a, int b) { return (int64_t)((mpanels + a - 1) / a) * ((npanels + b - 1) / b); };
        while (ntiles(tp_m, tp_n) < 2 * (int64_t)nthr && (tp_m > 1 || tp_n > 1)) {
            if (tp_n > 1 && (tp_n * NR >= tp_m * MR || tp_m == 1)) tp_n = (tp_n + 1) / 2;
            else tp_m = (tp_m + 1) / 2;
        }
        const int tiles_m = (mpanels + tp_m - 1) / tp_m, tiles_n = (npanels + tp_n - 1) / tp_n;
#pragma omp for schedule(dynamic) collapse(2)
        for (int tm = 0; tm < tiles_m; ++tm)
            for (int tn = 0; tn < tiles_n; ++tn) {
                const int ip0 = tm * tp_m, ip1 = std::min(mpanels, ip0 + tp_m);
                const int jp0 = tn * tp_n, jp1 = std::min(npanels, jp0 + tp_n);
                for (int pc = 0; pc < K; pc += KC) {
                    const int kc = std::min(KC, K - pc);
                    const float b_eff = pc == 0 ? beta : 1.f;       // later K chunks accumulate
                    for (int jp = jp0; jp < jp1; ++jp) {
                        const int col = jp * NR, nr = std::min(NR, N - col);
                        const float* bptr;
                        int ldbk;
                        if (pack_all_b || nr < NR) {
                            pack_b_chunk(B, ldb, transB, col, nr, pc, pc + kc, NR, bchunk);
                            bptr = bchunk;
                            ldbk = NR;
                        } else {
                            bptr = B + (size_t)pc * ldb + col;
                            ldbk = ldb;
                        }
                        const auto fn = (NR == 32 && nr <= 16) ? kern_avx512_n16 : kr.fn;
                        for (int ip = ip0; ip < ip1; ++ip) {
                            const int row = ip * MR;
                            fn(kc, ap_base + (size_t)ip * K * MR + (size_t)pc * MR, bptr, ldbk, C + (size_t)row * N + col, N,
                                  std::min(MR, M - row), nr, alpha, b_eff);
                        }
                    }
                }
            }
    }
    return 0;
}

API int hetu_cpu_batch_matmul(const float* A, const float* B, float* C, int batch, int M, int N, int K, int transA,
                              int transB) {
    for (int b = 0; b < batch; ++b)
        hetu_cpu_matmul(A + (size_t)b * M * K, B + (size_t)b * K * N, C + (size_t)b * M * N, M, N, K, transA, transB,
                        1.f, 0.f);
    return 0;
}

// ------------------------------------------------------------------------------------------------
// element-wise (reference AddConst/AddElewise/MultiplyConst/MultiplyElewise/DivideConst/DivideElewise/
// Opposite/Sqrt/Relu/Sigmoid/Tanh/Gelu .cpp).  op codes are private to this library.
// below this many elements an OpenMP region costs more than it saves (waking the pool takes tens of microseconds)
constexpr int64_t kParallelMin = 1 << 16;
enum { U_RELU = 0, U_SIGMOID, U_TANH, U_GELU, U_SQRT, U_RSQRT, U_NEG, U_EXP, U_LOG, U_ABS, U_ADDC, U_MULC, U_CDIV };
enum { B_ADD = 0, B_SUB, B_MUL, B_DIV, B_RELU_GRAD, B_SIGMOID_GRAD, B_TANH_GRAD, B_GELU_GRAD };

// the op switch sits outside the loop: each instantiation is a straight-line loop the compiler vectorises
namespace {
template <class F>
inline void map1(const float* x, float* y, int64_t n, F f) {
    if (n < kParallelMin) {
#pragma omp simd
        for (int64_t i = 0; i < n; ++i) y[i] = f(x[i]);
        return;
    }
#pragma omp parallel for simd
    for (int64_t i = 0; i < n; ++i) y[i] = f(x[i]);
}
template <class Body>
inline void for_n(int64_t n, Body body) {
    if (n < kParallelMin) {
#pragma omp simd
        for (int64_t i = 0; i < n; ++i) body(i);
        return;
    }
#pragma omp parallel for simd
    for (int64_t i = 0; i < n; ++i) body(i);
}
template <class F>
inline void map2(const float* x, const float* z, float* y, int64_t n, F f) {
    if (n < kParallelMin) {
#pragma omp simd
        for (int64_t i = 0; i < n; ++i) y[i] = f(x[i], z[i]);
        return;
    }
#pragma omp parallel for simd
    for (int64_t i = 0; i < n; ++i) y[i] = f(x[i], z[i]);
}
}  // namespace

API int hetu_cpu_unary(const float* x, float* y, int64_t n, int op, float c) {
    switch (op) {
        case U_RELU: map1(x, y, n, [](float v) { return v > 0.f ? v : 0.f; }); break;
        case U_SIGMOID: map1(x, y, n, [](float v) { return 1.f / (1.f + std::exp(-v)); }); break;
        case U_TANH: map1(x, y, n, [](float v) { return std::tanh(v); }); break;
        case U_GELU: map1(x, y, n, [](float v) { return 0.5f * v * (1.f + std::erf(v * 0.70710678118f)); }); break;
        case U_SQRT: map1(x, y, n, [](float v) { return std::sqrt(v); }); break;
        case U_RSQRT: map1(x, y, n, [](float v) { return 1.f / std::sqrt(v); }); break;
        case U_NEG: map1(x, y, n, [](float v) { return -v; }); break;
        case U_EXP: map1(x, y, n, [](float v) { return std::exp(v); }); break;
        case U_LOG: map1(x, y, n, [](float v) { return std::log(v); }); break;
        case U_ABS: map1(x, y, n, [](float v) { return std::fabs(v); }); break;
        case U_ADDC: map1(x, y, n, [c](float v) { return v + c; }); break;
        case U_MULC: map1(x, y, n, [c](float v) { return v * c; }); break;
        default: map1(x, y, n, [c](float v) { return c / v; }); break;
    }
    return 0;
}

// y = x (op) z ; for the *_GRAD codes x is the forward input (or output for sigmoid/tanh) and z the adjoint
API int hetu_cpu_binary(const float* x, const float* z, float* y, int64_t n, int op) {
    switch (op) {
        case B_ADD: map2(x, z, y, n, [](float a, float b) { return a + b; }); break;
        case B_SUB: map2(x, z, y, n, [](float a, float b) { return a - b; }); break;
        case B_MUL: map2(x, z, y, n, [](float a, float b) { return a * b; }); break;
        case B_DIV: map2(x, z, y, n, [](float a, float b) { return a / b; }); break;
        case B_RELU_GRAD: map2(x, z, y, n, [](float a, float b) { return a > 0.f ? b : 0.f; }); break;
        case B_SIGMOID_GRAD: map2(x, z, y, n, [](float a, float b) { return b * a * (1.f - a); }); break;
        case B_TANH_GRAD: map2(x, z, y, n, [](float a, float b) { return b * (1.f - a * a); }); break;
        default:
            map2(x, z, y, n, [](float a, float b) {
                float cdf = 0.5f * (1.f + std::erf(a * 0.70710678118f));
                float pdf = 0.3989422804f * std::exp(-0.5f * a * a);
                return b * (cdf + a * pdf);
            });
    }
    return 0;
}

API int hetu_cpu_array_set(float* x, int64_t n, float v) {
    for_n(n, [=](int64_t i) { x[i] = v; });
    return 0;
}

API int hetu_cpu_reshape(const float* x, float* y, int64_t n) {
    if (x != y) std::memcpy(y, x, sizeof(float) * n);
    return 0;
}

// bias-style broadcast of x[cols] (or x[rows/rep, cols]) over the leading dim (BroadcastTo.cpp)
API int hetu_cpu_broadcast_to(const float* x, float* y, int64_t in_n, int64_t out_n) {
#pragma omp parallel for if (out_n >= kParallelMin)
    for (int64_t i = 0; i < out_n; ++i) y[i] = x[i % in_n];
    return 0;
}

API int hetu_cpu_reduce_sum_axis0(const float* x, float* y, int64_t rows, int64_t cols) {
#pragma omp parallel for
    for (int64_t j = 0; j < cols; ++j) {
        float s = 0.f;
        for (int64_t i = 0; i < rows; ++i) s += x[i * cols + j];
        y[j] = s;
    }
    return 0;
}

API int hetu_cpu_transpose(const float* x, float* y, const int64_t* in_shape, const int* perm, int ndim) {
    int64_t in_stride[8], out_shape[8], n = 1;
    in_stride[ndim - 1] = 1;
    for (int d = ndim - 2; d >= 0; --d) in_stride[d] = in_stride[d + 1] * in_shape[d + 1];
    for (int d = 0; d < ndim; ++d) { out_shape[d] = in_shape[perm[d]]; n *= out_shape[d]; }
#pragma omp parallel for
    for (int64_t o = 0; o < n; ++o) {
        int64_t rem = o, src = 0;
        for (int d = ndim - 1; d >= 0; --d) {
            src += (rem % out_shape[d]) * in_stride[perm[d]];
            rem /= out_shape[d];
        }
        y[o] = x[src];
    }
    return 0;
}

API int hetu_cpu_concat(const float* a, const float* b, float* y, int64_t outer, int64_t a_inner, int64_t b_inner) {
#pragma omp parallel for
    for (int64_t o = 0; o < outer; ++o) {
        std::memcpy(y + o * (a_inner + b_inner), a + o * a_inner, sizeof(float) * a_inner);
        std::memcpy(y + o * (a_inner + b_inner) + a_inner, b + o * b_inner, sizeof(float) * b_inner);
    }
    return 0;
}

API int hetu_cpu_concat_grad(const float* g, float* ga, int64_t outer, int64_t a_inner, int64_t total_inner,
                             int64_t offset) {
#pragma omp parallel for
    for (int64_t o = 0; o < outer; ++o)
        std::memcpy(ga + o * a_inner, g + o * total_inner + offset, sizeof(float) * a_inner);
    return 0;
}

// NCHW zero padding of the two spatial dims (Pad.cpp)
API int hetu_cpu_pad(const float* x, float* y, int64_t nc, int H, int W, int pt, int pb, int pl, int pr, float value) {
    int Ho = H + pt + pb, Wo = W + pl + pr;
#pragma omp parallel for
    for (int64_t p = 0; p < nc; ++p)
        for (int h = 0; h < Ho; ++h)
            for (int w = 0; w < Wo; ++w) {
                int hi = h - pt, wi = w - pl;
                y[(p * Ho + h) * Wo + w] = (hi >= 0 && hi < H && wi >= 0 && wi < W) ? x[(p * H + hi) * W + wi] : value;
            }
    return 0;
}

API int hetu_cpu_pad_grad(const float* gy, float* gx, int64_t nc, int H, int W, int pt, int pb, int pl, int pr) {
    int Ho = H + pt + pb, Wo = W + pl + pr;
#pragma omp parallel for
    for (int64_t p = 0; p < nc; ++p)
        for (int h = 0; h < H; ++h)
            for (int w = 0; w < W; ++w) gx[(p * H + h) * W + w] = gy[(p * Ho + h + pt) * Wo + w + pl];
    return 0;
}

// ------------------------------------------------------------------------------------------------
// softmax family (Softmax.cpp, SoftmaxCrossEntropy.cpp)
API int hetu_cpu_softmax(const float* x, float* y, int64_t rows, int64_t cols) {
#pragma omp parallel for if (rows * cols >= kParallelMin)
    for (int64_t r = 0; r < rows; ++r) {
        const float* xr = x + r * cols;
        float* yr = y + r * cols;
        float m = xr[0];
        for (int64_t j = 1; j < cols; ++j) m = std::max(m, xr[j]);
        float s = 0.f;
        for (int64_t j = 0; j < cols; ++j) { yr[j] = std::exp(xr[j] - m); s += yr[j]; }
        float inv = 1.f / s;
        for (int64_t j = 0; j < cols; ++j) yr[j] *= inv;
    }
    return 0;
}

API int hetu_cpu_softmax_cross_entropy(const float* x, const float* label, float* loss, int64_t rows, int64_t cols) {
#pragma omp parallel for if (rows * cols >= kParallelMin)
    for (int64_t r = 0; r < rows; ++r) {
        const float* xr = x + r * cols;
        const float* lr = label + r * cols;
        float m = xr[0];
        for (int64_t j = 1; j < cols; ++j) m = std::max(m, xr[j]);
        float s = 0.f;
        for (int64_t j = 0; j < cols; ++j) s += std::exp(xr[j] - m);
        float lse = m + std::log(s), acc = 0.f;
        for (int64_t j = 0; j < cols; ++j) acc += lr[j] * (lse - xr[j]);
        loss[r] = acc;
    }
    return 0;
}

API int hetu_cpu_softmax_cross_entropy_grad(const float* x, const float* label, const float* g, float* gx,
                                            int64_t rows, int64_t cols) {
#pragma omp parallel for if (rows * cols >= kParallelMin)
    for (int64_t r = 0; r < rows; ++r) {
        const float* xr = x + r * cols;
        const float* lr = label + r * cols;
        float* gr = gx + r * cols;
        float m = xr[0];
        for (int64_t j = 1; j < cols; ++j) m = std::max(m, xr[j]);
        float s = 0.f, lsum = 0.f;
        for (int64_t j = 0; j < cols; ++j) { gr[j] = std::exp(xr[j] - m); s += gr[j]; lsum += lr[j]; }
        float inv = 1.f / s, gv = g[r];
        for (int64_t j = 0; j < cols; ++j) gr[j] = gv * (gr[j] * inv * lsum - lr[j]);
    }
    return 0;
}

// ------------------------------------------------------------------------------------------------
// convolution = im2col + GEMM (Conv2d.cpp uses dnnl::convolution_forward / backward_data / backward_weights)
namespace {
void im2col(const float* x, float* col, int C, int H, int W, int kh, int kw, int pad, int stride, int Ho, int Wo) {
#pragma omp parallel for
    for (int c = 0; c < C * kh * kw; ++c) {
        int wk = c % kw, hk = (c / kw) % kh, ci = c / (kw * kh);
        for (int ho = 0; ho < Ho; ++ho)
            for (int wo = 0; wo < Wo; ++wo) {
                int hi = ho * stride - pad + hk, wi = wo * stride - pad + wk;
                col[((size_t)c * Ho + ho) * Wo + wo] =
                    (hi >= 0 && hi < H && wi >= 0 && wi < W) ? x[((size_t)ci * H + hi) * W + wi] : 0.f;
            }
    }
}

void col2im(const float* col, float* x, int C, int H, int W, int kh, int kw, int pad, int stride, int Ho, int Wo) {
    std::memset(x, 0, sizeof(float) * (size_t)C * H * W);
#pragma omp parallel for
    for (int ci = 0; ci < C; ++ci)
        for (int hk = 0; hk < kh; ++hk)
            for (int wk = 0; wk < kw; ++wk) {
                int c = (ci * kh + hk) * kw + wk;
                for (int ho = 0; ho < Ho; ++ho)
                    for (int wo = 0; wo < Wo; ++wo) {
                        int hi = ho * stride - pad + hk, wi = wo * stride - pad + wk;
                        if (hi >= 0 && hi < H && wi >= 0 && wi < W)
                            x[((size_t)ci * H + hi) * W + wi] += col[((size_t)c * Ho + ho) * Wo + wo];
                    }
            }
}
}  // namespace

API int hetu_cpu_conv2d(const float* x, const float* w, float* y, int N, int C, int H, int W, int F, int kh, int kw,
                        int pad, int stride) {
    int Ho = (H + 2 * pad - kh) / stride + 1, Wo = (W + 2 * pad - kw) / stride + 1;
    std::vector<float> col((size_t)C * kh * kw * Ho * Wo);
    for (int n = 0; n < N; ++n) {
        im2col(x + (size_t)n * C * H * W, col.data(), C, H, W, kh, kw, pad, stride, Ho, Wo);
        hetu_cpu_matmul(w, col.data(), y + (size_t)n * F * Ho * Wo, F, Ho * Wo, C * kh * kw, 0, 0, 1.f, 0.f);
    }
    return 0;
}

API int hetu_cpu_conv2d_grad_data(const float* w, const float* gy, float* gx, int N, int C, int H, int W, int F,
                                  int kh, int kw, int pad, int stride) {
    int Ho = (H + 2 * pad - kh) / stride + 1, Wo = (W + 2 * pad - kw) / stride + 1;
    std::vector<float> col((size_t)C * kh * kw * Ho * Wo);
    for (int n = 0; n < N; ++n) {
        hetu_cpu_matmul(w, gy + (size_t)n * F * Ho * Wo, col.data(), C * kh * kw, Ho * Wo, F, 1, 0, 1.f, 0.f);
        col2im(col.data(), gx + (size_t)n * C * H * W, C, H, W, kh, kw, pad, stride, Ho, Wo);
    }
    return 0;
}

API int hetu_cpu_conv2d_grad_filter(const float* x, const float* gy, float* gw, int N, int C, int H, int W, int F,
                                    int kh, int kw, int pad, int stride) {
    int Ho = (H + 2 * pad - kh) / stride + 1, Wo = (W + 2 * pad - kw) / stride + 1;
    std::vector<float> col((size_t)C * kh * kw * Ho * Wo);
    for (int n = 0; n < N; ++n) {
        im2col(x + (size_t)n * C * H * W, col.data(), C, H, W, kh, kw, pad, stride, Ho, Wo);
        hetu_cpu_matmul(gy + (size_t)n * F * Ho * Wo, col.data(), gw, F, C * kh * kw, Ho * Wo, 0, 1, 1.f,
                        n == 0 ? 0.f : 1.f);
    }
    return 0;
}

// pooling (MaxPool.cpp / AvgPool.cpp).  mode 0 = max, 1 = average (padding counted, as oneDNN
// pooling_avg_include_padding which the reference selects)
API int hetu_cpu_pool2d(const float* x, float* y, int64_t nc, int H, int W, int kh, int kw, int pad, int stride,
                        int mode) {
    int Ho = (H + 2 * pad - kh) / stride + 1, Wo = (W + 2 * pad - kw) / stride + 1;
#pragma omp parallel for
    for (int64_t p = 0; p < nc; ++p)
        for (int ho = 0; ho < Ho; ++ho)
            for (int wo = 0; wo < Wo; ++wo) {
                float acc = mode == 0 ? -INFINITY : 0.f;
                for (int i = 0; i < kh; ++i)
                    for (int j = 0; j < kw; ++j) {
                        int hi = ho * stride - pad + i, wi = wo * stride - pad + j;
                        if (hi < 0 || hi >= H || wi < 0 || wi >= W) continue;
                        float v = x[(p * H + hi) * W + wi];
                        acc = mode == 0 ? std::max(acc, v) : acc + v;
                    }
                y[(p * Ho + ho) * Wo + wo] = mode == 0 ? acc : acc / (kh * kw);
            }
    return 0;
}

API int hetu_cpu_pool2d_grad(const float* x, const float* y, const float* gy, float* gx, int64_t nc, int H, int W,
                             int kh, int kw, int pad, int stride, int mode) {
    int Ho = (H + 2 * pad - kh) / stride + 1, Wo = (W + 2 * pad - kw) / stride + 1;
#pragma omp parallel for
    for (int64_t p = 0; p < nc; ++p) {
        float* gp = gx + p * H * W;
        std::memset(gp, 0, sizeof(float) * H * W);
        for (int ho = 0; ho < Ho; ++ho)
            for (int wo = 0; wo < Wo; ++wo) {
                float g = gy[(p * Ho + ho) * Wo + wo], out = y[(p * Ho + ho) * Wo + wo];
                bool done = false;
                for (int i = 0; i < kh; ++i)
                    for (int j = 0; j < kw; ++j) {
                        int hi = ho * stride - pad + i, wi = wo * stride - pad + j;
                        if (hi < 0 || hi >= H || wi < 0 || wi >= W) continue;
                        if (mode == 1) gp[hi * W + wi] += g / (kh * kw);
                        else if (!done && x[(p * H + hi) * W + wi] == out) { gp[hi * W + wi] += g; done = true; }
                    }
            }
    }
    return 0;
}

// batch norm over NCHW (BatchNorm.cpp).  Training mode: batch statistics, running stats updated with
// ``momentum``; save_mean / save_invstd are kept for the backward pass.
API int hetu_cpu_batch_norm(const float* x, const float* scale, const float* bias, float* y, float* running_mean,
                            float* running_var, float* save_mean, float* save_invstd, int N, int C, int HW,
                            float momentum, float eps, int training) {
#pragma omp parallel for
    for (int c = 0; c < C; ++c) {
        float mean, invstd;
        if (training) {
            double s = 0, ss = 0;
            for (int n = 0; n < N; ++n)
                for (int i = 0; i < HW; ++i) { double v = x[((size_t)n * C + c) * HW + i]; s += v; ss += v * v; }
            double cnt = (double)N * HW;
            mean = (float)(s / cnt);
            float var = (float)(ss / cnt - (s / cnt) * (s / cnt));
            invstd = 1.f / std::sqrt(var + eps);
            running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
            running_var[c] = (1.f - momentum) * running_var[c] + momentum * var;
            save_mean[c] = mean;
            save_invstd[c] = invstd;
        } else {
            mean = running_mean[c];
            invstd = 1.f / std::sqrt(running_var[c] + eps);
        }
        for (int n = 0; n < N; ++n)
            for (int i = 0; i < HW; ++i) {
                size_t k = ((size_t)n * C + c) * HW + i;
                y[k] = (x[k] - mean) * invstd * scale[c] + bias[c];
            }
    }
    return 0;
}

API int hetu_cpu_batch_norm_grad(const float* x, const float* gy, const float* scale, const float* save_mean,
                                 const float* save_invstd, float* gx, float* gscale, float* gbias, int N, int C,
                                 int HW) {
#pragma omp parallel for
    for (int c = 0; c < C; ++c) {
        double sg = 0, sgx = 0;
        float mean = save_mean[c], invstd = save_invstd[c];
        for (int n = 0; n < N; ++n)
            for (int i = 0; i < HW; ++i) {
                size_t k = ((size_t)n * C + c) * HW + i;
                sg += gy[k];
                sgx += gy[k] * (x[k] - mean) * invstd;
            }
        gbias[c] = (float)sg;
        gscale[c] = (float)sgx;
        float cnt = (float)N * HW;
        for (int n = 0; n < N; ++n)
            for (int i = 0; i < HW; ++i) {
                size_t k = ((size_t)n * C + c) * HW + i;
                float xh = (x[k] - mean) * invstd;
                gx[k] = scale[c] * invstd * (gy[k] - (float)sg / cnt - xh * (float)sgx / cnt);
            }
    }
    return 0;
}

// ------------------------------------------------------------------------------------------------
// dropout: counter-based hash so the backward pass regenerates the mask from (seed, index)
// (Dropout.cpp keeps the seed for the same purpose)
static inline uint32_t mix32(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return (uint32_t)((z ^ (z >> 31)) >> 32);
}

API int hetu_cpu_dropout(const float* x, float* y, int64_t n, float keep_prob, uint64_t seed) {
    float inv = 1.f / keep_prob;
    for_n(n, [=](int64_t i) {
        float u = mix32(seed * 0x100000001B3ull + (uint64_t)i) * (1.f / 4294967296.f);
        y[i] = u < keep_prob ? x[i] * inv : 0.f;
    });
    return 0;
}

// ------------------------------------------------------------------------------------------------
// embedding (EmbeddingLookup.cpp:15-35) and optimizers (Optimizers.cpp:19-60)
API int hetu_cpu_embedding_lookup(const float* table, const int64_t* ids, float* out, int64_t nids, int64_t width,
                                  int64_t nrows) {
#pragma omp parallel for
    for (int64_t i = 0; i < nids; ++i) {
        int64_t r = ids[i];
        if (r < 0 || r >= nrows) std::memset(out + i * width, 0, sizeof(float) * width);
        else std::memcpy(out + i * width, table + r * width, sizeof(float) * width);
    }
    return 0;
}

API int hetu_cpu_sgd_update(float* p, const float* g, int64_t n, float lr, float l2) {
    for_n(n, [=](int64_t i) { p[i] -= lr * (g[i] + l2 * p[i]); });
    return 0;
}

// rows may repeat: serial over ids, parallel over the row (matches the reference's accumulate semantics)
API int hetu_cpu_sgd_sparse_update(float* p, const int64_t* ids, const float* g, int64_t nids, int64_t width,
                                   float lr) {
    for (int64_t i = 0; i < nids; ++i) {
        float* row = p + ids[i] * width;
        const float* gr = g + i * width;
#pragma omp simd
        for (int64_t j = 0; j < width; ++j) row[j] -= lr * gr[j];
    }
    return 0;
}

API int hetu_cpu_momentum_update(float* p, const float* g, float* vel, int64_t n, float lr, float momentum,
                                 int nesterov, float l2) {
    for_n(n, [=](int64_t i) {
        float gi = g[i] + l2 * p[i];
        float v = momentum * vel[i] - lr * gi;
        vel[i] = v;
        p[i] += nesterov ? momentum * v - lr * gi : v;
    });
    return 0;
}

API int hetu_cpu_adagrad_update(float* p, const float* g, float* acc, int64_t n, float lr, float eps, float l2) {
    for_n(n, [=](int64_t i) {
        float gi = g[i] + l2 * p[i];
        acc[i] += gi * gi;
        p[i] -= lr * gi / (std::sqrt(acc[i]) + eps);
    });
    return 0;
}

API int hetu_cpu_adam_update(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1,
                             float beta2, float eps, int64_t step, float l2, int amsgrad, float* vmax) {
    float c1 = 1.f - std::pow(beta1, (float)step), c2 = 1.f - std::pow(beta2, (float)step);
    for_n(n, [=](int64_t i) {
        float gi = g[i] + l2 * p[i];
        m[i] = beta1 * m[i] + (1.f - beta1) * gi;
        v[i] = beta2 * v[i] + (1.f - beta2) * gi * gi;
        float vh = v[i];
        if (amsgrad) { vmax[i] = std::max(vmax[i], vh); vh = vmax[i]; }
        p[i] -= lr * (m[i] / c1) / (std::sqrt(vh / c2) + eps);
    });
    return 0;
}

// ------------------------------------------------------------------------------------------------
// initialisers (Initializers.cpp): one engine per thread, seeded from (seed, thread)
API int hetu_cpu_uniform_init(float* x, int64_t n, float lo, float hi, uint64_t seed) {
#pragma omp parallel
    {
        std::mt19937_64 gen(seed * 1000003ull + omp_get_thread_num());
        std::uniform_real_distribution<float> d(lo, hi);
#pragma omp for schedule(static)
        for (int64_t i = 0; i < n; ++i) x[i] = d(gen);
    }
    return 0;
}

API int hetu_cpu_normal_init(float* x, int64_t n, float mean, float stddev, uint64_t seed) {
#pragma omp parallel
    {
        std::mt19937_64 gen(seed * 1000003ull + omp_get_thread_num());
        std::normal_distribution<float> d(mean, stddev);
#pragma omp for schedule(static)
        for (int64_t i = 0; i < n; ++i) x[i] = d(gen);
    }
    return 0;
}

API int hetu_cpu_truncated_normal_init(float* x, int64_t n, float mean, float stddev, uint64_t seed) {
#pragma omp parallel
    {
        std::mt19937_64 gen(seed * 1000003ull + omp_get_thread_num());
        std::normal_distribution<float> d(mean, stddev);
#pragma omp for schedule(static)
        for (int64_t i = 0; i < n; ++i) {
            float v;
            do { v = d(gen); } while (std::fabs(v - mean) > 2.f * stddev);
            x[i] = v;
        }
    }
    return 0;
}

API int hetu_cpu_num_threads() { return omp_get_max_threads(); }
