// Flash-style attention forward on tcgen05 (sm_100a): O = softmax(Q K^T * scale + key mask) V, one kernel.
//
// reference: examples/nlp/bert/hetu_bert.py:249-264 builds attention from batch_matmul, a broadcast mask add, softmax
// and batch_matmul, so the [B, H, S, S] score tensor goes through HBM four times.  Here the scores never leave the SM:
//
//   one CTA per (128 query rows, batch*head); warp 4 lane 0 issues TMA and MMA, warps 0-3 own the TMEM lane quadrants
//   (thread = query row) and do the online softmax.
//
//   per block j of 64 keys:  K_j, V_j by TMA (fp32, 128B swizzle; V as an MN-major operand, so no transpose anywhere)
//                         S    = Q K_j^T          tcgen05.mma kind::tf32 -> TMEM columns [0, BN)
//                         p    = exp(S * scale + mask - m_new) in registers, row max / sum updated, O rescaled
//                         P   -> shared memory, written by the row's thread in the K-major swizzle-128B layout
//                         PV   = P V_j            second tcgen05.mma -> TMEM columns [BN, BN + D)
//                         O   += PV               (registers)
//   K_{j+1} is fetched as soon as S_j is complete and V_{j+1} as soon as PV_j is, so the loads hide behind the softmax.
//   epilogue: O / l -> global, lse = m + log(l) for the backward pass.
//
// Operands are fp32 in memory (the framework's storage type) and multiplied as tf32, like every other GEMM here.
#include "tcgen05.cuh"

namespace {

constexpr int FA_BM = 128;          // query rows per CTA
constexpr int FA_THREADS = 160;     // 4 softmax warps + 1 issue warp

template <int D, int FA_BN>
struct FaSmem {
    static constexpr int kQBytes = FA_BM * D * 4;                 // D/32 tiles of [128 rows][128 B]
    static constexpr int kKBytes = FA_BN * D * 4;                 // same shape
    static constexpr int kVBytes = FA_BN * D * 4;                 // FA_BN/32 key blocks x D/32 groups of [32 keys][128 B]
    static constexpr int kPBytes = FA_BM * FA_BN * 4;             // FA_BN/32 tiles of [128 rows][128 B]
    static constexpr int kTmemCols = (FA_BN + D) <= 128 ? 128 : 256;
    static constexpr int kTotal = kQBytes + kKBytes + kVBytes + kPBytes + 1024 /*alignment*/ + 1024 /*barriers, per-key bias*/;
};

__device__ __forceinline__ float fast_exp2(float x) {        // MUFU.EX2, flush-to-zero; exp2(-inf) = 0
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// wait of the many softmax threads: back off between polls so the spin does not take issue slots from working warps
__device__ __forceinline__ void mbar_wait_backoff(uint64_t* bar, uint32_t parity) {
    uint32_t done, addr = smem_u32(bar);
    for (;;) {
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t"
            "}" : "=r"(done) : "r"(addr), "r"(parity) : "memory");
        if (done) break;
        __nanosleep(40);
    }
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

template <int D, int FA_BN>
__global__ void __launch_bounds__(FA_THREADS, 1)
flash_attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                      const __grid_constant__ CUtensorMap tmV, const float* __restrict__ mask, float* __restrict__ O,
                      float* __restrict__ lse, int S, int Sk, int heads, float scale) {
    extern __shared__ uint8_t fa_smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(fa_smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* smQ = base;
    uint8_t* smK = smQ + FaSmem<D, FA_BN>::kQBytes;
    uint8_t* smV = smK + FaSmem<D, FA_BN>::kKBytes;
    uint8_t* smP = smV + FaSmem<D, FA_BN>::kVBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smP + FaSmem<D, FA_BN>::kPBytes);
    uint64_t* barQ = bars + 0;      // Q landed
    uint64_t* barK = bars + 1;      // K_j landed
    uint64_t* barV = bars + 2;      // V_j landed
    uint64_t* barS = bars + 3;      // S_j in TMEM (tcgen05.commit)
    uint64_t* barP = bars + 4;      // P_j in shared memory (128 arrivals)
    uint64_t* barO = bars + 5;      // PV_j in TMEM (tcgen05.commit)
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 6);
    float* bias_s = reinterpret_cast<float*>(bars + 16);        // [2][FA_BN] per-key additive term of the current block

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q0 = blockIdx.x * FA_BM;
    const int bh = blockIdx.y;
    const int nblk = (Sk + FA_BN - 1) / FA_BN;

    if (threadIdx.x == 0) {
        mbar_init(barQ, 1); mbar_init(barK, 1); mbar_init(barV, 1);
        mbar_init(barS, 1); mbar_init(barP, FA_BM); mbar_init(barO, 1);
        fence_barrier_init();
    }
    if (warp == 4) tmem_alloc(tmem_ptr, FaSmem<D, FA_BN>::kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    const uint32_t tmem_S = tmem_base, tmem_PV = tmem_base + FA_BN;

    if (warp == 4) {
        if (lane == 0) {
            prefetch_tmap(&tmQ); prefetch_tmap(&tmK); prefetch_tmap(&tmV);
            auto load_k = [&](int j) {
                mbar_expect_tx(barK, FaSmem<D, FA_BN>::kKBytes);
#pragma unroll
                for (int kb = 0; kb < D / 32; ++kb) tma_load_3d(smK + kb * (FA_BN * 128), &tmK, kb * 32, j * FA_BN, bh, barK);
            };
            auto load_v = [&](int j) {
                mbar_expect_tx(barV, FaSmem<D, FA_BN>::kVBytes);
#pragma unroll
                for (int kb = 0; kb < FA_BN / 32; ++kb)
#pragma unroll
                    for (int g = 0; g < D / 32; ++g)
                        tma_load_3d(smV + (kb * (D / 32) + g) * 4096, &tmV, g * 32, j * FA_BN + kb * 32, bh, barV);
            };
            mbar_expect_tx(barQ, FaSmem<D, FA_BN>::kQBytes);
#pragma unroll
            for (int kb = 0; kb < D / 32; ++kb) tma_load_3d(smQ + kb * (FA_BM * 128), &tmQ, kb * 32, q0, bh, barQ);
            load_k(0);
            load_v(0);
            // instruction descriptors: D = f32, A = B = tf32, N >> 3 @17, M >> 4 @24; bit 16 = B is MN-major
            const uint32_t idesc_s = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(FA_BN >> 3) << 17) | ((uint32_t)(FA_BM >> 4) << 24);
            const uint32_t idesc_o = (1u << 4) | (2u << 7) | (2u << 10) | (1u << 16) | ((uint32_t)(D >> 3) << 17) | ((uint32_t)(FA_BM >> 4) << 24);
            auto mma_s = [&]() {                          // S = Q K_j^T
#pragma unroll
                for (int kb = 0; kb < D / 32; ++kb)
#pragma unroll
                    for (int ki = 0; ki < 4; ++ki) {
                        const uint64_t a = make_smem_desc(smem_u32(smQ + kb * (FA_BM * 128)) + ki * 32, 16, 1024, 2);
                        const uint64_t b = make_smem_desc(smem_u32(smK + kb * (FA_BN * 128)) + ki * 32, 16, 1024, 2);
                        umma_tf32(tmem_S, a, b, idesc_s, (kb > 0 || ki > 0) ? 1u : 0u);
                    }
                umma_commit(barS);
            };
            mbar_wait(barQ, 0);
            mbar_wait(barK, 0);
            tc_fence_after();
            mma_s();
            for (int j = 0; j < nblk; ++j) {
                const uint32_t ph = j & 1;
                mbar_wait(barS, ph);                      // K_j consumed: fetch the next block under the softmax
                if (j + 1 < nblk) load_k(j + 1);
                mbar_wait(barP, ph);                      // P_j written, S_j read by every softmax thread
                mbar_wait(barV, ph);
                tc_fence_after();
#pragma unroll
                for (int kb = 0; kb < FA_BN / 32; ++kb)
#pragma unroll
                    for (int ki = 0; ki < 4; ++ki) {
                        const uint64_t a = make_smem_desc(smem_u32(smP + kb * (FA_BM * 128)) + ki * 32, 16, 1024, 2);
                        const uint64_t b = make_smem_desc(smem_u32(smV + kb * (D / 32) * 4096) + ki * 1024, 4096, 512, 1);
                        umma_tf32(tmem_PV, a, b, idesc_o, (kb > 0 || ki > 0) ? 1u : 0u);
                    }
                umma_commit(barO);
                if (j + 1 < nblk) {                       // S_{j+1} right behind PV_j: ready when the softmax warps return
                    mbar_wait(barK, ph ^ 1);
                    tc_fence_after();
                    mma_s();
                }
                mbar_wait(barO, ph);                      // V_j and P_j consumed
                if (j + 1 < nblk) load_v(j + 1);
            }
        }
    } else {
        // ---------------------------------------------------------------- online softmax: thread = query row
        const int r = warp * 32 + lane;                   // row inside the tile = TMEM lane
        const uint32_t lane_off = (uint32_t)(warp * 32) << 16;
        const int b = bh / heads;
        const float* mrow = mask ? mask + (size_t)b * Sk : nullptr;
        // scores are kept in the log2 domain: t = s * scale * log2(e) (+ mask * log2(e)), p = 2^(t - m)
        constexpr float kLog2e = 1.4426950408889634f;
        const float sl2 = scale * kLog2e;
        float m = -INFINITY, l = 0.f;
        float o[D];
#pragma unroll
        for (int c = 0; c < D; ++c) o[c] = 0.f;
        for (int j = 0; j < nblk; ++j) {
            const uint32_t ph = j & 1;
            const int kv0 = j * FA_BN;
            // additive term per key (mask, -inf beyond the last key), staged once per block instead of two compares
            // and a predicated load per element; blocks without one (the common case) skip it altogether
            const bool biased = (mrow != nullptr) || (kv0 + FA_BN > Sk) || !(sl2 > 0.f);
            const float* bj = bias_s + (j & 1) * FA_BN;
            if (biased) {
                if (r < FA_BN) {
                    const int kv = kv0 + r;
                    bias_s[(j & 1) * FA_BN + r] = kv < Sk ? (mrow != nullptr ? __ldg(mrow + kv) * kLog2e : 0.f) : -INFINITY;
                }
                asm volatile("bar.sync 1, 128;" ::: "memory");      // the four softmax warps only
            }
            mbar_wait_backoff(barS, ph);
            tc_fence_after();
            // pass 1 over the TMEM row: running maximum (32 columns in registers at a time)
            float mx = m;
            if (!biased) {
                float raw = -INFINITY;
#pragma unroll 1
                for (int c0 = 0; c0 < FA_BN; c0 += 32) {
                    float s[32];
                    tmem_ld32(tmem_S + lane_off + (uint32_t)c0, s);
#pragma unroll
                    for (int c = 0; c < 32; ++c) raw = fmaxf(raw, s[c]);
                }
                mx = fmaxf(mx, raw * sl2);                        // sl2 > 0 here
            } else {
#pragma unroll 1
                for (int c0 = 0; c0 < FA_BN; c0 += 32) {
                    float s[32];
                    tmem_ld32(tmem_S + lane_off + (uint32_t)c0, s);
#pragma unroll
                    for (int c = 0; c < 32; c += 4) {
                        const float4 b4 = *reinterpret_cast<const float4*>(bj + c0 + c);
                        mx = fmaxf(mx, fmaxf(fmaxf(fmaf(s[c], sl2, b4.x), fmaf(s[c + 1], sl2, b4.y)),
                                             fmaxf(fmaf(s[c + 2], sl2, b4.z), fmaf(s[c + 3], sl2, b4.w))));
                    }
                }
            }
            const float m_use = (mx == -INFINITY) ? 0.f : mx;     // fully masked so far: exp2(-inf - 0) = 0
            const float corr = fast_exp2(m - m_use);              // m = -inf -> 0
            const float neg_m = -m_use;
            // pass 2: p = 2^(t - m), row sum, P_j -> shared memory in the K-major swizzle-128B layout:
            // tile kb = 32 keys, row r at r * 128 B, 16-byte chunk index ^ (r & 7)
            float sum = 0.f;
#pragma unroll 1
            for (int c0 = 0; c0 < FA_BN; c0 += 32) {
                float s[32];
                tmem_ld32(tmem_S + lane_off + (uint32_t)c0, s);
                if (!biased) {
#pragma unroll
                    for (int c = 0; c < 32; ++c) s[c] = fast_exp2(fmaf(s[c], sl2, neg_m));
                } else {
#pragma unroll
                    for (int c = 0; c < 32; c += 4) {
                        const float4 b4 = *reinterpret_cast<const float4*>(bj + c0 + c);
                        s[c] = fast_exp2(fmaf(s[c], sl2, b4.x) + neg_m);
                        s[c + 1] = fast_exp2(fmaf(s[c + 1], sl2, b4.y) + neg_m);
                        s[c + 2] = fast_exp2(fmaf(s[c + 2], sl2, b4.z) + neg_m);
                        s[c + 3] = fast_exp2(fmaf(s[c + 3], sl2, b4.w) + neg_m);
                    }
                }
                uint8_t* trow = smP + (c0 >> 5) * (FA_BM * 128) + r * 128;
#pragma unroll
                for (int ch = 0; ch < 8; ++ch) {
                    sum += (s[4 * ch] + s[4 * ch + 1]) + (s[4 * ch + 2] + s[4 * ch + 3]);
                    *reinterpret_cast<float4*>(trow + ((ch ^ (r & 7)) << 4)) = make_float4(s[4 * ch], s[4 * ch + 1], s[4 * ch + 2], s[4 * ch + 3]);
                }
            }
            l = l * corr + sum;
            m = mx;
            fence_proxy_async();                          // generic-proxy stores -> visible to the tensor core's reads
            tc_fence_before();                            // and our TMEM loads of S_j are done before S_{j+1} overwrites it
            mbar_arrive(barP);
            if (__any_sync(0xffffffffu, corr != 1.f)) {       // the running maximum rarely moves after the first blocks
#pragma unroll
                for (int c = 0; c < D; ++c) o[c] *= corr;
            }
            mbar_wait_backoff(barO, ph);
            tc_fence_after();
#pragma unroll
            for (int c0 = 0; c0 < D; c0 += 32) {
                float pv[32];
                tmem_ld32(tmem_PV + lane_off + (uint32_t)c0, pv);
#pragma unroll
                for (int c = 0; c < 32; ++c) o[c0 + c] += pv[c];
            }
            tc_fence_before();
        }
        const int row = q0 + r;
        if (row < S) {
            const float inv = l > 0.f ? 1.f / l : 0.f;
            float4* dst = reinterpret_cast<float4*>(O + ((size_t)bh * S + row) * D);
#pragma unroll
            for (int c = 0; c < D; c += 4) dst[c >> 2] = make_float4(o[c] * inv, o[c + 1] * inv, o[c + 2] * inv, o[c + 3] * inv);
            if (lse != nullptr) lse[(size_t)bh * S + row] = (l > 0.f) ? (m + log2f(l)) * 0.6931471805599453f : -INFINITY;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 4) tmem_dealloc(tmem_base, FaSmem<D, FA_BN>::kTmemCols);
}

}  // namespace

template <int D, int FA_BN>
static int launch_flash_fwd(const float* q, const float* k, const float* v, const float* mask, float* out, float* lse,
                            int B, int H, int S, int Sk, float scale, cudaStream_t st) {
    const long long BH = (long long)B * H;
    CUtensorMap tmQ, tmK, tmV;
    if (make_tmap(&tmQ, q, D, S, BH, (long long)S * D, FA_BM, false)) return -1;
    if (make_tmap(&tmK, k, D, Sk, BH, (long long)Sk * D, FA_BN, false)) return -1;
    if (make_tmap(&tmV, v, D, Sk, BH, (long long)Sk * D, 32, true)) return -1;
    static bool attr_set = false;
    if (!attr_set) {
        HETU_CUDA(cudaFuncSetAttribute(flash_attn_fwd_kernel<D, FA_BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, FaSmem<D, FA_BN>::kTotal));
        attr_set = true;
    }
    dim3 grid((S + FA_BM - 1) / FA_BM, (unsigned)BH);
    flash_attn_fwd_kernel<D, FA_BN><<<grid, FA_THREADS, FaSmem<D, FA_BN>::kTotal, st>>>(tmQ, tmK, tmV, mask, out, lse, S, Sk, H, scale);
    return 0;
}

// q [B, H, S, D], k / v [B, H, Sk, D] contiguous fp32; mask: additive key mask [B, Sk] or null; lse [B, H, S] or null.
// D must be 32 or 64 (one thread keeps a whole output row in registers).
HETU_API int hetu_flash_attn_fwd(const float* q, const float* k, const float* v, const float* mask, float* out, float* lse,
                                 int B, int H, int S, int Sk, int D, float scale, cudaStream_t st) {
    if (B <= 0 || H <= 0 || S <= 0 || Sk <= 0) return 0;
    if ((reinterpret_cast<uintptr_t>(q) | reinterpret_cast<uintptr_t>(k) | reinterpret_cast<uintptr_t>(v) |
         reinterpret_cast<uintptr_t>(out)) & 15) {
        hetu_set_error("flash_attn_fwd: operands must be 16-byte aligned");
        return -1;
    }
    int rc;
    // 64 keys per block: 96 KB of shared memory, so two CTAs share an SM and one's softmax runs under the other's MMAs
    // (128 keys per block, one CTA per SM, measured 35% slower)
    if (D == 64) rc = launch_flash_fwd<64, 64>(q, k, v, mask, out, lse, B, H, S, Sk, scale, st);
    else if (D == 32) rc = launch_flash_fwd<32, 64>(q, k, v, mask, out, lse, B, H, S, Sk, scale, st);
    else { hetu_set_error("flash_attn_fwd: head dimension %d not supported (32 or 64)", D); return -1; }
    if (rc) return rc;
    HETU_CHECK_LAUNCH("hetu_flash_attn_fwd");
    return 0;
}
