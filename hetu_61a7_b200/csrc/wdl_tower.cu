// Wide&Deep dense tower: forward, loss and backward of one training step in ONE kernel.
//
// reference graph: examples/ctr/models/wdl_criteo.py:8-42 - three matmul (+relu) layers on the 13 dense
// features, concat with the 26 x 128 embedding block, a [3584, 1] head, sigmoid, binary cross entropy, mean;
// the reference runs it as ~45 separate kernels per step (python/hetu/gpu_ops/executor.py:1000-1056).
//
// Design (B200): the batch of 128 samples is exactly one tcgen05 M tile.  One thread-block cluster of 8 CTAs owns
// the step; CTA c owns output columns [32c, 32c+32) of every 256-wide layer.
//   * every 128 x 256 x 256 product is one UMMA chain (M=128, N=32 per CTA, kind::tf32, fp32 accumulator in TMEM),
//     operands brought in by TMA: activations K-major (SWIZZLE_128B), weights in place either MN-major
//     (forward, W[k][n]) or K-major (backward data, W[j][n] rows), activations MN-major for the weight gradients;
//   * layers hand their activations to each other through L2 (plain row-major buffers) with one cluster barrier per
//     layer - no grid-wide synchronisation, no host round trip, no re-launch;
//   * the head never materialises its input gradient: dL/dy4 is the outer product dz x W4, so the y3 branch needs
//     only two matrix-vector products (u = h2^T dz, v = W3 w4y) and the embedding branch is applied straight to
//     the table by wdl_sparse_sgd_rank1_kernel below;
//   * the per-sample logit is reduced across the cluster through distributed shared memory.
// Everything the kernel reads lives in this GPU's memory.
#include "tcgen05.cuh"

namespace {
constexpr int TB = 128;            // batch rows (one UMMA M tile)
constexpr int TH = 256;            // hidden width
constexpr int TCL = 8;             // CTAs per cluster
constexpr int TSL = TH / TCL;      // 32 output columns per CTA
constexpr int TTHREADS = 384;      // 12 warps: 0 = TMA + MMA issue, 1 = TMEM owner, 2..5 = epilogue, 6..11 = vector work
constexpr int TWARPS = TTHREADS / 32;
constexpr int kAArea = TB * TH * 4;          // 128 KB
constexpr int kBArea = TSL * TH * 4;         // 32 KB
constexpr int kKB = TH / 32;                 // 8 k-blocks of 32 fp32 (128 B)
constexpr int kMaxSlice4 = 160;              // float4 columns of the head input per CTA: pitch <= 5120

struct TowerParams {
    const float* x;        // [128, nd]
    float* y4;             // [128, pitch]: columns [0, emb_cols) filled by the gather, [emb_cols, pitch) written here
    const float* labels;   // [128]
    const float* W1;       // [nd, 256]
    const float* W2;       // [256, 256]
    const float* W3;       // [256, 256]
    const float* W4;       // [pitch]
    float* h1; float* h2; float* dh2;          // [128, 256] work buffers
    float* loss; float* pred; float* dz;       // [1], [128], [128]
    float* dW1; float* dW2; float* dW3; float* dW4;
    const float* zpf;      // [128, fields]: per-(sample, field) partial logits written by the gather kernel
    int nd, emb_cols, pitch, width;
    // data-parallel all-reduce + SGD inside the kernel (apply_sgd != 0).  Gradient slices are PUSHED, as they are
    // produced, into block [slot][rank] of every rank's arena (peer st.global over NVLink; world == 1: the local
    // buffer); one release flag per (source, destination) closes the step's pushes; after the flags of all sources
    // have arrived every rank sums the `world` blocks in rank order and applies the identical update.  Two slots
    // alternate by step parity, so no barrier is needed: a rank overwrites a slot only after it has seen every
    // peer's flag of the step in between.
    int apply_sgd, world, rank, total, o2, o3, o4;
    float* arena[8];       // [2][world][total] floats on every rank
    float* arena_mc;       // NVLS multicast mapping of the arenas (or null): one multimem.st reaches every rank
    unsigned* flags_mc;    // the same for the flag rows
    unsigned* flags[8];    // flags[dst][src] = number of complete steps pushed by src
    unsigned* epoch;       // this rank's completed-step counter (device memory: graph replays keep counting)
    const float* lr;
    long long* opt_step;
    // programmatic dependent launch: the kernel may start while the gather that feeds it is still running (pdl != 0:
    // griddepcontrol.wait before the first use of zpf / y4's embedding block) and releases the sparse SGD kernel as
    // soon as dz is out (dz_flag counts published dz vectors)
    int pdl;
    unsigned* dz_flag;
    int stamp_cta;
    unsigned long long* stamps;   // optional phase time stamps (ns), written by CTA 0 / thread 0
};

struct TowerSmem {
    uint64_t full[4][kKB];       // one set of TMA barriers per tensor-core phase (each used once, parity 0)
    uint64_t done[4];
    uint32_t tmem_ptr;
    float zpart[TB];
    float dz[TB];
    float w4y[TH];
    float u[TSL], v3[TSL];
    float red[TWARPS][TSL];
    float lossw[4];
    alignas(16) float x[TB * 16];   // rows padded to 16 floats
    float w1[16 * TSL];          // this CTA's 32 columns of W1
    float h2s[TB][TSL + 1];      // this CTA's 32 columns of h2 (kept for u = h2^T dz and the relu mask of dh2)
    float dh1[TB][TSL + 1];
    float4 colsum[2][kMaxSlice4];
};

__device__ __forceinline__ void stamp(const TowerParams& p, int i) {
    if (p.stamps != nullptr && threadIdx.x == 0 && (int)blockIdx.x == p.stamp_cta) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        p.stamps[i] = t;
    }
}
__device__ __forceinline__ void cluster_sync_all() {
    __syncwarp();
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ float ld_dsmem(const float* local, uint32_t rank) {
    uint32_t a = smem_u32(local), ra;
    float v;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(a), "r"(rank));
    asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(ra) : "memory");
    return v;
}
// Where a gradient slice goes: (a) apply_sgd == 0: the caller's gradient buffer; (b) one GPU: straight into the weights
// (w -= lr * g), every slice is applied by the CTA that produced it once no CTA reads the old weights any more;
// (c) data parallel: block [slot][rank] of every rank's arena (peer stores), summed and applied in the kernel's tail.
// `old` = the weights at `prm`, loaded by the caller ahead of time (only read when the update is applied in place): a
// load-modify-store per slice would serialise one L2 round trip per store.
__device__ __forceinline__ bool inplace(const TowerParams& p) { return p.apply_sgd && p.world == 1; }
__device__ __forceinline__ void grad_store4(const TowerParams& p, unsigned slot, float* own, const float* prm, int off,
                                            float4 v, float lr, float4 old) {
    if (!p.apply_sgd) { *reinterpret_cast<float4*>(own) = v; return; }
    if (p.world == 1) {
        old.x -= lr * v.x; old.y -= lr * v.y; old.z -= lr * v.z; old.w -= lr * v.w;
        *reinterpret_cast<float4*>(const_cast<float*>(prm)) = old;
        return;
    }
    const size_t at = ((size_t)slot * p.world + p.rank) * p.total + off;
    if (p.arena_mc != nullptr) {       // the switch replicates the store into every rank's arena
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p.arena_mc + at), "f"(v.x), "f"(v.y),
                     "f"(v.z), "f"(v.w) : "memory");
        return;
    }
    for (int r = 0; r < p.world; ++r) *reinterpret_cast<float4*>(p.arena[r] + at) = v;
}
__device__ __forceinline__ void grad_store1(const TowerParams& p, unsigned slot, float* own, const float* prm, int off,
                                            float v, float lr, float old) {
    if (!p.apply_sgd) { *own = v; return; }
    if (p.world == 1) { *const_cast<float*>(prm) = old - lr * v; return; }
    const size_t at = ((size_t)slot * p.world + p.rank) * p.total + off;
    if (p.arena_mc != nullptr) {
        asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(p.arena_mc + at), "f"(v) : "memory");
        return;
    }
    for (int r = 0; r < p.world; ++r) p.arena[r][at] = v;
}
// TMA tile load delivered to the same shared-memory offset of every CTA in `mask`; each destination's mbarrier (same
// offset) receives the byte count.  Only used on tensor maps over this GPU's own work buffers (h1, h2, dh2): TMA
// multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another GPU's memory
// hung the GPU.
__device__ __forceinline__ void tma_load_3d_mc(void* dst, const CUtensorMap* map, int c0, int c1, int c2, uint64_t* bar,
                                               uint16_t mask) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster "
        "[%0], [%1, {%3, %4, %5}], [%2], %6;"
        ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "h"(mask)
        : "memory");
}

// idesc: D = f32, A = B = tf32, M = 128, N = 32
__device__ __forceinline__ uint32_t tower_idesc(int a_mn, int b_mn) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) |
           ((uint32_t)(TSL >> 3) << 17) | ((uint32_t)(TB >> 4) << 24);
}

// C[128, 32c..32c+32) = A[128, 256] * Bslice, A K-major through tmA, B either MN-major ([k][n] memory, forward) or
// K-major ([j][k] memory, backward data).  Called by ONE thread of every CTA of the cluster, all CTAs past the same
// cluster barrier: CTA c fetches k-block c of A once and multicasts it to the whole cluster (each CTA needs all of A
// but reads only 1/8 of it from L2); B slices are private.
__device__ void tower_gemm_ka(const CUtensorMap* tmA, const CUtensorMap* tmB, bool b_mn, int c, uint8_t* sA, uint8_t* sB,
                              uint64_t* full, uint64_t* done, uint32_t tmem_d) {
    asm volatile("fence.proxy.async.global;" ::: "memory");        // activations were written with ordinary stores
    for (int kb = 0; kb < kKB; ++kb) mbar_expect_tx(&full[kb], 16384 + 4096);
    tma_load_3d_mc(sA + c * 16384, tmA, c * 32, 0, 0, &full[c], (uint16_t)0xFF);           // box {32 k, 128 rows}
    for (int kb = 0; kb < kKB; ++kb) {
        if (b_mn) tma_load_3d(sB + kb * 4096, tmB, c * TSL, kb * 32, 0, &full[kb]);      // box {32 n, 32 k}
        else tma_load_3d(sB + kb * 4096, tmB, kb * 32, c * TSL, 0, &full[kb]);           // box {32 k, 32 rows}
    }
    const uint32_t idesc = tower_idesc(0, b_mn ? 1 : 0);
    for (int kb = 0; kb < kKB; ++kb) {
        mbar_wait(&full[kb], 0);
        tc_fence_after();
        const uint32_t a_base = smem_u32(sA + kb * 16384), b_base = smem_u32(sB + kb * 4096);
#pragma unroll
        for (int ki = 0; ki < 4; ++ki) {
            uint64_t a_desc = make_smem_desc(a_base + ki * 32, 16, 1024, 2);
            uint64_t b_desc = b_mn ? make_smem_desc(b_base + ki * 1024, 4096, 512, 1) : make_smem_desc(b_base + ki * 32, 16, 1024, 2);
            umma_tf32(tmem_d, a_desc, b_desc, idesc, (kb > 0 || ki > 0) ? 1u : 0u);
        }
    }
    umma_commit(done);
}

__global__ void __cluster_dims__(TCL, 1, 1) __launch_bounds__(TTHREADS, 1)
wdl_tower_kernel(const __grid_constant__ CUtensorMap tm_h1k, const __grid_constant__ CUtensorMap tm_h2k,
                 const __grid_constant__ CUtensorMap tm_dh2k, const __grid_constant__ CUtensorMap tm_W2mn,
                 const __grid_constant__ CUtensorMap tm_W3mn, const __grid_constant__ CUtensorMap tm_W2k,
                 const __grid_constant__ CUtensorMap tm_h1mn, const __grid_constant__ CUtensorMap tm_dh2mn, TowerParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;
    uint8_t* sB = smem + kAArea;
    TowerSmem& S = *reinterpret_cast<TowerSmem*>(smem + kAArea + kBArea);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int c = (int)cluster_ctarank();
    const int n0 = c * TSL;
    const int hcs4 = p.pitch / (4 * TCL);                // float4 columns of the head gradient handled by this CTA
    const int fields = p.emb_cols / p.width;
    const unsigned epoch0 = p.apply_sgd ? *reinterpret_cast<volatile unsigned*>(p.epoch) : 0u;
    const unsigned slot = epoch0 & 1u;
    const float lr_now = p.apply_sgd ? *p.lr : 0.f;

    stamp(p, 0);
    if (tid == 0) {
        prefetch_tmap(&tm_h1k); prefetch_tmap(&tm_W2mn);
        for (int g = 0; g < 4; ++g) {
            for (int i = 0; i < kKB; ++i) mbar_init(&S.full[g][i], 1);
            mbar_init(&S.done[g], 1);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(&S.tmem_ptr, 128);
    // ---- P0: stage x, this CTA's W1 columns and the y3 part of W4; logits of the embedding block (CTA 0) ----------
    for (int i = tid; i < TB * 16; i += TTHREADS) S.x[i] = (i & 15) < p.nd ? p.x[(i >> 4) * p.nd + (i & 15)] : 0.f;   // rows padded to 16
    for (int i = tid; i < p.nd * TSL; i += TTHREADS) S.w1[i] = __ldg(p.W1 + (i >> 5) * TH + n0 + (i & 31));
    for (int i = tid; i < TH; i += TTHREADS) S.w4y[i] = __ldg(p.W4 + p.emb_cols + i);
    if (tid < TB) S.zpart[tid] = 0.f;
    const unsigned dz_seq = (p.dz_flag != nullptr) ? *reinterpret_cast<volatile unsigned*>(p.dz_flag) : 0u;
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = S.tmem_ptr;
    stamp(p, 1);
    // ---- P1: h1 = relu(x W1), K = nd (13): too thin for a tensor-core tile --------------------------------------
    {
        float w1r[16];                                   // this lane's column of W1 (zero beyond nd)
#pragma unroll
        for (int k = 0; k < 16; ++k) w1r[k] = k < p.nd ? S.w1[k * TSL + lane] : 0.f;
        for (int b = warp; b < TB; b += TWARPS) {
            const float4* xr = reinterpret_cast<const float4*>(&S.x[b * 16]);
            float acc = 0.f;
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4) {
                const float4 xv = xr[k4];
                acc += xv.x * w1r[4 * k4] + xv.y * w1r[4 * k4 + 1] + xv.z * w1r[4 * k4 + 2] + xv.w * w1r[4 * k4 + 3];
            }
            p.h1[b * TH + n0 + lane] = fmaxf(acc, 0.f);
        }
    }
    __syncthreads();
    cluster_sync_all();                 // arrive.release / wait.acquire: this thread's stores are visible cluster-wide                                                             // h1 complete
    stamp(p, 2);
    // ---- P2: h2 = relu(h1 W2) ------------------------------------------------------------------------------------
    if (tid == 0) tower_gemm_ka(&tm_h1k, &tm_W2mn, true, c, sA, sB, S.full[0], &S.done[0], tmem_base);
    if (warp >= 2 && warp < 6) {
        const int q = warp & 3, row = q * 32 + lane;
        mbar_wait(&S.done[0], 0);
        tc_fence_after();
        float v[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16), v);
        float4* dst = reinterpret_cast<float4*>(p.h2 + row * TH + n0);
#pragma unroll
        for (int j = 0; j < 32; ++j) { v[j] = fmaxf(v[j], 0.f); S.h2s[row][j] = v[j]; }
#pragma unroll
        for (int j = 0; j < 8; ++j) dst[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        tc_fence_before();
    }
    __syncthreads();
    cluster_sync_all();                 // arrive.release / wait.acquire: this thread's stores are visible cluster-wide                                                             // h2 complete
    stamp(p, 3);
    // ---- P3: y3 = h2 W3 -> y4[:, emb_cols + n], logit partial of the y3 block -------------------------------------
    if (tid == 0) { tc_fence_after(); tower_gemm_ka(&tm_h2k, &tm_W3mn, true, c, sA, sB, S.full[1], &S.done[1], tmem_base); }
    if (warp >= 2 && warp < 6) {
        const int q = warp & 3, row = q * 32 + lane;
        mbar_wait(&S.done[1], 0);
        tc_fence_after();
        float v[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16), v);
        float4* dst = reinterpret_cast<float4*>(p.y4 + (size_t)row * p.pitch + p.emb_cols + n0);
        float acc = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) dst[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
#pragma unroll
        for (int j = 0; j < 32; ++j) acc += v[j] * S.w4y[n0 + j];
        S.zpart[row] += acc;
        tc_fence_before();
    } else if (warp >= 6) {
        // v3 = W3 w4y for this CTA's 32 rows of W3: independent of the activations, done while the tensor core works
        for (int k = warp - 6; k < TSL; k += 6) {
            const float* wr = p.W3 + (size_t)(n0 + k) * TH;
            float a2 = 0.f;
#pragma unroll
            for (int j = 0; j < TH / 32; ++j) a2 += __ldg(wr + lane + 32 * j) * S.w4y[lane + 32 * j];
            a2 = warp_sum(a2);
            if (lane == 0) S.v3[k] = a2;
        }
    }
    __syncthreads();
    cluster_sync_all();                 // arrive.release / wait.acquire: this thread's stores are visible cluster-wide                                                             // every CTA's zpart and y3 columns complete
    stamp(p, 4);
    // ---- head: logits through distributed shared memory, sigmoid, BCE, dz -----------------------------------------
    if (p.pdl) asm volatile("griddepcontrol.wait;" ::: "memory");      // the gather's rows and partial logits are complete
    if (tid < TB) {
        float z = 0.f;
        const float* zp = p.zpf + (size_t)tid * fields;                 // per-(sample, field) dots left by the gather kernel
        for (int f = 0; f < fields; ++f) z += __ldcg(zp + f);
#pragma unroll
        for (int r = 0; r < TCL; ++r) z += ld_dsmem(&S.zpart[tid], (uint32_t)r);
        const float pr = 1.0f / (1.0f + expf(-z));
        const float y = p.labels[tid];
        const float g = (1.0f / TB) * (-y / pr + (1.f - y) / (1.f - pr));          // d mean(bce) / d pred
        const float dzv = g * pr * (1.f - pr);                                      // through the sigmoid
        S.dz[tid] = dzv;
        if (c == 0) {
            p.pred[tid] = pr;
            p.dz[tid] = dzv;
            float l = -y * logf(pr) - (1.f - y) * logf(1.f - pr);
            l = warp_sum(l);
            if (lane == 0) S.lossw[warp] = l;
        }
    }
    if (c == 0 && p.dz_flag != nullptr) __threadfence();
    __syncthreads();
    if (c == 0 && tid == 0) {
        p.loss[0] = (S.lossw[0] + S.lossw[1] + S.lossw[2] + S.lossw[3]) * (1.0f / TB);
        if (p.dz_flag != nullptr)                           // dz is out: the sparse SGD kernel may go ahead
            asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p.dz_flag), "r"(dz_seq + 1u) : "memory");
    }
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    stamp(p, 5);
    // ---- P5: u = h2^T dz (slice, from shared memory), dW3 = u x w4y, dh2 = relu'(h2) . (dz x v3) -------------------
    {
        float acc = 0.f;
        for (int b = warp; b < TB; b += TWARPS) acc += S.h2s[b][lane] * S.dz[b];
        S.red[warp][lane] = acc;
    }
    for (int idx = tid; idx < TB * TSL; idx += TTHREADS) {
        const int k = idx & (TSL - 1), b = idx >> 5;
        p.dh2[b * TH + n0 + k] = S.h2s[b][k] > 0.f ? S.dz[b] * S.v3[k] : 0.f;
    }
    __syncthreads();
    cluster_sync_all();                 // arrive.release / wait.acquire: this thread's stores are visible cluster-wide                                                             // dh2 complete
    stamp(p, 6);
    // ---- P6a: dh1 = (dh2 W2^T) . relu'(h1)   |  vector warps: dW4 = y4^T dz, dW3 --------------------------------
    if (tid == 0) {
        tc_fence_after();
        tower_gemm_ka(&tm_dh2k, &tm_W2k, false, c, sA, sB, S.full[2], &S.done[2], tmem_base);
        mbar_wait(&S.done[2], 0);                      // operands of P6a consumed: the A / B areas can be refilled
        // ---- P6b issue: dW2[:, slice] = h1^T dh2[:, slice]; M = 256 as two halves, K = batch = 128 -----------------
        asm volatile("fence.proxy.async.global;" ::: "memory");
        for (int h = 0; h < 2; ++h)
            for (int kb = 0; kb < 4; ++kb) {
                const int s = h * 4 + kb;
                mbar_expect_tx(&S.full[3][s], 16384 + (h == 0 ? 4096 : 0));
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    tma_load_3d(sA + s * 16384 + j * 4096, &tm_h1mn, h * 128 + j * 32, kb * 32, 0, &S.full[3][s]);
                if (h == 0) tma_load_3d(sB + kb * 4096, &tm_dh2mn, n0, kb * 32, 0, &S.full[3][s]);
            }
        const uint32_t idesc = tower_idesc(1, 1);
        for (int h = 0; h < 2; ++h)
            for (int kb = 0; kb < 4; ++kb) {
                const int s = h * 4 + kb;
                mbar_wait(&S.full[3][s], 0);
                tc_fence_after();
                const uint32_t a_base = smem_u32(sA + s * 16384), b_base = smem_u32(sB + kb * 4096);
#pragma unroll
                for (int ki = 0; ki < 4; ++ki) {
                    uint64_t a_desc = make_smem_desc(a_base + ki * 1024, 4096, 512, 1);
                    uint64_t b_desc = make_smem_desc(b_base + ki * 1024, 4096, 512, 1);
                    umma_tf32(tmem_base + 32 + h * 32, a_desc, b_desc, idesc, (kb > 0 || ki > 0) ? 1u : 0u);
                }
            }
        umma_commit(&S.done[3]);
    } else if (warp >= 2 && warp < 6) {
        const int q = warp & 3, row = q * 32 + lane;
        const float4* hp = reinterpret_cast<const float4*>(p.h1 + row * TH + n0);
        float4 h[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) h[j] = __ldcg(hp + j);                         // relu mask of h1, loaded under the MMA
        mbar_wait(&S.done[2], 0);
        tc_fence_after();
        float v[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16), v);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            S.dh1[row][4 * j + 0] = h[j].x > 0.f ? v[4 * j + 0] : 0.f;
            S.dh1[row][4 * j + 1] = h[j].y > 0.f ? v[4 * j + 1] : 0.f;
            S.dh1[row][4 * j + 2] = h[j].z > 0.f ? v[4 * j + 2] : 0.f;
            S.dh1[row][4 * j + 3] = h[j].w > 0.f ? v[4 * j + 3] : 0.f;
        }
    } else if (warp >= 6) {
        // dW4 slice: (float4 column, batch half) items, 16 independent loads in flight per thread
        const int t6 = tid - 6 * 32;
        if (warp == 6) {                                 // u = h2^T dz for this CTA's slice (partials from P5)
            float a = 0.f;
#pragma unroll
            for (int w = 0; w < TWARPS; ++w) a += S.red[w][lane];
            S.u[lane] = a;
        }
        for (int item = t6; item < 2 * hcs4; item += 6 * 32) {
            const int c4 = item % hcs4, half = item / hcs4;
            const float4* yp = reinterpret_cast<const float4*>(p.y4 + (size_t)(half * 64) * p.pitch) + (size_t)c * hcs4 + c4;
            const size_t rstride = (size_t)p.pitch >> 2;
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
            for (int b0 = 0; b0 < 64; b0 += 16) {
                float4 t[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) t[j] = __ldcg(yp + (size_t)(b0 + j) * rstride);
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const float d = S.dz[half * 64 + b0 + j];
                    acc.x += t[j].x * d; acc.y += t[j].y * d; acc.z += t[j].z * d; acc.w += t[j].w * d;
                }
            }
            S.colsum[half][c4] = acc;
        }
        asm volatile("bar.sync 1, 192;" ::: "memory");          // the six vector warps only: both batch halves are in
        for (int c4 = t6; c4 < hcs4; c4 += 6 * 32) {
            const float4 a = S.colsum[0][c4], b = S.colsum[1][c4];
            const int off = 4 * (c * hcs4 + c4);
            float4 old = make_float4(0.f, 0.f, 0.f, 0.f);
            if (inplace(p)) old = *reinterpret_cast<const float4*>(p.W4 + off);
            grad_store4(p, slot, p.dW4 + off, p.W4 + off, p.o4 + off, make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w), lr_now, old);
        }
        for (int base = t6; base < TSL * (TH / 4); base += 4 * 6 * 32) {
            float4 old[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int idx = base + u * 6 * 32;
                if (inplace(p) && idx < TSL * (TH / 4))
                    old[u] = *reinterpret_cast<const float4*>(p.W3 + (n0 + (idx >> 6)) * TH + 4 * (idx & (TH / 4 - 1)));
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int idx = base + u * 6 * 32;
                if (idx >= TSL * (TH / 4)) break;
                const int n4 = idx & (TH / 4 - 1), k = idx >> 6;
                const float uk = S.u[k];
                const int off = (n0 + k) * TH + 4 * n4;
                grad_store4(p, slot, p.dW3 + off, p.W3 + off, p.o3 + off,
                            make_float4(uk * S.w4y[4 * n4], uk * S.w4y[4 * n4 + 1], uk * S.w4y[4 * n4 + 2], uk * S.w4y[4 * n4 + 3]),
                            lr_now, old[u]);
            }
        }
    }
    __syncthreads();
    if (p.apply_sgd && p.world == 1) cluster_sync_all();   // every CTA's dh1 product has consumed the old W2
    stamp(p, 7);
    // ---- P6b epilogue: dW2 slice   |  vector warps: dW4 store, dW3 = u x w4y, dW1 = x^T dh1 (slice) ----------------
    if (warp >= 2 && warp < 6) {
        const int q = warp & 3;
        mbar_wait(&S.done[3], 0);
        tc_fence_after();
        if (p.stamps != nullptr && warp == 2 && lane == 0 && c == p.stamp_cta) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            p.stamps[9] = t;
        }
        // TMEM lane = row of dW2, so a thread holds 32 columns of ONE row: rows go through a warp-private tile in the
        // (now idle) A area and leave as 128-byte row segments, four rows per store instruction - over NVLink a
        // 16-byte store per lane would be one packet each
        float* tile = reinterpret_cast<float*>(sA) + (warp - 2) * (32 * 36);
        const int sub_row = lane >> 3, c4 = (lane & 7) * 4;
#pragma unroll 1
        for (int h = 0; h < 2; ++h) {
            float v[32];
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + 32 + h * 32, v);
#pragma unroll
            for (int j = 0; j < 8; ++j)
                *reinterpret_cast<float4*>(&tile[lane * 36 + 4 * j]) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            __syncwarp();
            float4 old[8];
            if (inplace(p)) {
#pragma unroll
                for (int r = 0; r < 8; ++r)
                    old[r] = *reinterpret_cast<const float4*>(p.W2 + (h * 128 + q * 32 + 4 * r + sub_row) * TH + n0 + c4);
            }
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const int off = (h * 128 + q * 32 + 4 * r + sub_row) * TH + n0 + c4;
                const float4 g = *reinterpret_cast<const float4*>(&tile[(4 * r + sub_row) * 36 + c4]);
                grad_store4(p, slot, p.dW2 + off, p.W2 + off, p.o2 + off, g, lr_now, old[r]);
            }
            __syncwarp();
        }
    } else if (warp >= 6) {
        // dW1[k, n0 + lane] = sum_b x[b, k] dh1[b, lane].  Shared-memory bandwidth sets the pace, so every load of
        // dh1 feeds all 16 (padded) k at once: warp w covers rows w, w + 6, ..., the six partial sums meet in the
        // (now idle) h2 slice buffer.
        {
            float acc[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) acc[k] = 0.f;
            for (int b = warp - 6; b < TB; b += 6) {
                const float d = S.dh1[b][lane];
                const float4* xr = reinterpret_cast<const float4*>(&S.x[b * 16]);
#pragma unroll
                for (int k4 = 0; k4 < 4; ++k4) {
                    const float4 xv = xr[k4];
                    acc[4 * k4] += xv.x * d; acc[4 * k4 + 1] += xv.y * d; acc[4 * k4 + 2] += xv.z * d; acc[4 * k4 + 3] += xv.w * d;
                }
            }
            float* part = &S.h2s[0][0];                  // [6][16][32]
#pragma unroll
            for (int k = 0; k < 16; ++k) part[((warp - 6) * 16 + k) * TSL + lane] = acc[k];
            asm volatile("bar.sync 1, 192;" ::: "memory");
            for (int idx = tid - 6 * 32; idx < p.nd * TSL; idx += 6 * 32) {
                const int n = idx & (TSL - 1), k = idx >> 5;
                const float old = inplace(p) ? p.W1[k * TH + n0 + n] : 0.f;
                float g = 0.f;
#pragma unroll
                for (int w = 0; w < 6; ++w) g += part[(w * 16 + k) * TSL + n];
                grad_store1(p, slot, p.dW1 + k * TH + n0 + n, p.W1 + k * TH + n0 + n, k * TH + n0 + n, g, lr_now, old);
            }
        }
        if (p.stamps != nullptr && warp == 6 && lane == 0 && c == p.stamp_cta) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            p.stamps[15] = t;
        }
    }
    tc_fence_before();
    __syncthreads();
    stamp(p, 8);
    if (warp == 1) {
        tmem_dealloc(tmem_base, 128);
        if (p.stamps != nullptr && lane == 0 && c == p.stamp_cta) {       // diagnostics: after the TMEM release
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            p.stamps[14] = t;
        }
    }
    if (!p.apply_sgd) return;
    if (p.world == 1) {                                  // the weights were updated where the gradients were produced
        if (c == 0 && tid == 0 && p.opt_step != nullptr) *p.opt_step += 1;
        return;
    }
    // ---- all-reduce + SGD: every gradient slice of this step has been pushed ---------------------------------------
    __syncthreads();
    if (tid == 0) __threadfence_system();                // this CTA's pushes (ordered before by the barrier) reach the peers
    cluster_sync_all();
    const unsigned target = epoch0 + 1u;
    if (c == 0 && tid == 0) {
        __threadfence_system();
        if (p.flags_mc != nullptr) {                     // one multicast store raises this rank's flag on every rank
            asm volatile("multimem.st.release.sys.global.u32 [%0], %1;" ::"l"(p.flags_mc + p.rank), "r"(target) : "memory");
        } else {
            for (int r = 0; r < p.world; ++r)
                if (r != p.rank)
                    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p.flags[r] + p.rank), "r"(target) : "memory");
        }
        *p.epoch = target;                               // hetu_wdl_dp_apply (next in the stream) reduces this step's slot
        if (p.opt_step != nullptr) *p.opt_step += 1;
    }
    stamp(p, 9);
}
// Second half of the data-parallel step: every rank's gradient blocks are (being) pushed into this rank's arena by the
// tower kernels; once the flags of all sources show this step, sum the blocks in rank order (bit-identical replicas)
// and apply SGD.  A full-width grid: the 8 SMs of the tower cluster would need ~6 us just to read world x 0.55 MB.
struct DpApplyParams {
    const float* arena;        // this rank's [2][world][total]
    const unsigned* flags;     // this rank's flag row [world]
    const unsigned* epoch;     // set by the tower kernel of this step
    const float* lr;
    float* prm[4];
    int off[5];
    int world, rank, total;
};
__global__ void __launch_bounds__(256) wdl_dp_apply_kernel(DpApplyParams p) {
    const unsigned target = *reinterpret_cast<const volatile unsigned*>(p.epoch);
    const unsigned slot = (target - 1u) & 1u;
    if ((int)threadIdx.x < p.world && (int)threadIdx.x != p.rank) {
        const unsigned* f = p.flags + threadIdx.x;
        unsigned v;
        do { asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory"); } while ((int)(v - target) < 0);
    }
    __syncthreads();
    const float lr = *p.lr;
    const float4* blk = reinterpret_cast<const float4*>(p.arena + (size_t)slot * p.world * p.total);
    const size_t rs = (size_t)p.total >> 2;
    const int n4 = p.total >> 2;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += gridDim.x * blockDim.x) {
        int sg = 0;
        while (sg < 3 && 4 * i >= p.off[sg + 1]) ++sg;
        float4* w = reinterpret_cast<float4*>(p.prm[sg] + (4 * i - p.off[sg]));
        float4 wv = *w;
        float4 g = __ldcg(blk + i);
        for (int r = 1; r < p.world; ++r) {
            const float4 t = __ldcg(blk + (size_t)r * rs + i);
            g.x += t.x; g.y += t.y; g.z += t.z; g.w += t.w;
        }
        wv.x -= lr * g.x; wv.y -= lr * g.y; wv.z -= lr * g.z; wv.w -= lr * g.w;
        *w = wv;
    }
}
}  // namespace

HETU_API int hetu_wdl_dp_apply(const float* arena_local, const void* flags_local, const void* epoch, const float* lr,
                               float* W1, float* W2, float* W3, float* W4, int nd, int emb_cols, int world, int rank,
                               cudaStream_t s) {
    DpApplyParams p;
    p.arena = arena_local; p.flags = (const unsigned*)flags_local; p.epoch = (const unsigned*)epoch; p.lr = lr;
    p.prm[0] = W1; p.prm[1] = W2; p.prm[2] = W3; p.prm[3] = W4;
    p.off[0] = 0; p.off[1] = nd * TH; p.off[2] = p.off[1] + TH * TH; p.off[3] = p.off[2] + TH * TH;
    p.off[4] = p.off[3] + emb_cols + TH;
    p.total = p.off[4]; p.world = world; p.rank = rank;
    wdl_dp_apply_kernel<<<kNumSMs, 256, 0, s>>>(p);
    HETU_CHECK_LAUNCH("hetu_wdl_dp_apply");
    return 0;
}

static unsigned long long* g_tower_stamps = nullptr;
static void ensure_stamps() {
    if (g_tower_stamps == nullptr && getenv("HETU_WDL_STAMPS") != nullptr) {
        cudaMalloc(&g_tower_stamps, 16 * sizeof(unsigned long long));
        const unsigned long long init[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, ~0ull, 0, ~0ull, 0, 0, 0};
        cudaMemcpy(g_tower_stamps, init, sizeof(init), cudaMemcpyHostToDevice);
    }
}

// work: 3 * 128 * 256 floats (h1, h2, dh2).  All pointers are this GPU's memory; y4 is the concat buffer whose
// first emb_cols columns hold the gathered embedding rows (hetu_gather_rows_pitched).
HETU_API int hetu_wdl_tower_step(const float* x, float* y4, const float* zpf, const float* labels, const float* W1,
                                 const float* W2, const float* W3, const float* W4, float* work, float* loss, float* pred,
                                 float* dz, float* dW1, float* dW2, float* dW3, float* dW4, int batch, int nd, int hidden,
                                 int emb_cols, int width, void* const* arenas, void* const* flags, int world, int rank,
                                 void* epoch, const float* lr, void* opt_step, int pdl, void* dz_flag, void* arena_mc, void* flags_mc,
                                 cudaStream_t s) {
    if (batch != TB || hidden != TH || nd < 1 || nd > 16 || emb_cols <= 0 || emb_cols % (4 * TCL) != 0 || width <= 0 ||
        emb_cols % width != 0 || (emb_cols + TH) / (4 * TCL) > kMaxSlice4) {
        hetu_set_error("hetu_wdl_tower_step: needs batch 128, hidden 256, 1..16 dense features, emb_cols %% 32 == 0, "
                       "emb_cols + 256 <= %d (got %d, %d, %d, %d, width %d)", kMaxSlice4 * 4 * TCL, batch, hidden, nd,
                       emb_cols, width);
        return -1;
    }
    TowerParams p;
    p.x = x; p.y4 = y4; p.zpf = zpf; p.width = width; p.labels = labels; p.W1 = W1; p.W2 = W2; p.W3 = W3; p.W4 = W4;
    p.h1 = work; p.h2 = work + TB * TH; p.dh2 = work + 2 * TB * TH;
    p.loss = loss; p.pred = pred; p.dz = dz; p.dW1 = dW1; p.dW2 = dW2; p.dW3 = dW3; p.dW4 = dW4;
    p.nd = nd; p.emb_cols = emb_cols; p.pitch = emb_cols + TH;
    p.o2 = nd * TH; p.o3 = p.o2 + TH * TH; p.o4 = p.o3 + TH * TH; p.total = p.o4 + p.pitch;
    p.apply_sgd = arenas != nullptr ? 1 : 0;
    p.world = 1; p.rank = 0; p.epoch = (unsigned*)epoch; p.lr = lr; p.opt_step = (long long*)opt_step;
    for (int i = 0; i < 8; ++i) { p.arena[i] = nullptr; p.flags[i] = nullptr; }
    p.arena_mc = (world > 1) ? (float*)arena_mc : nullptr;
    p.flags_mc = (world > 1) ? (unsigned*)flags_mc : nullptr;
    if (p.apply_sgd) {
        if (world < 1 || world > 8 || rank < 0 || rank >= world || epoch == nullptr || lr == nullptr ||
            (world > 1 && flags == nullptr)) {
            hetu_set_error("hetu_wdl_tower_step: in-kernel SGD needs 1 <= world <= 8, 0 <= rank < world, an epoch counter, "
                           "lr and (world > 1) the flag table (got world %d rank %d)", world, rank);
            return -1;
        }
        p.world = world; p.rank = rank;
        for (int i = 0; i < world; ++i) {
            p.arena[i] = (float*)arenas[i];
            p.flags[i] = world > 1 ? (unsigned*)flags[i] : nullptr;
        }
    }
    p.pdl = pdl ? 1 : 0; p.dz_flag = (unsigned*)dz_flag;
    p.stamps = nullptr; p.stamp_cta = getenv("HETU_WDL_STAMP_CTA") ? atoi(getenv("HETU_WDL_STAMP_CTA")) : 0;
    if (getenv("HETU_WDL_STAMPS") != nullptr) {          // diagnostics: phase boundaries
        ensure_stamps();
        p.stamps = g_tower_stamps;
    }
    CUtensorMap h1k, h2k, dh2k, W2mn, W3mn, W2k, h1mn, dh2mn;
    if (make_tmap(&h1k, p.h1, TH, TB, 1, 0, TB, false) || make_tmap(&h2k, p.h2, TH, TB, 1, 0, TB, false) ||
        make_tmap(&dh2k, p.dh2, TH, TB, 1, 0, TB, false) || make_tmap(&W2mn, W2, TH, TH, 1, 0, 32, true) ||
        make_tmap(&W3mn, W3, TH, TH, 1, 0, 32, true) || make_tmap(&W2k, W2, TH, TH, 1, 0, 32, false) ||
        make_tmap(&h1mn, p.h1, TH, TB, 1, 0, 32, true) || make_tmap(&dh2mn, p.dh2, TH, TB, 1, 0, 32, true))
        return -1;
    const int smem_bytes = 1024 + kAArea + kBArea + (int)sizeof(TowerSmem);
    static bool attr_set = false;
    if (!attr_set) {
        HETU_CUDA(cudaFuncSetAttribute(wdl_tower_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
        attr_set = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(TCL); cfg.blockDim = dim3(TTHREADS); cfg.dynamicSmemBytes = smem_bytes; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    HETU_CUDA(cudaLaunchKernelEx(&cfg, wdl_tower_kernel, h1k, h2k, dh2k, W2mn, W3mn, W2k, h1mn, dh2mn, p));
    HETU_CHECK_LAUNCH("hetu_wdl_tower_step");
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// Embedding side of the fused step.  Row r of a sharded table lives on GPU r % nshard at local row r / nshard
// (parallel/sharded_embedding.py); nshard == 1 is the plain local table.
namespace {
constexpr int kWdlMaxShards = 16;
struct WdlShards { float* base[kWdlMaxShards]; int nshard; };

// lookup + concat: row i = (sample b, field f) goes to out[b * pitch + f * width ...].  With w4 / zpf given the warp
// also leaves the row's share of the head logit, zpf[i] = <row, w4[f * width ...]>, so the tower kernel never re-reads
// the embedding block for its forward pass.
__global__ void __launch_bounds__(256) gather_rows_pitched_kernel(WdlShards tab, const void* __restrict__ idx, int idx_code,
                                                                  float* __restrict__ out, size_t n, int width, int fields,
                                                                  int pitch, size_t vocab, const float* __restrict__ w4,
                                                                  float* __restrict__ zpf, float* __restrict__ w4_snap,
                                                                  const unsigned* __restrict__ dz_flag,
                                                                  unsigned* __restrict__ dz_expect,
                                                                  unsigned long long* stamps) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");     // the tower kernel's first layers do not need us
    if (stamps != nullptr && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        atomicMin(stamps + 10, t);
    }
    if (dz_expect != nullptr && blockIdx.x == 0 && threadIdx.x == 0)    // what dz_flag will read once this step's dz is out
        *dz_expect = *reinterpret_cast<const volatile unsigned*>(dz_flag) + 1u;
    if (w4_snap != nullptr) {       // the head weights as of this step's forward pass (the tower kernel updates W4)
        const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
        if (gid < (size_t)fields * width) w4_snap[gid] = w4[gid];
    }
    size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t i = warp; i < n; i += nwarp) {
        const long long r = load_index(idx, idx_code, i);
        const int f = (int)(i % fields);
        float* dst = out + (i / fields) * (size_t)pitch + f * (size_t)width;
        const bool ok = r >= 0 && (size_t)r < vocab;
        const float* src = ok ? tab.base[(int)(r % tab.nshard)] + (size_t)(r / tab.nshard) * width : nullptr;
        float dot = 0.f;
        if ((width & 3) == 0) {
            for (int j = lane; j < (width >> 2); j += 32) {
                const float4 v = ok ? reinterpret_cast<const float4*>(src)[j] : make_float4(0.f, 0.f, 0.f, 0.f);
                reinterpret_cast<float4*>(dst)[j] = v;
                if (w4 != nullptr) {
                    const float4 w = __ldg(reinterpret_cast<const float4*>(w4 + f * (size_t)width) + j);
                    dot += v.x * w.x + v.y * w.y + v.z * w.z + v.w * w.w;
                }
            }
        } else {
            for (int j = lane; j < width; j += 32) {
                const float v = ok ? src[j] : 0.f;
                dst[j] = v;
                if (w4 != nullptr) dot += v * __ldg(w4 + f * (size_t)width + j);
            }
        }
        if (zpf != nullptr) {
            dot = warp_sum(dot);
            if (lane == 0) zpf[i] = dot;
        }
    }
    if (stamps != nullptr && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        atomicMax(stamps + 11, t);
    }
}

// SGD on the embedding rows with the head's rank-one gradient: d loss / d row(b, f) = dz[b] * W4[f * width ...].
// Duplicate ids accumulate through atomics, as the summed IndexedSlices of the unfused path do.
__global__ void __launch_bounds__(256) wdl_sparse_sgd_rank1_kernel(WdlShards tab, const void* __restrict__ idx, int idx_code,
                                                                   const float* __restrict__ dz, const float* __restrict__ w4,
                                                                   const float* __restrict__ lr_p, float scale, size_t n,
                                                                   int width, int fields, size_t vocab,
                                                                   const unsigned* __restrict__ dz_flag,
                                                                   const unsigned* __restrict__ dz_expect,
                                                                   unsigned long long* stamps) {
    if (stamps != nullptr && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        atomicMin(stamps + 12, t);
    }
    if (dz_flag != nullptr) {
        // launched programmatically behind the tower kernel, which is still running its backward pass: go once it
        // has published this step's dz.  (The tower kernel is resident before this grid can be scheduled.)
        if (threadIdx.x == 0) {
            const unsigned want = *reinterpret_cast<const volatile unsigned*>(dz_expect);
            unsigned v;
            do { asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(dz_flag) : "memory"); } while ((int)(v - want) < 0);
        }
        __syncthreads();
    }
    size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    const float k = -(*lr_p) * scale;
    for (size_t i = warp; i < n; i += nwarp) {
        const long long r = load_index(idx, idx_code, i);
        if (r < 0 || (size_t)r >= vocab) continue;
        float* row = tab.base[(int)(r % tab.nshard)] + (size_t)(r / tab.nshard) * width;
        const float g = k * dz[i / fields];
        const float* w = w4 + (i % fields) * (size_t)width;
        if ((width & 3) == 0) {
            for (int j = lane; j < (width >> 2); j += 32) {
                const float4 wv = __ldg(reinterpret_cast<const float4*>(w) + j);
                atomicAdd(reinterpret_cast<float4*>(row) + j, make_float4(g * wv.x, g * wv.y, g * wv.z, g * wv.w));
            }
        } else {
            for (int j = lane; j < width; j += 32) atomicAdd(row + j, g * w[j]);
        }
    }
    if (stamps != nullptr && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        atomicMax(stamps + 13, t);
    }
    // completion of this grid must imply completion of the tower kernel for whatever is queued behind it
    if (dz_flag != nullptr) asm volatile("griddepcontrol.wait;" ::: "memory");
}

int fill_shards(WdlShards& tab, void* const* bases, int nshard, const char* who) {
    if (nshard < 1 || nshard > kWdlMaxShards) { hetu_set_error("%s: nshard %d outside [1, %d]", who, nshard, kWdlMaxShards); return -1; }
    tab.nshard = nshard;
    for (int i = 0; i < kWdlMaxShards; ++i) tab.base[i] = i < nshard ? (float*)bases[i] : nullptr;
    return 0;
}
}  // namespace

HETU_API int hetu_gather_rows_pitched(void* const* bases, int nshard, const void* idx, int idx_code, float* out, size_t n,
                                      int width, int fields, int pitch, size_t vocab, const float* w4, float* zpf,
                                      float* w4_snap, const void* dz_flag, void* dz_expect, cudaStream_t s) {
    if (n == 0) return 0;
    WdlShards tab;
    ensure_stamps();
    if (fill_shards(tab, bases, nshard, "hetu_gather_rows_pitched")) return -1;
    static bool carve_set = false;
    if (!carve_set) {
        // the tower kernel that follows needs the SMs configured for maximal shared memory; a kernel in between that
        // prefers L1 makes the SMs it ran on re-partition twice per step
        if (getenv("HETU_WDL_NO_CARVEOUT") == nullptr) {
            HETU_CUDA(cudaFuncSetAttribute(gather_rows_pitched_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                           cudaSharedmemCarveoutMaxShared));
            HETU_CUDA(cudaFuncSetAttribute(wdl_sparse_sgd_rank1_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                           cudaSharedmemCarveoutMaxShared));
        }
        carve_set = true;
    }
    if (w4_snap != nullptr && (size_t)fields * width > (size_t)ew_grid(n * 32, 256, 16) * 256) {
        hetu_set_error("hetu_gather_rows_pitched: grid too small to snapshot %d head weights", fields * width);
        return -1;
    }
    gather_rows_pitched_kernel<<<ew_grid(n * 32, 256, 16), 256, 0, s>>>(tab, idx, idx_code, out, n, width, fields, pitch, vocab,
                                                                       w4, zpf, w4_snap, (const unsigned*)dz_flag,
                                                                       (unsigned*)dz_expect,
                                                                       getenv("HETU_WDL_STAMPS") != nullptr ? g_tower_stamps : nullptr);
    HETU_CHECK_LAUNCH("hetu_gather_rows_pitched");
    return 0;
}

HETU_API int hetu_wdl_sparse_sgd_rank1(void* const* bases, int nshard, const void* idx, int idx_code, const float* dz,
                                       const float* w4, const float* lr_p, float scale, size_t n, int width, int fields,
                                       size_t vocab, const void* dz_flag, const void* dz_expect, cudaStream_t s) {
    if (n == 0) return 0;
    WdlShards tab;
    if (fill_shards(tab, bases, nshard, "hetu_wdl_sparse_sgd_rank1")) return -1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(ew_grid(n * 32, 256, 16)); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = 0; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = dz_flag != nullptr ? 1 : 0;
    HETU_CUDA(cudaLaunchKernelEx(&cfg, wdl_sparse_sgd_rank1_kernel, tab, idx, idx_code, dz, w4, lr_p, scale, n, width, fields,
                                 vocab, (const unsigned*)dz_flag, (const unsigned*)dz_expect,
                                 getenv("HETU_WDL_STAMPS") != nullptr ? g_tower_stamps : nullptr));
    HETU_CHECK_LAUNCH("hetu_wdl_sparse_sgd_rank1");
    return 0;
}

// diagnostics: phase time stamps (ns) of the last hetu_wdl_tower_step launched with HETU_WDL_STAMPS set
HETU_API int hetu_wdl_tower_stamps(unsigned long long* out9 /* 15 values; [10..13] = gather start/end, sparse start/end, [14] = last CTA after its TMEM release */) {
    if (g_tower_stamps == nullptr) { hetu_set_error("hetu_wdl_tower_stamps: no stamped launch yet"); return -1; }
    HETU_CUDA(cudaDeviceSynchronize());
    HETU_CUDA(cudaMemcpy(out9, g_tower_stamps, 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    const unsigned long long reset[4] = {~0ull, 0ull, ~0ull, 0ull};
    HETU_CUDA(cudaMemcpy(g_tower_stamps + 10, reset, sizeof(reset), cudaMemcpyHostToDevice));
    return 0;
}
