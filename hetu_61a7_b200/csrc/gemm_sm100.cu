// GEMM for sm_100a: C = act(alpha * op(A) op(B) + beta * C + bias), fp32 storage.
//
// reference: src/ops/MatrixMult.cu:22-26, Linear.cu:35-54, BatchMatrixMult.cu:31-36,
// Addmm.cu, Baddbmm.cu - all cublasSgemm(StridedBatched) in fp32 on CUDA cores.
//
// Path 1 (tensor cores): warp-specialised tcgen05 kernel.
//   warp 0      : TMA producer  (cp.async.bulk.tensor.3d -> 128B-swizzled smem ring, mbarrier tx)
//   warp 1      : TMEM allocator + MMA issuer (one elected thread, tcgen05.mma kind::tf32,
//                 fp32 accumulator in TMEM, tcgen05.commit releases smem stages)
//   warps 2..5  : epilogue (tcgen05.ld 32x32b -> registers -> warp-private smem transpose ->
//                 coalesced global stores with alpha/beta/bias/activation fused)
//   Operands are consumed in place for all four transpose combinations: a row-major [rows, K]
//   operand is "K-major", a row-major [K, cols] operand is "MN-major"; both are expressible
//   in the shared-memory matrix descriptor, so nothing is transposed in memory.
// Path 2 (CUDA cores): register-tiled FFMA kernel for shapes TMA cannot describe (row pitch
//   not a multiple of 16 bytes), tiny problems, and the exact-fp32 precision mode.
#include "tcgen05.cuh"
#include <mutex>
#include <unordered_map>
#include <string>
#include <cstring>

// ACT_RELU_MASK (tcgen05 kernels only): C holds the forward activations h on entry; C = (h > 0) ? alpha * A B : 0 -
// the relu backward folded into the GEMM that produces dH, written over h.
enum { ACT_NONE = 0, ACT_RELU = 1, ACT_GELU = 2, ACT_SIGMOID = 3, ACT_RELU_MASK = 4 };

__device__ __forceinline__ float apply_act(float v, int act) {
    if (act == ACT_RELU) return fmaxf(v, 0.f);
    if (act == ACT_GELU) return v * 0.5f * (1.0f + erff(v * 0.70710678118654752f));
    if (act == ACT_SIGMOID) return 1.f / (1.f + __expf(-v));
    return v;
}

// ================================================================== tcgen05 GEMM kernel
constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 32;        // 32 fp32 = 128 bytes = one swizzle-128B row
constexpr int UMMA_K = 8;          // kind::tf32
constexpr int kGemmThreads = 192;  // 6 warps

template <int BLOCK_N>
struct GemmCfg {
    static constexpr int kStages = BLOCK_N == 128 ? 3 : 4;   // N=128: 3 stages so two CTAs share an SM
    static constexpr int kABytes = BLOCK_M * BLOCK_K * 4;     // 16 KB
    static constexpr int kBBytes = BLOCK_N * BLOCK_K * 4;
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kEpiBytes = 4 * 32 * 36 * 4;         // per-warp 32x36 transpose tiles (aliases stage 0)
    static_assert(kEpiBytes <= kStageBytes, "epilogue staging must fit in one stage");
    static constexpr int kSmem = kStages * kStageBytes + 1024 /*align*/ + 256 /*barriers*/;
    static constexpr int kTmemCols = BLOCK_N < 32 ? 32 : BLOCK_N;
};

struct GemmParams {
    float* C;
    const float* bias;
    int M, N, K;
    long long sC;
    int a_mn_major, b_mn_major;    // 1 = MN-major (A given as [K,M] / B given as [K,N])
    int a_batched, b_batched;
    // tensor-parallel fusions over the symmetric heap (2-SM kernel only):
    //  ag_world > 1: A is row-sharded over ag_world GPUs (all-gather -> GEMM).  Every rank holds a full-size
    //                gathered buffer; the epilogue warps first PUSH this rank's shard to all peers in
    //                128-row blocks (posted NVLink stores + a release flag per block) while the producer /
    //                MMA warps already work on the locally owned tiles; tiles of remote rows wait for
    //                their block's flag.  Each byte crosses NVLink once.
    //  rs_world > 1: row block r of the partial product is stored into slice `rs_rank` of rank r's staging
    //                buffer rs_out[r] ([world][rs_rows][N]) with plain 128-bit NVLink stores from the epilogue;
    //                the owner sums its `world` slices after the closing barrier (GEMM -> reduce-scatter)
    int ag_world, ag_rows, ag_rank, rs_world, rs_rows, rs_rank;
    float* ag_buf[8];              // gathered buffer of every rank (peer-mapped)
    unsigned* ag_flag[8];          // flags[world][ag_rows / 128] of every rank
    float* rs_out[8];
    int dbg;                       // diagnostics: 1 = no TMA after the first ring fill, 2 = no MMA
    int a_mn4, b_mn4;              // MN-major operand described by ONE 4-D box {32 mn, 32 k, 4 groups, 1}
    float alpha, beta;
    int act;
};

template <int BLOCK_N>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tf32_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, GemmParams p) {
    using Cfg = GemmCfg<BLOCK_N>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;                                               // [stages][16 KB]
    uint8_t* smem_b = smem + Cfg::kStages * Cfg::kABytes;                 // [stages][kBBytes]
    float* smem_epi = reinterpret_cast<float*>(smem);   // reused after the last MMA has retired
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
    uint64_t* empty_bar = full_bar + Cfg::kStages;
    uint64_t* tmem_full_bar = empty_bar + Cfg::kStages;
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.y * BLOCK_M, n0 = blockIdx.x * BLOCK_N, bz = blockIdx.z;
    const int num_kb = (p.K + BLOCK_K - 1) / BLOCK_K;

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
        // ------------------------------------------------------------ TMA producer
        if (lane == 0) {
            const int ba = p.a_batched ? bz : 0, bb = p.b_batched ? bz : 0;
            for (int kb = 0; kb < num_kb; ++kb) {
                const int st = kb % Cfg::kStages;
                if (kb >= Cfg::kStages) mbar_wait(&empty_bar[st], ((kb / Cfg::kStages) - 1) & 1);
                mbar_expect_tx(&full_bar[st], Cfg::kStageBytes);
                uint8_t* a_dst = smem_a + st * Cfg::kABytes;
                uint8_t* b_dst = smem_b + st * Cfg::kBBytes;
                const int k0 = kb * BLOCK_K;
                if (!p.a_mn_major) {
                    tma_load_3d(a_dst, &tmA, k0, m0, ba, &full_bar[st]);               // box {32k, 128m}
                } else {
#pragma unroll
                    for (int j = 0; j < BLOCK_M / 32; ++j)                              // box {32m, 32k} x4
                        tma_load_3d(a_dst + j * (BLOCK_K * 128), &tmA, m0 + j * 32, k0, ba, &full_bar[st]);
                }
                if (!p.b_mn_major) {
                    tma_load_3d(b_dst, &tmB, k0, n0, bb, &full_bar[st]);               // box {32k, BLOCK_N n}
                } else {
#pragma unroll
                    for (int j = 0; j < BLOCK_N / 32; ++j)
                        tma_load_3d(b_dst + j * (BLOCK_K * 128), &tmB, n0 + j * 32, k0, bb, &full_bar[st]);
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer
        if (lane == 0) {
            // instruction descriptor: D=f32, A=B=tf32, majors, N>>3 @17, M>>4 @24
            uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)p.a_mn_major << 15) |
                             ((uint32_t)p.b_mn_major << 16) | ((uint32_t)(BLOCK_N >> 3) << 17) |
                             ((uint32_t)(BLOCK_M >> 4) << 24);
            for (int kb = 0; kb < num_kb; ++kb) {
                const int st = kb % Cfg::kStages;
                mbar_wait(&full_bar[st], (kb / Cfg::kStages) & 1);
                tc_fence_after();
                const uint32_t a_base = smem_u32(smem_a + st * Cfg::kABytes);
                const uint32_t b_base = smem_u32(smem_b + st * Cfg::kBBytes);
#pragma unroll
                for (int ki = 0; ki < BLOCK_K / UMMA_K; ++ki) {
                    // K-major : rows of 128 B, 8-row groups 1024 B apart (SBO); K step = 32 B inside the row.
                    // MN-major: [k][128 B of MN] atoms, 4-k-row groups 512 B apart (SBO), next 32 MN
                    //           elements BLOCK_K*128 B further (LBO); K step = 8 rows = 1024 B.
                    uint64_t a_desc = p.a_mn_major ? make_smem_desc(a_base + ki * (UMMA_K * 128), BLOCK_K * 128, 512, 1)
                                                   : make_smem_desc(a_base + ki * (UMMA_K * 4), 16, 1024, 2);
                    uint64_t b_desc = p.b_mn_major ? make_smem_desc(b_base + ki * (UMMA_K * 128), BLOCK_K * 128, 512, 1)
                                                   : make_smem_desc(b_base + ki * (UMMA_K * 4), 16, 1024, 2);
                    umma_tf32(tmem_base, a_desc, b_desc, idesc, (kb > 0 || ki > 0) ? 1u : 0u);
                }
                umma_commit(&empty_bar[st]);       // smem stage reusable once these MMAs retire
            }
            umma_commit(tmem_full_bar);            // accumulator complete
        }
    } else {
        // ------------------------------------------------------------ epilogue (warps 2..5)
        // TMEM lane = output row, so a thread holds one row x 32 columns per tcgen05.ld.  Rows are
        // transposed through a warp-private smem tile (row pitch 36 floats: 16 B aligned and
        // conflict-free for 128-bit accesses), then each pass stores 4 rows x 128 B: lanes 0-7
        // cover one row with float4s.
        const int q = warp & 3;                    // TMEM lane quadrant this warp may access
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
                // C is read for beta / the relu mask: all eight loads of this warp pass are issued before the first store
                const bool need_c = p.beta != 0.f || p.act == ACT_RELU_MASK;
                float4 cin[8];
                if (need_c) {
#pragma unroll
                    for (int r = 0; r < 32; r += 4) {
                        const int row = row_base + r + sub_row;
                        cin[r >> 2] = row < p.M ? *reinterpret_cast<const float4*>(Cb + (size_t)row * p.N + col) : make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                }
#pragma unroll
                for (int r = 0; r < 32; r += 4) {
                    const int row = row_base + r + sub_row;
                    if (row < p.M) {
                        float4 a = *reinterpret_cast<const float4*>(&tile[(r + sub_row) * 36 + c4]);
                        float4 o = make_float4(p.alpha * a.x + bv.x, p.alpha * a.y + bv.y, p.alpha * a.z + bv.z, p.alpha * a.w + bv.w);
                        float4* dst = reinterpret_cast<float4*>(Cb + (size_t)row * p.N + col);
                        if (p.beta != 0.f) {
                            const float4 c = cin[r >> 2];
                            o.x += p.beta * c.x; o.y += p.beta * c.y; o.z += p.beta * c.z; o.w += p.beta * c.w;
                        }
                        if (p.act == ACT_RELU_MASK) {
                            const float4 h = cin[r >> 2];
                            o.x = h.x > 0.f ? o.x : 0.f; o.y = h.y > 0.f ? o.y : 0.f; o.z = h.z > 0.f ? o.z : 0.f; o.w = h.w > 0.f ? o.w : 0.f;
                        } else if (p.act != ACT_NONE) { o.x = apply_act(o.x, p.act); o.y = apply_act(o.y, p.act); o.z = apply_act(o.z, p.act); o.w = apply_act(o.w, p.act); }
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
                        *dst = p.act == ACT_RELU_MASK ? (*dst > 0.f ? o : 0.f) : apply_act(o, p.act);
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

template <int BLOCK_N>
static int launch_tc(const float* A, const float* B, GemmParams& p, int batch, long long sA, long long sB, cudaStream_t s) {
    using Cfg = GemmCfg<BLOCK_N>;
    CUtensorMap tmA, tmB;
    // A: K-major -> memory [M][K]: inner=K rows=M, box rows 128.  MN-major -> memory [K][M]: inner=M rows=K, box rows 32.
    if (!p.a_mn_major) { if (make_tmap(&tmA, A, p.K, p.M, p.a_batched ? batch : 1, sA, BLOCK_M, false)) return -1; }
    else { if (make_tmap(&tmA, A, p.M, p.K, p.a_batched ? batch : 1, sA, BLOCK_K, true)) return -1; }
    if (!p.b_mn_major) { if (make_tmap(&tmB, B, p.K, p.N, p.b_batched ? batch : 1, sB, BLOCK_N, false)) return -1; }
    else { if (make_tmap(&tmB, B, p.N, p.K, p.b_batched ? batch : 1, sB, BLOCK_K, true)) return -1; }
    static bool attr_set = false;
    if (!attr_set) {
        HETU_CUDA(cudaFuncSetAttribute(gemm_tf32_tcgen05_kernel<BLOCK_N>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmem));
        attr_set = true;
    }
    dim3 grid((p.N + BLOCK_N - 1) / BLOCK_N, (p.M + BLOCK_M - 1) / BLOCK_M, batch);
    gemm_tf32_tcgen05_kernel<BLOCK_N><<<grid, kGemmThreads, Cfg::kSmem, s>>>(tmA, tmB, p);
    return 0;
}

// ============================================================ 2-SM persistent tcgen05 GEMM
// Large problems: a CTA pair (cluster 2x1x1, the two SMs of a TPC) owns a 256x256 output tile.
// Each CTA stages its own 128 rows of A and its own 128 columns of B (32 KB per k-block instead
// of 48 KB for the 1-SM 128x256 tile, which is what bounds the tf32 kernel: L2 -> smem bytes);
// the leader CTA issues tcgen05.mma.cta_group::2 (M = 256), which reads both CTAs' shared memory
// and writes rows 0-127 to the leader's TMEM and rows 128-255 to the peer's.  The kernel is
// persistent (static round-robin over tiles, M fastest so concurrent tiles share a B panel in
// L2) and double-buffers the accumulator in TMEM (2 x 256 columns), so the epilogue of tile i
// overlaps the main loop of tile i+1.
//   barriers: full[s]   leader only, tx bytes of BOTH CTAs' TMA loads land here
//             empty[s]  per CTA, armed by tcgen05.commit ... multicast::cluster (mask 0b11)
//             tfull[a]  per CTA, multicast commit after the last k-block of a tile
//             tempty[a] leader only, 8 arrivals = 4 epilogue warps x 2 CTAs (peer arrives remotely)
constexpr int k2Stages = 6;
constexpr int k2ABytes = BLOCK_M * BLOCK_K * 4;            // 16 KB : 128 rows of A
constexpr int k2BBytes = 128 * BLOCK_K * 4;                // 16 KB : this CTA's 128 columns of B
constexpr int k2StageBytes = k2ABytes + k2BBytes;
constexpr int k2EpiBytes = 4 * 32 * 36 * 4;
constexpr int k2Smem = k2Stages * k2StageBytes + k2EpiBytes + 1024 + 256;
constexpr int k2TileN = 256;

__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma_load_3d_2sm(void* dst, const CUtensorMap* map, int c0, int c1, int c2, uint64_t* leader_bar) {
    // peer bit (bit 24 of the shared::cluster address) cleared: the bytes are counted on CTA 0's barrier
    asm volatile(
        "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(leader_bar) & 0xFEFFFFFFu), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
__device__ __forceinline__ void tma_load_4d_2sm(void* dst, const CUtensorMap* map, int c0, int c1, int c2, int c3, uint64_t* leader_bar) {
    asm volatile(
        "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
        ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(leader_bar) & 0xFEFFFFFFu), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
__device__ __forceinline__ void tmem_alloc_2sm(uint32_t* dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_2sm(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_tf32_2sm(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accum) : "memory");
}
__device__ __forceinline__ void umma_commit_2sm(uint64_t* bar) {
    asm volatile(
        "{\n\t"
        ".reg .b16 m;\n\t"
        "mov.b16 m, 3;\n\t"
        "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], m;\n\t"
        "}" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_leader(uint64_t* bar) {
    asm volatile(
        "{\n\t"
        ".reg .b32 ra;\n\t"
        "mapa.shared::cluster.u32 ra, %0, 0;\n\t"
        "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n\t"
        "}" ::"r"(smem_u32(bar)) : "memory");
}

// Tile order: bands of up to 8 N-tiles, N fastest inside a band, so the ~74 tiles in flight share a few
// A panels and one band of B in L2 and A is streamed from HBM once per band (M fastest re-read A
// once per N tile: 2.2 GB of DRAM reads for the 32768x2048x2048 case, measured with ncu).
__device__ __forceinline__ void decode_tile(int tile, int tiles_m, int tiles_n, int& bz, int& tm, int& tn) {
    const int tiles_mn = tiles_m * tiles_n;
    bz = tile / tiles_mn;
    int rem = tile - bz * tiles_mn;
    const int band_sz = 8 * tiles_m;
    const int band = rem / band_sz;
    rem -= band * band_sz;
    const int bw = min(8, tiles_n - band * 8);
    tm = rem / bw;
    tn = band * 8 + (rem - tm * bw);
}

// Work scheduler shared by the three warp roles.  Normal mode: whole tiles, round-robin over clusters.
// Stream-K mode (few tiles, long K - weight gradients): the k-blocks of all tiles form one sequence
// that is cut into equal contiguous ranges, one per cluster; a range that covers only part of a tile is
// added into the zero-initialised C with red.global.add.v4.f32.
struct TileSched {
    int tile, step, num_tiles, num_kb, streamk;
    long long pos, end;
    __device__ __forceinline__ TileSched(int cluster_id, int num_clusters, int num_tiles_, int num_kb_, int kb_per_cluster)
        : tile(cluster_id), step(num_clusters), num_tiles(num_tiles_), num_kb(num_kb_), streamk(kb_per_cluster > 0) {
        pos = (long long)cluster_id * kb_per_cluster;
        const long long total = (long long)num_tiles_ * num_kb_;
        end = pos + kb_per_cluster < total ? pos + kb_per_cluster : total;
    }
    __device__ __forceinline__ bool next(int& t, int& kb0, int& kb1) {
        if (!streamk) {
            if (tile >= num_tiles) return false;
            t = tile; kb0 = 0; kb1 = num_kb; tile += step;
            return true;
        }
        if (pos >= end) return false;
        t = (int)(pos / num_kb);
        kb0 = (int)(pos - (long long)t * num_kb);
        const long long left = end - pos;
        kb1 = (num_kb - kb0) < left ? num_kb : kb0 + (int)left;
        pos += kb1 - kb0;
        return true;
    }
};

// all-gather mode: start with the row tiles this rank owns (no waiting), then walk the peers' rows
__device__ __forceinline__ int ag_rotate(int tm, int tiles_m, const GemmParams& p) {
    const int first = (p.ag_rank * p.ag_rows) >> 8;
    const int t = tm + first;
    return t >= tiles_m ? t - tiles_m : t;
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1)
gemm_tf32_2sm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, GemmParams p,
                     int tiles_m, int tiles_n, int num_tiles, int kb_per_cluster) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + k2Stages * k2ABytes;
    float* smem_epi = reinterpret_cast<float*>(smem + k2Stages * k2StageBytes);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + k2Stages * k2StageBytes + k2EpiBytes);
    uint64_t* empty_bar = full_bar + k2Stages;
    uint64_t* tfull_bar = empty_bar + k2Stages;
    uint64_t* tempty_bar = tfull_bar + 2;
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tempty_bar + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;
    const int num_kb = (p.K + BLOCK_K - 1) / BLOCK_K;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmA);
        prefetch_tmap(&tmB);
        for (int i = 0; i < k2Stages; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], 8); }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc_2sm(tmem_ptr_smem, 512);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                     // peer barriers are initialised before anyone signals them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    if (warp == 0) {
        // ------------------------------------------------------------ TMA producer (both CTAs)
        if (lane == 0) {
            int it = 0, tile, kb0, kb1;
            TileSched sched(cluster_id, num_clusters, num_tiles, num_kb, kb_per_cluster);
            while (sched.next(tile, kb0, kb1)) {
                int bz, tm, tn;
                decode_tile(tile, tiles_m, tiles_n, bz, tm, tn);
                if (p.ag_world > 1) tm = ag_rotate(tm, tiles_m, p);
                const int m0 = tm * 256 + (int)rank * 128;
                const int nb0 = tn * k2TileN + (int)rank * 128;
                const int ba = p.a_batched ? bz : 0, bb = p.b_batched ? bz : 0;
                const CUtensorMap* mapA = &tmA;
                if (p.ag_world > 1 && m0 < p.M) {      // rows pushed by a peer: wait until that 128-row block has landed
                    const int owner = min(m0 / p.ag_rows, p.ag_world - 1);
                    if (owner != p.ag_rank) {
                        const unsigned* f = p.ag_flag[p.ag_rank] + owner * (p.ag_rows >> 7) + ((m0 - owner * p.ag_rows) >> 7);
                        unsigned v;
                        do { asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory"); } while (v == 0);
                        asm volatile("fence.proxy.async.global;" ::: "memory");     // generic-proxy writes -> TMA reads
                    }
                }
                for (int kb = kb0; kb < kb1; ++kb, ++it) {
                    const int st = it % k2Stages;
                    if (it >= k2Stages) mbar_wait(&empty_bar[st], ((it / k2Stages) - 1) & 1);
                    if (p.dbg == 1 && it >= k2Stages) {
                        if (rank == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&full_bar[st])) : "memory");
                        continue;
                    }
                    if (rank == 0) mbar_expect_tx(&full_bar[st], 2 * k2StageBytes);
                    uint8_t* a_dst = smem_a + st * k2ABytes;
                    uint8_t* b_dst = smem_b + st * k2BBytes;
                    const int k0 = kb * BLOCK_K;
                    if (!p.a_mn_major) {
                        tma_load_3d_2sm(a_dst, mapA, k0, m0, ba, &full_bar[st]);
                    } else if (p.a_mn4) {
                        tma_load_4d_2sm(a_dst, &tmA, 0, k0, m0 >> 5, ba, &full_bar[st]);
                    } else {
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            tma_load_3d_2sm(a_dst + j * (BLOCK_K * 128), &tmA, m0 + j * 32, k0, ba, &full_bar[st]);
                    }
                    if (!p.b_mn_major) {
                        tma_load_3d_2sm(b_dst, &tmB, k0, nb0, bb, &full_bar[st]);
                    } else if (p.b_mn4) {
                        tma_load_4d_2sm(b_dst, &tmB, 0, k0, nb0 >> 5, bb, &full_bar[st]);
                    } else {
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            tma_load_3d_2sm(b_dst + j * (BLOCK_K * 128), &tmB, nb0 + j * 32, k0, bb, &full_bar[st]);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer (leader CTA only)
        if (lane == 0 && rank == 0) {
            const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)p.a_mn_major << 15) |
                                   ((uint32_t)p.b_mn_major << 16) | ((uint32_t)(k2TileN >> 3) << 17) |
                                   ((uint32_t)(256 >> 4) << 24);
            int it = 0, t_iter = 0, tile, kb0, kb1;
            TileSched sched(cluster_id, num_clusters, num_tiles, num_kb, kb_per_cluster);
            for (; sched.next(tile, kb0, kb1); ++t_iter) {
                const int acc = t_iter & 1;
                if (t_iter >= 2) mbar_wait(&tempty_bar[acc], ((t_iter >> 1) - 1) & 1);   // epilogues drained this buffer
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * k2TileN);
                for (int kb = kb0; kb < kb1; ++kb, ++it) {
                    const int st = it % k2Stages;
                    mbar_wait(&full_bar[st], (it / k2Stages) & 1);
                    tc_fence_after();
                    const uint32_t a_base = smem_u32(smem_a + st * k2ABytes);
                    const uint32_t b_base = smem_u32(smem_b + st * k2BBytes);
#pragma unroll
                    for (int ki = 0; ki < BLOCK_K / UMMA_K; ++ki) {
                        uint64_t a_desc = p.a_mn_major ? make_smem_desc(a_base + ki * (UMMA_K * 128), BLOCK_K * 128, 512, 1)
                                                       : make_smem_desc(a_base + ki * (UMMA_K * 4), 16, 1024, 2);
                        uint64_t b_desc = p.b_mn_major ? make_smem_desc(b_base + ki * (UMMA_K * 128), BLOCK_K * 128, 512, 1)
                                                       : make_smem_desc(b_base + ki * (UMMA_K * 4), 16, 1024, 2);
                        if (p.dbg != 2) umma_tf32_2sm(d_tmem, a_desc, b_desc, idesc, (kb > kb0 || ki > 0) ? 1u : 0u);
                    }
                    umma_commit_2sm(&empty_bar[st]);
                }
                umma_commit_2sm(&tfull_bar[acc]);
            }
        }
    } else {
        // ------------------------------------------------------------ epilogue (warps 2..5, both CTAs)
        const int q = warp & 3;
        float* tile_s = smem_epi + (warp - 2) * (32 * 36);
        const int sub_row = lane >> 3, c4 = (lane & 7) * 4;
        if (p.ag_world > 1) {
            // ---- all-gather push phase (128 threads): my shard -> every peer's gathered buffer
            const int et = threadIdx.x - 64;
            const int blocks = p.ag_rows >> 7;
            const int items = blocks * (p.ag_world - 1);
            const size_t blk_f4 = (size_t)128 * p.K / 4;
            for (int item = blockIdx.x; item < items; item += gridDim.x) {
                const int pi = item / blocks, blk = item - pi * blocks;
                const int peer = pi + (pi >= p.ag_rank ? 1 : 0);
                const size_t off = ((size_t)p.ag_rank * p.ag_rows + (size_t)blk * 128) * p.K;
                const float4* src = reinterpret_cast<const float4*>(p.ag_buf[p.ag_rank] + off);
                float4* dst = reinterpret_cast<float4*>(p.ag_buf[peer] + off);
                for (size_t i = et; i < blk_f4; i += 128 * 8) {
                    float4 v[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) if (i + u * 128 < blk_f4) v[u] = __ldg(src + i + u * 128);
#pragma unroll
                    for (int u = 0; u < 8; ++u) if (i + u * 128 < blk_f4) dst[i + u * 128] = v[u];
                }
                __threadfence_system();
                asm volatile("bar.sync 1, 128;" ::: "memory");
                if (et == 0) {
                    unsigned* f = p.ag_flag[peer] + p.ag_rank * blocks + blk;
                    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(f), "r"(1u) : "memory");
                }
            }
        }
        int t_iter = 0, tile, kb0, kb1;
        TileSched sched(cluster_id, num_clusters, num_tiles, num_kb, kb_per_cluster);
        for (; sched.next(tile, kb0, kb1); ++t_iter) {
            const int acc = t_iter & 1;
            const bool partial = kb0 != 0 || kb1 != num_kb || p.rs_world > 1;   // accumulate into C (stream-K / reduce-scatter)
            int bz, tm, tn;
            decode_tile(tile, tiles_m, tiles_n, bz, tm, tn);
            if (p.ag_world > 1) tm = ag_rotate(tm, tiles_m, p);
            const int m0 = tm * 256 + (int)rank * 128, n0 = tn * k2TileN;
            mbar_wait(&tfull_bar[acc], (t_iter >> 1) & 1);
            tc_fence_after();
            float* Cb = p.C + (size_t)bz * p.sC;
            const int row_base = m0 + q * 32;
            const bool vec_ok = (p.N & 3) == 0 && ((reinterpret_cast<uintptr_t>(Cb) & 15) == 0);
#pragma unroll 1
            for (int c0 = 0; c0 < k2TileN; c0 += 32) {
                if (n0 + c0 >= p.N || row_base >= p.M) break;
                float v[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * k2TileN + c0), v);
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    *reinterpret_cast<float4*>(&tile_s[lane * 36 + j * 4]) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                __syncwarp();
                const int col = n0 + c0 + c4;
                if (partial) {
                    for (int r = 0; r < 32; r += 4) {
                        const int row = row_base + r + sub_row;
                        if (row >= p.M) continue;
                        float* dst = Cb + (size_t)row * p.N + col;
                        const float* src = &tile_s[(r + sub_row) * 36 + c4];
                        if (p.rs_world > 1) {              // my slice of the row owner's staging buffer, over NVLink
                            const int owner = min(row / p.rs_rows, p.rs_world - 1);
                            dst = p.rs_out[owner] + ((size_t)p.rs_rank * p.rs_rows + (row - owner * p.rs_rows)) * p.N + col;
                            if (vec_ok && col + 3 < p.N) {
                                *reinterpret_cast<float4*>(dst) = make_float4(p.alpha * src[0], p.alpha * src[1], p.alpha * src[2], p.alpha * src[3]);
                            } else {
                                for (int u = 0; u < 4 && col + u < p.N; ++u) dst[u] = p.alpha * src[u];
                            }
                            continue;
                        }
                        if (vec_ok && col + 3 < p.N) {
                            float4 o = make_float4(p.alpha * src[0], p.alpha * src[1], p.alpha * src[2], p.alpha * src[3]);
                            if (kb0 == 0 && p.bias != nullptr) {
                                const float4 bv = __ldg(reinterpret_cast<const float4*>(p.bias + col));
                                o.x += bv.x; o.y += bv.y; o.z += bv.z; o.w += bv.w;
                            }
                            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(o.x), "f"(o.y), "f"(o.z), "f"(o.w) : "memory");
                        } else {
                            for (int u = 0; u < 4 && col + u < p.N; ++u)
                                atomicAdd(dst + u, p.alpha * src[u] + ((kb0 == 0 && p.bias) ? __ldg(p.bias + col + u) : 0.f));
                        }
                    }
                } else if (vec_ok && col + 3 < p.N) {
                    float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (p.bias != nullptr) bv = __ldg(reinterpret_cast<const float4*>(p.bias + col));
                    // C is read for beta / the relu mask: all eight loads of this warp pass are issued before the first store
                    const bool need_c = p.beta != 0.f || p.act == ACT_RELU_MASK;
                    float4 cin[8];
                    if (need_c) {
#pragma unroll
                        for (int r = 0; r < 32; r += 4) {
                            const int row = row_base + r + sub_row;
                            cin[r >> 2] = row < p.M ? *reinterpret_cast<const float4*>(Cb + (size_t)row * p.N + col) : make_float4(0.f, 0.f, 0.f, 0.f);
                        }
                    }
#pragma unroll
                    for (int r = 0; r < 32; r += 4) {
                        const int row = row_base + r + sub_row;
                        if (row < p.M) {
                            float4 a = *reinterpret_cast<const float4*>(&tile_s[(r + sub_row) * 36 + c4]);
                            float4 o = make_float4(p.alpha * a.x + bv.x, p.alpha * a.y + bv.y, p.alpha * a.z + bv.z, p.alpha * a.w + bv.w);
                            float4* dst = reinterpret_cast<float4*>(Cb + (size_t)row * p.N + col);
                            if (p.beta != 0.f) {
                                const float4 c = cin[r >> 2];
                                o.x += p.beta * c.x; o.y += p.beta * c.y; o.z += p.beta * c.z; o.w += p.beta * c.w;
                            }
                            if (p.act == ACT_RELU_MASK) {
                                const float4 h = cin[r >> 2];
                                o.x = h.x > 0.f ? o.x : 0.f; o.y = h.y > 0.f ? o.y : 0.f; o.z = h.z > 0.f ? o.z : 0.f; o.w = h.w > 0.f ? o.w : 0.f;
                            } else if (p.act != ACT_NONE) { o.x = apply_act(o.x, p.act); o.y = apply_act(o.y, p.act); o.z = apply_act(o.z, p.act); o.w = apply_act(o.w, p.act); }
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
                            float o = p.alpha * tile_s[(r + sub_row) * 36 + c4 + u] + (p.bias ? __ldg(p.bias + cc) : 0.f);
                            float* dst = Cb + (size_t)row * p.N + cc;
                            if (p.beta != 0.f) o += p.beta * (*dst);
                            *dst = p.act == ACT_RELU_MASK ? (*dst > 0.f ? o : 0.f) : apply_act(o, p.act);
                        }
                    }
                }
                __syncwarp();
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_leader(&tempty_bar[acc]);      // this warp no longer reads accumulator `acc`
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                     // the peer's smem / TMEM stay valid until the leader's MMAs retired
    if (warp == 1) tmem_dealloc_2sm(tmem_base, 512);
}

static int launch_tc_2sm(const float* A, const float* B, GemmParams& p, int batch, long long sA, long long sB, cudaStream_t s) {
    CUtensorMap tmA, tmB;
    p.a_mn4 = p.b_mn4 = 0;
    if (!p.a_mn_major) { if (make_tmap(&tmA, A, p.K, p.M, p.a_batched ? batch : 1, sA, BLOCK_M, false)) return -1; }
    else if (p.M % 128 == 0 && make_tmap_mn4(&tmA, A, p.M, p.K, p.a_batched ? batch : 1, sA) == 0) p.a_mn4 = 1;
    else { if (make_tmap(&tmA, A, p.M, p.K, p.a_batched ? batch : 1, sA, BLOCK_K, true)) return -1; }
    if (!p.b_mn_major) { if (make_tmap(&tmB, B, p.K, p.N, p.b_batched ? batch : 1, sB, 128, false)) return -1; }
    else if (p.N % 128 == 0 && make_tmap_mn4(&tmB, B, p.N, p.K, p.b_batched ? batch : 1, sB) == 0) p.b_mn4 = 1;
    else { if (make_tmap(&tmB, B, p.N, p.K, p.b_batched ? batch : 1, sB, BLOCK_K, true)) return -1; }
    static int max_clusters[64] = {0};
    int dev = 0;
    HETU_CUDA(cudaGetDevice(&dev));
    if (max_clusters[dev & 63] == 0) {
        HETU_CUDA(cudaFuncSetAttribute(gemm_tf32_2sm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, k2Smem));
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(2, 1, 1);
        cfg.blockDim = dim3(kGemmThreads, 1, 1);
        cfg.dynamicSmemBytes = k2Smem;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        int n = 0;
        if (cudaOccupancyMaxActiveClusters(&n, gemm_tf32_2sm_kernel, &cfg) != cudaSuccess || n <= 0) {
            cudaGetLastError();
            int sms = 0;
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            n = sms / 2;
        }
        max_clusters[dev & 63] = n;
    }
    const int tiles_m = (p.M + 255) / 256, tiles_n = (p.N + k2TileN - 1) / k2TileN;
    const int num_tiles = tiles_m * tiles_n * batch;
    const int maxc = max_clusters[dev & 63];
    int clusters = num_tiles < maxc ? num_tiles : maxc;
    const int num_kb = (p.K + BLOCK_K - 1) / BLOCK_K;
    // stream-K when whole-tile scheduling would leave >10% of the machine idle in the last wave
    const int waves = (num_tiles + maxc - 1) / maxc;
    const double eff = (double)num_tiles / ((double)waves * maxc);
    int kb_per_cluster = 0;
    static int use_streamk = -1;
    if (use_streamk < 0) { const char* e = getenv("HETU_GEMM_STREAMK"); use_streamk = (e && e[0] == '1') ? 1 : 0; }   // opt-in: measured slower than whole tiles (atomics + lost L2 lockstep)
    // few output tiles but a very long K (weight gradients of convolutions / small layers): without
    // splitting K only num_tiles CTA pairs would work, so stream-K is always used there
    // (round 2: from 64 k-blocks on, and also when a single wave fills at most 60% of the CTA pairs - the weight
    // gradients of transformer layers, e.g. 1024 x 1024 x 4096: 32 one-SM CTAs took 74 us, cuBLAS 25 us)
    static int sk_policy = -1;
    if (sk_policy < 0) { const char* e = getenv("HETU_GEMM_STREAMK_POLICY"); sk_policy = (e && e[0] == '0') ? 0 : 1; }
    const int min_kb = sk_policy ? 64 : 256;
    const bool starved = (num_tiles * 4 <= maxc || (sk_policy && waves == 1 && eff <= 0.6)) && num_kb >= min_kb;
    if ((starved || (use_streamk && eff < 0.9 && num_kb >= 128)) && p.beta == 0.f && p.act == ACT_NONE && p.rs_world <= 1 &&
        p.ag_world <= 1 && (long long)num_tiles * num_kb >= 8LL * maxc) {
        clusters = maxc;
        const long long total = (long long)num_tiles * num_kb;
        kb_per_cluster = (int)((total + clusters - 1) / clusters);
        HETU_CUDA(cudaMemsetAsync(p.C, 0, sizeof(float) * (size_t)(batch > 1 ? p.sC * batch : (long long)p.M * p.N), s));
    }
    gemm_tf32_2sm_kernel<<<dim3(2 * clusters), kGemmThreads, k2Smem, s>>>(tmA, tmB, p, tiles_m, tiles_n, num_tiles, kb_per_cluster);
    return 0;
}

// ==================================================================== SIMT fp32 kernel
// 64x64 tile, BK = 16, 256 threads, 4x4 micro-tile per thread; all transposes by indexing.
__global__ void __launch_bounds__(256) gemm_fp32_simt_kernel(const float* __restrict__ A, const float* __restrict__ B,
                                                             float* __restrict__ C, const float* __restrict__ bias, int M,
                                                             int N, int K, int tA, int tB, long long sA, long long sB,
                                                             long long sC, float alpha, float beta, int act) {
    __shared__ float As[16][64 + 4];
    __shared__ float Bs[16][64 + 4];
    const int bz = blockIdx.z;
    A += (size_t)bz * sA; B += (size_t)bz * sB; C += (size_t)bz * sC;
    const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    float acc[4][4] = {};
    for (int k0 = 0; k0 < K; k0 += 16) {
        for (int i = threadIdx.x; i < 16 * 64; i += 256) {
            int kk, mm;
            if (!tA) { kk = i & 15; mm = i >> 4; } else { mm = i & 63; kk = i >> 6; }   // keep global reads contiguous
            int gm = m0 + mm, gk = k0 + kk;
            float v = 0.f;
            if (gm < M && gk < K) v = tA ? A[(size_t)gk * M + gm] : A[(size_t)gm * K + gk];
            As[kk][mm] = v;
        }
        for (int i = threadIdx.x; i < 16 * 64; i += 256) {
            int kk, nn;
            if (!tB) { nn = i & 63; kk = i >> 6; } else { kk = i & 15; nn = i >> 4; }
            int gn = n0 + nn, gk = k0 + kk;
            float v = 0.f;
            if (gn < N && gk < K) v = tB ? B[(size_t)gn * K + gk] : B[(size_t)gk * N + gn];
            Bs[kk][nn] = v;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < 16; ++kk) {
            float a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = Bs[kk][tx * 4 + j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        int gm = m0 + ty * 4 + i;
        if (gm >= M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int gn = n0 + tx * 4 + j;
            if (gn >= N) continue;
            float o = alpha * acc[i][j] + (bias ? bias[gn] : 0.f);
            float* dst = C + (size_t)gm * N + gn;
            if (beta != 0.f) o += beta * (*dst);
            *dst = apply_act(o, act);
        }
    }
}

// Skinny GEMV-like case N == 1 (WDL's final [B, 3584] x [3584, 1]) and its transposes are
// handled well enough by the SIMT kernel; a split-K variant keeps dW-style products
// (tiny M x N, huge K) from running on a single CTA.
__global__ void __launch_bounds__(256) gemm_fp32_splitk_kernel(const float* __restrict__ A, const float* __restrict__ B,
                                                               float* __restrict__ C, int M, int N, int K, int tA, int tB,
                                                               float alpha, int kchunk) {
    // each CTA: 16x16 output tile, K range [blockIdx.z*kchunk, +kchunk); atomicAdd into zeroed C
    __shared__ float As[16][17], Bs[16][17];
    int m0 = blockIdx.y * 16, n0 = blockIdx.x * 16;
    int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    int kbeg = blockIdx.z * kchunk, kend = min(K, kbeg + kchunk);
    float acc = 0.f;
    for (int k0 = kbeg; k0 < kend; k0 += 16) {
        int gm = m0 + ty, gk = k0 + tx;
        As[ty][tx] = (gm < M && gk < kend) ? (tA ? A[(size_t)gk * M + gm] : A[(size_t)gm * K + gk]) : 0.f;
        int gn = n0 + tx; gk = k0 + ty;
        Bs[ty][tx] = (gn < N && gk < kend) ? (tB ? B[(size_t)gn * K + gk] : B[(size_t)gk * N + gn]) : 0.f;
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < 16; ++kk) acc = fmaf(As[ty][kk], Bs[kk][tx], acc);
        __syncthreads();
    }
    int gm = m0 + ty, gn = n0 + tx;
    if (gm < M && gn < N) atomicAdd(C + (size_t)gm * N + gn, alpha * acc);
}

__global__ void __launch_bounds__(256) gemm_finish_kernel(float* __restrict__ C, const float* __restrict__ bias, size_t n,
                                                          int N, int act) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        C[i] = apply_act(C[i] + (bias ? bias[i % N] : 0.f), act);
}

// ===================================================================== small-GEMM kernel
// Latency-optimised path for the tiny products of recommendation models (batch 128 x 256 x 256,
// 128 x 3584 x 1 ...): 16 x 64 output tile per CTA so even a 128 x 256 output spreads over 32
// SMs, K consumed in chunks of 32 with register prefetch of the next chunk, optional split-K
// (gridDim.z) with atomic accumulation for long reductions feeding a handful of tiles.
constexpr int KC = 128;     // K elements per chunk: 40 loads in flight per thread hide the L2/HBM latency
template <int TA, int TB>
__global__ void __launch_bounds__(256) gemm_small_kernel(const float* __restrict__ A, const float* __restrict__ B,
                                                         float* __restrict__ C, const float* __restrict__ bias, int M, int N,
                                                         int K, long long sA, long long sB, long long sC, float alpha,
                                                         float beta, int act, int kchunk, int splits, int batch) {
    __shared__ float As[KC][16 + 1];
    __shared__ __align__(16) float Bs[KC][64];
    const int bz = blockIdx.z / splits, split = blockIdx.z % splits;
    A += (size_t)bz * sA; B += (size_t)bz * sB; C += (size_t)bz * sC;
    const int m0 = blockIdx.y * 16, n0 = blockIdx.x * 64;
    const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
    const int kbeg = split * kchunk, kend = min(K, kbeg + kchunk);
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    constexpr int NA = 16 * KC / 256, NB = 64 * KC / 256;
    float ra[NA], rb[NB];
    auto load_chunk = [&](int k0) {
#pragma unroll
        for (int u = 0; u < NA; ++u) {
            int e = tid + u * 256;
            int kk, mm;
            if (!TA) { kk = e % KC; mm = e / KC; } else { mm = e & 15; kk = e >> 4; }
            int gm = m0 + mm, gk = k0 + kk;
            ra[u] = (gm < M && gk < kend) ? (TA ? A[(size_t)gk * M + gm] : A[(size_t)gm * K + gk]) : 0.f;
        }
#pragma unroll
        for (int u = 0; u < NB; ++u) {
            int e = tid + u * 256;
            int kk, nn;
            if (!TB) { nn = e & 63; kk = e >> 6; } else { kk = e % KC; nn = e / KC; }
            int gn = n0 + nn, gk = k0 + kk;
            rb[u] = (gn < N && gk < kend) ? (TB ? B[(size_t)gn * K + gk] : B[(size_t)gk * N + gn]) : 0.f;
        }
    };
    auto store_chunk = [&]() {
#pragma unroll
        for (int u = 0; u < NA; ++u) {
            int e = tid + u * 256;
            int kk, mm;
            if (!TA) { kk = e % KC; mm = e / KC; } else { mm = e & 15; kk = e >> 4; }
            As[kk][mm] = ra[u];
        }
#pragma unroll
        for (int u = 0; u < NB; ++u) {
            int e = tid + u * 256;
            int kk, nn;
            if (!TB) { nn = e & 63; kk = e >> 6; } else { kk = e % KC; nn = e / KC; }
            Bs[kk][nn] = rb[u];
        }
    };
    if (kbeg < kend) load_chunk(kbeg);
    for (int k0 = kbeg; k0 < kend; k0 += KC) {
        store_chunk();
        __syncthreads();
        if (k0 + KC < kend) load_chunk(k0 + KC);      // prefetch while computing
        const int kmax = min(KC, kend - k0);
#pragma unroll 8
        for (int kk = 0; kk < kmax; ++kk) {
            float a = As[kk][ty];
            float4 b = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
            acc[0] = fmaf(a, b.x, acc[0]); acc[1] = fmaf(a, b.y, acc[1]);
            acc[2] = fmaf(a, b.z, acc[2]); acc[3] = fmaf(a, b.w, acc[3]);
        }
        __syncthreads();
    }
    const int gm = m0 + ty;
    if (gm >= M) return;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int gn = n0 + tx * 4 + j;
        if (gn >= N) continue;
        float* dst = C + (size_t)gm * N + gn;
        if (splits > 1) {
            atomicAdd(dst, alpha * acc[j]);           // C pre-zeroed; bias/act applied by the finish kernel
        } else {
            float o = alpha * acc[j] + (bias ? bias[gn] : 0.f);
            if (beta != 0.f) o += beta * (*dst);
            *dst = apply_act(o, act);
        }
    }
}

static void launch_small(const float* A, const float* B, float* C, const float* bias, int M, int N, int K, int tA, int tB,
                         int batch, long long sA, long long sB, long long sC, float alpha, float beta, int act,
                         cudaStream_t s) {
    int tiles = ((M + 15) / 16) * ((N + 63) / 64) * batch;
    int splits = 1;
    if (beta == 0.f && tiles < kNumSMs && K >= 2 * KC) {
        splits = (kNumSMs + tiles - 1) / tiles;
        int maxs = K / KC;
        if (splits > maxs) splits = maxs;
        if (splits < 1) splits = 1;
    }
    int kchunk = ((K + splits - 1) / splits + KC - 1) / KC * KC;
    splits = (K + kchunk - 1) / kchunk;
    dim3 grid((N + 63) / 64, (M + 15) / 16, batch * splits);
    if (splits > 1) cudaMemsetAsync(C, 0, sizeof(float) * (size_t)M * N * batch, s);
#define SMALL(TA_, TB_) gemm_small_kernel<TA_, TB_><<<grid, 256, 0, s>>>(A, B, C, bias, M, N, K, sA, sB, sC, alpha, beta, act, kchunk, splits, batch)
    if (!tA && !tB) SMALL(0, 0); else if (!tA && tB) SMALL(0, 1); else if (tA && !tB) SMALL(1, 0); else SMALL(1, 1);
#undef SMALL
    if (splits > 1 && (bias || act != ACT_NONE))
        gemm_finish_kernel<<<ew_grid((size_t)M * N * batch, 256), 256, 0, s>>>(C, bias, (size_t)M * N * batch, N, act);
}

// ================================================================= skinny GEMM (N <= 8)
// MoE gate / logistic-regression shapes: [tokens, d] @ [d, n_experts] and its weight gradient.  These are
// bandwidth-bound reads of the big operand; a tile kernel wastes 8-32x of its columns on them.
constexpr int kSkinnyN = 8;
constexpr size_t kSkinnySmemMax = 96 * 1024;     // B table of the row kernel in shared memory (dynamic, opt-in above 48 KB)

// A row-major [M, K] (not transposed): one warp per output row, lanes stride K with 128-bit loads.
template <int TB>
__global__ void __launch_bounds__(256) gemm_skinny_rows_kernel(const float* __restrict__ A, const float* __restrict__ B,
                                                               float* __restrict__ C, const float* __restrict__ bias,
                                                               int M, int N, int K, float alpha, float beta, int act) {
    extern __shared__ float4 skinny_smem[];
    float* Bs = reinterpret_cast<float*>(skinny_smem);      // [N][K] when it fits (bs_ok), else unused
    const int lane = threadIdx.x & 31;
    const bool bs_ok = (size_t)N * K * 4 <= kSkinnySmemMax && (K & 3) == 0;     // must match launch_skinny
    if (bs_ok) {
        for (int i = threadIdx.x; i < N * K; i += 256) {
            const int n = i / K, k = i - n * K;
            Bs[i] = TB ? __ldg(B + i) : __ldg(B + (size_t)k * N + n);
        }
        __syncthreads();
    }
    for (int row = blockIdx.x * 8 + (threadIdx.x >> 5); row < M; row += gridDim.x * 8) {
    const float* a = A + (size_t)row * K;
    float acc[kSkinnyN];
#pragma unroll
    for (int n = 0; n < kSkinnyN; ++n) acc[n] = 0.f;
    const bool vec = (K & 3) == 0 && ((reinterpret_cast<uintptr_t>(A) & 15) == 0);
    if (vec && bs_ok) {
        for (int kb = lane * 4; kb < K; kb += 512) {
            float4 av[4];
#pragma unroll
            for (int u = 0; u < 4; ++u)
                av[u] = (kb + u * 128 < K) ? __ldcs(reinterpret_cast<const float4*>(a + kb + u * 128)) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int k = kb + u * 128;
                if (k >= K) break;
#pragma unroll
                for (int n = 0; n < kSkinnyN; ++n) {
                    if (n < N) {
                        const float4 bv = *reinterpret_cast<const float4*>(Bs + n * K + k);
                        acc[n] += av[u].x * bv.x + av[u].y * bv.y + av[u].z * bv.z + av[u].w * bv.w;
                    }
                }
            }
        }
    } else if (vec) {
        for (int kb = lane * 4; kb < K; kb += 512) {
            float4 av[4];
#pragma unroll
            for (int u = 0; u < 4; ++u)             // four independent 128-bit loads in flight per lane
                av[u] = (kb + u * 128 < K) ? __ldcs(reinterpret_cast<const float4*>(a + kb + u * 128)) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int k = kb + u * 128;
                if (k >= K) break;
#pragma unroll
                for (int n = 0; n < kSkinnyN; ++n) {
                    if (n < N) {
                        if (TB) {   // B stored [N, K]
                            const float4 bv = __ldg(reinterpret_cast<const float4*>(B + (size_t)n * K + k));
                            acc[n] += av[u].x * bv.x + av[u].y * bv.y + av[u].z * bv.z + av[u].w * bv.w;
                        } else {    // B stored [K, N]
                            const float* b = B + (size_t)k * N + n;
                            acc[n] += av[u].x * __ldg(b) + av[u].y * __ldg(b + N) + av[u].z * __ldg(b + 2 * N) + av[u].w * __ldg(b + 3 * N);
                        }
                    }
                }
            }
        }
    } else {
        for (int k = lane; k < K; k += 32) {
            const float av = __ldg(a + k);
#pragma unroll
            for (int n = 0; n < kSkinnyN; ++n)
                if (n < N) acc[n] += av * __ldg(TB ? B + (size_t)n * K + k : B + (size_t)k * N + n);
        }
    }
#pragma unroll
    for (int n = 0; n < kSkinnyN; ++n) {
        float v = acc[n];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0 && n < N) {
            float o = alpha * v + (bias ? __ldg(bias + n) : 0.f);
            float* dst = C + (size_t)row * N + n;
            if (beta != 0.f) o += beta * (*dst);
            *dst = apply_act(o, act);
        }
    }
    }   // row loop
}

// A stored [K, M] (transposed operand): threads run along M (coalesced), 4 k-lanes per block, K split over
// blockIdx.y; partial sums are added atomically into the zero-initialised C.
template <int TB>
__global__ void __launch_bounds__(256) gemm_skinny_cols_kernel(const float* __restrict__ A, const float* __restrict__ B,
                                                               float* __restrict__ C, int M, int N, int K, int kchunk,
                                                               float alpha) {
    __shared__ float red[4][64][kSkinnyN + 1];
    const int mt = threadIdx.x & 63, kl = threadIdx.x >> 6;
    const int m = blockIdx.x * 64 + mt;
    const int k_beg = blockIdx.y * kchunk, k_end = min(K, k_beg + kchunk);
    float acc[kSkinnyN];
#pragma unroll
    for (int n = 0; n < kSkinnyN; ++n) acc[n] = 0.f;
    if (m < M) {
#pragma unroll 8
        for (int k = k_beg + kl; k < k_end; k += 4) {
            const float av = __ldcs(A + (size_t)k * M + m);
#pragma unroll
            for (int n = 0; n < kSkinnyN; ++n)
                if (n < N) acc[n] += av * __ldg(TB ? B + (size_t)n * K + k : B + (size_t)k * N + n);
        }
    }
#pragma unroll
    for (int n = 0; n < kSkinnyN; ++n) red[kl][mt][n] = acc[n];
    __syncthreads();
    if (kl == 0 && m < M) {
#pragma unroll
        for (int n = 0; n < kSkinnyN; ++n)
            if (n < N) atomicAdd(C + (size_t)m * N + n, alpha * (red[0][mt][n] + red[1][mt][n] + red[2][mt][n] + red[3][mt][n]));
    }
}

// Same product when M % 4 == 0 and A is 16-byte aligned: every thread owns four consecutive m (one 128-bit load per
// k), 32 lanes cover 128 m, 8 k-lanes per block with four loads in flight each - enough bytes in flight per SM to
// stream A at HBM speed (the scalar kernel above stalls at ~1.3 TB/s on the MoE gate's weight gradient).
template <int TB>
__global__ void __launch_bounds__(256) gemm_skinny_cols_vec4_kernel(const float* __restrict__ A, const float* __restrict__ B,
                                                                    float* __restrict__ C, int M, int N, int K, int kchunk,
                                                                    float alpha) {
    __shared__ float red[8][32][4 * kSkinnyN + 1];
    const int lane = threadIdx.x & 31, kl = threadIdx.x >> 5;
    const int m = (blockIdx.x * 32 + lane) * 4;
    const int k_beg = blockIdx.y * kchunk, k_end = min(K, k_beg + kchunk);
    float acc[kSkinnyN][4];
#pragma unroll
    for (int n = 0; n < kSkinnyN; ++n) acc[n][0] = acc[n][1] = acc[n][2] = acc[n][3] = 0.f;
    if (m < M) {
        for (int k0 = k_beg + kl; k0 < k_end; k0 += 32) {
            float4 av[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int k = k0 + u * 8;
                av[u] = k < k_end ? __ldcs(reinterpret_cast<const float4*>(A + (size_t)k * M + m)) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int k = k0 + u * 8;
                if (k >= k_end) break;
#pragma unroll
                for (int n = 0; n < kSkinnyN; ++n) {
                    if (n < N) {
                        const float b = __ldg(TB ? B + (size_t)n * K + k : B + (size_t)k * N + n);
                        acc[n][0] += av[u].x * b; acc[n][1] += av[u].y * b; acc[n][2] += av[u].z * b; acc[n][3] += av[u].w * b;
                    }
                }
            }
        }
    }
#pragma unroll
    for (int n = 0; n < kSkinnyN; ++n)
#pragma unroll
        for (int j = 0; j < 4; ++j) red[kl][lane][n * 4 + j] = acc[n][j];
    __syncthreads();
    // 128 m x N outputs per block: thread -> (m offset, n)
    for (int o = threadIdx.x; o < 128 * N; o += 256) {
        const int mo = o / N, n = o - mo * N;
        const int mm = blockIdx.x * 128 + mo;
        if (mm >= M) continue;
        float v = 0.f;
#pragma unroll
        for (int q = 0; q < 8; ++q) v += red[q][mo >> 2][n * 4 + (mo & 3)];
        atomicAdd(C + (size_t)mm * N + n, alpha * v);
    }
}

// K <= 8: a sum of a few outer products (gradient of a [*, d] x [d, 1] projection); one thread per output.
__global__ void __launch_bounds__(256) gemm_tinyk_kernel(const float* __restrict__ A, const float* __restrict__ B,
                                                         float* __restrict__ C, const float* __restrict__ bias, int M, int N,
                                                         int K, int tA, int tB, float alpha, float beta, int act) {
    const size_t total = (size_t)M * N;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int m = (int)(i / N), n = (int)(i - (size_t)m * N);
        float acc = 0.f;
        for (int k = 0; k < K; ++k)
            acc += __ldg(tA ? A + (size_t)k * M + m : A + (size_t)m * K + k) * __ldg(tB ? B + (size_t)n * K + k : B + (size_t)k * N + n);
        float o = alpha * acc + (bias ? __ldg(bias + n) : 0.f);
        if (beta != 0.f) o += beta * C[i];
        C[i] = apply_act(o, act);
    }
}

// K <= 8 (measured: from K = 9 on the tiled kernels win), A not transposed, N % 4 == 0: every thread owns four consecutive columns and keeps their K x 4 block of B in
// registers while it walks the 64 rows of its row block (32-bit index arithmetic, one 128-bit store per row).  The
// per-element kernel above spends a 64-bit div / mod per output and ran at 0.4 TB/s on the MoE gate's input gradient
// [tokens, E] x [E, d].
constexpr int kTinyKRows = 64;
template <int TB>
__global__ void __launch_bounds__(256) gemm_tinyk_rows_kernel(const float* __restrict__ A, const float* __restrict__ B,
                                                              float* __restrict__ C, const float* __restrict__ bias, int M, int N,
                                                              int K, float alpha, float beta, int act) {
    const int n = (blockIdx.x * 256 + threadIdx.x) * 4;
    if (n >= N) return;
    float4 b[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        b[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (k < K) {
            if (TB) b[k] = make_float4(__ldg(B + (size_t)n * K + k), __ldg(B + (size_t)(n + 1) * K + k), __ldg(B + (size_t)(n + 2) * K + k),
                                       __ldg(B + (size_t)(n + 3) * K + k));
            else b[k] = __ldg(reinterpret_cast<const float4*>(B + (size_t)k * N + n));
        }
    }
    float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
    if (bias != nullptr) bv = __ldg(reinterpret_cast<const float4*>(bias + n));
    const int m_end = min(M, (int)(blockIdx.y + 1) * kTinyKRows);
    for (int m = blockIdx.y * kTinyKRows; m < m_end; ++m) {
        const float* a = A + (size_t)m * K;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            if (k < K) {
                const float av = __ldg(a + k);                        // same address for the whole block: broadcast
                acc.x += av * b[k].x; acc.y += av * b[k].y; acc.z += av * b[k].z; acc.w += av * b[k].w;
            }
        }
        float4 o = make_float4(alpha * acc.x + bv.x, alpha * acc.y + bv.y, alpha * acc.z + bv.z, alpha * acc.w + bv.w);
        float4* dst = reinterpret_cast<float4*>(C + (size_t)m * N + n);
        if (beta != 0.f) { const float4 c = *dst; o.x += beta * c.x; o.y += beta * c.y; o.z += beta * c.z; o.w += beta * c.w; }
        *dst = make_float4(apply_act(o.x, act), apply_act(o.y, act), apply_act(o.z, act), apply_act(o.w, act));
    }
}

static void launch_skinny(const float* A, const float* B, float* C, const float* bias, int M, int N, int K, int tA, int tB,
                          float alpha, float beta, int act, cudaStream_t s) {
    if (!tA) {
        dim3 grid(std::min((M + 7) / 8, kNumSMs * 6));
        // B in shared memory up to 96 KB (opt-in above 48 KB): with 8 experts and d = 2048 the table is 64 KB, and without it
        // every lane fetched B from global memory per k (1 ms for the MoE gate at 4 GPUs - the N = 4 anomaly of round 1)
        const size_t sm = ((size_t)N * K * 4 <= kSkinnySmemMax && (K & 3) == 0) ? (size_t)N * K * 4 : 0;
        static bool attr_done = false;
        if (!attr_done) {
            cudaFuncSetAttribute(gemm_skinny_rows_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSkinnySmemMax);
            cudaFuncSetAttribute(gemm_skinny_rows_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSkinnySmemMax);
            attr_done = true;
        }
        if (tB) gemm_skinny_rows_kernel<1><<<grid, 256, sm, s>>>(A, B, C, bias, M, N, K, alpha, beta, act);
        else gemm_skinny_rows_kernel<0><<<grid, 256, sm, s>>>(A, B, C, bias, M, N, K, alpha, beta, act);
        return;
    }
    if ((M & 3) == 0 && (reinterpret_cast<uintptr_t>(A) & 15) == 0 && K >= 256) {
        const int mb = (M + 127) / 128;
        int sp = (4 * kNumSMs + mb - 1) / mb;
        int kc = (K + sp - 1) / sp;
        kc = (kc + 31) / 32 * 32;
        sp = (K + kc - 1) / kc;
        cudaMemsetAsync(C, 0, sizeof(float) * (size_t)M * N, s);
        dim3 g(mb, sp);
        if (tB) gemm_skinny_cols_vec4_kernel<1><<<g, 256, 0, s>>>(A, B, C, M, N, K, kc, alpha);
        else gemm_skinny_cols_vec4_kernel<0><<<g, 256, 0, s>>>(A, B, C, M, N, K, kc, alpha);
        if (bias || act != ACT_NONE)
            gemm_finish_kernel<<<ew_grid((size_t)M * N, 256), 256, 0, s>>>(C, bias, (size_t)M * N, N, act);
        return;
    }
    const int mblocks = (M + 63) / 64;
    int splits = (4 * kNumSMs + mblocks - 1) / mblocks;
    int kchunk = (K + splits - 1) / splits;
    kchunk = (kchunk + 63) / 64 * 64;
    splits = (K + kchunk - 1) / kchunk;
    cudaMemsetAsync(C, 0, sizeof(float) * (size_t)M * N, s);
    dim3 grid(mblocks, splits);
    if (tB) gemm_skinny_cols_kernel<1><<<grid, 256, 0, s>>>(A, B, C, M, N, K, kchunk, alpha);
    else gemm_skinny_cols_kernel<0><<<grid, 256, 0, s>>>(A, B, C, M, N, K, kchunk, alpha);
    if (bias || act != ACT_NONE)
        gemm_finish_kernel<<<ew_grid((size_t)M * N, 256), 256, 0, s>>>(C, bias, (size_t)M * N, N, act);
}

// ============================================================================ dispatch
static bool aligned16p(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

HETU_API int hetu_gemm(const float* A, const float* B, float* C, const float* bias, int M, int N, int K, int transA,
                       int transB, int batch, long long sA, long long sB, long long sC, float alpha, float beta, int act,
                       int mode, cudaStream_t s) {
    if (M <= 0 || N <= 0 || batch <= 0) return 0;
    if (K <= 0) {
        hetu_set_error("hetu_gemm: K must be positive");
        return -1;
    }
    // TMA needs 16-byte aligned bases and row pitches: the contiguous dim of each operand % 4 == 0
    long long a_inner = transA ? M : K, b_inner = transB ? K : N;
    bool tma_ok = aligned16p(A) && aligned16p(B) && (a_inner % 4 == 0) && (b_inner % 4 == 0) &&
                  (sA % 4 == 0) && (sB % 4 == 0);
    const double macs = (double)M * N * K * batch;
    bool big = macs > (double)(1 << 25) && M >= 64 && N >= 8 && K >= 32;
    if (act == ACT_RELU_MASK && !(mode != 1 && tma_ok && big && K > 8 && N > kSkinnyN && beta == 0.f)) {
        hetu_set_error("hetu_gemm: the relu-mask epilogue exists in the tensor-core kernels only (M %d N %d K %d)", M, N, K);
        return -1;
    }
    if (K <= 8 && batch == 1 && !transA && (N & 3) == 0 && N >= 256 && M <= 65535 * kTinyKRows && act != ACT_RELU_MASK &&
        ((reinterpret_cast<uintptr_t>(B) | reinterpret_cast<uintptr_t>(C) | reinterpret_cast<uintptr_t>(bias)) & 15) == 0) {
        dim3 grid((N / 4 + 255) / 256, (M + kTinyKRows - 1) / kTinyKRows);
        if (transB) gemm_tinyk_rows_kernel<1><<<grid, 256, 0, s>>>(A, B, C, bias, M, N, K, alpha, beta, act);
        else gemm_tinyk_rows_kernel<0><<<grid, 256, 0, s>>>(A, B, C, bias, M, N, K, alpha, beta, act);
        HETU_CHECK_LAUNCH("hetu_gemm(tiny-k rows)");
        return 0;
    }
    if (K <= 8 && batch == 1) {
        gemm_tinyk_kernel<<<ew_grid((size_t)M * N, 256), 256, 0, s>>>(A, B, C, bias, M, N, K, transA, transB, alpha, beta, act);
        HETU_CHECK_LAUNCH("hetu_gemm(tiny-k)");
        return 0;
    }
    if (N <= kSkinnyN && batch == 1 && (double)M * K >= (double)(1 << 16) && (!transA || beta == 0.f)) {
        launch_skinny(A, B, C, bias, M, N, K, transA, transB, alpha, beta, act, s);
        HETU_CHECK_LAUNCH("hetu_gemm(skinny)");
        return 0;
    }
    static int force_simt = -1;
    if (force_simt < 0) { const char* e = getenv("HETU_GEMM_FORCE_SIMT"); force_simt = (e && e[0] == '1') ? 1 : 0; }
    if (mode != 1 && tma_ok && big && (!force_simt || act == ACT_RELU_MASK)) {
        GemmParams p;
        p.C = C; p.bias = bias; p.M = M; p.N = N; p.K = K; p.sC = sC;
        p.a_mn_major = transA ? 1 : 0;
        p.b_mn_major = transB ? 0 : 1;
        p.a_batched = (batch > 1 && sA != 0) ? 1 : 0;
        p.b_batched = (batch > 1 && sB != 0) ? 1 : 0;
        p.alpha = alpha; p.beta = beta; p.act = act;
        p.a_mn4 = p.b_mn4 = 0;
        p.ag_world = p.rs_world = 0; p.ag_rows = p.rs_rows = 1;
        { const char* e = getenv("HETU_GEMM_DBG"); p.dbg = e ? atoi(e) : 0; }
        int rc;
        static int use_2sm = -1;
        if (use_2sm < 0) { const char* e = getenv("HETU_GEMM_2SM"); use_2sm = (e && e[0] == '0') ? 0 : 1; }
        // persistent 2-SM kernel: large problems; tall problems with few output columns (conv forward: the
        // per-CTA set-up of the one-shot kernel costs more than the MMAs wasted on the empty part of the
        // 256-wide tile); and few-tile / long-K products, which it runs as stream-K
        const long long tiles_256 = (long long)((M + 255) / 256) * ((N + 255) / 256) * batch;
        static double min2 = -1.0;      // smallest product (MACs) sent to the 2-SM kernel; HETU_GEMM_2SM_MIN_MACS overrides
        if (min2 < 0) { const char* e = getenv("HETU_GEMM_2SM_MIN_MACS"); min2 = e ? atof(e) : 74.0 * 256 * 256 * 1024; }
        const bool big2 = M >= 512 && N >= 256 && (double)M * N * batch * K >= min2;
        const bool tall = M >= 8192 && N >= 64 && K >= 64;
        static int sk_policy = -1;
        if (sk_policy < 0) { const char* e = getenv("HETU_GEMM_STREAMK_POLICY"); sk_policy = (e && e[0] == '0') ? 0 : 1; }
        const bool starved = sk_policy ? ((tiles_256 * 4 <= 74 || (tiles_256 <= 44 && M >= 256 && N >= 256)) && K >= 2048 && M >= 32 &&
                                          N > kSkinnyN && beta == 0.f && act == ACT_NONE)   // e.g. conv1's weight gradient: 64 x 28 x 131072
                                       : (tiles_256 * 4 <= 74 && K >= 8192 && M >= 32 && N >= 32 && beta == 0.f && act == ACT_NONE);
        if (use_2sm && (big2 || tall || starved))
            rc = launch_tc_2sm(A, B, p, batch, sA, sB, s);
        else if (N > 128) rc = launch_tc<256>(A, B, p, batch, sA, sB, s);
        else if (N > 64) rc = launch_tc<128>(A, B, p, batch, sA, sB, s);
        else if (N > 32) rc = launch_tc<64>(A, B, p, batch, sA, sB, s);
        else rc = launch_tc<32>(A, B, p, batch, sA, sB, s);
        if (rc) return rc;
        HETU_CHECK_LAUNCH("hetu_gemm(tcgen05)");
        return 0;
    }
    if (macs <= (double)(1 << 27)) {
        launch_small(A, B, C, bias, M, N, K, transA, transB, batch, sA, sB, sC, alpha, beta, act, s);
        HETU_CHECK_LAUNCH("hetu_gemm(small)");
        return 0;
    }
    dim3 grid((N + 63) / 64, (M + 63) / 64, batch);
    gemm_fp32_simt_kernel<<<grid, 256, 0, s>>>(A, B, C, bias, M, N, K, transA, transB, sA, sB, sC, alpha, beta, act);
    HETU_CHECK_LAUNCH("hetu_gemm(simt)");
    return 0;
}

// ============================================================ tensor-parallel fused GEMMs
// reference: the model-parallel split/datatransfer ops of BASELINE.json (all-gather -> GEMM and
// GEMM -> reduce-scatter).  Both reuse the persistent 2-SM kernel; the collective is not a separate
// step but the operand fetch (TMA over peer-mapped memory) / the epilogue (NVLink reductions).
// The caller brackets each call with hetu_symm_barrier (inputs complete / outputs complete).

// C[M_total, N] = allgather_rows(A) @ op(B).  gathered[r] = rank r's [M_total, K] gathered buffer (this rank's
// own rows already in place at row rank*M_total/world), flags[r] = rank r's zeroed flag array
// [world][M_total/world/128].  One opening barrier (flags zeroed + shards in place everywhere) must precede.
HETU_API int hetu_gemm_ag(void* const* gathered, void* const* flags, int rank, int world, const float* B, float* C,
                          const float* bias, int M_total, int N, int K, int transB, float alpha, int act, cudaStream_t s) {
    if (world < 1 || world > 8 || M_total % world) { hetu_set_error("hetu_gemm_ag: bad world / M"); return -1; }
    const int rows = M_total / world;
    if (rows % 128 || K % 4 || (transB ? K : N) % 4) { hetu_set_error("hetu_gemm_ag: shard rows must be a multiple of 128"); return -1; }
    GemmParams p = {};
    p.C = C; p.bias = bias; p.M = M_total; p.N = N; p.K = K; p.sC = 0;
    p.a_mn_major = 0; p.b_mn_major = transB ? 0 : 1;
    p.alpha = alpha; p.beta = 0.f; p.act = act;
    p.ag_world = world; p.ag_rows = rows; p.ag_rank = rank; p.rs_rows = 1;
    for (int r = 0; r < world; ++r) { p.ag_buf[r] = (float*)gathered[r]; p.ag_flag[r] = (unsigned*)flags[r]; }
    if (launch_tc_2sm((const float*)gathered[rank], B, p, 1, 0, 0, s)) return -1;
    HETU_CHECK_LAUNCH("hetu_gemm_ag");
    return 0;
}

// partial = op(A) @ op(B) [M, N]; rows [r*M/world, (r+1)*M/world) are stored into slice `rank` of out_peers[r]
// (a [world][M/world][N] staging buffer on rank r).  The caller brackets the call with two barriers and sums
// its own staging slices afterwards.
HETU_API int hetu_gemm_rs(const float* A, const float* B, void* const* out_peers, int rank, int world, int M, int N, int K,
                          int transA, int transB, float alpha, cudaStream_t s) {
    if (world < 1 || world > 8 || M % world) { hetu_set_error("hetu_gemm_rs: bad world / M"); return -1; }
    if ((transA ? M : K) % 4 || (transB ? K : N) % 4 || N % 4) { hetu_set_error("hetu_gemm_rs: dims must be multiples of 4"); return -1; }
    GemmParams p = {};
    p.C = (float*)out_peers[0]; p.bias = nullptr; p.M = M; p.N = N; p.K = K; p.sC = 0;
    p.a_mn_major = transA ? 1 : 0; p.b_mn_major = transB ? 0 : 1;
    p.alpha = alpha; p.beta = 0.f; p.act = ACT_NONE;
    p.rs_world = world; p.rs_rows = M / world; p.rs_rank = rank; p.ag_rows = 1;
    for (int r = 0; r < world; ++r) p.rs_out[r] = (float*)out_peers[r];
    if (launch_tc_2sm(A, B, p, 1, 0, 0, s)) return -1;
    HETU_CHECK_LAUNCH("hetu_gemm_rs");
    return 0;
}

// ==================================================================== CSR SpMM (GNN)
// reference: src/ops/CuSparseCsrmm.cu (hand spmm_kernel / spmm_T_kernel).  One warp per CSR
// row; lanes stride over the dense columns.  trans: scatter with atomics into out[col].
__global__ void __launch_bounds__(256) csr_spmm_kernel(const float* __restrict__ val, const int* __restrict__ rowptr,
                                                       const int* __restrict__ colidx, const float* __restrict__ dense,
                                                       float* __restrict__ out, int nrow, int n, int trans) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    int nwarp = (gridDim.x * blockDim.x) >> 5;
    for (int r = warp; r < nrow; r += nwarp) {
        int beg = rowptr[r], end = rowptr[r + 1];
        if (!trans) {
            for (int c = lane; c < n; c += 32) {
                float acc = 0.f;
                for (int e = beg; e < end; ++e) acc += val[e] * dense[(size_t)colidx[e] * n + c];
                out[(size_t)r * n + c] = acc;
            }
        } else {
            for (int e = beg; e < end; ++e) {
                float v = val[e];
                int col = colidx[e];
                for (int c = lane; c < n; c += 32) atomicAdd(out + (size_t)col * n + c, v * dense[(size_t)r * n + c]);
            }
        }
    }
}
HETU_API int hetu_csr_spmm(const float* val, const int* rowptr, const int* colidx, const float* dense, float* out,
                           int nrow, int n, int trans, cudaStream_t s) {
    if (nrow == 0 || n == 0) return 0;
    csr_spmm_kernel<<<ew_grid((size_t)nrow * 32, 256, 16), 256, 0, s>>>(val, rowptr, colidx, dense, out, nrow, n, trans);
    HETU_CHECK_LAUNCH("hetu_csr_spmm");
    return 0;
}
