// Fused MoE gate: softmax + top-k + capacity-slot assignment + GShard balance loss, and its backward.
//
// reference: the gate is ~30 small ops there - cudnnSoftmax, TopKIdx.cu (k serial scans by ONE thread
// per row), OneHot.cu, CumSum.cu (one thread per column walking all tokens), ReduceSum / ReduceMean,
// element-wise products (python/hetu/layers/TopGate.py:15-79).  Here the forward is two launches and
// the backward one.
//
// Packed output P, planar float [3k * T + 1 + E]: plane i in [0,k) = expert index of route i for every token,
// plane k+i = slot inside the expert's capacity buffer, plane 2k+i = gate value softmax(logits)[t, idx_i];
// P[3kT] = balance loss E * sum_i sum_e mean_t(softmax)[e] * mean_t(route_i == e); P[3kT+1 .. 3kT+E] = number of
// routes per expert (the backward needs it).  Every plane is contiguous, so the graph hands the pieces on as views.
// Slots follow token order: route 0 of every token comes first, route i starts after all routes < i of the same expert.
#include "common.cuh"

constexpr int GATE_TOK_PER_BLOCK = 256;
constexpr int GATE_MAX_K = 4;
constexpr int GATE_SMALL_E = 32;             // up to here one THREAD handles a token (row in registers)

// counts one (route, expert) hit per lane into a shared table with one atomic per distinct expert in the warp
__device__ __forceinline__ void warp_count(int* table, int e, bool on) {
    unsigned act = __ballot_sync(0xffffffffu, on);
    if (on) {
        unsigned same = __match_any_sync(act, e);
        if ((int)(__ffs(same) - 1) == (int)(threadIdx.x & 31)) atomicAdd(&table[e], __popc(same));
    }
}

// E <= 32: one thread per token, 256 consecutive tokens per block
template <int EMAX>
__global__ void __launch_bounds__(GATE_TOK_PER_BLOCK) moe_gate_topk_small_kernel(const float* __restrict__ logits, float* __restrict__ P,
                                                                                 int* __restrict__ blk_cnt, float* __restrict__ me,
                                                                                 int T, int E, int k) {
    extern __shared__ float sm[];
    float* me_blk = sm;                          // [E]
    int* cnt_blk = (int*)(sm + E);               // [k][E]
    for (int j = threadIdx.x; j < E * (k + 1); j += blockDim.x) sm[j] = 0.f;
    __syncthreads();
    const int t = blockIdx.x * GATE_TOK_PER_BLOCK + threadIdx.x;
    const bool on = t < T;
    float v[EMAX];
    float mx = -INFINITY;
#pragma unroll
    for (int j = 0; j < EMAX; ++j) {
        v[j] = (on && j < E) ? logits[(size_t)t * E + j] : -INFINITY;
        mx = fmaxf(mx, v[j]);
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < EMAX; ++j) s += (j < E && on) ? __expf(v[j] - mx) : 0.f;
    const float inv = on ? 1.f / s : 0.f;
#pragma unroll
    for (int j = 0; j < EMAX; ++j) {
        if (j < E) {                                                     // uniform branch
            float p = on ? __expf(v[j] - mx) * inv : 0.f;
            p = warp_sum(p);
            if ((threadIdx.x & 31) == 0) atomicAdd(&me_blk[j], p);
        }
    }
    unsigned taken = 0;
    for (int i = 0; i < k; ++i) {
        float best = -INFINITY;
        int bidx = 0;
#pragma unroll
        for (int j = 0; j < EMAX; ++j)
            if (j < E && !((taken >> j) & 1u) && v[j] > best) { best = v[j]; bidx = j; }          // first maximum wins ties
        if (best == -INFINITY)                                                                     // all remaining are -inf
            for (int j = E - 1; j >= 0; --j) if (!((taken >> j) & 1u)) bidx = j;
        taken |= 1u << bidx;
        if (on) {
            P[(size_t)i * T + t] = (float)bidx;
            P[(size_t)(2 * k + i) * T + t] = __expf(best - mx) * inv;
        }
        warp_count(cnt_blk + i * E, bidx, on);
    }
    __syncthreads();
    for (int j = threadIdx.x; j < E; j += blockDim.x) atomicAdd(&me[j], me_blk[j]);
    for (int j = threadIdx.x; j < k * E; j += blockDim.x) {
        int i = j / E, e = j - i * E;
        blk_cnt[((size_t)i * gridDim.x + blockIdx.x) * E + e] = cnt_blk[j];
    }
}

// any E: one warp per token (tokens of a block are consecutive)
__global__ void moe_gate_topk_kernel(const float* __restrict__ logits, float* __restrict__ P, int* __restrict__ blk_cnt,
                                     float* __restrict__ me, int T, int E, int k) {
    extern __shared__ float sm[];
    float* me_blk = sm;                          // [E]
    int* cnt_blk = (int*)(sm + E);               // [k][E]
    for (int j = threadIdx.x; j < E * (k + 1); j += blockDim.x) sm[j] = 0.f;      // zero bits for both views
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    const int t0 = blockIdx.x * GATE_TOK_PER_BLOCK;
    const int t1 = min(T, t0 + GATE_TOK_PER_BLOCK);
    for (int t = t0 + warp; t < t1; t += nwarp) {
        const float* row = logits + (size_t)t * E;
        float mx = -INFINITY;
        for (int j = lane; j < E; j += 32) mx = fmaxf(mx, row[j]);
        mx = warp_max(mx);
        float s = 0.f;
        for (int j = lane; j < E; j += 32) s += __expf(row[j] - mx);
        s = warp_sum(s);
        const float inv = 1.f / s;
        for (int j = lane; j < E; j += 32) atomicAdd(&me_blk[j], __expf(row[j] - mx) * inv);
        int chosen[GATE_MAX_K];
        for (int i = 0; i < k; ++i) {
            float best = -INFINITY;
            int bidx = 0x7fffffff;
            for (int j = lane; j < E; j += 32) {
                bool taken = false;
                for (int q = 0; q < i; ++q) taken |= (chosen[q] == j);
                float v = row[j];
                if (!taken && (v > best || (v == best && j < bidx))) { best = v; bidx = j; }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                float ob = __shfl_xor_sync(0xffffffffu, best, o);
                int oi = __shfl_xor_sync(0xffffffffu, bidx, o);
                if (ob > best || (ob == best && oi < bidx)) { best = ob; bidx = oi; }
            }
            chosen[i] = bidx;
            if (lane == 0) {
                P[(size_t)i * T + t] = (float)bidx;
                P[(size_t)(2 * k + i) * T + t] = __expf(best - mx) * inv;
                atomicAdd(&cnt_blk[i * E + bidx], 1);
            }
        }
    }
    __syncthreads();
    for (int j = threadIdx.x; j < E; j += blockDim.x) atomicAdd(&me[j], me_blk[j]);
    for (int j = threadIdx.x; j < k * E; j += blockDim.x) {
        int i = j / E, e = j - i * E;
        blk_cnt[((size_t)i * gridDim.x + blockIdx.x) * E + e] = cnt_blk[j];
    }
}

// slot of every route: (tokens of earlier blocks + routes of lower rank) with the same expert, plus the
// rank among the earlier tokens of this block.  Block 0 also writes the balance loss and the per-expert totals.
__global__ void moe_gate_slot_kernel(float* __restrict__ P, const int* __restrict__ blk_cnt, const float* __restrict__ me,
                                     int T, int E, int k) {
    extern __shared__ int smi[];
    int* base = smi;                              // [k][E]
    int* idx_s = smi + k * E;                     // [k][GATE_TOK_PER_BLOCK]
    const int nblk = gridDim.x;
    for (int j = threadIdx.x; j < k * E; j += blockDim.x) {
        int i = j / E, e = j - i * E;
        int b = 0;
        for (int q = 0; q < i; ++q)
            for (int bb = 0; bb < nblk; ++bb) b += blk_cnt[((size_t)q * nblk + bb) * E + e];      // all routes of lower rank
        for (int bb = 0; bb < blockIdx.x; ++bb) b += blk_cnt[((size_t)i * nblk + bb) * E + e];    // earlier blocks, this rank
        base[j] = b;
    }
    const int t = blockIdx.x * GATE_TOK_PER_BLOCK + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = 0; i < k; ++i) idx_s[i * GATE_TOK_PER_BLOCK + threadIdx.x] = t < T ? (int)P[(size_t)i * T + t] : -1;
    __syncthreads();
    for (int i = 0; i < k; ++i) {
        const int e = idx_s[i * GATE_TOK_PER_BLOCK + threadIdx.x];
        // rank among the earlier tokens of the block: whole earlier warps by a strided scan, own warp by match
        int rank = 0;
        for (int j = 0; j < warp * 32; ++j) rank += (idx_s[i * GATE_TOK_PER_BLOCK + j] == e);
        unsigned same = __match_any_sync(0xffffffffu, e);
        rank += __popc(same & ((1u << lane) - 1u));
        if (t < T) P[(size_t)(k + i) * T + t] = (float)(base[i * E + e] + rank);
    }
    if (blockIdx.x == 0) {                        // balance loss and the per-expert route counts
        __shared__ float red[32];
        float acc = 0.f;
        for (int j = threadIdx.x; j < k * E; j += blockDim.x) {
            int i = j / E, e = j - i * E;
            int c = 0;
            for (int bb = 0; bb < nblk; ++bb) c += blk_cnt[((size_t)i * nblk + bb) * E + e];
            acc += me[e] * (float)c;
        }
        acc = block_sum(acc, red);
        float* tail = P + (size_t)3 * k * T;
        if (threadIdx.x == 0) tail[0] = acc * (float)E / ((float)T * (float)T);
        for (int e = threadIdx.x; e < E; e += blockDim.x) {
            int c = 0;
            for (int i = 0; i < k; ++i)
                for (int bb = 0; bb < nblk; ++bb) c += blk_cnt[((size_t)i * nblk + bb) * E + e];
            tail[1 + e] = (float)c;
        }
    }
}

struct GateGrads { const float* dgate[GATE_MAX_K]; const float* daux; };

// dlogits[t, j] = sum_i dgate_i * p_sel_i * (delta(j, idx_i) - p_j)  +  dl * sum_e c_e p_e (delta(e, j) - p_j)
// with c_e = E / T^2 * cnt[e];  one warp per token.  Missing gradients (null pointers) count as zero.
__global__ void moe_gate_bwd_kernel(const float* __restrict__ logits, const float* __restrict__ P, GateGrads gg,
                                    float* __restrict__ dlogits, int T, int E, int k) {
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const float* cnt = P + (size_t)3 * k * T + 1;
    const float dl = gg.daux ? gg.daux[0] : 0.f;
    const float cscale = dl * (float)E / ((float)T * (float)T);
    for (long long t = warp; t < T; t += nwarps) {
        const float* row = logits + (size_t)t * E;
        float mx = -INFINITY;
        for (int j = lane; j < E; j += 32) mx = fmaxf(mx, row[j]);
        mx = warp_max(mx);
        float s = 0.f, dot = 0.f;
        for (int j = lane; j < E; j += 32) { float ex = __expf(row[j] - mx); s += ex; dot += ex * cnt[j]; }
        s = warp_sum(s);
        dot = warp_sum(dot);
        const float inv = 1.f / s;
        const float cp = cscale * dot * inv;                 // sum_e c_e p_e
        float dg[GATE_MAX_K];
        int sel[GATE_MAX_K];
        float gsel = 0.f;                                    // sum_i dgate_i * p_sel_i
#pragma unroll
        for (int i = 0; i < GATE_MAX_K; ++i) {
            dg[i] = 0.f; sel[i] = -1;
            if (i < k) {
                dg[i] = gg.dgate[i] ? gg.dgate[i][t] : 0.f;
                sel[i] = (int)P[(size_t)i * T + t];
                gsel += dg[i] * P[(size_t)(2 * k + i) * T + t];
            }
        }
        for (int j = lane; j < E; j += 32) {
            float p = __expf(row[j] - mx) * inv;
            float g = -p * (gsel + cp) + cscale * cnt[j] * p;
#pragma unroll
            for (int i = 0; i < GATE_MAX_K; ++i)
                if (sel[i] == j) g += dg[i] * p;
            dlogits[(size_t)t * E + j] = g;
        }
    }
}

// blk_cnt: int [k * nblk * E] scratch, me: float [E] scratch (zeroed here), P: planar [3kT + 1 + E]
HETU_API int hetu_moe_gate_fwd(const float* logits, float* P, int* blk_cnt, float* me, int T, int E, int k, cudaStream_t s) {
    if (k < 1 || k > GATE_MAX_K || k > E) { hetu_set_error("moe_gate: k must be in [1, %d] and <= E", GATE_MAX_K); return -1; }
    if (T <= 0) return 0;
    const int nblk = (T + GATE_TOK_PER_BLOCK - 1) / GATE_TOK_PER_BLOCK;
    size_t sm1 = (size_t)E * (k + 1) * sizeof(float);
    size_t sm2 = ((size_t)k * E + (size_t)k * GATE_TOK_PER_BLOCK) * sizeof(int);
    if (sm1 > 48 * 1024 || sm2 > 48 * 1024) { hetu_set_error("moe_gate: too many experts (%d) for the shared-memory tables", E); return -1; }
    HETU_CUDA(cudaMemsetAsync(me, 0, (size_t)E * sizeof(float), s));
    if (E <= 8) moe_gate_topk_small_kernel<8><<<nblk, GATE_TOK_PER_BLOCK, sm1, s>>>(logits, P, blk_cnt, me, T, E, k);
    else if (E <= 16) moe_gate_topk_small_kernel<16><<<nblk, GATE_TOK_PER_BLOCK, sm1, s>>>(logits, P, blk_cnt, me, T, E, k);
    else if (E <= GATE_SMALL_E) moe_gate_topk_small_kernel<32><<<nblk, GATE_TOK_PER_BLOCK, sm1, s>>>(logits, P, blk_cnt, me, T, E, k);
    else moe_gate_topk_kernel<<<nblk, 256, sm1, s>>>(logits, P, blk_cnt, me, T, E, k);
    HETU_CHECK_LAUNCH("moe_gate_topk");
    moe_gate_slot_kernel<<<nblk, GATE_TOK_PER_BLOCK, sm2, s>>>(P, blk_cnt, me, T, E, k);
    HETU_CHECK_LAUNCH("moe_gate_slot");
    return 0;
}

// dgates: k device pointers ([T] each, may be null), daux: [1] or null
HETU_API int hetu_moe_gate_bwd(const float* logits, const float* P, void* const* dgates, const float* daux, float* dlogits,
                               int T, int E, int k, cudaStream_t s) {
    if (T <= 0) return 0;
    if (k < 1 || k > GATE_MAX_K) { hetu_set_error("moe_gate_bwd: k must be in [1, %d]", GATE_MAX_K); return -1; }
    GateGrads gg;
    for (int i = 0; i < GATE_MAX_K; ++i) gg.dgate[i] = i < k ? (const float*)dgates[i] : nullptr;
    gg.daux = daux;
    long long blocks = ((long long)T * 32 + 255) / 256;
    long long cap = (long long)kNumSMs * 8;
    moe_gate_bwd_kernel<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, s>>>(logits, P, gg, dlogits, T, E, k);
    HETU_CHECK_LAUNCH("moe_gate_bwd");
    return 0;
}
