// Fused MoE gate: softmax + top-k + capacity-slot assignment + GShard balance loss, and its backward.
//
// reference: the gate is ~30 small ops there - cudnnSoftmax, TopKIdx.cu (k serial scans by ONE thread
// per row), OneHot.cu, CumSum.cu (one thread per column walking all tokens), ReduceSum / ReduceMean,
// element-wise products (python/hetu/layers/TopGate.py:15-79).  Here the forward is two launches and
// the backward two.
//
// Packed output P [T + 1, 3k] (float): columns [0,k) expert index of route i, [k,2k) slot inside the
// expert's capacity buffer, [2k,3k) gate value softmax(logits)[t, idx_i]; P[T][0] = balance loss
// E * sum_i sum_e mean_t(softmax)[e] * mean_t(route_i == e).  Slots follow token order: route 0 of
// every token comes first, route i starts after all routes < i of the same expert.
#include "common.cuh"

constexpr int GATE_TOK_PER_BLOCK = 256;      // 8 warps x 32 tokens each
constexpr int GATE_MAX_K = 4;

// one warp per token (tokens of a block are consecutive): softmax statistics, top-k, gate values;
// per-block per-route expert counts -> blk_cnt[k][nblk][E]; sum of probabilities -> me[E]
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
    const int W = 3 * k;
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
                P[(size_t)t * W + i] = (float)bidx;
                P[(size_t)t * W + 2 * k + i] = __expf(best - mx) * inv;
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
// rank among the earlier tokens of this block.  Block 0 also writes the balance loss.
__global__ void moe_gate_slot_kernel(float* __restrict__ P, const int* __restrict__ blk_cnt, const float* __restrict__ me,
                                     int T, int E, int k) {
    extern __shared__ int smi[];
    int* base = smi;                              // [k][E]
    int* idx_s = smi + k * E;                     // [k][GATE_TOK_PER_BLOCK]
    const int nblk = gridDim.x, W = 3 * k;
    for (int j = threadIdx.x; j < k * E; j += blockDim.x) {
        int i = j / E, e = j - i * E;
        int b = 0;
        for (int q = 0; q < i; ++q)
            for (int bb = 0; bb < nblk; ++bb) b += blk_cnt[((size_t)q * nblk + bb) * E + e];      // all routes of lower rank
        for (int bb = 0; bb < blockIdx.x; ++bb) b += blk_cnt[((size_t)i * nblk + bb) * E + e];    // earlier blocks, this rank
        base[j] = b;
    }
    const int t = blockIdx.x * GATE_TOK_PER_BLOCK + threadIdx.x;
    for (int i = 0; i < k; ++i) idx_s[i * GATE_TOK_PER_BLOCK + threadIdx.x] = t < T ? (int)P[(size_t)t * W + i] : -1;
    __syncthreads();
    if (t < T) {
        for (int i = 0; i < k; ++i) {
            const int e = idx_s[i * GATE_TOK_PER_BLOCK + threadIdx.x];
            int rank = 0;
            for (int j = 0; j < (int)threadIdx.x; ++j) rank += (idx_s[i * GATE_TOK_PER_BLOCK + j] == e);
            P[(size_t)t * W + k + i] = (float)(base[i * E + e] + rank);
        }
    }
    if (blockIdx.x == 0) {                        // balance loss (and zero the rest of the extra row)
        __shared__ float red[32];
        float acc = 0.f;
        for (int j = threadIdx.x; j < k * E; j += blockDim.x) {
            int i = j / E, e = j - i * E;
            int c = 0;
            for (int bb = 0; bb < nblk; ++bb) c += blk_cnt[((size_t)i * nblk + bb) * E + e];
            acc += me[e] * (float)c;
        }
        acc = block_sum(acc, red);
        if (threadIdx.x == 0) P[(size_t)T * W] = acc * (float)E / ((float)T * (float)T);
        for (int j = threadIdx.x + 1; j < W; j += blockDim.x) P[(size_t)T * W + j] = 0.f;
    }
}

// cnt[e] = number of routes (all ranks) that chose expert e, from the index columns of P
__global__ void moe_gate_count_kernel(const float* __restrict__ P, float* __restrict__ cnt, int T, int k) {
    const int W = 3 * k;
    for (long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x; j < (long long)T * k; j += (long long)gridDim.x * blockDim.x) {
        long long t = j / k;
        int i = (int)(j - t * k);
        atomicAdd(&cnt[(int)P[t * W + i]], 1.f);
    }
}

// dlogits[t, j] = sum_i dgate_i * p_sel_i * (delta(j, idx_i) - p_j)  +  dl * sum_e c_e p_e (delta(e, j) - p_j)
// with c_e = E / T^2 * cnt[e];  one warp per token
__global__ void moe_gate_bwd_kernel(const float* __restrict__ logits, const float* __restrict__ P, const float* __restrict__ dP,
                                    const float* __restrict__ cnt, float* __restrict__ dlogits, int T, int E, int k) {
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const int W = 3 * k;
    const float dl = dP[(size_t)T * W];
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
        float gsel = 0.f;                                    // sum_i dgate_i * p_sel_i
        for (int i = 0; i < k; ++i) gsel += dP[(size_t)t * W + 2 * k + i] * P[(size_t)t * W + 2 * k + i];
        for (int j = lane; j < E; j += 32) {
            float p = __expf(row[j] - mx) * inv;
            float g = -p * (gsel + cp) + cscale * cnt[j] * p;
            for (int i = 0; i < k; ++i)
                if ((int)P[(size_t)t * W + i] == j) g += dP[(size_t)t * W + 2 * k + i] * p;
            dlogits[(size_t)t * E + j] = g;
        }
    }
}

// blk_cnt: int [k * nblk * E] scratch, me: float [E] scratch (zeroed here), P: [T + 1, 3k]
HETU_API int hetu_moe_gate_fwd(const float* logits, float* P, int* blk_cnt, float* me, int T, int E, int k, cudaStream_t s) {
    if (k < 1 || k > GATE_MAX_K || k > E) { hetu_set_error("moe_gate: k must be in [1, %d] and <= E", GATE_MAX_K); return -1; }
    if (T <= 0) return 0;
    const int nblk = (T + GATE_TOK_PER_BLOCK - 1) / GATE_TOK_PER_BLOCK;
    size_t sm1 = (size_t)E * (k + 1) * sizeof(float);
    size_t sm2 = ((size_t)k * E + (size_t)k * GATE_TOK_PER_BLOCK) * sizeof(int);
    if (sm1 > 48 * 1024 || sm2 > 48 * 1024) { hetu_set_error("moe_gate: too many experts (%d) for the shared-memory tables", E); return -1; }
    HETU_CUDA(cudaMemsetAsync(me, 0, (size_t)E * sizeof(float), s));
    moe_gate_topk_kernel<<<nblk, 256, sm1, s>>>(logits, P, blk_cnt, me, T, E, k);
    HETU_CHECK_LAUNCH("moe_gate_topk");
    moe_gate_slot_kernel<<<nblk, GATE_TOK_PER_BLOCK, sm2, s>>>(P, blk_cnt, me, T, E, k);
    HETU_CHECK_LAUNCH("moe_gate_slot");
    return 0;
}

// cnt: float [E] scratch (zeroed here)
HETU_API int hetu_moe_gate_bwd(const float* logits, const float* P, const float* dP, float* cnt, float* dlogits, int T, int E, int k,
                               cudaStream_t s) {
    if (T <= 0) return 0;
    HETU_CUDA(cudaMemsetAsync(cnt, 0, (size_t)E * sizeof(float), s));
    moe_gate_count_kernel<<<ew_grid((size_t)T * k, 256), 256, 0, s>>>(P, cnt, T, k);
    HETU_CHECK_LAUNCH("moe_gate_count");
    long long blocks = ((long long)T * 32 + 255) / 256;
    long long cap = (long long)kNumSMs * 8;
    moe_gate_bwd_kernel<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, s>>>(logits, P, dP, cnt, dlogits, T, E, k);
    HETU_CHECK_LAUNCH("moe_gate_bwd");
    return 0;
}
