// In-kernel NVLink collectives over a CUDA-IPC symmetric heap.
//
// reference: src/communication/mpi_nccl_communication.cu:137-277 - every collective is a
// host-launched NCCL call on a side stream (ncclAllReduce per gradient tensor, grouped
// ncclSend/ncclRecv for all-to-all), followed by separate optimizer launches
// (python/hetu/optimizer.py:184-196).
//
// Here each rank maps every peer's buffer (cudaIpcOpenMemHandle) and the kernels read / write
// peer HBM directly with ld/st.global over NVLink 5 / NVSwitch.  Ordering between GPUs uses one
// monotonically increasing epoch per rank and system-scope release/acquire flags, so a
// collective is ONE kernel (plus a pack kernel where the input is scattered), allocates
// nothing, and can be captured in a CUDA graph.  The gradient all-reduce is fused with the
// gradient scale and the optimizer update.  Each collective kernel opens
// with a barrier (inputs ready / buffers free) and closes with one (reads done / stores landed).
#include "common.cuh"
#include <cstring>

constexpr int kMaxPeers = 16;
struct PeerPtrs { float* p[kMaxPeers]; };
struct FlagPtrs { unsigned* f[kMaxPeers]; };

// ------------------------------------------------------------------ symmetric heap (host)
HETU_API int hetu_symm_alloc(size_t bytes, void** ptr, void* handle64) {
    HETU_CUDA(cudaMalloc(ptr, bytes));
    HETU_CUDA(cudaMemset(*ptr, 0, bytes));
    cudaIpcMemHandle_t h;
    HETU_CUDA(cudaIpcGetMemHandle(&h, *ptr));
    memcpy(handle64, &h, sizeof(h));
    return 0;
}
HETU_API int hetu_symm_open(const void* handle64, void** ptr) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, sizeof(h));
    HETU_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}
HETU_API int hetu_symm_close(void* ptr) { HETU_CUDA(cudaIpcCloseMemHandle(ptr)); return 0; }
HETU_API int hetu_symm_free(void* ptr) { HETU_CUDA(cudaFree(ptr)); return 0; }

// ------------------------------------------------------------------ device-side primitives
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float4 ld_sys_f4(const float* p) {
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float ld_sys_f1(const float* p) {
    float v;
    asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}

// All ranks call this the same number of times.  Thread p of the calling CTA publishes the new
// epoch into peer p's flag slot [rank] and waits until its own slot [p] reaches it.
__device__ __forceinline__ unsigned cross_gpu_barrier(FlagPtrs flags, int rank, int world, unsigned* epoch) {
    __shared__ unsigned s_epoch;
    if (threadIdx.x == 0) {
        __threadfence_system();
        s_epoch = *epoch + 1;
        *epoch = s_epoch;
    }
    __syncthreads();
    unsigned e = s_epoch;
    if ((int)threadIdx.x < world) {
        st_release_sys(flags.f[threadIdx.x] + rank, e);
        while ((int)(ld_acquire_sys(flags.f[rank] + threadIdx.x) - e) < 0) { __nanosleep(40); }
    }
    __syncthreads();
    return e;
}

__global__ void __launch_bounds__(32) symm_barrier_kernel(FlagPtrs flags, int rank, int world, unsigned* epoch) {
    cross_gpu_barrier(flags, rank, world, epoch);
}

static int fill_flags(FlagPtrs& fp, void* const* flag_ptrs, int world) {
    if (world > kMaxPeers) { hetu_set_error("symm: world %d > %d", world, kMaxPeers); return -1; }
    for (int i = 0; i < kMaxPeers; ++i) fp.f[i] = i < world ? (unsigned*)flag_ptrs[i] : nullptr;
    return 0;
}
static void fill_peers(PeerPtrs& pp, void* const* ptrs, int world, size_t byte_off) {
    for (int i = 0; i < kMaxPeers; ++i) pp.p[i] = i < world ? (float*)((char*)ptrs[i] + byte_off) : nullptr;
}

HETU_API int hetu_symm_barrier(void* const* flag_ptrs, int rank, int world, void* epoch, cudaStream_t s) {
    FlagPtrs fp;
    if (fill_flags(fp, flag_ptrs, world)) return -1;
    symm_barrier_kernel<<<1, 32, 0, s>>>(fp, rank, world, (unsigned*)epoch);
    HETU_CHECK_LAUNCH("hetu_symm_barrier");
    return 0;
}

// Opening barrier of the multi-CTA collectives: "every rank's send data is in place".  It runs as the 1-CTA
// symm_barrier_kernel queued right before the collective on the same stream (open_collective below), so no CTA of the
// collective ever spins on another CTA of its own grid - the grid makes progress whatever else occupies the SMs.
// Every CTA calls this last.  The final CTA to leave runs a second cross-GPU barrier - "every
// rank has finished reading / its stores have landed" - so the buffers may be reused by the next
// collective without double buffering (CUDA-graph replays bake buffer addresses in), then
// commits the gate for the next collective kernel.
__device__ __forceinline__ void gate_exit(FlagPtrs flags, int rank, int world, unsigned* epoch, unsigned* gate) {
    __shared__ bool s_last;
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) s_last = (atomicAdd(gate + 2, 1u) == gridDim.x * gridDim.y - 1);
    __syncthreads();
    if (s_last) {
        cross_gpu_barrier(flags, rank, world, epoch);
        if (threadIdx.x == 0) {
            gate[2] = 0;
            gate[1] = ((volatile unsigned*)gate)[0];
            __threadfence();
        }
    }
}

// -------------------------------------------------------------- pack (scatter -> flat arena)
constexpr int kMaxTensors = 48;
struct PackList { const float* src[kMaxTensors]; long long off[kMaxTensors]; long long n[kMaxTensors]; };

__global__ void __launch_bounds__(256) pack_kernel(PackList pl, float* __restrict__ dst, float scale) {
    int t = blockIdx.y;
    long long n = pl.n[t];
    const float* __restrict__ src = pl.src[t];
    float* __restrict__ d = dst + pl.off[t];
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        d[i] = src[i] * scale;
}
HETU_API int hetu_pack_tensors(int ntensors, void* const* srcs, const long long* offsets, const long long* sizes,
                               float* dst, float scale, cudaStream_t s) {
    for (int base = 0; base < ntensors; base += kMaxTensors) {
        int cnt = ntensors - base < kMaxTensors ? ntensors - base : kMaxTensors;
        PackList pl;
        long long maxn = 0;
        for (int i = 0; i < kMaxTensors; ++i) {
            bool on = i < cnt;
            pl.src[i] = on ? (const float*)srcs[base + i] : nullptr;
            pl.off[i] = on ? offsets[base + i] : 0;
            pl.n[i] = on ? sizes[base + i] : 0;
            if (on && sizes[base + i] > maxn) maxn = sizes[base + i];
        }
        int gx = (int)((maxn + 1023) / 1024);
        int cap = (kNumSMs * 8 + cnt - 1) / cnt;
        if (gx > cap) gx = cap;
        if (gx < 1) gx = 1;
        pack_kernel<<<dim3(gx, cnt), 256, 0, s>>>(pl, dst, scale);
    }
    HETU_CHECK_LAUNCH("hetu_pack_tensors");
    return 0;
}

// ----------------------------------------------------- one-shot all-reduce (+ fused optimizer)
// Every rank has packed its contribution into its own arena (parity slot chosen by the host).
// After the barrier each rank sums the `world` arenas element-wise straight out of peer HBM.
// MODE 0: write the sum to `out`.  MODE 1: apply the optimizer step to the parameter arena
// described by `tl` (flat offsets into the arena) - every rank performs the identical update,
// so parameters stay bit-identical without a second exchange.
enum { OPT_SGD = 0, OPT_MOMENTUM, OPT_NESTEROV, OPT_ADAGRAD, OPT_ADAM, OPT_ADAMW };
struct ParamList {
    float* p[kMaxTensors]; float* s1[kMaxTensors]; float* s2[kMaxTensors];
    long long off[kMaxTensors]; long long n[kMaxTensors]; int count;
};

__device__ __forceinline__ float opt_apply(int kind, float pv, float gv, float* s1, float* s2, long long i, float lr,
                                           float h0, float h1, float h2, float h3, float b1t, float b2t) {
    if (kind == OPT_SGD) return pv - lr * gv;
    if (kind == OPT_MOMENTUM) { float v = h0 * s1[i] - lr * gv; s1[i] = v; return pv + v; }
    if (kind == OPT_NESTEROV) { float tmp = lr * gv; float v = h0 * (s1[i] - tmp); s1[i] = v; return pv + v - tmp; }
    if (kind == OPT_ADAGRAD) { float a = s1[i] + gv * gv; s1[i] = a; return pv - lr * gv / (sqrtf(a) + h0); }
    float m = h0 * s1[i] + (1.f - h0) * gv;
    float v = h1 * s2[i] + (1.f - h1) * gv * gv;
    s1[i] = m; s2[i] = v;
    float upd = (m / b1t) / (sqrtf(v / b2t) + h2);
    return kind == OPT_ADAM ? pv - lr * upd : pv - lr * (upd + h3 * pv);
}

__global__ void __launch_bounds__(256) symm_allreduce_kernel(PeerPtrs arenas, FlagPtrs flags, int rank, int world,
                                                             unsigned* epoch, unsigned* gate, long long n,
                                                             float* __restrict__ out) {
    long long n4 = n >> 2;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int p = 0; p < world; ++p) {
            float4 v = ld_sys_f4(arenas.p[p] + (i << 2));
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
        reinterpret_cast<float4*>(out)[i] = acc;
    }
    for (long long i = (n4 << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        float acc = 0.f;
        for (int p = 0; p < world; ++p) acc += ld_sys_f1(arenas.p[p] + i);
        out[i] = acc;
    }
    gate_exit(flags, rank, world, epoch, gate);
}

__global__ void __launch_bounds__(256) symm_allreduce_opt_kernel(PeerPtrs arenas, FlagPtrs flags, int rank, int world,
                                                                 unsigned* epoch, unsigned* gate, ParamList tl, int kind,
                                                                 const float* __restrict__ lr_p,
                                                                 const long long* __restrict__ step_p, float h0, float h1,
                                                                 float h2, float h3, float l2reg) {
    float lr = *lr_p;
    float b1t = 1.f, b2t = 1.f;
    if (kind >= OPT_ADAM) {
        float st = (float)(*step_p);
        b1t = 1.f - powf(h0, st);
        b2t = 1.f - powf(h1, st);
    }
    for (int t = 0; t < tl.count; ++t) {
        long long n = tl.n[t], off = tl.off[t];
        float* __restrict__ p = tl.p[t];
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
            float g = 0.f;
            for (int r = 0; r < world; ++r) g += ld_sys_f1(arenas.p[r] + off + i);
            float pv = p[i];
            p[i] = opt_apply(kind, pv, g + l2reg * pv, tl.s1[t], tl.s2[t], i, lr, h0, h1, h2, h3, b1t, b2t);
        }
    }
    gate_exit(flags, rank, world, epoch, gate);
}

static int coop_grid(long long work_items) {
    long long need = (work_items + 255) / 256;
    long long cap = kNumSMs * 2;
    return (int)(need < 1 ? 1 : (need < cap ? need : cap));
}

HETU_API int hetu_symm_allreduce(void* const* arena_ptrs, size_t byte_off, void* const* flag_ptrs, int rank, int world,
                                 void* epoch, void* gate, long long n, float* out, cudaStream_t s) {
    FlagPtrs fp;
    if (fill_flags(fp, flag_ptrs, world)) return -1;
    PeerPtrs pp;
    fill_peers(pp, arena_ptrs, world, byte_off);
    symm_barrier_kernel<<<1, 32, 0, s>>>(fp, rank, world, (unsigned*)epoch);   // opening barrier (see above)
    symm_allreduce_kernel<<<coop_grid((n + 3) / 4), 256, 0, s>>>(pp, fp, rank, world, (unsigned*)epoch, (unsigned*)gate, n, out);
    HETU_CHECK_LAUNCH("hetu_symm_allreduce");
    return 0;
}

HETU_API int hetu_symm_allreduce_opt(void* const* arena_ptrs, size_t byte_off, void* const* flag_ptrs, int rank,
                                     int world, void* epoch, void* gate, int kind, int ntensors, void* const* params,
                                     void* const* s1, void* const* s2, const long long* offsets, const long long* sizes,
                                     const float* lr_p, const void* step_p, float h0, float h1, float h2, float h3,
                                     float l2reg, cudaStream_t s) {
    if (ntensors > kMaxTensors) { hetu_set_error("symm_allreduce_opt: %d tensors > %d", ntensors, kMaxTensors); return -1; }
    if (kind > OPT_ADAMW) { hetu_set_error("symm_allreduce_opt: optimizer kind %d has no fused path", kind); return -1; }
    FlagPtrs fp;
    if (fill_flags(fp, flag_ptrs, world)) return -1;
    PeerPtrs pp;
    fill_peers(pp, arena_ptrs, world, byte_off);
    ParamList tl;
    tl.count = ntensors;
    long long maxn = 0;
    for (int i = 0; i < kMaxTensors; ++i) {
        bool on = i < ntensors;
        tl.p[i] = on ? (float*)params[i] : nullptr;
        tl.s1[i] = on ? (float*)s1[i] : nullptr;
        tl.s2[i] = on ? (float*)s2[i] : nullptr;
        tl.off[i] = on ? offsets[i] : 0;
        tl.n[i] = on ? sizes[i] : 0;
        if (on && sizes[i] > maxn) maxn = sizes[i];
    }
    long long total = 0;
    for (int i = 0; i < ntensors; ++i) total += sizes[i];
    symm_barrier_kernel<<<1, 32, 0, s>>>(fp, rank, world, (unsigned*)epoch);   // opening barrier (see above)
    symm_allreduce_opt_kernel<<<coop_grid(total), 256, 0, s>>>(pp, fp, rank, world, (unsigned*)epoch, (unsigned*)gate, tl,
                                                                kind, lr_p, (const long long*)step_p, h0, h1, h2, h3, l2reg);
    HETU_CHECK_LAUNCH("hetu_symm_allreduce_opt");
    return 0;
}

// ------------------------------------------------------- all-to-all / all-gather (peer stores)
// send: local [world][chunk] (all-to-all) or local [chunk] (all-gather).  Each rank stores chunk
// p into peer p's receive buffer at slot [rank]; the last CTA to finish its stores fences,
// publishes the epoch to every peer and waits for theirs, so when the kernel completes the local
// receive buffer holds every peer's chunk.
__global__ void __launch_bounds__(256) symm_exchange_kernel(const float* __restrict__ send, PeerPtrs recv, FlagPtrs flags,
                                                            int rank, int world, unsigned* epoch, unsigned* gate,
                                                            long long chunk, int same_chunk) {
    // barrier 1: every peer has consumed the previous contents of its receive buffer
    long long chunk4 = chunk >> 2;
    long long total4 = chunk4 * world;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += (long long)gridDim.x * blockDim.x) {
        int p = (int)(i / chunk4);
        long long j = i % chunk4;
        const float4* src = reinterpret_cast<const float4*>(send + (same_chunk ? 0 : (long long)p * chunk));
        reinterpret_cast<float4*>(recv.p[p] + (long long)rank * chunk)[j] = src[j];     // st.global over NVLink
    }
    // barrier 2 (by the last CTA, after a system fence): every peer's stores have landed here
    gate_exit(flags, rank, world, epoch, gate);
}
HETU_API int hetu_symm_exchange(const float* send, void* const* recv_ptrs, size_t byte_off, void* const* flag_ptrs,
                                int rank, int world, void* epoch, void* gate, long long chunk, int same_chunk,
                                cudaStream_t s) {
    if (chunk & 3) { hetu_set_error("symm_exchange: chunk (%lld floats) must be a multiple of 4", chunk); return -1; }
    FlagPtrs fp;
    if (fill_flags(fp, flag_ptrs, world)) return -1;
    PeerPtrs pp;
    fill_peers(pp, recv_ptrs, world, byte_off);
    int grid = coop_grid(chunk / 4 * world);
    if (grid > kNumSMs) grid = kNumSMs * 2;
    symm_barrier_kernel<<<1, 32, 0, s>>>(fp, rank, world, (unsigned*)epoch);   // opening barrier (see above)
    symm_exchange_kernel<<<grid, 256, 0, s>>>(send, pp, fp, rank, world, (unsigned*)epoch, (unsigned*)gate, chunk, same_chunk);
    HETU_CHECK_LAUNCH("hetu_symm_exchange");
    return 0;
}

// ------------------------------------------------------------- reduce-scatter (peer loads)
// every rank packed its full [world][chunk] contribution into its arena; out[chunk] = sum over
// peers of their slot [rank].
__global__ void __launch_bounds__(256) symm_reduce_scatter_kernel(PeerPtrs arenas, FlagPtrs flags, int rank, int world,
                                                                  unsigned* epoch, unsigned* gate, long long chunk,
                                                                  float* __restrict__ out) {
    long long n4 = chunk >> 2;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int p = 0; p < world; ++p) {
            float4 v = ld_sys_f4(arenas.p[p] + (long long)rank * chunk + (i << 2));
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
        reinterpret_cast<float4*>(out)[i] = acc;
    }
    gate_exit(flags, rank, world, epoch, gate);
}
HETU_API int hetu_symm_reduce_scatter(void* const* arena_ptrs, size_t byte_off, void* const* flag_ptrs, int rank,
                                      int world, void* epoch, void* gate, long long chunk, float* out, cudaStream_t s) {
    if (chunk & 3) { hetu_set_error("symm_reduce_scatter: chunk must be a multiple of 4 floats"); return -1; }
    FlagPtrs fp;
    if (fill_flags(fp, flag_ptrs, world)) return -1;
    PeerPtrs pp;
    fill_peers(pp, arena_ptrs, world, byte_off);
    symm_barrier_kernel<<<1, 32, 0, s>>>(fp, rank, world, (unsigned*)epoch);   // opening barrier (see above)
    symm_reduce_scatter_kernel<<<coop_grid(chunk / 4), 256, 0, s>>>(pp, fp, rank, world, (unsigned*)epoch, (unsigned*)gate, chunk, out);
    HETU_CHECK_LAUNCH("hetu_symm_reduce_scatter");
    return 0;
}

// ============================================================================================
// NVLS collectives: the buffers are mapped through an NVSwitch multicast object (parallel/symm.py
// MulticastBuffer), so ONE instruction addresses all ranks' copies - multimem.ld_reduce returns the sum over
// the ranks computed in the switch, multimem.st writes every copy.  reference call sites:
// python/hetu/gpu_ops/AllReduceCommunicate.py:15-44, BroadcastCommunicate.py:17-20, ReduceCommunicate.py:15-19
// (ncclAllReduce / ncclBroadcast / ncclReduce on a side stream).
//   all-reduce: rank r reduces slice r of the buffer in the switch and multicasts the result back into every rank's
//               copy (two NVLink traversals per element, no intermediate buffer, every replica bit-identical);
//   broadcast : the root multicasts its data into every copy;
//   reduce    : the root reads the switch-reduced sum.
// Each kernel is bracketed by the 1-CTA cross-GPU barrier kernel (launched by the host entry on the same stream).
__device__ __forceinline__ float4 mm_ld_reduce4(const float* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void mm_st4(float* mc, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}
// mode 0: all-reduce in place (slice `rank` of [0, n4)); 1: broadcast src -> every copy; 2: reduce -> out (root only)
__global__ void __launch_bounds__(256) nvls_collective_kernel(float* mc, const float* __restrict__ src, float* __restrict__ out,
                                                              long long n4, int rank, int world, int mode, float scale) {
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nth = (long long)gridDim.x * blockDim.x;
    if (mode == 0) {
        const long long per = (n4 + world - 1) / world, lo = per * rank, hi = lo + per < n4 ? lo + per : n4;
        for (long long i = lo + tid; i < hi; i += nth) {
            float4 v = mm_ld_reduce4(mc + 4 * i);
            v.x *= scale; v.y *= scale; v.z *= scale; v.w *= scale;
            mm_st4(mc + 4 * i, v);
        }
    } else if (mode == 1) {
        for (long long i = tid; i < n4; i += nth) mm_st4(mc + 4 * i, __ldg(reinterpret_cast<const float4*>(src) + i));
    } else {
        for (long long i = tid; i < n4; i += nth) {
            float4 v = mm_ld_reduce4(mc + 4 * i);
            v.x *= scale; v.y *= scale; v.z *= scale; v.w *= scale;
            reinterpret_cast<float4*>(out)[i] = v;
        }
    }
    __threadfence_system();
}

HETU_API int hetu_nvls_collective(void* mc_ptr, const float* src, float* out, long long n, int mode, int root, float scale,
                                  void* const* flag_ptrs, int rank, int world, void* epoch, cudaStream_t s) {
    if (n % 4 != 0 || mc_ptr == nullptr) { hetu_set_error("hetu_nvls_collective: needs a multicast mapping and n %% 4 == 0"); return -1; }
    FlagPtrs fp;
    if (fill_flags(fp, flag_ptrs, world)) return -1;
    const long long n4 = n / 4;
    symm_barrier_kernel<<<1, 32, 0, s>>>(fp, rank, world, (unsigned*)epoch);          // every rank's input is in its copy
    const bool active = mode == 0 || rank == root;
    if (active) {
        const long long work = mode == 0 ? (n4 + world - 1) / world : n4;
        nvls_collective_kernel<<<coop_grid(work), 256, 0, s>>>((float*)mc_ptr, src, out, n4, rank, world, mode, scale);
    }
    symm_barrier_kernel<<<1, 32, 0, s>>>(fp, rank, world, (unsigned*)epoch);          // results have landed everywhere
    HETU_CHECK_LAUNCH("hetu_nvls_collective");
    return 0;
}
