// MoE dispatch / combine layout transforms and the gate gradient.
// reference: src/ops/LayoutTransform.cu:3-116 (128 CTAs x 1024 threads, block-stride over
// tokens, one launch per top-k route, combine through float atomics).
//
// One warp per token, float4 row copies, all k routes in one launch.  The peer variants
// take a table of per-GPU expert-buffer base pointers: dispatch stores each token row
// straight into the owner GPU's [E_local, world, C, d] buffer (st.global over NVLink) and
// combine reads the expert outputs back the same way - the all-to-all IS the layout
// transform.
#include "common.cuh"

struct Routes { const float* idx[4]; const float* loc[4]; const float* gate[4]; int k; int has_gate; };
constexpr int kMaxPeers = 16;
struct PeerBufs { float* base[kMaxPeers]; int world; int rank; };

__device__ __forceinline__ bool route_ok(const Routes& r, int j, int t, int capacity, int total_experts, int& e, int& slot) {
    float fi = r.idx[j][t], fl = r.loc[j][t];
    e = (int)fi; slot = (int)fl;
    return fi >= 0.f && fl >= 0.f && slot < capacity && e < total_experts;
}

__global__ void __launch_bounds__(256) moe_dispatch_kernel(const float* __restrict__ x, Routes r, float* __restrict__ out,
                                                           int T, int d, int capacity, int total_experts) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    int nwarp = (gridDim.x * blockDim.x) >> 5;
    for (int t = warp; t < T; t += nwarp) {
        const float4* src = reinterpret_cast<const float4*>(x + (size_t)t * d);
        for (int j = 0; j < r.k; ++j) {
            int e, slot;
            if (!route_ok(r, j, t, capacity, total_experts, e, slot)) continue;
            float g = r.has_gate ? r.gate[j][t] : 1.f;
            float4* dst = reinterpret_cast<float4*>(out + ((size_t)e * capacity + slot) * d);
            for (int c = lane; c < (d >> 2); c += 32) {
                float4 v = src[c];
                if (r.has_gate) { v.x *= g; v.y *= g; v.z *= g; v.w *= g; }
                dst[c] = v;
            }
        }
    }
}
__global__ void __launch_bounds__(256) moe_dispatch_kernel_s(const float* __restrict__ x, Routes r, float* __restrict__ out,
                                                             int T, int d, int capacity, int total_experts) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    int nwarp = (gridDim.x * blockDim.x) >> 5;
    for (int t = warp; t < T; t += nwarp)
        for (int j = 0; j < r.k; ++j) {
            int e, slot;
            if (!route_ok(r, j, t, capacity, total_experts, e, slot)) continue;
            float g = r.has_gate ? r.gate[j][t] : 1.f;
            for (int c = lane; c < d; c += 32) out[((size_t)e * capacity + slot) * d + c] = g * x[(size_t)t * d + c];
        }
}

static Routes make_routes(void* const* idx, void* const* loc, void* const* gate, int k, int has_gate) {
    Routes r;
    r.k = k; r.has_gate = has_gate;
    for (int i = 0; i < 4; ++i) {
        r.idx[i] = i < k ? (const float*)idx[i] : nullptr;
        r.loc[i] = i < k ? (const float*)loc[i] : nullptr;
        r.gate[i] = (has_gate && i < k) ? (const float*)gate[i] : nullptr;
    }
    return r;
}

// `out` must be zero-filled by the caller (dropped slots stay zero)
HETU_API int hetu_moe_dispatch(const float* x, void* const* idx, void* const* loc, void* const* gate, int k, int has_gate,
                               float* out, int T, int d, int capacity, int total_experts, cudaStream_t s) {
    if (T == 0) return 0;
    Routes r = make_routes(idx, loc, gate, k, has_gate);
    int grid = ew_grid((size_t)T * 32, 256, 16);
    bool vec = (d & 3) == 0 && ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
    if (vec) moe_dispatch_kernel<<<grid, 256, 0, s>>>(x, r, out, T, d, capacity, total_experts);
    else moe_dispatch_kernel_s<<<grid, 256, 0, s>>>(x, r, out, T, d, capacity, total_experts);
    HETU_CHECK_LAUNCH("hetu_moe_dispatch");
    return 0;
}

// out[t] = sum_j gate_j[t] * y[idx_j[t]*C + loc_j[t]]   (fixed order => deterministic)
__global__ void __launch_bounds__(256) moe_combine_kernel(const float* __restrict__ y, Routes r, float* __restrict__ out,
                                                          int T, int d, int capacity, int rows) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    int nwarp = (gridDim.x * blockDim.x) >> 5;
    for (int t = warp; t < T; t += nwarp) {
        int row[4];
        float g[4];
        for (int j = 0; j < 4; ++j) {
            row[j] = -1; g[j] = 0.f;
            if (j < r.k) {
                int e, slot;
                if (route_ok(r, j, t, capacity, 0x7fffffff, e, slot) && e * capacity + slot < rows) {
                    row[j] = e * capacity + slot;
                    g[j] = r.has_gate ? r.gate[j][t] : 1.f;
                }
            }
        }
        for (int c = lane; c < d; c += 32) {
            float acc = 0.f;
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (row[j] >= 0) acc += g[j] * y[(size_t)row[j] * d + c];
            out[(size_t)t * d + c] = acc;
        }
    }
}
HETU_API int hetu_moe_combine(const float* y, void* const* idx, void* const* loc, void* const* gate, int k, int has_gate,
                              float* out, int T, int d, int capacity, int rows, cudaStream_t s) {
    if (T == 0) return 0;
    Routes r = make_routes(idx, loc, gate, k, has_gate);
    moe_combine_kernel<<<ew_grid((size_t)T * 32, 256, 16), 256, 0, s>>>(y, r, out, T, d, capacity, rows);
    HETU_CHECK_LAUNCH("hetu_moe_combine");
    return 0;
}

__global__ void __launch_bounds__(256) moe_gate_grad_kernel(const float* __restrict__ dout, const float* __restrict__ y,
                                                            const float* __restrict__ idx, const float* __restrict__ loc,
                                                            float* __restrict__ out, int T, int d, int capacity, int rows) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    int nwarp = (gridDim.x * blockDim.x) >> 5;
    for (int t = warp; t < T; t += nwarp) {
        float fi = idx[t], fl = loc[t];
        int row = (int)fi * capacity + (int)fl;
        float acc = 0.f;
        if (fi >= 0.f && fl >= 0.f && (int)fl < capacity && row < rows)
            for (int c = lane; c < d; c += 32) acc += dout[(size_t)t * d + c] * y[(size_t)row * d + c];
        acc = warp_sum(acc);
        if (lane == 0) out[t] = acc;
    }
}
HETU_API int hetu_moe_gate_grad(const float* dout, const float* y, const float* idx, const float* loc, float* out, int T,
                                int d, int capacity, int rows, cudaStream_t s) {
    if (T == 0) return 0;
    moe_gate_grad_kernel<<<ew_grid((size_t)T * 32, 256, 16), 256, 0, s>>>(dout, y, idx, loc, out, T, d, capacity, rows);
    HETU_CHECK_LAUNCH("hetu_moe_gate_grad");
    return 0;
}

// ----------------------------------------------------------------- NVLink peer variants
// Expert e lives on GPU e / E_local.  The owner's receive buffer is laid out
// [E_local][world(src rank)][C][d]; token t of this rank with route (e, slot) is stored at
// peer[e / E_local] + (((e % E_local) * world + rank) * C + slot) * d.
__global__ void __launch_bounds__(256) moe_dispatch_peer_kernel(const float* __restrict__ x, Routes r, PeerBufs pb, int T,
                                                                int d, int capacity, int e_local) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    int nwarp = (gridDim.x * blockDim.x) >> 5;
    int total = e_local * pb.world;
    for (int t = warp; t < T; t += nwarp) {
        const float4* src = reinterpret_cast<const float4*>(x + (size_t)t * d);
        for (int j = 0; j < r.k; ++j) {
            int e, slot;
            if (!route_ok(r, j, t, capacity, total, e, slot)) continue;
            float g = r.has_gate ? r.gate[j][t] : 1.f;
            float* base = pb.base[e / e_local];
            float4* dst = reinterpret_cast<float4*>(base + ((((size_t)(e % e_local)) * pb.world + pb.rank) * capacity + slot) * d);
            for (int c = lane; c < (d >> 2); c += 32) {
                float4 v = src[c];
                if (r.has_gate) { v.x *= g; v.y *= g; v.z *= g; v.w *= g; }
                dst[c] = v;                                  // st.global to a peer-mapped address
            }
        }
    }
}
HETU_API int hetu_moe_dispatch_peer(const float* x, void* const* idx, void* const* loc, void* const* gate, int k,
                                    int has_gate, void* const* peer_bases, int world, int rank, int T, int d,
                                    int capacity, int e_local, cudaStream_t s) {
    if (T == 0) return 0;
    if (world > kMaxPeers || (d & 3)) { hetu_set_error("moe_dispatch_peer: world %d > %d or d %% 4 != 0", world, kMaxPeers); return -1; }
    Routes r = make_routes(idx, loc, gate, k, has_gate);
    PeerBufs pb;
    pb.world = world; pb.rank = rank;
    for (int i = 0; i < kMaxPeers; ++i) pb.base[i] = i < world ? (float*)peer_bases[i] : nullptr;
    moe_dispatch_peer_kernel<<<ew_grid((size_t)T * 32, 256, 16), 256, 0, s>>>(x, r, pb, T, d, capacity, e_local);
    HETU_CHECK_LAUNCH("hetu_moe_dispatch_peer");
    return 0;
}
__global__ void __launch_bounds__(256) moe_combine_peer_kernel(Routes r, PeerBufs pb, float* __restrict__ out, int T, int d,
                                                               int capacity, int e_local) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    int nwarp = (gridDim.x * blockDim.x) >> 5;
    int total = e_local * pb.world;
    for (int t = warp; t < T; t += nwarp) {
        const float* rowp[4];
        float g[4];
        for (int j = 0; j < 4; ++j) {
            rowp[j] = nullptr; g[j] = 0.f;
            if (j < r.k) {
                int e, slot;
                if (route_ok(r, j, t, capacity, total, e, slot)) {
                    rowp[j] = pb.base[e / e_local] + ((((size_t)(e % e_local)) * pb.world + pb.rank) * capacity + slot) * d;
                    g[j] = r.has_gate ? r.gate[j][t] : 1.f;
                }
            }
        }
        for (int c = lane; c < (d >> 2); c += 32) {
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (rowp[j]) {
                    float4 v = reinterpret_cast<const float4*>(rowp[j])[c];   // ld.global from the expert's GPU
                    acc.x += g[j] * v.x; acc.y += g[j] * v.y; acc.z += g[j] * v.z; acc.w += g[j] * v.w;
                }
            reinterpret_cast<float4*>(out + (size_t)t * d)[c] = acc;
        }
    }
}
HETU_API int hetu_moe_combine_peer(void* const* idx, void* const* loc, void* const* gate, int k, int has_gate,
                                   void* const* peer_bases, int world, int rank, float* out, int T, int d, int capacity,
                                   int e_local, cudaStream_t s) {
    if (T == 0) return 0;
    if (world > kMaxPeers || (d & 3)) { hetu_set_error("moe_combine_peer: world %d > %d or d %% 4 != 0", world, kMaxPeers); return -1; }
    Routes r = make_routes(idx, loc, gate, k, has_gate);
    PeerBufs pb;
    pb.world = world; pb.rank = rank;
    for (int i = 0; i < kMaxPeers; ++i) pb.base[i] = i < world ? (float*)peer_bases[i] : nullptr;
    moe_combine_peer_kernel<<<ew_grid((size_t)T * 32, 256, 16), 256, 0, s>>>(r, pb, out, T, d, capacity, e_local);
    HETU_CHECK_LAUNCH("hetu_moe_combine_peer");
    return 0;
}

// d gate_j[t] = < dout[t, :], y[route_j(t)] > with the expert output row read from its owner GPU
__global__ void __launch_bounds__(256) moe_gate_grad_peer_kernel(const float* __restrict__ dout, const float* __restrict__ idx,
                                                                 const float* __restrict__ loc, PeerBufs pb,
                                                                 float* __restrict__ out, int T, int d, int capacity,
                                                                 int e_local) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    int nwarp = (gridDim.x * blockDim.x) >> 5;
    int total = e_local * pb.world;
    for (int t = warp; t < T; t += nwarp) {
        float fi = idx[t], fl = loc[t];
        int e = (int)fi, slot = (int)fl;
        float acc = 0.f;
        if (fi >= 0.f && fl >= 0.f && slot < capacity && e < total) {
            const float* row = pb.base[e / e_local] + ((((size_t)(e % e_local)) * pb.world + pb.rank) * capacity + slot) * d;
            for (int c = lane; c < d; c += 32) acc += dout[(size_t)t * d + c] * row[c];
        }
        acc = warp_sum(acc);
        if (lane == 0) out[t] = acc;
    }
}
HETU_API int hetu_moe_gate_grad_peer(const float* dout, const float* idx, const float* loc, void* const* peer_bases,
                                     int world, int rank, float* out, int T, int d, int capacity, int e_local,
                                     cudaStream_t s) {
    if (T == 0) return 0;
    if (world > kMaxPeers) { hetu_set_error("moe_gate_grad_peer: world %d > %d", world, kMaxPeers); return -1; }
    PeerBufs pb;
    pb.world = world; pb.rank = rank;
    for (int i = 0; i < kMaxPeers; ++i) pb.base[i] = i < world ? (float*)peer_bases[i] : nullptr;
    moe_gate_grad_peer_kernel<<<ew_grid((size_t)T * 32, 256, 16), 256, 0, s>>>(dout, idx, loc, pb, out, T, d, capacity, e_local);
    HETU_CHECK_LAUNCH("hetu_moe_gate_grad_peer");
    return 0;
}

// ---- chunked variants: the exchange of one (local expert e, capacity part h) chunk runs while the GEMMs of another
// chunk occupy the tensor cores (ops/moe_ffn.py puts them on a side stream).  Buffer layout [E_local, parts, world,
// C / parts, d]: every chunk is a contiguous block of rows, and so is every expert.  e_filter / h_filter select the
// routes handled by this launch.  combine: accumulate != 0 adds to `out` (the chunks of one token's k routes arrive
// one after the other).
struct Chunk { int e_local, e_filter, parts, h_filter, cpart; };

__device__ __forceinline__ bool chunk_row(const Chunk& ck, int e, int slot, int world, int rank, size_t& row) {
    if ((e % ck.e_local) != ck.e_filter) return false;
    const int h = slot / ck.cpart;
    if (ck.h_filter >= 0 && h != ck.h_filter) return false;
    row = (((size_t)ck.e_filter * ck.parts + h) * world + rank) * ck.cpart + (slot - h * ck.cpart);
    return true;
}

__global__ void __launch_bounds__(256) moe_dispatch_peer_x_kernel(const float* __restrict__ x, Routes r, PeerBufs pb, int T, int d,
                                                                  int capacity, Chunk ck) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    int nwarp = (gridDim.x * blockDim.x) >> 5;
    int total = ck.e_local * pb.world;
    for (int t = warp; t < T; t += nwarp) {
        const float4* src = reinterpret_cast<const float4*>(x + (size_t)t * d);
        for (int j = 0; j < r.k; ++j) {
            int e, slot;
            size_t row;
            if (!route_ok(r, j, t, capacity, total, e, slot) || !chunk_row(ck, e, slot, pb.world, pb.rank, row)) continue;
            float g = r.has_gate ? r.gate[j][t] : 1.f;
            float4* dst = reinterpret_cast<float4*>(pb.base[e / ck.e_local] + row * d);
            for (int c = lane; c < (d >> 2); c += 32) {
                float4 v = src[c];
                if (r.has_gate) { v.x *= g; v.y *= g; v.z *= g; v.w *= g; }
                dst[c] = v;
            }
        }
    }
}
__global__ void __launch_bounds__(256) moe_combine_peer_x_kernel(Routes r, PeerBufs pb, float* __restrict__ out, int T, int d, int capacity,
                                                                 Chunk ck, int accumulate) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    int nwarp = (gridDim.x * blockDim.x) >> 5;
    int total = ck.e_local * pb.world;
    for (int t = warp; t < T; t += nwarp) {
        const float* rowp[4];
        float g[4];
        bool any = false;
        for (int j = 0; j < 4; ++j) {
            rowp[j] = nullptr; g[j] = 0.f;
            if (j < r.k) {
                int e, slot;
                size_t row;
                if (route_ok(r, j, t, capacity, total, e, slot) && chunk_row(ck, e, slot, pb.world, pb.rank, row)) {
                    rowp[j] = pb.base[e / ck.e_local] + row * d;
                    g[j] = r.has_gate ? r.gate[j][t] : 1.f;
                    any = true;
                }
            }
        }
        if (!any && accumulate) continue;
        for (int c = lane; c < (d >> 2); c += 32) {
            float4 acc = accumulate ? reinterpret_cast<const float4*>(out + (size_t)t * d)[c] : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (rowp[j]) {
                    float4 v = reinterpret_cast<const float4*>(rowp[j])[c];
                    acc.x += g[j] * v.x; acc.y += g[j] * v.y; acc.z += g[j] * v.z; acc.w += g[j] * v.w;
                }
            reinterpret_cast<float4*>(out + (size_t)t * d)[c] = acc;
        }
    }
}
// d gate[t] = <dout[t], expert output row of the route>, the row read at its owner, chunked buffer layout
__global__ void __launch_bounds__(256) moe_gate_grad_peer_x_kernel(const float* __restrict__ dout, const float* __restrict__ idx,
                                                                   const float* __restrict__ loc, PeerBufs pb,
                                                                   float* __restrict__ out, int T, int d, int capacity, int e_local,
                                                                   int parts) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    int nwarp = (gridDim.x * blockDim.x) >> 5;
    int total = e_local * pb.world;
    const int cpart = capacity / parts;
    for (int t = warp; t < T; t += nwarp) {
        float fi = idx[t], fl = loc[t];
        int e = (int)fi, slot = (int)fl;
        float acc = 0.f;
        if (fi >= 0.f && fl >= 0.f && slot < capacity && e < total) {
            const int h = slot / cpart;
            const float4* row = reinterpret_cast<const float4*>(
                pb.base[e / e_local] + ((((size_t)(e % e_local) * parts + h) * pb.world + pb.rank) * cpart + (slot - h * cpart)) * d);
            const float4* g = reinterpret_cast<const float4*>(dout + (size_t)t * d);
            for (int c = lane; c < (d >> 2); c += 32) {
                float4 a = g[c], b = row[c];
                acc += a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w;
            }
        }
        acc = warp_sum(acc);
        if (lane == 0) out[t] = acc;
    }
}

static int fill_chunk(Chunk& ck, PeerBufs& pb, void* const* peer_bases, int world, int rank, int d, int capacity, int e_local,
                      int e_filter, int parts, int h_filter, const char* who) {
    if (world > kMaxPeers || (d & 3) || parts < 1 || capacity % parts || e_filter < 0 || e_filter >= e_local || h_filter >= parts) {
        hetu_set_error("%s: world %d > %d, d %% 4 != 0, capacity %d not divisible into %d parts, or chunk (%d, %d) out of range", who,
                       world, kMaxPeers, capacity, parts, e_filter, h_filter);
        return -1;
    }
    ck.e_local = e_local; ck.e_filter = e_filter; ck.parts = parts; ck.h_filter = h_filter; ck.cpart = capacity / parts;
    pb.world = world; pb.rank = rank;
    for (int i = 0; i < kMaxPeers; ++i) pb.base[i] = i < world ? (float*)peer_bases[i] : nullptr;
    return 0;
}
// h_filter < 0: all capacity parts of expert e_filter in one launch
HETU_API int hetu_moe_dispatch_peer_x(const float* x, void* const* idx, void* const* loc, void* const* gate, int k, int has_gate,
                                      void* const* peer_bases, int world, int rank, int T, int d, int capacity, int e_local,
                                      int e_filter, int parts, int h_filter, cudaStream_t s) {
    if (T == 0) return 0;
    Chunk ck; PeerBufs pb;
    if (fill_chunk(ck, pb, peer_bases, world, rank, d, capacity, e_local, e_filter, parts, h_filter, "moe_dispatch_peer_x")) return -1;
    Routes r = make_routes(idx, loc, gate, k, has_gate);
    moe_dispatch_peer_x_kernel<<<ew_grid((size_t)T * 32, 256, 16), 256, 0, s>>>(x, r, pb, T, d, capacity, ck);
    HETU_CHECK_LAUNCH("hetu_moe_dispatch_peer_x");
    return 0;
}
HETU_API int hetu_moe_combine_peer_x(void* const* idx, void* const* loc, void* const* gate, int k, int has_gate,
                                     void* const* peer_bases, int world, int rank, float* out, int T, int d, int capacity,
                                     int e_local, int e_filter, int parts, int h_filter, int accumulate, cudaStream_t s) {
    if (T == 0) return 0;
    Chunk ck; PeerBufs pb;
    if (fill_chunk(ck, pb, peer_bases, world, rank, d, capacity, e_local, e_filter, parts, h_filter, "moe_combine_peer_x")) return -1;
    Routes r = make_routes(idx, loc, gate, k, has_gate);
    moe_combine_peer_x_kernel<<<ew_grid((size_t)T * 32, 256, 16), 256, 0, s>>>(r, pb, out, T, d, capacity, ck, accumulate);
    HETU_CHECK_LAUNCH("hetu_moe_combine_peer_x");
    return 0;
}
HETU_API int hetu_moe_gate_grad_peer_x(const float* dout, const float* idx, const float* loc, void* const* peer_bases, int world,
                                       int rank, float* out, int T, int d, int capacity, int e_local, int parts, cudaStream_t s) {
    if (T == 0) return 0;
    Chunk ck; PeerBufs pb;
    if (fill_chunk(ck, pb, peer_bases, world, rank, d, capacity, e_local, 0, parts, -1, "moe_gate_grad_peer_x")) return -1;
    moe_gate_grad_peer_x_kernel<<<ew_grid((size_t)T * 32, 256, 16), 256, 0, s>>>(dout, idx, loc, pb, out, T, d, capacity, e_local, parts);
    HETU_CHECK_LAUNCH("hetu_moe_gate_grad_peer_x");
    return 0;
}
