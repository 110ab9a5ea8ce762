// Steady-state replay of a captured training step in ONE native call, with the input copy of step t+1 and the loss
// read-back of step t running beside the graph of step t.
//
// reference: Executor.run walks the topological order in Python every step and the dataloader hands each batch over
// with its own copy (python/hetu/gpu_ops/executor.py:1000-1056, dataloader.py:88-140); on one stream that is
// H2D -> compute -> D2H, strictly one after the other.
//
// Here a step whose inputs all come from dataloaders is captured as
//     feed_select (copy staging slot s -> the graph's static input buffer, s <- s + 1)  ->  the step  ->
//     result_publish (small results -> result ring slot s)
// with the slot index living on the DEVICE, so the same graph can be replayed while the host fills the next staging
// slot.  hetu_step_plan_run(plan, batch pointers) then does per step: host slices -> pinned slot, H2D into device
// staging slot t % S on a copy stream (ordered after the graph that last read that slot), graph launch on the compute
// stream (ordered after that copy), D2H of result ring slot t % S on a third stream (ordered after the graph).  The
// compute stream sees graph after graph; graph/executor.py builds the plan once and calls this per step.
#include "common.cuh"
#include <cstring>

// state[0] = slot the NEXT replay reads, state[1] = slot of the replay in flight.  One CTA: every thread reads the slot
// before thread 0 advances it.
__global__ void __launch_bounds__(1024) feed_select_kernel(const uint4* __restrict__ stage, uint4* __restrict__ dst, long long total16,
                                                           int nslots, int* __restrict__ state) {
    const int s = state[0];
    __syncthreads();
    const uint4* src = stage + (long long)s * total16;
    for (long long i = threadIdx.x; i < total16; i += blockDim.x) dst[i] = src[i];
    if (threadIdx.x == 0) { state[1] = s; state[0] = (s + 1 == nslots) ? 0 : s + 1; }
}
__global__ void __launch_bounds__(256) result_publish_kernel(const float* __restrict__ src, float* __restrict__ ring, int n,
                                                             const int* __restrict__ state) {
    const int s = state[1];
    for (int i = threadIdx.x; i < n; i += blockDim.x) ring[(long long)s * n + i] = src[i];
}
// stage: nslots x total bytes (total % 16 == 0), dst: the static input buffer of the captured graph
HETU_API int hetu_feed_select(const void* stage, void* dst, long long total, int nslots, int* state, cudaStream_t s) {
    if (total % 16 || nslots < 1) { hetu_set_error("feed_select: total %lld must be a multiple of 16", total); return -1; }
    feed_select_kernel<<<1, 1024, 0, s>>>((const uint4*)stage, (uint4*)dst, total / 16, nslots, state);
    HETU_CHECK_LAUNCH("hetu_feed_select");
    return 0;
}
HETU_API int hetu_result_publish(const float* src, float* ring, int n, const int* state, cudaStream_t s) {
    result_publish_kernel<<<1, 256, 0, s>>>(src, ring, n, state);
    HETU_CHECK_LAUNCH("hetu_result_publish");
    return 0;
}

namespace {
constexpr int kPlanMaxFeeds = 16;
constexpr int kPlanMaxSlots = 8;
struct StepPlan {
    cudaGraphExec_t exec;
    cudaStream_t stream, copy_stream, rb_stream;
    int n, nslots, dslots;
    long long off[kPlanMaxFeeds], nbytes[kPlanMaxFeeds];
    char* host[kPlanMaxSlots];
    cudaEvent_t ev[kPlanMaxSlots];             // H2D out of pinned slot i done
    bool used[kPlanMaxSlots];
    char* stage;                               // device staging: dslots x total
    cudaEvent_t graph_done[kPlanMaxSlots];     // the graph that read device slot d has finished
    cudaEvent_t rb_read[kPlanMaxSlots];        // the read-back of result ring slot d has finished
    bool dused[kPlanMaxSlots], rused[kPlanMaxSlots];
    long long total;
    unsigned turn, dturn;
    // read-back of one small result: device ring [dslots][rb_bytes] -> ring of pinned slots
    const char* rb_ring;
    long long rb_bytes;
    int rb_slots;
    char* rb_host[kPlanMaxSlots];
    cudaEvent_t rb_ev[kPlanMaxSlots];
    unsigned rb_turn;
};
}  // namespace

// host_slots: nslots pinned staging buffers of `total` bytes; stage: device staging of dslots x total bytes read by the
// graph's feed_select node; dturn0: the device slot the next replay will read.
HETU_API void* hetu_step_plan_create(void* graph_exec, cudaStream_t stream, int n, const long long* off, const long long* nbytes,
                                     void* const* host_slots, int nslots, void* stage, int dslots, long long total, int dturn0) {
    if (n < 1 || n > kPlanMaxFeeds || nslots < 1 || nslots > kPlanMaxSlots || dslots < 1 || dslots > kPlanMaxSlots ||
        graph_exec == nullptr) {
        hetu_set_error("step_plan_create: %d feeds (max %d), %d / %d slots (max %d)", n, kPlanMaxFeeds, nslots, dslots, kPlanMaxSlots);
        return nullptr;
    }
    StepPlan* p = new StepPlan();
    memset(p, 0, sizeof(*p));
    p->exec = (cudaGraphExec_t)graph_exec;
    p->stream = stream;
    p->n = n; p->nslots = nslots; p->dslots = dslots; p->stage = (char*)stage; p->total = total;
    p->dturn = (unsigned)dturn0;
    for (int i = 0; i < n; ++i) { p->off[i] = off[i]; p->nbytes[i] = nbytes[i]; }
    bool ok = cudaStreamCreateWithFlags(&p->copy_stream, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&p->rb_stream, cudaStreamNonBlocking) == cudaSuccess;
    for (int i = 0; i < kPlanMaxSlots && ok; ++i) {
        if (i < nslots) p->host[i] = (char*)host_slots[i];
        ok = cudaEventCreateWithFlags(&p->ev[i], cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&p->graph_done[i], cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&p->rb_read[i], cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&p->rb_ev[i], cudaEventDisableTiming) == cudaSuccess;
    }
    if (!ok) {
        hetu_set_error("step_plan_create: stream / event creation failed");
        delete p;
        return nullptr;
    }
    return p;
}

// every step also copies `nbytes` of result ring slot (device, dslots x nbytes, written by the graph's result_publish
// node) into the next of `nslots` pinned host slots; host slot of step s (counted from this call) = s % nslots.
HETU_API int hetu_step_plan_set_readback(void* plan, const void* dev_ring, long long nbytes, void* const* host_slots, int nslots) {
    StepPlan* p = (StepPlan*)plan;
    if (nslots < 1 || nslots > kPlanMaxSlots || p->rb_slots != 0) { hetu_set_error("step_plan_set_readback: bad slot count %d", nslots); return -1; }
    for (int i = 0; i < nslots; ++i) p->rb_host[i] = (char*)host_slots[i];
    p->rb_ring = (const char*)dev_ring; p->rb_bytes = nbytes; p->rb_slots = nslots; p->rb_turn = 0;
    return 0;
}

// src[i]: host pointer of feed i for this step (pinned or pageable).  Returns after everything is queued.
HETU_API int hetu_step_plan_run(void* plan, void* const* src) {
    StepPlan* p = (StepPlan*)plan;
    const int slot = (int)(p->turn++ % (unsigned)p->nslots);
    const int d = (int)(p->dturn % (unsigned)p->dslots);
    p->dturn++;
    if (p->used[slot]) HETU_CUDA(cudaEventSynchronize(p->ev[slot]));      // the copy that last used this pinned slot has left the host
    char* h = p->host[slot];
    for (int i = 0; i < p->n; ++i) memcpy(h + p->off[i], src[i], (size_t)p->nbytes[i]);
    if (p->dused[d]) HETU_CUDA(cudaStreamWaitEvent(p->copy_stream, p->graph_done[d], 0));   // last reader of device slot d
    HETU_CUDA(cudaMemcpyAsync(p->stage + (long long)d * p->total, h, (size_t)p->total, cudaMemcpyHostToDevice, p->copy_stream));
    HETU_CUDA(cudaEventRecord(p->ev[slot], p->copy_stream));
    p->used[slot] = true;
    HETU_CUDA(cudaStreamWaitEvent(p->stream, p->ev[slot], 0));
    if (p->rused[d]) HETU_CUDA(cudaStreamWaitEvent(p->stream, p->rb_read[d], 0));           // result ring slot d has been read back
    HETU_CUDA(cudaGraphLaunch(p->exec, p->stream));
    HETU_CUDA(cudaEventRecord(p->graph_done[d], p->stream));
    p->dused[d] = true;
    if (p->rb_slots > 0) {
        const int rs = (int)(p->rb_turn++ % (unsigned)p->rb_slots);
        HETU_CUDA(cudaStreamWaitEvent(p->rb_stream, p->graph_done[d], 0));
        HETU_CUDA(cudaMemcpyAsync(p->rb_host[rs], p->rb_ring + (long long)d * p->rb_bytes, (size_t)p->rb_bytes, cudaMemcpyDeviceToHost,
                                  p->rb_stream));
        HETU_CUDA(cudaEventRecord(p->rb_ev[rs], p->rb_stream));
        HETU_CUDA(cudaEventRecord(p->rb_read[d], p->rb_stream));
        p->rused[d] = true;
    }
    return 0;
}

HETU_API int hetu_step_plan_wait(void* plan, int slot) {
    StepPlan* p = (StepPlan*)plan;
    if (slot < 0 || slot >= p->rb_slots) { hetu_set_error("step_plan_wait: slot %d out of range", slot); return -1; }
    HETU_CUDA(cudaEventSynchronize(p->rb_ev[slot]));
    return 0;
}

// 1: the read-back in host slot `slot` has landed, 0: still in flight
HETU_API int hetu_step_plan_ready(void* plan, int slot) {
    StepPlan* p = (StepPlan*)plan;
    if (slot < 0 || slot >= p->rb_slots) return 0;
    return cudaEventQuery(p->rb_ev[slot]) == cudaSuccess ? 1 : 0;
}

// device slot the next replay reads (the general path continues from here after the plan is dropped)
HETU_API int hetu_step_plan_dturn(void* plan) { return (int)(((StepPlan*)plan)->dturn % (unsigned)((StepPlan*)plan)->dslots); }

HETU_API void hetu_step_plan_destroy(void* plan) {
    StepPlan* p = (StepPlan*)plan;
    if (p == nullptr) return;
    cudaStreamSynchronize(p->copy_stream);
    cudaStreamSynchronize(p->rb_stream);
    for (int i = 0; i < kPlanMaxSlots; ++i) {
        cudaEventDestroy(p->ev[i]); cudaEventDestroy(p->graph_done[i]); cudaEventDestroy(p->rb_read[i]); cudaEventDestroy(p->rb_ev[i]);
    }
    cudaStreamDestroy(p->copy_stream);
    cudaStreamDestroy(p->rb_stream);
    delete p;
}
