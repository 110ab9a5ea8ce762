/* hetu_b200.h - C ABI of libhetu_b200.so (hand-written sm_100a kernels).
 *
 * Every function returns 0 on success; on failure a message is available from
 * hetu_last_error().  All pointers are device pointers unless noted; `s` is the
 * CUDA stream to launch on.  Tensors are dense row-major fp32 unless noted; row ids
 * are passed with an `idx_code` (0 = float32-encoded like the reference, 1 = int32,
 * 2 = int64).  The comment above each group names the reference code it replaces
 * (paths under the Hetu tree, /root/reference).
 */
#ifndef HETU_B200_H_
#define HETU_B200_H_

#include <cuda_runtime.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- runtime ------------------------------------------------------------------ */
/* Array / stream / event ABI: reference src/common/dlarray.h:22-65, c_runtime_api.cc:93-142,
 * src/cuda_common/gpu_runtime.cc:61-118.  Arrays are dense row-major fp32; device arrays come from
 * the device's stream-ordered memory pool, host arrays are pinned. */
typedef enum { kDLCPU = 1, kDLGPU = 2 } DLDeviceType;
typedef struct { int device_id; int device_type; } DLContext;
typedef struct { void* data; DLContext ctx; int ndim; long long* shape; long long* stride; } DLArray;
typedef DLArray* DLArrayHandle;
typedef struct { int device_id; void* handle; } DLStream;
typedef DLStream* DLStreamHandle;
typedef struct { int device_id; void* handle; } DLEvent;
typedef DLEvent* DLEventHandle;
int DLArrayAlloc(const long long* shape, int ndim, DLContext ctx, DLArrayHandle* out);
int DLArrayFree(DLArrayHandle handle);
/* any of host->device, device->host, device->device (same or peer GPU), host->host; stream NULL = blocking */
int DLArrayCopyFromTo(DLArrayHandle from, DLArrayHandle to, DLStreamHandle stream);
int DLStreamCreate(int device_id, DLStreamHandle* out);
int DLStreamDestroy(DLStreamHandle handle);
int DLStreamSync(DLStreamHandle handle);
int DLEventCreate(int device_id, DLEventHandle* out);
int DLEventDestroy(DLEventHandle handle);
int DLEventRecord(DLStreamHandle stream, DLEventHandle event);
int DLEventSync(DLEventHandle event);
int DLStreamWaitEvent(DLStreamHandle stream, DLEventHandle event);
int DLEventElapsedTime(DLEventHandle start, DLEventHandle end, float* ms);
/* workspace chunk cache: reference src/cuda_common/gpu_chunk.cc:18-86 (find_chunk / del_chunk).
 * hetu_find_chunk returns NULL on failure; hetu_del_chunk returns the chunk to the cache. */
void* hetu_find_chunk(size_t bytes, int device_id);
int hetu_del_chunk(void* ptr, int device_id);
size_t hetu_chunk_cache_bytes(int device_id);
int hetu_chunk_cache_release(int device_id);

const char* hetu_last_error(void);
int hetu_version(void);
/* *p += 1 on the device (step counters read by dropout / Adam; graph-capturable). */
int hetu_increment_i64(void* p, cudaStream_t s);

/* ---- element-wise: reference src/ops/{Abs,AddConst,...,Where}.cu ---------------- */
int hetu_unary(int op, const float* x, float* y, size_t n, float a, float b, cudaStream_t s);
int hetu_binary(int op, const float* x, const float* y, float* o, size_t n, float a, cudaStream_t s);
/* mode 0: y[inner] broadcast over rows (bias); mode 1: y[outer] broadcast along each row. */
int hetu_binary_bcast(int op, const float* x, const float* y, float* o, size_t n, size_t inner, int mode, cudaStream_t s);
/* hi = nearest tf32 value of x, lo = x - hi (operands of the 3xTF32 matmul mode) */
int hetu_split_tf32(const float* x, float* hi, float* lo, size_t n, cudaStream_t s);
/* src [B, R, C] -> dst [B, C, R] (shared-memory tiles); the NCHW <-> NHWC changes around the convolution GEMMs */
int hetu_transpose_batched(const float* src, float* dst, int B, int R, int C, cudaStream_t s);
int hetu_where(const float* c, const float* a, const float* b, float* o, size_t n, cudaStream_t s);
/* generic strided copy (transpose/slice/concat/pad/broadcast): reference
 * src/ops/{Transpose,Slice,Concat,Concatenate,Pad,Broadcast,BroadcastShape}.cu */
int hetu_strided_copy(const float* src, float* dst, int nd, const long long* shape, const long long* ss,
                      const long long* ds, size_t n, cudaStream_t s);
int hetu_one_hot(const float* idx, float* out, size_t n, int classes, cudaStream_t s); /* OneHot.cu */

/* ---- reductions: ReduceSum.cu ReduceMean.cu ReduceSumAxisZero.cu Max.cu Min.cu
 *      Argmax.cu TopKIdx.cu TopKVal.cu CumSum.cu -------------------------------------- */
int hetu_reduce_mid(const float* x, float* out, size_t A, size_t R, size_t B, int mode, cudaStream_t s);
int hetu_argmax_rows(const float* x, float* out, size_t rows, size_t cols, cudaStream_t s);
int hetu_topk_rows(const float* x, float* idx, float* val, size_t rows, size_t cols, int k, cudaStream_t s);
int hetu_cumsum_cols(const float* x, float* out, size_t T, size_t E, float bias, cudaStream_t s);

/* ---- GEMM (tcgen05 + TMA; FFMA fallback): MatrixMult.cu Linear.cu Addmm.cu
 *      BatchMatrixMult.cu Baddbmm.cu (cublasSgemm*) ; CSR SpMM: CuSparseCsrmm.cu --------
 * C = act(alpha * op(A) op(B) + beta * C + bias); mode 0 = tf32 tensor cores when
 * eligible, 1 = exact fp32 FFMA; act 0 none, 1 relu, 2 gelu, 3 sigmoid. */
int hetu_gemm(const float* A, const float* B, float* C, const float* bias, int M, int N, int K, int transA,
              int transB, int batch, long long sA, long long sB, long long sC, float alpha, float beta, int act,
              int mode, cudaStream_t s);
/* 16-bit operand products (csrc/gemm_f16_sm100.cu): A / B fp32 in memory, converted (and made K-major) into the caller's
 * scratch a16 / b16 (hetu_gemm_f16_scratch_elems(rows, K, batch) 2-byte elements each), multiplied with tcgen05.mma
 * kind::f16 (half_kind 0 = fp16, 1 = bf16), fp32 accumulation and result.  No half GEMM exists in the reference
 * (src/ops/MatrixMult.cu is cublasSgemm). */
long long hetu_gemm_f16_scratch_elems(int rows, int K, int batch);
int hetu_gemm_f16(const float* A, const float* B, float* C, const float* bias, void* a16, void* b16, int M, int N, int K,
                  int transA, int transB, int batch, long long sA, long long sB, long long sC, float alpha, float beta,
                  int act, int half_kind, cudaStream_t s);
int hetu_csr_spmm(const float* val, const int* rowptr, const int* colidx, const float* dense, float* out,
                  int nrow, int n, int trans, cudaStream_t s);

/* ---- NN: CudnnSoftmax.cu Softmax.cu LayerNorm.cu InstanceNorm2d.cu CudnnBn.cu Dropout.cu
 *      Conv2d.cu (im2col) CudnnConv2d.cu MaxPool.cu AvgPool.cu Cudnn{Max,Avg}Pool.cu ------ */
int hetu_softmax_rows(const float* x, float* y, int rows, int cols, cudaStream_t s);
int hetu_softmax_grad_rows(const float* y, const float* dy, float* dx, int rows, int cols, cudaStream_t s);
int hetu_layernorm_fwd(const float* x, const float* gamma, const float* beta, float* y, float* mean, float* rstd,
                       int rows, int cols, float eps, cudaStream_t s);
int hetu_layernorm_bwd(const float* dy, const float* x, const float* gamma, const float* mean, const float* rstd,
                       float* dx, float* dgamma, float* dbeta, int rows, int cols, cudaStream_t s);
int hetu_batchnorm_fwd(const float* x, const float* gamma, const float* beta, float* y, float* run_mean,
                       float* run_var, float* save_mean, float* save_rstd, int N, int C, int HW, float momentum,
                       float eps, int inference, cudaStream_t s);
int hetu_batchnorm_bwd(const float* dy, const float* x, const float* gamma, const float* save_mean,
                       const float* save_rstd, float* dx, float* dgamma, float* dbeta, int N, int C, int HW,
                       cudaStream_t s);
int hetu_dropout(const float* x, float* y, size_t n, float p, unsigned long long seed, const void* step, cudaStream_t s);
int hetu_im2col(const float* x, float* col, int N, int C, int H, int W, int kh, int kw, int ph, int pw, int sh,
                int sw, int OH, int OW, cudaStream_t s);
int hetu_col2im(const float* col, float* x, int N, int C, int H, int W, int kh, int kw, int ph, int pw, int sh,
                int sw, int OH, int OW, cudaStream_t s);
/* same from colT [N, C*kh*kw, OH*OW] (batched GEMM W^T dY_n on NCHW dY): coalesced reads */
int hetu_col2im_t(const float* colT, float* x, int N, int C, int H, int W, int kh, int kw, int ph, int pw,
                  int sh, int sw, int OH, int OW, cudaStream_t s);
int hetu_pool2d_fwd(const float* x, float* y, int NC, int H, int W, int kh, int kw, int pad, int stride, int OH,
                    int OW, int is_max, cudaStream_t s);
int hetu_pool2d_bwd(const float* x, const float* y, const float* dy, float* dx, int NC, int H, int W, int kh, int kw,
                    int pad, int stride, int OH, int OW, int is_max, cudaStream_t s);

/* ---- losses: SoftmaxCrossEntropy.cu CudnnSoftmaxEntropy.cu SoftmaxCrossEntropySparse.cu
 *      BinaryCrossEntropy.cu ------------------------------------------------------------ */
int hetu_softmax_xent(const float* x, const float* y, float* loss, int rows, int cols, cudaStream_t s);
int hetu_softmax_xent_grad(const float* x, const float* y, const float* g, float* dx, int rows, int cols, cudaStream_t s);
int hetu_softmax_xent_sparse(const float* x, const float* lab, float* loss, int rows, int cols, int ignored, cudaStream_t s);
int hetu_softmax_xent_sparse_grad(const float* x, const float* lab, const float* g, float* dx, int rows, int cols,
                                  int ignored, cudaStream_t s);
int hetu_bce(const float* p, const float* y, float* o, size_t n, cudaStream_t s);
int hetu_bce_grad(const float* p, const float* y, const float* g, float* o, size_t n, cudaStream_t s);

/* ---- embedding + sparse optimizers: EmbeddingLookup.cu OptimizersSparse.cu IndexedSlices.cu;
 *      the *_sharded variants replace the PS SparsePull / SparsePush RPCs
 *      (ps-lite/include/ps/server/PSFHandle.h:101-164) with NVLink peer loads / red.adds ---- */
int hetu_gather_rows(const float* table, const void* idx, int idx_code, float* out, size_t n, int width,
                     size_t vocab, cudaStream_t s);
int hetu_gather_rows_sharded(void* const* bases /*host array*/, int nshard, const void* idx, int idx_code,
                             float* out, size_t n, int width, size_t vocab, cudaStream_t s);
int hetu_dedup_sorted_rows(const void* sidx, const void* perm, const float* vals, void* didx, float* dval, size_t n,
                           int width, cudaStream_t s);
int hetu_sparse_opt(int kind, float* param, const void* idx, const float* grad, float* s1, float* s2, size_t n,
                    int width, const float* lr_p, const void* step_p, float h0, float h1, float h2, float h3,
                    float l2reg, int use_atomic, cudaStream_t s);
int hetu_sparse_sgd_sharded(void* const* bases /*host array*/, int nshard, const void* idx, int idx_code,
                            const float* grad, size_t n, int width, size_t vocab, const float* lr_p, float scale,
                            cudaStream_t s);
/* sharded table, optimizers with state applied at the owner: phase 1 adds every gradient row into the owner's accumulation
 * shard (remote atomics; the first toucher lists the row), phase 2 (after a cross-GPU barrier) applies Momentum / Nesterov /
 * AdaGrad / Adam / AdamW to the listed rows.  reference: ps-lite/include/ps/server/optimizer.h:36-357,
 * src/ops/OptimizersSparse.cu:54-67,287-330 */
int hetu_sparse_push_sharded(void* const* acc, void* const* flag, void* const* list, void* const* count /*host arrays*/,
                             int nshard, int cap, const void* idx, int idx_code, const float* grad, size_t n, int width,
                             size_t vocab, float scale, cudaStream_t s);
int hetu_sparse_apply_owner(int kind, float* table, float* acc, void* flag, const void* list, void* count, void* done,
                            float* s1, float* s2, int width, int cap, const float* lr_p, const void* step_p, float h0,
                            float h1, float h2, float h3, float l2reg, cudaStream_t s);
int hetu_scatter_add_rows(float* dense, const void* idx, int idx_code, const float* vals, size_t n, int width,
                          size_t vocab, cudaStream_t s);

/* ---- Wide&Deep training step fused (csrc/wdl_tower.cu; graph of examples/ctr/models/wdl_criteo.py:8-42) -----
 * gather_rows_pitched: lookup + concat, row (b, f) -> out[b * pitch + f * width]; bases = host array of the nshard
 *   table shards (row r on shard r % nshard, local row r / nshard; nshard == 1: plain table).
 * wdl_tower_step: ONE cluster kernel (8 CTAs, tcgen05 kind::tf32) for the dense tower forward, the head, sigmoid +
 *   BCE + mean, and the whole backward pass; y4 [128, emb_cols + 256] holds the gathered rows in its first emb_cols
 *   columns; work = 3 * 128 * 256 floats; writes loss[1], pred[128], dz[128] and the four weight gradients.
 * wdl_sparse_sgd_rank1: table rows -= lr * scale * dz[b] * W4[f * width ...] (the head's gradient is rank one). */
int hetu_gather_rows_pitched(void* const* bases /*host array*/, int nshard, const void* idx, int idx_code, float* out,
                             size_t n, int width, int fields, int pitch, size_t vocab,
                             const float* w4 /*optional*/, float* zpf /*optional [n]: <row, w4 block of its field>*/,
                             float* w4_snap /*optional: copy of w4[0 .. fields * width)*/,
                             const void* dz_flag /*optional*/, void* dz_expect /*optional: *dz_flag + 1*/, cudaStream_t s);
int hetu_wdl_tower_step(const float* x, float* y4, const float* zpf /*[128, fields] from the gather*/,
                        const float* labels, const float* W1, const float* W2, const float* W3, const float* W4,
                        float* work, float* loss, float* pred, float* dz, float* dW1, float* dW2, float* dW3,
                        float* dW4, int batch, int nd, int hidden, int emb_cols, int width,
                        /* arenas != NULL: data-parallel all-reduce + SGD inside the kernel.  arenas[r] = rank r's
                         * [2][world][total] float arena (peer-mapped; total = nd*256 + 2*65536 + emb_cols + 256),
                         * flags[r] = rank r's unsigned[world] flag row (zeroed once), epoch = this rank's device
                         * counter (zeroed once), lr / opt_step = device scalars.  The gradient slices are pushed to
                         * all ranks as the backward pass produces them; W1..W4 are updated in place. */
                        void* const* arenas /*host array or NULL*/, void* const* flags /*host array*/, int world,
                        int rank, void* epoch, const float* lr, void* opt_step,
                        /* pdl != 0: launched as a programmatic dependent of the gather (the first layers overlap
                         * it); dz_flag (optional, device unsigned, zeroed once) is bumped when dz is out, which is
                         * what hetu_wdl_sparse_sgd_rank1 waits for when it is given the same flag */
                        int pdl, void* dz_flag,
                        /* NVLS multicast mappings of the arenas / flag rows (or NULL: plain peer stores): the pushes
                         * become multimem.st, one store per value instead of one per rank */
                        void* arena_mc, void* flags_mc, cudaStream_t s);
/* data-parallel W&D: after hetu_wdl_tower_step has pushed this step's gradient blocks, wait for every source's flag,
 * sum the blocks of this rank's arena in rank order and apply SGD to W1..W4 (full-width grid) */
int hetu_wdl_dp_apply(const float* arena_local, const void* flags_local, const void* epoch, const float* lr, float* W1,
                      float* W2, float* W3, float* W4, int nd, int emb_cols, int world, int rank, cudaStream_t s);
int hetu_wdl_tower_stamps(unsigned long long* out9);   /* diagnostics, see benchmarks/wdl_tower_phases.py */
int hetu_wdl_sparse_sgd_rank1(void* const* bases /*host array*/, int nshard, const void* idx, int idx_code,
                              const float* dz, const float* w4, const float* lr_p, float scale, size_t n, int width,
                              int fields, size_t vocab, const void* dz_flag /*optional*/,
                              const void* dz_expect /*optional*/, cudaStream_t s);

/* ---- HET embedding cache in HBM (csrc/het_cache.cu; reference src/hetu_cache/{include/embedding.h,src/cache.cc,
 *      src/lfuopt_cache.cc}).  8-way set-associative table: tags [nsets*8] i64 (-1 empty), data / grad [nsets*8, width],
 *      version i64, updates i32, score u32, queued i32, lock [nsets] i32; policy 0 LRU, 1 LFU, 2 LFUOpt.
 *      probe: slot / cached version per id; install: rows from the server (miss) or refresh (stale), dirty victims to the
 *      eviction list; gather: rows out; update: cached row and kept gradient += delta, lines over push_bound queued;
 *      drain: queued (or all dirty) lines hand their gradients over.  Host side: parallel/gpu_cache.py. ---- */
#define HETU_HET_ARGS void* tags, float* data, float* grad, void* version, void* updates, void* score, void* queued, \
                      void* lock, int nsets, int width, int policy
int hetu_het_probe(HETU_HET_ARGS, const void* ids, size_t n, int* slot, void* ver, unsigned tick, int touch, void* stats,
                   cudaStream_t s);
int hetu_het_install(HETU_HET_ARGS, const void* ids, const void* vers, const float* rows, size_t m, unsigned tick,
                     void* ev_ids, float* ev_grads, void* ev_upd, void* ev_count, int ev_cap, void* stats, cudaStream_t s);
int hetu_het_gather(HETU_HET_ARGS, const int* slot, size_t n, float* out, const float* fallback, const int* fb_index,
                    cudaStream_t s);
int hetu_het_update(HETU_HET_ARGS, const int* slot, const float* delta, size_t n, int push_bound, int* push_slots,
                    void* push_count, int* direct, void* direct_count, cudaStream_t s);
int hetu_het_drain(HETU_HET_ARGS, const int* push_slots, const void* push_count, int all_dirty, void* out_ids,
                   float* out_grads, void* out_upd, void* out_count, int out_cap, cudaStream_t s);

/* ---- dense optimizers, all tensors in one launch: Optimizers.cu ------------------------
 * kind: 0 SGD 1 Momentum 2 Nesterov 3 AdaGrad 4 Adam 5 AdamW 6 Lamb.  The pointer arrays are
 * HOST arrays of device pointers (<= 48 entries).  Lamb needs `scratch`: 2 floats per tensor of
 * device memory, zeroed by the caller, for the parameter / update norms. */
int hetu_multi_tensor_opt(int kind, int ntensors, void* const* params, void* const* grads, void* const* s1,
                          void* const* s2, const long long* sizes, const float* lr_p, const void* step_p, float h0,
                          float h1, float h2, float h3, float l2reg, float* scratch, cudaStream_t s);

/* Lamb step for callers that cannot allocate device memory (plugins inside another framework's
 * callback, services that forbid cudaMalloc after start-up, thin FFI clients that own only the
 * tensors they pass).  Same arithmetic as hetu_multi_tensor_opt(kind = 6) but the per-tensor norm
 * scratch is a static device array inside the library: 768 bytes held on every GPU that loads the
 * library, at most 96 tensors per call (more is refused with an error).  The array is shared: the
 * call clears the slots it uses, and calls that can overlap on one GPU race - keep them on one
 * stream.  The first call on a device resolves the array's address; make it before any stream
 * capture.  Calls no cudaMalloc/cudaFree.  This entry point has no counterpart in the reference. */
int hetu_lamb_step_noalloc(int ntensors, void* const* params, void* const* grads, void* const* s1,
                           void* const* s2, const long long* sizes, const float* lr_p, const void* step_p,
                           float beta1, float beta2, float eps, float weight_decay, float l2reg, cudaStream_t s);

/* ---- MoE layout transforms: LayoutTransform.cu; *_peer: AllToAll fused with the transform
 *      (python/hetu/gpu_ops/AllToAll.py:13-29 + mpi_nccl_communication.cu:245-277) --------- */
int hetu_moe_dispatch(const float* x, void* const* idx, void* const* loc, void* const* gate, int k, int has_gate,
                      float* out, int T, int d, int capacity, int total_experts, cudaStream_t s);
int hetu_moe_combine(const float* y, void* const* idx, void* const* loc, void* const* gate, int k, int has_gate,
                     float* out, int T, int d, int capacity, int rows, cudaStream_t s);
int hetu_moe_gate_grad(const float* dout, const float* y, const float* idx, const float* loc, float* out, int T,
                       int d, int capacity, int rows, cudaStream_t s);
int hetu_moe_dispatch_peer(const float* x, void* const* idx, void* const* loc, void* const* gate, int k,
                           int has_gate, void* const* peer_bases, int world, int rank, int T, int d, int capacity,
                           int e_local, cudaStream_t s);
int hetu_moe_combine_peer(void* const* idx, void* const* loc, void* const* gate, int k, int has_gate,
                          void* const* peer_bases, int world, int rank, float* out, int T, int d, int capacity,
                          int e_local, cudaStream_t s);

/* chunked exchanges: buffer layout [E_local, parts, world, capacity / parts, d]; one launch moves the routes of local
 * expert e_filter whose slot falls into capacity part h_filter (h_filter < 0: every part); combine may accumulate into
 * out.  ops/moe_ffn.py runs them on a side stream under the GEMMs of the other chunks. */
int hetu_moe_dispatch_peer_x(const float* x, void* const* idx, void* const* loc, void* const* gate, int k, int has_gate,
                             void* const* peer_bases, int world, int rank, int T, int d, int capacity, int e_local,
                             int e_filter, int parts, int h_filter, cudaStream_t s);
int hetu_moe_combine_peer_x(void* const* idx, void* const* loc, void* const* gate, int k, int has_gate,
                            void* const* peer_bases, int world, int rank, float* out, int T, int d, int capacity,
                            int e_local, int e_filter, int parts, int h_filter, int accumulate, cudaStream_t s);
int hetu_moe_gate_grad_peer_x(const float* dout, const float* idx, const float* loc, void* const* peer_bases, int world,
                              int rank, float* out, int T, int d, int capacity, int e_local, int parts, cudaStream_t s);

int hetu_moe_gate_grad_peer(const float* dout, const float* idx, const float* loc, void* const* peer_bases,
                            int world, int rank, float* out, int T, int d, int capacity, int e_local,
                            cudaStream_t s);

/* ---- attention: examples/nlp/bert/hetu_bert.py:249-264 builds it from 5 graph ops; here the two
 *      batched GEMMs run on hetu_gemm and everything between them is one in-place row kernel ---- */
int hetu_attn_softmax(float* s, const float* mask, float* lse, long long rows, int cols, long long rows_per_batch,
                      float scale, cudaStream_t st);
int hetu_attn_softmax_bwd(const float* p, float* dp, long long rows, int cols, float scale, cudaStream_t st);
/* s (raw scores q k^T) <- exp(s * scale + mask - lse[row]): probabilities recomputed for the backward pass */
int hetu_attn_probs_from_lse(float* s, const float* mask, const float* lse, long long rows, int cols,
                             long long rows_per_batch, float scale, cudaStream_t st);
/* flash-style forward in one tcgen05 kernel (csrc/attention_sm100.cu): q [B,H,S,D], k/v [B,H,Sk,D] fp32 contiguous,
 * additive key mask [B,Sk] or null, lse [B,H,S] or null; D = 32 or 64.  The reference builds attention from five ops
 * (examples/nlp/bert/hetu_bert.py:249-264). */
int hetu_flash_attn_fwd(const float* q, const float* k, const float* v, const float* mask, float* out, float* lse,
                        int B, int H, int S, int Sk, int D, float scale, cudaStream_t st);

/* ---- symmetric heap + in-kernel NVLink collectives (csrc/collectives.cu) ----------------------
 * reference: src/communication/mpi_nccl_communication.cu:31-483 (NCCL wrappers) and
 * gpu_ops/{AllReduceCommunicate,AllGatherCommunicate,ReduceScatterCommunicate,AllToAll}.py.
 * Buffers come from hetu_symm_alloc (cudaMalloc + CUDA IPC handle); every rank opens its peers'
 * handles and passes the table of base pointers (`*_ptrs`, up to 16 entries).  `flag_ptrs` is the
 * table of each rank's flag page, `epoch`/`gate` are device counters owned by the caller.  Every
 * collective opens and closes with a cross-GPU barrier, so buffers are reusable on return and the
 * launches are CUDA-graph replay safe. */
int hetu_symm_alloc(size_t bytes, void** ptr, void* handle64);
int hetu_symm_open(const void* handle64, void** ptr);
int hetu_symm_close(void* ptr);
int hetu_symm_free(void* ptr);
/* NVLS collectives on a multicast-mapped symmetric buffer (parallel/symm.py MulticastBuffer): mode 0 all-reduce in place
 * (multimem.ld_reduce of the rank's slice + multimem.st of the result), 1 broadcast of src from root, 2 reduce into out on
 * root.  reference: gpu_ops/AllReduceCommunicate.py:15-44, BroadcastCommunicate.py:17-20, ReduceCommunicate.py:15-19. */
int hetu_nvls_collective(void* mc_ptr, const float* src, float* out, long long n, int mode, int root, float scale,
                         void* const* flag_ptrs /*host array*/, int rank, int world, void* epoch, cudaStream_t s);
int hetu_symm_barrier(void* const* flag_ptrs, int rank, int world, void* epoch, cudaStream_t s);
int hetu_pack_tensors(int ntensors, void* const* srcs, const long long* offsets, const long long* sizes, float* dst,
                      float scale, cudaStream_t s);
int hetu_symm_allreduce(void* const* arena_ptrs, size_t byte_off, void* const* flag_ptrs, int rank, int world,
                        void* epoch, void* gate, long long n, float* out, cudaStream_t s);
/* all-reduce fused with the optimizer update of `ntensors` parameters (kind = OPT_* of multi_tensor_opt) */
int hetu_symm_allreduce_opt(void* const* arena_ptrs, size_t byte_off, void* const* flag_ptrs, int rank, int world,
                            void* epoch, void* gate, int kind, int ntensors, void* const* params, void* const* s1,
                            void* const* s2, const long long* offsets, const long long* sizes, const float* lr_p,
                            const void* step_p, float h0, float h1, float h2, float h3, float l2reg, cudaStream_t s);
/* same_chunk = 1: all-gather (every peer receives `chunk` floats from me); 0: all-to-all of `chunk` floats per peer */
int hetu_symm_exchange(const float* send, void* const* recv_ptrs, size_t byte_off, void* const* flag_ptrs, int rank,
                       int world, void* epoch, void* gate, long long chunk, int same_chunk, cudaStream_t s);
int hetu_symm_reduce_scatter(void* const* arena_ptrs, size_t byte_off, void* const* flag_ptrs, int rank, int world,
                             void* epoch, void* gate, long long chunk, float* out, cudaStream_t s);

/* ---- tensor-parallel GEMMs fused with their collective (persistent 2-SM tcgen05 kernel) ---------
 * reference: the model-parallel split / datatransfer ops (gpu_ops/Dispatch.py marker; lowering absent
 * upstream).  hetu_gemm_ag: C = allgather_rows(A) @ op(B); the epilogue warps push this rank's shard to
 * every peer's gathered buffer in 128-row blocks while the MMA warps start on the local rows; remote
 * tiles wait for their block's flag.  hetu_gemm_rs: partial = op(A) @ op(B); row block r is stored into
 * slice `rank` of rank r's [world][M/world][N] staging buffer straight from the epilogue. */
int hetu_gemm_ag(void* const* gathered, void* const* flags, int rank, int world, const float* B, float* C,
                 const float* bias, int M_total, int N, int K, int transB, float alpha, int act, cudaStream_t s);
int hetu_gemm_rs(const float* A, const float* B, void* const* out_peers, int rank, int world, int M, int N, int K,
                 int transA, int transB, float alpha, cudaStream_t s);

/* ---- parameter initialisation: Initializers.cu ------------------------------------------
 * kind 0 constant(a) 1 uniform[a,b) 2 normal(mean a, std b) 3 truncated normal (|z| <= 2). */
int hetu_init_random(float* out, size_t n, int kind, float a, float b, unsigned long long seed, cudaStream_t s);

/* ---- hetuml GBDT on the device: reference third_party/HetuML/hetuml/cc/ml/include/model/gbdt/{hist,helper}/ ----
 * One launch per tree level.  hist [num_slots, total_bins, 2] (g, h) must be zeroed by the caller; only stored
 * CSR entries are visited (bins[q] = bin of stored entry q, rows sorted by column).  node_of[i] = node of
 * instance i (-1 = ignore); slot_of_node maps node id -> histogram slot (-1 = inactive). */
int hetu_gbdt_hist(const long long* indptr, const int* indices, const unsigned char* bins, const int* node_of,
                   const int* slot_of_node, const float* g, const float* h, const long long* bin_off,
                   const unsigned char* feat_on, float* hist, long long total_bins, long long n, cudaStream_t s);
/* move instances of split nodes to left_child[node] (bin <= split_bin[node]) or left_child[node] + 1 */
int hetu_gbdt_route(const long long* indptr, const int* indices, const unsigned char* bins, int* node_of,
                    const int* split_feat, const int* split_bin, const int* zero_bin, const int* left_child,
                    long long n, cudaStream_t s);

/* ---- fused MoE gate: reference python/hetu/layers/TopGate.py:15-79 lowered through cudnnSoftmax, TopKIdx.cu,
 * OneHot.cu, CumSum.cu, ReduceSum/ReduceMean (about 30 launches).  P planar [3k * T + 1 + E]: k planes of expert
 * indices, k of capacity slots, k of gate values, then the balance loss and the per-expert route counts.
 * Scratch: blk_cnt int[k * ceil(T / 256) * E], me float[E].  bwd: dgates = k pointers to [T] (null = no gradient). */
int hetu_moe_gate_fwd(const float* logits, float* P, int* blk_cnt, float* me, int T, int E, int k, cudaStream_t s);
int hetu_moe_gate_bwd(const float* logits, const float* P, void* const* dgates, const float* daux, float* dlogits,
                      int T, int E, int k, cudaStream_t s);

/* ---- steady-state replay of a captured step (csrc/step_plan.cu).  The captured graph starts with hetu_feed_select
 * (device staging slot s -> static input buffer, s advances on the device) and ends with hetu_result_publish (small
 * result -> device ring slot s); hetu_step_plan_run then overlaps the H2D of the next step and the D2H of the last
 * result with the graph: host slices -> pinned slot -> H2D on a copy stream -> cudaGraphLaunch -> D2H on a third stream. */
int hetu_feed_select(const void* stage, void* dst, long long total, int nslots, int* state, cudaStream_t s);
int hetu_result_publish(const float* src, float* ring, int n, const int* state, cudaStream_t s);
void* hetu_step_plan_create(void* graph_exec, cudaStream_t stream, int n, const long long* off, const long long* nbytes,
                            void* const* host_slots, int nslots, void* stage, int dslots, long long total, int dturn0);
int hetu_step_plan_set_readback(void* plan, const void* dev_ring, long long nbytes, void* const* host_slots, int nslots);
int hetu_step_plan_run(void* plan, void* const* src);
int hetu_step_plan_wait(void* plan, int slot);
int hetu_step_plan_ready(void* plan, int slot);
int hetu_step_plan_dturn(void* plan);
void hetu_step_plan_destroy(void* plan);

#ifdef __cplusplus
}
#endif
#endif /* HETU_B200_H_ */
