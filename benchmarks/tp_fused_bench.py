"""Fused TP kernels vs the NCCL baseline (all_gather + GEMM, GEMM + reduce_scatter), BERT-large shapes.
Run: torchrun --nproc-per-node N benchmarks/tp_fused_bench.py   (device-timed, max over ranks)"""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import hetu as ht
from hetu_61a7_b200.parallel.comm import device_max_ms
from hetu_61a7_b200.parallel import mp_lowering as mp
from hetu_61a7_b200.ops.linalg import gemm

comm = ht.wrapped_mpi_nccl_init()
w, r = comm.nrank, comm.rank
dev = torch.device('cuda', torch.cuda.current_device())
group = mp._MPComm([ht.gpu(i) for i in range(w)])
tokens, hidden, ffn = 32 * 128 * max(1, w // 2), 1024, 4096           # BERT-large, batch 32 x seq 128 per 2 GPUs


def timeit(fn, iters=20):
    for _ in range(3):
        fn()
    comm.barrier(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record(); torch.cuda.synchronize()
    return device_max_ms(s.elapsed_time(e) / iters)


x_shard = torch.randn(tokens // w, hidden, device=dev)
w_col = torch.randn(hidden, ffn // w, device=dev) * 0.02
a, b = ht.Variable('a'), ht.Variable('b')
ag = mp.MPAllGatherMatMulOp(a, b, False, w, group)
gathered = torch.empty(tokens, hidden, device=dev)


def nccl_ag():
    dist.all_gather_into_tensor(gathered, x_shard)
    return gemm(gathered, w_col)


t_f = timeit(lambda: ag.compute([x_shard, w_col]))
t_n = timeit(nccl_ag)
t_g = timeit(lambda: gemm(gathered, w_col))
ref = nccl_ag(); got = ag.compute([x_shard, w_col])
err = float((got - ref).abs().max() / ref.abs().max())
flops = 2.0 * tokens * hidden * (ffn // w)
res = {'world': w, 'ag_gemm': {'M': tokens, 'K': hidden, 'N': ffn // w, 'fused_ms': t_f, 'nccl_ms': t_n, 'gemm_only_ms': t_g,
                               'fused_tflops': flops / t_f / 1e9, 'relerr_vs_nccl': err, 'fused_kernel': ag.fused_launches > 0}}

h_local = torch.randn(tokens, ffn // w, device=dev) * 0.1
w_row = torch.randn(ffn // w, hidden, device=dev) * 0.02
rs = mp.MPMatMulReduceScatterOp(a, b, False, False, w, group)
out_shard = torch.empty(tokens // w, hidden, device=dev)


def nccl_rs():
    full = gemm(h_local, w_row)
    dist.reduce_scatter_tensor(out_shard, full)
    return out_shard


t_f = timeit(lambda: rs.compute([h_local, w_row]))
t_n = timeit(nccl_rs)
t_g = timeit(lambda: gemm(h_local, w_row))
ref = nccl_rs().clone(); got = rs.compute([h_local, w_row])
err = float((got - ref).abs().max() / ref.abs().max())
flops = 2.0 * tokens * hidden * (ffn // w)
res['gemm_rs'] = {'M': tokens, 'K': ffn // w, 'N': hidden, 'fused_ms': t_f, 'nccl_ms': t_n, 'gemm_only_ms': t_g,
                  'fused_tflops': flops / t_f / 1e9, 'relerr_vs_nccl': err, 'fused_kernel': rs.fused_launches > 0}
if r == 0:
    print(json.dumps(res), flush=True)
