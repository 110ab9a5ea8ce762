#!/usr/bin/env python
"""Reference arm of bench.py: runs the UNMODIFIED reference (a copy of /root/reference under
baseline/_ref, built by baseline/build_reference.sh) through its own public API and stock code path.

This process never imports hetu_61a7_b200 and never maps libhetu_*.so: `hetu` here is
baseline/_ref/python/hetu, its native code is baseline/_ref/build/lib/*.so.

Workloads (the configs BASELINE.json names):
  wdl : examples/ctr/models/wdl_criteo.py driven like examples/ctr/run_hetu.py:120-135
        (N = 1: ht.Executor(eval_nodes, ctx=ht.gpu(0)); N > 1: same graph stated in the arm, table on the GPUs, comm_mode='AllReduce' - see bench_wdl)
  moe : examples/moe/test_moe_top.py:16-97 (TopKGate + MoELayer, AllToAll over NCCL)

Environment shims (none touches the reference tree): numpy 2.x removed np.int / np.float / np.bool which the
reference uses; `turtle` (Tk) and `boto3` are imported by the reference but never called on these paths.
Multi-rank rendezvous: baseline/mpi_stub implements the MPI calls of the reference's communicator over TCP and
takes RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT from torchrun's environment.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time
import types
import warnings

warnings.filterwarnings('ignore')
HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, '_ref')


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=200)
    ap.add_argument('--warmup', type=int, default=10)
    ap.add_argument('--workload', default='both', choices=['wdl', 'moe', 'both'])
    ap.add_argument('--batch', type=int, default=128)
    ap.add_argument('--moe-steps', type=int, default=0)
    ap.add_argument('--ctx', default='gpu', choices=['gpu', 'cpu'], help='cpu: plumbing check without a GPU')
    return ap.parse_args()


def shim_environment():
    import numpy as np
    for n, t in (('int', int), ('float', float), ('bool', bool)):
        if not hasattr(np, n):
            setattr(np, n, t)

    def stub(name, **attrs):
        if name not in sys.modules:
            m = types.ModuleType(name)
            m.__dict__.update(attrs)
            sys.modules[name] = m
    try:
        import turtle  # noqa: F401
    except Exception:
        stub('turtle', forward=None)
    try:
        import boto3  # noqa: F401
    except Exception:
        stub('boto3')
        stub('botocore')
        stub('botocore.exceptions', ClientError=type('ClientError', (Exception,), {}))
    sys.path.insert(0, os.path.join(REF, 'python'))


class Cuda(object):
    """cudart through ctypes: device sync and CUDA events on the legacy default stream (which orders after every
    blocking stream of the process - the reference creates its streams with cudaStreamCreate)."""

    def __init__(self):
        self.rt = ctypes.CDLL('libcudart.so.12', mode=ctypes.RTLD_GLOBAL)

    def sync(self):
        assert self.rt.cudaDeviceSynchronize() == 0

    def event(self):
        e = ctypes.c_void_p()
        assert self.rt.cudaEventCreate(ctypes.byref(e)) == 0
        return e

    def record(self, e):
        assert self.rt.cudaEventRecord(e, None) == 0

    def elapsed(self, a, b):
        ms = ctypes.c_float()
        assert self.rt.cudaEventElapsedTime(ctypes.byref(ms), a, b) == 0
        return ms.value


class ClockSampler(object):
    Q = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,'
         'clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, gpu_index=0):
        self.rows, self.proc, self.gpu_index = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ['nvidia-smi', '--query-gpu=' + self.Q, '--format=csv,noheader,nounits', '-lms', '50',
                 '-i', str(self.gpu_index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
            t0 = time.time()
            while not self.rows and time.time() - t0 < 5.0:
                time.sleep(0.01)
            self.rows = []
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(',')])

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        self.proc.terminate()
        sm = sorted(float(r[1]) for r in self.rows if len(r) > 2 and r[1].replace('.', '').isdigit())
        mx = [float(r[2]) for r in self.rows if len(r) > 2 and r[2].replace('.', '').isdigit()]
        reasons = set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for r in self.rows:
            for i, n in enumerate(names):
                if len(r) > 5 + i and r[5 + i].lower().startswith('active'):
                    reasons.add(n)
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'reasons': sorted(reasons), 'samples': len(self.rows)}


def criteo_synthetic(np, n, num_dense=13, num_sparse=26, vocab=33762577, seed=0, zipf_a=1.05):
    """Same generator as the other arm (hetu_61a7_b200/utils/data.py:criteo_synthetic), restated here so that this
    process imports nothing of the new framework."""
    rng = np.random.RandomState(seed)
    dense = rng.rand(n, num_dense).astype(np.float32)
    field = vocab // num_sparse
    ranks = rng.zipf(zipf_a, size=(n, num_sparse)).astype(np.int64)
    ids = (ranks - 1) % field + np.arange(num_sparse, dtype=np.int64)[None, :] * field
    labels = (rng.rand(n, 1) < 0.25).astype(np.float32)
    return dense, ids, labels


class Timer(object):
    def __init__(self, cuda, comm):
        self.cuda, self.comm = cuda, comm

    def barrier(self):
        self.max_over_ranks(0.0)              # an NCCL all-reduce on the communicator's stream + a stream sync

    def max_over_ranks(self, v):
        if self.comm is None or self.comm.nrank == 1:
            return v
        import hetu as ht
        import numpy as np
        from hetu.communicator.mpi_nccl_comm import ncclDataType_t, ncclRedOp_t
        if self.cuda is not None:
            self.cuda.sync()
        a = ht.array(np.array([v], dtype=np.float32), ctx=ht.gpu(self.comm.device_id.value))
        self.comm.dlarrayNcclAllReduce(a, a, ncclDataType_t.ncclFloat32, ncclRedOp_t.ncclMax)
        self.comm.stream.sync()
        return float(a.asnumpy()[0])

    def run(self, fn, steps, min_seconds=0.5, max_blocks=40):
        """Same rule as the other arm: blocks of exactly `steps` steps, each bracketed by a barrier and a device
        synchronize, repeated until >= 3 blocks and >= min_seconds; median block, max over ranks."""
        c = self.cuda
        per_step, total = [], 0.0
        while True:
            self.barrier()
            if c is None:
                t0 = time.perf_counter()
                for i in range(steps):
                    fn(i)
                ms = (time.perf_counter() - t0) * 1e3
            else:
                c.sync()
                a, b = c.event(), c.event()
                c.record(a)
                for i in range(steps):
                    fn(i)
                c.record(b)
                c.sync()
                self.barrier()
                ms = self.max_over_ranks(c.elapsed(a, b))
            per_step.append(ms / steps)
            total += ms
            if (len(per_step) >= 3 and total >= min_seconds * 1e3) or len(per_step) >= max_blocks:
                break
        per_step.sort()
        return per_step[len(per_step) // 2]


def bench_wdl(args, ht, np, comm, cuda, world, rank, local_rank):
    sys.path.insert(0, os.path.join(REF, 'examples', 'ctr'))
    from models.wdl_criteo import wdl_criteo       # the reference's own model file
    B = args.batch
    npool = 32
    pool = [criteo_synthetic(np, B, seed=1000 * rank + i) for i in range(npool)]
    dense = np.concatenate([p[0] for p in pool])
    sparse = np.concatenate([p[1] for p in pool])
    labels = np.concatenate([p[2] for p in pool])
    dense_input = ht.dataloader_op([[dense, B, 'train']])
    sparse_input = ht.dataloader_op([[sparse, B, 'train']])
    y_ = ht.dataloader_op([[labels, B, 'train']])
    if world == 1:
        loss, prediction, y_, train_op = wdl_criteo(dense_input, sparse_input, y_)
    else:
        from hetu import init

        def wdl_criteo_gpu_table(dense_input, sparse_input, y_):
            feature_dimension, embedding_size, learning_rate = 33762577, 128, 0.01
            Embedding = init.random_normal([feature_dimension, embedding_size], stddev=0.01, name="snd_order_embedding")
            sparse = ht.embedding_lookup_op(Embedding, sparse_input)
            sparse = ht.array_reshape_op(sparse, (-1, 26 * embedding_size))
            W1 = init.random_normal([13, 256], stddev=0.01, name="W1")
            W2 = init.random_normal([256, 256], stddev=0.01, name="W2")
            W3 = init.random_normal([256, 256], stddev=0.01, name="W3")
            W4 = init.random_normal([256 + 26 * embedding_size, 1], stddev=0.01, name="W4")
            relu1 = ht.relu_op(ht.matmul_op(dense_input, W1))
            relu2 = ht.relu_op(ht.matmul_op(relu1, W2))
            y3 = ht.matmul_op(relu2, W3)
            y = ht.sigmoid_op(ht.matmul_op(ht.concat_op(sparse, y3, axis=1), W4))
            loss = ht.reduce_mean_op(ht.binarycrossentropy_op(y, y_), [0])
            train_op = ht.optim.SGDOptimizer(learning_rate=learning_rate).minimize(loss)
            return loss, y, y_, train_op
        loss, prediction, y_, train_op = wdl_criteo_gpu_table(dense_input, sparse_input, y_)
    eval_nodes = {'train': [loss, prediction, y_, train_op]}
    if world == 1:
        ctx = ht.gpu(0) if args.ctx == 'gpu' else ht.cpu(0)
        executor = ht.Executor(eval_nodes, ctx=ctx, seed=123)                       # run_hetu.py:120-122 (--comm None)
    else:
        # N > 1.  The reference ships multi-worker W&D only in PS / Hybrid mode (examples/ctr/tests/{ps,hybrid}_wdl_criteo.sh),
        # which needs ps-lite (ZeroMQ + protobuf: neither can be had offline).  Its AllReduce mode does not run with the
        # stock model file: wdl_criteo.py pins the table with ctx=ht.cpu(0), the gradient node inherits that as a
        # DeviceGroup and OptimizerOp.backward_hook -> AllReduceCommunicateOp.__init__ fails in is_gpu_ctx
        # ("'DeviceGroup' object has no attribute 'device_type'", python/hetu/gpu_ops/AllReduceCommunicate.py:10) before
        # forward_hook would have resolved it.  So for N > 1 the arm states the same graph (wdl_criteo.py:8-42) itself,
        # without the two ctx=ht.cpu(0) arguments: table replicated on every GPU, sparse gradients all-gathered by the
        # reference's AllReduceCommunicateOp, everything else as in the model file.  Kernels, executor, NCCL path: the
        # reference's own.
        comm = ht.wrapped_mpi_nccl_init()
        executor = ht.Executor(eval_nodes, ctx=ht.gpu(comm.local_rank), comm_mode='AllReduce', seed=123)
    timer = Timer(cuda, executor.config.nccl_comm if world > 1 else None)
    last = {}

    def step_dev(i):
        executor.run('train')                                                       # results stay on the device

    def step_e2e(i):
        out = executor.run('train', convert_to_numpy_ret_vals=True)                 # run_hetu.py:20-21
        last['loss'] = out[0]

    W = max(args.warmup, 3)
    for i in range(W):
        step_e2e(i)
    sampler = ClockSampler(local_rank)
    sampler.start()
    ms_dev = timer.run(step_dev, args.steps)
    ms_e2e = timer.run(step_e2e, args.steps)
    clocks = sampler.stop()
    h2d = pool[0][0].nbytes + pool[0][1].astype(np.float32).nbytes + pool[0][2].nbytes
    return {'ms_dev': ms_dev, 'ms_e2e': ms_e2e, 'clocks': clocks, 'h2d': h2d,
            'd2h': 4 + 2 * B * 4, 'global_batch': B * world, 'loss': float(np.asarray(last['loss']).reshape(-1)[0])}


def bench_moe(args, ht, np, cuda, world, rank, local_rank, steps):
    from hetu import layers as htl
    bs, ntok, dim, hid, nle, top = 16, 1024, 2048, 2048, 2, 2
    comm = ht.wrapped_mpi_nccl_init()
    device_id = comm.local_rank
    ctx = ht.gpu(device_id % 8) if args.ctx == 'gpu' else ht.cpu(0)
    # examples/moe/test_moe_top.py:16-38, with num_gpus = world
    # Naming: the reference script calls its experts "expert_<i>".  In this snapshot OptimizerOp.backward_hook
    # (python/hetu/optimizer.py:149-151) drops the gradient of every parameter whose name contains "expert" from
    # the optimizer's inputs, and Optimizer.update (optimizer.py:187) then fails `assert params_size == len(grads)`
    # on the first step - the script cannot train as written.  The arm therefore only picks another name; the
    # reference code itself stays untouched.  Consequence at N > 1: the expert weights take the reference's
    # ordinary AllReduce path (4 x 16.8 MB per step) like every other dense parameter.
    experts = [htl.Expert(embed_dim=dim, ffn_dim=hid, dropout_rate=0.1, activation='relu',
                          name='ffn_%d' % (device_id * nle + i)) for i in range(nle)]
    gate = htl.TopKGate(embed_dim=dim, num_tokens=bs * ntok, num_experts=nle * world, k=top)
    model = htl.MoELayer(gate=gate, experts=experts, num_tokens=ntok, embed_dim=dim, all2all_size=world, top=top)
    x = ht.Variable(name='x', ctx=ctx, trainable=False)
    y_ = ht.Variable(name='y_', ctx=ctx, trainable=False)
    y, l_aux = model(x)
    y = ht.array_reshape_op(y, [-1, ntok, dim])
    y = ht.reduce_sum_op(y, axes=2)
    y = ht.softmax_op(y)
    loss = ht.nll_loss_op(y, y_, ntok)
    loss = loss + l_aux
    opt = ht.optim.SGDOptimizer(learning_rate=0.125)
    train_op = opt.minimize(loss)
    executor = ht.Executor({'train': [loss, y, y_, train_op]}, ctx=ctx, comm_mode='AllReduce')
    rng = np.random.RandomState(rank)
    x_np = rng.normal(size=(bs, ntok, dim)).astype(np.float32)
    x_val = ht.array(arr=x_np, ctx=ctx)
    y_np = np.zeros((bs,), dtype=np.float32)
    y_val = ht.array(arr=y_np, ctx=ctx)
    timer = Timer(cuda, comm if world > 1 else None)
    last = {}

    def step_dev(i):                # test_moe_top.py:83-90: inputs resident on the GPU, loss read every step
        out = executor.run('train', eval_node_list=[loss, y, y_, train_op], feed_dict={x: x_val, y_: y_val})
        last['loss'] = out[0].asnumpy()

    def step_e2e(i):                # the step's batch comes from host memory (numpy feed -> H2D by the executor)
        out = executor.run('train', eval_node_list=[loss, y, y_, train_op], feed_dict={x: x_np, y_: y_np})
        last['loss'] = out[0].asnumpy()

    for i in range(max(3, min(args.warmup, 4))):
        step_dev(i)
    ms_dev = timer.run(step_dev, steps)
    for i in range(2):
        step_e2e(i)
    ms_e2e = timer.run(step_e2e, steps)
    return {'ms_dev': ms_dev, 'ms_e2e': ms_e2e, 'tokens': bs * ntok * world,
            'h2d': x_np.nbytes + y_np.nbytes, 'd2h': 4, 'loss': float(np.asarray(last['loss']).reshape(-1)[0])}


def main():
    args = parse()
    lib = os.path.join(REF, 'build', 'lib', 'libc_runtime_api.so')
    if not os.path.exists(lib):
        print(json.dumps({'impl': 'reference', 'unavailable': 'baseline/_ref/build/lib/libc_runtime_api.so is not '
                          'built (run baseline/build_reference.sh)'}))
        return 0
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    os.environ.setdefault('OMP_NUM_THREADS', str(max(1, (os.cpu_count() or 8) // max(world, 1))))
    shim_environment()
    import numpy as np
    import hetu as ht
    cuda = None
    if args.ctx == 'gpu':
        cuda = Cuda()
        assert cuda.rt.cudaSetDevice(local_rank) == 0
    out = {'impl': 'reference', 'metric': 'wdl_criteo_samples_per_s', 'unit': 'samples/s', 'n_gpus': world,
           'steps': args.steps, 'warmup': max(args.warmup, 3), 'higher_is_better': True, 'scaling': 'weak',
           'vs_baseline': None, 'dtype': 'fp32 (cuBLAS SGEMM, the reference default)',
           'data': 'synthetic (Criteo-shaped zipf ids; random-init weights)',
           'reference_build': 'baseline/build_reference.sh: unmodified src/{common,cuda_common,ops} + plain-C++ '
                              'files of src/dnnl_ops, sm_100; MPI through baseline/mpi_stub'}
    if args.workload in ('wdl', 'both'):
        w = bench_wdl(args, ht, np, None, cuda, world, rank, local_rank)
        out.update({
            'value': w['global_batch'] / (w['ms_dev'] / 1e3), 'ms_per_step': w['ms_dev'],
            'e2e': {'value': w['global_batch'] / (w['ms_e2e'] / 1e3), 'unit': 'samples/s', 'ms_per_step': w['ms_e2e'],
                    'h2d_bytes_per_step': w['h2d'], 'd2h_bytes_per_step': w['d2h']},
            'clocks': w['clocks'],
            'config': {'model': 'Wide&Deep (13 dense + 26 sparse, 33762577 x 128 fp32 table, MLP 13-256-256-256)',
                       'global_batch': w['global_batch'], 'seq_len': None,
                       'parallelism': 'dp%d' % world if world > 1 else 'single',
                       'embedding': ('host table (ctx=ht.cpu(0) in examples/ctr/models/wdl_criteo.py), C++ lookup + '
                                     'sparse SGD on the CPU') if world == 1 else
                                    'replicated on every GPU (graph of wdl_criteo.py without its ctx=ht.cpu(0) pin, which crashes the '
                                    'reference in AllReduce mode), sparse grads all-gathered over NCCL',
                       'optimizer': 'SGD lr 0.01', 'final_loss': w['loss'],
                       'l2_policy': 'working set (17.3 GB table, random rows) exceeds L2'},
        })
    if args.workload in ('moe', 'both'):
        steps = args.moe_steps or args.steps
        m = bench_moe(args, ht, np, cuda, world, rank, local_rank, steps)
        moe_out = {'metric': 'moe_tokens_per_s', 'value': m['tokens'] / (m['ms_dev'] / 1e3), 'unit': 'tokens/s',
                   'ms_per_step': m['ms_dev'], 'steps': steps,
                   'e2e': {'value': m['tokens'] / (m['ms_e2e'] / 1e3), 'unit': 'tokens/s', 'ms_per_step': m['ms_e2e'],
                           'h2d_bytes_per_step': m['h2d'], 'd2h_bytes_per_step': m['d2h']},
                   'final_loss': m['loss'],
                   'config': {'model': 'MoE layer top-2, 16x1024 tokens/GPU, d=h=2048, 2 experts/GPU',
                              'global_tokens': m['tokens'], 'parallelism': 'ep%d' % world}}
        if args.workload == 'moe':
            out.update({'metric': 'moe_tokens_per_s', 'value': moe_out['value'], 'unit': 'tokens/s',
                        'ms_per_step': m['ms_dev'], 'config': moe_out['config'], 'e2e': moe_out['e2e']})
        else:
            out['moe'] = moe_out
    if rank == 0:
        print(json.dumps(out))
        sys.stdout.flush()
    os._exit(0)          # the reference's executors own worker threads / NCCL comms that do not always unwind cleanly


if __name__ == '__main__':
    sys.exit(main())
