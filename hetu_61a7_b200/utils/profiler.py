"""Profilers.  reference: python/hetu/profiler.py:55-606.

``HetuProfiler`` micro-benchmarks every op of a graph with synthetic inputs (a
cost model for strategy search); ``NCCLProfiler`` tabulates collective latency
over device subsets.  Timing is on the device (CUDA events) where a GPU exists."""
from __future__ import annotations

import time
from typing import Dict, List, Optional

import numpy as np
import torch

from ..ndarray import DLContext, cpu, gpu


def _timeit(fn, device, warmup, iters):
    if device.type == 'cuda':
        for _ in range(warmup):
            fn()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(device)
        s.record()
        for _ in range(iters):
            fn()
        e.record()
        e.synchronize()
        return s.elapsed_time(e) / iters
    for _ in range(warmup):
        fn()
    t0 = time.perf_counter()
    for _ in range(iters):
        fn()
    return (time.perf_counter() - t0) * 1000.0 / iters


class NCCLOP(object):
    """reference profiler.py:18-23 (an Enum there; plain ints keep ``NCCLOP.AllReduce.value``-free call sites simple)."""
    AllReduce, AllGather, ReduceScatter, Reduce, Broadcast = range(5)
    names = ('allreduce', 'allgather', 'reducescatter', 'reduce', 'broadcast')


class BaseProfiler(object):
    """Shared device / timing plumbing of the profilers (reference profiler.py:26-52)."""

    def __init__(self):
        self.idx = 0
        self.ctx = None
        self.use_gpu = False

    def set_ctx(self, ctx):
        from ..ndarray import is_gpu_ctx
        self.ctx = ctx
        self.use_gpu = is_gpu_ctx(ctx) and torch.cuda.is_available()
        self.device = ctx.torch_device if self.use_gpu else torch.device('cpu')

    def renew_instances(self):
        raise NotImplementedError

    def init_memory(self, shape, seed=0):
        g = torch.Generator().manual_seed(int(seed))
        return torch.randn(tuple(shape), generator=g).to(getattr(self, 'device', 'cpu'))

    def time_fn(self, fn, warmup=5, iters=100):
        return _timeit(fn, getattr(self, 'device', torch.device('cpu')), warmup, iters)


class HetuProfiler(object):
    def __init__(self, computing_nodes, feed_shapes, node_to_shape_map, ctx=None, rank=-1,
                 warmup=5, iters=100):
        self.computing_nodes = list(computing_nodes)
        self.feed_shapes = feed_shapes
        self.node_to_shape_map = node_to_shape_map
        self.ctx = ctx if ctx is not None else (gpu(0) if torch.cuda.is_available() else cpu(0))
        self.device = self.ctx.torch_device
        self.warmup, self.iters = warmup, iters
        self.timer: Dict = {}

    def _synth(self, node, shape, consumer):
        """Synthetic input: ids for index-like inputs (zipf), N(0,1) otherwise."""
        from ..ops.embedding import EmbeddingLookUp
        if shape is None:
            return None
        if isinstance(consumer, EmbeddingLookUp) and node is consumer.inputs[1]:
            vocab = self.node_to_shape_map.get(consumer.inputs[0], (1000,))[0]
            ids = (np.random.zipf(1.2, size=shape) - 1) % vocab
            return torch.from_numpy(ids.astype(np.float32)).to(self.device)
        return torch.randn(tuple(shape), device=self.device)

    def profile(self, profiler='gpu'):
        from ..optimizer import OptimizerOp
        for node in self.computing_nodes:
            if isinstance(node, OptimizerOp) or not getattr(node, 'graph_safe', True):
                continue
            shapes = [self.node_to_shape_map.get(i) or self.feed_shapes.get(i) for i in node.inputs]
            if any(s is None for s in shapes):
                continue
            try:
                ins = [self._synth(i, s, node) for i, s in zip(node.inputs, shapes)]
                node_ctx, node.ctx = node.ctx, node.ctx if isinstance(node.ctx, DLContext) else self.ctx
                kw = {'inference': True} if getattr(node, 'takes_inference', False) else {}
                self.timer[node] = _timeit(lambda: node.compute(ins, None, None, **kw), self.device,
                                           self.warmup, self.iters)
                node.ctx = node_ctx
            except Exception:                      # op needs state the synthetic feed lacks
                continue
        return self.timer

    def profile_n_log(self, log_file, profiler='gpu'):
        self.profile(profiler)
        with open(log_file, 'w') as fw:
            for node, ms in self.timer.items():
                print(node.name, node.op_type, '%.6f ms' % ms, file=fw)
        return self.timer


class NCCLProfiler(object):
    """Latency tables for allreduce / allgather / reducescatter / reduce / broadcast /
    sendrecv over device subsets."""

    def __init__(self):
        from ..parallel.comm import world
        self.comm = world()
        self.device = (torch.device('cuda', torch.cuda.current_device())
                       if torch.cuda.is_available() else torch.device('cpu'))

    def _run(self, prim, numel, comm, warmup=3, iters=10, **kw):
        x = torch.ones(numel, device=self.device)
        n = comm.nrank
        if prim == 'allreduce':
            fn = lambda: comm.allreduce(x)
        elif prim == 'allgather':
            out = torch.empty(numel * n, device=self.device)
            fn = lambda: comm.allgather(x, out)
        elif prim == 'reducescatter':
            big = torch.ones(numel * n, device=self.device)
            fn = lambda: comm.reducescatter(big, x)
        elif prim == 'reduce':
            fn = lambda: comm.reduce(x, kw.get('root', 0))
        elif prim == 'broadcast':
            fn = lambda: comm.broadcast(x, kw.get('root', 0))
        elif prim == 'alltoall':
            out = torch.empty_like(x)
            fn = lambda: comm.alltoall(x, out)
        else:
            raise ValueError(prim)
        ms = _timeit(fn, self.device, warmup, iters)
        from ..parallel.comm import device_max_ms
        return device_max_ms(ms)

    def profile(self, prim, sizes, ranks: Optional[List[int]] = None, **kw):
        from ..parallel.comm import new_group_comm
        comm = self.comm if ranks is None else new_group_comm(list(ranks))
        out = {}
        if ranks is not None and self.comm.world_rank not in ranks:
            return out
        for numel in sizes:
            out[numel] = self._run(prim, int(numel), comm, **kw)
        return out

    def profile_allreduce(self, size, devices=None, num_iters=10):
        return self.profile('allreduce', [size], devices)[size] if True else None

    def profile_sendrecv(self, size, devices, num_iters=10):
        a, b = devices
        x = torch.ones(int(size), device=self.device)
        comm = self.comm
        if comm.world_rank not in (a, b):
            return None

        def fn():
            if comm.world_rank == a:
                comm.send(x, b)
            else:
                comm.recv(x, a)
        return _timeit(fn, self.device, 2, num_iters)
