"""Parameter-server communication ops.

reference: python/hetu/gpu_ops/ParameterServerCommunicate.py:38-330.  A gradient
that goes to the PS is scaled by -lr on the worker (the server only does ``+=``),
copied D2H into a pinned buffer on the d2h stream and pushed; the fresh value is
pulled back (dense: DDPushPull; sparse: SparsePush + SparsePull / SSPushPull; with
``cstable_policy`` through the HET cache).  ``bsp`` = -1 ASP, 0 BSP (barrier),
k > 0 SSP with staleness k.  These ops touch the host every step, so graphs that
contain them run eagerly (``graph_safe = False``).
"""
from __future__ import annotations

import numpy as np
import torch

from ..graph.node import Op
from ..ndarray import IndexedSlices, NDArray


class ParameterServerCommunicateOp(Op):
    graph_safe = False

    def __init__(self, nodeA, parameter, optimizer):
        super().__init__(ParameterServerCommunicateOp, [nodeA], None)
        self.parameter = parameter
        self.optimizer = optimizer          # (kind, [lr, hyper...])
        self.use_indexed_slices = nodeA.use_indexed_slices
        self.comm = None
        self.cache = None
        self.bsp = -1
        self._ssp_inited = False

    @property
    def learning_rate(self):
        return float(self.optimizer[1][0])

    def forward_hook(self, config):
        self.ctx = config.context
        self.comm = config.ps_comm
        self.bsp = config.bsp
        self.config = config
        param = self.parameter
        width = param.shape[-1] if param.is_embed else 1
        if param.is_embed and config.cstable_policy is not None:
            import os
            limit = max(param.shape[0] // 10, 1)
            dev = config.context.torch_device
            if dev.type == 'cuda' and torch.cuda.is_available() and os.environ.get('HETU_HET_GPU', '1') == '1':
                # the cache lines live in HBM (csrc/het_cache.cu); only staleness checks, misses and pushes reach the host
                from ..parallel.gpu_cache import GPUHetCache
                self.cache = GPUHetCache(limit, param.shape[0], param.shape[1], param.id, config.cstable_policy,
                                         config.cache_bound, comm=self.comm, device=dev)
            else:
                from ..parallel.cstable import CacheSparseTable
                self.cache = CacheSparseTable(limit, param.shape[0], param.shape[1], param.id,
                                              config.cstable_policy, config.cache_bound, comm=self.comm)
            param.cache = self.cache
        if self.bsp > 0 and not self._ssp_inited:
            self.comm.ssp_init(self.comm.rank() // 1, self.comm.nrank(), self.bsp)
            self._ssp_inited = True

    def _sync(self):
        if self.bsp == 0:
            self.comm.BarrierWorker()
        elif self.bsp > 0:
            self.comm.ssp_sync(self.comm.rank(), self.config.ps_step if hasattr(self.config, 'ps_step') else 0)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        g = input_vals[0]
        param = self.parameter
        lr = self.learning_rate
        dev = self.config.context.torch_device
        if isinstance(g, IndexedSlices):
            ind, val = g._iv()
            if self.cache is not None and hasattr(self.cache, 'device_update'):
                self.cache.device_update(ind.detach(), val.detach() * (-lr))
                self._sync()
                return None
            ind_h = ind.detach().cpu().numpy().astype(np.int64)
            val_h = (val.detach() * (-lr)).cpu().numpy().astype(np.float32)
            if self.cache is not None:
                self.cache.embedding_update(ind_h, val_h, sync=True)
            else:
                self.comm.SparsePush(param.id, ind_h, val_h)
                self.comm.Wait(param.id)
            self._sync()
            return None
        g_h = (g.detach() * (-lr)).cpu().numpy().astype(np.float32).reshape(-1)
        fresh = self.comm.DDPushPull(param.id, g_h)
        self._sync()
        cur = self.config.placeholder_to_arr_map.get(param)
        fresh_t = torch.from_numpy(fresh.reshape(param.shape))
        if cur is None:
            self.config.placeholder_to_arr_map[param] = fresh_t.to(dev)
        else:
            cur.copy_(fresh_t.to(cur.device))
        return None

    def gradient(self, g):
        raise NotImplementedError

    def infer_shape(self, input_shapes):
        return None


class ParameterServerSparsePullOp(Op):
    """Pull the embedding rows that ``deps_node`` (the ids) will need."""
    graph_safe = False

    def __init__(self, parameter, deps_node):
        super().__init__(ParameterServerSparsePullOp, [deps_node], None)
        self.parameter = parameter
        self.comm = None

    def forward_hook(self, config):
        self.ctx = config.context
        self.comm = config.ps_comm
        self.config = config

    def compute(self, input_vals, output_val=None, stream_handle=None):
        ids = input_vals[0]
        ids_h = ids.detach().cpu().numpy().astype(np.int64).reshape(-1)
        cache = getattr(self.parameter, 'cache', None)
        if cache is not None:
            rows = cache.embedding_lookup(ids_h, sync=True)
        else:
            rows = self.comm.SparsePull(self.parameter.id, ids_h, self.parameter.shape[1])
        t = torch.from_numpy(rows.reshape(tuple(ids.shape) + (self.parameter.shape[1],)))
        return t.to(self.config.context.torch_device, non_blocking=True)

    def gradient(self, g):
        return [None]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0]) + (self.parameter.shape[1],)


def parameterServerCommunicate_op(node, parameter, optimizer):
    return ParameterServerCommunicateOp(node, parameter, optimizer)


def parameterServerSparsePull_op(parameter, deps_node):
    return ParameterServerSparsePullOp(parameter, deps_node)
