"""Variables / placeholders.  reference: python/hetu/gpu_ops/Variable.py:8-140."""
from __future__ import annotations

import numpy as np
import torch

from ..graph.node import Op
from ..ndarray import NDArray, DLContext, cpu, to_tensor
from .. import stream as _stream


class PlaceholderOp(Op):
    """A leaf: a trainable parameter (value or initializer given), a constant
    (``trainable=False`` with a value) or a feed (nothing given)."""

    def __init__(self, name, value=None, initializer=None, trainable=True, dtype=np.float32, ctx=None):
        super().__init__(PlaceholderOp, [], ctx)
        self.name = name
        self.is_embed = False
        self.shape = None
        if value is None and initializer is None:
            trainable = False
        elif value is not None:
            assert initializer is None, 'Value already specified, initializer should be None.'
            assert isinstance(value, (np.ndarray, NDArray, torch.Tensor)), \
                'Value data type %s not valid.' % str(type(value))
            self.shape = tuple(value.shape)
        else:
            self.shape = tuple(initializer.shape)
        self.tensor_value = value
        self.initializer = initializer
        self.trainable = trainable
        self.dtype = dtype
        self.reshaped = False
        self.embedding_offsets = None
        self.parts = {}
        self.cur_part = {}
        self.ps_resident = False
        self.sharded = False

    @property
    def is_feed(self) -> bool:
        return self.tensor_value is None and self.initializer is None and not self.ps_resident

    def compute(self, input_vals, output_val=None, stream_handle=None):
        assert self.shape, "placeholder %s values provided by feed_dict" % self.name
        return self.tensor_value.tensor if isinstance(self.tensor_value, NDArray) else self.tensor_value

    def gradient(self, output_grad):
        return None

    def infer_shape(self, input_shapes):
        assert self.shape, "placeholder %s shape provided by feed_shape" % self.name
        return tuple(self.shape)

    def forward_hook(self, config):
        pass

    def backward_hook(self, config):
        """Materialise the value on this process's device - or hand the
        parameter to the parameter server / sharded embedding tier."""
        if not isinstance(self.ctx, DLContext):
            self.resolve_ctx(config)
        strategy = config.node_strategy.get(self, config.comm_mode)
        on_ps = self.trainable and (strategy == 'PS' or (strategy == 'Hybrid' and self.is_embed))
        if on_ps and config.ps_comm is not None:
            self._init_on_ps(config)
            return
        from ..parallel.sharded_embedding import should_shard, ShardedTable
        if should_shard(self, config) and self.shape is not None and len(self.shape) == 2:
            table = ShardedTable(config.symm, self.shape, self.initializer, config.seed + self.id,
                                 name='%s_%d' % (self.name, self.id), bsp=config.bsp)
            if self.initializer is None and self.tensor_value is not None:
                tv = self.tensor_value
                table.import_dense(tv.asnumpy() if isinstance(tv, NDArray) else tv)
            self.initializer = None
            self.tensor_value = table
            self.sharded = True
            return
        if self.initializer is not None:
            self.initializer(self, config.seed, config.np_rand, config.comp_stream)
            self.initializer = None
        elif self.tensor_value is not None:
            value = self.tensor_value
            dev = self.ctx.torch_device
            if isinstance(value, NDArray):
                t = value.tensor
            else:
                t = to_tensor(value)
            if t.dtype == torch.float64:
                t = t.float()
            if t.device != dev:
                t = t.to(dev)
            elif not isinstance(value, NDArray):
                t = t.clone()
            self.tensor_value = NDArray(t, self.ctx)

    def _init_on_ps(self, config):
        """Hand the parameter to the servers: ParamInit on every shard (server-side RNG), or
        zero-init + a push of the user's value from rank 0.  Dense parameters keep a device
        replica refreshed by DDPushPull; embedding tables stay remote and are read row-wise
        (SparsePull / HET cache).  reference: executor.py:1191-1217, initializers.py:28-38."""
        comm = config.ps_comm
        self.ps_resident = True
        ptype = 0 if not self.is_embed else (2 if config.cstable_policy is not None else 1)
        if ptype == 0:
            length, width = int(np.prod(self.shape)), 1
        else:
            length, width = int(self.shape[0]), int(np.prod(self.shape[1:]))
        if self.initializer is not None:
            self.initializer.init_on_ps(comm, self.id, ptype, seed=config.seed + self.id)
            self.initializer = None
        else:
            comm.InitTensor(self.id, ptype, length, width, 0, 0.0, 1.0, 0)
            if comm.rank() == 0:
                v = self.tensor_value
                v = v.asnumpy() if isinstance(v, NDArray) else (v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else np.asarray(v))
                if ptype == 0:
                    comm.Push(self.id, v.astype(np.float32).reshape(-1))
                else:
                    comm.SparsePush(self.id, np.arange(length), v.astype(np.float32).reshape(length, width))
            comm.BarrierWorker()
        self.event = (_stream.CSEvent(comm, self.id) if ptype == 2 else _stream.PSEvent(comm, self.id))
        if self.is_embed:
            self.tensor_value = None
        else:
            flat = comm.Pull(self.id)
            t = torch.from_numpy(flat.reshape(self.shape).copy()).to(self.ctx.torch_device)
            self.tensor_value = NDArray(t, self.ctx)

    # -- model-parallel sharding (reference Variable.py:83-123) ----------------
    def reshape_in_mp(self, cur_part, parts):
        self.cur_part, self.parts = dict(cur_part), dict(parts)
        if self.reshaped:
            return
        self.reshaped = True
        if self.shape is None:
            return
        ori_shape = list(self.shape)
        for i, pts in self.parts.items():
            assert ori_shape[i] % pts == 0
            ori_shape[i] //= pts
        self.full_shape = tuple(self.shape)
        self.shape = tuple(ori_shape)
        if self.initializer is not None:
            self.initializer.set_shard(self.full_shape, self.cur_part, self.parts)
        elif self.tensor_value is not None:
            self.tensor_value = self.reshape_tensor(self.tensor_value)
            assert tuple(self.shape) == tuple(self.tensor_value.shape)

    def reshape_tensor(self, tensor):
        if isinstance(tensor, NDArray):
            tensor = tensor.asnumpy()
        elif isinstance(tensor, torch.Tensor):
            tensor = tensor.detach().cpu().numpy()
        if self.embedding_offsets is not None:
            offset, length = self.embedding_offsets
            tensor = tensor - offset
            tensor[tensor < 0] = -1
            tensor[tensor >= length] = -1
        if not self.parts:
            return tensor
        slcs = []
        for i, n in enumerate(tensor.shape):
            if i in self.parts:
                part = n // self.parts[i]
                st = part * self.cur_part.get(i, 0)
                slcs.append(slice(st, st + part))
            else:
                slcs.append(slice(0, n))
        return np.ascontiguousarray(tensor[tuple(slcs)])


def Variable(name, value=None, initializer=None, trainable=True, dtype=np.float32, ctx=None):
    return PlaceholderOp(name, value, initializer, trainable, dtype, ctx)


def placeholder_op(name, value=None, initializer=None, trainable=True, dtype=np.float32, ctx=None):
    return PlaceholderOp(name, value, initializer, trainable, dtype, ctx)
