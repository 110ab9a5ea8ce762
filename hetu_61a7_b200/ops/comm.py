"""Communication ops as first-class graph nodes.

reference op files: AllReduceCommunicate AllGatherCommunicate
ReduceScatterCommunicate BroadcastCommunicate ReduceCommunicate AllToAll
HAllToAll PipelineSend PipelineReceive DataTransfer Dispatch
(python/hetu/gpu_ops/*.py; every collective there is a host-launched NCCL call
on a side stream, executor.py:654-668).

Each op has two engines:
 * ``nccl``  - ``torch.distributed`` (baseline; also the only engine for
   pipeline send/recv, which BASELINE.json keeps on NCCL p2p);
 * ``symm``  - the in-kernel NVLink path over the symmetric heap
   (parallel/symm.py, csrc/collectives.cu): peers are read / written with plain
   ``ld.global`` / ``st.global`` on IPC-mapped pointers and ordered by
   system-scope flags, so the op is one kernel and CUDA-graph capturable.
The engine is picked per executor (``HetuConfig.use_nccl_collectives``).
"""
from __future__ import annotations

import numpy as np
import torch

from ..graph.node import Op
from ..ndarray import IndexedSlices


def _comm_of(op):
    from ..parallel.comm import world
    return op.comm if getattr(op, 'comm', None) is not None else world()


def _symm_engine(op):
    """The symmetric-heap engine spans the WORLD; a collective over a sub-group (data-parallel replicas of
    one pipeline stage, a tensor-parallel worker) has to stay on its own NCCL communicator."""
    cfg = getattr(op, '_config', None)
    eng = getattr(cfg, 'symm', None) if cfg is not None else None
    if eng is None:
        return None
    comm = getattr(op, 'comm', None)
    if comm is not None and (comm.nrank != eng.world or getattr(comm, 'group', None) is not None):
        return None
    return eng


class _CommOp(Op):
    def forward_hook(self, config):
        super().forward_hook(config)
        self._config = config
        if getattr(self, 'comm', None) is None:
            self.comm = config.nccl_comm

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


class AllReduceCommunicateOp(_CommOp):
    """Sum-all-reduce of a gradient (no averaging, like the reference:
    AllReduceCommunicate.py:15-44).  IndexedSlices are all-gathered."""

    def __init__(self, nodeA, comm=None, ctx=None):
        super().__init__(AllReduceCommunicateOp, [nodeA], ctx)
        self.comm = comm
        self.use_indexed_slices = nodeA.use_indexed_slices

    def compute(self, input_vals, output_val=None, stream_handle=None):
        v = input_vals[0]
        comm = _comm_of(self)
        if isinstance(v, IndexedSlices):
            ind, val = v._iv()
            if comm.multi:
                ind = comm.allgather(ind.contiguous())
                val = comm.allgather(val.contiguous())
            return IndexedSlices(ind, val, v.dense_shape)
        if v.device.type == 'meta' or not comm.multi:
            return v
        eng = _symm_engine(self)
        if eng is not None and v.is_cuda and v.dtype == torch.float32:
            return eng.allreduce(v, key=('op', self.id))
        out = v.clone()
        comm.allreduce(out)
        return out

    def gradient(self, g):
        return [allreduceCommunicate_op(g, self.comm)]


class AllReduceCommunicateP2POp(AllReduceCommunicateOp):
    def __init__(self, nodeA, comm=None, ctx=None):
        super().__init__(nodeA, comm, ctx)
        self.op_type = 'AllReduceCommunicateP2POp'


class GroupAllReduceCommunicateOp(AllReduceCommunicateOp):
    def __init__(self, nodeA, group_comm, ctx=None):
        super().__init__(nodeA, group_comm, ctx)
        self.op_type = 'GroupAllReduceCommunicateOp'


class AllGatherCommunicateOp(_CommOp):
    """Concatenate every rank's tensor along dim 0."""

    def __init__(self, nodeA, comm=None, ctx=None):
        super().__init__(AllGatherCommunicateOp, [nodeA], ctx)
        self.comm = comm

    def compute(self, input_vals, output_val=None, stream_handle=None):
        v = input_vals[0]
        comm = _comm_of(self)
        if v.device.type == 'meta':
            return torch.empty((v.shape[0] * comm.nrank,) + tuple(v.shape[1:]), device='meta')
        eng = _symm_engine(self)
        if eng is not None and comm.multi and v.is_cuda and v.dtype == torch.float32:
            return eng.allgather(v, key=('op', self.id))
        return comm.allgather(v)

    def gradient(self, g):
        return [reducescatterCommunicate_op(g, self.comm)]

    def infer_shape(self, input_shapes):
        s = list(input_shapes[0])
        s[0] *= _comm_of(self).nrank
        return tuple(s)


class ReduceScatterCommunicateOp(_CommOp):
    def __init__(self, nodeA, comm=None, ctx=None):
        super().__init__(ReduceScatterCommunicateOp, [nodeA], ctx)
        self.comm = comm

    def compute(self, input_vals, output_val=None, stream_handle=None):
        v = input_vals[0]
        comm = _comm_of(self)
        if v.device.type == 'meta':
            return torch.empty((v.shape[0] // comm.nrank,) + tuple(v.shape[1:]), device='meta')
        eng = _symm_engine(self)
        if eng is not None and comm.multi and v.is_cuda and v.dtype == torch.float32:
            return eng.reducescatter(v, key=('op', self.id))
        return comm.reducescatter(v)

    def gradient(self, g):
        return [allgatherCommunicate_op(g, self.comm)]

    def infer_shape(self, input_shapes):
        s = list(input_shapes[0])
        s[0] //= _comm_of(self).nrank
        return tuple(s)


class BroadcastCommunicateOp(_CommOp):
    def __init__(self, nodeA, comm=None, root=0, ctx=None):
        super().__init__(BroadcastCommunicateOp, [nodeA], ctx)
        self.comm, self.root = comm, root

    def compute(self, input_vals, output_val=None, stream_handle=None):
        v = input_vals[0]
        comm = _comm_of(self)
        if v.device.type == 'meta' or not comm.multi:
            return v
        eng = _symm_engine(self)
        if eng is not None and v.is_cuda:                # NVLS: the root's multimem.st writes every rank's copy
            out = eng.broadcast(v, self.root, key=('bcast', self.id))
            if out is not None:
                return out
        out = v.clone()
        comm.broadcast(out, self.root)
        return out

    def gradient(self, g):
        return [reduceCommunicate_op(g, self.comm, self.root)]


class ReduceCommunicateOp(_CommOp):
    def __init__(self, nodeA, comm=None, root=0, ctx=None):
        super().__init__(ReduceCommunicateOp, [nodeA], ctx)
        self.comm, self.root = comm, root

    def compute(self, input_vals, output_val=None, stream_handle=None):
        v = input_vals[0]
        comm = _comm_of(self)
        if v.device.type == 'meta' or not comm.multi:
            return v
        eng = _symm_engine(self)
        if eng is not None and v.is_cuda:                # NVLS: the root reads the sum the switch computed
            out = eng.reduce(v, self.root, key=('reduce', self.id))
            if out is not None:
                return out
        out = v.clone()
        comm.reduce(out, self.root)
        return out

    def gradient(self, g):
        return [broadcastCommunicate_op(g, self.comm, self.root, self.raw_ctx)]


class AllToAllOp(_CommOp):
    """Equal-split all-to-all over dim 0 (MoE dispatch / combine exchange)."""

    def __init__(self, node_A, comm=None, ctx=None):
        super().__init__(AllToAllOp, [node_A], ctx)
        self.comm = comm

    def compute(self, input_vals, output_val=None, stream_handle=None):
        v = input_vals[0]
        comm = _comm_of(self)
        if v.device.type == 'meta' or not comm.multi:
            return v
        eng = _symm_engine(self)
        if eng is not None and v.is_cuda and v.dtype == torch.float32:
            return eng.alltoall(v, key=('op', self.id))
        return comm.alltoall(v.contiguous())

    def gradient(self, g):
        return [alltoall_op(g, ctx=self.raw_ctx)]


class HAllToAllOp(_CommOp):
    """Hierarchical all-to-all, API-compatible with reference HAllToAll.py:9-99
    (gather to the node leader -> leader-only inter-node exchange -> scatter).
    Inside one NVSwitch domain every GPU reaches every peer at full bandwidth, so
    the three-phase hierarchy buys nothing; the op keeps its
    ``(num_nodes, num_local_gpus)`` signature and result layout and is executed
    as one direct exchange (the result of an all-to-all does not depend on the
    route the chunks take)."""

    def __init__(self, node_A, num_nodes, num_local_gpus, comm=None, ctx=None):
        super().__init__(HAllToAllOp, [node_A], ctx)
        self.num_nodes, self.num_local_gpus, self.comm = num_nodes, num_local_gpus, comm

    def compute(self, input_vals, output_val=None, stream_handle=None):
        v = input_vals[0]
        comm = _comm_of(self)
        if v.device.type == 'meta' or not comm.multi:
            return v
        assert self.num_nodes * self.num_local_gpus == comm.nrank, \
            'halltoall: num_nodes*num_local_gpus must equal the group size'
        eng = _symm_engine(self)
        if eng is not None and v.is_cuda and v.dtype == torch.float32:
            return eng.alltoall(v, key=('op', self.id))
        return comm.alltoall(v.contiguous())

    def gradient(self, g):
        return [halltoall_op(g, self.num_nodes, self.num_local_gpus, ctx=self.raw_ctx)]


def ha2a_layout_transform(x: torch.Tensor, num_nodes: int, num_local_gpus: int) -> torch.Tensor:
    """Leader-side permutation [gpu][node][gpu'][off] -> [node][gpu'][gpu][off]
    (reference src/ops/H_A2A_LayoutTransform.cu:3-47); kept for parity and tests."""
    G, Nn = num_local_gpus, num_nodes
    rows = x.shape[0] // (G * Nn * G)
    return x.reshape(G, Nn, G, rows, -1).permute(1, 2, 0, 3, 4).reshape(x.shape).contiguous()


def ha2a_reverse_layout_transform(x: torch.Tensor, num_nodes: int, num_local_gpus: int) -> torch.Tensor:
    G, Nn = num_local_gpus, num_nodes
    rows = x.shape[0] // (G * Nn * G)
    return x.reshape(Nn, G, G, rows, -1).permute(2, 0, 1, 3, 4).reshape(x.shape).contiguous()


class PipelineSendOp(Op):
    """NCCL p2p send of an activation / gradient to the next stage.  With several
    destinations (unequal DP degrees) targets rotate round-robin
    (reference PipelineSend.py:38-39)."""
    graph_safe = False

    def __init__(self, node_A, destination, comm=None, ctx=None):
        super().__init__(PipelineSendOp, [node_A], ctx)
        self.const_attr = (destination,) if isinstance(destination, int) else tuple(destination)
        self.comm = comm
        self.index = 0
        self.ntargets = len(self.const_attr)
        self.use_indexed_slices = node_A.use_indexed_slices

    @property
    def desc(self):
        return self.name + '(send node %s to rank %s)' % (self.inputs[0].name, str(self.const_attr))

    def compute(self, input_vals, output_val=None, stream_handle=None):
        from ..parallel.comm import isend
        v = input_vals[0]
        dst = self.const_attr[self.index]
        d = getattr(self, 'direction', 0)
        if isinstance(v, IndexedSlices):
            ind, val = v._iv()
            isend(torch.tensor([ind.numel(), val.shape[1]], dtype=torch.int64, device=val.device), dst, d)
            isend(ind, dst, d)
            isend(val, dst, d)
        elif v.device.type != 'meta':
            hdr = torch.zeros(8, dtype=torch.int64)
            hdr[0] = v.dim()
            hdr[1:1 + v.dim()] = torch.tensor(list(v.shape), dtype=torch.int64)
            isend(hdr.to(v.device), dst, d)
            isend(v, dst, d)
        self.step_index()
        return None

    def step_index(self):
        self.index = (self.index + 1) % self.ntargets

    def gradient(self, g):
        return [None]

    def infer_shape(self, input_shapes):
        return None


class PipelineReceiveOp(Op):
    graph_safe = False

    def __init__(self, source, comm=None, use_indexed_slices=False, ctx=None):
        super().__init__(PipelineReceiveOp, [], ctx)
        self.const_attr = (source,) if isinstance(source, int) else tuple(source)
        self.comm = comm
        self.index = 0
        self.ntargets = len(self.const_attr)
        self.use_indexed_slices = use_indexed_slices
        self.static_shape_value = None

    @property
    def desc(self):
        return self.name + '(receive from rank %s)' % str(self.const_attr)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        from ..parallel.comm import recv_now, world
        src = self.const_attr[self.index]
        d = getattr(self, 'direction', 0)
        dev = self.ctx.torch_device
        if world().active and torch.distributed.get_backend() == 'gloo':
            dev = torch.device('cpu')
        if self.use_indexed_slices:
            hdr = recv_now(torch.zeros(2, dtype=torch.int64, device=dev), src, d)
            n, w = [int(x) for x in hdr.tolist()]
            ind = recv_now(torch.empty(n, dtype=torch.int64, device=dev), src, d)
            val = recv_now(torch.empty((n, w), dtype=torch.float32, device=dev), src, d)
            out = IndexedSlices(ind, val, None)
        else:
            h = recv_now(torch.zeros(8, dtype=torch.int64, device=dev), src, d).tolist()
            shape = tuple(int(x) for x in h[1:1 + int(h[0])])
            out = recv_now(torch.empty(shape, dtype=torch.float32, device=dev), src, d)
        self.step_index()
        return out

    def step_index(self):
        self.index = (self.index + 1) % self.ntargets

    def gradient(self, g):
        return []

    def infer_shape(self, input_shapes):
        return self.static_shape_value


class DispatchOp(Op):
    """Marks how ``node`` is partitioned for model parallelism: ``parts`` maps
    dim -> number of partitions (tuple form: one entry per dim).  Lowered by
    parallel/mp_lowering.py into local slices and all-gather / reduce-scatter /
    all-reduce ops (the reference ships only the marker: Dispatch.py:5-65)."""

    def __init__(self, node, parts):
        super().__init__(DispatchOp, [node], None)
        if isinstance(parts, dict):
            self.parts = dict(parts)
        else:
            self.parts = {i: p for i, p in enumerate(parts) if p != 1}

    def compute(self, input_vals, output_val=None, stream_handle=None):
        # un-lowered (single device): identity
        return input_vals[0]

    def gradient(self, g):
        return [dispatch_gradient(g, self.inputs[0])]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


class DispatchGradientOp(Op):
    def __init__(self, node, forward_input):
        super().__init__(DispatchGradientOp, [node, forward_input], None)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        return input_vals[0]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


# ---------------------------------------------------------------- public API
def allreduceCommunicate_op(node, comm=None):
    return AllReduceCommunicateOp(node, comm)


def allreduceCommunicatep2p_op(node, comm=None):
    return AllReduceCommunicateP2POp(node, comm)


def groupallreduceCommunicate_op(node, group_comm):
    return GroupAllReduceCommunicateOp(node, group_comm)


def allgatherCommunicate_op(node, comm=None):
    return AllGatherCommunicateOp(node, comm)


def reducescatterCommunicate_op(node, comm=None):
    return ReduceScatterCommunicateOp(node, comm)


def broadcastCommunicate_op(node, comm=None, root=0, ctx=None):
    return BroadcastCommunicateOp(node, comm, root, ctx)


def reduceCommunicate_op(node, comm=None, rank=0):
    return ReduceCommunicateOp(node, comm, rank)


def alltoall_op(node, ctx=None):
    return AllToAllOp(node, ctx=ctx)


def halltoall_op(node, num_nodes, num_local_gpus, ctx=None):
    return HAllToAllOp(node, num_nodes, num_local_gpus, ctx=ctx)


def pipeline_send_op(node, destination, comm=None, ctx=None):
    return PipelineSendOp(node, destination, comm, ctx=ctx)


def pipeline_receive_op(source, comm=None, use_indexed_slices=False, ctx=None):
    return PipelineReceiveOp(source, comm, use_indexed_slices, ctx=ctx)


def dispatch(node, parts={}):
    """Declare how ``node`` is partitioned over the devices of the enclosing model-parallel
    worker (tuple context).  ``parts`` = tuple with one entry per dim or {dim: n}."""
    from ..context import get_current_context
    from ..parallel.mp_lowering import lower_dispatch
    pdict = dict(parts) if isinstance(parts, dict) else {i: p for i, p in enumerate(parts)}
    lowered = lower_dispatch(node, pdict, get_current_context())
    if lowered is not None:
        return lowered
    return DispatchOp(node, parts)


def dispatch_gradient(node, forward_input):
    return DispatchGradientOp(node, forward_input)
