"""Model-parallel (tensor-parallel) lowering of ``ht.dispatch``.

The reference ships only the marker op (``DispatchOp.compute`` asserts "should be replaced in
preprocessing phase", python/hetu/gpu_ops/Dispatch.py:10-11) - the pass that replaces it is
missing from the snapshot.  The contract is taken from
examples/runner/parallel/test_mlp_mp.py:60-96: inside ``with ht.context((gpu0, gpu1, ...))`` a
tuple context is ONE model-parallel worker; ``ht.dispatch(node, parts)`` says how ``node`` is
partitioned over the devices of that worker (``parts[d]`` partitions along dim ``d``; all ones
= replicated), and ops in between compute on the local shards.

Design: lowering happens at graph-construction time, so the symbolic autodiff that runs later
differentiates the real, lowered graph.  Every node carries ``mp_state``:
    None                      replicated on the worker's devices
    ('split', dim, n)         this device holds slice ``mp_index`` of ``n`` along ``dim``
    ('partial', n)            this device holds a partial sum (needs a reduction)
``dispatch`` converts between states with first-class comm ops (local slice, all-gather,
all-reduce, reduce-scatter - each with its own gradient); ``Op.__init__`` calls ``adapt``
which propagates states through matmul / element-wise ops and inserts the conversions an op
needs (e.g. an all-reduce before a non-linear op that receives a partial sum).  On B200 the
all-gather -> GEMM and GEMM -> reduce-scatter pairs run over the symmetric heap.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from ..graph.node import Op


def _group_of(op) -> Optional[tuple]:
    """The tuple context (one MP worker) an op was built under, if any."""
    rc = op.raw_ctx
    if rc is None or not rc.is_mp:
        return None
    for w in rc.workers:
        if isinstance(w, tuple):
            return w
    return None


class _MPComm(object):
    """Lazily resolved communicator of one MP worker (group of device ids on this node)."""

    def __init__(self, devices):
        self.devices = tuple(devices)
        self._comm = None

    def resolve(self):
        if self._comm is None:
            from .comm import new_group_comm, world
            w = world()
            if not w.multi:
                self._comm = w
            else:
                ranks = sorted(d.device_id for d in self.devices)
                self._comm = new_group_comm(ranks)
        return self._comm

    @property
    def index(self) -> int:
        c = self.resolve()
        return max(c.rank, 0)

    @property
    def size(self) -> int:
        c = self.resolve()
        return c.nrank if c.multi else len(self.devices) if False else c.nrank


class MPSliceOp(Op):
    """replicated -> split: keep slice ``mp_index`` along ``dim`` (no communication)."""

    def __init__(self, node, dim, n, group, ctx=None):
        super().__init__(MPSliceOp, [node], ctx, _adapt=False)
        self.dim, self.n, self.group = dim, n, group
        self.mp_group = group
        self.mp_state = ('split', dim, n)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        n = self.n
        part = x.shape[self.dim] // n
        idx = 0 if x.device.type == 'meta' else (self.group.index % n)
        from ..ops.shape import materialize
        return materialize(x.narrow(self.dim, idx * part, part))

    def gradient(self, g):
        return [MPAllGatherOp(g, self.dim, self.n, self.group, ctx=self.raw_ctx)]


class MPAllGatherOp(Op):
    """split -> replicated: concatenate the group's slices along ``dim``."""

    def __init__(self, node, dim, n, group, ctx=None):
        super().__init__(MPAllGatherOp, [node], ctx, _adapt=False)
        self.dim, self.n, self.group = dim, n, group
        self.mp_group = group
        self.mp_state = None

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        if x.device.type == 'meta':
            shp = list(x.shape)
            shp[self.dim] *= self.n
            return torch.empty(shp, device='meta')
        comm = self.group.resolve()
        if not comm.multi:
            return torch.cat([x] * self.n, dim=self.dim) if self.n > 1 and False else x
        xt = x.transpose(0, self.dim).contiguous() if self.dim != 0 else x.contiguous()
        out = comm.allgather(xt)
        return out.transpose(0, self.dim).contiguous() if self.dim != 0 else out

    def gradient(self, g):
        return [MPSliceOp(g, self.dim, self.n, self.group, ctx=self.raw_ctx)]


class MPAllReduceOp(Op):
    """partial -> replicated.  Gradient: identity (every device already holds the full grad)."""

    def __init__(self, node, group, ctx=None):
        super().__init__(MPAllReduceOp, [node], ctx, _adapt=False)
        self.group = group
        self.mp_group = group
        self.mp_state = None

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        if x.device.type == 'meta':
            return x
        comm = self.group.resolve()
        if not comm.multi:
            return x
        out = x.clone()
        comm.allreduce(out)
        return out

    def gradient(self, g):
        return [g]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


class MPReduceScatterOp(Op):
    """partial -> split(dim)."""

    def __init__(self, node, dim, n, group, ctx=None):
        super().__init__(MPReduceScatterOp, [node], ctx, _adapt=False)
        self.dim, self.n, self.group = dim, n, group
        self.mp_group = group
        self.mp_state = ('split', dim, n)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        if x.device.type == 'meta':
            shp = list(x.shape)
            shp[self.dim] //= self.n
            return torch.empty(shp, device='meta')
        comm = self.group.resolve()
        if not comm.multi:
            return x
        xt = x.transpose(0, self.dim).contiguous() if self.dim != 0 else x.contiguous()
        out = comm.reducescatter(xt)
        return out.transpose(0, self.dim).contiguous() if self.dim != 0 else out

    def gradient(self, g):
        return [MPAllGatherOp(g, self.dim, self.n, self.group, ctx=self.raw_ctx)]


_GROUPS = {}


def _group(devices) -> _MPComm:
    key = tuple(devices)
    if key not in _GROUPS:
        _GROUPS[key] = _MPComm(devices)
    return _GROUPS[key]


def to_replicated(node, group, ctx=None):
    st = getattr(node, 'mp_state', None)
    if st is None:
        return node
    if st[0] == 'split':
        return MPAllGatherOp(node, st[1], st[2], group, ctx=ctx)
    return MPAllReduceOp(node, group, ctx=ctx)


def lower_dispatch(node, parts, ctx_group):
    """Implementation of ``ht.dispatch`` inside an MP context."""
    devices = None
    if ctx_group is not None and ctx_group.is_mp:
        devices = next(w for w in ctx_group.workers if isinstance(w, tuple))
    src_group = getattr(node, 'mp_group', None)
    if devices is None and src_group is None:
        return None                                       # not model parallel: marker only
    group = _group(devices) if devices is not None else src_group
    split = [(d, p) for d, p in sorted(parts.items()) if p != 1]
    if len(split) > 1:
        raise NotImplementedError('dispatch over more than one dimension (%s) - 2-D device meshes '
                                  'are not lowered yet' % (parts,))
    st = getattr(node, 'mp_state', None)
    # a conversion out of an MP worker issued from a later (non-MP) stage still has to run on
    # every device of the source worker
    ctx = ctx_group if devices is not None else node.raw_ctx
    if not split:                                        # -> replicated
        out = to_replicated(node, src_group or group, ctx=ctx)
    else:
        d, n = split[0]
        if st is None:
            out = MPSliceOp(node, d, n, group, ctx=ctx)
        elif st[0] == 'split' and st[1] == d and st[2] == n:
            out = node
        elif st[0] == 'partial':
            out = MPReduceScatterOp(node, d, n, group, ctx=ctx)
        else:
            out = MPSliceOp(to_replicated(node, src_group or group, ctx=ctx), d, n, group, ctx=ctx)
    if out is not node or getattr(out, 'mp_group', None) is None:
        out.mp_group = group
    return out


_LINEAR_UNARY = {'AddByConstOp', 'MulByConstOp', 'OppositeOp', 'ArrayReshapeOp', 'TransposeOp'}
_ELEMENTWISE_UNARY = {'ReluOp', 'GeluOp', 'SigmoidOp', 'TanhOp', 'ExpOp', 'AbsOp', 'LeakyReluOp', 'SqrtOp',
                      'DropoutOp', 'MulByConstOp', 'AddByConstOp', 'OppositeOp'}


def adapt(op) -> None:
    """Called at the end of ``Op.__init__``: derive ``op.mp_state`` from the inputs' states,
    inserting the conversions the op needs."""
    ins = [i for i in op.inputs if i is not None]
    states = [getattr(i, 'mp_state', None) for i in ins]
    if all(s is None for s in states):
        return
    group = next(getattr(i, 'mp_group', None) for i in ins if getattr(i, 'mp_state', None) is not None)
    op.mp_group = group
    name = op.op_type
    ctx = op.raw_ctx
    if name == 'MatMulOp' and len(ins) == 2:
        a, b = states
        tA, tB = op.matmul_attr_trans_A, op.matmul_attr_trans_B
        a_k, a_m = (0, 1) if tA else (1, 0)             # which dim of A is K / M
        b_k, b_n = (1, 0) if tB else (0, 1)
        if a and b and a[0] == 'split' and b[0] == 'split' and a[1] == a_k and b[1] == b_k and a[2] == b[2]:
            op.mp_state = ('partial', a[2])              # contraction dim is sharded on both sides
            return
        if a is None and b and b[0] == 'split' and b[1] == b_n:
            op.mp_state = ('split', 1, b[2])             # column parallel
            return
        if b is None and a and a[0] == 'split' and a[1] == a_m:
            op.mp_state = ('split', 0, a[2])             # row (batch) parallel
            return
    if len(ins) == 1 and name in _ELEMENTWISE_UNARY:
        s = states[0]
        if s[0] == 'split' or name in _LINEAR_UNARY and name != 'AddByConstOp':
            op.mp_state = s
            return
    if name in ('AddOp', 'MinusOp', 'MulOp', 'DivOp') and states[0] == states[1] and states[0] is not None \
            and (states[0][0] == 'split' or name in ('AddOp', 'MinusOp')):
        op.mp_state = states[0]
        return
    from ..ops.elementwise import BinaryOp
    if isinstance(op, BinaryOp) and len(ins) == 2 and states[0] == states[1] and states[0][0] == 'split' \
            and name not in ('BoolMatrixOp',):
        op.mp_state = states[0]                          # same-shape element-wise on local shards
        return
    # everything else needs replicated inputs
    for k, i in enumerate(op.inputs):
        if i is not None and getattr(i, 'mp_state', None) is not None:
            op.inputs[k] = to_replicated(i, getattr(i, 'mp_group', group), ctx=i.raw_ctx)   # runs on the whole worker
    op.mp_state = None


def match_grad_state(grad, node):
    """Autodiff hook: the adjoint of ``node`` must be laid out like ``node`` itself (split ->
    same split; replicated or partial-sum forward value -> replicated adjoint).  Gradient ops
    that fell back to replicated inputs, or contractions that produced partial sums, are
    converted here so every ``gradient()`` implementation can assume a matching layout."""
    have = getattr(grad, 'mp_state', None)
    want = getattr(node, 'mp_state', None)
    if want is not None and want[0] == 'partial':
        want = None
    if have == want:
        return grad
    group = getattr(node, 'mp_group', None) or getattr(grad, 'mp_group', None)
    if group is None:
        return grad
    # the conversion is a collective of the whole MP worker: run it under the tuple context
    ctx = grad.raw_ctx if (grad.raw_ctx is not None and grad.raw_ctx.is_mp) else node.raw_ctx
    if want is None:
        out = to_replicated(grad, group, ctx=ctx)
    elif have is not None and have[0] == 'partial':
        out = MPReduceScatterOp(grad, want[1], want[2], group, ctx=ctx)
    else:
        out = MPSliceOp(to_replicated(grad, group, ctx=ctx), want[1], want[2], group, ctx=ctx)
    out.mp_group = group
    return out
