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
import os


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
    """Lazily resolved communicator of the model-parallel worker THIS rank belongs to.  A context with
    several tuple workers (data-parallel replicas of a tensor-parallel group) yields one process group
    per tuple; every rank creates all of them (group creation is collective) and keeps its own."""

    def __init__(self, workers):
        workers = list(workers)
        if workers and not isinstance(workers[0], tuple):
            workers = [tuple(workers)]                     # a single tuple was passed
        self.workers = [tuple(w) for w in workers]
        self.devices = self.workers[0]
        self._comm = None

    def resolve(self):
        if self._comm is None:
            from .comm import new_group_comm, world
            w = world()
            if not w.multi:
                self._comm = w
            else:
                mine = None
                for tup in self.workers:
                    ranks = sorted(d.device_id for d in tup)
                    comm = new_group_comm(ranks)
                    if w.world_rank in ranks:
                        mine = comm
                self._comm = mine if mine is not None else new_group_comm(sorted(d.device_id for d in self.workers[0]))
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


def _symm_for(group):
    """Symmetric-heap engine when the MP worker spans the whole (single-node) world, else None."""
    comm = group.resolve()
    if not comm.multi or not torch.cuda.is_available():
        return None
    from .comm import world
    if comm is not world() or comm.nrank > 8:
        return None
    from .symm import SymmEngine
    return SymmEngine.get(comm)


class MPAllGatherMatMulOp(Op):
    """Y = allgather_rows(X_shard) @ op(W) - sequence-parallel input of a column-parallel (or replicated)
    linear layer.  On B200 one kernel (csrc/gemm_sm100.cu hetu_gemm_ag): the TMA producer of the
    persistent 2-SM GEMM reads each 128-row block straight from its owner's HBM over NVLink, so the
    gather overlaps the MMAs tile by tile and the gathered activation is never materialised."""

    def __init__(self, shard, weight, trans_B, n, group, ctx=None):
        super().__init__(MPAllGatherMatMulOp, [shard, weight], ctx, _adapt=False)
        self.trans_B, self.n, self.group = trans_B, n, group
        self.mp_group = group
        self.mp_state = getattr(weight, 'mp_state', None)       # col-parallel weight -> split(1)
        if self.mp_state is not None:
            self.mp_state = ('split', 1, self.mp_state[2])
        self.fused_launches = 0

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x, w = input_vals
        if x.device.type == 'meta':
            return torch.empty((x.shape[0] * self.n, w.shape[0] if self.trans_B else w.shape[1]), device='meta')
        from ..ops.linalg import gemm
        comm = self.group.resolve()
        if not comm.multi:
            return gemm(x, w, False, self.trans_B)
        eng = _symm_for(self.group) if x.is_cuda else None
        m, k = x.shape
        nn = w.shape[0] if self.trans_B else w.shape[1]
        if eng is not None and m % 128 == 0 and k % 4 == 0 and nn % 4 == 0 and x.dtype == torch.float32 \
                and (m * self.n) >= 512 and nn >= 256:
            from .. import _lib as L
            rank = comm.rank
            buf = eng.buffer(('tp_ag', self.id), m * self.n * k * 4)           # full gathered activation
            flg = eng.buffer(('tp_ag_flags', self.id), self.n * (m // 128) * 4)
            buf.tensor(numel=m * self.n * k)[rank * m * k:(rank + 1) * m * k].copy_(x.reshape(-1))
            flg.tensor(dtype=torch.int32, numel=self.n * (m // 128)).zero_()
            out = torch.empty((m * self.n, nn), dtype=torch.float32, device=x.device)
            eng.barrier()                                         # shards in place, flags zeroed, last round's readers done
            L.call('gemm_ag', buf.ptrs, flg.ptrs, L.c_int(rank), L.c_int(self.n), L.p(w.contiguous()), L.p(out),
                   L.p(None), L.c_int(m * self.n), L.c_int(nn), L.c_int(k), L.c_int(int(self.trans_B)), L.c_f32(1.0),
                   L.c_int(0), L.stream_ptr(x.device))
            self.fused_launches += 1
            return out
        return gemm(comm.allgather(x.contiguous()), w, False, self.trans_B)

    def gradient(self, g):
        from ..ops.linalg import matmul_op
        shard, w = self.inputs
        c = self.raw_ctx
        d_full = matmul_op(g, w, False, not self.trans_B, ctx=c)          # partial sum or replicated
        gathered = MPAllGatherOp(shard, 0, self.n, self.group, ctx=c)
        if self.trans_B:
            d_w = matmul_op(g, gathered, True, False, ctx=c)
        else:
            d_w = matmul_op(gathered, g, True, False, ctx=c)
        return [d_full, d_w]                     # autodiff re-lays d_full out as split(0) (reduce-scatter / slice)


class MPMatMulReduceScatterOp(Op):
    """Y_shard = reduce_scatter_rows(op(A) @ op(B)) - output of a row-parallel linear layer handed to a
    sequence-parallel region.  On B200 one kernel (hetu_gemm_rs): the epilogue of the 2-SM GEMM adds each
    row block into its owner's buffer with red.global.add.v4.f32 over NVLink."""

    def __init__(self, a, b, trans_A, trans_B, n, group, ctx=None):
        super().__init__(MPMatMulReduceScatterOp, [a, b], ctx, _adapt=False)
        self.trans_A, self.trans_B, self.n, self.group = trans_A, trans_B, n, group
        self.mp_group = group
        self.mp_state = ('split', 0, n)
        self.fused_launches = 0

    def compute(self, input_vals, output_val=None, stream_handle=None):
        a, b = input_vals
        m = a.shape[1] if self.trans_A else a.shape[0]
        k = a.shape[0] if self.trans_A else a.shape[1]
        nn = b.shape[0] if self.trans_B else b.shape[1]
        if a.device.type == 'meta':
            return torch.empty((m // self.n, nn), device='meta')
        from ..ops.linalg import gemm
        comm = self.group.resolve()
        if not comm.multi:
            return gemm(a, b, self.trans_A, self.trans_B)
        eng = _symm_for(self.group) if a.is_cuda else None
        if eng is not None and m % self.n == 0 and nn % 4 == 0 and k % 4 == 0 and m % 4 == 0 and a.dtype == torch.float32 \
                and m >= 512 and nn >= 256:
            from .. import _lib as L
            rows = m // self.n
            buf = eng.buffer(('tp_rs', self.id), self.n * rows * nn * 4)        # [world][rows][N] staging
            eng.barrier()                                         # last round's staging has been consumed everywhere
            L.call('gemm_rs', L.p(a.contiguous()), L.p(b.contiguous()), buf.ptrs, L.c_int(comm.rank), L.c_int(self.n),
                   L.c_int(m), L.c_int(nn), L.c_int(k), L.c_int(int(self.trans_A)), L.c_int(int(self.trans_B)),
                   L.c_f32(1.0), L.stream_ptr(a.device))
            eng.barrier()                                         # every rank's partial rows have landed
            from ..ops.reduce import _reduce
            acc = _reduce(buf.tensor(numel=self.n * rows * nn).view(self.n, rows, nn), [0], False, 0)
            self.fused_launches += 1
            return acc
        return comm.reducescatter(gemm(a, b, self.trans_A, self.trans_B))

    def gradient(self, g):
        from ..ops.linalg import matmul_op
        a, b = self.inputs
        c = self.raw_ctx
        if not self.trans_A:
            d_a = MPAllGatherMatMulOp(g, b, not self.trans_B, self.n, self.group, ctx=c)     # AG(g) @ op(B)^T
            d_a.mp_state = getattr(a, 'mp_state', None)
        else:
            full = MPAllGatherOp(g, 0, self.n, self.group, ctx=c)
            d_a = matmul_op(b, full, self.trans_B, True, ctx=c)
        full = MPAllGatherOp(g, 0, self.n, self.group, ctx=c)
        if not self.trans_A and not self.trans_B:
            d_b = matmul_op(a, full, True, False, ctx=c)
        elif not self.trans_A and self.trans_B:
            d_b = matmul_op(full, a, True, False, ctx=c)
        elif self.trans_A and not self.trans_B:
            d_b = matmul_op(a, full, False, False, ctx=c)
        else:
            d_b = matmul_op(full, a, True, True, ctx=c)
        return [d_a, d_b]


def reduce_scatter_of(node, d, n, group, ctx=None):
    """partial -> split(d): fused into the producing GEMM when that is a plain matmul and d == 0."""
    if d == 0 and node.op_type == 'MatMulOp' and len(node.inputs) == 2 and os.environ.get('HETU_TP_FUSE', '1') == '1':
        a, b = node.inputs
        return MPMatMulReduceScatterOp(a, b, node.matmul_attr_trans_A, node.matmul_attr_trans_B, n, group,
                                       ctx=ctx if ctx is not None else node.raw_ctx)
    return MPReduceScatterOp(node, d, n, group, ctx=ctx)


_GROUPS = {}


def _group(workers) -> _MPComm:
    key = tuple(tuple(w) if isinstance(w, tuple) else w for w in workers)
    if key not in _GROUPS:
        _GROUPS[key] = _MPComm(workers)
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
        devices = [w for w in ctx_group.workers if isinstance(w, tuple)]
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
            out = reduce_scatter_of(node, d, n, group, ctx=ctx)
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
        if a and a[0] == 'split' and a[1] == a_m and not tA and b and b[0] == 'split' and b[1] == b_n \
                and a[2] == b[2] and os.environ.get('HETU_TP_FUSE', '1') == '1':
            # sequence-parallel activation into a column-parallel weight: all-gather -> GEMM, one kernel
            return MPAllGatherMatMulOp(ins[0], ins[1], tB, a[2], group, ctx=ctx)
    if len(ins) == 1 and name in _ELEMENTWISE_UNARY:
        s = states[0]
        if s[0] == 'split' or name in _LINEAR_UNARY and name != 'AddByConstOp':
            op.mp_state = s
            return
    if name in ('AddOp', 'MinusOp', 'MulOp', 'DivOp') and states[0] == states[1] and states[0] is not None \
            and (states[0][0] == 'split' or name in ('AddOp', 'MinusOp')):
        op.mp_state = states[0]
        return
    if name == 'BroadcastToOp' and len(ins) == 2:
        a, like = states
        if like is not None and like[0] == 'split' and (
                (a is None and like[1] == 0) or (a is not None and a[0] == 'split' and a[2] == like[2] and like[1] != 0)):
            op.mp_state = like                 # bias shard broadcast over a column shard / replicated bias over row shards
            return
    if name == 'ReduceToShapeOp' and len(ins) == 2:
        g, like = states
        if g is not None and g[0] == 'split':
            if like is not None and like[0] == 'split' and g[1] != 0 and g[2] == like[2]:
                op.mp_state = like             # column-sharded gradient summed over rows -> shard of the bias gradient
                return
            if like is None and g[1] == 0:
                op.mp_state = ('partial', g[2])    # rows are sharded: every member holds a partial sum
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
        out = reduce_scatter_of(grad, want[1], want[2], group, ctx=ctx)
    else:
        out = MPSliceOp(to_replicated(grad, group, ctx=ctx), want[1], want[2], group, ctx=ctx)
    out.mp_group = group
    return out
