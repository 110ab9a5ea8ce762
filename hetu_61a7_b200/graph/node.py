"""Graph node base class.

reference: python/hetu/gpu_ops/Node.py:18-220.  A node owns no storage; the
executor threads ``torch.Tensor`` values through ``compute`` (functional: the
result is returned, memory comes from the caching allocator / the CUDA-graph
private pool).  ``infer_shape`` defaults to running ``compute`` on meta tensors
so the ~170 ops do not each need a hand-written shape rule.
"""
from __future__ import annotations

from copy import copy, deepcopy
from typing import List, Optional, Sequence, Tuple

import torch

from ..context import DeviceGroup, get_current_context
from ..ndarray import DLContext, IndexedSlices, ND_Sparse_Array, NDArray

G_NODE_ID = 0


def _next_id() -> int:
    global G_NODE_ID
    G_NODE_ID += 1
    return G_NODE_ID - 1


class OpMeta(type):
    """Runs the model-parallel state propagation once an op is fully constructed (sub-class
    constructors set their attributes after ``Op.__init__``)."""

    def __call__(cls, *args, **kwargs):
        obj = super().__call__(*args, **kwargs)
        if not obj._no_adapt and any(getattr(i, 'mp_state', None) is not None for i in obj.inputs if i is not None):
            from ..parallel.mp_lowering import adapt
            repl = adapt(obj)
            if repl is not None:          # a fused communication+compute node takes the op's place
                return repl
        return obj


class Op(object, metaclass=OpMeta):
    """Basic unit of the dataflow graph."""

    #: False for nodes that must touch the host every step (PS RPC, dataloaders,
    #: data-dependent output shapes); such graphs run eagerly instead of being
    #: captured into one CUDA graph.
    graph_safe = True

    def __init__(self, op_type, inputs: List['Op'], ctx=None, _adapt=True) -> None:
        self.inputs: List[Op] = list(inputs)
        self._no_adapt = not _adapt
        self.mp_state = None          # see parallel/mp_lowering.py
        self.mp_group = None
        self.raw_ctx: Optional[DeviceGroup] = (
            get_current_context() if ctx is None else DeviceGroup(ctx))
        self.ctx = ctx
        self.const_attr = None
        self.dtype = None
        self.inplace = False
        self.lazy_execution = False
        self.event = None
        self.use_indexed_slices = False
        self.op_type: str = op_type.__name__ if isinstance(op_type, type) else str(op_type)
        self.id: int = _next_id()
        self.name: str = self.op_type + str(self.id)

    # ------------------------------------------------------------------ sugar
    @property
    def desc(self) -> str:
        return self.name + '(' + ', '.join(inp.name for inp in self.inputs) + ')'

    def __add__(self, other):
        from ..ops.elementwise import add_op, addbyconst_op
        return add_op(self, other) if isinstance(other, Op) else addbyconst_op(self, other)

    def __mul__(self, other):
        from ..ops.elementwise import mul_op, mul_byconst_op
        return mul_op(self, other) if isinstance(other, Op) else mul_byconst_op(self, other)

    def __sub__(self, other):
        from ..ops.elementwise import minus_op, addbyconst_op
        return minus_op(self, other) if isinstance(other, Op) else addbyconst_op(self, -other)

    def __rsub__(self, other):
        from ..ops.elementwise import minus_byconst_op
        return minus_byconst_op(self, other)

    def __neg__(self):
        from ..ops.elementwise import opposite_op
        return opposite_op(self)

    def __truediv__(self, other):
        from ..ops.elementwise import div_op, mul_byconst_op
        return div_op(self, other) if isinstance(other, Op) else mul_byconst_op(self, 1.0 / other)

    def __rtruediv__(self, other):
        from ..ops.elementwise import div_const_op
        return div_const_op(other, self)

    __radd__ = __add__
    __rmul__ = __mul__

    def __str__(self) -> str:
        return self.name

    __repr__ = __str__

    # identity semantics, taken from object so that the executor's per-node dict lookups stay in C
    __hash__ = object.__hash__
    __eq__ = object.__eq__

    def __deepcopy__(self, memo):
        if id(self) not in memo:
            new_op = copy(self)
            memo[id(self)] = new_op
            for k, v in self.__dict__.items():
                if k in ('inputs', 'grad_nodes'):
                    new_op.__dict__[k] = [deepcopy(n, memo) for n in v]
                elif k in ('grad_node', 'forward_node', 'optimizer'):
                    new_op.__dict__[k] = deepcopy(v, memo)
        return memo[id(self)]

    # ---------------------------------------------------------------- contract
    def compute(self, input_vals, output_val=None, stream_handle=None):
        """Return the value of this node given its inputs' values."""
        raise NotImplementedError('%s.compute' % self.op_type)

    def gradient(self, output_grad: 'Op') -> List[Optional['Op']]:
        raise NotImplementedError('%s.gradient' % self.op_type)

    def infer_shape(self, input_shapes: Sequence[Tuple[int, ...]]) -> Tuple[int, ...]:
        """Default: trace ``compute`` on the meta device (no data, no kernels)."""
        metas = [torch.empty(tuple(s), device='meta') if s is not None else None
                 for s in input_shapes]
        out = self.compute(metas)
        if isinstance(out, IndexedSlices):
            return out.get_dense_shape()
        return tuple(out.shape)

    naive_infer_shape = infer_shape

    # ------------------------------------------------------------------ hooks
    @property
    def on_gpu(self) -> bool:
        return isinstance(self.ctx, DLContext) and self.ctx.device_type == 2

    @property
    def on_cpu(self) -> bool:
        return not self.on_gpu

    def resolve_ctx(self, config) -> None:
        """Pick the concrete device this process runs the node on."""
        rc = self.raw_ctx
        if rc is None or len(rc) == 0:
            self.ctx = config.context
            return
        unit = None
        if rc.worker_num > 0:
            if config.context in rc:
                unit = config.context
            else:
                unit = rc.cur_worker(getattr(config, 'dp_rank', 0) or 0)
            if isinstance(unit, tuple):
                unit = unit[getattr(config, 'mp_index', 0) % len(unit)]
        else:
            unit = rc.servers[0]
        self.ctx = unit

    def forward_hook(self, config) -> None:
        """Post-order pass: resolve the context (reference Node.py:192-213; the
        H2D/D2H transfer ops the reference inserts are implicit here - the
        executor moves a value when producer and consumer devices differ)."""
        if not isinstance(self.ctx, DLContext):
            self.resolve_ctx(config)

    def backward_hook(self, config) -> None:
        """Pre-order pass; comm-inserting ops override (optimizer, PS ops)."""
        pass

    def add_transfer_op(self, src_node, dst_ctx, h2d_ops=None, d2h_ops=None):
        from ..ops.transfer import datah2d_op, datad2h_op
        if src_node.ctx == dst_ctx or src_node.ctx is None:
            return src_node
        return datah2d_op(src_node, dst_ctx) if dst_ctx.device_type == 2 else datad2h_op(src_node)

    def get_default_state(self, status=None, enforce_order=False):
        return None

    def reset_status(self) -> None:
        """Forget the model-parallel layout derived for this node (reference Node.py:218-220 placeholder)."""
        self.mp_state = None
        self.mp_group = None


# ---------------------------------------------------------------------------
# helpers shared by op modules
def as_index(t: torch.Tensor) -> torch.Tensor:
    """Float-encoded ids (reference convention) -> int64."""
    if t.dtype in (torch.int64,):
        return t
    return t.long()


def is_meta(*ts) -> bool:
    return any(isinstance(t, torch.Tensor) and t.device.type == 'meta' for t in ts)


def unwrap(v):
    return v.tensor if isinstance(v, NDArray) else v
