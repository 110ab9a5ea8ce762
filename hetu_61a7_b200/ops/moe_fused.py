"""Fused MoE exchange: layout transform + all-to-all in one peer-memory kernel (K3b/K3c).

reference chain (python/hetu/layers/moe_layer.py:60-88): ``layout_transform_op`` (scatter
into a local [E*C, d] buffer) -> ``alltoall_op`` (grouped ncclSend/ncclRecv, 256 MiB per GPU)
-> experts -> ``alltoall_op`` -> ``reverse_layout_transform_op``.

Here the dispatch kernel stores every token row directly into the owner GPU's expert buffer
(``hetu_moe_dispatch_peer``: st.global over NVLink into [E_local, world, C, d]) and the
combine kernel reads the expert outputs back from their owners
(``hetu_moe_combine_peer``) - the all-to-all IS the layout transform, no staging copy, and
the expert GEMMs see an expert-major contiguous buffer.  Without a symmetric heap (single
GPU, CPU, NCCL baseline) the ops fall back to the unfused chain with the same result layout.
"""
from __future__ import annotations

import ctypes

import torch

from .. import _lib as L
from ..graph.node import Op
from .moe import moe_dispatch, moe_combine, moe_gate_grad, _ptrs


def _engine(op):
    cfg = getattr(op, '_config', None)
    return getattr(cfg, 'symm', None) if cfg is not None else None


def _world(op):
    cfg = getattr(op, '_config', None)
    comm = getattr(cfg, 'nccl_comm', None) if cfg is not None else None
    return comm if comm is not None and comm.multi else None


class FusedMoEDispatchOp(Op):
    """(x [T, d], routes) -> expert-major receive buffer [E_local, world, C, d]."""

    def __init__(self, x, indices_s, location_s, capacity, e_local, world, gates=None, ctx=None):
        gates = list(gates) if gates is not None else []
        super().__init__(FusedMoEDispatchOp, [x] + list(indices_s) + list(location_s) + gates, ctx)
        self.capacity, self.e_local, self.world, self.k = int(capacity), int(e_local), int(world), len(indices_s)
        self.has_gate = len(gates) > 0

    def forward_hook(self, config):
        super().forward_hook(config)
        self._config = config

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        k = self.k
        idxs, locs = input_vals[1:1 + k], input_vals[1 + k:1 + 2 * k]
        gates = input_vals[1 + 2 * k:1 + 3 * k] if self.has_gate else None
        E, W, C, d = self.e_local, self.world, self.capacity, x.shape[-1]
        if x.device.type == 'meta':
            return torch.empty((E, W, C, d), device='meta')
        eng, comm = _engine(self), _world(self)
        if eng is not None and comm is not None and x.is_cuda and d % 4 == 0:
            buf = eng.buffer(('moe_recv', self.id), E * W * C * d * 4)
            recv = buf.tensor(numel=E * W * C * d)
            recv.zero_()
            eng.barrier()                      # receive buffers zeroed and free on every rank
            g = [t.contiguous().reshape(-1) for t in gates] if gates else []
            L.call('moe_dispatch_peer', L.p(x.contiguous()), _ptrs([t.contiguous().reshape(-1) for t in idxs]),
                   _ptrs([t.contiguous().reshape(-1) for t in locs]), _ptrs(g), L.c_int(k),
                   L.c_int(int(self.has_gate)), buf.ptrs, L.c_int(W), L.c_int(eng.rank), L.c_int(x.shape[0]),
                   L.c_int(d), L.c_int(C), L.c_int(E), L.stream_ptr(x.device))
            eng.barrier()                      # every peer's rows have landed
            return recv.view(E, W, C, d)
        local = moe_dispatch(x, idxs, locs, C, E * W, gates=gates)           # [W*E*C, d] (dest-rank major)
        if comm is not None:
            local = comm.alltoall(local.contiguous())                       # [W(src), E, C, d]
        return local.view(W, E, C, d).permute(1, 0, 2, 3).contiguous()

    def gradient(self, g):
        k = self.k
        idxs, locs = self.inputs[1:1 + k], self.inputs[1 + k:1 + 2 * k]
        gates = self.inputs[1 + 2 * k:1 + 3 * k] if self.has_gate else None
        gx = FusedMoECombineOp(g, idxs, locs, self.capacity, self.e_local, self.world, gates=gates,
                               ctx=self.raw_ctx)
        return [gx] + [None] * (len(self.inputs) - 1)

    def infer_shape(self, input_shapes):
        return (self.e_local, self.world, self.capacity, input_shapes[0][-1])


class FusedMoECombineOp(Op):
    """(y [E_local, world, C, d], routes, gates) -> tokens [T, d] (sum over the k routes)."""

    def __init__(self, y, indices_s, location_s, capacity, e_local, world, gates=None, ctx=None):
        gates = list(gates) if gates is not None else []
        super().__init__(FusedMoECombineOp, [y] + list(indices_s) + list(location_s) + gates, ctx)
        self.capacity, self.e_local, self.world, self.k = int(capacity), int(e_local), int(world), len(indices_s)
        self.has_gate = len(gates) > 0
        self.send_buf = None
        self.local_rows = None

    def forward_hook(self, config):
        super().forward_hook(config)
        self._config = config

    def compute(self, input_vals, output_val=None, stream_handle=None):
        y = input_vals[0]
        k = self.k
        idxs, locs = input_vals[1:1 + k], input_vals[1 + k:1 + 2 * k]
        gates = input_vals[1 + 2 * k:1 + 3 * k] if self.has_gate else None
        E, W, C = self.e_local, self.world, self.capacity
        d = y.shape[-1]
        T = idxs[0].numel()
        if y.device.type == 'meta':
            return torch.empty((T, d), device='meta')
        eng, comm = _engine(self), _world(self)
        if eng is not None and comm is not None and y.is_cuda and d % 4 == 0:
            buf = eng.buffer(('moe_send', self.id), E * W * C * d * 4)
            self.send_buf = buf
            buf.tensor(numel=E * W * C * d).copy_(y.reshape(-1))
            eng.barrier()                      # every rank's expert outputs are in place
            out = torch.empty((T, d), dtype=torch.float32, device=y.device)
            g = [t.contiguous().reshape(-1) for t in gates] if gates else []
            L.call('moe_combine_peer', _ptrs([t.contiguous().reshape(-1) for t in idxs]),
                   _ptrs([t.contiguous().reshape(-1) for t in locs]), _ptrs(g), L.c_int(k),
                   L.c_int(int(self.has_gate)), buf.ptrs, L.c_int(W), L.c_int(eng.rank), L.p(out), L.c_int(T),
                   L.c_int(d), L.c_int(C), L.c_int(E), L.stream_ptr(y.device))
            return out
        rows = y.reshape(E, W, C, d).permute(1, 0, 2, 3).contiguous().view(W * E * C, d)   # [W(src), E, C, d]
        if comm is not None:
            rows = comm.alltoall(rows)                                                    # back at the source
        self.local_rows = rows
        return moe_combine(rows, idxs, locs, C, gates=gates)

    def gradient(self, g):
        k = self.k
        idxs, locs = self.inputs[1:1 + k], self.inputs[1 + k:1 + 2 * k]
        gates = self.inputs[1 + 2 * k:1 + 3 * k] if self.has_gate else None
        gy = FusedMoEDispatchOp(g, idxs, locs, self.capacity, self.e_local, self.world, gates=gates,
                                ctx=self.raw_ctx)
        out = [gy] + [None] * (2 * k)
        if self.has_gate:
            out += [FusedMoEGateGradOp(g, self, idxs[j], locs[j], ctx=self.raw_ctx) for j in range(k)]
        return out

    def infer_shape(self, input_shapes):
        return (input_shapes[1][0], input_shapes[0][-1])


class FusedMoEGateGradOp(Op):
    def __init__(self, dout, combine_node, idx, loc, ctx=None):
        super().__init__(FusedMoEGateGradOp, [dout, combine_node, idx, loc], ctx)
        self.forward_node = combine_node

    def forward_hook(self, config):
        super().forward_hook(config)
        self._config = config

    def compute(self, input_vals, output_val=None, stream_handle=None):
        dout, _, idx, loc = input_vals
        fwd = self.forward_node
        T, d = dout.shape
        if dout.device.type == 'meta':
            return torch.empty((T,), device='meta')
        eng = _engine(self)
        if fwd.send_buf is not None and eng is not None:
            out = torch.empty((T,), dtype=torch.float32, device=dout.device)
            L.call('moe_gate_grad_peer', L.p(dout.contiguous()), L.p(idx.contiguous().reshape(-1)),
                   L.p(loc.contiguous().reshape(-1)), fwd.send_buf.ptrs, L.c_int(fwd.world), L.c_int(eng.rank),
                   L.p(out), L.c_int(T), L.c_int(d), L.c_int(fwd.capacity), L.c_int(fwd.e_local),
                   L.stream_ptr(dout.device))
            return out
        return moe_gate_grad(dout, fwd.local_rows, idx.reshape(-1), loc.reshape(-1), fwd.capacity)

    def infer_shape(self, input_shapes):
        return (input_shapes[0][0],)


def fused_moe_dispatch_op(x, indices_s, location_s, capacity, e_local, world, gates=None, ctx=None):
    return FusedMoEDispatchOp(x, indices_s, location_s, capacity, e_local, world, gates, ctx=ctx)


def fused_moe_combine_op(y, indices_s, location_s, capacity, e_local, world, gates=None, ctx=None):
    return FusedMoECombineOp(y, indices_s, location_s, capacity, e_local, world, gates, ctx=ctx)
