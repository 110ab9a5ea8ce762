"""Fused top-k MoE gate: softmax + top-k + capacity-slot assignment + balance loss as ONE op.

reference: python/hetu/layers/TopGate.py:15-79 builds the gate from ~30 graph ops (softmax,
topk_idx, one_hot, cumsum_with_bias, reduce_sum / reduce_mean, element-wise products); each is a
kernel launch forward and most have one backward.  ``fused_topk_gate_op(logits, k)`` returns the
planar tensor P [3k * T + 1 + E] of csrc/moe_gate.cu (k planes of expert indices, k of slots, k of gate
values, the balance loss, the per-expert route counts).  ``gate_piece_op`` hands one plane (or the loss) on as
a view; the gradients that arrive at the pieces go straight into ONE backward op - no padded copies, no sums.
``layers.moe.topkgating`` uses it when ``HETU_MOE_FUSED_GATE=1`` (default) - same values as the composed gate."""
from __future__ import annotations

import ctypes

import torch

from .. import _lib as L
from ..graph.node import Op


def _reference_forward(logits: torch.Tensor, k: int) -> torch.Tensor:
    T, E = logits.shape
    p = torch.softmax(logits.float(), dim=1)
    idx = torch.topk(logits, k, dim=1).indices                                   # largest first
    P = torch.zeros(3 * k * T + 1 + E, dtype=torch.float32, device=logits.device)
    planes = P[:3 * k * T].view(3 * k, T)
    base = torch.zeros(E, dtype=torch.float32, device=logits.device)
    aux = 0.0
    me = p.mean(dim=0)
    for i in range(k):
        onehot = torch.nn.functional.one_hot(idx[:, i], E).float()
        loc = (torch.cumsum(onehot, dim=0) - 1 + base) * onehot
        planes[i] = idx[:, i].float()
        planes[k + i] = loc.sum(dim=1)
        planes[2 * k + i] = p.gather(1, idx[:, i:i + 1]).squeeze(1)
        base = base + onehot.sum(dim=0)
        aux = aux + (me * onehot.mean(dim=0)).sum() * E
    P[3 * k * T] = aux
    P[3 * k * T + 1:] = base
    return P


def _reference_backward(logits: torch.Tensor, P: torch.Tensor, dgates, daux, k: int) -> torch.Tensor:
    T, E = logits.shape
    p = torch.softmax(logits.float(), dim=1)
    planes = P[:3 * k * T].view(3 * k, T)
    idx = planes[:k].t().long()                                                   # [T, k]
    cnt = P[3 * k * T + 1:]
    dl = daux.reshape(-1)[0] if daux is not None else 0.0
    dp = (dl * E / float(T * T) * cnt).unsqueeze(0).expand(T, E).clone()          # d aux / d p[t, e]
    dg = torch.stack([d.reshape(-1) if d is not None else torch.zeros(T, device=logits.device) for d in dgates], 1)
    dp.scatter_add_(1, idx, dg.to(dp.dtype))
    return p * (dp - (dp * p).sum(dim=1, keepdim=True))


class FusedTopKGateOp(Op):
    def __init__(self, logits, k, ctx=None):
        super().__init__(FusedTopKGateOp, [logits], ctx)
        self.k = int(k)
        self.fused_launches = 0
        #: set by layers.moe.topkgating: the row count its pieces were built for
        self.expected_tokens = None

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        T, E = x.shape
        k = self.k
        if self.expected_tokens is not None and int(self.expected_tokens) != int(T):
            raise ValueError('gate was built for num_tokens=%d but the logits have %d rows (pass the total number of tokens '
                             'per step, batch_size * tokens, to the gate)' % (self.expected_tokens, T))
        if x.device.type == 'meta':
            return torch.empty((3 * k * T + 1 + E,), device='meta')
        if L.kernels_on(x) and x.dtype == torch.float32 and k <= 4 and E * (k + 1) * 4 <= 48 * 1024:
            x = x.contiguous()
            P = torch.empty((3 * k * T + 1 + E,), dtype=torch.float32, device=x.device)
            nblk = (T + 255) // 256
            blk_cnt = torch.empty(k * nblk * E, dtype=torch.int32, device=x.device)
            me = torch.empty(E, dtype=torch.float32, device=x.device)
            L.call('moe_gate_fwd', L.p(x), L.p(P), L.p(blk_cnt), L.p(me), L.c_int(T), L.c_int(E), L.c_int(k), L.stream_ptr(x.device))
            self.fused_launches += 1
            return P
        return _reference_forward(x, k)

    def gradient(self, g):
        # g is the sum of the tokens the pieces returned (or a single token): read the piece gradients off it
        # instead of executing that sum
        tokens = [g] if isinstance(g, _GateGradToken) else list(getattr(g, 'inputs', []))
        if not tokens or not all(isinstance(t, _GateGradToken) for t in tokens):
            raise NotImplementedError('the packed gate tensor is consumed through gate_piece_op only')
        dgates, daux = [None] * self.k, None
        for t in tokens:
            if t.which == 'aux':
                daux = t.inputs[0] if daux is None else daux + t.inputs[0]
            else:
                i = t.which[1]
                dgates[i] = t.inputs[0] if dgates[i] is None else dgates[i] + t.inputs[0]
        return [FusedTopKGateGradOp(self.inputs[0], self, dgates, daux, self.k, ctx=self.raw_ctx)]

    def infer_shape(self, input_shapes):
        T, E = input_shapes[0]
        return (3 * self.k * T + 1 + E,)


class GatePieceOp(Op):
    """One plane of the packed gate tensor, as a view: ('idx' | 'loc' | 'gate', i) -> [T]; 'aux' -> [1]."""

    def __init__(self, packed, which, num_tokens, ctx=None):
        super().__init__(GatePieceOp, [packed], ctx)
        self.which, self.T = which, int(num_tokens)
        self.k = packed.k

    def compute(self, input_vals, output_val=None, stream_handle=None):
        P, T, k = input_vals[0], self.T, self.k
        if self.which == 'aux':
            return P[3 * k * T:3 * k * T + 1]
        plane = {'idx': 0, 'loc': k, 'gate': 2 * k}[self.which[0]] + self.which[1]
        return P[plane * T:(plane + 1) * T]

    def gradient(self, g):
        if self.which != 'aux' and self.which[0] != 'gate':
            return [None]
        return [_GateGradToken(g, self.which, ctx=self.raw_ctx)]

    def infer_shape(self, input_shapes):
        return (1,) if self.which == 'aux' else (self.T,)


class _GateGradToken(Op):
    """Carries the gradient of one piece to FusedTopKGateOp.gradient; never part of an executed graph."""

    def __init__(self, g, which, ctx=None):
        super().__init__(_GateGradToken, [g], ctx)
        self.which = which

    def compute(self, input_vals, output_val=None, stream_handle=None):
        raise RuntimeError('gate gradient token executed: the packed gate tensor has a consumer other than gate_piece_op')

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


class FusedTopKGateGradOp(Op):
    def __init__(self, logits, packed, dgates, daux, k, ctx=None):
        self.k = int(k)
        self.slots = [i for i in range(self.k) if dgates[i] is not None]
        self.has_aux = daux is not None
        super().__init__(FusedTopKGateGradOp, [logits, packed] + [dgates[i] for i in self.slots] + ([daux] if self.has_aux else []), ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x, P = input_vals[0], input_vals[1]
        if x.device.type == 'meta':
            return torch.empty(tuple(x.shape), device='meta')
        T, E = x.shape
        k = self.k
        dgates = [None] * k
        for j, i in enumerate(self.slots):
            dgates[i] = input_vals[2 + j]
        daux = input_vals[2 + len(self.slots)] if self.has_aux else None
        if L.kernels_on(x, P) and x.dtype == torch.float32:
            x, P = x.contiguous(), P.contiguous()
            dgates = [d.contiguous().reshape(-1) if d is not None else None for d in dgates]
            daux = daux.contiguous() if daux is not None else None
            out = torch.empty_like(x)
            ptrs = (ctypes.c_void_p * k)(*[d.data_ptr() if d is not None else None for d in dgates])
            L.call('moe_gate_bwd', L.p(x), L.p(P), ptrs, L.p(daux), L.p(out), L.c_int(T), L.c_int(E), L.c_int(k),
                   L.stream_ptr(x.device))
            return out
        return _reference_backward(x, P, dgates, daux, k)

    def gradient(self, g):
        raise NotImplementedError('second-order gradients of the gate are not defined')

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


def fused_topk_gate_op(logits, k, ctx=None):
    return FusedTopKGateOp(logits, k, ctx=ctx)


def gate_piece_op(packed, which, num_tokens, ctx=None):
    return GatePieceOp(packed, which, num_tokens, ctx=ctx)
