"""Fused top-k MoE gate: softmax + top-k + capacity-slot assignment + balance loss as ONE op.

reference: python/hetu/layers/TopGate.py:15-79 builds the gate from ~30 graph ops (softmax,
topk_idx, one_hot, cumsum_with_bias, reduce_sum / reduce_mean, element-wise products); each is a
kernel launch forward and most have one backward.  ``fused_topk_gate_op(logits, k)`` returns the
packed tensor P [T + 1, 3k] of csrc/moe_gate.cu (expert index / slot / gate value per route,
P[T][0] = balance loss); its gradient is one op as well.  ``layers.moe.topkgating`` slices the
pieces out of P when ``HETU_MOE_FUSED_GATE=1`` (default) - same values as the composed gate."""
from __future__ import annotations

import torch

from .. import _lib as L
from ..graph.node import Op


def _reference_forward(logits: torch.Tensor, k: int) -> torch.Tensor:
    T, E = logits.shape
    p = torch.softmax(logits.float(), dim=1)
    idx = torch.topk(logits, k, dim=1).indices                                   # largest first
    P = torch.zeros((T + 1, 3 * k), dtype=torch.float32, device=logits.device)
    base = torch.zeros(E, dtype=torch.float32, device=logits.device)
    aux = 0.0
    me = p.mean(dim=0)
    for i in range(k):
        onehot = torch.nn.functional.one_hot(idx[:, i], E).float()
        loc = (torch.cumsum(onehot, dim=0) - 1 + base) * onehot
        P[:T, i] = idx[:, i].float()
        P[:T, k + i] = loc.sum(dim=1)
        P[:T, 2 * k + i] = p.gather(1, idx[:, i:i + 1]).squeeze(1)
        base = base + onehot.sum(dim=0)
        aux = aux + (me * onehot.mean(dim=0)).sum() * E
    P[T, 0] = aux
    return P


def _reference_backward(logits: torch.Tensor, P: torch.Tensor, dP: torch.Tensor, k: int) -> torch.Tensor:
    T, E = logits.shape
    p = torch.softmax(logits.float(), dim=1)
    idx = P[:T, :k].long()
    cnt = torch.zeros(E, dtype=torch.float32, device=logits.device)
    cnt.index_add_(0, idx.reshape(-1), torch.ones(T * k, device=logits.device))
    c = dP[T, 0] * E / float(T * T) * cnt                                         # d aux / d p[t, e]
    dp = c.unsqueeze(0).expand(T, E).clone()
    dp.scatter_add_(1, idx, dP[:T, 2 * k:3 * k])
    return p * (dp - (dp * p).sum(dim=1, keepdim=True))


class FusedTopKGateOp(Op):
    def __init__(self, logits, k, ctx=None):
        super().__init__(FusedTopKGateOp, [logits], ctx)
        self.k = int(k)
        self.fused_launches = 0
        #: set by layers.moe.topkgating: the row count its column slices were built for
        self.expected_tokens = None

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        T, E = x.shape
        if self.expected_tokens is not None and int(self.expected_tokens) != int(T):
            raise ValueError('gate was built for num_tokens=%d but the logits have %d rows (pass the total number of tokens '
                             'per step, batch_size * tokens, to the gate)' % (self.expected_tokens, T))
        if x.device.type == 'meta':
            return torch.empty((T + 1, 3 * self.k), device='meta')
        if L.kernels_on(x) and x.dtype == torch.float32 and self.k <= 4 and E * (self.k + 1) * 4 <= 48 * 1024:
            x = x.contiguous()
            P = torch.empty((T + 1, 3 * self.k), dtype=torch.float32, device=x.device)
            nblk = (T + 255) // 256
            blk_cnt = torch.empty(self.k * nblk * E, dtype=torch.int32, device=x.device)
            me = torch.empty(E, dtype=torch.float32, device=x.device)
            L.call('moe_gate_fwd', L.p(x), L.p(P), L.p(blk_cnt), L.p(me), L.c_int(T), L.c_int(E), L.c_int(self.k), L.stream_ptr(x.device))
            self.fused_launches += 1
            return P
        return _reference_forward(x, self.k)

    def gradient(self, g):
        return [FusedTopKGateGradOp(self.inputs[0], self, g, self.k, ctx=self.raw_ctx)]

    def infer_shape(self, input_shapes):
        return (input_shapes[0][0] + 1, 3 * self.k)


class FusedTopKGateGradOp(Op):
    def __init__(self, logits, packed, grad, k, ctx=None):
        super().__init__(FusedTopKGateGradOp, [logits, packed, grad], ctx)
        self.k = int(k)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x, P, dP = input_vals
        if x.device.type == 'meta':
            return torch.empty(tuple(x.shape), device='meta')
        T, E = x.shape
        if L.kernels_on(x, P, dP) and x.dtype == torch.float32:
            x, P, dP = x.contiguous(), P.contiguous(), dP.contiguous()
            out = torch.empty_like(x)
            cnt = torch.empty(E, dtype=torch.float32, device=x.device)
            L.call('moe_gate_bwd', L.p(x), L.p(P), L.p(dP), L.p(cnt), L.p(out), L.c_int(T), L.c_int(E), L.c_int(self.k),
                   L.stream_ptr(x.device))
            return out
        return _reference_backward(x, P, dP, self.k)

    def gradient(self, g):
        raise NotImplementedError('second-order gradients of the gate are not defined')

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


def fused_topk_gate_op(logits, k, ctx=None):
    return FusedTopKGateOp(logits, k, ctx=ctx)
