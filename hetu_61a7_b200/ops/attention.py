"""Multi-head attention as one graph op (new capability).

The reference has no attention op: BERT builds it from five ops - batch_matmul, additive mask,
softmax, dropout, batch_matmul (examples/nlp/bert/hetu_bert.py:220-270) - and
autodiff adds six more for the backward pass.  ``fused_attention_op(q, k, v, mask)`` computes
softmax(q k^T * scale + mask) v with two batched tcgen05 GEMMs and ONE row kernel in between
(csrc/attention.cu: scale + key mask + softmax + log-sum-exp in place on the score buffer); the
backward pass is four batched GEMMs and one row kernel (ds = p * (dp - <dp, p>) * scale in place).
"""
from __future__ import annotations

import math
import os

import torch

from .. import _lib as L
from ..graph.node import Op


def _k(*ts):
    return L.kernels_on(*ts) and all(t.dtype == torch.float32 for t in ts)


class FusedAttentionOp(Op):
    """q, k, v: [B, H, S, D]; mask (optional): additive, broadcastable to [B, 1, 1, S]."""

    def __init__(self, q, k, v, mask=None, scale=None, ctx=None):
        super().__init__(FusedAttentionOp, [q, k, v] + ([mask] if mask is not None else []), ctx)
        self.scale = scale
        self.lse = None
        self.probs = None

    def compute(self, input_vals, output_val=None, stream_handle=None):
        q, k, v = input_vals[:3]
        mask = input_vals[3] if len(input_vals) > 3 else None
        if q.device.type == 'meta':
            return torch.empty_like(q)
        scale = self.scale if self.scale is not None else 1.0 / math.sqrt(q.shape[-1])
        B, H, S, D = q.shape
        Sk = k.shape[2]
        if _k(q, k, v) and (mask is None or mask.numel() == B * Sk) and D in (32, 64) \
                and os.environ.get('HETU_FLASH_ATTN', '1') == '1':
            # one kernel, scores stay in TMEM / shared memory (csrc/attention_sm100.cu)
            q, k, v = q.contiguous(), k.contiguous(), v.contiguous()
            out = torch.empty_like(q)
            self.lse = torch.empty((B, H, S), dtype=torch.float32, device=q.device)
            self.probs = None
            m = mask.contiguous() if mask is not None else None
            L.call('flash_attn_fwd', L.p(q), L.p(k), L.p(v), L.p(m), L.p(out), L.p(self.lse), L.c_int(B), L.c_int(H),
                   L.c_int(S), L.c_int(Sk), L.c_int(D), L.c_f32(scale), L.stream_ptr(q.device))
            self.flash_launches = getattr(self, 'flash_launches', 0) + 1
            return out
        if _k(q, k, v) and (mask is None or mask.numel() == B * Sk):
            from .linalg import gemm
            q3, k3, v3 = (t.contiguous().view(B * H, t.shape[2], D) for t in (q, k, v))
            p = gemm(q3, k3, False, True)                       # [B*H, S, Sk] scores
            self.lse = torch.empty((B, H, S), dtype=torch.float32, device=q.device)
            m = mask.contiguous() if mask is not None else None
            L.call('attn_softmax', L.p(p), L.p(m), L.p(self.lse), L.c_i64(B * H * S), L.c_int(Sk),
                   L.c_i64(H * S), L.c_f32(scale), L.stream_ptr(q.device))
            self.probs = p
            return gemm(p, v3).view(B, H, S, D)
        s = torch.matmul(q, k.transpose(-1, -2)) * scale
        if mask is not None:
            s = s + mask.reshape(B, 1, 1, Sk) if mask.numel() == B * Sk else s + mask
        self.lse = torch.logsumexp(s, dim=-1)
        return torch.matmul(torch.softmax(s, dim=-1), v)

    def gradient(self, g):
        grad = FusedAttentionGradOp(self, g, ctx=self.raw_ctx)
        outs = [_Pick(grad, self.inputs[i], i, ctx=self.raw_ctx) for i in range(3)]
        return outs + ([None] if len(self.inputs) > 3 else [])

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


class FusedAttentionGradOp(Op):
    def __init__(self, fwd, grad, ctx=None):
        super().__init__(FusedAttentionGradOp, [grad, fwd] + list(fwd.inputs), ctx)
        self.forward_node = fwd
        self.results = None

    def compute(self, input_vals, output_val=None, stream_handle=None):
        do, o, q, k, v = input_vals[:5]
        mask = input_vals[5] if len(input_vals) > 5 else None
        if q.device.type == 'meta':
            return torch.empty_like(q)
        fwd = self.forward_node
        scale = fwd.scale if fwd.scale is not None else 1.0 / math.sqrt(q.shape[-1])
        B, H, S, D = q.shape
        Sk = k.shape[2]
        have_p = getattr(fwd, 'probs', None) is not None
        if _k(q, k, v, do, o) and (have_p or fwd.lse is not None) and (mask is None or mask.numel() == B * Sk):
            from .linalg import gemm
            q3, k3, v3, do3 = (t.contiguous().view(B * H, t.shape[2], D) for t in (q, k, v, do))
            if have_p:
                p = fwd.probs
            else:           # flash forward kept only the log-sum-exp: p = exp(q k^T * scale + mask - lse)
                p = gemm(q3, k3, False, True)
                m = mask.contiguous() if mask is not None else None
                L.call('attn_probs_from_lse', L.p(p), L.p(m), L.p(fwd.lse), L.c_i64(B * H * S), L.c_int(Sk),
                       L.c_i64(H * S), L.c_f32(scale), L.stream_ptr(q.device))
            dv = gemm(p, do3, True, False).view(B, H, Sk, D)     # P^T dO
            ds = gemm(do3, v3, False, True)                      # dP = dO V^T, turned into dS in place
            L.call('attn_softmax_bwd', L.p(p), L.p(ds), L.c_i64(B * H * S), L.c_int(Sk), L.c_f32(scale),
                   L.stream_ptr(q.device))
            dq = gemm(ds, k3).view(B, H, S, D)
            dk = gemm(ds, q3, True, False).view(B, H, Sk, D)
        else:
            s = torch.matmul(q, k.transpose(-1, -2)) * scale
            if mask is not None:
                s = s + (mask.reshape(B, 1, 1, Sk) if mask.numel() == B * Sk else mask)
            p = torch.exp(s - fwd.lse.unsqueeze(-1))
            dv = torch.matmul(p.transpose(-1, -2), do)
            dp = torch.matmul(do, v.transpose(-1, -2))
            delta = (do * o).sum(dim=-1, keepdim=True)
            ds = p * (dp - delta) * scale
            dq = torch.matmul(ds, k)
            dk = torch.matmul(ds.transpose(-1, -2), q)
        self.results = (dq, dk, dv)
        return dq

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[2])


class _Pick(Op):
    def __init__(self, grad_node, like, which, ctx=None):
        super().__init__(_Pick, [grad_node, like], ctx)
        self.which = which
        self.op_type = 'AttentionGradPick'

    def compute(self, input_vals, output_val=None, stream_handle=None):
        like = input_vals[1]
        if like.device.type == 'meta':
            return torch.empty_like(like)
        return self.inputs[0].results[self.which]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[1])


def fused_attention_op(q, k, v, mask=None, scale=None, ctx=None):
    return FusedAttentionOp(q, k, v, mask, scale, ctx=ctx)
