"""Reductions, broadcasts, n-ary sum, arg/top-k, cumulative sum.

reference op files: ReduceSum ReduceMean ReduceSumAxisZero Sum Norm Max Min
Argmax Argsort TopKIdx TopKVal Cumsum Broadcast BroadcastShape
(python/hetu/gpu_ops/*.py; kernels src/ops/ReduceSum.cu, ReduceSumAxisZero.cu,
TopKIdx.cu, CumSum.cu ...).

A reduction over a contiguous run of axes views the tensor as [A, R, B] and runs
``hetu_reduce_mid`` (warp-shuffle tree when B == 1, coalesced column walk when
B > 1).  The gate's cumulative sum is a block-parallel scan (the reference walks
T tokens serially per thread: src/ops/CumSum.cu:3-18).
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import numpy as np
import torch

from .. import _lib as L
from ..graph.node import Op, as_index
from ..ndarray import IndexedSlices

R_SUM, R_MEAN, R_MAX, R_MIN, R_SQSUM = 0, 1, 2, 3, 4


def _norm_axes(axes, ndim):
    if axes is None:
        return list(range(ndim))
    if isinstance(axes, int):
        axes = [axes]
    return sorted(a % ndim for a in axes)


def _reduce(x: torch.Tensor, axes: List[int], keepdims, mode: int) -> torch.Tensor:
    """Shared implementation for sum/mean/max/min over ``axes``."""
    nd = x.dim()
    if isinstance(keepdims, (list, tuple)):
        kd_list = list(keepdims) + [False] * (len(axes) - len(keepdims))
    else:
        kd_list = [bool(keepdims)] * len(axes)
    out_shape = []
    for i in range(nd):
        if i in axes:
            if kd_list[axes.index(i)]:
                out_shape.append(1)
        else:
            out_shape.append(x.shape[i])
    contiguous_run = len(axes) > 0 and axes == list(range(axes[0], axes[-1] + 1))
    if (L.kernels_on(x) and x.dtype == torch.float32 and contiguous_run and x.numel() > 0):
        x = x.contiguous()
        A = int(np.prod(x.shape[:axes[0]])) if axes[0] > 0 else 1
        Rn = int(np.prod(x.shape[axes[0]:axes[-1] + 1]))
        Bn = int(np.prod(x.shape[axes[-1] + 1:])) if axes[-1] + 1 < nd else 1
        out = torch.empty(out_shape, dtype=x.dtype, device=x.device)
        L.call('reduce_mid', L.p(x), L.p(out), L.c_sz(A), L.c_sz(Rn), L.c_sz(Bn), L.c_int(mode),
               L.stream_ptr(x.device))
        return out
    if len(axes) == 0:
        return x.reshape(out_shape)
    if mode == R_SUM:
        y = x.sum(dim=axes)
    elif mode == R_MEAN:
        y = x.mean(dim=axes)
    elif mode == R_MAX:
        y = x.amax(dim=axes)
    elif mode == R_MIN:
        y = x.amin(dim=axes)
    else:
        y = (x * x).sum(dim=axes)
    return y.reshape(out_shape)


class ReduceSumOp(Op):
    mode = R_SUM

    def __init__(self, node, axes, keepdims=False, ctx=None):
        super().__init__(type(self), [node], ctx)
        self.axes = axes if axes is None or isinstance(axes, (list, tuple)) else [axes]
        self.keepdims = keepdims

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        return _reduce(x, _norm_axes(self.axes, x.dim()), self.keepdims, self.mode)

    def gradient(self, g):
        from .shape import broadcast_like_reduced_op
        scale = None
        return [broadcast_like_reduced_op(g, self.inputs[0], self.axes, self.keepdims,
                                          mean=(self.mode == R_MEAN), ctx=self.raw_ctx)]


class ReduceMeanOp(ReduceSumOp):
    mode = R_MEAN


class ReduceSumAxisZeroOp(Op):
    def __init__(self, node, ctx=None):
        super().__init__(ReduceSumAxisZeroOp, [node], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        if x.dim() == 1:
            return _reduce(x, [0], True, R_SUM)
        return _reduce(x, [0], False, R_SUM)

    def gradient(self, g):
        from .shape import broadcastto_op
        return [broadcastto_op(g, self.inputs[0], ctx=self.raw_ctx)]


class ReduceToShapeOp(Op):
    """Sum ``node`` down to the shape of ``like`` (inverse of numpy broadcasting).
    Used for the gradients of broadcasting binary ops; a no-op when shapes agree."""

    def __init__(self, node, like, ctx=None):
        super().__init__(ReduceToShapeOp, [node, like], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        g, like = input_vals
        if tuple(g.shape) == tuple(like.shape):
            return g
        tgt = tuple(like.shape)
        diff = g.dim() - len(tgt)
        if diff > 0:
            g = _reduce(g, list(range(diff)), False, R_SUM)
        axes = [i for i in range(len(tgt)) if tgt[i] == 1 and g.shape[i] != 1]
        if axes:
            # handle each contiguous run of axes with one kernel
            for ax in reversed(axes):
                g = _reduce(g, [ax], True, R_SUM)
        return g.reshape(tgt)

    def gradient(self, g):
        from .shape import broadcastto_op
        return [broadcastto_op(g, self.inputs[0], ctx=self.raw_ctx), None]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[1])


def reduce_to_shape_op(node, like, ctx=None):
    return ReduceToShapeOp(node, like, ctx=ctx)


class SumOp(Op):
    """n-ary add; IndexedSlices inputs stay sparse when every input is sparse
    (reference Sum.py:13-172)."""

    def __init__(self, node_list, ctx=None, sparse=False):
        super().__init__(SumOp, list(node_list), ctx)
        self.sparse = sparse

    def compute(self, input_vals, output_val=None, stream_handle=None):
        sparse = [v for v in input_vals if isinstance(v, IndexedSlices)]
        dense = [v for v in input_vals if not isinstance(v, IndexedSlices)]
        if sparse and not dense:
            acc = sparse[0]
            for s in sparse[1:]:
                acc = acc.merge(s)
            return acc
        shape = torch.broadcast_shapes(*[tuple(d.shape) for d in dense])
        if len(dense) >= 2 and L.kernels_on(*dense) and all(
                tuple(d.shape) == tuple(shape) and d.dtype == torch.float32 for d in dense):
            out = None
            from .elementwise import _binary_kernel, B
            out = _binary_kernel(B['ADD'], dense[0], dense[1])
            for d in dense[2:]:
                out = _binary_kernel(B['ADD'], out, d)
        else:
            out = dense[0]
            for d in dense[1:]:
                out = out + d
            if len(dense) == 1:
                out = out.clone() if sparse else out
        for s in sparse:
            ind, val = s._iv()
            out = out.index_add(0, ind, val)
        return out

    def gradient(self, g):
        from .elementwise import _unbroadcast
        return [_unbroadcast(g, n, self.raw_ctx) for n in self.inputs]

    def infer_shape(self, input_shapes):
        return tuple(torch.broadcast_shapes(*[tuple(s) for s in input_shapes]))


SparseSumOp = SumOp


class NormOp(Op):
    def __init__(self, node, axis, p=2, keepdims=True, ctx=None):
        super().__init__(NormOp, [node], ctx)
        self.axis, self.p, self.keepdims = axis, p, keepdims

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        return torch.linalg.vector_norm(x, ord=self.p, dim=self.axis, keepdim=self.keepdims)

    def gradient(self, g):
        return [norm_gradient_op(self.inputs[0], self, g, self.axis, self.p, ctx=self.raw_ctx)]


class NormGradientOp(Op):
    def __init__(self, node, node_y, grad_y, axis, p, ctx=None):
        super().__init__(NormGradientOp, [node, node_y, grad_y], ctx)
        self.axis, self.p = axis, p

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x, y, gy = input_vals
        if y.dim() < x.dim():
            y = y.unsqueeze(self.axis)
            gy = gy.unsqueeze(self.axis)
        p = self.p
        if p == 1:
            return gy * x.sign()
        if p == 2:
            return gy * x / y.clamp_min(1e-12)
        return gy * x * x.abs().pow(p - 2) / y.clamp_min(1e-12).pow(p - 1)

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


class _MaxMinOp(Op):
    mode = R_MAX

    def __init__(self, node_A, node_B=None, dim=0, keepdim=False, ctx=None):
        super().__init__(type(self), [node_A] + ([node_B] if node_B is not None else []), ctx)
        self.dim, self.keepdim = dim, keepdim

    def compute(self, input_vals, output_val=None, stream_handle=None):
        if len(input_vals) == 2:
            from .elementwise import _binary_kernel, B, _k
            a, b = input_vals
            if _k(a, b) and a.shape == b.shape:
                return _binary_kernel(B['MAX'] if self.mode == R_MAX else B['MIN'], a, b)
            return torch.maximum(a, b) if self.mode == R_MAX else torch.minimum(a, b)
        x = input_vals[0]
        return _reduce(x, [self.dim % x.dim()], self.keepdim, self.mode)

    def gradient(self, g):
        raise NotImplementedError


class MaxOp(_MaxMinOp):
    mode = R_MAX


class MinOp(_MaxMinOp):
    mode = R_MIN


class ArgmaxOp(Op):
    def __init__(self, node, dim=0, ctx=None):
        super().__init__(ArgmaxOp, [node], ctx)
        self.dim = dim

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        if x.device.type == 'meta':
            return torch.empty([s for i, s in enumerate(x.shape) if i != self.dim % x.dim()],
                               device='meta')
        if L.kernels_on(x) and x.dtype == torch.float32 and self.dim % x.dim() == x.dim() - 1:
            x = x.contiguous()
            out = torch.empty(x.shape[:-1], dtype=torch.float32, device=x.device)
            L.call('argmax_rows', L.p(x), L.p(out), L.c_sz(out.numel()), L.c_sz(x.shape[-1]),
                   L.stream_ptr(x.device))
            return out
        return x.argmax(dim=self.dim).to(torch.float32)

    def gradient(self, g):
        return [None]


class ArgsortOp(Op):
    def __init__(self, node, dim=1, descending=False, ctx=None):
        super().__init__(ArgsortOp, [node], ctx)
        self.dim, self.descending = dim, descending

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        if x.device.type == 'meta':
            return torch.empty_like(x)
        return torch.argsort(x, dim=self.dim, descending=self.descending, stable=True).to(torch.float32)

    def gradient(self, g):
        return [None]


class TopKIdxOp(Op):
    """Indices of the k largest entries of every row, largest first (float-encoded
    for API parity; reference TopKIdx.py / src/ops/TopKIdx.cu:3-26)."""

    def __init__(self, node, topk, ctx=None):
        super().__init__(TopKIdxOp, [node], ctx)
        self.k = topk

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        if x.device.type == 'meta':
            return torch.empty((x.shape[0], self.k), device='meta')
        if L.kernels_on(x) and x.dtype == torch.float32 and x.dim() == 2 and self.k <= 8:
            x = x.contiguous()
            idx = torch.empty((x.shape[0], self.k), dtype=torch.float32, device=x.device)
            L.call('topk_rows', L.p(x), L.p(idx), L.p(None), L.c_sz(x.shape[0]), L.c_sz(x.shape[1]),
                   L.c_int(self.k), L.stream_ptr(x.device))
            return idx
        return torch.topk(x, self.k, dim=-1).indices.to(torch.float32)

    def gradient(self, g):
        return [None]


class TopKValOp(Op):
    def __init__(self, node, node_idx, ctx=None):
        super().__init__(TopKValOp, [node, node_idx], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x, idx = input_vals
        if x.device.type == 'meta':
            return torch.empty(tuple(idx.shape), device='meta')
        return torch.gather(x, -1, as_index(idx))

    def gradient(self, g):
        return [None, None]


class CumsumOp(Op):
    """Inclusive prefix sum along ``dim`` plus ``bias`` (MoE slot index = cumsum-1)."""

    def __init__(self, node, bias=-1, dim=0, ctx=None):
        super().__init__(CumsumOp, [node], ctx)
        self.bias, self.dim = bias, dim

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        if x.device.type == 'meta':
            return torch.empty_like(x)
        if (L.kernels_on(x) and x.dtype == torch.float32 and x.dim() == 2 and self.dim == 0
                and x.shape[1] <= 1024):
            x = x.contiguous()
            out = torch.empty_like(x)
            L.call('cumsum_cols', L.p(x), L.p(out), L.c_sz(x.shape[0]), L.c_sz(x.shape[1]),
                   L.c_f32(float(self.bias)), L.stream_ptr(x.device))
            return out
        return torch.cumsum(x, dim=self.dim) + self.bias

    def gradient(self, g):
        return [None]


# ---------------------------------------------------------------- public API
def reduce_sum_op(node, axes, keepdims=False, ctx=None):
    return ReduceSumOp(node, axes, keepdims, ctx=ctx)


def reduce_mean_op(node, axes, keepdims=False, ctx=None):
    return ReduceMeanOp(node, axes, keepdims, ctx=ctx)


def reducesumaxiszero_op(node, ctx=None):
    return ReduceSumAxisZeroOp(node, ctx=ctx)


def sum_op(node_list, ctx=None, sparse=False):
    return SumOp(node_list, ctx=ctx, sparse=sparse)


def norm_op(node, axis, p=2, keepdims=True, ctx=None):
    return NormOp(node, axis, p, keepdims, ctx=ctx)


def norm_gradient_op(node, node_y, grad_y, axis, p, ctx=None):
    return NormGradientOp(node, node_y, grad_y, axis, p, ctx=ctx)


def max_op(node_A, node_B=None, dim=0, keepdim=False, ctx=None):
    return MaxOp(node_A, node_B, dim, keepdim, ctx=ctx)


def min_op(node_A, node_B=None, dim=0, keepdim=False, ctx=None):
    return MinOp(node_A, node_B, dim, keepdim, ctx=ctx)


def argmax_op(node, dim=0, ctx=None):
    return ArgmaxOp(node, dim, ctx=ctx)


def argsort_op(node, dim=1, descending=False, ctx=None):
    return ArgsortOp(node, dim, descending, ctx=ctx)


def topk_idx_op(node, topk, ctx=None):
    return TopKIdxOp(node, topk, ctx=ctx)


def topk_val_op(nodeA, nodeB, ctx=None):
    return TopKValOp(nodeA, nodeB, ctx=ctx)


def cumsum_with_bias_op(node, bias=-1, dim=0, ctx=None):
    return CumsumOp(node, bias=bias, dim=dim, ctx=ctx)
