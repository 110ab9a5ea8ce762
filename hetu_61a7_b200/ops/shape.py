"""Shape / layout / indexing ops.

reference op files: Reshape Transpose Broadcast BroadcastShape Concat
Concatenate Split Slice SliceAssign SliceByMatrix Gather Scatter Scatter1D
Indexing Repeat Roll OneHot Pad Interpolate (python/hetu/gpu_ops/*.py, kernels
src/ops/<same>.cu - one index-arithmetic kernel per op with device-side shape
buffers).

Here every layout op is expressed as a *strided view* (free) followed, when a
dense result is required, by ONE generic kernel, ``hetu_strided_copy`` (shape
and strides passed by value, innermost dimension vectorised when both sides are
unit-stride).  Concatenate is n launches into slices of one output.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional, Sequence

import numpy as np
import torch

from .. import _lib as L
from ..graph.node import Op, as_index

_I64x8 = ctypes.c_int64 * 8


def strided_copy(src: torch.Tensor, dst: torch.Tensor) -> None:
    """dst[...] = src[...] for equal-shaped (possibly non-contiguous) fp32 views."""
    assert tuple(src.shape) == tuple(dst.shape), (src.shape, dst.shape)
    if src.numel() == 0:
        return
    if not (L.kernels_on(src, dst) and src.dtype == torch.float32 and dst.dtype == torch.float32
            and src.dim() <= 8):
        dst.copy_(src)
        return
    # drop size-1 dims and merge adjacent dims that are contiguous on both sides
    dims = [(s, a, b) for s, a, b in zip(src.shape, src.stride(), dst.stride()) if s != 1]
    merged = []
    for s, a, b in dims:
        if merged and merged[-1][1] == a * s and merged[-1][2] == b * s:
            ps, pa, pb = merged[-1]
            merged[-1] = (ps * s, a, b)
        else:
            merged.append((s, a, b))
    if not merged:
        merged = [(1, 1, 1)]
    nd = len(merged)
    # batched 2-D transpose (NCHW <-> NHWC around the convolution GEMMs): dst [B, X, Y] contiguous, src laid out [B, Y, X]
    m3 = merged if nd == 3 else ([(1, 0, 0)] + merged if nd == 2 else None)
    if m3 is not None:
        (bn, ba, bb), (x, xa, xb), (y, ya, yb) = m3
        if xb == y and yb == 1 and xa == 1 and ya == x and (bn == 1 or (ba == x * y and bb == x * y)) and bn <= 65535 \
                and x >= 8 and y >= 8:
            L.call('transpose_batched', L.p(src), L.p(dst), L.c_int(bn), L.c_int(y), L.c_int(x), L.stream_ptr(src.device))
            return
    shp = _I64x8(*([m[0] for m in merged] + [1] * (8 - nd)))
    ss = _I64x8(*([m[1] for m in merged] + [0] * (8 - nd)))
    ds = _I64x8(*([m[2] for m in merged] + [0] * (8 - nd)))
    L.call('strided_copy', L.p(src), L.p(dst), L.c_int(nd), shp, ss, ds, L.c_sz(src.numel()),
           L.stream_ptr(src.device))


def materialize(view: torch.Tensor) -> torch.Tensor:
    if view.is_contiguous():
        return view
    if view.device.type != 'cuda':
        return view.contiguous()
    out = torch.empty(tuple(view.shape), dtype=view.dtype, device=view.device)
    strided_copy(view, out)
    return out


def _norm_shape(shape, numel):
    shape = [int(s) for s in shape]
    if -1 in shape:
        known = int(np.prod([s for s in shape if s != -1])) or 1
        shape[shape.index(-1)] = numel // known
    return tuple(shape)


class ArrayReshapeOp(Op):
    def __init__(self, node, output_shape, ctx=None):
        super().__init__(ArrayReshapeOp, [node], ctx)
        self.output_shape = list(output_shape)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        return materialize(x).reshape(_norm_shape(self.output_shape, x.numel()))

    def gradient(self, g):
        return [array_reshape_gradient_op(self.inputs[0], g, ctx=self.raw_ctx)]

    def infer_shape(self, input_shapes):
        return _norm_shape(self.output_shape, int(np.prod(input_shapes[0])))


class ArrayReshapeGradientOp(Op):
    def __init__(self, node_in, node_out, ctx=None):
        super().__init__(ArrayReshapeGradientOp, [node_in, node_out], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x, g = input_vals
        return materialize(g).reshape(tuple(x.shape))

    def gradient(self, g):
        raise NotImplementedError

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


class TransposeOp(Op):
    def __init__(self, node_A, perm=None, ctx=None):
        super().__init__(TransposeOp, [node_A], ctx)
        self.perm = list(perm) if perm is not None else None

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        perm = self.perm if self.perm is not None else list(range(x.dim()))[::-1]
        return materialize(x.permute(*perm))

    def gradient(self, g):
        if self.perm is None:
            return [transpose_op(g, None, ctx=self.raw_ctx)]
        inv = [0] * len(self.perm)
        for i, p in enumerate(self.perm):
            inv[p] = i
        return [transpose_op(g, inv, ctx=self.raw_ctx)]


class BroadcastToOp(Op):
    """Broadcast node_A to the shape of node_B (numpy rules)."""

    def __init__(self, node_A, node_B, ctx=None):
        super().__init__(BroadcastToOp, [node_A, node_B], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        a, b = input_vals
        return materialize(a.expand(tuple(b.shape)))

    def gradient(self, g):
        from .reduce import reduce_to_shape_op
        return [reduce_to_shape_op(g, self.inputs[0], ctx=self.raw_ctx), None]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[1])


class BroadcastShapeOp(Op):
    """Broadcast to an explicit shape; ``add_axes`` lists the output axes that
    are inserted (size-1) before expanding.  reference BroadcastShape.py."""

    def __init__(self, node_A, shape, add_axes=(), ctx=None):
        super().__init__(BroadcastShapeOp, [node_A], ctx)
        self.shape, self.add_axes = tuple(shape), tuple(add_axes)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        for ax in sorted(self.add_axes):
            x = x.unsqueeze(ax)
        return materialize(x.expand(self.shape))

    def gradient(self, g):
        from .reduce import reduce_sum_op, reduce_to_shape_op
        if self.add_axes:
            return [reduce_sum_op(g, list(self.add_axes), keepdims=False, ctx=self.raw_ctx)]
        return [reduce_to_shape_op(g, self.inputs[0], ctx=self.raw_ctx)]

    def infer_shape(self, input_shapes):
        return self.shape


class BroadcastLikeReducedOp(Op):
    """Gradient of a sum/mean reduction: re-insert reduced axes and expand to
    the shape of the forward input (scaled by 1/count for mean)."""

    def __init__(self, grad, like, axes, keepdims, mean=False, ctx=None):
        super().__init__(BroadcastLikeReducedOp, [grad, like], ctx)
        self.axes, self.keepdims, self.mean = axes, keepdims, mean

    def compute(self, input_vals, output_val=None, stream_handle=None):
        g, like = input_vals
        nd = like.dim()
        axes = list(range(nd)) if self.axes is None else sorted(a % nd for a in (
            self.axes if isinstance(self.axes, (list, tuple)) else [self.axes]))
        shp = [1 if i in axes else like.shape[i] for i in range(nd)]
        g = g.reshape(shp)
        if self.mean:
            cnt = 1
            for a in axes:
                cnt *= like.shape[a]
            from .elementwise import _unary_kernel, U, _k
            g = _unary_kernel(U['MUL_CONST'], g, 1.0 / cnt) if _k(g) else g * (1.0 / cnt)
        return materialize(g.expand(tuple(like.shape)))

    def gradient(self, g):
        raise NotImplementedError

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[1])


def broadcast_like_reduced_op(grad, like, axes, keepdims, mean=False, ctx=None):
    return BroadcastLikeReducedOp(grad, like, axes, keepdims, mean, ctx=ctx)


def _cat(vals: Sequence[torch.Tensor], axis: int) -> torch.Tensor:
    if vals[0].device.type != 'cuda' or not all(v.dtype == torch.float32 for v in vals):
        return torch.cat(list(vals), dim=axis)
    shape = list(vals[0].shape)
    shape[axis] = sum(v.shape[axis] for v in vals)
    out = torch.empty(shape, dtype=vals[0].dtype, device=vals[0].device)
    off = 0
    for v in vals:
        strided_copy(v, out.narrow(axis, off, v.shape[axis]))
        off += v.shape[axis]
    return out


class ConcatOp(Op):
    def __init__(self, node_A, node_B, axis=0, ctx=None):
        super().__init__(ConcatOp, [node_A, node_B], ctx)
        self.axis = axis

    def compute(self, input_vals, output_val=None, stream_handle=None):
        return _cat(input_vals, self.axis)

    def gradient(self, g):
        return [concat_gradient_op(g, self.inputs[0], self.axis, 0, ctx=self.raw_ctx),
                concat_gradient_op(g, self.inputs[1], self.axis, 1, ctx=self.raw_ctx)]


class ConcatGradientOp(Op):
    def __init__(self, grad_node, input_node, axis, idx, ctx=None):
        super().__init__(ConcatGradientOp, [grad_node, input_node], ctx)
        self.axis, self.idx = axis, idx

    def compute(self, input_vals, output_val=None, stream_handle=None):
        g, x = input_vals
        n = x.shape[self.axis]
        start = 0 if self.idx == 0 else g.shape[self.axis] - n
        return materialize(g.narrow(self.axis, start, n))

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[1])


class ConcatenateOp(Op):
    def __init__(self, node_list, axis=0, ctx=None):
        super().__init__(ConcatenateOp, list(node_list), ctx)
        self.axis = axis

    def compute(self, input_vals, output_val=None, stream_handle=None):
        return _cat(input_vals, self.axis)

    def gradient(self, g):
        return [ConcatenateGradientOp(g, self.inputs, i, self.axis, ctx=self.raw_ctx)
                for i in range(len(self.inputs))]


class ConcatenateGradientOp(Op):
    def __init__(self, grad_node, input_nodes, idx, axis, ctx=None):
        super().__init__(ConcatenateGradientOp, [grad_node] + list(input_nodes), ctx)
        self.axis, self.idx = axis, idx

    def compute(self, input_vals, output_val=None, stream_handle=None):
        g, xs = input_vals[0], input_vals[1:]
        off = sum(x.shape[self.axis] for x in xs[:self.idx])
        return materialize(g.narrow(self.axis, off, xs[self.idx].shape[self.axis]))

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[1 + self.idx])


def _split_bounds(shape, axes, indices, splits):
    begin = [0] * len(shape)
    size = list(shape)
    for ax, ind, spl in zip(axes, indices, splits):
        part = shape[ax] // spl
        begin[ax] = ind * part
        size[ax] = part if ind != spl - 1 else shape[ax] - begin[ax]
    return begin, size


class SplitOp(Op):
    """Take part ``indices[i]`` of ``splits[i]`` equal parts along ``axes[i]``."""

    def __init__(self, node_A, axes, indices, splits, ctx=None):
        super().__init__(SplitOp, [node_A], ctx)
        self.axes, self.indices, self.splits = list(axes), list(indices), list(splits)
        assert len(self.axes) == len(self.splits)
        assert all(0 <= x < self.splits[i] for i, x in enumerate(self.indices))

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        begin, size = _split_bounds(list(x.shape), self.axes, self.indices, self.splits)
        for ax in self.axes:
            x = x.narrow(ax, begin[ax], size[ax])
        return materialize(x)

    def gradient(self, g):
        return [SplitGradientOp(g, self.inputs[0], self.axes, self.indices, self.splits,
                                ctx=self.raw_ctx)]


class SplitGradientOp(Op):
    def __init__(self, node, like, axes, indices, splits, ctx=None):
        super().__init__(SplitGradientOp, [node] + ([like] if like is not None else []), ctx)
        self.axes, self.indices, self.splits = list(axes), list(indices), list(splits)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        g = input_vals[0]
        if len(input_vals) > 1:
            full = list(input_vals[1].shape)
        else:
            full = list(g.shape)
            for ax, spl in zip(self.axes, self.splits):
                full[ax] = g.shape[ax] * spl
        begin, size = _split_bounds(full, self.axes, self.indices, self.splits)
        out = torch.zeros(full, dtype=g.dtype, device=g.device)
        view = out
        for ax in self.axes:
            view = view.narrow(ax, begin[ax], size[ax])
        strided_copy(g, view)
        return out


class SliceOp(Op):
    def __init__(self, node, begin, size, ctx=None):
        super().__init__(SliceOp, [node], ctx)
        self.begin, self.size = list(begin), list(size)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        for d, (b, s) in enumerate(zip(self.begin, self.size)):
            n = x.shape[d] - b if s == -1 else s
            x = x.narrow(d, b, n)
        return materialize(x)

    def gradient(self, g):
        return [SliceGradientOp(g, self.begin, None, like=self.inputs[0], ctx=self.raw_ctx)]


class SliceGradientOp(Op):
    def __init__(self, node, begin, size=None, like=None, ctx=None):
        super().__init__(SliceGradientOp, [node] + ([like] if like is not None else []), ctx)
        self.begin, self.size = list(begin), (list(size) if size is not None else None)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        g = input_vals[0]
        full = list(input_vals[1].shape) if len(input_vals) > 1 else self.size
        out = torch.zeros(full, dtype=g.dtype, device=g.device)
        view = out
        for d, b in enumerate(self.begin):
            view = view.narrow(d, b, g.shape[d])
        strided_copy(g, view)
        return out


class SliceAssignOp(Op):
    def __init__(self, node, begin, size, val, ctx=None):
        super().__init__(SliceAssignOp, [node], ctx)
        self.begin, self.size, self.val = list(begin), list(size), float(val)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        out = x.clone()
        view = out
        for d, (b, s) in enumerate(zip(self.begin, self.size)):
            view = view.narrow(d, b, x.shape[d] - b if s == -1 else s)
        if x.device.type != 'meta':
            view.fill_(self.val)
        return out

    def gradient(self, g):
        return [g]


class SliceAssignMatrixOp(Op):
    def __init__(self, node_A, node_B, begin_A, size_A, begin_B, size_B, ctx=None):
        super().__init__(SliceAssignMatrixOp, [node_A, node_B], ctx)
        self.begin_A, self.size_A = list(begin_A), list(size_A)
        self.begin_B, self.size_B = list(begin_B), list(size_B)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        a, b = input_vals
        out = a.clone()
        va, vb = out, b
        for d in range(a.dim()):
            na = a.shape[d] - self.begin_A[d] if self.size_A[d] == -1 else self.size_A[d]
            nb = b.shape[d] - self.begin_B[d] if self.size_B[d] == -1 else self.size_B[d]
            va = va.narrow(d, self.begin_A[d], na)
            vb = vb.narrow(d, self.begin_B[d], nb)
        strided_copy(vb, va) if a.device.type != 'meta' else None
        return out

    def gradient(self, g):
        return [g, None]


class SliceByMatrixOp(Op):
    """out[r, :] = A[index1[r], index2[r], :] for a 3-D A."""

    def __init__(self, node_A, index1, index2, ctx=None):
        super().__init__(SliceByMatrixOp, [node_A, index1, index2], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        a, i1, i2 = input_vals
        if a.device.type == 'meta':
            return torch.empty((i1.shape[0], a.shape[2]), device='meta')
        return a[as_index(i1).reshape(-1), as_index(i2).reshape(-1)]

    def gradient(self, g):
        return [slice_by_matrix_gradient_op(self.inputs[0], g, self.inputs[1], self.inputs[2],
                                            ctx=self.raw_ctx), None, None]


class SliceByMatrixGradientOp(Op):
    def __init__(self, input, grad, index1, index2, ctx=None):
        super().__init__(SliceByMatrixGradientOp, [input, grad, index1, index2], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        a, g, i1, i2 = input_vals
        if a.device.type == 'meta':
            return torch.empty_like(a)
        out = torch.zeros_like(a)
        out.index_put_((as_index(i1).reshape(-1), as_index(i2).reshape(-1)), g, accumulate=True)
        return out

    def gradient(self, g):
        return [g]


class GatherOp(Op):
    def __init__(self, node, dim, index, ctx=None):
        super().__init__(GatherOp, [node, index], ctx)
        self.dim = dim

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x, idx = input_vals
        if x.device.type == 'meta':
            return torch.empty(tuple(idx.shape), device='meta')
        return torch.gather(x, self.dim, as_index(idx))

    def gradient(self, g):
        return [gather_gradient_op(self.inputs[0], g, self.dim, self.inputs[1], ctx=self.raw_ctx),
                None]


class GatherGradientOp(Op):
    def __init__(self, input, grad, dim, index, ctx=None):
        super().__init__(GatherGradientOp, [input, grad, index], ctx)
        self.dim = dim

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x, g, idx = input_vals
        if x.device.type == 'meta':
            return torch.empty_like(x)
        return torch.zeros_like(x).scatter_add_(self.dim, as_index(idx), g)


class ScatterOp(Op):
    """target.scatter_(dim=0 by default, index, src) - returns the updated copy."""

    def __init__(self, node_target, node_index, node_src, dim=0, ctx=None):
        super().__init__(ScatterOp, [node_target, node_index, node_src], ctx)
        self.dim = dim

    def compute(self, input_vals, output_val=None, stream_handle=None):
        t, idx, src = input_vals
        if t.device.type == 'meta':
            return torch.empty_like(t)
        return t.scatter(self.dim, as_index(idx), src)

    def gradient(self, g):
        return [g, None, None]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


class Scatter1DOp(Op):
    """out[index[i]] = input[i]"""

    def __init__(self, input, index, ctx=None):
        super().__init__(Scatter1DOp, [input, index], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x, idx = input_vals
        if x.device.type == 'meta':
            return torch.empty_like(x)
        out = torch.zeros_like(x)
        out[as_index(idx)] = x
        return out

    def gradient(self, g):
        return [scatter1d_grad_op(g, self.inputs[1], ctx=self.raw_ctx), None]


class Scatter1DGradOp(Op):
    def __init__(self, input, index, ctx=None):
        super().__init__(Scatter1DGradOp, [input, index], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        g, idx = input_vals
        if g.device.type == 'meta':
            return torch.empty_like(g)
        return g[as_index(idx)]


def _gather_rows(x: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    if L.kernels_on(x) and x.dtype == torch.float32 and x.dim() == 2:
        from .embedding import gather_rows_kernel
        return gather_rows_kernel(x, idx)
    return x[as_index(idx)]


class IndexingOp(Op):
    """out[i, :] = input[index[i], :]"""

    def __init__(self, input, index, ctx=None):
        super().__init__(IndexingOp, [input, index], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x, idx = input_vals
        if x.device.type == 'meta':
            return torch.empty_like(x)
        return _gather_rows(x, idx.reshape(-1))

    def gradient(self, g):
        return [indexing_grad_op(g, self.inputs[1], ctx=self.raw_ctx), None]


class IndexingGradOp(Op):
    """grad_in[index[i], :] = grad_out[i, :]"""

    def __init__(self, input, index, ctx=None):
        super().__init__(IndexingGradOp, [input, index], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        g, idx = input_vals
        if g.device.type == 'meta':
            return torch.empty_like(g)
        out = torch.zeros_like(g)
        out[as_index(idx).reshape(-1)] = g
        return out


class RepeatOp(Op):
    def __init__(self, node_A, reps, ctx=None):
        super().__init__(RepeatOp, [node_A], ctx)
        self.reps = reps

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        reps = (self.reps,) if isinstance(self.reps, int) else tuple(self.reps)
        if len(reps) < x.dim():
            reps = (1,) * (x.dim() - len(reps)) + reps
        return x.repeat(*reps)

    def gradient(self, g):
        return [repeat_gradient_op(self.inputs[0], g, ctx=self.raw_ctx)]


class RepeatGradientOp(Op):
    def __init__(self, node_input, node_grad, ctx=None):
        super().__init__(RepeatGradientOp, [node_input, node_grad], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x, g = input_vals
        if x.device.type == 'meta':
            return torch.empty_like(x)
        shp = list(x.shape)
        shp = [1] * (g.dim() - len(shp)) + shp
        view_shape, red = [], []
        for i, (gs, xs) in enumerate(zip(g.shape, shp)):
            view_shape += [gs // xs, xs]
            red.append(2 * i)
        return g.reshape(view_shape).sum(dim=red).reshape(tuple(x.shape))

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


class RollOp(Op):
    def __init__(self, node, shift=0, axis=None, ctx=None):
        super().__init__(RollOp, [node], ctx)
        self.shift, self.axis = shift, axis

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        if self.axis is None:
            return torch.roll(x.reshape(-1), self.shift).reshape(x.shape)
        return torch.roll(x, self.shift, self.axis)

    def gradient(self, g):
        neg = [-s for s in self.shift] if isinstance(self.shift, (list, tuple)) else -self.shift
        return [roll_op(g, neg, self.axis, ctx=self.raw_ctx)]


class OneHotOp(Op):
    def __init__(self, node, num_classes, ctx=None):
        super().__init__(OneHotOp, [node], ctx)
        self.num_classes = int(num_classes)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        if x.device.type == 'meta':
            return torch.empty(tuple(x.shape) + (self.num_classes,), device='meta')
        if L.kernels_on(x) and x.dtype == torch.float32:
            x = x.contiguous()
            out = torch.empty(tuple(x.shape) + (self.num_classes,), dtype=torch.float32,
                              device=x.device)
            L.call('one_hot', L.p(x), L.p(out), L.c_sz(x.numel()), L.c_int(self.num_classes),
                   L.stream_ptr(x.device))
            return out
        return torch.nn.functional.one_hot(as_index(x), self.num_classes).to(torch.float32)

    def gradient(self, g):
        return [None]


class PadOp(Op):
    def __init__(self, node_A, paddings, mode="CONSTANT", constant_values=0, ctx=None):
        super().__init__(PadOp, [node_A], ctx)
        self.paddings, self.mode, self.constant_values = paddings, mode, constant_values

    def _flat(self, nd):
        pads = list(self.paddings)
        pads = [(0, 0)] * (nd - len(pads)) + [tuple(p) for p in pads]
        flat = []
        for lo, hi in reversed(pads):
            flat += [lo, hi]
        return pads, flat

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        pads, flat = self._flat(x.dim())
        mode = {'CONSTANT': 'constant', 'REFLECT': 'reflect', 'SYMMETRIC': 'replicate'}[self.mode.upper()]
        if mode == 'constant':
            return torch.nn.functional.pad(x, flat, mode=mode, value=float(self.constant_values))
        return torch.nn.functional.pad(x, flat[:2 * (x.dim() - 1)], mode=mode)

    def gradient(self, g):
        return [pad_gradient_op(g, self.paddings, self.mode, ctx=self.raw_ctx)]


class PadGradientOp(Op):
    def __init__(self, node_A, paddings, mode="CONSTANT", ctx=None):
        super().__init__(PadGradientOp, [node_A], ctx)
        self.paddings, self.mode = paddings, mode

    def compute(self, input_vals, output_val=None, stream_handle=None):
        g = input_vals[0]
        pads = [(0, 0)] * (g.dim() - len(self.paddings)) + [tuple(p) for p in self.paddings]
        for d, (lo, hi) in enumerate(pads):
            g = g.narrow(d, lo, g.shape[d] - lo - hi)
        return materialize(g)


class InterpolateOp(Op):
    def __init__(self, input, size=None, scale_factor=None, mode='bicubic', align_corners=False,
                 ctx=None):
        super().__init__(InterpolateOp, [input], ctx)
        self.size, self.scale_factor, self.mode, self.align_corners = size, scale_factor, mode, align_corners

    def compute(self, input_vals, output_val=None, stream_handle=None):
        return torch.nn.functional.interpolate(
            input_vals[0], size=self.size, scale_factor=self.scale_factor, mode=self.mode,
            align_corners=self.align_corners)

    def gradient(self, g):
        return [interpolate_grad_op(g, self.inputs[0], self.mode, self.align_corners, ctx=self.raw_ctx)]


class InterpolateGradOp(Op):
    def __init__(self, grad, input, mode='bicubic', align_corners=False, ctx=None):
        super().__init__(InterpolateGradOp, [grad, input], ctx)
        self.mode, self.align_corners = mode, align_corners

    def compute(self, input_vals, output_val=None, stream_handle=None):
        g, x = input_vals
        if x.device.type == 'meta':
            return torch.empty_like(x)
        with torch.enable_grad():
            xx = x.detach().requires_grad_(True)
            y = torch.nn.functional.interpolate(xx, size=tuple(g.shape[2:]), mode=self.mode,
                                                align_corners=self.align_corners)
            (gx,) = torch.autograd.grad(y, xx, g)
        return gx


# ---------------------------------------------------------------- public API
def array_reshape_op(node, output_shape, ctx=None):
    return ArrayReshapeOp(node, output_shape, ctx=ctx)


def array_reshape_gradient_op(node_in, node_out, ctx=None):
    return ArrayReshapeGradientOp(node_in, node_out, ctx=ctx)


def transpose_op(node_A, perm=None, ctx=None):
    return TransposeOp(node_A, perm, ctx=ctx)


def broadcastto_op(node_A, node_B, ctx=None):
    return BroadcastToOp(node_A, node_B, ctx=ctx)


def broadcast_shape_op(node_A, shape, add_axes=(), ctx=None):
    return BroadcastShapeOp(node_A, shape, add_axes, ctx=ctx)


def concat_op(node_A, node_B, axis=0, ctx=None):
    return ConcatOp(node_A, node_B, axis, ctx=ctx)


def concat_gradient_op(grad_node, input_node, axis, idx, ctx=None):
    return ConcatGradientOp(grad_node, input_node, axis, idx, ctx=ctx)


def concatenate_op(node_list, axis=0, ctx=None):
    return ConcatenateOp(node_list, axis, ctx=ctx)


def concatenate_gradient_op(grad_node, input_node, axis, ctx=None):
    """Gradient for one input of a concatenate; ``input_node`` carries its
    position in ``input_node.concat_index`` (set by the caller) or is looked up
    at run time by shape when used standalone."""
    return ConcatGradientOp(grad_node, input_node, axis, 0, ctx=ctx)


def split_op(node, axes, indices, splits, ctx=None):
    return SplitOp(node, axes, indices, splits, ctx=ctx)


def split_gradient_op(node, axes, indices, splits, ctx=None):
    return SplitGradientOp(node, None, axes, indices, splits, ctx=ctx)


def slice_op(node, begin, size, ctx=None):
    return SliceOp(node, begin, size, ctx=ctx)


def slice_gradient_op(node, begin, size=None, ctx=None):
    return SliceGradientOp(node, begin, size, ctx=ctx)


def slice_assign_op(node, begin, size, val, ctx=None):
    return SliceAssignOp(node, begin, size, val, ctx=ctx)


def slice_assign_matrix_op(node_A, node_B, begin_A, size_A, begin_B, size_B, ctx=None):
    return SliceAssignMatrixOp(node_A, node_B, begin_A, size_A, begin_B, size_B, ctx=ctx)


def slice_by_matrix_op(node_A, index1, index2, ctx=None):
    return SliceByMatrixOp(node_A, index1, index2, ctx=ctx)


def slice_by_matrix_gradient_op(input, grad, index1, index2, ctx=None):
    return SliceByMatrixGradientOp(input, grad, index1, index2, ctx=ctx)


def gather_op(node, dim, index, ctx=None):
    return GatherOp(node, dim, index, ctx=ctx)


def gather_gradient_op(input, grad, dim, index, ctx=None):
    return GatherGradientOp(input, grad, dim, index, ctx=ctx)


def scatter_op(node1, node2, node3, ctx=None):
    return ScatterOp(node1, node2, node3, ctx=ctx)


def scatter1d_op(input_mat, index_mat, ctx=None):
    return Scatter1DOp(input_mat, index_mat, ctx=ctx)


def scatter1d_grad_op(output_grad_mat, index_mat, ctx=None):
    return Scatter1DGradOp(output_grad_mat, index_mat, ctx=ctx)


def indexing_op(input_mat, index_mat, ctx=None):
    return IndexingOp(input_mat, index_mat, ctx=ctx)


def indexing_grad_op(output_grad, index, ctx=None):
    return IndexingGradOp(output_grad, index, ctx=ctx)


def repeat_op(node, reps, ctx=None):
    return RepeatOp(node, reps, ctx=ctx)


def repeat_gradient_op(node_input, node_grad, ctx=None):
    return RepeatGradientOp(node_input, node_grad, ctx=ctx)


def roll_op(node, shift=0, axis=None, ctx=None):
    return RollOp(node, shift, axis, ctx=ctx)


def one_hot_op(node, num_classes, ctx=None):
    return OneHotOp(node, num_classes, ctx=ctx)


def pad_op(node_A, paddings, mode="CONSTANT", constant_values=0, ctx=None):
    return PadOp(node_A, paddings, mode, constant_values, ctx=ctx)


def pad_gradient_op(node_A, paddings, mode="CONSTANT", ctx=None):
    return PadGradientOp(node_A, paddings, mode, ctx=ctx)


def interpolate_op(input, size=None, scale_factor=None, mode='bicubic', align_corners=False,
                   ctx=None):
    return InterpolateOp(input, size, scale_factor, mode, align_corners, ctx=ctx)


def interpolate_grad_op(grad, input, mode='bicubic', align_corners=False, ctx=None):
    return InterpolateGradOp(grad, input, mode, align_corners, ctx=ctx)
