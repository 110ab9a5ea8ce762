"""Loss ops.

reference op files: SoftmaxCrossEntropy SoftmaxCrossEntropySparse CrossEntropy
CrossEntropySparse BinaryCrossEntropy NllLoss (python/hetu/gpu_ops/*.py;
kernels src/ops/SoftmaxCrossEntropy.cu:5-30 - one thread per row, serial over
classes - and CudnnSoftmaxEntropy.cu).

B200 path: ``hetu_softmax_xent`` is a one-pass online-softmax row kernel (one
warp per row up to 1024 classes, one CTA per row beyond) that emits the loss;
the backward kernel emits ``(softmax - y) * g`` without materialising softmax.
"""
from __future__ import annotations

import torch

from .. import _lib as L
from ..graph.node import Op, as_index


def _k(*ts):
    return L.kernels_on(*ts) and all(t.dtype == torch.float32 for t in ts)


from .elementwise import _cpu_native  # noqa: E402


class SoftmaxCrossEntropyOp(Op):
    """loss[b] = -sum_c y_[b,c] * log softmax(x)[b,c]"""

    def __init__(self, node_A, node_B, use_cudnn=True, ctx=None):
        super().__init__(SoftmaxCrossEntropyOp, [node_A, node_B], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x, y_ = input_vals
        if x.device.type == 'meta':
            return torch.empty(tuple(x.shape[:-1]), device='meta')
        if _k(x, y_):
            x, y_ = x.contiguous(), y_.contiguous()
            cols = x.shape[-1]
            rows = x.numel() // cols
            loss = torch.empty(x.shape[:-1], dtype=x.dtype, device=x.device)
            L.call('softmax_xent', L.p(x), L.p(y_), L.p(loss), L.c_int(rows), L.c_int(cols),
                   L.stream_ptr(x.device))
            return loss
        if x.device.type == 'cpu' and x.dtype == torch.float32 and y_.dtype == torch.float32 \
                and x.shape == y_.shape and _cpu_native():
            from .. import cpu_links
            loss = torch.empty(x.shape[:-1], dtype=x.dtype)
            cpu_links.softmax_crossentropy(x.contiguous(), y_.contiguous(), loss)
            return loss
        return -(y_ * torch.log_softmax(x, dim=-1)).sum(dim=-1)

    def gradient(self, g):
        return [SoftmaxCrossEntropyGradientOp(self.inputs[0], self.inputs[1], g, ctx=self.raw_ctx),
                None]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0][:-1])


class SoftmaxCrossEntropyGradientOp(Op):
    def __init__(self, node_A, node_B, node_C, use_cudnn=True, ctx=None):
        super().__init__(SoftmaxCrossEntropyGradientOp, [node_A, node_B, node_C], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x, y_, g = input_vals
        if x.device.type == 'meta':
            return torch.empty_like(x)
        if _k(x, y_, g):
            x, y_, g = x.contiguous(), y_.contiguous(), g.contiguous()
            cols = x.shape[-1]
            dx = torch.empty_like(x)
            L.call('softmax_xent_grad', L.p(x), L.p(y_), L.p(g), L.p(dx), L.c_int(x.numel() // cols),
                   L.c_int(cols), L.stream_ptr(x.device))
            return dx
        return (torch.softmax(x, dim=-1) * y_.sum(dim=-1, keepdim=True) - y_) * g.unsqueeze(-1)

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


class SoftmaxCrossEntropySparseOp(Op):
    """Integer labels; rows whose label == ignored_index contribute 0."""

    def __init__(self, node_A, node_B, ignored_index=-1, ctx=None):
        super().__init__(SoftmaxCrossEntropySparseOp, [node_A, node_B], ctx)
        self.ignored_index = ignored_index

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x, lab = input_vals
        if x.device.type == 'meta':
            return torch.empty(tuple(x.shape[:-1]), device='meta')
        if _k(x) and lab.is_cuda and lab.dtype == torch.float32:
            x, lab = x.contiguous(), lab.contiguous()
            cols = x.shape[-1]
            loss = torch.empty(x.shape[:-1], dtype=x.dtype, device=x.device)
            L.call('softmax_xent_sparse', L.p(x), L.p(lab), L.p(loss), L.c_int(x.numel() // cols),
                   L.c_int(cols), L.c_int(self.ignored_index), L.stream_ptr(x.device))
            return loss
        idx = as_index(lab)
        valid = (idx != self.ignored_index) & (idx >= 0)
        lsm = torch.log_softmax(x, dim=-1)
        picked = torch.gather(lsm, -1, idx.clamp_min(0).unsqueeze(-1)).squeeze(-1)
        return torch.where(valid, -picked, torch.zeros_like(picked))

    def gradient(self, g):
        return [SoftmaxCrossEntropySparseGradientOp(self.inputs[0], self.inputs[1], g,
                                                    self.ignored_index, ctx=self.raw_ctx), None]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0][:-1])


class SoftmaxCrossEntropySparseGradientOp(Op):
    def __init__(self, node_A, node_B, node_C, ignored_index=-1, ctx=None):
        super().__init__(SoftmaxCrossEntropySparseGradientOp, [node_A, node_B, node_C], ctx)
        self.ignored_index = ignored_index

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x, lab, g = input_vals
        if x.device.type == 'meta':
            return torch.empty_like(x)
        if _k(x, g) and lab.is_cuda and lab.dtype == torch.float32:
            x, lab, g = x.contiguous(), lab.contiguous(), g.contiguous()
            cols = x.shape[-1]
            dx = torch.empty_like(x)
            L.call('softmax_xent_sparse_grad', L.p(x), L.p(lab), L.p(g), L.p(dx),
                   L.c_int(x.numel() // cols), L.c_int(cols), L.c_int(self.ignored_index),
                   L.stream_ptr(x.device))
            return dx
        idx = as_index(lab)
        valid = ((idx != self.ignored_index) & (idx >= 0)).to(x.dtype).unsqueeze(-1)
        onehot = torch.zeros_like(x).scatter_(-1, idx.clamp_min(0).unsqueeze(-1), 1.0)
        return (torch.softmax(x, dim=-1) - onehot) * valid * g.unsqueeze(-1)

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


class CrossEntropyOp(Op):
    """loss[b] = -sum_c y_[b,c] * log(y[b,c]) on probabilities."""

    def __init__(self, node_y, node_y_, ctx=None):
        super().__init__(CrossEntropyOp, [node_y, node_y_], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        y, y_ = input_vals
        if y.device.type == 'meta':
            return torch.empty(tuple(y.shape[:-1]), device='meta')
        return -(y_ * torch.log(y)).sum(dim=-1)

    def gradient(self, g):
        return [CrossEntropyGradientOp(g, self.inputs[0], self.inputs[1], ctx=self.raw_ctx), None]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0][:-1])


class CrossEntropyGradientOp(Op):
    def __init__(self, grad, node_y, node_y_, ctx=None):
        super().__init__(CrossEntropyGradientOp, [grad, node_y, node_y_], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        g, y, y_ = input_vals
        if y.device.type == 'meta':
            return torch.empty_like(y)
        return -(y_ / y) * g.unsqueeze(-1)

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[1])


class CrossEntropySparseOp(Op):
    def __init__(self, node_y, node_y_, ignored_index=-1, ctx=None):
        super().__init__(CrossEntropySparseOp, [node_y, node_y_], ctx)
        self.ignored_index = ignored_index

    def compute(self, input_vals, output_val=None, stream_handle=None):
        y, lab = input_vals
        if y.device.type == 'meta':
            return torch.empty(tuple(y.shape[:-1]), device='meta')
        idx = as_index(lab)
        valid = (idx != self.ignored_index) & (idx >= 0)
        picked = torch.gather(y, -1, idx.clamp_min(0).unsqueeze(-1)).squeeze(-1)
        return torch.where(valid, -torch.log(picked), torch.zeros_like(picked))

    def gradient(self, g):
        return [CrossEntropySparseGradientOp(g, self.inputs[0], self.inputs[1], self.ignored_index,
                                             ctx=self.raw_ctx), None]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0][:-1])


class CrossEntropySparseGradientOp(Op):
    def __init__(self, grad, node_y, node_y_, ignored_index=-1, ctx=None):
        super().__init__(CrossEntropySparseGradientOp, [grad, node_y, node_y_], ctx)
        self.ignored_index = ignored_index

    def compute(self, input_vals, output_val=None, stream_handle=None):
        g, y, lab = input_vals
        if y.device.type == 'meta':
            return torch.empty_like(y)
        idx = as_index(lab)
        valid = ((idx != self.ignored_index) & (idx >= 0)).to(y.dtype)
        picked = torch.gather(y, -1, idx.clamp_min(0).unsqueeze(-1))
        val = (-(g * valid).unsqueeze(-1) / picked)
        return torch.zeros_like(y).scatter_(-1, idx.clamp_min(0).unsqueeze(-1), val)

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[1])


class BinaryCrossEntropyOp(Op):
    """loss = -y_ log(p) - (1 - y_) log(1 - p), element-wise."""

    def __init__(self, node_A, node_B, ctx=None):
        super().__init__(BinaryCrossEntropyOp, [node_A, node_B], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        p, y_ = input_vals
        if p.device.type == 'meta':
            return torch.empty_like(p)
        if _k(p, y_) and p.shape == y_.shape:
            p, y_ = p.contiguous(), y_.contiguous()
            out = torch.empty_like(p)
            L.call('bce', L.p(p), L.p(y_), L.p(out), L.c_sz(p.numel()), L.stream_ptr(p.device))
            return out
        return -y_ * torch.log(p) - (1 - y_) * torch.log(1 - p)

    def gradient(self, g):
        return [BinaryCrossEntropyGradientOp(self.inputs[0], self.inputs[1], g, ctx=self.raw_ctx), None]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


class BinaryCrossEntropyGradientOp(Op):
    def __init__(self, node_A, node_B, node_C, ctx=None):
        super().__init__(BinaryCrossEntropyGradientOp, [node_A, node_B, node_C], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        p, y_, g = input_vals
        if p.device.type == 'meta':
            return torch.empty_like(p)
        if _k(p, y_, g) and p.shape == y_.shape == g.shape:
            p, y_, g = p.contiguous(), y_.contiguous(), g.contiguous()
            out = torch.empty_like(p)
            L.call('bce_grad', L.p(p), L.p(y_), L.p(g), L.p(out), L.c_sz(p.numel()),
                   L.stream_ptr(p.device))
            return out
        return g * (-y_ / p + (1 - y_) / (1 - p))

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


class NllLossOp(Op):
    """-mean_b input[b, target[b]] -> shape (1,)   (reference NllLoss.cu)."""

    def __init__(self, input, target, cols, ctx=None):
        super().__init__(NllLossOp, [input, target], ctx)
        self.cols = cols

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x, t = input_vals
        if x.device.type == 'meta':
            return torch.empty((1,), device='meta')
        idx = as_index(t).reshape(-1)
        picked = torch.gather(x, 1, idx.unsqueeze(1)).squeeze(1)
        return (-picked.sum() / x.shape[0]).reshape(1)

    def gradient(self, g):
        return [nll_loss_grad_op(g, self.inputs[1], self.cols, ctx=self.raw_ctx), None]

    def infer_shape(self, input_shapes):
        return (1,)


class NllLossGradOp(Op):
    def __init__(self, output_grad, target, cols, ctx=None):
        super().__init__(NllLossGradOp, [output_grad, target], ctx)
        self.cols = cols

    def compute(self, input_vals, output_val=None, stream_handle=None):
        g, t = input_vals
        rows = t.shape[0]
        if g.device.type == 'meta':
            return torch.empty((rows, self.cols), device='meta')
        idx = as_index(t).reshape(-1)
        out = torch.zeros((rows, self.cols), dtype=g.dtype, device=g.device)
        val = (-g.reshape(-1)[0] / rows).expand(rows, 1)
        return out.scatter_(1, idx.unsqueeze(1), val)

    def infer_shape(self, input_shapes):
        return (input_shapes[1][0], self.cols)


def softmaxcrossentropy_op(node_A, node_B, use_cudnn=True, ctx=None):
    return SoftmaxCrossEntropyOp(node_A, node_B, use_cudnn, ctx=ctx)


def softmaxcrossentropy_sparse_op(node_A, node_B, ignored_index=-1, ctx=None):
    return SoftmaxCrossEntropySparseOp(node_A, node_B, ignored_index, ctx=ctx)


def crossentropy_op(node_y, node_y_, ctx=None):
    return CrossEntropyOp(node_y, node_y_, ctx=ctx)


def crossentropy_sparse_op(node_y, node_y_, ignored_index=-1, ctx=None):
    return CrossEntropySparseOp(node_y, node_y_, ignored_index, ctx=ctx)


def binarycrossentropy_op(node_A, node_B, ctx=None):
    return BinaryCrossEntropyOp(node_A, node_B, ctx=ctx)


def nll_loss_op(input, target, cols, ctx=None):
    return NllLossOp(input, target, cols, ctx=ctx)


def nll_loss_grad_op(output_grad, target, cols, ctx=None):
    return NllLossGradOp(output_grad, target, cols, ctx=ctx)
