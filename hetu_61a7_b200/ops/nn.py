"""Neural-network ops: convolution, pooling, normalisation, dropout, softmax.

reference op files: Conv2d Conv2dAddBias Conv2dBroadcast Conv2dReduceSum
MaxPool AvgPool BatchNorm LayerNorm InstanceNorm2d Dropout Dropout2d Softmax
(python/hetu/gpu_ops/*.py; the reference goes to cuDNN for conv/pool/BN/softmax
- src/ops/Cudnn*.cu - and has a 5-launch LayerNorm backward with three
full-size workspaces, src/ops/LayerNorm.cu:64-255).

B200 path: convolution = im2col/col2im kernels around the tcgen05 GEMM;
softmax / LayerNorm forward are one-pass row kernels (warp-shuffle Welford);
LayerNorm backward is a single pass for dX plus a two-stage column reduction for
dgamma/dbeta; dropout draws Philox bits in-kernel from (op id, step counter) so
the mask is recomputed - never stored - in backward and is CUDA-graph safe.
"""
from __future__ import annotations

import os

import math

import numpy as np
import torch
import torch.nn.functional as F

from .. import _lib as L
from ..graph.node import Op
from .linalg import gemm
from .shape import materialize, strided_copy


def _k(*ts):
    return L.kernels_on(*ts) and all(t.dtype == torch.float32 for t in ts)


def _pair(v):
    return (v, v) if isinstance(v, int) else tuple(v)


# ------------------------------------------------------------------ conv2d
def _conv_out(h, k, p, s):
    return (h + 2 * p - k) // s + 1


def im2col(x: torch.Tensor, kh, kw, ph, pw, sh, sw) -> torch.Tensor:
    """[N,C,H,W] -> [N*OH*OW, C*kh*kw] (K-major rows for the GEMM A operand)."""
    n, c, h, w = x.shape
    oh, ow = _conv_out(h, kh, ph, sh), _conv_out(w, kw, pw, sw)
    if _k(x):
        x = x.contiguous()
        col = torch.empty((n * oh * ow, c * kh * kw), dtype=x.dtype, device=x.device)
        L.call('im2col', L.p(x), L.p(col), L.c_int(n), L.c_int(c), L.c_int(h), L.c_int(w),
               L.c_int(kh), L.c_int(kw), L.c_int(ph), L.c_int(pw), L.c_int(sh), L.c_int(sw),
               L.c_int(oh), L.c_int(ow), L.stream_ptr(x.device))
        return col
    col = F.unfold(x, (kh, kw), padding=(ph, pw), stride=(sh, sw))      # [N, C*kh*kw, OH*OW]
    return col.transpose(1, 2).reshape(n * oh * ow, c * kh * kw)


def col2im(col: torch.Tensor, n, c, h, w, kh, kw, ph, pw, sh, sw) -> torch.Tensor:
    oh, ow = _conv_out(h, kh, ph, sh), _conv_out(w, kw, pw, sw)
    if _k(col):
        col = col.contiguous()
        x = torch.empty((n, c, h, w), dtype=col.dtype, device=col.device)
        L.call('col2im', L.p(col), L.p(x), L.c_int(n), L.c_int(c), L.c_int(h), L.c_int(w),
               L.c_int(kh), L.c_int(kw), L.c_int(ph), L.c_int(pw), L.c_int(sh), L.c_int(sw),
               L.c_int(oh), L.c_int(ow), L.stream_ptr(col.device))
        return x
    col = col.reshape(n, oh * ow, c * kh * kw).transpose(1, 2)
    return F.fold(col, (h, w), (kh, kw), padding=(ph, pw), stride=(sh, sw))


# ---- per-step caches: the backward pass of a convolution needs im2col(x) and the NHWC copy of dY again (weight and data
# gradient); recomputing them was a quarter of the ResNet-18 step.  Entries are valid for the tensor object they were
# built from; the executor clears them at the start and end of every pass over the graph (reset_conv_cache).
_CONV_CACHE = {'col': {}, 'dy': None, 'bytes': 0}


def reset_conv_cache():
    _CONV_CACHE['col'].clear()
    _CONV_CACHE['dy'] = None
    _CONV_CACHE['bytes'] = 0


def _cached_im2col(x, kh, kw, ph, pw, sh, sw, store):
    key = (id(x), kh, kw, ph, pw, sh, sw)
    hit = _CONV_CACHE['col'].get(key)
    if hit is not None and hit[0] is x and hit[1] == x._version:
        if not store:                                    # the weight gradient is the last user: let the buffer go
            del _CONV_CACHE['col'][key]
            _CONV_CACHE['bytes'] -= hit[2].numel() * hit[2].element_size()
        return hit[2]
    col = im2col(x, kh, kw, ph, pw, sh, sw)
    limit = float(os.environ.get('HETU_CONV_COL_CACHE_MB', '8192')) * (1 << 20)
    nbytes = col.numel() * col.element_size()
    if store and _CONV_CACHE['bytes'] + nbytes <= limit:
        _CONV_CACHE['col'][key] = (x, x._version, col)
        _CONV_CACHE['bytes'] += nbytes
    return col


def _nhwc_rows(dy, oc):
    hit = _CONV_CACHE['dy']
    if hit is not None and hit[0] is dy and hit[1] == dy._version:
        return hit[2]
    dy2 = materialize(dy.permute(0, 2, 3, 1)).reshape(-1, oc)       # [N*OH*OW, OC]
    _CONV_CACHE['dy'] = (dy, dy._version, dy2)
    return dy2


def conv2d_forward(x, w, bias, padding, stride, act=0):
    ph, pw = _pair(padding)
    sh, sw = _pair(stride)
    if x.device.type != 'cuda':
        return F.conv2d(x, w, bias, stride=(sh, sw), padding=(ph, pw))
    n, c, h, wd = x.shape
    oc, _, kh, kw = w.shape
    oh, ow = _conv_out(h, kh, ph, sh), _conv_out(wd, kw, pw, sw)
    col = _cached_im2col(x, kh, kw, ph, pw, sh, sw, store=kh * kw > 1)   # [N*OH*OW, C*kh*kw]; kept for the weight gradient
    y = gemm(col, w.reshape(oc, -1), False, True, bias=bias, act=act)   # [N*OH*OW, OC]
    return materialize(y.reshape(n, oh, ow, oc).permute(0, 3, 1, 2))


def conv2d_grad_data(dy, w, x_shape, padding, stride):
    ph, pw = _pair(padding)
    sh, sw = _pair(stride)
    if dy.device.type != 'cuda':
        return torch.nn.grad.conv2d_input(x_shape, w, dy, stride=(sh, sw), padding=(ph, pw))
    n, c, h, wd = x_shape
    oc, _, kh, kw = w.shape
    oh, ow = dy.shape[2], dy.shape[3]
    if _k(dy, w) and oh * ow >= 256 and (oh * ow) % 4 == 0 and oc % 4 == 0 and kh * kw > 1 and sh == sw and sh <= 2 \
            and n * c <= 65535 and os.environ.get('HETU_CONV_DGRAD_BATCHED', '1') == '1':
        # dY stays NCHW: per image colT_n [C*kh*kw, OH*OW] = W^T [C*kh*kw, OC] @ dY_n [OC, OH*OW] (one batched GEMM), and the
        # gather that folds it back reads rows along ow - no NHWC copy of dY, coalesced col2im
        w3 = w.reshape(1, oc, c * kh * kw).expand(n, oc, c * kh * kw).contiguous()
        colT = gemm(w3, dy.contiguous().view(n, oc, oh * ow), True, False)          # [N, C*kh*kw, OH*OW]
        x = torch.empty((n, c, h, wd), dtype=dy.dtype, device=dy.device)
        L.call('col2im_t', L.p(colT), L.p(x), L.c_int(n), L.c_int(c), L.c_int(h), L.c_int(wd), L.c_int(kh), L.c_int(kw),
               L.c_int(ph), L.c_int(pw), L.c_int(sh), L.c_int(sw), L.c_int(oh), L.c_int(ow), L.stream_ptr(dy.device))
        return x
    dy2 = _nhwc_rows(dy, oc)                                        # [N*OH*OW, OC]
    col = gemm(dy2, w.reshape(oc, -1), False, False)               # [N*OH*OW, C*kh*kw]
    return col2im(col, n, c, h, wd, kh, kw, ph, pw, sh, sw)


def conv2d_grad_filter(x, dy, w_shape, padding, stride):
    ph, pw = _pair(padding)
    sh, sw = _pair(stride)
    if x.device.type != 'cuda':
        return torch.nn.grad.conv2d_weight(x, w_shape, dy, stride=(sh, sw), padding=(ph, pw))
    oc, c, kh, kw = w_shape
    col = _cached_im2col(x, kh, kw, ph, pw, sh, sw, store=False)    # [N*OH*OW, C*kh*kw]: the forward pass's, when it is still there
    dy2 = _nhwc_rows(dy, oc)                                        # [N*OH*OW, OC]
    return gemm(dy2, col, True, False).reshape(w_shape)             # [OC, C*kh*kw]


class Conv2dOp(Op):
    def __init__(self, node_A, node_B, padding=0, stride=1, ctx=None):
        super().__init__(Conv2dOp, [node_A, node_B], ctx)
        self.padding, self.stride = padding, stride

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x, w = input_vals
        if x.device.type == 'meta':
            return F.conv2d(x, w, None, stride=_pair(self.stride), padding=_pair(self.padding))
        return conv2d_forward(x, w, None, self.padding, self.stride)

    def gradient(self, g):
        x, w = self.inputs
        return [conv2d_gradient_of_data_op(w, g, x, self.padding, self.stride, ctx=self.raw_ctx),
                conv2d_gradient_of_filter_op(x, g, w, self.padding, self.stride, ctx=self.raw_ctx)]


class Conv2dGradientOfDataOp(Op):
    """inputs: filter, dY, X(for its shape)."""

    def __init__(self, node_A, node_B, node_C, padding=0, stride=1, ctx=None):
        super().__init__(Conv2dGradientOfDataOp, [node_A, node_B, node_C], ctx)
        self.padding, self.stride = padding, stride

    def compute(self, input_vals, output_val=None, stream_handle=None):
        w, dy, x = input_vals
        if dy.device.type == 'meta':
            return torch.empty_like(x)
        return conv2d_grad_data(dy, w, tuple(x.shape), self.padding, self.stride)

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[2])


class Conv2dGradientOfFilterOp(Op):
    """inputs: X, dY, filter(for its shape)."""

    def __init__(self, input_X, gradient_Y, input_filter, padding=0, stride=1, ctx=None):
        super().__init__(Conv2dGradientOfFilterOp, [input_X, gradient_Y, input_filter], ctx)
        self.padding, self.stride = padding, stride

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x, dy, w = input_vals
        if dy.device.type == 'meta':
            return torch.empty_like(w)
        return conv2d_grad_filter(x, dy, tuple(w.shape), self.padding, self.stride)

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[2])


class Conv2dAddBiasOp(Op):
    def __init__(self, node_A, node_B, bias, padding=0, stride=1, ctx=None):
        super().__init__(Conv2dAddBiasOp, [node_A, node_B, bias], ctx)
        self.padding, self.stride = padding, stride

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x, w, b = input_vals
        if x.device.type == 'meta':
            return F.conv2d(x, w, None, stride=_pair(self.stride), padding=_pair(self.padding))
        return conv2d_forward(x, w, b.reshape(-1), self.padding, self.stride)

    def gradient(self, g):
        x, w, b = self.inputs
        return [conv2d_gradient_of_data_op(w, g, x, self.padding, self.stride, ctx=self.raw_ctx),
                conv2d_gradient_of_filter_op(x, g, w, self.padding, self.stride, ctx=self.raw_ctx),
                conv2d_reducesum_op(g, ctx=self.raw_ctx)]


class Conv2dBroadcastToOp(Op):
    """bias [C] -> [N,C,H,W] of node_B."""

    def __init__(self, node_A, node_B, ctx=None):
        super().__init__(Conv2dBroadcastToOp, [node_A, node_B], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        b, like = input_vals
        return materialize(b.reshape(1, -1, 1, 1).expand(tuple(like.shape)))

    def gradient(self, g):
        return [conv2d_reducesum_op(g, ctx=self.raw_ctx), None]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[1])


class Conv2dReduceSumOp(Op):
    """[N,C,H,W] -> [C]"""

    def __init__(self, node, ctx=None):
        super().__init__(Conv2dReduceSumOp, [node], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        from .reduce import _reduce, R_SUM
        x = input_vals[0]
        if x.device.type == 'meta':
            return torch.empty((x.shape[1],), device='meta')
        y = _reduce(x, [2, 3], False, R_SUM)        # [N, C]
        return _reduce(y, [0], False, R_SUM)

    def gradient(self, g):
        return [conv2d_broadcastto_op(g, self.inputs[0], ctx=self.raw_ctx)]


# ------------------------------------------------------------------ pooling
class _Pool2dOp(Op):
    is_max = True

    def __init__(self, node_A, kernel_H, kernel_W, padding, stride, ctx=None):
        super().__init__(type(self), [node_A], ctx)
        self.kernel_H, self.kernel_W, self.padding, self.stride = kernel_H, kernel_W, padding, stride

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        if _k(x):
            x = x.contiguous()
            n, c, h, w = x.shape
            oh = _conv_out(h, self.kernel_H, self.padding, self.stride)
            ow = _conv_out(w, self.kernel_W, self.padding, self.stride)
            y = torch.empty((n, c, oh, ow), dtype=x.dtype, device=x.device)
            L.call('pool2d_fwd', L.p(x), L.p(y), L.c_int(n * c), L.c_int(h), L.c_int(w),
                   L.c_int(self.kernel_H), L.c_int(self.kernel_W), L.c_int(self.padding),
                   L.c_int(self.stride), L.c_int(oh), L.c_int(ow), L.c_int(int(self.is_max)),
                   L.stream_ptr(x.device))
            return y
        if self.is_max:
            return F.max_pool2d(x, (self.kernel_H, self.kernel_W), self.stride, self.padding)
        return F.avg_pool2d(x, (self.kernel_H, self.kernel_W), self.stride, self.padding,
                            count_include_pad=True)

    def gradient(self, g):
        cls = MaxPool2dGradientOp if self.is_max else AvgPool2dGradientOp
        return [cls(self, g, self.inputs[0], self.kernel_H, self.kernel_W, self.padding,
                    self.stride, ctx=self.raw_ctx)]


class MaxPool2dOp(_Pool2dOp):
    is_max = True


class AvgPool2dOp(_Pool2dOp):
    is_max = False


class _Pool2dGradientOp(Op):
    is_max = True

    def __init__(self, node_out, node_out_gradient, node_in, kernel_H, kernel_W, padding, stride,
                 ctx=None):
        super().__init__(type(self), [node_out, node_out_gradient, node_in], ctx)
        self.kernel_H, self.kernel_W, self.padding, self.stride = kernel_H, kernel_W, padding, stride

    def compute(self, input_vals, output_val=None, stream_handle=None):
        y, dy, x = input_vals
        if x.device.type == 'meta':
            return torch.empty_like(x)
        if _k(x, dy, y):
            x, dy, y = x.contiguous(), dy.contiguous(), y.contiguous()
            n, c, h, w = x.shape
            dx = torch.empty_like(x)
            L.call('pool2d_bwd', L.p(x), L.p(y), L.p(dy), L.p(dx), L.c_int(n * c), L.c_int(h),
                   L.c_int(w), L.c_int(self.kernel_H), L.c_int(self.kernel_W), L.c_int(self.padding),
                   L.c_int(self.stride), L.c_int(y.shape[2]), L.c_int(y.shape[3]),
                   L.c_int(int(self.is_max)), L.stream_ptr(x.device))
            return dx
        with torch.enable_grad():
            xx = x.detach().requires_grad_(True)
            if self.is_max:
                yy = F.max_pool2d(xx, (self.kernel_H, self.kernel_W), self.stride, self.padding)
            else:
                yy = F.avg_pool2d(xx, (self.kernel_H, self.kernel_W), self.stride, self.padding,
                                  count_include_pad=True)
            (dx,) = torch.autograd.grad(yy, xx, dy)
        return dx

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[2])


class MaxPool2dGradientOp(_Pool2dGradientOp):
    is_max = True


class AvgPool2dGradientOp(_Pool2dGradientOp):
    is_max = False


# ------------------------------------------------------------- batch norm
class BatchNormOp(Op):
    """Spatial batch norm over NCHW (or NC).  Running statistics live on the op
    like the reference (BatchNorm.py:24-27) but ARE exposed to save/load here
    through ``state_tensors`` (the reference loses them at checkpoint time)."""
    takes_inference = True

    def __init__(self, node_in, bn_scale, bn_bias, momentum=0.1, eps=1e-5, ctx=None):
        super().__init__(BatchNormOp, [node_in, bn_scale, bn_bias], ctx)
        self.momentum, self.eps = momentum, eps
        self.running_mean = self.running_var = None
        self.save_mean = self.save_rstd = None

    def state_tensors(self):
        return {'running_mean': self.running_mean, 'running_var': self.running_var}

    def compute(self, input_vals, output_val=None, stream_handle=None, inference=False):
        x, gamma, beta = input_vals
        if x.device.type == 'meta':
            return torch.empty_like(x)
        c = x.shape[1]
        gamma, beta = gamma.reshape(-1), beta.reshape(-1)
        if self.running_mean is None or self.running_mean.device != x.device:
            self.running_mean = torch.zeros(c, dtype=x.dtype, device=x.device)
            self.running_var = torch.ones(c, dtype=x.dtype, device=x.device)
            self.save_mean = torch.zeros(c, dtype=x.dtype, device=x.device)
            self.save_rstd = torch.ones(c, dtype=x.dtype, device=x.device)
        if _k(x):
            x = x.contiguous()
            y = torch.empty_like(x)
            hw = x.numel() // (x.shape[0] * c)
            L.call('batchnorm_fwd', L.p(x), L.p(gamma), L.p(beta), L.p(y), L.p(self.running_mean),
                   L.p(self.running_var), L.p(self.save_mean), L.p(self.save_rstd),
                   L.c_int(x.shape[0]), L.c_int(c), L.c_int(hw), L.c_f32(self.momentum),
                   L.c_f32(self.eps), L.c_int(int(inference)), L.stream_ptr(x.device))
            return y
        if inference:
            return F.batch_norm(x, self.running_mean, self.running_var, gamma, beta, False, 0.0, self.eps)
        dims = [0] + list(range(2, x.dim()))
        mean = x.mean(dim=dims)
        var = x.var(dim=dims, unbiased=False)
        nel = x.numel() // c
        self.running_mean.mul_(1 - self.momentum).add_(mean * self.momentum)
        self.running_var.mul_(1 - self.momentum).add_(var * (nel / max(nel - 1, 1)) * self.momentum)
        self.save_mean.copy_(mean)
        self.save_rstd.copy_((var + self.eps).rsqrt())
        shp = [1, c] + [1] * (x.dim() - 2)
        return (x - mean.reshape(shp)) * self.save_rstd.reshape(shp) * gamma.reshape(shp) + beta.reshape(shp)

    def gradient(self, g):
        bn_grad = batch_normalization_gradient_op(g, self.inputs[0], self.inputs[1], self, self.eps,
                                                  ctx=self.raw_ctx)
        return [batch_normalization_gradient_of_data_op(bn_grad, self.inputs[0], ctx=self.raw_ctx),
                batch_normalization_gradient_of_scale_op(bn_grad, self.inputs[1], ctx=self.raw_ctx),
                batch_normalization_gradient_of_bias_op(bn_grad, self.inputs[2], ctx=self.raw_ctx)]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


class BatchNormGradientOp(Op):
    """Computes (dX, dgamma, dbeta) once; the three ``gradient_of_*`` nodes pick."""

    def __init__(self, out_gradient, in_node, bn_scale, forward_node, eps, ctx=None):
        super().__init__(BatchNormGradientOp, [out_gradient, in_node, bn_scale], ctx)
        self.forward_node, self.eps = forward_node, eps
        self.results = None

    def compute(self, input_vals, output_val=None, stream_handle=None):
        dy, x, gamma = input_vals
        if x.device.type == 'meta':
            return torch.empty_like(x)
        fwd = self.forward_node
        c = x.shape[1]
        gamma = gamma.reshape(-1)
        if _k(x, dy):
            x, dy = x.contiguous(), dy.contiguous()
            dx = torch.empty_like(x)
            dgamma = torch.empty(c, dtype=x.dtype, device=x.device)
            dbeta = torch.empty(c, dtype=x.dtype, device=x.device)
            hw = x.numel() // (x.shape[0] * c)
            L.call('batchnorm_bwd', L.p(dy), L.p(x), L.p(gamma), L.p(fwd.save_mean), L.p(fwd.save_rstd),
                   L.p(dx), L.p(dgamma), L.p(dbeta), L.c_int(x.shape[0]), L.c_int(c), L.c_int(hw),
                   L.stream_ptr(x.device))
        else:
            dims = [0] + list(range(2, x.dim()))
            shp = [1, c] + [1] * (x.dim() - 2)
            xhat = (x - fwd.save_mean.reshape(shp)) * fwd.save_rstd.reshape(shp)
            dbeta = dy.sum(dim=dims)
            dgamma = (dy * xhat).sum(dim=dims)
            nel = x.numel() // c
            dx = (gamma * fwd.save_rstd).reshape(shp) / nel * (
                nel * dy - dbeta.reshape(shp) - xhat * dgamma.reshape(shp))
        self.results = (dx, dgamma, dbeta)
        return dx

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[1])


class _BNPick(Op):
    which = 0

    def __init__(self, bn_gradient, like, ctx=None):
        super().__init__(type(self), [bn_gradient, like], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        like = input_vals[1]
        if like.device.type == 'meta':
            return torch.empty_like(like)
        return self.inputs[0].results[self.which].reshape(tuple(like.shape))

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[1])


class BatchNormGradientOfDataOp(_BNPick):
    which = 0


class BatchNormGradientOfScaleOp(_BNPick):
    which = 1


class BatchNormGradientOfBiasOp(_BNPick):
    which = 2


# ------------------------------------------------------------- layer norm
class LayerNormOp(Op):
    def __init__(self, node_in, ln_scale, ln_bias, eps=0.01, ctx=None):
        super().__init__(LayerNormOp, [node_in, ln_scale, ln_bias], ctx)
        self.eps = eps
        self.save_mean = self.save_rstd = None

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x, gamma, beta = input_vals
        if x.device.type == 'meta':
            return torch.empty_like(x)
        cols = x.shape[-1]
        rows = x.numel() // cols
        if _k(x, gamma, beta):
            x = x.contiguous()
            y = torch.empty_like(x)
            self.save_mean = torch.empty(rows, dtype=x.dtype, device=x.device)
            self.save_rstd = torch.empty(rows, dtype=x.dtype, device=x.device)
            L.call('layernorm_fwd', L.p(x), L.p(gamma.contiguous()), L.p(beta.contiguous()), L.p(y),
                   L.p(self.save_mean), L.p(self.save_rstd), L.c_int(rows), L.c_int(cols),
                   L.c_f32(self.eps), L.stream_ptr(x.device))
            return y
        mean = x.mean(dim=-1, keepdim=True)
        var = x.var(dim=-1, unbiased=False, keepdim=True)
        rstd = (var + self.eps).rsqrt()
        self.save_mean, self.save_rstd = mean.reshape(-1), rstd.reshape(-1)
        return (x - mean) * rstd * gamma + beta

    def gradient(self, g):
        ln_grad = LayerNormGradientOp(g, self.inputs[0], self.inputs[1], self, self.eps, ctx=self.raw_ctx)
        return [LayerNormGradientOfDataOp(ln_grad, self.inputs[0], ctx=self.raw_ctx),
                LayerNormGradientOfScaleOp(ln_grad, self.inputs[1], ctx=self.raw_ctx),
                LayerNormGradientOfBiasOp(ln_grad, self.inputs[2], ctx=self.raw_ctx)]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


class LayerNormGradientOp(Op):
    def __init__(self, out_gradient, in_node, ln_scale, forward_node, eps, ctx=None):
        super().__init__(LayerNormGradientOp, [out_gradient, in_node, ln_scale], ctx)
        self.forward_node, self.eps = forward_node, eps
        self.results = None

    def compute(self, input_vals, output_val=None, stream_handle=None):
        dy, x, gamma = input_vals
        if x.device.type == 'meta':
            return torch.empty_like(x)
        fwd = self.forward_node
        cols = x.shape[-1]
        rows = x.numel() // cols
        if _k(x, dy, gamma):
            x, dy = x.contiguous(), dy.contiguous()
            dx = torch.empty_like(x)
            dgamma = torch.empty(cols, dtype=x.dtype, device=x.device)
            dbeta = torch.empty(cols, dtype=x.dtype, device=x.device)
            L.call('layernorm_bwd', L.p(dy), L.p(x), L.p(gamma.contiguous()), L.p(fwd.save_mean),
                   L.p(fwd.save_rstd), L.p(dx), L.p(dgamma), L.p(dbeta), L.c_int(rows), L.c_int(cols),
                   L.stream_ptr(x.device))
        else:
            x2, dy2 = x.reshape(rows, cols), dy.reshape(rows, cols)
            xhat = (x2 - fwd.save_mean.reshape(-1, 1)) * fwd.save_rstd.reshape(-1, 1)
            dbeta = dy2.sum(0)
            dgamma = (dy2 * xhat).sum(0)
            dxhat = dy2 * gamma.reshape(1, -1)
            dx = fwd.save_rstd.reshape(-1, 1) * (
                dxhat - dxhat.mean(1, keepdim=True) - xhat * (dxhat * xhat).mean(1, keepdim=True))
            dx = dx.reshape(x.shape)
        self.results = (dx, dgamma, dbeta)
        return dx

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[1])


class LayerNormGradientOfDataOp(_BNPick):
    which = 0


class LayerNormGradientOfScaleOp(_BNPick):
    which = 1


class LayerNormGradientOfBiasOp(_BNPick):
    which = 2


class InstanceNorm2dOp(Op):
    def __init__(self, node_in, eps=0.01, ctx=None):
        super().__init__(InstanceNorm2dOp, [node_in], ctx)
        self.eps = eps
        self.save_mean = self.save_rstd = None

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        if x.device.type == 'meta':
            return torch.empty_like(x)
        n, c = x.shape[:2]
        hw = x.numel() // (n * c)
        if _k(x):
            x = x.contiguous()
            y = torch.empty_like(x)
            self.save_mean = torch.empty(n * c, dtype=x.dtype, device=x.device)
            self.save_rstd = torch.empty(n * c, dtype=x.dtype, device=x.device)
            L.call('layernorm_fwd', L.p(x), L.p(None), L.p(None), L.p(y), L.p(self.save_mean),
                   L.p(self.save_rstd), L.c_int(n * c), L.c_int(hw), L.c_f32(self.eps),
                   L.stream_ptr(x.device))
            return y
        x2 = x.reshape(n * c, hw)
        mean = x2.mean(1, keepdim=True)
        rstd = (x2.var(1, unbiased=False, keepdim=True) + self.eps).rsqrt()
        self.save_mean, self.save_rstd = mean.reshape(-1), rstd.reshape(-1)
        return ((x2 - mean) * rstd).reshape(x.shape)

    def gradient(self, g):
        return [InstanceNorm2dGradientOp(g, self.inputs[0], self, ctx=self.raw_ctx)]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


class InstanceNorm2dGradientOp(Op):
    def __init__(self, out_gradient, in_node, forward_node, ctx=None):
        super().__init__(InstanceNorm2dGradientOp, [out_gradient, in_node], ctx)
        self.forward_node = forward_node

    def compute(self, input_vals, output_val=None, stream_handle=None):
        dy, x = input_vals
        if x.device.type == 'meta':
            return torch.empty_like(x)
        fwd = self.forward_node
        n, c = x.shape[:2]
        hw = x.numel() // (n * c)
        x2, dy2 = x.reshape(n * c, hw), dy.reshape(n * c, hw)
        xhat = (x2 - fwd.save_mean.reshape(-1, 1)) * fwd.save_rstd.reshape(-1, 1)
        dx = fwd.save_rstd.reshape(-1, 1) * (
            dy2 - dy2.mean(1, keepdim=True) - xhat * (dy2 * xhat).mean(1, keepdim=True))
        return dx.reshape(x.shape)

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[1])


# ---------------------------------------------------------------- dropout
_step_counters = {}


def step_counter(device) -> torch.Tensor:
    """Per-device int64 step counter; the executor bumps it once per run so that
    dropout masks differ between steps even when the step is a replayed CUDA graph."""
    key = str(device)
    if key not in _step_counters:
        _step_counters[key] = torch.zeros(1, dtype=torch.int64, device=device)
    return _step_counters[key]


class DropoutOp(Op):
    takes_inference = True
    uses_step_counter = True

    def __init__(self, node_in, keep_prob, recompute=True, inplace=False, ctx=None):
        super().__init__(DropoutOp, [node_in], ctx)
        self.keep_prob, self.recompute, self.inplace = keep_prob, recompute, inplace
        self.mask = None
        self.seed = 0x9E3779B97F4A7C15 ^ (self.id * 0x100000001B3)

    def _apply(self, x):
        if _k(x):
            x = x.contiguous()
            y = torch.empty_like(x)
            L.call('dropout', L.p(x), L.p(y), L.c_sz(x.numel()), L.c_f32(1.0 - self.keep_prob),
                   L.c_u64(self.seed & 0xFFFFFFFFFFFFFFFF), L.p(step_counter(x.device)),
                   L.stream_ptr(x.device))
            return y
        return x * self.mask / self.keep_prob

    def compute(self, input_vals, output_val=None, stream_handle=None, inference=False):
        x = input_vals[0]
        if x.device.type == 'meta' or inference or self.keep_prob >= 1.0:
            return x
        if not _k(x):
            self.mask = (torch.rand(x.shape, device=x.device) < self.keep_prob).to(x.dtype)
        return self._apply(x)

    def gradient(self, g):
        return [dropout_gradient_op(g, self.keep_prob, self, ctx=self.raw_ctx)]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


class DropoutGradientOp(Op):
    def __init__(self, node_in, keep_prob, forward_node, ctx=None):
        super().__init__(DropoutGradientOp, [node_in], ctx)
        self.keep_prob, self.forward_node = keep_prob, forward_node

    def compute(self, input_vals, output_val=None, stream_handle=None):
        g = input_vals[0]
        if g.device.type == 'meta' or self.keep_prob >= 1.0:
            return g
        return self.forward_node._apply(g)       # same (seed, step) => same mask

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


class Dropout2dOp(DropoutOp):
    """Channel-wise dropout on NCHW: one Bernoulli draw per (n, c) plane."""

    def __init__(self, node_in, keep_prob, ctx=None):
        super().__init__(node_in, keep_prob, ctx=ctx)
        self.op_type = 'Dropout2dOp'

    def _apply(self, x):
        return x * self.mask / self.keep_prob

    def compute(self, input_vals, output_val=None, stream_handle=None, inference=False):
        x = input_vals[0]
        if x.device.type == 'meta' or inference or self.keep_prob >= 1.0:
            return x
        m = (torch.rand(x.shape[:2], device=x.device) < self.keep_prob).to(x.dtype)
        self.mask = m.reshape(tuple(x.shape[:2]) + (1,) * (x.dim() - 2))
        return self._apply(x)

    def gradient(self, g):
        return [dropout2d_gradient_op(g, self.keep_prob, self, ctx=self.raw_ctx)]


# ---------------------------------------------------------------- softmax
def softmax_func(y):
    """numpy helper kept for API parity (reference Softmax.py:10)."""
    b = y - np.max(y, axis=-1, keepdims=True)
    e = np.exp(b)
    return e / np.sum(e, axis=-1, keepdims=True)


def softmax_rows(x: torch.Tensor) -> torch.Tensor:
    if _k(x):
        x = x.contiguous()
        y = torch.empty_like(x)
        cols = x.shape[-1]
        L.call('softmax_rows', L.p(x), L.p(y), L.c_int(x.numel() // cols), L.c_int(cols),
               L.stream_ptr(x.device))
        return y
    return torch.softmax(x, dim=-1)


class SoftmaxOp(Op):
    def __init__(self, node_A, ctx=None):
        super().__init__(SoftmaxOp, [node_A], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        if x.device.type == 'meta':
            return torch.empty_like(x)
        return softmax_rows(x)

    def gradient(self, g):
        return [SoftmaxGradientOp(self, g, ctx=self.raw_ctx)]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


class SoftmaxGradientOp(Op):
    """dx = y * (dy - sum(dy * y))"""

    def __init__(self, node_y, grad, ctx=None):
        super().__init__(SoftmaxGradientOp, [node_y, grad], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        y, dy = input_vals
        if y.device.type == 'meta':
            return torch.empty_like(y)
        if _k(y, dy):
            y, dy = y.contiguous(), dy.contiguous()
            dx = torch.empty_like(y)
            cols = y.shape[-1]
            L.call('softmax_grad_rows', L.p(y), L.p(dy), L.p(dx), L.c_int(y.numel() // cols),
                   L.c_int(cols), L.stream_ptr(y.device))
            return dx
        return y * (dy - (dy * y).sum(dim=-1, keepdim=True))

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


# ---------------------------------------------------------------- public API
def conv2d_op(node_A, node_B, padding=0, stride=1, ctx=None):
    return Conv2dOp(node_A, node_B, padding, stride, ctx=ctx)


def conv2d_gradient_of_data_op(node_A, node_B, node_C, padding=0, stride=1, ctx=None):
    return Conv2dGradientOfDataOp(node_A, node_B, node_C, padding, stride, ctx=ctx)


def conv2d_gradient_of_filter_op(input_X, gradient_Y, input_filter, padding=0, stride=1, ctx=None):
    return Conv2dGradientOfFilterOp(input_X, gradient_Y, input_filter, padding, stride, ctx=ctx)


def conv2d_add_bias_op(node_A, node_B, bias, padding=0, stride=1, ctx=None):
    return Conv2dAddBiasOp(node_A, node_B, bias, padding, stride, ctx=ctx)


def conv2d_broadcastto_op(node_A, node_B, ctx=None):
    return Conv2dBroadcastToOp(node_A, node_B, ctx=ctx)


def conv2d_reducesum_op(node, ctx=None):
    return Conv2dReduceSumOp(node, ctx=ctx)


def max_pool2d_op(node_A, kernel_H, kernel_W, padding, stride, ctx=None):
    return MaxPool2dOp(node_A, kernel_H, kernel_W, padding, stride, ctx=ctx)


def max_pool2d_gradient_op(node_out, node_out_gradient, node_in, kernel_H, kernel_W, padding,
                           stride, ctx=None):
    return MaxPool2dGradientOp(node_out, node_out_gradient, node_in, kernel_H, kernel_W, padding,
                               stride, ctx=ctx)


def avg_pool2d_op(node_A, kernel_H, kernel_W, padding, stride, ctx=None):
    return AvgPool2dOp(node_A, kernel_H, kernel_W, padding, stride, ctx=ctx)


def avg_pool2d_gradient_op(node_out, node_out_gradient, node_in, kernel_H, kernel_W, padding,
                           stride, ctx=None):
    return AvgPool2dGradientOp(node_out, node_out_gradient, node_in, kernel_H, kernel_W, padding,
                               stride, ctx=ctx)


def batch_normalization_op(node_in, bn_scale, bn_bias, momentum=0.1, eps=1e-5, ctx=None):
    return BatchNormOp(node_in, bn_scale, bn_bias, momentum, eps, ctx=ctx)


def batch_normalization_gradient_op(out_gradient, in_node, bn_scale, forward_node, eps, ctx=None):
    return BatchNormGradientOp(out_gradient, in_node, bn_scale, forward_node, eps, ctx=ctx)


def batch_normalization_gradient_of_data_op(bn_gradient, in_arr, ctx=None):
    return BatchNormGradientOfDataOp(bn_gradient, in_arr, ctx=ctx)


def batch_normalization_gradient_of_scale_op(bn_gradient, in_scale, ctx=None):
    return BatchNormGradientOfScaleOp(bn_gradient, in_scale, ctx=ctx)


def batch_normalization_gradient_of_bias_op(bn_gradient, in_bias, ctx=None):
    return BatchNormGradientOfBiasOp(bn_gradient, in_bias, ctx=ctx)


def layer_normalization_op(node_in, ln_scale, ln_bias, eps=0.01, ctx=None):
    return LayerNormOp(node_in, ln_scale, ln_bias, eps, ctx=ctx)


def instance_normalization2d_op(node_in, eps=0.01, ctx=None):
    return InstanceNorm2dOp(node_in, eps, ctx=ctx)


def dropout_op(node_in, keep_prob, recompute=True, inplace=False, ctx=None):
    return DropoutOp(node_in, keep_prob, recompute, inplace, ctx=ctx)


def dropout_gradient_op(node_in, keep_prob, forward_node, ctx=None):
    return DropoutGradientOp(node_in, keep_prob, forward_node, ctx=ctx)


def dropout2d_op(node_in, keep_prob, ctx=None):
    return Dropout2dOp(node_in, keep_prob, ctx=ctx)


def dropout2d_gradient_op(node_in, keep_prob, forward_node, ctx=None):
    return DropoutGradientOp(node_in, keep_prob, forward_node, ctx=ctx)


def softmax_op(node, ctx=None):
    return SoftmaxOp(node, ctx=ctx)
