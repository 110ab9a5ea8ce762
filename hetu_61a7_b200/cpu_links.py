"""ctypes bindings of the CPU op library (``lib/libhetu_cpu.so``, csrc_cpu/cpu_ops.cc).

reference: python/hetu/cpu_links/dnnl_op.py:9-391 - the same 47 entry points, same argument
order (outputs are written in place), so scripts and tests written against ``ht.cpu_links``
keep working.  Arguments may be ``NDArray``, ``torch.Tensor`` (CPU, fp32, contiguous) or numpy
arrays; what is behind the ABI is an OpenMP library instead of oneDNN (not in this image).
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from . import _lib as L
from .ndarray import NDArray

c_int, c_i64, c_f32, c_u64 = ctypes.c_int, ctypes.c_int64, ctypes.c_float, ctypes.c_uint64

U = dict(RELU=0, SIGMOID=1, TANH=2, GELU=3, SQRT=4, RSQRT=5, NEG=6, EXP=7, LOG=8, ABS=9, ADDC=10, MULC=11, CDIV=12)
B = dict(ADD=0, SUB=1, MUL=2, DIV=3, RELU_GRAD=4, SIGMOID_GRAD=5, TANH_GRAD=6, GELU_GRAD=7)


def available() -> bool:
    return L.cpu_lib() is not None


def _lib():
    lib = L.cpu_lib()
    if lib is None:
        raise RuntimeError('libhetu_cpu.so is not built; run `python -c "import __graft_entry__ as g; g.build()"`')
    return lib


def _t(a) -> torch.Tensor:
    if isinstance(a, NDArray):
        a = a.tensor
    if isinstance(a, np.ndarray):
        a = torch.from_numpy(a)
    assert a.device.type == 'cpu' and a.is_contiguous(), 'cpu_links need contiguous CPU tensors'
    return a


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


def _f(*arrs):
    ts = [_t(a) for a in arrs]
    for t in ts:
        assert t.dtype == torch.float32, 'cpu_links operate on fp32'
    return ts


def _unary(code, x, y, c=0.0):
    x, y = _f(x, y)
    _lib().hetu_cpu_unary(_p(x), _p(y), c_i64(x.numel()), c_int(code), c_f32(c))


def _binary(code, a, b, y):
    a, b, y = _f(a, b, y)
    if b.numel() != a.numel():
        b = b.expand_as(a).contiguous()
    _lib().hetu_cpu_binary(_p(a), _p(b), _p(y), c_i64(a.numel()), c_int(code))


def matrix_multiply(matA, transposeA, matB, transposeB, matC):
    a, b, c = _f(matA, matB, matC)
    M = a.shape[1] if transposeA else a.shape[0]
    K = a.shape[0] if transposeA else a.shape[1]
    N = b.shape[0] if transposeB else b.shape[1]
    _lib().hetu_cpu_matmul(_p(a), _p(b), _p(c), c_int(M), c_int(N), c_int(K), c_int(int(transposeA)),
                           c_int(int(transposeB)), c_f32(1.0), c_f32(0.0))


def batch_matrix_multiply(matA, transposeA, matB, transposeB, matC):
    a, b, c = _f(matA, matB, matC)
    batch = int(np.prod(a.shape[:-2]))
    M = a.shape[-1] if transposeA else a.shape[-2]
    K = a.shape[-2] if transposeA else a.shape[-1]
    N = b.shape[-2] if transposeB else b.shape[-1]
    _lib().hetu_cpu_batch_matmul(_p(a), _p(b), _p(c), c_int(batch), c_int(M), c_int(N), c_int(K),
                                 c_int(int(transposeA)), c_int(int(transposeB)))


def matrix_elementwise_multiply_by_const(mat, val, output):
    _unary(U['MULC'], mat, output, val)


def matrix_elementwise_multiply(matA, matB, output):
    _binary(B['MUL'], matA, matB, output)


def matrix_elementwise_add_by_const(mat, val, output):
    _unary(U['ADDC'], mat, output, val)


def matrix_elementwise_add(matA, matB, output):
    _binary(B['ADD'], matA, matB, output)


def matrix_elementwise_divide_by_const(mat, val, output):
    _unary(U['MULC'], mat, output, 1.0 / val)


def matrix_elementwise_divide(matA, matB, output):
    _binary(B['DIV'], matA, matB, output)


def broadcast_to(in_arr, out_arr):
    x, y = _f(in_arr, out_arr)
    _lib().hetu_cpu_broadcast_to(_p(x), _p(y), c_i64(x.numel()), c_i64(y.numel()))


def reduce_sum_axis_zero(in_arr, out_arr):
    x, y = _f(in_arr, out_arr)
    _lib().hetu_cpu_reduce_sum_axis0(_p(x), _p(y), c_i64(x.shape[0]), c_i64(x.numel() // max(x.shape[0], 1)))


def array_set(output, value):
    (y,) = _f(output)
    _lib().hetu_cpu_array_set(_p(y), c_i64(y.numel()), c_f32(value))


def reshape(in_arr, out_arr):
    x, y = _f(in_arr, out_arr)
    _lib().hetu_cpu_reshape(_p(x), _p(y), c_i64(x.numel()))


def softmax(mat, output):
    x, y = _f(mat, output)
    _lib().hetu_cpu_softmax(_p(x), _p(y), c_i64(x.numel() // x.shape[-1]), c_i64(x.shape[-1]))


def softmax_crossentropy(matA, matB, output):
    x, lab, y = _f(matA, matB, output)
    _lib().hetu_cpu_softmax_cross_entropy(_p(x), _p(lab), _p(y), c_i64(x.numel() // x.shape[-1]), c_i64(x.shape[-1]))


def softmax_crossentropy_gradient(matA, matB, grad, output):
    x, lab, g, y = _f(matA, matB, grad, output)
    _lib().hetu_cpu_softmax_cross_entropy_grad(_p(x), _p(lab), _p(g), _p(y), c_i64(x.numel() // x.shape[-1]),
                                               c_i64(x.shape[-1]))


def sqrt(in_arr, out_arr):
    _unary(U['SQRT'], in_arr, out_arr)


def rsqrt(in_arr, out_arr):
    _unary(U['RSQRT'], in_arr, out_arr)


def tanh(in_arr, out_arr):
    _unary(U['TANH'], in_arr, out_arr)


def opposite(in_arr, out_arr):
    _unary(U['NEG'], in_arr, out_arr)


def sigmoid(in_arr, out_arr):
    _unary(U['SIGMOID'], in_arr, out_arr)


def relu(in_arr, out_arr):
    _unary(U['RELU'], in_arr, out_arr)


def relu_gradient(input, in_grad, output):
    _binary(B['RELU_GRAD'], input, in_grad, output)


def gelu(in_arr, out_arr):
    _unary(U['GELU'], in_arr, out_arr)


def gelu_gradient(input, in_grad, output):
    _binary(B['GELU_GRAD'], input, in_grad, output)


def _conv_dims(x, f):
    N, C, H, W = x.shape
    F, _, kh, kw = f.shape
    return [c_int(v) for v in (N, C, H, W, F, kh, kw)]


def conv2d(input_x, input_f, output, padding=0, stride=1):
    x, f, y = _f(input_x, input_f, output)
    _lib().hetu_cpu_conv2d(_p(x), _p(f), _p(y), *_conv_dims(x, f), c_int(padding), c_int(stride))


def conv2d_gradient_of_data(input_f, gradient_y, gradient_x, padding=0, stride=1):
    f, gy, gx = _f(input_f, gradient_y, gradient_x)
    _lib().hetu_cpu_conv2d_grad_data(_p(f), _p(gy), _p(gx), *_conv_dims(gx, f), c_int(padding), c_int(stride))


def conv2d_gradient_of_filter(input_x, gradient_y, gradient_f, padding=0, stride=1):
    x, gy, gf = _f(input_x, gradient_y, gradient_f)
    _lib().hetu_cpu_conv2d_grad_filter(_p(x), _p(gy), _p(gf), *_conv_dims(x, gf), c_int(padding), c_int(stride))


def _pool(mode, x, kh, kw, y, padding, stride):
    x, y = _f(x, y)
    N, C, H, W = x.shape
    _lib().hetu_cpu_pool2d(_p(x), _p(y), c_i64(N * C), c_int(H), c_int(W), c_int(kh), c_int(kw), c_int(padding),
                           c_int(stride), c_int(mode))


def avg_pool(input, H, W, output, padding=0, stride=1):
    _pool(1, input, H, W, output, padding, stride)


def max_pool(input, H, W, output, padding=0, stride=1):
    _pool(0, input, H, W, output, padding, stride)


def avg_pool_gradient(gradient_Y, H, W, gradient_X, padding=0, stride=1):
    gy, gx = _f(gradient_Y, gradient_X)
    N, C, Hi, Wi = gx.shape
    _lib().hetu_cpu_pool2d_grad(_p(gx), _p(gy), _p(gy), _p(gx), c_i64(N * C), c_int(Hi), c_int(Wi), c_int(H), c_int(W),
                                c_int(padding), c_int(stride), c_int(1))


def max_pool_gradient(input, input_grad, H, W, output, padding=0, stride=1):
    x, gy, gx = _f(input, input_grad, output)
    N, C, Hi, Wi = x.shape
    y = torch.empty_like(gy)
    _pool(0, x, H, W, y, padding, stride)
    _lib().hetu_cpu_pool2d_grad(_p(x), _p(y), _p(gy), _p(gx), c_i64(N * C), c_int(Hi), c_int(Wi), c_int(H), c_int(W),
                                c_int(padding), c_int(stride), c_int(0))


def batch_norm(input, bn_scale, bn_bias, output, running_mean, running_var, save_mean, save_var, momentum=0.99,
               eps=0.01):
    x, s, b, y, rm, rv, sm, sv = _f(input, bn_scale, bn_bias, output, running_mean, running_var, save_mean, save_var)
    N, C = x.shape[0], x.shape[1]
    _lib().hetu_cpu_batch_norm(_p(x), _p(s), _p(b), _p(y), _p(rm), _p(rv), _p(sm), _p(sv), c_int(N), c_int(C),
                               c_int(x.numel() // (N * C)), c_f32(momentum), c_f32(eps), c_int(1))


def batch_norm_inference(input, bn_scale, bn_bias, output, mean, var, eps):
    x, s, b, y, rm, rv = _f(input, bn_scale, bn_bias, output, mean, var)
    N, C = x.shape[0], x.shape[1]
    _lib().hetu_cpu_batch_norm(_p(x), _p(s), _p(b), _p(y), _p(rm), _p(rv), None, None, c_int(N), c_int(C),
                               c_int(x.numel() // (N * C)), c_f32(0.0), c_f32(eps), c_int(0))


def batch_norm_gradient(gradient_Y, input_X, bn_scale, gradient_X, gradient_bn_scale, gradient_bn_bias, mean, var,
                        eps=0.01):
    """``mean`` / ``var`` are the save_mean / save_invstd written by ``batch_norm``."""
    gy, x, s, gx, gs, gb, m, v = _f(gradient_Y, input_X, bn_scale, gradient_X, gradient_bn_scale, gradient_bn_bias,
                                    mean, var)
    N, C = x.shape[0], x.shape[1]
    _lib().hetu_cpu_batch_norm_grad(_p(x), _p(gy), _p(s), _p(m), _p(v), _p(gx), _p(gs), _p(gb), c_int(N), c_int(C),
                                    c_int(x.numel() // (N * C)))


def concat(input_x, input_y, output, axis=0):
    a, b, y = _f(input_x, input_y, output)
    outer = int(np.prod(a.shape[:axis])) if axis > 0 else 1
    _lib().hetu_cpu_concat(_p(a), _p(b), _p(y), c_i64(outer), c_i64(a.numel() // outer), c_i64(b.numel() // outer))


def concat_gradient(output_gradient, input_gradient, axis=0, id=0):
    g, ga = _f(output_gradient, input_gradient)
    outer = int(np.prod(g.shape[:axis])) if axis > 0 else 1
    total, mine = g.numel() // outer, ga.numel() // outer
    _lib().hetu_cpu_concat_grad(_p(g), _p(ga), c_i64(outer), c_i64(mine), c_i64(total),
                                c_i64(0 if id == 0 else total - mine))


_DROPOUT_SEED = [0x5EED]


def dropout(in_arr, dropout, out_arr, seed=None):
    """``dropout`` is the KEEP probability (reference Dropout.py:19-31 passes keep_prob)."""
    x, y = _f(in_arr, out_arr)
    if seed is None:
        _DROPOUT_SEED[0] += 1
        seed = _DROPOUT_SEED[0]
    _lib().hetu_cpu_dropout(_p(x), _p(y), c_i64(x.numel()), c_f32(dropout), c_u64(seed))
    return seed


def dropout_gradient(in_gradient_y, dropout, out_gradient_x, seed=None):
    return globals()['dropout'](in_gradient_y, dropout, out_gradient_x, _DROPOUT_SEED[0] if seed is None else seed)


def _pads(paddings, ndim):
    assert len(paddings) == ndim and all(tuple(p) == (0, 0) for p in paddings[:-2]), \
        'only the two innermost dims can be padded'
    (pt, pb), (pl, pr) = paddings[-2], paddings[-1]
    return [c_int(v) for v in (pt, pb, pl, pr)]


def pad(in_arr, out_arr, paddings, mode='CONSTANT', constant_values=0):
    assert mode == 'CONSTANT'
    x, y = _f(in_arr, out_arr)
    H, W = x.shape[-2:]
    _lib().hetu_cpu_pad(_p(x), _p(y), c_i64(x.numel() // (H * W)), c_int(H), c_int(W), *_pads(paddings, x.dim()),
                        c_f32(constant_values))


def pad_gradient(out_grad_arr, in_grad_arr, paddings, mode='CONSTANT'):
    gy, gx = _f(out_grad_arr, in_grad_arr)
    H, W = gx.shape[-2:]
    _lib().hetu_cpu_pad_grad(_p(gy), _p(gx), c_i64(gx.numel() // (H * W)), c_int(H), c_int(W),
                             *_pads(paddings, gx.dim()))


def transpose(in_arr, out_arr, perm):
    x, y = _f(in_arr, out_arr)
    nd = x.dim()
    shp = (c_i64 * nd)(*x.shape)
    pm = (c_int * nd)(*perm)
    _lib().hetu_cpu_transpose(_p(x), _p(y), shp, pm, c_int(nd))


def embedding_lookup(in_mat, ids, out_mat):
    tab, out = _f(in_mat, out_mat)
    idx = _t(ids)
    idx = idx.long().contiguous()
    _lib().hetu_cpu_embedding_lookup(_p(tab), _p(idx), _p(out), c_i64(idx.numel()), c_i64(tab.shape[-1]),
                                     c_i64(tab.shape[0]))


def add_l2_regularization(param, grad, l2reg):
    p, g = _f(param, grad)
    g.add_(p, alpha=l2reg)


def sgd_update(param, grad, lr, l2reg=0.0):
    p, g = _f(param, grad)
    _lib().hetu_cpu_sgd_update(_p(p), _p(g), c_i64(p.numel()), c_f32(lr), c_f32(l2reg))


def sgd_update_sparse(param, grad_ind, grad_value, lr):
    p, gv = _f(param, grad_value)
    idx = _t(grad_ind).long().contiguous()
    _lib().hetu_cpu_sgd_sparse_update(_p(p), _p(idx), _p(gv), c_i64(idx.numel()), c_i64(p.shape[-1]), c_f32(lr))


def momentum_update(param, grad, velocity, lr, momentum, nesterov, l2reg=0.0):
    p, g, v = _f(param, grad, velocity)
    _lib().hetu_cpu_momentum_update(_p(p), _p(g), _p(v), c_i64(p.numel()), c_f32(lr), c_f32(momentum),
                                    c_int(int(nesterov)), c_f32(l2reg))


def adagrad_update(param, grad, accumulation, lr, eps, l2reg=0.0):
    p, g, a = _f(param, grad, accumulation)
    _lib().hetu_cpu_adagrad_update(_p(p), _p(g), _p(a), c_i64(p.numel()), c_f32(lr), c_f32(eps), c_f32(l2reg))


def adam_update(param, grad, expavg, expavgsq, lr, beta1, beta2, beta1t, beta2t, eps, step=None, l2reg=0.0):
    """``beta1t`` / ``beta2t`` (beta**t, reference signature) are used to recover ``t`` when
    ``step`` is not given."""
    p, g, m, v = _f(param, grad, expavg, expavgsq)
    if step is None:
        step = max(1, int(round(np.log(max(beta1t, 1e-30)) / np.log(beta1)))) if 0 < beta1 < 1 else 1
    _lib().hetu_cpu_adam_update(_p(p), _p(g), _p(m), _p(v), c_i64(p.numel()), c_f32(lr), c_f32(beta1), c_f32(beta2),
                                c_f32(eps), c_i64(step), c_f32(l2reg), c_int(0), None)


def normal_init(param, mean, stddev, seed):
    (p,) = _f(param)
    _lib().hetu_cpu_normal_init(_p(p), c_i64(p.numel()), c_f32(mean), c_f32(stddev), c_u64(seed))


def uniform_init(param, lb, ub, seed):
    (p,) = _f(param)
    _lib().hetu_cpu_uniform_init(_p(p), c_i64(p.numel()), c_f32(lb), c_f32(ub), c_u64(seed))


def truncated_normal_init(param, mean, stddev, seed):
    (p,) = _f(param)
    _lib().hetu_cpu_truncated_normal_init(_p(p), c_i64(p.numel()), c_f32(mean), c_f32(stddev), c_u64(seed))
