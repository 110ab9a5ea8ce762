"""Imperative kernel API: ``ht.gpu_links.<kernel>(inputs..., output, stream=None)``.

reference: python/hetu/gpu_links/*.py (113 files, ~180 functions) - thin ctypes wrappers, one per
CUDA kernel of src/ops/*.cu, that write their result into a caller-provided output ``NDArray``;
tests/test_gpu_op.py is written against this surface.

Same names, same argument order, same in-place-output convention.  Each call runs the framework's
own op implementation (``Op.compute``), so arrays that live on a B200 go through the sm_100a
kernels of ``lib/libhetu_b200.so`` and host arrays through the CPU library / torch - the answer
does not depend on where the arrays are.  Arguments may be ``NDArray``, torch tensors or numpy
arrays (numpy outputs are written through a temporary).  ``stream`` is accepted for source
compatibility; work is issued on the current CUDA stream.
"""
from __future__ import annotations

import builtins as _b

import numpy as np
import torch

from . import ops as _ops
from .ndarray import IndexedSlices, ND_Sparse_Array, NDArray
from .ops.variable import placeholder_op as _placeholder


# --------------------------------------------------------------------------------- plumbing
def _t(a):
    if isinstance(a, NDArray):
        return a.tensor
    if isinstance(a, np.ndarray):
        return torch.from_numpy(a)
    return a


def _store(out, res):
    """Write ``res`` into the caller's output array (shape of the output wins when sizes agree)."""
    o = _t(out)
    res = res if torch.is_tensor(res) else torch.as_tensor(res)
    if res.numel() == o.numel():
        res = res.reshape(o.shape)
    o.copy_(res.to(device=o.device, dtype=o.dtype))
    return out


_ph_count = [0]


def _node(t):
    _ph_count[0] += 1
    n = _placeholder('gpu_links_in%d' % _ph_count[0], trainable=False)
    if torch.is_tensor(t):
        n.shape = tuple(t.shape)
    return n


def _run(ctor, ins, *args, **kw):
    """Build the op over stand-in nodes and evaluate it on the given arrays."""
    vals = [a if isinstance(a, (ND_Sparse_Array, IndexedSlices)) else _t(a) for a in ins]
    op = ctor(*[_node(v) for v in vals], *args, **kw)
    try:
        op.infer_shape([tuple(v.shape) for v in vals])
    except Exception:
        pass
    return op.compute(vals, None, None)


def _go(ctor, ins, out, *args, **kw):
    return _store(out, _run(ctor, ins, *args, **kw))


def _pair(v):
    return tuple(v) if isinstance(v, (tuple, list)) else (v, v)


# ------------------------------------------------------------------- element-wise / scalar
def abs_val(in_mat, out_mat, stream=None):
    return _go(_ops.abs_op, [in_mat], out_mat)


def abs_gradient(in_mat, grad_mat, out_mat, stream=None):
    return _store(out_mat, _t(grad_mat) * torch.sign(_t(in_mat)))


def matrix_elementwise_add_by_const(in_mat, val, out_mat, stream=None):
    return _go(_ops.addbyconst_op, [in_mat], out_mat, val)


def matrix_elementwise_add(matA, matB, matC, lazy_input=False, stream=None):
    return _go(_ops.add_op, [matA, matB], matC)


matrix_elementwise_add_simple = matrix_elementwise_add


def matrix_elementwise_add_lazy(matA, matB, matC, gpu_buf=None, stream=None):
    return matrix_elementwise_add(matA, matB, matC)


def matrix_elementwise_minus(input1, input2, output, stream=None):
    return _go(_ops.minus_op, [input1, input2], output)


def minus_by_const(input, output, val, stream=None):
    """output = val - input (reference MinusByConst.cu)."""
    return _go(_ops.minus_byconst_op, [input], output, val)


def matrix_elementwise_multiply_by_const(in_mat, val, out_mat, stream=None):
    return _go(_ops.mul_byconst_op, [in_mat], out_mat, val)


def matrix_elementwise_multiply(matA, matB, matC, stream=None):
    return _go(_ops.mul_op, [matA, matB], matC)


def matrix_elementwise_divide_const(val, in_mat, out_mat, stream=None):
    """out = val / in."""
    return _go(lambda n: _ops.div_const_op(val, n), [in_mat], out_mat)


def matrix_elementwise_divide(matA, matB, matC, stream=None):
    return _go(_ops.div_op, [matA, matB], matC)


def matrix_opposite(in_arr, out_arr, stream=None):
    return _go(_ops.opposite_op, [in_arr], out_arr)


def matrix_sqrt(in_arr, out_arr, stream=None):
    return _go(_ops.sqrt_op, [in_arr], out_arr)


def matrix_rsqrt(in_arr, out_arr, stream=None):
    return _go(_ops.rsqrt_op, [in_arr], out_arr)


def exp(in_arr, out_arr, stream=None):
    return _go(_ops.exp_op, [in_arr], out_arr)


def floor(in_arr, out_arr, stream=None):
    return _go(_ops.floor_op, [in_arr], out_arr)


def log_link(input, output, stream=None):
    return _go(_ops.log_op, [input], output)


def log_grad_link(output_grad, input, input_grad, stream=None):
    return _go(_ops.log_grad_op, [output_grad, input], input_grad)


def sin(in_arr, out_arr, stream=None):
    return _go(_ops.sin_op, [in_arr], out_arr)


def cos(in_arr, out_arr, stream=None):
    return _go(_ops.cos_op, [in_arr], out_arr)


def sigmoid(in_arr, out_arr, stream=None):
    return _go(_ops.sigmoid_op, [in_arr], out_arr)


def tanh(in_arr, out_arr, stream=None):
    return _go(_ops.tanh_op, [in_arr], out_arr)


def tanh_gradient(forward_arr, grad_arr, out_arr, stream=None):
    return _go(_ops.tanh_gradient_op, [forward_arr, grad_arr], out_arr)


def relu(in_arr, out_arr, stream=None):
    return _go(_ops.relu_op, [in_arr], out_arr)


def relu_gradient(in_arr, in_grad_arr, out_arr, stream=None):
    return _go(_ops.relu_gradient_op, [in_arr, in_grad_arr], out_arr)


def leaky_relu(in_arr, alpha, out_arr, stream=None):
    return _go(_ops.leaky_relu_op, [in_arr], out_arr, alpha)


def leaky_relu_gradient(in_arr, in_grad_arr, alpha, out_arr, stream=None):
    return _go(_ops.leaky_relu_gradient_op, [in_arr, in_grad_arr], out_arr, alpha)


def gelu(in_arr, out_arr, stream=None):
    return _go(_ops.gelu_op, [in_arr], out_arr)


def gelu_gradient(in_arr, in_grad_arr, out_arr, stream=None):
    return _go(_ops.gelu_gradient_op, [in_arr, in_grad_arr], out_arr)


def pow_matrix(in_arr, out_arr, eps, stream=None):
    """out = in ** eps."""
    return _go(_ops.pow_op, [in_arr], out_arr, eps)


def pow_gradient(in_arr, in_grad_arr, out_arr, eps, stream=None):
    return _go(_ops.pow_gradient_op, [in_arr, in_grad_arr], out_arr, eps)


def const_pow(in_arr, out_arr, val, stream=None):
    """out = val ** in."""
    return _go(_ops.const_pow_op, [in_arr], out_arr, val)


def const_pow_gradient(in_arr, grad_arr, out_arr, val, stream=None):
    return _go(_ops.const_pow_gradient_op, [in_arr, grad_arr], out_arr, val)


def bool(input, output, stream=None):
    """output = 1 where input > 0, else 0 (reference Bool.cu: bool_kernel)."""
    return _store(output, (_t(input) > 0).to(_t(output).dtype))


_CMP = {0: torch.eq, 1: torch.lt, 2: torch.gt, 3: torch.le, 4: torch.ge}


def bool_val(input, output, val, cond, stream=None):
    """cond: 0 ==, 1 <, 2 >, 3 <=, 4 >= against the scalar ``val``."""
    x = _t(input)
    return _store(output, _CMP[int(cond)](x, torch.full_like(x, float(val))).to(_t(output).dtype))


def bool_matrix(input_A, input_B, output, cond, stream=None):
    return _store(output, _CMP[int(cond)](_t(input_A), _t(input_B)).to(_t(output).dtype))


def clamp(input_mat, output_mat, mmin=None, mmax=None, min_mat=None, max_mat=None, stream=None):
    lo = _t(min_mat) if min_mat is not None else mmin
    hi = _t(max_mat) if max_mat is not None else mmax
    return _store(output_mat, torch.clamp(_t(input_mat), min=lo, max=hi))


def masked_fill(input, mask, output, val, stream=None):
    return _go(_ops.masked_fill_op, [input, mask], output, val)


def where(cond, arr1, arr2, out_arr, stream=None):
    return _go(_ops.where_op, [cond, arr1, arr2], out_arr)


def where_const(cond, arr1, const_attr, out_arr, stream=None):
    return _go(_ops.where_const_op, [cond, arr1], out_arr, const_attr)


def max_mat(matA, matB, out_mat, stream=None):
    return _store(out_mat, torch.maximum(_t(matA), _t(matB)))


def min_mat(matA, matB, out_mat, stream=None):
    return _store(out_mat, torch.minimum(_t(matA), _t(matB)))


def array_set(arr, value, stream=None):
    _t(arr).fill_(value)
    return arr


def arange(start, end, step, out_mat, stream=None):
    o = _t(out_mat)
    return _store(out_mat, torch.arange(start, end, step, dtype=o.dtype, device=o.device))


def clone(input_mat, output_mat, stream=None):
    return _store(output_mat, _t(input_mat))


def one_hot(in_arr, out_arr, stream=None):
    return _go(_ops.one_hot_op, [in_arr], out_arr, int(_t(out_arr).shape[-1]))


# ---------------------------------------------------------------------------- linear algebra
def matrix_multiply(matA, transA, matB, transB, matC, stream=None):
    return _store(matC, _run(_ops.matmul_op, [matA, matB], _b.bool(transA), _b.bool(transB)))


def matmul_with_bias(matA, transA, matB, transB, bias, matC, stream=None):
    return _store(matC, _run(_ops.linear_op, [matA, matB, bias], _b.bool(transA), _b.bool(transB)))


def batch_matrix_multiply(matA, transA, matB, transB, matC, stream=None):
    return _store(matC, _run(_ops.batch_matmul_op, [matA, matB], _b.bool(transA), _b.bool(transB)))


def addmm(input_mat, matA, matB, matC, alpha, beta, stream=None):
    """C = beta * input + alpha * A @ B."""
    return _store(matC, _run(_ops.addmm_op, [input_mat, matA, matB], alpha, beta))


def addmm_gradient(input_mat, output_mat, axis, beta, stream=None):
    g = _t(input_mat) * beta
    o = _t(output_mat)
    return _store(output_mat, g.sum(dim=int(axis)) if g.numel() != o.numel() else g)


def baddbmm(input_mat, matA, matB, matC, alpha, beta, stream=None):
    return _store(matC, _run(_ops.baddbmm_op, [input_mat, matA, matB], alpha, beta))


def matrix_dot(matA, matB, matC, stream=None):
    """Row-wise broadcasted product used by MatrixDot.cu: C = A * B with B broadcast over rows."""
    return _store(matC, _t(matA) * _t(matB))


def outer(matA, matB, matC, stream=None):
    return _store(matC, torch.outer(_t(matA).reshape(-1), _t(matB).reshape(-1)))


def CuSparse_Csrmv(mat, trans, in_arr, out_arr, stream=None):
    return _store(out_arr, _run(_ops.csrmv_op, [mat, in_arr], _b.bool(trans)))


def CuSparse_Csrmm(matA, transA, matB, transB, matC, stream=None, start_pos=-1, end_pos=-1):
    return _store(matC, _run(_ops.csrmm_op, [matA, matB], _b.bool(transA), _b.bool(transB)))


# ------------------------------------------------------------------------------- reductions
def reduce_sum(in_arr, out_arr, axes, stream=None):
    axes = [axes] if isinstance(axes, int) else list(axes)
    return _store(out_arr, _run(_ops.reduce_sum_op, [in_arr], axes))


def reduce_mean(in_arr, out_arr, axes, stream=None):
    axes = [axes] if isinstance(axes, int) else list(axes)
    return _store(out_arr, _run(_ops.reduce_mean_op, [in_arr], axes))


def reduce_sum_axis_zero(in_arr, out_arr, workspace_arr=None, stream=None):
    return _go(_ops.reducesumaxiszero_op, [in_arr], out_arr)


_reduce_sum_axis_zero = reduce_sum_axis_zero


def norm(in_arr, out_arr, axis, p, stream=None):
    return _store(out_arr, _run(_ops.norm_op, [in_arr], axis, p))


def norm_gradient(input, input_y, grad_y, output, axis, p, stream=None):
    return _store(output, _run(_ops.norm_gradient_op, [input, input_y, grad_y], axis, p))


def max(in_mat, out_mat_val, dim, stream=None):
    return _store(out_mat_val, torch.amax(_t(in_mat), dim=int(dim)))


def min(in_mat, out_mat_val, dim, stream=None):
    return _store(out_mat_val, torch.amin(_t(in_mat), dim=int(dim)))


def argmax(in_mat, out_mat, dim, stream=None):
    return _store(out_mat, _run(_ops.argmax_op, [in_mat], dim))


def argsort(input, output, index, output_index, dim, descending, stream=None):
    """Sorted values -> ``output``, their source positions -> ``output_index`` (``index`` is the
    reference's scratch array of positions)."""
    vals, idx = torch.sort(_t(input), dim=int(dim), descending=_b.bool(descending), stable=True)
    _store(output, vals)
    return _store(output_index, idx)


def topk_idx(in_mat, out_mat_idx, k, stream=None):
    return _store(out_mat_idx, _run(_ops.topk_idx_op, [in_mat], int(k)))


def topk_val(in_mat, out_mat_idx, out_mat_val, k, stream=None):
    vals, idx = torch.topk(_t(in_mat), int(k), dim=-1)
    _store(out_mat_idx, idx)
    return _store(out_mat_val, vals)


def cumsum_with_bias(input, output, bias, dim, stream=None):
    return _store(output, _run(_ops.cumsum_with_bias_op, [input], bias, dim))


# ----------------------------------------------------------------------- shape / indexing
def array_reshape(in_arr, out_arr, stream=None):
    return _store(out_arr, _t(in_arr).reshape(_t(out_arr).shape))


def matrix_transpose(in_mat, out_mat, perm, stream=None):
    return _store(out_mat, _run(_ops.transpose_op, [in_mat], list(perm)))


def matrix_transpose_simple(in_mat, out_mat, gpu_buf=None, stream=None):
    x = _t(in_mat)
    return _store(out_mat, x.permute(*reversed(range(x.dim()))))


def broadcast_to(in_arr, out_arr, stream=None):
    x, o = _t(in_arr), _t(out_arr)
    return _store(out_arr, x.expand(o.shape) if x.dim() == o.dim() else x.reshape((1,) * (o.dim() - x.dim()) + tuple(x.shape)).expand(o.shape))


def broadcast_shape(in_arr, out_arr, add_axes=None, stream=None):
    x, o = _t(in_arr), _t(out_arr)
    if add_axes:
        for ax in sorted(add_axes):
            x = x.unsqueeze(int(ax))
    elif x.dim() < o.dim():
        x = x.reshape((1,) * (o.dim() - x.dim()) + tuple(x.shape))
    return _store(out_arr, x.expand(o.shape))


def broadcast_shape_simple(in_arr, out_arr, out_strides=None, in_dims=None, stream=None):
    return broadcast_shape(in_arr, out_arr)


def conv2d_broadcast_to(in_arr, out_arr, stream=None):
    """Per-channel bias [C] -> NCHW."""
    return _store(out_arr, _t(in_arr).reshape(1, -1, 1, 1).expand(_t(out_arr).shape))


def conv2d_reduce_sum(in_arr, out_arr, stream=None):
    return _store(out_arr, _t(in_arr).sum(dim=(0, 2, 3)))


def concat(in_arr1, in_arr2, out_arr, axis=0, stream=None):
    return _store(out_arr, torch.cat([_t(in_arr1), _t(in_arr2)], dim=int(axis)))


def concat_gradient(out_grad_arr, in_arr, axis=0, idx=0, stream=None):
    """Slice of the concatenated gradient that belongs to input ``idx`` (0 = first)."""
    g, x = _t(out_grad_arr), _t(in_arr)
    n = x.shape[axis]
    start = 0 if idx == 0 else g.shape[axis] - n
    return _store(in_arr, g.narrow(int(axis), start, n))


def concatenate(in_arrs, out_arr, axis=0, stream=None):
    return _store(out_arr, torch.cat([_t(a) for a in in_arrs], dim=int(axis)))


def concatenate_gradient(out_grad_arr, in_arr, axis, offset, stream=None):
    x = _t(in_arr)
    return _store(in_arr, _t(out_grad_arr).narrow(int(axis), int(offset), x.shape[axis]))


def matrix_slice(in_arr, out_arr, begin_pos, stream=None):
    x, o = _t(in_arr), _t(out_arr)
    idx = tuple(slice(int(b), int(b) + s) for b, s in zip(begin_pos, o.shape))
    return _store(out_arr, x[idx])


def matrix_slice_gradient(in_arr, out_arr, begin_pos, stream=None):
    g, o = _t(in_arr), _t(out_arr)
    o.zero_()
    idx = tuple(slice(int(b), int(b) + s) for b, s in zip(begin_pos, g.shape))
    o[idx] = g.to(o.dtype)
    return out_arr


def matrix_slice_simple(in_arr, out_arr, gpu_buf, stream=None):
    return matrix_slice(in_arr, out_arr, [int(v) for v in np.asarray(_t(gpu_buf).cpu()).reshape(-1)[:_t(out_arr).dim()]])


def matrix_slice_gradient_simple(in_arr, out_arr, gpu_buf, stream=None):
    return matrix_slice_gradient(in_arr, out_arr, [int(v) for v in np.asarray(_t(gpu_buf).cpu()).reshape(-1)[:_t(out_arr).dim()]])


def slice_assign(in_arr, out_arr, val, begin_pos, end_pos, stream=None):
    x = _t(in_arr).clone()
    x[tuple(slice(int(b), int(e)) for b, e in zip(begin_pos, end_pos))] = val
    return _store(out_arr, x)


def slice_assign_matrix(arr_A, arr_B, out_arr, begin_pos_A, end_pos_A, begin_pos_B, stream=None):
    a, b = _t(arr_A).clone(), _t(arr_B)
    sa = tuple(slice(int(s), int(e)) for s, e in zip(begin_pos_A, end_pos_A))
    sb = tuple(slice(int(s), int(s) + (int(e) - int(s0))) for s, s0, e in zip(begin_pos_B, begin_pos_A, end_pos_A))
    a[sa] = b[sb]
    return _store(out_arr, a)


def slice_by_matrix(in_arr, index1, index2, out_arr, stream=None):
    return _store(out_arr, _t(in_arr)[_t(index1).long(), _t(index2).long()])


def slice_by_matrix_gradient(in_arr, index1, index2, out_arr, stream=None):
    o = _t(out_arr)
    o.zero_()
    o.index_put_((_t(index1).long(), _t(index2).long()), _t(in_arr).to(o.dtype), accumulate=True)
    return out_arr


def gather(in_arr, index, out_arr, dim, stream=None):
    return _store(out_arr, torch.gather(_t(in_arr), int(dim), _t(index).long()))


def gather_gradient(in_arr, index, out_arr, dim, stream=None):
    o = _t(out_arr)
    o.zero_()
    o.scatter_add_(int(dim), _t(index).long(), _t(in_arr).to(o.dtype))
    return out_arr


def scatter(target_mat, dim, index_mat, src_mat, stream=None):
    _t(target_mat).scatter_(int(dim), _t(index_mat).long(), _t(src_mat).to(_t(target_mat).dtype))
    return target_mat


def scatter1d(input_mat, index_mat, output_mat, stream=None):
    o = _t(output_mat)
    o[_t(index_mat).long().reshape(-1)] = _t(input_mat).to(o.dtype)
    return output_mat


def scatter1d_grad(output_grad_mat, index_mat, input_grad_mat, stream=None):
    return _store(input_grad_mat, _t(output_grad_mat)[_t(index_mat).long().reshape(-1)])


def indexing(input_mat, index_mat, output_mat, stream=None):
    return _store(output_mat, _t(input_mat)[_t(index_mat).long().reshape(-1)])


def indexing_grad(output_grad, index, input_grad, stream=None):
    o = _t(input_grad)
    o.zero_()
    o.index_add_(0, _t(index).long().reshape(-1), _t(output_grad).to(o.dtype))
    return input_grad


def repeat(in_arr, out_arr, stream=None):
    x, o = _t(in_arr), _t(out_arr)
    x = x.reshape((1,) * (o.dim() - x.dim()) + tuple(x.shape))
    return _store(out_arr, x.repeat(*[os // xs for os, xs in zip(o.shape, x.shape)]))


def repeat_gradient(in_arr, out_arr, stream=None):
    g, o = _t(in_arr), _t(out_arr)
    shape = (1,) * (g.dim() - o.dim()) + tuple(o.shape)
    view = []
    for gs, os in zip(g.shape, shape):
        view += [gs // os, os]
    return _store(out_arr, g.reshape(view).sum(dim=tuple(range(0, 2 * len(shape), 2))))


def roll(input_mat, output_mat, shift, axis, stream=None):
    return _store(output_mat, torch.roll(_t(input_mat), shift, axis))


def pad(in_arr, out_arr, paddings, mode='CONSTANT', constant_values=0, stream=None):
    return _store(out_arr, _run(_ops.pad_op, [in_arr], paddings, mode, constant_values))


def pad_gradient(out_grad_arr, in_grad_arr, paddings, mode='CONSTANT', stream=None):
    return _store(in_grad_arr, _run(_ops.pad_gradient_op, [out_grad_arr], paddings, mode))


def bicubic_interpolate(in_arr, out_arr, align_corners, stream=None):
    o = _t(out_arr)
    return _store(out_arr, torch.nn.functional.interpolate(_t(in_arr), size=tuple(o.shape[-2:]), mode='bicubic',
                                                           align_corners=_b.bool(align_corners)))


def bicubic_interpolate_gradient(input_grad, output, align_corners, stream=None):
    o = _t(output)
    x = torch.zeros_like(o, requires_grad=True)
    y = torch.nn.functional.interpolate(x, size=tuple(_t(input_grad).shape[-2:]), mode='bicubic', align_corners=_b.bool(align_corners))
    (g,) = torch.autograd.grad(y, x, _t(input_grad).to(o.dtype))
    return _store(output, g)


# -------------------------------------------------------------------------------------- nn
def CuDNN_conv2d(in_arr_x, in_arr_f, out_arr, padding=(0, 0), stride=(1, 1), stream=None):
    return _store(out_arr, _run(_ops.conv2d_op, [in_arr_x, in_arr_f], _pair(padding), _pair(stride)))


def conv2d(in_arr_x, in_arr_f, out_arr, workspace_arr=None, padding=0, stride=1, stream=None):
    return CuDNN_conv2d(in_arr_x, in_arr_f, out_arr, padding, stride)


def CuDNN_conv2d_with_bias(in_arr_x, in_arr_f, bias, out_arr, padding=(0, 0), stride=(1, 1), stream=None):
    return _store(out_arr, _run(_ops.conv2d_add_bias_op, [in_arr_x, in_arr_f, bias], _pair(padding), _pair(stride)))


def CuDNN_conv2d_gradient_of_data(in_arr_f, in_gradient_y, out_gradient_x, padding=(0, 0), stride=(1, 1), stream=None):
    return _store(out_gradient_x, torch.nn.grad.conv2d_input(tuple(_t(out_gradient_x).shape), _t(in_arr_f), _t(in_gradient_y),
                                                             stride=_pair(stride), padding=_pair(padding))
                  if not _on_gpu(in_arr_f) else
                  _run(_ops.conv2d_gradient_of_data_op, [in_arr_f, in_gradient_y, out_gradient_x], _pair(padding), _pair(stride)))


def CuDNN_conv2d_gradient_of_filter(in_arr_x, in_gradient_y, out_gradient_f, padding=(0, 0), stride=(1, 1), stream=None):
    return _store(out_gradient_f, torch.nn.grad.conv2d_weight(_t(in_arr_x), tuple(_t(out_gradient_f).shape), _t(in_gradient_y),
                                                              stride=_pair(stride), padding=_pair(padding))
                  if not _on_gpu(in_arr_x) else
                  _run(_ops.conv2d_gradient_of_filter_op, [in_arr_x, in_gradient_y, out_gradient_f], _pair(padding), _pair(stride)))


def conv2d_gradient_of_data(in_arr_f, in_gradient_y, out_gradient_x, workspace_im2col=None, padding=0, stride=1, stream=None):
    return CuDNN_conv2d_gradient_of_data(in_arr_f, in_gradient_y, out_gradient_x, padding, stride)


def conv2d_gradient_of_filter(in_arr_x, in_gradient_y, out_gradient_f, workspace_im2col=None, workspace_batch_filter=None,
                              padding=0, stride=1, stream=None):
    return CuDNN_conv2d_gradient_of_filter(in_arr_x, in_gradient_y, out_gradient_f, padding, stride)


def _on_gpu(a):
    t = _t(a)
    return torch.is_tensor(t) and t.device.type == 'cuda'


def CuDNN_max_pooling2d(in_arr, kernel_H, kernel_W, pooled_layer, padding=0, stride=1, stream=None):
    return _store(pooled_layer, _run(_ops.max_pool2d_op, [in_arr], kernel_H, kernel_W, padding, stride))


max_pooling2d = CuDNN_max_pooling2d


def CuDNN_max_pooling2d_gradient(in_arr_y, in_gradient_y, in_arr_x, kernel_H, kernel_W, out_gradient_x, padding=0, stride=1,
                                 stream=None):
    x = _t(in_arr_x).detach().clone().requires_grad_(True)
    y = torch.nn.functional.max_pool2d(x, (kernel_H, kernel_W), stride=stride, padding=padding)
    (g,) = torch.autograd.grad(y, x, _t(in_gradient_y))
    return _store(out_gradient_x, g)


def max_pooling2d_gradient(in_arr, in_grad_arr, kernel_H, kernel_W, out_grad_arr, padding=0, stride=1, stream=None):
    return CuDNN_max_pooling2d_gradient(None, in_grad_arr, in_arr, kernel_H, kernel_W, out_grad_arr, padding, stride)


def CuDNN_average_pooling2d(in_arr, kernel_H, kernel_W, pooled_layer, padding=0, stride=1, stream=None):
    return _store(pooled_layer, _run(_ops.avg_pool2d_op, [in_arr], kernel_H, kernel_W, padding, stride))


average_pooling2d = CuDNN_average_pooling2d


def average_pooling2d_gradient(in_gradient_y, kernel_H, kernel_W, out_gradient_x, padding=0, stride=1, stream=None):
    x = torch.zeros_like(_t(out_gradient_x), requires_grad=True)
    y = torch.nn.functional.avg_pool2d(x, (kernel_H, kernel_W), stride=stride, padding=padding)
    (g,) = torch.autograd.grad(y, x, _t(in_gradient_y))
    return _store(out_gradient_x, g)


def CuDNN_average_pooling2d_gradient(in_arr_y, in_gradient_y, in_arr_x, kernel_H, kernel_W, out_gradient_x, padding=0, stride=1,
                                     stream=None):
    return average_pooling2d_gradient(in_gradient_y, kernel_H, kernel_W, out_gradient_x, padding, stride)


def CuDNN_Batch_Normalization(in_arr, bn_scale_arr, bn_bias_arr, out_arr, running_mean, running_var, save_mean, save_var,
                              momentum, eps, stream=None):
    """Training-mode spatial batch norm: writes the output, the batch statistics (``save_mean`` /
    ``save_var`` = biased variance) and updates the running statistics in place."""
    x = _t(in_arr)
    mean = x.mean(dim=(0, 2, 3))
    var = x.var(dim=(0, 2, 3), unbiased=False)
    y = (x - mean.view(1, -1, 1, 1)) / torch.sqrt(var.view(1, -1, 1, 1) + eps)
    _store(out_arr, y * _t(bn_scale_arr).view(1, -1, 1, 1) + _t(bn_bias_arr).view(1, -1, 1, 1))
    _store(save_mean, mean)
    _store(save_var, var)
    n = x.numel() / x.shape[1]
    rm, rv = _t(running_mean), _t(running_var)
    rm.mul_(1 - momentum).add_(momentum * mean.reshape(rm.shape))
    rv.mul_(1 - momentum).add_(momentum * (var * n / _b.max(n - 1, 1)).reshape(rv.shape))
    return out_arr


def CuDNN_Batch_Normalization_inference(in_arr, bn_scale_arr, bn_bias_arr, out_arr, estimated_mean, estimated_var, eps, stream=None):
    x = _t(in_arr)
    y = (x - _t(estimated_mean).view(1, -1, 1, 1)) / torch.sqrt(_t(estimated_var).view(1, -1, 1, 1) + eps)
    return _store(out_arr, y * _t(bn_scale_arr).view(1, -1, 1, 1) + _t(bn_bias_arr).view(1, -1, 1, 1))


def CuDNN_Batch_Normalization_gradient(in_gradient_y, in_arr_x, in_bn_scale, out_gradient_x, out_gradient_bn_scale,
                                       out_gradient_bn_bias, save_mean, save_var, eps, stream=None):
    g, x = _t(in_gradient_y), _t(in_arr_x)
    mean, var = _t(save_mean).view(1, -1, 1, 1), _t(save_var).view(1, -1, 1, 1)
    inv = torch.rsqrt(var + eps)
    xhat = (x - mean) * inv
    dbias = g.sum(dim=(0, 2, 3))
    dscale = (g * xhat).sum(dim=(0, 2, 3))
    m = x.numel() / x.shape[1]
    dx = _t(in_bn_scale).view(1, -1, 1, 1) * inv * (g - dbias.view(1, -1, 1, 1) / m - xhat * dscale.view(1, -1, 1, 1) / m)
    _store(out_gradient_bn_scale, dscale)
    _store(out_gradient_bn_bias, dbias)
    return _store(out_gradient_x, dx)


def layer_normalization(in_arr, ln_scale, ln_bias, mean, var, out_arr, eps, stream=None):
    x = _t(in_arr)
    mu = x.mean(dim=-1, keepdim=True)
    v = x.var(dim=-1, keepdim=True, unbiased=False)
    _store(mean, mu)
    _store(var, v)
    res = _run(_ops.layer_normalization_op, [in_arr, ln_scale, ln_bias], eps)
    return _store(out_arr, res[0] if isinstance(res, (tuple, list)) else res)


def layer_normalization_gradient(out_grads, in_arr, ln_scale, grad_arr, grad_scale, grad_bias, mean_arr, var_arr, eps, stream=None):
    g, x = _t(out_grads), _t(in_arr)
    mu, v = _t(mean_arr).reshape(*x.shape[:-1], 1), _t(var_arr).reshape(*x.shape[:-1], 1)
    inv = torch.rsqrt(v + eps)
    xhat = (x - mu) * inv
    red = tuple(range(x.dim() - 1))
    _store(grad_bias, g.sum(dim=red))
    _store(grad_scale, (g * xhat).sum(dim=red))
    gs = g * _t(ln_scale)
    dx = inv * (gs - gs.mean(dim=-1, keepdim=True) - xhat * (gs * xhat).mean(dim=-1, keepdim=True))
    return _store(grad_arr, dx)


def instance_normalization2d(in_arr, mean, var, out_arr, eps, stream=None):
    x = _t(in_arr)
    mu = x.mean(dim=(2, 3), keepdim=True)
    v = x.var(dim=(2, 3), keepdim=True, unbiased=False)
    _store(mean, mu)
    _store(var, v)
    return _store(out_arr, (x - mu) * torch.rsqrt(v + eps))


def instance_normalization2d_gradient(out_grads, in_arr, grad_arr, mean_arr, var_arr, eps, stream=None):
    g, x = _t(out_grads), _t(in_arr)
    mu, v = _t(mean_arr).reshape(x.shape[0], x.shape[1], 1, 1), _t(var_arr).reshape(x.shape[0], x.shape[1], 1, 1)
    inv = torch.rsqrt(v + eps)
    xhat = (x - mu) * inv
    dx = inv * (g - g.mean(dim=(2, 3), keepdim=True) - xhat * (g * xhat).mean(dim=(2, 3), keepdim=True))
    return _store(grad_arr, dx)


def softmax(in_arr, out_arr, stream=None):
    return _go(_ops.softmax_op, [in_arr], out_arr)


CuDNN_softmax = softmax


def CuDNN_softmax_gradient(y_arr, grad_arr, out_arr, stream=None):
    y, g = _t(y_arr), _t(grad_arr)
    return _store(out_arr, y * (g - (g * y).sum(dim=-1, keepdim=True)))


def softmax_cross_entropy(in_arr_a, in_arr_b, out_arr, stream=None):
    """out[i] = -sum_j b[i, j] * log softmax(a)[i, j]."""
    return _go(_ops.softmaxcrossentropy_op, [in_arr_a, in_arr_b], out_arr)


CuDNN_softmax_cross_entropy = softmax_cross_entropy


def softmax_cross_entropy_gradient(in_arr_a, in_arr_b, in_arr_c, out_arr, stream=None):
    """(softmax(a) - b) * c with the upstream gradient c per row."""
    a, b, c = _t(in_arr_a), _t(in_arr_b), _t(in_arr_c)
    return _store(out_arr, (torch.softmax(a, dim=-1) - b) * c.reshape(*a.shape[:-1], 1))


def CuDNN_softmax_cross_entropy_gradient(grad_arr, y_arr, label, out_arr, stream=None):
    return softmax_cross_entropy_gradient(y_arr, label, grad_arr, out_arr)


def softmax_cross_entropy_sparse(in_arr_a, in_arr_b, ignored_index, out_arr, stream=None):
    return _store(out_arr, _run(_ops.softmaxcrossentropy_sparse_op, [in_arr_a, in_arr_b], int(ignored_index)))


def softmax_cross_entropy_sparse_gradient(in_arr_a, in_arr_b, in_arr_c, ignored_index, out_arr, stream=None):
    a, lab, c = _t(in_arr_a), _t(in_arr_b).long().reshape(-1), _t(in_arr_c)
    p = torch.softmax(a.reshape(-1, a.shape[-1]), dim=-1)
    keep = lab != int(ignored_index)
    onehot = torch.zeros_like(p)
    onehot[keep, lab[keep]] = 1
    g = (p - onehot) * keep.unsqueeze(1).to(p.dtype) * c.reshape(-1, 1)
    return _store(out_arr, g)


def cross_entropy(y, y_, out, stream=None):
    return _go(_ops.crossentropy_op, [y, y_], out)


def cross_entropy_gradient(grad_arr, y_arr, label, out_arr, stream=None):
    y = _t(y_arr)
    return _store(out_arr, -_t(label) / y * _t(grad_arr).reshape(*y.shape[:-1], 1))


def cross_entropy_sparse(y, y_, ignored_index, out, stream=None):
    return _store(out, _run(_ops.crossentropy_sparse_op, [y, y_], int(ignored_index)))


def cross_entropy_sparse_gradient(grad_arr, y_arr, label, ignored_index, out_arr, stream=None):
    y, lab = _t(y_arr), _t(label).long().reshape(-1)
    y2 = y.reshape(-1, y.shape[-1])
    g = torch.zeros_like(y2)
    keep = lab != int(ignored_index)
    rows = torch.nonzero(keep).reshape(-1)
    g[rows, lab[keep]] = -1.0 / y2[rows, lab[keep]] * _t(grad_arr).reshape(-1)[rows]
    return _store(out_arr, g)


def binary_cross_entropy(in_arr_a, in_arr_b, out_arr, stream=None):
    return _go(_ops.binarycrossentropy_op, [in_arr_a, in_arr_b], out_arr)


def binary_cross_entropy_gradient(in_arr_a, in_arr_b, in_arr_c, out_arr, stream=None):
    p, y = _t(in_arr_a), _t(in_arr_b)
    return _store(out_arr, (-(y / p) + (1 - y) / (1 - p)) * _t(in_arr_c))


def nll_loss_link(input, target, output, stream=None):
    x, t = _t(input), _t(target).long().reshape(-1)
    return _store(output, -x[torch.arange(x.shape[0], device=x.device), t].mean())


def nll_loss_grad_link(output_grad, target, input_grad, stream=None):
    o, t = _t(input_grad), _t(target).long().reshape(-1)
    o.zero_()
    o[torch.arange(o.shape[0], device=o.device), t] = -_t(output_grad).reshape(-1)[0].to(o.dtype) / o.shape[0]
    return input_grad


def dropout(in_arr, dropout_rate, out_arr, seed, stream=None):
    """Keep probability ``1 - dropout_rate``; the mask is a pure function of ``seed`` so the gradient
    can regenerate it (reference Dropout.cu)."""
    x = _t(in_arr)
    gen = torch.Generator(device='cpu').manual_seed(int(_seed_value(seed)))
    mask = (torch.rand(x.shape, generator=gen) >= dropout_rate).to(device=x.device, dtype=x.dtype)
    return _store(out_arr, x * mask / (1 - dropout_rate))


def dropout_gradient_recompute(grad_arr, dropout_rate, out_arr, seed, stream=None):
    return dropout(grad_arr, dropout_rate, out_arr, seed)


def dropout_gradient(grad_arr, fw_output_arr, dropout_rate, out_arr, stream=None):
    g = _t(grad_arr)
    return _store(out_arr, g * (_t(fw_output_arr) != 0).to(g.dtype) / (1 - dropout_rate))


def dropout2d(in_arr, dropout2d_rate, out_arr, seed, stream=None):
    x = _t(in_arr)
    gen = torch.Generator(device='cpu').manual_seed(int(_seed_value(seed)))
    mask = (torch.rand(x.shape[:2], generator=gen) >= dropout2d_rate).to(device=x.device, dtype=x.dtype)
    return _store(out_arr, x * mask.reshape(*x.shape[:2], 1, 1) / (1 - dropout2d_rate))


def dropout2d_gradient(grad_arr, dropout2d_rate, out_arr, seed, stream=None):
    return dropout2d(grad_arr, dropout2d_rate, out_arr, seed)


def _seed_value(seed):
    return seed.value if hasattr(seed, 'value') else seed


def CuDNN_Dropout(in_arr, dropout, out_arr, reserve_size=None, reserve_space=None, firstTime=True, stream=None):
    return globals()['dropout'](in_arr, dropout, out_arr, 0)


def CuDNN_Dropout_gradient(in_gradient_y, dropout, out_gradient_x, reserve_size=None, reserve_space=None, stream=None):
    return globals()['dropout'](in_gradient_y, dropout, out_gradient_x, 0)


def embedding_lookup(in_mat, ids, out_mat, stream=None):
    return _store(out_mat, _run(_ops.embedding_lookup_op, [in_mat, ids]))


def embedding_lookup_gradient(grad_out, ids, grad_in, stream=None):
    """Dense scatter-add of the row gradients into the (zeroed) table gradient."""
    o = _t(grad_in)
    o.zero_()
    g = _t(grad_out)
    o.index_add_(0, _t(ids).long().reshape(-1), g.reshape(-1, g.shape[-1]).to(o.dtype))
    return grad_in


def indexedslice_oneside_add(indslice, output, stream=None):
    o = _t(output)
    idx, val = _t(indslice.indices).long().reshape(-1), _t(indslice.values)
    o.index_add_(0, idx, val.reshape(-1, val.shape[-1]).to(o.dtype))
    return output


# -------------------------------------------------------------------------------- initialisers
def _init(cls, arr, a, b, seed):
    t = _t(arr)
    init = cls(a, b, tuple(t.shape))
    init.seed = int(_seed_value(seed))
    rng = np.random.RandomState(init.seed & 0x7fffffff)
    return _store(arr, init._draw(init.shape, t.device, rng))


def normal_init(arr, mean, stddev, seed, stream=None):
    from .initializers import NormalInit
    return _init(NormalInit, arr, mean, stddev, seed)


def uniform_init(arr, lb, ub, seed, stream=None):
    from .initializers import UniformInit
    return _init(UniformInit, arr, lb, ub, seed)


def truncated_normal_init(arr, mean, stddev, seed, stream=None):
    from .initializers import TruncatedNormalInit
    return _init(TruncatedNormalInit, arr, mean, stddev, seed)


# ---------------------------------------------------------------------------------- optimizers
def add_l2_regularization(param, grad, l2reg, stream=None):
    _t(grad).add_(_t(param), alpha=l2reg)
    return grad


def sgd_update(param, grad, lr, l2reg, stream=None):
    if isinstance(grad, IndexedSlices):
        p = _t(param)
        idx, val = _t(grad.indices).long().reshape(-1), _t(grad.values)
        p.index_add_(0, idx, val.reshape(-1, val.shape[-1]).to(p.dtype), alpha=-lr)
        return param
    p, g = _t(param), _t(grad)
    p.add_(g + l2reg * p if l2reg else g, alpha=-lr)
    return param


def momentum_update(param, grad, velocity, lr, momentum, nesterov, l2reg, stream=None):
    p, g, v = _t(param), _t(grad), _t(velocity)
    g = g + l2reg * p if l2reg else g
    v.mul_(momentum).add_(g, alpha=-lr)
    if nesterov:
        p.add_(momentum * v - lr * g)
    else:
        p.add_(v)
    return param


def adagrad_update(param, grad, accumulation, lr, eps, l2reg, stream=None):
    p, g, a = _t(param), _t(grad), _t(accumulation)
    g = g + l2reg * p if l2reg else g
    a.add_(g * g)
    p.add_(-lr * g / (torch.sqrt(a) + eps))
    return param


def adam_update(param, grad, expavg, expavgsq, lr, beta1, beta2, beta1t, beta2t, eps, l2reg, stream=None):
    p, g, m, v = _t(param), _t(grad), _t(expavg), _t(expavgsq)
    g = g + l2reg * p if l2reg else g
    m.mul_(beta1).add_(g, alpha=1 - beta1)
    v.mul_(beta2).add_(g * g, alpha=1 - beta2)
    p.add_(-lr * (m / (1 - beta1t)) / (torch.sqrt(v / (1 - beta2t)) + eps))
    return param


def adamw_update(param, grad, expavg, expavgsq, lr, beta1, beta2, beta1t, beta2t, eps, weight_decay, stream=None):
    p, g, m, v = _t(param), _t(grad), _t(expavg), _t(expavgsq)
    m.mul_(beta1).add_(g, alpha=1 - beta1)
    v.mul_(beta2).add_(g * g, alpha=1 - beta2)
    p.add_(-lr * ((m / (1 - beta1t)) / (torch.sqrt(v / (1 - beta2t)) + eps) + weight_decay * p))
    return param


def lamb_update(param, grad, expavg, expavgsq, lr, beta1, beta2, beta1t, beta2t, eps, weight_decay, stream=None):
    p, g, m, v = _t(param), _t(grad), _t(expavg), _t(expavgsq)
    m.mul_(beta1).add_(g, alpha=1 - beta1)
    v.mul_(beta2).add_(g * g, alpha=1 - beta2)
    upd = (m / (1 - beta1t)) / (torch.sqrt(v / (1 - beta2t)) + eps) + weight_decay * p
    pn, un = torch.linalg.vector_norm(p), torch.linalg.vector_norm(upd)
    trust = torch.where((pn > 0) & (un > 0), pn / un, torch.ones_like(pn))
    p.add_(-lr * trust * upd)
    return param


# ----------------------------------------------------------------------------------------- MoE
def layout_transform_top1(input, indices_s, location_s, output, capacity, stream=None):
    """Dispatch: out[idx * capacity + loc] = in[token] for tokens whose slot is inside the capacity."""
    x, o = _t(input), _t(output)
    idx, loc = _t(indices_s).long().reshape(-1), _t(location_s).long().reshape(-1)
    o.zero_()
    flat = o.reshape(-1, x.shape[-1])
    ok = (loc < capacity) & (loc >= 0)
    flat[idx[ok] * capacity + loc[ok]] = x.reshape(-1, x.shape[-1])[ok].to(o.dtype)
    return output


def layout_transform_top2(input, indices_s1, indices_s2, location_s1, location_s2, output, capacity, stream=None):
    x, o = _t(input), _t(output)
    o.zero_()
    flat = o.reshape(-1, x.shape[-1])
    for ind, locs in ((indices_s1, location_s1), (indices_s2, location_s2)):
        idx, loc = _t(ind).long().reshape(-1), _t(locs).long().reshape(-1)
        ok = (loc < capacity) & (loc >= 0)
        flat[idx[ok] * capacity + loc[ok]] = x.reshape(-1, x.shape[-1])[ok].to(o.dtype)
    return output


def layout_transform_top1_gradient(input, indice, location, output, capacity, stream=None):
    g, o = _t(input), _t(output)
    idx, loc = _t(indice).long().reshape(-1), _t(location).long().reshape(-1)
    o.zero_()
    ok = (loc < capacity) & (loc >= 0)
    flat = g.reshape(-1, g.shape[-1])
    o.reshape(-1, g.shape[-1])[ok] = flat[idx[ok] * capacity + loc[ok]].to(o.dtype)
    return output


def reverse_layout_transform_top1(input, indices_s, location_s, gates, output, capacity, stream=None):
    """Combine: out[token] = gate[token] * expert_out[idx * capacity + loc]."""
    layout_transform_top1_gradient(input, indices_s, location_s, output, capacity)
    o = _t(output)
    o.mul_(_t(gates).reshape(-1, 1).to(o.dtype))
    return output


def reverse_layout_transform_top2(input, indices_s1, indices_s2, location_s1, location_s2, gates_1, gates_2, output, capacity,
                                  stream=None):
    o = _t(output)
    tmp = torch.empty_like(o)
    reverse_layout_transform_top1(input, indices_s1, location_s1, gates_1, output, capacity)
    reverse_layout_transform_top1(input, indices_s2, location_s2, gates_2, tmp, capacity)
    o.add_(tmp)
    return output


def reverse_layout_transform_no_gate(input, indices_s, location_s, output, capacity, stream=None):
    return layout_transform_top1_gradient(input, indices_s, location_s, output, capacity)


def reverse_layout_transform_no_gate_gradient(input, indice, location, output, capacity, stream=None):
    return layout_transform_top1(input, indice, location, output, capacity)


def reverse_layout_transform_top1_gradient_data(input, indice, location, gate, output, capacity, stream=None):
    g = _t(input) * _t(gate).reshape(-1, 1)
    return layout_transform_top1(g, indice, location, output, capacity)


def reverse_layout_transform_top2_gradient_data(input, indice_1, indice_2, location_1, location_2, gate_1, gate_2, output, capacity,
                                                stream=None):
    x, o = _t(input), _t(output)
    o.zero_()
    flat = o.reshape(-1, x.shape[-1])
    for ind, locs, gate in ((indice_1, location_1, gate_1), (indice_2, location_2, gate_2)):
        idx, loc = _t(ind).long().reshape(-1), _t(locs).long().reshape(-1)
        ok = (loc < capacity) & (loc >= 0)
        flat[idx[ok] * capacity + loc[ok]] = (x * _t(gate).reshape(-1, 1))[ok].to(o.dtype)
    return output


def reverse_layout_transform_top1_gradient_gate(combined_output, expert_output, indice, location, output, capacity, stream=None):
    """d gate[token] = <upstream grad of the combined row, expert output row of that token>."""
    g, e = _t(combined_output), _t(expert_output)
    idx, loc = _t(indice).long().reshape(-1), _t(location).long().reshape(-1)
    ok = (loc < capacity) & (loc >= 0)
    rows = torch.zeros_like(g)
    rows[ok] = e.reshape(-1, e.shape[-1])[idx[ok] * capacity + loc[ok]].to(g.dtype)
    return _store(output, (g * rows).sum(dim=-1))


def ha2a_layout_transform(input, output, num_nodes, num_local_gpus, stream=None):
    """[gpu][node][gpu'][rows] -> [node][gpu'][gpu][rows] row-block permutation on the leader GPU."""
    x = _t(input)
    G, N = int(num_local_gpus), int(num_nodes)
    blk = x.reshape(G, N, G, -1, x.shape[-1])
    return _store(output, blk.permute(1, 2, 0, 3, 4).reshape(x.shape))


def ha2a_reverse_layout_transform(input, output, num_nodes, num_local_gpus, stream=None):
    x = _t(input)
    G, N = int(num_local_gpus), int(num_nodes)
    blk = x.reshape(N, G, G, -1, x.shape[-1])
    return _store(output, blk.permute(2, 0, 1, 3, 4).reshape(x.shape))


def sam_group_sum_link(gate_mat, out_mat, num_local_gpus, stream=None):
    return _store(out_mat, _run(_ops.sam_group_sum_op, [gate_mat], int(num_local_gpus)))


def group_topk_idx(in_mat, top1_group, out_mat_idx, k, num_local_gpus, stream=None):
    return _store(out_mat_idx, _run(_ops.group_topk_idx_op, [in_mat, top1_group], int(k), int(num_local_gpus)))


def sammax_link(in_mat, top1_group, topk_indice, out_mat, num_local_gpus, stream=None):
    return _store(out_mat, _run(_ops.sam_max_op, [in_mat, top1_group, topk_indice], int(num_local_gpus)))


def sammax_grad_link(output_grad, in_mat, top1_group, topk_indice, out_mat, num_local_gpus, stream=None):
    """Gradient of sammax: passes where the forward value was kept (non-zero)."""
    tmp = torch.empty_like(_t(in_mat))
    sammax_link(in_mat, top1_group, topk_indice, tmp, num_local_gpus)
    return _store(out_mat, _t(output_grad) * (tmp != 0).to(_t(output_grad).dtype))
