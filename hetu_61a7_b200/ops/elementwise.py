"""Element-wise and scalar ops.

reference op files (python/hetu/gpu_ops/): Abs AddConst AddElewise Bool Clamp
ConstPow Division Exp Floor Gelu LeakyRelu LogElewise MaskedFill MinusByConst
MinusElewise MultiplyConst MultiplyElewise Opposite Pow Relu Sigmoid Sin Sqrt
Tanh Where OnesLike ZerosLike Full Rand Arange; kernels src/ops/<same>.cu.

On CUDA tensors every op below is one launch of the vectorised functor kernels
in csrc/elementwise.cu (``hetu_unary`` / ``hetu_binary`` / ``hetu_ternary``,
float4 loads, grid-stride over 148xk CTAs); on CPU/meta tensors the torch
expression is the oracle.
"""
from __future__ import annotations

import math

import numpy as np
import torch

from .. import _lib as L
from ..graph.node import Op, as_index

# opcode tables - must match csrc/elementwise.cu
U = dict(ABS=0, ADD_CONST=1, MUL_CONST=2, EXP=3, FLOOR=4, GELU=5, LEAKY_RELU=6, LOG=7,
         MINUS_BY_CONST=8, OPPOSITE=9, POW=10, CONST_POW=11, RELU=12, SIGMOID=13, SIN=14,
         COS=15, SQRT=16, RSQRT=17, TANH=18, DIV_CONST=19, CLAMP=20, BOOL=21, BOOL_EQ=22,
         BOOL_LT=23, BOOL_GT=24, BOOL_LE=25, BOOL_GE=26, FILL=27, COPY=28, SIGN=29,
         CLAMP_MIN=30, CLAMP_MAX=31)
B = dict(ADD=0, SUB=1, MUL=2, DIV=3, RELU_GRAD=4, GELU_GRAD=5, LEAKY_RELU_GRAD=6, TANH_GRAD=7,
         ABS_GRAD=8, POW_GRAD=9, CONST_POW_GRAD=10, LOG_GRAD=11, EQ=12, LT=13, GT=14, LE=15,
         GE=16, MASKED_FILL=17, WHERE_CONST=18, MAX=19, MIN=20, SIGMOID_GRAD=21)


def _unary_kernel(code, x, a=0.0, b=0.0):
    x = x.contiguous()
    y = torch.empty_like(x)
    L.call('unary', L.c_int(code), L.p(x), L.p(y), L.c_sz(x.numel()), L.c_f32(a), L.c_f32(b),
           L.stream_ptr(x.device))
    return y


def _binary_kernel(code, x, y, a=0.0):
    x = x.contiguous()
    y = y.contiguous()
    out = torch.empty_like(x)
    L.call('binary', L.c_int(code), L.p(x), L.p(y), L.p(out), L.c_sz(x.numel()), L.c_f32(a),
           L.stream_ptr(x.device))
    return out


def _bcast_kernel(code, x, y):
    """x [outer, inner] (+,-,*,/) y broadcast along rows (y [inner]) or cols (y [outer,1])."""
    out = torch.empty_like(x)
    inner = y.numel() if y.dim() == 1 or y.shape[-1] != 1 else 1
    mode = 0 if (y.dim() >= 1 and y.shape[-1] == x.shape[-1] and y.numel() == x.shape[-1]) else 1
    if mode == 0:
        L.call('binary_bcast', L.c_int(code), L.p(x), L.p(y), L.p(out), L.c_sz(x.numel()),
               L.c_sz(x.shape[-1]), L.c_int(0), L.stream_ptr(x.device))
    else:
        L.call('binary_bcast', L.c_int(code), L.p(x), L.p(y), L.p(out), L.c_sz(x.numel()),
               L.c_sz(x.numel() // y.numel()), L.c_int(1), L.stream_ptr(x.device))
    return out


_F32 = torch.float32


def _k(*ts):
    return L.kernels_on(*ts) and all(t.dtype == _F32 for t in ts)


#: GPU opcode -> CPU library opcode for the activations the CPU executor runs natively
_CPU_UNARY = {U['RELU']: 0, U['SIGMOID']: 1, U['TANH']: 2, U['SQRT']: 4, U['RSQRT']: 5, U['OPPOSITE']: 6}
_CPU_BINARY = {B['RELU_GRAD']: 4}
_cpu_state = {}


def _cpu_native() -> bool:
    if 'on' not in _cpu_state:
        import os
        from .. import cpu_links
        _cpu_state['on'] = os.environ.get('HETU_CPU_NATIVE', '1') == '1' and cpu_links.available()
    return _cpu_state['on']


# --------------------------------------------------------------------- unary
class UnaryOp(Op):
    """y = f(x; a, b).  Sub-classes set ``code`` / ``fn`` / ``grad``."""
    code = None

    def __init__(self, node, a=0.0, b=0.0, ctx=None):
        super().__init__(type(self), [node], ctx)
        self.a, self.b = a, b

    def fn(self, x):
        raise NotImplementedError

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        if _k(x):
            return _unary_kernel(self.code, x, float(self.a), float(self.b))
        cpu_code = _CPU_UNARY.get(self.code)
        if cpu_code is not None and x.device.type == 'cpu' and x.dtype == _F32 and _cpu_native():
            from .. import cpu_links
            x = x.contiguous()
            y = torch.empty_like(x)
            cpu_links._unary(cpu_code, x, y)
            return y
        return self.fn(x)

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


def _unary(name, code, fn, grad=None):
    def gradient(self, g):
        return [grad(self, g)] if grad is not None else [None]
    cls = type(name, (UnaryOp,), {'code': code, 'fn': lambda self, x: fn(self, x),
                                  'gradient': gradient})
    return cls


class BinaryOp(Op):
    """out = f(x, y; a) for same-shape operands."""
    code = None

    def __init__(self, node_A, node_B, a=0.0, ctx=None):
        super().__init__(type(self), [node_A, node_B], ctx)
        self.a = a

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x, y = input_vals
        if _k(x, y) and x.shape == y.shape:
            return _binary_kernel(self.code, x, y, float(self.a))
        cpu_code = _CPU_BINARY.get(self.code)
        if cpu_code is not None and x.device.type == 'cpu' and x.shape == y.shape and x.dtype == _F32 \
                and y.dtype == _F32 and _cpu_native():
            from .. import cpu_links
            x, y = x.contiguous(), y.contiguous()
            out = torch.empty_like(x)
            cpu_links._binary(cpu_code, x, y, out)
            return out
        return self.fn(x, y)

    def infer_shape(self, input_shapes):
        a, b = input_shapes
        return tuple(a) if len(a) >= len(b) else tuple(b)


_SQRT1_2 = 0.7071067811865476

AbsOp = _unary('AbsOp', U['ABS'], lambda s, x: x.abs(),
               lambda s, g: abs_gradient_op(s.inputs[0], g, ctx=s.raw_ctx))
AddByConstOp = _unary('AddByConstOp', U['ADD_CONST'], lambda s, x: x + s.a, lambda s, g: g)
MulByConstOp = _unary('MulByConstOp', U['MUL_CONST'], lambda s, x: x * s.a,
                      lambda s, g: mul_byconst_op(g, s.a, ctx=s.raw_ctx))
ExpOp = _unary('ExpOp', U['EXP'], lambda s, x: x.exp(),
               lambda s, g: mul_op(s, g, ctx=s.raw_ctx))
FloorOp = _unary('FloorOp', U['FLOOR'], lambda s, x: x.floor())
GeluOp = _unary('GeluOp', U['GELU'], lambda s, x: x * 0.5 * (1.0 + torch.erf(x * _SQRT1_2)),
                lambda s, g: gelu_gradient_op(s.inputs[0], g, ctx=s.raw_ctx))
LeakyReluOp = _unary('LeakyReluOp', U['LEAKY_RELU'],
                     lambda s, x: torch.where(x > 0, x, x * s.a),
                     lambda s, g: leaky_relu_gradient_op(s.inputs[0], g, s.a, ctx=s.raw_ctx))
LogOp = _unary('LogOp', U['LOG'], lambda s, x: x.log(),
               lambda s, g: log_grad_op(g, s.inputs[0], ctx=s.raw_ctx))
MinusByConstOp = _unary('MinusByConstOp', U['MINUS_BY_CONST'], lambda s, x: s.a - x,
                        lambda s, g: opposite_op(g, ctx=s.raw_ctx))
OppositeOp = _unary('OppositeOp', U['OPPOSITE'], lambda s, x: -x,
                    lambda s, g: opposite_op(g, ctx=s.raw_ctx))
PowOp = _unary('PowOp', U['POW'], lambda s, x: x.pow(s.a),
               lambda s, g: pow_gradient_op(s.inputs[0], g, s.a, ctx=s.raw_ctx))
ConstPowOp = _unary('ConstPowOp', U['CONST_POW'],
                    lambda s, x: torch.pow(torch.tensor(float(s.a), device=x.device), x)
                    if x.device.type != 'meta' else torch.empty_like(x),
                    lambda s, g: const_pow_gradient_op(s, g, s.a, ctx=s.raw_ctx))
ReluOp = _unary('ReluOp', U['RELU'], lambda s, x: x.relu(),
                lambda s, g: relu_gradient_op(s.inputs[0], g, ctx=s.raw_ctx))
SigmoidOp = _unary('SigmoidOp', U['SIGMOID'], lambda s, x: x.sigmoid(),
                   lambda s, g: sigmoid_gradient_op(s, g, ctx=s.raw_ctx))
SinOp = _unary('SinOp', U['SIN'], lambda s, x: x.sin(),
               lambda s, g: mul_op(cos_op(s.inputs[0], ctx=s.raw_ctx), g, ctx=s.raw_ctx))
CosOp = _unary('CosOp', U['COS'], lambda s, x: x.cos(),
               lambda s, g: mul_op(mul_byconst_op(sin_op(s.inputs[0], ctx=s.raw_ctx), -1.0,
                                                  ctx=s.raw_ctx), g, ctx=s.raw_ctx))
SqrtOp = _unary('SqrtOp', U['SQRT'], lambda s, x: x.sqrt(),
                lambda s, g: mul_op(mul_byconst_op(rsqrt_op(s.inputs[0], ctx=s.raw_ctx), 0.5,
                                                   ctx=s.raw_ctx), g, ctx=s.raw_ctx))
ReciprocalSqrtOp = _unary(
    'ReciprocalSqrtOp', U['RSQRT'], lambda s, x: x.rsqrt(),
    lambda s, g: mul_op(mul_byconst_op(div_op(s, s.inputs[0], ctx=s.raw_ctx), -0.5,
                                       ctx=s.raw_ctx), g, ctx=s.raw_ctx))
TanhOp = _unary('TanhOp', U['TANH'], lambda s, x: x.tanh(),
                lambda s, g: tanh_gradient_op(s, g, ctx=s.raw_ctx))
DivConstOp = _unary(
    'DivConstOp', U['DIV_CONST'], lambda s, x: s.a / x,
    lambda s, g: mul_op(div_op(opposite_op(s, ctx=s.raw_ctx), s.inputs[0], ctx=s.raw_ctx), g,
                        ctx=s.raw_ctx))
BoolOp = _unary('BoolOp', U['BOOL'], lambda s, x: (x > 0).to(x.dtype))
SignOp = _unary('SignOp', U['SIGN'], lambda s, x: x.sign())


class BoolValOp(UnaryOp):
    """cond: 0 ==, 1 <, 2 >, 3 <=, 4 >= against a constant (1e-6 tolerance like
    src/ops/Bool.cu)."""

    def __init__(self, node, val, cond=0, ctx=None):
        super().__init__(node, float(val), 0.0, ctx)
        self.val, self.cond = val, cond
        self.code = U['BOOL_EQ'] + cond

    def fn(self, x):
        d = x - self.val
        r = [d.abs() < 1e-6, d < -1e-6, d > 1e-6, d < 1e-6, d > -1e-6][self.cond]
        return r.to(x.dtype)

    def gradient(self, g):
        return [None]


class BoolMatrixOp(BinaryOp):
    def __init__(self, node_A, node_B, cond, ctx=None):
        super().__init__(node_A, node_B, 0.0, ctx)
        self.cond = cond
        self.code = B['EQ'] + cond

    def fn(self, x, y):
        if x.dim() == 1 and y.dim() == 2:
            x = x.unsqueeze(0).expand(y.shape[0], -1)
            y = y.expand(-1, x.shape[1])
        d = x - y
        r = [d.abs() < 1e-6, d < -1e-6, d > 1e-6, d < 1e-6, d > -1e-6][self.cond]
        return r.to(x.dtype)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x, y = input_vals
        if x.dim() == 1 and y.dim() == 2:
            return self.fn(x, y)
        return super().compute(input_vals)

    def infer_shape(self, input_shapes):
        a, b = input_shapes
        if len(a) == len(b):
            return tuple(a)
        return (a[0], a[0])

    def gradient(self, g):
        return [None, None]


class ClampOp(Op):
    def __init__(self, node_A, node_B=None, node_C=None, mmin=None, mmax=None, ctx=None):
        inputs = [node_A] + [n for n in (node_B, node_C) if n is not None]
        super().__init__(ClampOp, inputs, ctx)
        self.mmin, self.mmax = mmin, mmax
        self.min_mat, self.max_mat = node_B is not None, node_C is not None

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        i = 1
        lo = hi = None
        if self.min_mat:
            lo = input_vals[i]
            i += 1
        if self.max_mat:
            hi = input_vals[i]
        if lo is None and hi is None and _k(x):
            if self.mmin is not None and self.mmax is not None:
                return _unary_kernel(U['CLAMP'], x, float(self.mmin), float(self.mmax))
            if self.mmin is not None:
                return _unary_kernel(U['CLAMP_MIN'], x, float(self.mmin))
            if self.mmax is not None:
                return _unary_kernel(U['CLAMP_MAX'], x, float(self.mmax))
        y = x
        if lo is not None:
            y = torch.maximum(y, lo) if not (_k(y, lo) and y.shape == lo.shape) \
                else _binary_kernel(B['MAX'], y, lo)
        elif self.mmin is not None:
            y = y.clamp(min=self.mmin)
        if hi is not None:
            y = torch.minimum(y, hi) if not (_k(y, hi) and y.shape == hi.shape) \
                else _binary_kernel(B['MIN'], y, hi)
        elif self.mmax is not None:
            y = y.clamp(max=self.mmax)
        return y

    def gradient(self, g):
        raise NotImplementedError

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


# -------------------------------------------------------------------- binary
def _bshape(a, b):
    return tuple(torch.broadcast_shapes(tuple(a), tuple(b)))


class _ArithOp(BinaryOp):
    """+,-,*,/ with numpy broadcasting; the row/column-broadcast cases (bias
    add, per-row scale) have their own kernel variant."""
    tfn = None

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x, y = input_vals
        if _k(x, y):
            if x.shape == y.shape:
                return _binary_kernel(self.code, x, y)
            if x.dim() >= 1 and y.dim() >= 1 and x.is_contiguous() and y.is_contiguous():
                if y.numel() == x.shape[-1] and y.shape[-1] == x.shape[-1] and x.numel() > y.numel():
                    return _bcast_kernel(self.code, x, y)
                if (y.dim() == x.dim() and y.shape[-1] == 1 and x.shape[-1] > 1
                        and y.shape[:-1] == x.shape[:-1]):
                    return _bcast_kernel(self.code, x, y)
        return type(self).tfn(x, y)

    def infer_shape(self, input_shapes):
        return _bshape(*input_shapes)


def _unbroadcast(grad_node, like_node, ctx):
    """Reduce a broadcast gradient back to ``like_node``'s shape at run time."""
    from .reduce import reduce_to_shape_op
    return reduce_to_shape_op(grad_node, like_node, ctx=ctx)


class AddOp(_ArithOp):
    code = B['ADD']
    tfn = staticmethod(torch.add)

    def gradient(self, g):
        return [_unbroadcast(g, self.inputs[0], self.raw_ctx),
                _unbroadcast(g, self.inputs[1], self.raw_ctx)]


class MinusOp(_ArithOp):
    code = B['SUB']
    tfn = staticmethod(torch.sub)

    def gradient(self, g):
        return [_unbroadcast(g, self.inputs[0], self.raw_ctx),
                _unbroadcast(opposite_op(g, ctx=self.raw_ctx), self.inputs[1], self.raw_ctx)]


class MulOp(_ArithOp):
    code = B['MUL']
    tfn = staticmethod(torch.mul)

    def gradient(self, g):
        a, b = self.inputs
        return [_unbroadcast(mul_op(b, g, ctx=self.raw_ctx), a, self.raw_ctx),
                _unbroadcast(mul_op(a, g, ctx=self.raw_ctx), b, self.raw_ctx)]


class DivOp(_ArithOp):
    code = B['DIV']
    tfn = staticmethod(torch.div)

    def gradient(self, g):
        a, b = self.inputs
        ga = div_op(g, b, ctx=self.raw_ctx)
        gb = opposite_op(mul_op(div_op(self, b, ctx=self.raw_ctx), g, ctx=self.raw_ctx),
                         ctx=self.raw_ctx)
        return [_unbroadcast(ga, a, self.raw_ctx), _unbroadcast(gb, b, self.raw_ctx)]


def _bin(name, code, fn):
    def gradient(self, g):
        raise NotImplementedError(name + '.gradient (second-order autodiff is not supported)')
    return type(name, (BinaryOp,), {'code': code, 'fn': lambda self, x, y: fn(self, x, y),
                                    'gradient': gradient})


_INV_SQRT_2PI = 0.3989422804014327
# grad ops take (input_or_output, upstream_grad) like the reference
ReluGradientOp = _bin('ReluGradientOp', B['RELU_GRAD'], lambda s, x, g: g * (x > 0).to(g.dtype))
GeluGradientOp = _bin('GeluGradientOp', B['GELU_GRAD'], lambda s, x, g: g * (
    0.5 + 0.5 * torch.erf(x * _SQRT1_2) + x * _INV_SQRT_2PI * torch.exp(-0.5 * x * x)))
LeakyReluGradientOp = _bin('LeakyReluGradientOp', B['LEAKY_RELU_GRAD'],
                           lambda s, x, g: torch.where(x > 0, g, g * s.a))
TanhGradientOp = _bin('TanhGradientOp', B['TANH_GRAD'], lambda s, y, g: g * (1 - y * y))
SigmoidGradientOp = _bin('SigmoidGradientOp', B['SIGMOID_GRAD'], lambda s, y, g: g * y * (1 - y))
AbsGradientOp = _bin('AbsGradientOp', B['ABS_GRAD'], lambda s, x, g: g * x.sign())
PowGradientOp = _bin('PowGradientOp', B['POW_GRAD'], lambda s, x, g: s.a * x.pow(s.a - 1) * g)
ConstPowGradientOp = _bin('ConstPowGradientOp', B['CONST_POW_GRAD'],
                          lambda s, y, g: y * g * math.log(s.a))
LogGradOp = _bin('LogGradOp', B['LOG_GRAD'], lambda s, g, x: g / x)


class MaskedFillOp(BinaryOp):
    code = B['MASKED_FILL']

    def __init__(self, input, mask, val, ctx=None):
        super().__init__(input, mask, float(val), ctx)
        self.val = val

    def fn(self, x, m):
        return torch.where(m != 0, torch.full_like(x, float(self.val)), x) \
            if x.device.type != 'meta' else torch.empty_like(x)

    def gradient(self, g):
        return [None, None]


class WhereOp(Op):
    def __init__(self, cond, node_A, node_B, ctx=None):
        super().__init__(WhereOp, [cond, node_A, node_B], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        c, a, b = input_vals
        if _k(c, a, b) and c.shape == a.shape == b.shape:
            out = torch.empty_like(a)
            L.call('where', L.p(c.contiguous()), L.p(a.contiguous()), L.p(b.contiguous()), L.p(out),
                   L.c_sz(a.numel()), L.stream_ptr(a.device))
            return out
        return torch.where(c != 0, a, b)

    def gradient(self, g):
        zeros = zeroslike_op(self.inputs[0], ctx=self.raw_ctx)
        return [None, where_op(self.inputs[0], g, zeros, ctx=self.raw_ctx),
                where_op(self.inputs[0], zeros, g, ctx=self.raw_ctx)]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[1])


class WhereConstOp(BinaryOp):
    code = B['WHERE_CONST']

    def __init__(self, cond, node_A, const_attr, ctx=None):
        super().__init__(cond, node_A, float(const_attr), ctx)
        self.const_attr = const_attr

    def fn(self, c, a):
        return torch.where(c != 0, a, torch.full_like(a, float(self.const_attr))) \
            if a.device.type != 'meta' else torch.empty_like(a)

    def gradient(self, g):
        return [None, where_const_op(self.inputs[0], g, 0.0, ctx=self.raw_ctx)]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[1])


# ------------------------------------------------------------- constructors
class _FillLikeOp(Op):
    def __init__(self, node, value, ctx=None):
        super().__init__(type(self), [node], ctx)
        self.value = float(value)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        if isinstance(x, torch.Tensor) and _k(x):
            return _unary_kernel(U['FILL'], x, self.value)
        return torch.full(tuple(x.shape), self.value, dtype=torch.float32, device=x.device)

    def gradient(self, g):
        return [None]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


class OnesLikeOp(_FillLikeOp):
    def __init__(self, node, ctx=None):
        super().__init__(node, 1.0, ctx)


class ZerosLikeOp(_FillLikeOp):
    def __init__(self, node, ctx=None):
        super().__init__(node, 0.0, ctx)


class FullLikeOp(_FillLikeOp):
    pass


class _SourceOp(Op):
    """Ops with no inputs produce their value on the executor's device."""

    def _device(self):
        from ..ndarray import DLContext
        return self.ctx.torch_device if isinstance(self.ctx, DLContext) else torch.device('cpu')


class FullOp(_SourceOp):
    def __init__(self, size, fill_value, ctx=None):
        super().__init__(FullOp, [], ctx)
        self.size, self.fill_value = tuple(size), float(fill_value)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        return torch.full(self.size, self.fill_value, dtype=torch.float32, device=self._device())

    def gradient(self, g):
        return []

    def infer_shape(self, input_shapes):
        return self.size


class RandOp(_SourceOp):
    graph_safe = True

    def __init__(self, size, ctx=None):
        super().__init__(RandOp, [], ctx)
        self.size = tuple(size)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        return torch.rand(self.size, dtype=torch.float32, device=self._device())

    def gradient(self, g):
        return []

    def infer_shape(self, input_shapes):
        return self.size


class ArangeOp(_SourceOp):
    def __init__(self, start, end, step=1.0, ctx=None):
        super().__init__(ArangeOp, [], ctx)
        self.start, self.end, self.step = start, end, step

    def compute(self, input_vals, output_val=None, stream_handle=None):
        return torch.arange(self.start, self.end, self.step, dtype=torch.float32,
                            device=self._device())

    def gradient(self, g):
        return []

    def infer_shape(self, input_shapes):
        return (int(math.ceil((self.end - self.start) / self.step)),)


# ---------------------------------------------------------------- public API
def abs_op(node_A, ctx=None):
    return AbsOp(node_A, ctx=ctx)


def abs_gradient_op(node_A, node_B, ctx=None):
    return AbsGradientOp(node_A, node_B, ctx=ctx)


def addbyconst_op(node, const_val, ctx=None):
    return AddByConstOp(node, float(const_val), ctx=ctx)


def add_op(node_A, node_B, ctx=None):
    return AddOp(node_A, node_B, ctx=ctx)


def minus_op(node_A, node_B, ctx=None):
    return MinusOp(node_A, node_B, ctx=ctx)


def minus_byconst_op(node, const_val, ctx=None):
    return MinusByConstOp(node, float(const_val), ctx=ctx)


def mul_op(node_A, node_B, ctx=None):
    return MulOp(node_A, node_B, ctx=ctx)


def mul_byconst_op(node_A, const_val, ctx=None):
    return MulByConstOp(node_A, float(const_val), ctx=ctx)


def div_op(node_A, node_B, ctx=None):
    return DivOp(node_A, node_B, ctx=ctx)


def div_const_op(const_val, node_A, ctx=None):
    return DivConstOp(node_A, float(const_val), ctx=ctx)


def opposite_op(node, ctx=None):
    return OppositeOp(node, ctx=ctx)


def exp_op(node, ctx=None):
    return ExpOp(node, ctx=ctx)


def log_op(input, ctx=None):
    return LogOp(input, ctx=ctx)


def log_grad_op(output_grad, input, ctx=None):
    return LogGradOp(output_grad, input, ctx=ctx)


def floor_op(node, ctx=None):
    return FloorOp(node, ctx=ctx)


def gelu_op(node, ctx=None):
    return GeluOp(node, ctx=ctx)


def gelu_gradient_op(node_A, node_B, ctx=None):
    return GeluGradientOp(node_A, node_B, ctx=ctx)


def leaky_relu_op(node, alpha, ctx=None):
    return LeakyReluOp(node, float(alpha), ctx=ctx)


def leaky_relu_gradient_op(node_A, node_B, alpha, ctx=None):
    return LeakyReluGradientOp(node_A, node_B, float(alpha), ctx=ctx)


def relu_op(node, ctx=None):
    return ReluOp(node, ctx=ctx)


def relu_gradient_op(node_A, node_B, ctx=None):
    return ReluGradientOp(node_A, node_B, ctx=ctx)


def sigmoid_op(node, ctx=None):
    return SigmoidOp(node, ctx=ctx)


def sigmoid_gradient_op(forward_node, output_grad, ctx=None):
    return SigmoidGradientOp(forward_node, output_grad, ctx=ctx)


def tanh_op(node, ctx=None):
    return TanhOp(node, ctx=ctx)


def tanh_gradient_op(forward_node, output_grad, ctx=None):
    return TanhGradientOp(forward_node, output_grad, ctx=ctx)


def sin_op(node, ctx=None):
    return SinOp(node, ctx=ctx)


def cos_op(node, ctx=None):
    return CosOp(node, ctx=ctx)


def sqrt_op(node, ctx=None):
    return SqrtOp(node, ctx=ctx)


def rsqrt_op(node, ctx=None):
    return ReciprocalSqrtOp(node, ctx=ctx)


def pow_op(node, eps, ctx=None):
    return PowOp(node, float(eps), ctx=ctx)


def pow_gradient_op(node_A, node_B, eps, ctx=None):
    return PowGradientOp(node_A, node_B, float(eps), ctx=ctx)


def const_pow_op(node, val, ctx=None):
    return ConstPowOp(node, float(val), ctx=ctx)


def const_pow_gradient_op(input_node, grad_node, val, ctx=None):
    return ConstPowGradientOp(input_node, grad_node, float(val), ctx=ctx)


def bool_op(node, input=None, val=None, cond=0, ctx=None):
    if input is None and val is None:
        return BoolOp(node, ctx=ctx)
    if input is not None and val is None:
        return BoolMatrixOp(node, input, cond=cond, ctx=ctx)
    if input is None and val is not None:
        return BoolValOp(node, val, cond=cond, ctx=ctx)
    raise AssertionError('bool_op: give either `input` or `val`, not both')


def clamp_op(node_A, min_mat=None, max_mat=None, min=None, max=None, ctx=None):
    return ClampOp(node_A, min_mat, max_mat, min, max, ctx=ctx)


def masked_fill_op(input, mask, val, ctx=None):
    return MaskedFillOp(input, mask, val, ctx=ctx)


def where_op(cond, node_A, node_B, ctx=None):
    return WhereOp(cond, node_A, node_B, ctx=ctx)


def where_const_op(cond, node_A, const_attr, ctx=None):
    return WhereConstOp(cond, node_A, const_attr, ctx=ctx)


def oneslike_op(node, ctx=None):
    return OnesLikeOp(node, ctx=ctx)


def zeroslike_op(node, ctx=None):
    return ZerosLikeOp(node, ctx=ctx)


def full_op(size, fill_value, ctx=None):
    return FullOp(size, fill_value, ctx=ctx)


def full_like_op(node, fill_value, ctx=None):
    return FullLikeOp(node, fill_value, ctx=ctx)


def rand_op(size, ctx=None):
    return RandOp(size, ctx=ctx)


def arange_op(start, end, step=1.0, ctx=None):
    return ArangeOp(start, end, step, ctx=ctx)
