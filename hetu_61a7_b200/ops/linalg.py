"""Matrix products.

reference op files: MatrixMult Linear Addmm BatchMatrixMult Baddbmm MatrixDot
Outer(kernel only) CuSparse (python/hetu/gpu_ops/*.py); the reference calls
cublasSgemm / cublasSgemmStridedBatched / hand CSR kernels
(src/ops/MatrixMult.cu:22-26, Linear.cu:35-54, BatchMatrixMult.cu:31-36,
CuSparseCsrmm.cu).

Here every dense product is ``hetu_gemm`` (csrc/gemm_sm100.cu): a
warp-specialised tcgen05 kernel - TMA loads into 128B-swizzled shared memory,
``tcgen05.mma kind::tf32`` accumulating in TMEM, epilogue (alpha/beta/bias/activation) read back with
``tcgen05.ld``.  ``set_matmul_precision('bf16' | 'fp16')`` routes large products through
``hetu_gemm_f16`` (csrc/gemm_f16_sm100.cu): operands converted to 16 bit (and to K-major) in one pass,
``tcgen05.mma kind::f16`` with fp32 accumulation, fp32 result - storage and master weights stay fp32.  All
four transpose combinations are served by K-major / MN-major shared-memory
descriptors, no operand is physically transposed.  Shapes TMA cannot describe
(row pitch not a multiple of 16 B, e.g. the 13-feature WDL input) and the
``fp32`` precision mode use the register-tiled FFMA kernel in the same file.
"""
from __future__ import annotations

import os

import numpy as np
import torch

from .. import _lib as L
from ..graph.node import Op, as_index
from ..ndarray import ND_Sparse_Array

ACT_NONE, ACT_RELU, ACT_GELU, ACT_SIGMOID = 0, 1, 2, 3
ACT_RELU_MASK = 4      # tensor-core kernels only: out holds h on entry, out = (h > 0) ? A @ B : 0  (csrc/gemm_sm100.cu)
_PRECISION = {'mode': os.environ.get('HETU_MATMUL_PRECISION', 'tf32')}
_MODE_CODE = {'tf32': 0, 'fp32': 1, 'bf16': 3, 'fp16': 4, '3xtf32': 1}     # 3xtf32: small products run as exact fp32


def set_matmul_precision(mode: str) -> None:
    """'tf32' (default: tcgen05 kind::tf32, fp32 accumulate), 'fp32' (FFMA, exact), 'bf16' / 'fp16' (operands
    rounded to 16 bit on the fly, tcgen05 kind::f16, fp32 accumulate; products too small for the tensor-core
    path run as tf32), '3xtf32' (fp32-faithful on the tensor cores: every operand is split into its nearest tf32
    value and the remainder, A B = A_lo B_hi + A_hi B_lo + A_hi B_hi, three tcgen05 products with fp32 accumulation;
    products too small for the tensor-core path run as exact fp32)."""
    assert mode in _MODE_CODE
    _PRECISION['mode'] = mode


def get_matmul_precision() -> str:
    return _PRECISION['mode']


def gemm(a: torch.Tensor, b: torch.Tensor, trans_a=False, trans_b=False, bias=None,
         c_in=None, alpha=1.0, beta=0.0, act=ACT_NONE, out=None) -> torch.Tensor:
    """C = act(alpha * op(A) @ op(B) + beta * C_in + bias); 2-D or batched 3-D."""
    if not (L.kernels_on(a, b) and a.dtype == torch.float32 and b.dtype == torch.float32):
        return _gemm_torch(a, b, trans_a, trans_b, bias, c_in, alpha, beta, act)
    a = a if a.is_contiguous() else a.contiguous()
    b = b if b.is_contiguous() else b.contiguous()
    if _PRECISION['mode'] == '3xtf32' and act != ACT_RELU_MASK:
        r = _gemm_3xtf32(a, b, trans_a, trans_b, bias, c_in, alpha, beta, act, out)
        if r is not None:
            return r
    if a.dim() == 2 and b.dim() == 2 and _PRECISION['mode'] not in ('fp32', '3xtf32'):
        padded = _pad_for_tma(a, b, trans_a, trans_b, bias, c_in, alpha, beta, act)
        if padded is not None:
            return padded
    batched = a.dim() == 3
    if batched:
        batch = a.shape[0]
        am, ak = (a.shape[2], a.shape[1]) if trans_a else (a.shape[1], a.shape[2])
        bk, bn = (b.shape[-1], b.shape[-2]) if trans_b else (b.shape[-2], b.shape[-1])
        sa = a.shape[1] * a.shape[2]
        sb = b.shape[-2] * b.shape[-1] if b.dim() == 3 else 0
    else:
        batch = 1
        am, ak = (a.shape[1], a.shape[0]) if trans_a else (a.shape[0], a.shape[1])
        bk, bn = (b.shape[1], b.shape[0]) if trans_b else (b.shape[0], b.shape[1])
        sa = sb = 0
    assert ak == bk, 'gemm inner dims differ: %s x %s (tA=%s tB=%s)' % (
        tuple(a.shape), tuple(b.shape), trans_a, trans_b)
    out_shape = (batch, am, bn) if batched else (am, bn)
    if c_in is not None and beta != 0.0:
        if out is not None and out.data_ptr() == c_in.data_ptr() and out.is_contiguous() and tuple(c_in.shape) == out_shape:
            c = out                                          # accumulate in place
        else:
            c = c_in.expand(out_shape).contiguous() if tuple(c_in.shape) != out_shape else c_in.clone()
    elif out is not None and out.is_contiguous() and out.numel() == int(np.prod(out_shape)) and out.dtype == torch.float32:
        c = out.view(out_shape)                          # caller-provided destination (e.g. a symmetric-heap buffer)
        beta = 0.0
    else:
        c = torch.empty(out_shape, dtype=torch.float32, device=a.device)
        beta = 0.0
    if bias is not None:
        bias = bias.reshape(-1).contiguous()
        assert bias.numel() == bn
    mode = _PRECISION['mode']
    if mode in ('bf16', 'fp16') and float(am) * bn * ak * batch >= float(1 << 25) and am >= 64 and bn >= 8 and ak >= 32 \
            and (batch == 1 or (sa != 0 and sb != 0)):
        kp = (ak + 7) // 8 * 8
        h = torch.bfloat16 if mode == 'bf16' else torch.float16
        a16 = torch.empty(batch * am * kp, dtype=h, device=a.device)
        b16 = torch.empty(batch * bn * kp, dtype=h, device=a.device)
        L.call('gemm_f16', L.p(a), L.p(b), L.p(c), L.p(bias), L.p(a16), L.p(b16), L.c_int(am), L.c_int(bn), L.c_int(ak),
               L.c_int(int(trans_a)), L.c_int(int(trans_b)), L.c_int(batch), L.c_i64(sa), L.c_i64(sb), L.c_i64(am * bn),
               L.c_f32(alpha), L.c_f32(beta), L.c_int(act), L.c_int(1 if mode == 'bf16' else 0), L.stream_ptr(a.device))
        return c
    L.call('gemm', L.p(a), L.p(b), L.p(c), L.p(bias), L.c_int(am), L.c_int(bn), L.c_int(ak),
           L.c_int(int(trans_a)), L.c_int(int(trans_b)), L.c_int(batch), L.c_i64(sa), L.c_i64(sb),
           L.c_i64(am * bn), L.c_f32(alpha), L.c_f32(beta), L.c_int(act),
           L.c_int(_MODE_CODE[_PRECISION['mode']]), L.stream_ptr(a.device))
    return c


def _split_tf32(t):
    hi, lo = torch.empty_like(t), torch.empty_like(t)
    L.call('split_tf32', L.p(t), L.p(hi), L.p(lo), L.c_sz(t.numel()), L.stream_ptr(t.device))
    return hi, lo


def _gemm_3xtf32(a, b, trans_a, trans_b, bias, c_in, alpha, beta, act, out):
    """Three tf32 tensor-core products of the split operands; None when the product would not take the tensor-core
    path anyway (too small, or a row pitch TMA cannot describe) - the caller then runs it as exact fp32."""
    am, ak = ((a.shape[-1], a.shape[-2]) if trans_a else (a.shape[-2], a.shape[-1]))
    bn = b.shape[-2] if trans_b else b.shape[-1]
    batch = a.shape[0] if a.dim() == 3 else 1
    a_inner, b_inner = (am if trans_a else ak), (ak if trans_b else bn)
    if not (float(am) * bn * ak * batch > float(1 << 25) and am >= 64 and bn > 8 and ak >= 32
            and a_inner % 4 == 0 and b_inner % 4 == 0):
        return None
    a_hi, a_lo = _split_tf32(a)
    b_hi, b_lo = _split_tf32(b)
    saved = _PRECISION['mode']
    _PRECISION['mode'] = 'tf32'
    try:
        c = gemm(a_lo, b_hi, trans_a, trans_b, c_in=c_in, alpha=alpha, beta=beta, out=out if c_in is None else None)
        c = gemm(a_hi, b_lo, trans_a, trans_b, c_in=c, alpha=alpha, beta=1.0, out=c)
        c = gemm(a_hi, b_hi, trans_a, trans_b, bias=bias, c_in=c, alpha=alpha, beta=1.0, act=act, out=c)
    finally:
        _PRECISION['mode'] = saved
    return c


def _cpu_native(*ts) -> bool:
    """CPU executor: route through the OpenMP op library (csrc_cpu) when it is built."""
    from .. import cpu_links
    return (os.environ.get('HETU_CPU_NATIVE', '1') == '1' and cpu_links.available()
            and all(t.device.type == 'cpu' and t.dtype == torch.float32 for t in ts))


def _pad4(t, dim):
    n = t.shape[dim]
    if n % 4 == 0:
        return t
    shp = list(t.shape)
    shp[dim] = (n + 3) // 4 * 4
    out = torch.zeros(shp, dtype=t.dtype, device=t.device)
    out.narrow(dim, 0, n).copy_(t)
    return out


def _pad_for_tma(a, b, trans_a, trans_b, bias, c_in, alpha, beta, act):
    """TMA needs 16-byte row pitches.  Large products whose contiguous dimension is not a multiple of
    four floats (first conv layer: K = 27; BERT vocabulary 30522) would otherwise fall to the CUDA-core
    kernel at a few percent of tensor-core speed: zero-pad the odd dimension instead (a copy of one
    operand) and slice the result."""
    m, k = (a.shape[1], a.shape[0]) if trans_a else (a.shape[0], a.shape[1])
    n = b.shape[0] if trans_b else b.shape[1]
    a_inner, b_inner = (m if trans_a else k), (k if trans_b else n)
    if a_inner % 4 == 0 and b_inner % 4 == 0:
        return None                                  # already eligible
    if float(m) * n * k < float(1 << 27) or m < 64 or k < 8 or n <= 8:
        return None                                  # small / skinny: dedicated CUDA-core kernels
    a2 = _pad4(_pad4(a, 0), 1)
    b2 = _pad4(_pad4(b, 0), 1)
    n2 = b2.shape[0] if trans_b else b2.shape[1]
    m2 = a2.shape[1] if trans_a else a2.shape[0]
    bias2 = None
    if bias is not None:
        bias2 = _pad4(bias.reshape(-1), 0)
    c2 = None
    if c_in is not None and beta != 0.0:
        c2 = _pad4(_pad4(c_in.expand(m, n).contiguous(), 0), 1)
    out = gemm(a2, b2, trans_a, trans_b, bias=bias2, c_in=c2, alpha=alpha, beta=beta, act=act)
    if m2 != m or n2 != n:
        out = out[:m, :n].contiguous()
    return out


def _gemm_torch(a, b, trans_a, trans_b, bias, c_in, alpha, beta, act):
    if a.dim() == 2 and b.dim() == 2 and _cpu_native(a, b):
        from .. import cpu_links
        m = a.shape[1] if trans_a else a.shape[0]
        n = b.shape[0] if trans_b else b.shape[1]
        y = torch.empty((m, n), dtype=torch.float32)
        cpu_links.matrix_multiply(a.contiguous(), trans_a, b.contiguous(), trans_b, y)
    else:
        if trans_a:
            a = a.transpose(-1, -2)
        if trans_b:
            b = b.transpose(-1, -2)
        y = torch.matmul(a, b)
    if alpha != 1.0:
        y = y * alpha
    if c_in is not None and beta != 0.0:
        y = y + beta * c_in
    if bias is not None:
        y = y + bias.reshape(-1)
    if act == ACT_RELU:
        y = y.relu()
    elif act == ACT_GELU:
        y = torch.nn.functional.gelu(y)
    elif act == ACT_SIGMOID:
        y = y.sigmoid()
    return y


class MatMulOp(Op):
    def __init__(self, node_A, node_B, trans_A=False, trans_B=False, ctx=None):
        super().__init__(MatMulOp, [node_A, node_B], ctx)
        self.matmul_attr_trans_A = trans_A
        self.matmul_attr_trans_B = trans_B
        self.act = ACT_NONE          # set by the executor's epilogue-fusion pass (graph/fusion.py)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        return gemm(input_vals[0], input_vals[1], self.matmul_attr_trans_A, self.matmul_attr_trans_B, act=self.act)

    def gradient(self, g):
        A, B = self.inputs
        tA, tB = self.matmul_attr_trans_A, self.matmul_attr_trans_B
        c = self.raw_ctx
        if not tA and not tB:      # C = A B
            return [matmul_op(g, B, False, True, ctx=c), matmul_op(A, g, True, False, ctx=c)]
        if tA and not tB:          # C = A^T B
            return [matmul_op(B, g, False, True, ctx=c), matmul_op(A, g, False, False, ctx=c)]
        if not tA and tB:          # C = A B^T
            return [matmul_op(g, B, False, False, ctx=c), matmul_op(g, A, True, False, ctx=c)]
        return [matmul_op(B, g, True, True, ctx=c), matmul_op(g, A, True, True, ctx=c)]

    def infer_shape(self, input_shapes):
        a, b = input_shapes[:2]                      # a fused bias may ride along as a third input (graph/fusion.py)
        m = a[1] if self.matmul_attr_trans_A else a[0]
        n = b[0] if self.matmul_attr_trans_B else b[1]
        return (m, n)


class LinearOp(Op):
    """matmul + bias in the GEMM epilogue (reference Linear.cu does a bias
    broadcast launch followed by SGEMM beta=1)."""

    def __init__(self, node_A, node_B, bias, trans_A=False, trans_B=False, act=ACT_NONE, ctx=None):
        super().__init__(LinearOp, [node_A, node_B, bias], ctx)
        self.matmul_attr_trans_A, self.matmul_attr_trans_B, self.act = trans_A, trans_B, act

    def compute(self, input_vals, output_val=None, stream_handle=None):
        a, b, bias = input_vals
        return gemm(a, b, self.matmul_attr_trans_A, self.matmul_attr_trans_B, bias=bias, act=self.act)

    def gradient(self, g):
        from .reduce import reduce_sum_op
        A, B, _ = self.inputs
        tA, tB = self.matmul_attr_trans_A, self.matmul_attr_trans_B
        c = self.raw_ctx
        if self.act == ACT_RELU:
            from .elementwise import relu_gradient_op
            g = relu_gradient_op(self, g, ctx=c)
        elif self.act != ACT_NONE:
            raise NotImplementedError('fused activation gradient')
        if not tA and not tB:
            ga, gb = matmul_op(g, B, False, True, ctx=c), matmul_op(A, g, True, False, ctx=c)
        elif tA and not tB:
            ga, gb = matmul_op(B, g, False, True, ctx=c), matmul_op(A, g, False, False, ctx=c)
        elif not tA and tB:
            ga, gb = matmul_op(g, B, False, False, ctx=c), matmul_op(g, A, True, False, ctx=c)
        else:
            ga, gb = matmul_op(B, g, True, True, ctx=c), matmul_op(g, A, True, True, ctx=c)
        return [ga, gb, reduce_sum_op(g, [0], keepdims=False, ctx=c)]

    def infer_shape(self, input_shapes):
        a, b = input_shapes[0], input_shapes[1]
        m = a[1] if self.matmul_attr_trans_A else a[0]
        n = b[0] if self.matmul_attr_trans_B else b[1]
        return (m, n)


class AddmmOp(Op):
    """beta * input + alpha * (A @ B)"""

    def __init__(self, node_A, node_B, node_C, alpha=1.0, beta=1.0, ctx=None):
        super().__init__(AddmmOp, [node_A, node_B, node_C], ctx)
        self.alpha, self.beta = alpha, beta

    def compute(self, input_vals, output_val=None, stream_handle=None):
        inp, a, b = input_vals
        return gemm(a, b, c_in=inp, alpha=self.alpha, beta=self.beta)

    def gradient(self, g):
        from .elementwise import mul_byconst_op
        inp, A, B = self.inputs
        c = self.raw_ctx
        return [addmm_gradient_op(inp, g, self.beta, ctx=c),
                mul_byconst_op(matmul_op(g, B, False, True, ctx=c), self.alpha, ctx=c),
                mul_byconst_op(matmul_op(A, g, True, False, ctx=c), self.alpha, ctx=c)]

    def infer_shape(self, input_shapes):
        return (input_shapes[1][0], input_shapes[2][1])


class AddmmGradientOp(Op):
    def __init__(self, node_input, node_grad, beta=1.0, ctx=None):
        super().__init__(AddmmGradientOp, [node_input, node_grad], ctx)
        self.beta = beta

    def compute(self, input_vals, output_val=None, stream_handle=None):
        from .reduce import ReduceToShapeOp
        inp, g = input_vals
        red = ReduceToShapeOp.compute(self, [g, inp])
        return red * self.beta if self.beta != 1.0 else red

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


class BatchMatMulOp(Op):
    def __init__(self, node_A, node_B, trans_A=False, trans_B=False, ctx=None):
        super().__init__(BatchMatMulOp, [node_A, node_B], ctx)
        self.trans_A, self.trans_B = trans_A, trans_B

    def compute(self, input_vals, output_val=None, stream_handle=None):
        a, b = input_vals
        lead = a.shape[:-2]
        if a.device.type == 'meta' or not L.kernels_on(a, b):
            return _gemm_torch(a, b, self.trans_A, self.trans_B, None, None, 1.0, 0.0, ACT_NONE)
        a3 = a.reshape(-1, a.shape[-2], a.shape[-1])
        b3 = b.reshape(-1, b.shape[-2], b.shape[-1])
        y = gemm(a3, b3, self.trans_A, self.trans_B)
        return y.reshape(tuple(lead) + tuple(y.shape[-2:]))

    def gradient(self, g):
        A, B = self.inputs
        tA, tB = self.trans_A, self.trans_B
        c = self.raw_ctx
        bm = batch_matmul_op
        if not tA and not tB:
            return [bm(g, B, False, True, ctx=c), bm(A, g, True, False, ctx=c)]
        if tA and not tB:
            return [bm(B, g, False, True, ctx=c), bm(A, g, False, False, ctx=c)]
        if not tA and tB:
            return [bm(g, B, False, False, ctx=c), bm(g, A, True, False, ctx=c)]
        return [bm(B, g, True, True, ctx=c), bm(g, A, True, True, ctx=c)]


class BaddbmmOp(Op):
    """beta * input + alpha * bmm(A, B)"""

    def __init__(self, node_A, node_B, node_C, alpha=1.0, beta=1.0, ctx=None):
        super().__init__(BaddbmmOp, [node_A, node_B, node_C], ctx)
        self.alpha, self.beta = alpha, beta

    def compute(self, input_vals, output_val=None, stream_handle=None):
        inp, a, b = input_vals
        if a.device.type == 'meta' or not L.kernels_on(a, b, inp):
            return self.beta * inp + self.alpha * torch.matmul(a, b)
        return gemm(a, b, c_in=inp, alpha=self.alpha, beta=self.beta)

    def gradient(self, g):
        from .elementwise import mul_byconst_op
        inp, A, B = self.inputs
        c = self.raw_ctx
        return [mul_byconst_op(g, self.beta, ctx=c),
                mul_byconst_op(batch_matmul_op(g, B, False, True, ctx=c), self.alpha, ctx=c),
                mul_byconst_op(batch_matmul_op(A, g, True, False, ctx=c), self.alpha, ctx=c)]


class MatrixDotOp(Op):
    """Row-wise / axis-wise dot: sum(A * B, axis=axes)... reference MatrixDot.py
    multiplies A by B broadcast along ``axes`` and keeps A's shape."""

    def __init__(self, node_A, node_B, axes=0, ctx=None):
        super().__init__(MatrixDotOp, [node_A, node_B], ctx)
        self.axes = axes

    def compute(self, input_vals, output_val=None, stream_handle=None):
        a, b = input_vals
        return a * b

    def gradient(self, g):
        from .elementwise import mul_op
        from .reduce import reduce_to_shape_op
        A, B = self.inputs
        c = self.raw_ctx
        return [mul_op(g, B, ctx=c), reduce_to_shape_op(mul_op(g, A, ctx=c), B, ctx=c)]


class OuterOp(Op):
    def __init__(self, node_A, node_B, ctx=None):
        super().__init__(OuterOp, [node_A, node_B], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        a, b = input_vals
        return gemm(a.reshape(-1, 1), b.reshape(1, -1))

    def gradient(self, g):
        from .reduce import reduce_sum_op
        from .elementwise import mul_op
        from .shape import broadcastto_op, array_reshape_op
        A, B = self.inputs
        c = self.raw_ctx
        ga = array_reshape_op(matmul_op(g, array_reshape_op(B, [-1, 1], ctx=c), ctx=c), [-1], ctx=c)
        gb = array_reshape_op(matmul_op(array_reshape_op(A, [1, -1], ctx=c), g, ctx=c), [-1], ctx=c)
        return [ga, gb]


def _csr_dense(sp) -> torch.Tensor:
    return sp.to_dense() if isinstance(sp, ND_Sparse_Array) else sp


class CsrmvOp(Op):
    """CSR matrix x dense vector (GNN).  Device path: ``hetu_csr_spmm`` with one column."""

    def __init__(self, node_A, node_B, trans=False, ctx=None):
        super().__init__(CsrmvOp, [node_A, node_B], ctx)
        self.trans = trans

    def compute(self, input_vals, output_val=None, stream_handle=None):
        sp, v = input_vals
        y = csr_spmm(sp, v.reshape(-1, 1), self.trans)
        return y.reshape(-1)

    def gradient(self, g):
        return [None, csrmv_op(self.inputs[0], g, not self.trans, ctx=self.raw_ctx)]

    def infer_shape(self, input_shapes):
        return (input_shapes[0][1] if self.trans else input_shapes[0][0],)


class CsrmmOp(Op):
    def __init__(self, node_A, node_B, trans_A=False, trans_B=False, ctx=None):
        super().__init__(CsrmmOp, [node_A, node_B], ctx)
        self.trans_A, self.trans_B = trans_A, trans_B

    def compute(self, input_vals, output_val=None, stream_handle=None):
        sp, d = input_vals
        if self.trans_B:
            d = d.t().contiguous()
        return csr_spmm(sp, d, self.trans_A)

    def gradient(self, g):
        return [None, csrmm_op(self.inputs[0], g, not self.trans_A, False, ctx=self.raw_ctx)]

    def infer_shape(self, input_shapes):
        a, b = input_shapes
        m = a[1] if self.trans_A else a[0]
        n = b[0] if self.trans_B else b[1]
        return (m, n)


def csr_spmm(sp, dense: torch.Tensor, trans: bool) -> torch.Tensor:
    """(CSR or CSR^T) @ dense.  Warp-per-row gather kernel; the transposed
    product scatters with atomics like the reference's spmm_T_kernel."""
    if not isinstance(sp, ND_Sparse_Array):
        m = sp.t() if trans else sp
        return m @ dense
    if L.kernels_on(dense) and dense.dtype == torch.float32:
        dense = dense.contiguous()
        nrow, ncol = sp.shape
        n = dense.shape[1]
        out_rows = ncol if trans else nrow
        out = torch.zeros((out_rows, n), dtype=torch.float32, device=dense.device)
        L.call('csr_spmm', L.p(sp.data.tensor), L.p(sp.row.tensor), L.p(sp.col.tensor), L.p(dense),
               L.p(out), L.c_int(nrow), L.c_int(n), L.c_int(int(trans)), L.stream_ptr(dense.device))
        return out
    m = sp.to_dense().to(dense.device)
    return (m.t() if trans else m) @ dense


# ---------------------------------------------------------------- public API
def matmul_op(node_A, node_B, trans_A=False, trans_B=False, ctx=None):
    return MatMulOp(node_A, node_B, trans_A, trans_B, ctx=ctx)


def linear_op(node_A, node_B, bias, trans_A=False, trans_B=False, ctx=None):
    return LinearOp(node_A, node_B, bias, trans_A, trans_B, ctx=ctx)


def addmm_op(node_A, node_B, node_C, alpha=1.0, beta=1.0, ctx=None):
    return AddmmOp(node_A, node_B, node_C, alpha, beta, ctx=ctx)


def addmm_gradient_op(node_input, node_grad, beta=1.0, ctx=None):
    return AddmmGradientOp(node_input, node_grad, beta, ctx=ctx)


def batch_matmul_op(node_A, node_B, trans_A=False, trans_B=False, ctx=None):
    return BatchMatMulOp(node_A, node_B, trans_A, trans_B, ctx=ctx)


def baddbmm_op(node_A, node_B, node_C, alpha=1.0, beta=1.0, ctx=None):
    return BaddbmmOp(node_A, node_B, node_C, alpha, beta, ctx=ctx)


def matrix_dot_op(node_A, node_B, axes=0, ctx=None):
    return MatrixDotOp(node_A, node_B, axes, ctx=ctx)


def outer_op(node_A, node_B, ctx=None):
    return OuterOp(node_A, node_B, ctx=ctx)


def csrmv_op(node_A, node_B, trans=False, ctx=None):
    return CsrmvOp(node_A, node_B, trans, ctx=ctx)


def csrmm_op(node_A, node_B, trans_A=False, trans_B=False, ctx=None):
    return CsrmmOp(node_A, node_B, trans_A, trans_B, ctx=ctx)
