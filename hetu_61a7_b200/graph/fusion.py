"""Graph-level epilogue fusion run by the executor before the first step.

matmul / linear followed by relu: the activation moves into the GEMM epilogue (tcgen05 kernel:
applied on the TMEM->register path; small/SIMT kernels: in the store loop) and the ReluOp becomes
a pass-through.  Safe without rebuilding the backward graph because relu is idempotent and its
gradient only needs the sign: every consumer of the GEMM output must be ``ReluOp`` or
``ReluGradientOp`` (which then see y = relu(x) instead of x: relu(y) == y, (y > 0) == (x > 0)).
The reference launches the activation as its own kernel (gpu_ops/Relu.py over src/ops/Relu.cu).
"""
from __future__ import annotations

import os
from typing import Dict, List

ACT_RELU = 1


def fuse_matmul_bias(topo: List) -> int:
    """``matmul(x, w) + broadcastto(b, .)`` (the reference's spelling of a dense layer, examples/cnn/models/MLP.py)
    becomes one GEMM with the bias in its epilogue: the MatMulOp takes the bias as a third input, the AddOp turns
    into a pass-through and the BroadcastToOp produces nothing.  The backward graph is untouched (the adjoint of
    the sum goes to both addends whatever their values; the matmul gradients read x and w, not the product).
    Host executors only unless ``HETU_FUSE_BIAS=all`` (the device path keeps its measured kernel sequence)."""
    import torch
    mode = os.environ.get('HETU_FUSE_BIAS', 'cpu')
    if mode == '0' or (mode != 'all' and torch.cuda.is_available()):
        return 0
    consumers: Dict = {}
    for n in topo:
        for k, i in enumerate(n.inputs):
            if i is not None:
                consumers.setdefault(i, []).append((n, k))
    fused = 0
    present = set(id(n) for n in topo)
    for a in topo:
        if a.op_type != 'AddOp' or len(a.inputs) != 2 or getattr(a, 'mp_state', None) is not None:
            continue
        mm, bc = a.inputs
        if mm is None or bc is None or mm.op_type != 'MatMulOp' or bc.op_type != 'BroadcastToOp':
            continue
        if id(mm) not in present or id(bc) not in present:          # a requested output keeps its own value
            continue
        if getattr(mm, 'act', 0) != 0 or len(mm.inputs) != 2 or len(bc.inputs) != 2 or bc.inputs[1] is not mm:
            continue
        if getattr(mm, 'mp_state', None) is not None or getattr(bc, 'mp_state', None) is not None:
            continue
        bias = bc.inputs[0]
        if len(getattr(bias, 'shape', None) or ()) != 1:
            continue
        # besides the add (and the broadcast that uses the product as its shape reference) only shape-only readers
        # may consume the two nodes: ReduceToShapeOp(.., like) looks at like.shape and nothing else
        shape_only = lambda c, k: c.op_type == 'ReduceToShapeOp' and k == 1                      # noqa: E731
        if not all((c is a and k == 0) or (c is bc and k == 1) or shape_only(c, k) for c, k in consumers.get(mm, [])):
            continue
        if not all((c is a and k == 1) or shape_only(c, k) for c, k in consumers.get(bc, [])):
            continue
        from ..ops.linalg import gemm
        mm.inputs = list(mm.inputs) + [bias]
        if hasattr(mm, 'global_order') and getattr(bias, 'global_order', None) is not None \
                and bias.global_order > mm.global_order:
            bias.global_order = mm.global_order - 0.5   # placed graphs walk nodes in global order: the new input comes first
        mm.compute = (lambda input_vals, output_val=None, stream_handle=None, _m=mm:
                      gemm(input_vals[0], input_vals[1], _m.matmul_attr_trans_A, _m.matmul_attr_trans_B, bias=input_vals[2], act=_m.act))
        a.compute = (lambda input_vals, output_val=None, stream_handle=None: input_vals[0])
        # value unused (the add ignores it); shape readers get a zero-copy broadcast view
        bc.compute = (lambda input_vals, output_val=None, stream_handle=None: input_vals[0].expand_as(input_vals[1]))
        a.bias_fused_into = mm
        fused += 1
    return fused


def fuse_matmul_relu(topo: List) -> int:
    if os.environ.get('HETU_FUSE_EPILOGUE', '1') != '1':
        return 0
    consumers: Dict = {}
    for n in topo:
        for k, i in enumerate(n.inputs):
            if i is not None:
                consumers.setdefault(i, []).append((n, k))
    fused = 0
    for n in topo:
        if n.op_type not in ('MatMulOp', 'LinearOp') or getattr(n, 'act', 0) != 0:
            continue
        cons = consumers.get(n, [])
        through = [c for c, _ in cons if getattr(c, 'bias_fused_into', None) is n]
        if through:                                   # matmul + bias already folded: the activation follows the pass-through add
            cons = consumers.get(through[0], [])
        if not cons or not any(c.op_type == 'ReluOp' for c, _ in cons):
            continue
        if all((c.op_type == 'ReluOp') or (c.op_type == 'ReluGradientOp' and k == 0) for c, k in cons):
            n.act = ACT_RELU
            for c, _ in cons:
                if c.op_type == 'ReluOp':
                    c.compute = (lambda input_vals, output_val=None, stream_handle=None: input_vals[0])
                    c.fused_into = n
            fused += 1
    return fused
