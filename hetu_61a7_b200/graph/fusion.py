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
