"""ONNX -> Hetu graph.  reference: python/hetu/onnx/onnx2hetu.py:32 ``load_onnx(path)`` returning
(input placeholders, output nodes); handlers mirror the exporter's op set."""
from __future__ import annotations

from typing import Dict

import numpy as np

from . import proto as P


def load_onnx(path):
    import hetu as ht
    with open(path, 'rb') as f:
        m = P.read_model(f.read())
    env: Dict[str, object] = {}
    consts: Dict[str, np.ndarray] = dict(m['initializers'])
    shape_of: Dict[str, str] = {}            # tensor produced by Shape(x) -> name of x
    inputs = []
    for name, shape, _ in m['inputs']:
        if name in consts:
            continue
        v = ht.Variable(name)
        env[name] = v
        inputs.append(v)

    def get(name):
        if name not in env:
            arr = consts[name]
            env[name] = ht.Variable(name, value=np.asarray(arr, dtype=np.float32), trainable=arr.ndim > 0 and arr.dtype == np.float32)
        return env[name]

    simple = {'Relu': ht.relu_op, 'Sigmoid': ht.sigmoid_op, 'Tanh': ht.tanh_op, 'Exp': ht.exp_op, 'Log': ht.log_op,
              'Sqrt': ht.sqrt_op, 'Neg': ht.opposite_op, 'Abs': ht.abs_op, 'Gelu': ht.gelu_op, 'Floor': ht.floor_op,
              'Sin': ht.sin_op, 'Cos': ht.cos_op}
    binary = {'Add': (ht.add_op, ht.addbyconst_op), 'Mul': (ht.mul_op, ht.mul_byconst_op), 'Sub': (ht.minus_op, None),
              'Div': (ht.div_op, None)}
    for op, ins, outs, attrs, _ in m['nodes']:
        o = outs[0]
        if op in simple:
            env[o] = simple[op](get(ins[0]))
        elif op == 'Softmax':
            env[o] = ht.softmax_op(get(ins[0]))
        elif op == 'LeakyRelu':
            env[o] = ht.leaky_relu_op(get(ins[0]), attrs.get('alpha', 0.01))
        elif op in binary:
            a, b = ins
            scalar = consts.get(b) if b in consts and np.ndim(consts[b]) == 0 else None
            if scalar is not None and binary[op][1] is not None:
                env[o] = binary[op][1](get(a), float(scalar))
            else:
                env[o] = binary[op][0](get(a), get(b))
        elif op == 'MatMul':
            a, b = get(ins[0]), get(ins[1])
            env[o] = ht.matmul_op(a, b) if (ins[1] in consts and consts[ins[1]].ndim == 2) else ht.batch_matmul_op(a, b)
        elif op == 'Gemm':
            a, b = get(ins[0]), get(ins[1])
            tA, tB = bool(attrs.get('transA', 0)), bool(attrs.get('transB', 0))
            env[o] = ht.linear_op(a, b, get(ins[2]), trans_A=tA, trans_B=tB) if len(ins) > 2 else \
                ht.matmul_op(a, b, trans_A=tA, trans_B=tB)
        elif op == 'Conv':
            pads, strides = attrs.get('pads', [0, 0, 0, 0]), attrs.get('strides', [1, 1])
            y = ht.conv2d_op(get(ins[0]), get(ins[1]), padding=int(pads[0]), stride=int(strides[0]))
            if len(ins) > 2:
                y = y + ht.conv2d_broadcastto_op(get(ins[2]), y)
            env[o] = y
        elif op in ('MaxPool', 'AveragePool'):
            k, pads, strides = attrs['kernel_shape'], attrs.get('pads', [0, 0, 0, 0]), attrs.get('strides', [1, 1])
            fn = ht.max_pool2d_op if op == 'MaxPool' else ht.avg_pool2d_op
            env[o] = fn(get(ins[0]), int(k[0]), int(k[1]), padding=int(pads[0]), stride=int(strides[0]))
        elif op == 'BatchNormalization':
            node = ht.batch_normalization_op(get(ins[0]), get(ins[1]), get(ins[2]), momentum=1.0 - attrs.get('momentum', 0.9),
                                             eps=attrs.get('epsilon', 1e-5))
            import torch
            node.running_mean = torch.from_numpy(np.asarray(consts[ins[3]], dtype=np.float32).copy())
            node.running_var = torch.from_numpy(np.asarray(consts[ins[4]], dtype=np.float32).copy())
            env[o] = node
        elif op == 'LayerNormalization':
            env[o] = ht.layer_normalization_op(get(ins[0]), get(ins[1]), get(ins[2]), eps=attrs.get('epsilon', 1e-5))
        elif op == 'Reshape':
            env[o] = ht.array_reshape_op(get(ins[0]), [int(v) for v in consts[ins[1]]])
        elif op == 'Flatten':
            env[o] = ht.array_reshape_op(get(ins[0]), [0, -1] if False else [-1, 1])  # resolved by Reshape in practice
        elif op == 'Transpose':
            env[o] = ht.transpose_op(get(ins[0]), attrs.get('perm'))
        elif op == 'Concat':
            env[o] = ht.concatenate_op([get(i) for i in ins], axis=int(attrs.get('axis', 0)))
        elif op == 'Slice':
            starts, ends = [int(v) for v in consts[ins[1]]], [int(v) for v in consts[ins[2]]]
            size = [-1 if e >= (1 << 60) else e - s for s, e in zip(starts, ends)]
            env[o] = ht.slice_op(get(ins[0]), starts, size)
        elif op == 'Shape':
            shape_of[o] = ins[0]
        elif op == 'Expand':
            like = shape_of[ins[1]]
            src = get(ins[0])
            env[o] = ht.conv2d_broadcastto_op(src.inputs[0], get(like)) if getattr(src, '_bias4d', False) else \
                ht.broadcastto_op(src, get(like))
        elif op == 'ReduceSum':
            axes = [int(v) for v in consts[ins[1]]] if len(ins) > 1 else attrs.get('axes')
            env[o] = ht.reduce_sum_op(get(ins[0]), axes, keepdims=bool(attrs.get('keepdims', 1)))
        elif op == 'ReduceMean':
            env[o] = ht.reduce_mean_op(get(ins[0]), attrs.get('axes'), keepdims=bool(attrs.get('keepdims', 1)))
        elif op in ('Identity', 'Dropout'):
            env[o] = get(ins[0])
        elif op == 'Cast':
            env[o] = get(ins[0])
        elif op == 'OneHot':
            env[o] = ht.one_hot_op(get(ins[0]), int(np.asarray(consts[ins[1]]).reshape(-1)[0]))
        elif op == 'Where':
            env[o] = ht.where_op(get(ins[0]), get(ins[1]), get(ins[2]))
        elif op == 'Reciprocal':
            env[o] = ht.div_const_op(1.0, get(ins[0]))
        elif op == 'Gather':
            env[o] = ht.embedding_lookup_op(get(ins[0]), get(ins[1]))
        elif op == 'Pad':
            pads = [int(v) for v in consts[ins[1]]]
            n = len(pads) // 2
            env[o] = ht.pad_op(get(ins[0]), [[pads[i], pads[i + n]] for i in range(n)],
                               constant_values=float(consts[ins[2]]) if len(ins) > 2 else 0.0)
        else:
            raise NotImplementedError('no Hetu handler for ONNX op %s' % op)
        if op == 'Reshape' and list(consts[ins[1]]) == [1, -1, 1, 1]:
            env[o]._bias4d = True
    return inputs, [env[name] for name, _, _ in m['outputs']]
