"""Hetu graph -> ONNX (opset 17).

reference: python/hetu/onnx/hetu2onnx.py:27 ``export(executor, inputs, outputs, path)`` and its
per-op handlers under python/hetu/onnx/onnx_opset/.  The inference sub-graph reaching ``outputs``
is walked in topological order; trainable placeholders become initializers (current values read
from the executor), fed placeholders become graph inputs.
"""
from __future__ import annotations

from typing import Dict, List

import numpy as np

from . import proto as P
from ..graph.autodiff import find_topo_sort
from ..ndarray import NDArray
from ..ops.variable import PlaceholderOp

_UNARY = {'ReluOp': 'Relu', 'SigmoidOp': 'Sigmoid', 'TanhOp': 'Tanh', 'ExpOp': 'Exp', 'LogOp': 'Log', 'SqrtOp': 'Sqrt',
          'OppositeOp': 'Neg', 'AbsOp': 'Abs', 'SoftmaxOp': 'Softmax', 'FloorOp': 'Floor', 'SinOp': 'Sin', 'CosOp': 'Cos'}
_BINARY = {'AddOp': 'Add', 'MinusOp': 'Sub', 'MulOp': 'Mul', 'DivOp': 'Div'}


class _Ctx(object):
    def __init__(self):
        self.nodes: List[bytes] = []
        self.inits: List[bytes] = []
        self.n = 0

    def const(self, arr, hint='const'):
        name = '%s_%d' % (hint, self.n)
        self.n += 1
        self.inits.append(P.tensor(name, np.asarray(arr)))
        return name

    def add(self, op_type, inputs, outputs, **attrs):
        self.nodes.append(P.node(op_type, inputs, outputs, **attrs))

    def tmp(self, hint='t'):
        self.n += 1
        return '%s_%d' % (hint, self.n)


def _value_of(executor, node):
    v = executor.config.placeholder_to_arr_map.get(node)
    if v is None:
        v = node.tensor_value
    if isinstance(v, NDArray):
        return v.asnumpy()
    if hasattr(v, 'detach'):
        return v.detach().cpu().numpy()
    return np.asarray(v)


def _emit(c: _Ctx, node, ins: List[str], out: str):
    t = node.op_type
    if t in _UNARY:
        c.add(_UNARY[t], ins, [out], **({'axis': -1} if t == 'SoftmaxOp' else {}))
    elif t in _BINARY:
        c.add(_BINARY[t], ins, [out])
    elif t == 'GeluOp':
        c.add('Gelu', ins, [out])                                     # opset 20; decomposed by older runtimes
    elif t == 'LeakyReluOp':
        c.add('LeakyRelu', ins, [out], alpha=float(node.a))
    elif t == 'AddByConstOp':
        c.add('Add', [ins[0], c.const(np.float32(node.a))], [out])
    elif t == 'MulByConstOp':
        c.add('Mul', [ins[0], c.const(np.float32(node.a))], [out])
    elif t == 'MatMulOp' or t == 'LinearOp':
        tA, tB = int(node.matmul_attr_trans_A), int(node.matmul_attr_trans_B)
        act = getattr(node, 'act', 0)               # epilogue-fused activation (graph/fusion.py)
        if t == 'MatMulOp' and len(ins) == 3:
            # bias folded into the GEMM by fuse_matmul_bias: the graph still holds the Add (and the Relu) that follow,
            # and they are exported as themselves, so the product is exported bare
            ins, act = ins[:2], 0
        mm_out = c.tmp('mm') if act else out
        if t == 'LinearOp' or tA or tB:
            c.add('Gemm', ins, [mm_out], transA=tA, transB=tB, alpha=1.0, beta=1.0)
        else:
            c.add('MatMul', ins, [mm_out])
        if act:
            c.add({1: 'Relu', 2: 'Gelu', 3: 'Sigmoid'}[act], [mm_out], [out])
    elif t == 'BatchMatMulOp':
        a, b = ins
        if getattr(node, 'trans_A', False) or getattr(node, 'trans_B', False):
            raise NotImplementedError('transposed batch_matmul export')
        c.add('MatMul', [a, b], [out])
    elif t == 'Conv2dOp':
        p, s = node.padding, node.stride
        p = (p, p) if isinstance(p, int) else tuple(p)
        s = (s, s) if isinstance(s, int) else tuple(s)
        c.add('Conv', ins, [out], pads=[p[0], p[1], p[0], p[1]], strides=list(s))
    elif t in ('MaxPool2dOp', 'AvgPool2dOp'):
        p, s = node.padding, node.stride
        attrs = dict(kernel_shape=[node.kernel_H, node.kernel_W], pads=[p, p, p, p], strides=[s, s])
        if t == 'AvgPool2dOp':
            attrs['count_include_pad'] = 1
        c.add('MaxPool' if t == 'MaxPool2dOp' else 'AveragePool', ins, [out], **attrs)
    elif t == 'BatchNormOp':
        C = None
        mean = node.running_mean.detach().cpu().numpy() if node.running_mean is not None else None
        var = node.running_var.detach().cpu().numpy() if node.running_var is not None else None
        if mean is None:
            raise ValueError('batch norm %s has no running statistics yet: run the executor once' % node.name)
        c.add('BatchNormalization', ins[:3] + [c.const(mean, 'bn_mean'), c.const(var, 'bn_var')], [out],
              epsilon=float(node.eps), momentum=float(1.0 - node.momentum))
    elif t == 'LayerNormOp':
        c.add('LayerNormalization', ins, [out], axis=-1, epsilon=float(node.eps))
    elif t == 'ArrayReshapeOp':
        c.add('Reshape', [ins[0], c.const(np.asarray(node.output_shape, dtype=np.int64), 'shape')], [out])
    elif t == 'TransposeOp':
        c.add('Transpose', ins, [out], **({'perm': list(node.perm)} if node.perm is not None else {}))
    elif t in ('ConcatOp', 'ConcatenateOp'):
        c.add('Concat', ins, [out], axis=int(node.axis))
    elif t == 'SliceOp':
        begin = list(node.begin)
        ends = [(b + s) if s >= 0 else np.iinfo(np.int64).max for b, s in zip(begin, node.size)]
        c.add('Slice', [ins[0], c.const(np.asarray(begin, np.int64), 'starts'), c.const(np.asarray(ends, np.int64), 'ends'),
                        c.const(np.arange(len(begin), dtype=np.int64), 'axes')], [out])
    elif t == 'BroadcastToOp':
        shp = c.tmp('shape')
        c.add('Shape', [ins[1]], [shp])
        c.add('Expand', [ins[0], shp], [out])
    elif t == 'Conv2dBroadcastToOp':                                   # bias [C] -> like [N, C, H, W]
        r, shp = c.tmp('bias4d'), c.tmp('shape')
        c.add('Reshape', [ins[0], c.const(np.asarray([1, -1, 1, 1], np.int64), 'shape')], [r])
        c.add('Shape', [ins[1]], [shp])
        c.add('Expand', [r, shp], [out])
    elif t in ('ReduceSumOp', 'ReduceMeanOp'):
        axes = list(node.axes) if node.axes is not None else None
        if t == 'ReduceSumOp':
            c.add('ReduceSum', [ins[0]] + ([c.const(np.asarray(axes, np.int64), 'axes')] if axes is not None else []),
                  [out], keepdims=int(bool(node.keepdims)))
        else:
            c.add('ReduceMean', [ins[0]], [out], keepdims=int(bool(node.keepdims)),
                  **({'axes': axes} if axes is not None else {}))
    elif t == 'DropoutOp':
        c.add('Identity', [ins[0]], [out])                             # inference graph
    elif t == 'EmbeddingLookUp':
        idx = c.tmp('ids')
        c.add('Cast', [ins[1]], [idx], to=P.INT64)
        c.add('Gather', [ins[0], idx], [out], axis=0)
    elif t == 'PadOp':
        pads = np.asarray(node.paddings, dtype=np.int64)
        c.add('Pad', [ins[0], c.const(np.concatenate([pads[:, 0], pads[:, 1]]), 'pads'),
                      c.const(np.float32(node.constant_values), 'pad_value')], [out], mode='constant')
    elif t == 'OneHotOp':
        idx = c.tmp('ids')
        c.add('Cast', [ins[0]], [idx], to=P.INT64)
        c.add('OneHot', [idx, c.const(np.asarray([int(node.num_classes)], np.int64), 'depth'),
                         c.const(np.asarray([0.0, 1.0], np.float32), 'values')], [out], axis=-1)
    elif t == 'WhereOp':
        cond = c.tmp('cond')
        c.add('Cast', [ins[0]], [cond], to=P.BOOL)
        c.add('Where', [cond, ins[1], ins[2]], [out])
    elif t in ('RsqrtOp', 'ReciprocalSqrtOp'):
        r = c.tmp('sqrt')
        c.add('Sqrt', [ins[0]], [r])
        c.add('Reciprocal', [r], [out])
    elif t == 'DivConstOp':
        c.add('Div', [c.const(np.float32(node.a if hasattr(node, 'a') else node.const_attr), 'const'), ins[0]], [out])
    elif t == 'MinusByConstOp':
        c.add('Sub', [c.const(np.float32(node.a if hasattr(node, 'a') else node.const_attr), 'const'), ins[0]], [out])
    elif t == 'FusedAttentionOp':
        raise NotImplementedError('export fused attention with fused_attention=False models')
    else:
        raise NotImplementedError('no ONNX handler for %s' % t)


def export(executor, inputs, outputs, path, name='default'):
    """Write the inference graph computing ``outputs`` from ``inputs`` (lists of nodes) to ``path``."""
    c = _Ctx()
    names: Dict = {}
    g_inputs, g_outputs = [], []
    input_rank = {n: k for k, n in enumerate(inputs)}
    g_input_keys = []
    sub = executor.subexecutor[name] if name in executor.subexecutor else next(iter(executor.subexecutor.values()))
    shapes = getattr(sub, 'node_to_shape_map', {}) or {}
    for node in find_topo_sort(list(outputs)):
        nm = node.name.replace('(', '_').replace(')', '_')
        names[node] = nm
        if isinstance(node, PlaceholderOp):
            if node in inputs or (node.tensor_value is None and executor.config.placeholder_to_arr_map.get(node) is None):
                shp = shapes.get(node) or node.shape or ()
                g_inputs.append(P.value_info(nm, ['N'] + list(shp[1:]) if shp else ['N']))
                g_input_keys.append(input_rank.get(node, len(inputs) + len(g_input_keys)))
            else:
                c.inits.append(P.tensor(nm, _value_of(executor, node).astype(np.float32)))
            continue
        if node in inputs:
            shp = shapes.get(node) or ()
            g_inputs.append(P.value_info(nm, ['N'] + list(shp[1:]) if shp else ['N']))
            g_input_keys.append(input_rank.get(node, len(inputs) + len(g_input_keys)))
            continue
        _emit(c, node, [names[i] for i in node.inputs if i is not None], nm)
    for o in outputs:
        shp = shapes.get(o) or ()
        g_outputs.append(P.value_info(names[o], ['N'] + list(shp[1:]) if shp else ['N']))
    # graph inputs in the order the caller listed them (then any further free placeholders in graph order)
    g_inputs = [gi for _, gi in sorted(zip(g_input_keys, g_inputs), key=lambda kv: kv[0])]
    buf = P.model(c.nodes, g_inputs, g_outputs, c.inits)
    with open(path, 'wb') as f:
        f.write(buf)
    return path
