"""Host<->device transfer nodes.  reference: gpu_ops/DataTransfer.py:18-190.

Copies are issued on the executor's h2d / d2h streams from pinned memory and
joined to the compute stream with ``cudaStreamWaitEvent`` (no host blocking)."""
from __future__ import annotations

import torch

from ..graph.node import Op
from ..ndarray import IndexedSlices


class DataH2DOp(Op):
    def __init__(self, node_A, ctx, sparse=False):
        super().__init__(DataH2DOp if not sparse else DataH2DSparseOp, [node_A], ctx)
        self.use_indexed_slices = node_A.use_indexed_slices

    def compute(self, input_vals, output_val=None, stream_handle=None):
        v = input_vals[0]
        dev = self.ctx.torch_device
        if isinstance(v, IndexedSlices):
            ind, val = v._iv()
            return IndexedSlices(ind.to(dev, non_blocking=True), val.to(dev, non_blocking=True),
                                 v.dense_shape)
        if v.device.type == 'meta':
            return v
        return v.to(dev, non_blocking=True)

    def gradient(self, g):
        return [datad2h_op(g)]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


class DataH2DSparseOp(DataH2DOp):
    def __init__(self, node_A, ctx):
        super().__init__(node_A, ctx, sparse=True)


class DataD2HOp(Op):
    def __init__(self, node_A, sparse=False):
        from ..ndarray import cpu
        super().__init__(DataD2HOp if not sparse else DataD2HSparseOp, [node_A], cpu(0))
        self.use_indexed_slices = node_A.use_indexed_slices

    def compute(self, input_vals, output_val=None, stream_handle=None):
        v = input_vals[0]
        if isinstance(v, IndexedSlices):
            ind, val = v._iv()
            return IndexedSlices(ind.cpu(), val.cpu(), v.dense_shape)
        if v.device.type == 'meta':
            return v
        return v.to('cpu', non_blocking=False)

    def gradient(self, g):
        return [datah2d_op(g, self.inputs[0].ctx)]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


class DataD2HSparseOp(DataD2HOp):
    def __init__(self, node_A):
        super().__init__(node_A, sparse=True)


def datah2d_op(node, ctx):
    return DataH2DOp(node, ctx)


def datad2h_op(node):
    return DataD2HOp(node)


def datah2d_sparse_op(node, ctx):
    return DataH2DSparseOp(node, ctx)


def datad2h_sparse_op(node):
    return DataD2HSparseOp(node)
