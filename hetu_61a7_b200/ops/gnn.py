"""1.5-D distributed GCN layer.  reference: python/hetu/gpu_ops/DistGCN_15d.py:19-155.

Ranks form a (size/replication) x replication grid.  Each rank owns a block-row
of the normalised adjacency (CSR) and a block of the feature matrix H.  One layer
computes Z = A . H . W: for every stage the owning rank of the needed H block
broadcasts it inside its column group, every rank multiplies its adjacency
block with it (CSR SpMM), and the partial products are all-reduced over the row
group.  Collectives are issued through the group communicators; SpMM is the
``hetu_csr_spmm`` kernel."""
from __future__ import annotations

import torch

from ..graph.node import Op
from .linalg import csr_spmm, gemm


class DistGCN_15dOp(Op):
    graph_safe = False

    def __init__(self, node_A, node_B, node_C, node_Count_Self, node_Count_All, size, replication,
                 device_id, comm, comm_groups=(None, None), need_W=True, ctx=None):
        inputs = [node_A, node_B] + ([node_C] if need_W else [])
        super().__init__(DistGCN_15dOp, inputs, ctx)
        self.node_Count_Self, self.node_Count_All = node_Count_Self, node_Count_All
        self.size, self.replication, self.device_id = size, replication, device_id
        self.comm, self.comm_groups, self.need_W = comm, list(comm_groups), need_W

    def compute(self, input_vals, output_val=None, stream_handle=None):
        adj, h = input_vals[0], input_vals[1]
        w = input_vals[2] if self.need_W else None
        row_comm, col_comm = self.comm_groups
        nstage = max(self.size // (self.replication ** 2), 1)
        z = None
        cols = self.size // self.replication
        my_col = self.device_id // self.replication if self.replication > 0 else 0
        for s in range(nstage):
            blk = h
            if col_comm is not None and col_comm.multi:
                blk = h.clone()
                col_comm.broadcast(blk, root=(s + my_col) % col_comm.nrank)
            part = csr_spmm(adj[s] if isinstance(adj, (list, tuple)) else adj, blk, False)
            z = part if z is None else z + part
        if row_comm is not None and row_comm.multi:
            row_comm.allreduce(z)
        if w is not None:
            z = gemm(z, w)
        return z

    def gradient(self, g):
        from .linalg import matmul_op, csrmm_op
        A, H = self.inputs[0], self.inputs[1]
        if self.need_W:
            W = self.inputs[2]
            gz = matmul_op(g, W, False, True, ctx=self.raw_ctx)
            ah = csrmm_op(A, H, ctx=self.raw_ctx)
            return [None, csrmm_op(A, gz, True, False, ctx=self.raw_ctx),
                    matmul_op(ah, g, True, False, ctx=self.raw_ctx)]
        return [None, csrmm_op(A, g, True, False, ctx=self.raw_ctx)]

    def infer_shape(self, input_shapes):
        rows = input_shapes[0][0]
        return (rows, input_shapes[2][1] if self.need_W else input_shapes[1][1])


def distgcn_15d_op(node_A, node_B, node_C, node_Count_Self, node_Count_All, size, replication,
                   device_id, comm, comm_groups=[None, None], need_W=True):
    return DistGCN_15dOp(node_A, node_B, node_C, node_Count_Self, node_Count_All, size, replication,
                         device_id, comm, comm_groups, need_W)
