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


def _block_rows(node_count, block, nblocks):
    """Rows of block ``block`` when ``node_count`` rows are cut into ``nblocks`` pieces of ceil(n / nblocks)
    (reference DistGCN_15d.py:9-16)."""
    per = -(-node_count // nblocks)
    if node_count % nblocks == 0:
        return node_count // nblocks
    return per if block < nblocks - 1 else node_count - per * (nblocks - 1)


def _pick(groups, index):
    """The reference passes lists of group communicators indexed by grid coordinate; a single communicator is accepted
    as well."""
    if isinstance(groups, (list, tuple)):
        return groups[index] if groups else None
    return groups


class DistGCN_15dOp(Op):
    """Z = A . H (. W) on a (size / c) x c grid of ranks, c = ``replication`` (reference DistGCN_15d.py:19-141).

    Rank r sits at (r // c, r % c).  It holds block-row r // c of the feature matrix H and the block of the adjacency
    with those rows and the columns of its column part (stage blocks r % c * stages .. + stages, stages = size / c^2).
    Per stage the owner of the needed H block broadcasts it inside the column group; the received blocks are
    concatenated and multiplied by the local adjacency block in ONE SpMM (``hetu_csr_spmm``); the partial products
    of the c column parts are all-reduced over the row group.  The backward pass is the same distributed product applied
    to the output gradient (A is the symmetric normalised adjacency, A^T = A), followed by the local GEMMs and an
    all-reduce of the weight gradient over the ranks that share W."""
    graph_safe = False

    def __init__(self, node_A, node_B, node_C, node_Count_Self, node_Count_All, size, replication,
                 device_id, comm, comm_groups=(None, None), need_W=True, ctx=None):
        inputs = [node_A, node_B] + ([node_C] if need_W and node_C is not None else [])
        super().__init__(DistGCN_15dOp, inputs, ctx)
        self.node_Count_Self, self.node_Count_All = node_Count_Self, node_Count_All
        self.size, self.replication, self.device_id = size, max(int(replication), 1), device_id
        self.comm, self.comm_groups = comm, list(comm_groups)
        self.need_W = need_W and node_C is not None

    def _grid(self):
        c = self.replication
        rank = self.comm.rank if self.comm is not None and getattr(self.comm, 'multi', False) else 0
        return rank, rank // c, rank % c, max(self.size // (c * c), 1), max(self.size // c, 1)

    def _spmm_distributed(self, adj, h):
        rank, rank_c, rank_col, stages, nblocks = self._grid()
        c = self.replication
        col_comm = _pick(self.comm_groups[1] if len(self.comm_groups) > 1 else None, rank_col) if c > 1 else self.comm
        row_comm = _pick(self.comm_groups[0] if self.comm_groups else None, rank_c) if c > 1 else None
        if col_comm is None or not getattr(col_comm, 'multi', False):
            blocks = [h]
        else:
            if rank_col == c - 1:
                stages = nblocks - (c - 1) * stages          # the last column part takes the remainder
            first = rank_col * max(self.size // (c * c), 1)
            blocks = []
            for i in range(stages):
                q_c = first + i                               # block-row whose features this stage needs
                if q_c == rank_c:
                    blk = h.contiguous().clone()
                else:
                    blk = torch.empty((_block_rows(self.node_Count_All, q_c, nblocks), h.shape[1]), dtype=h.dtype, device=h.device)
                col_comm.broadcast(blk, root=q_c)            # the column group is ordered by block-row
                blocks.append(blk)
        feats = blocks[0] if len(blocks) == 1 else torch.cat(blocks, 0)
        if isinstance(adj, (list, tuple)):                    # adjacency already cut into per-stage column blocks
            z = None
            for a_blk, f_blk in zip(adj, blocks):
                part = csr_spmm(a_blk, f_blk, False)
                z = part if z is None else z + part
        else:
            z = csr_spmm(adj, feats, False)
        if row_comm is not None and getattr(row_comm, 'multi', False):
            row_comm.allreduce(z)
        return z

    def compute(self, input_vals, output_val=None, stream_handle=None):
        adj, h = input_vals[0], input_vals[1]
        w = input_vals[2] if self.need_W else None
        if w is not None and w.shape[1] < h.shape[1]:         # project first: fewer columns cross the wire
            return self._spmm_distributed(adj, gemm(h, w))
        z = self._spmm_distributed(adj, h)
        return gemm(z, w) if w is not None else z

    def gradient(self, g):
        from .linalg import matmul_op
        from .comm import groupallreduceCommunicate_op
        A, H = self.inputs[0], self.inputs[1]
        ag = DistGCN_15dOp(A, g, None, self.node_Count_Self, self.node_Count_All, self.size, self.replication,
                           self.device_id, self.comm, self.comm_groups, need_W=False, ctx=self.raw_ctx)
        if not self.need_W:
            return [None, ag]
        W = self.inputs[2]
        grad_H = matmul_op(ag, W, False, True, ctx=self.raw_ctx)
        grad_W = matmul_op(H, ag, True, False, ctx=self.raw_ctx)
        rank, rank_c, rank_col, _, _ = self._grid()
        if self.comm is not None and getattr(self.comm, 'multi', False):
            if self.replication > 1:
                groups = self.comm_groups[2] if len(self.comm_groups) == 3 else self.comm_groups[1]
                grad_W = groupallreduceCommunicate_op(grad_W, _pick(groups, rank_col))
            else:
                grad_W = groupallreduceCommunicate_op(grad_W, self.comm)
        return [None, grad_H, grad_W]

    def infer_shape(self, input_shapes):
        rows = input_shapes[0][0]
        return (rows, input_shapes[2][1] if self.need_W else input_shapes[1][1])


def distgcn_15d_op(node_A, node_B, node_C, node_Count_Self, node_Count_All, size, replication,
                   device_id, comm, comm_groups=[None, None], need_W=True):
    return DistGCN_15dOp(node_A, node_B, node_C, node_Count_Self, node_Count_All, size, replication,
                         device_id, comm, comm_groups, need_W)
