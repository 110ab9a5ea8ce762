"""Data parallelism.  reference: distributed_strategies/simple.py:6-39."""
from __future__ import annotations

from ..context import DeviceGroup
from ..graph.autodiff import find_topo_sort
from ..ops.variable import PlaceholderOp
from .base import Strategy


class DataParallel(Strategy):
    """aggregate in {'allreduce', 'ps', 'hybrid'}: dense parameters are
    synchronised by all-reduce (fused with the optimizer step on B200), or by
    the parameter server; with 'hybrid' embeddings go to the PS / sharded
    embedding tier and everything else is all-reduced."""

    def __init__(self, aggregate=None):
        super().__init__()
        if aggregate is None:
            aggregate = 'ps' if self.settings.enable_PS else 'allreduce'
        aggregate = aggregate.lower()
        assert aggregate in ('allreduce', 'ps', 'hybrid')
        self.aggregate = aggregate
        devices = self.worker_devices()
        embedding_ctxs = (['cpu:0'] if aggregate != 'allreduce' else []) + devices
        ctxs = (['cpu:0'] if aggregate == 'ps' else []) + devices
        self.embedding_raw_ctx = DeviceGroup(embedding_ctxs)
        self.raw_ctx = DeviceGroup(ctxs)

    def set_raw_ctxs_n_states(self, node_list, memory_pool=None):
        for node in find_topo_sort(node_list):
            if isinstance(node, PlaceholderOp) and node.trainable and not node.is_embed:
                node.raw_ctx = self.raw_ctx
            else:
                node.raw_ctx = self.embedding_raw_ctx
        return self.raw_ctx
