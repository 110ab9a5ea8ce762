"""Embedding lookup and its sparse gradient.

reference: python/hetu/gpu_ops/EmbeddingLookUp.py:10-149; kernel
src/ops/EmbeddingLookup.cu:3-19 (one thread per id, serial copy of the row) and
:67-79 (dense scatter with per-element atomicAdd after zeroing the whole table).

B200 path (csrc/embedding.cu): one warp per row, 128-bit loads; ids may be
float32 (reference convention), int32 or int64.  When the table is sharded over
the NVSwitch domain (parallel/sharded_embedding.py) the same kernel receives a
table of peer base pointers and reads rows straight out of the owner GPU's HBM.
The gradient stays sparse (``IndexedSlices``) and is consumed by the fused
sparse optimizer kernels.
"""
from __future__ import annotations

import torch

from .. import _lib as L
from ..graph.node import Op, as_index
from ..ndarray import IndexedSlices

_IDX_CODE = {torch.float32: 0, torch.int32: 1, torch.int64: 2}


def gather_rows_kernel(table: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    idx = idx.contiguous()
    if idx.dtype not in _IDX_CODE:
        idx = idx.long()
    width = table.shape[-1]
    out = torch.empty(tuple(idx.shape) + (width,), dtype=table.dtype, device=table.device)
    L.call('gather_rows', L.p(table), L.p(idx), L.c_int(_IDX_CODE[idx.dtype]), L.p(out),
           L.c_sz(idx.numel()), L.c_int(width), L.c_sz(table.shape[0]), L.stream_ptr(table.device))
    return out


class EmbeddingLookUp(Op):
    def __init__(self, embedding, index, ctx=None):
        super().__init__(EmbeddingLookUp, [embedding, index], ctx)
        self.grad_node = None
        if hasattr(embedding, 'is_embed'):
            embedding.is_embed = True

    def forward_hook(self, config):
        super().forward_hook(config)
        emb = self.inputs[0]
        if getattr(emb, 'ps_resident', False) and getattr(emb, 'is_embed', False):
            # PS / Hybrid: rows come from the parameter server (through the HET cache when one
            # is configured) every step -> host round trip, so this graph runs eagerly
            self._ps_comm = config.ps_comm
            self._config = config
            self.graph_safe = False

    def _ps_lookup(self, idx):
        import numpy as np
        emb = self.inputs[0]
        cache = getattr(emb, 'cache', None)
        if cache is not None and hasattr(cache, 'device_lookup'):       # GPU-resident HET cache: rows never visit the host
            return cache.device_lookup(idx.detach())
        ids = idx.detach().cpu().numpy().astype(np.int64).reshape(-1)
        width = int(emb.shape[-1])
        if cache is not None:
            rows = cache.embedding_lookup(ids, sync=True)
        else:
            rows = self._take_prefetched(emb, ids, width)
            if rows is None:
                rows = self._ps_comm.SparsePull(emb.id, ids, width)
            self._prefetch_next(emb, width)
        t = torch.from_numpy(np.asarray(rows).reshape(tuple(idx.shape) + (width,)))
        dev = self.ctx.torch_device if self.ctx is not None else idx.device
        return t.to(dev, non_blocking=True)

    # ---- PS prefetch (reference executor.py prefetch=True + dataloader.get_next_arr): while this step
    # computes, the rows of the NEXT batch are pulled on the worker's I/O thread.  ASP/SSP only - under
    # BSP the pull has to see this step's update.
    def _take_prefetched(self, emb, ids, width):
        import numpy as np
        pf = getattr(self, '_pf', None)
        self._pf = None
        if pf is None:
            return None
        next_ids, buf = pf
        self._ps_comm.Wait(emb.id)
        if next_ids.shape == ids.shape and np.array_equal(next_ids, ids):
            self.prefetch_hits = getattr(self, 'prefetch_hits', 0) + 1
            return buf
        return None

    def _prefetch_next(self, emb, width):
        import numpy as np
        cfg = getattr(self, '_config', None)
        src = self.inputs[1]
        if cfg is None or not getattr(cfg, 'prefetch', False) or getattr(cfg, 'bsp', -1) == 0 \
                or not hasattr(src, 'get_next_arr'):
            return
        name = getattr(cfg, '_running_name', None) or next(iter(src.dataloaders))
        if name not in src.dataloaders:
            return
        nxt = src.get_next_arr(name)
        next_ids = np.asarray(nxt).astype(np.int64).reshape(-1)
        buf = np.zeros((next_ids.size, width), dtype=np.float32)
        self._ps_comm.SparsePullAsync(emb.id, next_ids, buf)
        self._pf = (next_ids, buf)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        table, idx = input_vals
        if table is None and getattr(self, '_ps_comm', None) is not None:
            return self._ps_lookup(idx)
        if hasattr(table, 'lookup'):                       # ShardedTable / cache handle
            return table.lookup(idx)
        if table.device.type == 'meta':
            return torch.empty(tuple(idx.shape) + (table.shape[-1],), device='meta')
        if L.kernels_on(table) and table.dtype == torch.float32 and idx.is_cuda:
            return gather_rows_kernel(table, idx)
        ids = as_index(idx).to(table.device)
        return table[ids]

    def gradient(self, g):
        self.grad_node = embedding_lookup_gradient_op(g, self.inputs[1], None, ctx=self.raw_ctx)
        self.grad_node.embed_node = self.inputs[0]
        return [self.grad_node, None]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[1]) + (input_shapes[0][-1],)


class EmbeddingLookUp_Gradient(Op):
    """(vectors, index) -> IndexedSlices over the embedding table."""

    def __init__(self, vectors, index, embed_shape, ctx=None):
        super().__init__(EmbeddingLookUp_Gradient, [vectors, index], ctx)
        self.embed_shape = embed_shape
        self.embed_node = None
        self.use_indexed_slices = True

    def compute(self, input_vals, output_val=None, stream_handle=None):
        vec, idx = input_vals
        shape = self.embed_shape
        if shape is None and self.embed_node is not None:
            shape = getattr(self.embed_node, 'shape', None)
        return IndexedSlices(idx, vec, shape)

    def infer_shape(self, input_shapes):
        if self.embed_shape is not None:
            return tuple(self.embed_shape)
        return tuple(self.embed_node.shape)

    def gradient(self, g):
        raise NotImplementedError


def embedding_lookup_op(embedding, index, ctx=None):
    return EmbeddingLookUp(embedding, index, ctx=ctx)


def embedding_lookup_gradient_op(vectors, index, embed_shape, ctx=None):
    return EmbeddingLookUp_Gradient(vectors, index, embed_shape, ctx=ctx)
