"""Memory planning.

reference: python/hetu/memory_pool.py:20-194 (``HetuMemoryPool``: reuse buffers of equal shape once
their last consumer has run; ``test_memory`` dry-runs a graph to predict peak memory) and the
uncompiled C++ BFC allocator in src/memory_pool.

Here tensors come from torch's caching allocator / the CUDA-graph private pool, and the executor
drops a value at its last use (graph/executor.py ``_compute_all(release=True)``), so there is no
separate buffer-assignment step at run time.  What remains useful - and is implemented here - is
(1) the native best-fit-with-coalescing allocator (csrc_cpu/bfc_allocator.cc) that sub-allocates the
symmetric-heap arena, and (2) the planner that replays a step's buffer lifetimes through that
allocator to predict peak memory / fragmentation before anything is launched (strategy search,
"will this batch size fit in 180 GB").
"""
from __future__ import annotations

import ctypes
from typing import Dict, Iterable, List, Optional, Sequence

import numpy as np

from . import _lib as L


class BFCAllocator(object):
    """Offset allocator over ``capacity`` bytes (best fit, split on allocate, coalesce on free)."""

    def __init__(self, capacity: int, alignment: int = 256):
        lib = L.cpu_lib()
        if lib is None:
            raise RuntimeError('libhetu_cpu.so is not built')
        lib.hetu_bfc_create.restype = ctypes.c_void_p
        lib.hetu_bfc_alloc.restype = ctypes.c_int64
        self._lib = lib
        self.capacity, self.alignment = int(capacity), int(alignment)
        self._h = ctypes.c_void_p(lib.hetu_bfc_create(ctypes.c_uint64(self.capacity), ctypes.c_uint64(self.alignment)))

    def alloc(self, nbytes: int) -> int:
        """Byte offset of a block of at least ``nbytes`` (rounded up to the alignment); -1 when nothing fits."""
        return int(self._lib.hetu_bfc_alloc(self._h, ctypes.c_uint64(int(nbytes))))

    def free(self, offset: int) -> None:
        if self._lib.hetu_bfc_free(self._h, ctypes.c_uint64(int(offset))) != 0:
            raise ValueError('offset %d is not a live block' % offset)

    def stats(self) -> Dict[str, int]:
        out = (ctypes.c_uint64 * 6)()
        self._lib.hetu_bfc_stats(self._h, out)
        return dict(zip(('in_use', 'peak', 'live_blocks', 'free_fragments', 'largest_free', 'allocations'),
                        (int(v) for v in out)))

    def __del__(self):
        try:
            if self._h:
                self._lib.hetu_bfc_destroy(self._h)
                self._h = None
        except Exception:
            pass


def _nbytes(shape, itemsize=4) -> int:
    if shape is None:
        return 0
    return int(np.prod(shape, dtype=np.int64)) * itemsize if len(shape) else itemsize


class HetuMemoryPool(object):
    """Lifetime-based memory plan of one sub-executor's step."""

    def __init__(self, capacity: int = 180 << 30, alignment: int = 512):
        self.capacity, self.alignment = capacity, alignment
        self.plan: Dict = {}
        self.peak_bytes = 0
        self.param_bytes = 0

    def memory_plan(self, topo_order: Sequence, node_to_shape_map: Dict, eval_nodes: Iterable = (),
                    persistent: Iterable = ()):
        """Walk ``topo_order``: allocate every node's output when it is computed, free it after its last
        consumer (parameters, feeds, eval nodes and ``persistent`` nodes live for the whole step).
        Returns {node: (offset, nbytes)}; ``self.peak_bytes`` is the high-water mark of activations,
        ``self.param_bytes`` the resident parameters."""
        from .ops.variable import PlaceholderOp
        from .optimizer import OptimizerOp
        keep = set(eval_nodes) | set(persistent)
        last_use: Dict = {}
        for i, n in enumerate(topo_order):
            for inp in n.inputs:
                if inp is not None:
                    last_use[inp] = i
        alloc = BFCAllocator(self.capacity, self.alignment)
        live: Dict = {}
        self.plan, self.param_bytes = {}, 0
        for i, n in enumerate(topo_order):
            if isinstance(n, OptimizerOp):
                continue
            size = _nbytes(node_to_shape_map.get(n))
            if isinstance(n, PlaceholderOp) and not n.is_feed:
                self.param_bytes += size
                continue
            if size:
                off = alloc.alloc(size)
                if off < 0:
                    raise MemoryError('step does not fit in %d bytes (node %s needs %d)' % (self.capacity, n.name, size))
                live[n] = off
                self.plan[n] = (off, size)
            for inp in n.inputs:
                if inp is not None and last_use.get(inp) == i and inp in live and inp not in keep:
                    alloc.free(live.pop(inp))
        st = alloc.stats()
        self.peak_bytes = st['peak']
        self.stats = st
        return self.plan

    def test_memory(self, subexecutor, feed_shapes: Dict) -> Dict[str, int]:
        """Dry run: predicted bytes for one step of ``subexecutor`` with the given feed shapes
        (reference memory_pool.test_memory).  Nothing is launched."""
        shapes = subexecutor.infer_shape(feed_shapes)
        self.memory_plan(subexecutor.topo_order, shapes, eval_nodes=subexecutor.eval_node_list)
        return {'activation_peak_bytes': self.peak_bytes, 'parameter_bytes': self.param_bytes,
                'total_bytes': self.peak_bytes + self.param_bytes, 'buffers': len(self.plan),
                'fragments_at_end': self.stats['free_fragments']}
