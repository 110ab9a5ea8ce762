"""``CacheSparseTable`` - Python facade of the HET embedding cache (csrc_ps/cache.cc).

reference: python/hetu/cstable.py:19-211 over the ``hetu_cache`` pybind module.  Calls are
asynchronous by default (a thread runs the blocking C call, ctypes drops the GIL) and return a
future-like object with ``wait()``; ``sync=True`` blocks."""
from __future__ import annotations

import ctypes
from concurrent.futures import ThreadPoolExecutor

import numpy as np

from .. import _lib as L

_POLICY = {'LRU': 0, 'LFU': 1, 'LFUOPT': 2}
_I64P = ctypes.POINTER(ctypes.c_int64)
_F32P = ctypes.POINTER(ctypes.c_float)


class _Future(object):
    def __init__(self, fut, result=None):
        self._fut, self._result = fut, result

    def wait(self):
        if self._fut is not None:
            self._fut.result()
            self._fut = None
        return self._result

    result = wait


class CacheSparseTable(object):
    def __init__(self, limit, length, width, node_id, policy="LRU", bound=100, comm=None, push_bound=None):
        lib = L.ps_lib()
        if lib is None:
            raise RuntimeError('libhetu_ps.so is not built; run `python __graft_entry__.py`')
        self.lib = lib
        lib.hcache_create.restype = ctypes.c_void_p
        lib.hcache_last_error.restype = ctypes.c_char_p
        lib.hcache_size.restype = ctypes.c_uint64
        lib.hcache_keys.restype = ctypes.c_uint64
        self._limit, self._length, self._width, self.node_id = int(limit), int(length), int(width), int(node_id)
        self.policy = policy.upper()
        assert self.policy in _POLICY, 'policy must be LRU, LFU or LFUOpt'
        self.bound = int(bound)
        self.handle = ctypes.c_void_p(lib.hcache_create(
            ctypes.c_uint64(self._limit), ctypes.c_uint64(self._length), ctypes.c_uint64(self._width),
            ctypes.c_uint64(self.node_id), _POLICY[self.policy], ctypes.c_int64(self.bound),
            ctypes.c_int64(self.bound if push_bound is None else int(push_bound))))
        self.pool = ThreadPoolExecutor(max_workers=1)      # calls on one table are ordered
        self._perf_on = False

    def __del__(self):
        try:
            self.pool.shutdown(wait=True)
            self.lib.hcache_destroy(self.handle)
        except Exception:
            pass

    def _chk(self, rc):
        if rc != 0:
            raise RuntimeError('HET cache: ' + self.lib.hcache_last_error().decode())

    def _submit(self, fn, result, sync):
        if sync:
            fn()
            return _Future(None, result)
        return _Future(self.pool.submit(fn), result)

    def embedding_lookup(self, keys, dest=None, sync=False):
        k = np.ascontiguousarray(np.asarray(keys).reshape(-1), dtype=np.int64)
        if dest is None:
            dest = np.empty((k.size, self._width), dtype=np.float32)
        d = dest if isinstance(dest, np.ndarray) else dest.numpy()
        fut = self._submit(lambda: self._chk(self.lib.hcache_lookup(
            self.handle, k.ctypes.data_as(_I64P), ctypes.c_uint64(k.size), d.ctypes.data_as(_F32P))), dest, sync)
        return dest if sync else fut

    def embedding_update(self, keys, grads, sync=False):
        k = np.ascontiguousarray(np.asarray(keys).reshape(-1), dtype=np.int64)
        g = np.ascontiguousarray(grads, dtype=np.float32).reshape(k.size, self._width)
        return self._submit(lambda: self._chk(self.lib.hcache_update(
            self.handle, k.ctypes.data_as(_I64P), ctypes.c_uint64(k.size), g.ctypes.data_as(_F32P))), None, sync)

    def embedding_push_pull(self, pullkeys, dest, pushkeys, grads, sync=False):
        pk = np.ascontiguousarray(np.asarray(pullkeys).reshape(-1), dtype=np.int64)
        uk = np.ascontiguousarray(np.asarray(pushkeys).reshape(-1), dtype=np.int64)
        g = np.ascontiguousarray(grads, dtype=np.float32).reshape(uk.size, self._width)
        if dest is None:
            dest = np.empty((pk.size, self._width), dtype=np.float32)
        d = dest if isinstance(dest, np.ndarray) else dest.numpy()
        fut = self._submit(lambda: self._chk(self.lib.hcache_push_pull(
            self.handle, pk.ctypes.data_as(_I64P), ctypes.c_uint64(pk.size), d.ctypes.data_as(_F32P),
            uk.ctypes.data_as(_I64P), ctypes.c_uint64(uk.size), g.ctypes.data_as(_F32P))), dest, sync)
        return dest if sync else fut

    @property
    def width(self):
        return self._width

    @property
    def limit(self):
        return self._limit

    def perf_enabled(self, enable=True):
        self._perf_on = enable

    @property
    def perf(self):
        out = (ctypes.c_int64 * 6)()
        self.lib.hcache_perf(self.handle, out)
        names = ['num_lookup', 'num_miss', 'num_evict', 'num_transfered', 'num_pushed', 'num_requests']
        return dict(zip(names, [int(v) for v in out]))

    def flush(self) -> int:
        """Push every local update that is still below the push bound (and evicted dirty rows) to the servers;
        ``Executor.save`` calls this so that a checkpoint contains them (the reference does not flush on save)."""
        self.lib.hcache_flush.restype = ctypes.c_int64
        n = int(self.lib.hcache_flush(self.handle))
        if n < 0:
            self._chk(-1)
        return n

    def invalidate(self) -> None:
        """Drop every cached row and every un-pushed local update (after the server copy was replaced, e.g.
        ``Executor.load``); the next lookup pulls fresh rows."""
        self.pool.submit(lambda: None).result()            # queued calls on this table finish first
        self.lib.hcache_invalidate(self.handle)

    def bypass(self):
        self.lib.hcache_bypass(self.handle, 1)

    def undobypass(self):
        self.lib.hcache_bypass(self.handle, 0)

    def lookup(self, key):
        row = np.empty(self._width, dtype=np.float32)
        ok = self.lib.hcache_peek(self.handle, ctypes.c_int64(int(key)), row.ctypes.data_as(_F32P))
        return row if ok else None

    def count(self, key):
        return int(self.lib.hcache_count(self.handle, ctypes.c_int64(int(key))))

    def insert(self, key, embedding):
        e = np.ascontiguousarray(embedding, dtype=np.float32)
        self._chk(self.lib.hcache_put(self.handle, ctypes.c_int64(int(key)), e.ctypes.data_as(_F32P)))

    def keys(self):
        cap = max(int(self.lib.hcache_size(self.handle)), 1)
        out = np.empty(cap, dtype=np.int64)
        n = int(self.lib.hcache_keys(self.handle, out.ctypes.data_as(_I64P), ctypes.c_uint64(cap)))
        return out[:min(n, cap)]

    def overall_miss_rate(self, include_cold_start=False):
        p = self.perf
        if p['num_lookup'] == 0:
            return 0.0
        miss = p['num_miss'] if include_cold_start else max(p['num_miss'] - min(self._limit, p['num_miss']), 0)
        return miss / p['num_lookup']

    def overall_data_rate(self, include_cold_start=False):
        p = self.perf
        return p['num_transfered'] / max(p['num_lookup'], 1)

    def debug_keys(self):
        return sorted(self.keys().tolist())
