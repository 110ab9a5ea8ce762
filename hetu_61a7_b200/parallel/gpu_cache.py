"""GPU tier of the HET embedding cache (host side of csrc/het_cache.cu).

reference: python/hetu/cstable.py (CacheSparseTable over src/hetu_cache) - a host-memory cache in front of the parameter
server with bounded staleness: a cached row may lag the server by at most ``pull_bound`` versions, a local update may stay
un-pushed for at most ``push_bound`` updates of its row.  Same contract here, but the lines live in HBM:

    lookup(ids):   probe kernel (slots + cached versions) -> the versions of the batch's unique ids go to the server
                   (one request: ``hps_sync_embedding``) -> rows that are missing or too old come back, go up through a pinned
                   staging buffer with ONE cudaMemcpyAsync per array on the side stream -> install kernel (evicting by
                   LRU / LFU / LFUOpt inside an 8-way set; dirty victims leave through the eviction list) -> gather kernel.
    update(ids, delta): probe -> update kernel (cached row += delta, kept gradient += delta, update count) -> lines over
                   ``push_bound`` (and rows that were not cached) are drained and pushed (``hps_push_embedding``).

Everything that does not have to cross to the host tier stays on the device: a step whose rows are all cached and fresh
moves 8 bytes per unique id up and nothing down.
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from .. import _lib as L

_POLICY = {'LRU': 0, 'LFU': 1, 'LFUOPT': 2}
WAYS = 8


class GPUHetCache(object):
    def __init__(self, limit, length, width, node_id, policy='LFUOpt', bound=100, comm=None, push_bound=None, device=None):
        self.device = device or torch.device('cuda', torch.cuda.current_device())
        self.length, self.width, self.node_id = int(length), int(width), int(node_id)
        self.policy = _POLICY[policy.upper()]
        self.pull_bound = int(bound)
        self.push_bound = int(bound if push_bound is None else push_bound)
        self.comm = comm
        self.nsets = max((int(limit) + WAYS - 1) // WAYS, 1)
        n = self.nsets * WAYS
        d = self.device
        self.tags = torch.full((n,), -1, dtype=torch.int64, device=d)
        self.data = torch.zeros((n, self.width), dtype=torch.float32, device=d)
        self.grad = torch.zeros((n, self.width), dtype=torch.float32, device=d)
        self.version = torch.full((n,), -1, dtype=torch.int64, device=d)
        self.updates = torch.zeros(n, dtype=torch.int32, device=d)
        self.score = torch.zeros(n, dtype=torch.int32, device=d)
        self.queued = torch.zeros(n, dtype=torch.int32, device=d)
        self.lock = torch.zeros(self.nsets, dtype=torch.int32, device=d)
        self.stats = torch.zeros(4, dtype=torch.int64, device=d)          # lookups, misses, evictions
        self.tick = 0
        self.side = torch.cuda.Stream(device=d)                           # host-tier traffic
        self.pulled_rows = self.pushed_rows = self.requests = 0
        self._pin = {}

    # ------------------------------------------------------------------------------------------ helpers
    def _args(self):
        return (L.p(self.tags), L.p(self.data), L.p(self.grad), L.p(self.version), L.p(self.updates), L.p(self.score),
                L.p(self.queued), L.p(self.lock), L.c_int(self.nsets), L.c_int(self.width), L.c_int(self.policy))

    def _pinned(self, name, shape, dtype):
        """Pinned staging buffer, grown on demand and reused (one per role)."""
        n = int(np.prod(shape))
        cur = self._pin.get(name)
        if cur is None or cur.numel() < n or cur.dtype != dtype:
            cur = self._pin[name] = torch.empty(max(n, 1024), dtype=dtype).pin_memory()
        return cur[:n].view(shape)

    def _to_device(self, name, arr, dtype):
        """numpy -> pinned staging -> device with one asynchronous copy on the side stream."""
        t = torch.from_numpy(np.ascontiguousarray(arr))
        pin = self._pinned(name, tuple(t.shape), dtype)
        pin.copy_(t)
        with torch.cuda.stream(self.side):
            out = pin.to(self.device, non_blocking=True)
        return out

    def _st(self):
        return L.stream_ptr(self.device)

    # ------------------------------------------------------------------------------------------ lookup
    def device_lookup(self, ids: torch.Tensor) -> torch.Tensor:
        shape = tuple(ids.shape)
        flat = ids.reshape(-1).to(self.device).long().contiguous()
        n = flat.numel()
        self.tick += 1
        self.requests += 1
        slot = torch.empty(n, dtype=torch.int32, device=self.device)
        ver = torch.empty(n, dtype=torch.int64, device=self.device)
        L.call('het_probe', *self._args(), L.p(flat), L.c_sz(n), L.p(slot), L.p(ver), ctypes.c_uint(self.tick), L.c_int(1),
               L.p(self.stats), self._st())
        # the staleness check: unique ids + the versions held here (-1 = not cached) go to the server
        ids_h = flat.cpu().numpy()
        ver_h = ver.cpu().numpy()
        uniq, first = np.unique(ids_h, return_index=True)
        fb_index = None
        fetched = None
        if self.comm is not None:
            idx, nver, rows = self.comm.sync_embedding(self.node_id, uniq, ver_h[first], self.pull_bound, self.width)
            if idx.size:
                self.pulled_rows += int(idx.size)
                new_ids = uniq[idx]
                ids_d = self._to_device('ids', new_ids, torch.int64)
                ver_d = self._to_device('ver', nver, torch.int64)
                fetched = self._to_device('rows', rows, torch.float32)
                torch.cuda.current_stream(self.device).wait_stream(self.side)
                ev = self._evict_buffers(int(idx.size))
                L.call('het_install', *self._args(), L.p(ids_d), L.p(ver_d), L.p(fetched), L.c_sz(int(idx.size)),
                       ctypes.c_uint(self.tick), L.p(ev['ids']), L.p(ev['grads']), L.p(ev['upd']), L.p(ev['count']),
                       L.c_int(ev['cap']), L.p(self.stats), self._st())
                L.call('het_probe', *self._args(), L.p(flat), L.c_sz(n), L.p(slot), L.p(None), ctypes.c_uint(self.tick),
                       L.c_int(0), L.p(None), self._st())
                # positions whose row could not be cached (set full) are served from the fetched rows
                pos_of = {int(r): k for k, r in enumerate(new_ids)}
                fb = np.fromiter((pos_of.get(int(r), -1) for r in ids_h), dtype=np.int32, count=n)
                fb_index = self._to_device('fb', fb, torch.int32)
                torch.cuda.current_stream(self.device).wait_stream(self.side)
                self._push_evicted(ev)
        out = torch.empty((n, self.width), dtype=torch.float32, device=self.device)
        L.call('het_gather', *self._args(), L.p(slot), L.c_sz(n), L.p(out), L.p(fetched), L.p(fb_index), self._st())
        return out.view(shape + (self.width,))

    def _evict_buffers(self, cap):
        ev = getattr(self, '_ev', None)
        if ev is None or ev['cap'] < cap:
            d = self.device
            ev = self._ev = dict(cap=max(cap, 1024), ids=torch.empty(max(cap, 1024), dtype=torch.int64, device=d),
                                 grads=torch.empty((max(cap, 1024), self.width), dtype=torch.float32, device=d),
                                 upd=torch.empty(max(cap, 1024), dtype=torch.int64, device=d),
                                 count=torch.zeros(1, dtype=torch.int32, device=d))
        ev['count'].zero_()
        return ev

    def _push_rows(self, ids_d, grads_d, upd_d, k):
        """k rows of kept gradients leave for the server: device -> pinned host on the side stream, then one request."""
        if k == 0 or self.comm is None:
            return
        self.side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(self.side):
            ih = self._pinned('p_ids', (k,), torch.int64)
            gh = self._pinned('p_grads', (k, self.width), torch.float32)
            uh = self._pinned('p_upd', (k,), torch.int64)
            ih.copy_(ids_d[:k], non_blocking=True)
            gh.copy_(grads_d[:k], non_blocking=True)
            uh.copy_(upd_d[:k], non_blocking=True)
        self.side.synchronize()
        self.comm.push_embedding(self.node_id, ih.numpy(), gh.numpy(), uh.numpy())
        self.pushed_rows += k

    def _push_evicted(self, ev):
        k = min(int(ev['count'].item()), ev['cap'])
        self._push_rows(ev['ids'], ev['grads'], ev['upd'], k)

    # ------------------------------------------------------------------------------------------ update
    def device_update(self, ids: torch.Tensor, delta: torch.Tensor) -> None:
        """``delta`` = -lr * grad rows (the reference's workers scale before they push, ParameterServerCommunicate.py:38-57)."""
        flat = ids.reshape(-1).to(self.device).long().contiguous()
        delta = delta.reshape(-1, self.width).to(self.device).contiguous()
        n = flat.numel()
        slot = torch.empty(n, dtype=torch.int32, device=self.device)
        L.call('het_probe', *self._args(), L.p(flat), L.c_sz(n), L.p(slot), L.p(None), ctypes.c_uint(self.tick), L.c_int(0),
               L.p(None), self._st())
        push_slots = torch.empty(n, dtype=torch.int32, device=self.device)
        direct = torch.empty(n, dtype=torch.int32, device=self.device)
        counts = torch.zeros(2, dtype=torch.int32, device=self.device)
        L.call('het_update', *self._args(), L.p(slot), L.p(delta), L.c_sz(n), L.c_int(self.push_bound), L.p(push_slots),
               L.p(counts[0:1]), L.p(direct), L.p(counts[1:2]), self._st())
        npush, ndirect = [int(v) for v in counts.tolist()]
        if npush:
            self._drain(push_slots, counts[0:1], False, npush)
        if ndirect:                                      # rows without a line: their update goes straight to the server
            pos = direct[:ndirect].long()
            self._push_rows(flat[pos], delta[pos], torch.ones(ndirect, dtype=torch.int64, device=self.device), ndirect)

    def _drain(self, push_slots, push_count, all_dirty, cap):
        d = self.device
        out_ids = torch.empty(cap, dtype=torch.int64, device=d)
        out_grads = torch.empty((cap, self.width), dtype=torch.float32, device=d)
        out_upd = torch.empty(cap, dtype=torch.int64, device=d)
        out_count = torch.zeros(1, dtype=torch.int32, device=d)
        L.call('het_drain', *self._args(), L.p(push_slots), L.p(push_count), L.c_int(int(all_dirty)), L.p(out_ids),
               L.p(out_grads), L.p(out_upd), L.p(out_count), L.c_int(cap), self._st())
        k = min(int(out_count.item()), cap)
        self._push_rows(out_ids, out_grads, out_upd, k)
        return k

    def flush(self) -> int:
        """Push every kept gradient (checkpoints, end of training); returns the number of rows pushed."""
        return self._drain(None, None, True, self.nsets * WAYS)

    def invalidate(self) -> None:
        """Forget everything (the server copy was replaced, e.g. by Executor.load)."""
        self.tags.fill_(-1)
        self.version.fill_(-1)
        for t in (self.grad, self.updates, self.score, self.queued):
            t.zero_()

    # ---- CacheSparseTable-compatible surface (numpy in / out), used by code written against the host cache ---------
    def embedding_lookup(self, keys, dest=None, sync=True):
        out = self.device_lookup(torch.from_numpy(np.asarray(keys, dtype=np.int64))).cpu().numpy()
        if dest is not None:
            np.copyto(dest.reshape(out.shape), out)
            return dest
        return out

    def embedding_update(self, keys, grads, sync=True):
        self.device_update(torch.from_numpy(np.asarray(keys, dtype=np.int64)), torch.from_numpy(np.asarray(grads, dtype=np.float32)))

    def perf(self):
        s = [int(v) for v in self.stats.tolist()]
        return {'num_lookup': s[0], 'num_miss': s[1], 'num_evict': s[2], 'num_transfered': self.pulled_rows,
                'num_pushed': self.pushed_rows, 'num_requests': self.requests}

    def overall_miss_rate(self, include_cold_start=False):
        p = self.perf()
        return p['num_miss'] / max(p['num_lookup'], 1)

    def keys(self):
        t = self.tags[self.tags >= 0]
        return t.cpu().numpy()

    def size(self):
        return int((self.tags >= 0).sum().item())
