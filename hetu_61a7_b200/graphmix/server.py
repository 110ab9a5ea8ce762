"""Graph server process: owns one shard, runs the samplers, serves workers and peer servers.

reference: the server side of third_party/GraphMix (graphmix/src/graph/graph_handle.cc,
sampler.cc; configured from Python in ``server_init(server)``: ``server.init_cache(ratio, policy)``,
``server.add_sampler(kind, **kwargs)``, ``server.is_ready()``)."""
from __future__ import annotations

import ctypes
import json

import numpy as np

from . import _core as C
from .shard import Shard


class Server(object):
    def __init__(self, shard: Shard, rank: int, addresses, num_workers: int = 0):
        """``addresses``: [(host, port)] of every graph server, in rank order."""
        meta = shard.meta
        self._rank = rank
        self.meta = meta
        self.addresses = list(addresses)
        indptr, nbr = shard.csr()
        f = np.ascontiguousarray(shard.f_feat, dtype=np.float32)
        i = np.ascontiguousarray(shard.i_feat, dtype=np.int32)
        off = np.ascontiguousarray(meta['partition']['offset'], dtype=np.int64)
        self._keep = (indptr, nbr, f, i, off)
        self._h = C.lib().hgm_server_create(rank, int(meta['num_part']), off.ctypes.data_as(C.I64P),
                                            ctypes.c_int64(f.shape[1]), ctypes.c_int64(i.shape[1]),
                                            f.ctypes.data_as(C.F32P), i.ctypes.data_as(C.I32P), indptr.ctypes.data_as(C.I64P),
                                            nbr.ctypes.data_as(C.I64P), json.dumps(meta).encode(), int(num_workers))
        if not self._h:
            raise RuntimeError('graph server: ' + C.lib().hgm_last_error().decode())
        C.check(C.lib().hgm_server_set_peers(ctypes.c_void_p(self._h), ','.join('%s:%d' % a for a in self.addresses).encode()),
                'set_peers')
        C.check(C.lib().hgm_server_listen(ctypes.c_void_p(self._h), int(self.addresses[rank][1])), 'listen')

    def rank(self) -> int:
        return self._rank

    def port(self) -> int:
        return int(C.lib().hgm_server_port(ctypes.c_void_p(self._h)))

    def init_cache(self, ratio: float, policy: int = C.cache.LFUOpt):
        """Cache up to ``ratio`` x (number of remote nodes) remote node records."""
        C.check(C.lib().hgm_server_init_cache(ctypes.c_void_p(self._h), ctypes.c_double(ratio), int(policy)), 'init_cache')

    def add_sampler(self, kind: int, batch_size: int = 128, depth: int = 2, width: int = 2, rw_head: int = 128,
                    rw_length: int = 2, index: int = -1, thread: int = 1, queue: int = 4, seed: int = 1, **_):
        for t in range(max(1, int(thread))):
            C.check(C.lib().hgm_server_add_sampler(ctypes.c_void_p(self._h), int(kind), int(batch_size), int(depth), int(width),
                                                   int(rw_head), int(rw_length), int(queue), ctypes.c_uint64(seed + 7919 * t)),
                    'add_sampler')

    def is_ready(self):
        C.check(C.lib().hgm_server_ready(ctypes.c_void_p(self._h)), 'ready')

    def cache_stats(self):
        out = (ctypes.c_int64 * 3)()
        C.check(C.lib().hgm_server_cache_stats(ctypes.c_void_p(self._h), out), 'cache_stats')
        return {'lookups': out[0], 'hits': out[1], 'size': out[2]}

    def wait(self):
        C.lib().hgm_server_wait(ctypes.c_void_p(self._h))

    def stop(self):
        if self._h:
            C.lib().hgm_server_stop(ctypes.c_void_p(self._h))

    def close(self):
        if self._h:
            C.lib().hgm_server_free(ctypes.c_void_p(self._h))
            self._h = None
