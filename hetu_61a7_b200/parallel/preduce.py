"""Partial reduce (heterogeneity-aware DP).  reference: python/hetu/preduce.py:8-42 and
ps-lite/src/preduce_handler.cc:6-57: the PS scheduler groups whichever workers
are ready within ``wait_time`` ms; the group then averages its weights over a
lazily created sub-communicator."""
from __future__ import annotations

import numpy as np

from .comm import new_group_comm, world


class PartialReduce(object):
    def __init__(self, reduce_key=0, ps_comm=None):
        from ..graph.executor import get_worker_communicate
        self._reduce_key = reduce_key
        self.ps_comm = ps_comm if ps_comm is not None else get_worker_communicate()
        self.comm = world()
        self.rank, self.nrank = self.comm.rank, self.comm.nrank
        self._comm_map = {}

    def get_partner(self, max_worker=-1, wait_time=1.0):
        """Ranks (sorted tuple) of the group formed for this call."""
        if max_worker < 0:
            max_worker = self.nrank
        if self.ps_comm is None or self.nrank == 1:
            return tuple(range(self.nrank))
        partners = self.ps_comm.preduce_get_partner(self._reduce_key, self.rank, max_worker, wait_time)
        return tuple(sorted(int(p) for p in partners))

    def preduce(self, array, partner, stream=None):
        from ..ndarray import NDArray
        t = array.tensor if isinstance(array, NDArray) else array
        if len(partner) <= 1:
            return array
        if partner not in self._comm_map:
            self._comm_map[partner] = new_group_comm(list(partner))
        self._comm_map[partner].allreduce(t, op='avg')
        return array
