"""Cluster bring-up and the cross-worker reductions used by the ``parallel=True`` models.

reference: third_party/HetuML/hetuml/python/hetuml/cluster.py:8-55 (roles scheduler / server /
worker selected through DMLC_* variables, ``start_server(data_type)``, ``join()``) on top of the
HetuML-private ps-lite fork (hetuml/cc/ps-lite).  Here the roles run on the framework's own
parameter server (csrc_ps/ps.cc): one server implementation for both libraries."""
from __future__ import annotations

import os
from typing import Optional

import numpy as np

_COMM: Optional['PSComm'] = None


def current_comm() -> 'PSComm':
    if _COMM is None:
        raise RuntimeError('parallel=True needs a worker-role Cluster(...) to be created first')
    return _COMM


class PSComm(object):
    """Sum / max / mean over the workers, built from PS push + barrier + pull.  Every reduction
    uses a fresh zero-initialised server tensor (so the result is exact, no running residue);
    the previous one is dropped once everybody has left it."""

    _KEY_BASE = 1 << 40

    def __init__(self, worker):
        self.w = worker
        self.rank, self.nrank = worker.rank(), worker.nrank()
        self._seq = 0
        self._prev = None

    def allreduce(self, buf: np.ndarray) -> np.ndarray:
        """In-place sum over workers of a float64 vector (transported as float32 pairs hi + lo so that
        integer counts up to 2^48 and histogram sums keep their precision)."""
        flat = buf.reshape(-1)
        n = flat.size
        hi = flat.astype(np.float32)
        lo = (flat - hi.astype(np.float64)).astype(np.float32)
        key = self._KEY_BASE + self._seq
        self._seq += 1
        self.w.InitTensor(key, 0, 2 * n, 1, 0, 0.0, 0.0, 0)          # includes a worker barrier
        if self._prev is not None and self.rank == 0:
            self.w.Clear(self._prev)
        self.w.Push(key, np.concatenate([hi, lo]))
        self.w.BarrierWorker()
        out = self.w.Pull(key)
        self._prev = key
        flat[:] = out[:n].astype(np.float64) + out[n:].astype(np.float64)
        return buf

    def allgather(self, vec: np.ndarray) -> np.ndarray:
        vec = np.asarray(vec, dtype=np.float64).reshape(-1)
        buf = np.zeros((self.nrank, vec.size), dtype=np.float64)
        buf[self.rank] = vec
        return self.allreduce(buf)

    def allreduce_max(self, vec: np.ndarray) -> np.ndarray:
        return self.allgather(vec).max(axis=0)

    def average(self, vec: np.ndarray) -> np.ndarray:
        buf = np.asarray(vec, dtype=np.float64).copy()
        self.allreduce(buf)
        return np.ascontiguousarray((buf / self.nrank).astype(np.float32)).reshape(np.shape(vec))

    def barrier(self):
        self.w.BarrierWorker()


class Cluster(object):
    _ROLES = ('scheduler', 'server', 'worker')

    def __init__(self, scheduler: str, num_servers: int, num_workers: int, role: str):
        assert role in Cluster._ROLES, 'Unrecognizable role: %s, should be one of %s' % (role, Cluster._ROLES)
        splits = scheduler.split(':')
        assert len(splits) == 2, 'Invalid address: ' + scheduler
        self.scheduler_ip, self.scheduler_port = splits
        self.role, self.num_servers, self.num_workers = role, num_servers, num_workers
        self._setup_env()
        from ..parallel import ps
        self._ps = ps
        self.rank = 0
        self.worker = None
        global _COMM
        if role == 'scheduler':
            ps.start_scheduler_from_env(block=False)
        elif role == 'worker':
            self.worker = ps.PSWorker.from_env()
            self.rank = self.worker.rank()
            _COMM = PSComm(self.worker)

    def start_server(self, data_type='float'):
        """Run this process's server shard (all value types share the float server; 64-bit sums are
        carried as float pairs by ``PSComm``)."""
        if self.role == 'server':
            if data_type.lower() not in ('int', 'float', 'double'):
                raise Exception('Data type %s is not supported' % data_type)
            self._ps.start_server_from_env(block=False)

    def join(self):
        global _COMM
        if self.role == 'worker' and self.worker is not None:
            self.worker.BarrierWorker()
            self.worker.finalize()
            _COMM = None
        elif self.role == 'server':
            self._ps._lib().hps_wait_servers()
        elif self.role == 'scheduler':
            self._ps._lib().hps_wait_scheduler()

    def shutdown(self):
        """Stop a scheduler / server role started in this process."""
        if self.role == 'server':
            self._ps.stop_local_servers()
        elif self.role == 'scheduler':
            self._ps.stop_scheduler()

    def _setup_env(self):
        os.environ['DMLC_PS_ROOT_URI'] = self.scheduler_ip
        os.environ['DMLC_PS_ROOT_PORT'] = self.scheduler_port
        os.environ['DMLC_NUM_SERVER'] = str(self.num_servers)
        os.environ['DMLC_NUM_WORKER'] = str(self.num_workers)
        os.environ['DMLC_ROLE'] = self.role
