"""CPU ring all-reduce over TCP for host arrays (the reference's legacy CPU data-parallel path).

reference: src/communication/c_communication_nthread.cc:32-62 (``DL_Connect_Init(worker_number,
worker_id, source_ip, target_ip)``), :82-112 (``DL_Communicate_Init``), :497-506
(``DL_Communicate_Server/Client`` = averaged ring all-reduce), :489-495 (``DL_Communicate_Close``).
Native part: csrc_ps/ring.cc."""
from __future__ import annotations

import ctypes

import numpy as np

from .ps import _lib

_F32P = ctypes.POINTER(ctypes.c_float)


def _chk(rc, what):
    if rc != 0:
        lib = _lib()
        lib.hring_last_error.restype = ctypes.c_char_p
        raise RuntimeError('%s failed: %s' % (what, lib.hring_last_error().decode()))


class RingCommunicator(object):
    """``addresses``: [(host, port)] of all ranks; rank r listens on its port and sends to rank r+1."""

    def __init__(self, rank: int, addresses):
        self.rank, self.nrank = int(rank), len(addresses)
        nh, np_ = addresses[(self.rank + 1) % self.nrank]
        _chk(_lib().hring_connect_init(self.nrank, self.rank, int(addresses[self.rank][1]), str(nh).encode(), int(np_)),
             'ring connect')

    def allreduce(self, arr: np.ndarray, average: bool = True) -> np.ndarray:
        """In place on a contiguous float32 array."""
        if arr.dtype != np.float32 or not arr.flags['C_CONTIGUOUS']:
            raise TypeError('ring all-reduce works in place on contiguous float32 arrays')
        _chk(_lib().hring_allreduce(arr.ctypes.data_as(_F32P), ctypes.c_int64(arr.size), int(bool(average))), 'ring all-reduce')
        return arr

    def allgather(self, arr: np.ndarray) -> np.ndarray:
        if arr.dtype != np.float32 or not arr.flags['C_CONTIGUOUS']:
            raise TypeError('ring all-gather works in place on contiguous float32 arrays')
        _chk(_lib().hring_allgather(arr.ctypes.data_as(_F32P), ctypes.c_int64(arr.size)), 'ring all-gather')
        return arr

    def segment(self, length: int, rank=None):
        r = self.rank if rank is None else rank
        return length * r // self.nrank, length * (r + 1) // self.nrank

    def close(self):
        _lib().hring_close()


_ring = None


# reference spellings -------------------------------------------------------------------------------
def DL_Connect_Init(worker_number, worker_id, source_ip, target_ip):
    """``source_ip`` = this rank's ``host:port`` (listen), ``target_ip`` = the next rank's."""
    global _ring
    sh, sp = str(source_ip).replace('tcp://', '').rsplit(':', 1)
    th, tp = str(target_ip).replace('tcp://', '').rsplit(':', 1)
    addrs = [None] * int(worker_number)
    addrs[int(worker_id)] = (sh, int(sp))
    addrs[(int(worker_id) + 1) % int(worker_number)] = (th, int(tp))
    ring = RingCommunicator.__new__(RingCommunicator)
    ring.rank, ring.nrank = int(worker_id), int(worker_number)
    _chk(_lib().hring_connect_init(ring.nrank, ring.rank, int(sp), th.encode(), int(tp)), 'ring connect')
    _ring = ring
    return ring


def DL_Communicate_Init(arr=None):
    return _ring


def DL_Communicate(arr):
    """Averaged ring all-reduce of a numpy array / NDArray (reference DL_Communicate_Server/Client)."""
    a = arr if isinstance(arr, np.ndarray) else arr.asnumpy()
    a = np.ascontiguousarray(a, dtype=np.float32)
    _ring.allreduce(a, average=True)
    if not isinstance(arr, np.ndarray):
        arr[:] = a
    return a


DL_Communicate_Server = DL_Communicate
DL_Communicate_Client = DL_Communicate


def DL_Communicate_Close():
    global _ring
    if _ring is not None:
        _ring.close()
        _ring = None
