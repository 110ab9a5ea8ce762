"""ctypes loader for lib/libhetu_graph.so (csrc_graph/graph.cc)."""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_lib = None

I64P = ctypes.POINTER(ctypes.c_int64)
I32P = ctypes.POINTER(ctypes.c_int32)
F32P = ctypes.POINTER(ctypes.c_float)


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        path = os.path.join(_HERE, 'lib', 'libhetu_graph.so')
        if not os.path.exists(path):
            raise RuntimeError('libhetu_graph.so is not built; run `python -c "import __graft_entry__ as g; g.build()"`')
        _lib = ctypes.CDLL(path)
        _lib.hgm_last_error.restype = ctypes.c_char_p
        _lib.hgm_server_create.restype = ctypes.c_void_p
        _lib.hgm_client_create.restype = ctypes.c_void_p
        _lib.hgm_client_pull_graph.restype = ctypes.c_int64
        _lib.hgm_client_pull_node.restype = ctypes.c_int64
    return _lib


def check(rc, what):
    if rc != 0:
        raise RuntimeError('%s failed: %s' % (what, lib().hgm_last_error().decode()))


class sampler(object):
    """Sampler kinds (reference ``graphmix.sampler``)."""
    GraphSage, RandomWalk, GlobalNode, LocalNode = 0, 1, 2, 3
    names = {0: 'GraphSage', 1: 'RandomWalk', 2: 'GlobalNode', 3: 'LocalNode'}


class cache(object):
    """Remote-node cache policies (reference ``graphmix.cache``)."""
    LRU, LFU, LFUOpt = 0, 1, 2
