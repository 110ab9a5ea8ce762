"""ctypes loader for the in-tree native libraries.

 * ``lib/libhetu_b200.so``  - sm_100a CUDA kernels, C ABI declared in
   ``include/hetu_b200.h`` (built by ``__graft_entry__.build()`` with
   ``nvcc -gencode arch=compute_100a,code=sm_100a``).
 * ``lib/libhetu_cpu.so``   - OpenMP CPU op library (oracle + CPU executor).
 * ``lib/libhetu_ps.so``    - parameter server, HET cache, TCP van.

The reference binds its ops the same way (ctypes over a C ABI,
python/hetu/_base.py:61-68); what is behind the ABI is new.  On a box with a
GPU the CUDA library is mandatory: ops raise instead of silently falling back.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_DIR = os.path.join(_HERE, 'lib')

_cuda_lib: Optional[ctypes.CDLL] = None
_cpu_lib: Optional[ctypes.CDLL] = None
_ps_lib: Optional[ctypes.CDLL] = None
_tried = {'cuda': False, 'cpu': False, 'ps': False}

#: number of native kernel launches issued through this module (bench.py reports it)
launch_count = 0


def _load(name: str) -> Optional[ctypes.CDLL]:
    path = os.path.join(LIB_DIR, name)
    if not os.path.exists(path):
        return None
    return ctypes.CDLL(path, mode=ctypes.RTLD_GLOBAL)


def cuda_lib() -> Optional[ctypes.CDLL]:
    global _cuda_lib
    if not _tried['cuda']:
        _tried['cuda'] = True
        if torch.cuda.is_available():
            _cuda_lib = _load('libhetu_b200.so')
            if _cuda_lib is not None:
                _cuda_lib.hetu_last_error.restype = ctypes.c_char_p
    return _cuda_lib


def cpu_lib() -> Optional[ctypes.CDLL]:
    global _cpu_lib
    if not _tried['cpu']:
        _tried['cpu'] = True
        _cpu_lib = _load('libhetu_cpu.so')
    return _cpu_lib


def ps_lib() -> Optional[ctypes.CDLL]:
    global _ps_lib
    if not _tried['ps']:
        _tried['ps'] = True
        _ps_lib = _load('libhetu_ps.so')
    return _ps_lib


def require_cuda_lib() -> ctypes.CDLL:
    lib = cuda_lib()
    if lib is None:
        raise RuntimeError(
            'libhetu_b200.so is not built or no GPU is visible; run '
            '`python -c "import __graft_entry__ as g; g.build()"` from the repo root')
    return lib


_FORCE_TORCH = os.environ.get('HETU_FORCE_TORCH', '0') == '1'


def kernels_on(*tensors) -> bool:
    """True when every tensor is a real CUDA tensor, i.e. the native kernel path
    must be taken.  Raises if a GPU is present but the library is not."""
    if _FORCE_TORCH:
        return False
    ok = False
    for t in tensors:
        if t is None:
            continue
        if not isinstance(t, torch.Tensor) or t.device.type != 'cuda':
            return False
        ok = True
    if ok:
        require_cuda_lib()
    return ok


def stream_ptr(device=None) -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def p(t: Optional[torch.Tensor]) -> ctypes.c_void_p:
    return ctypes.c_void_p(t.data_ptr() if t is not None else 0)


def check(rc: int, what: str = '') -> None:
    if rc != 0:
        msg = _cuda_lib.hetu_last_error().decode() if _cuda_lib is not None else ''
        raise RuntimeError('native kernel %s failed (rc=%d): %s' % (what, rc, msg))


def call(fn_name: str, *args) -> None:
    """Invoke ``hetu_<fn_name>`` on the current stream of the current device."""
    global launch_count
    lib = require_cuda_lib()
    fn = getattr(lib, 'hetu_' + fn_name)
    rc = fn(*args)
    launch_count += 1
    if rc != 0:
        check(rc, fn_name)


c_int = ctypes.c_int
c_i64 = ctypes.c_int64
c_f32 = ctypes.c_float
c_u64 = ctypes.c_uint64
c_sz = ctypes.c_size_t
