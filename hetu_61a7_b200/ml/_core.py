"""ctypes loader for lib/libhetu_ml.so (the classical-ML core, csrc_ml/ml.cc)."""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_lib = None

I64P = ctypes.POINTER(ctypes.c_int64)
I32P = ctypes.POINTER(ctypes.c_int32)
F32P = ctypes.POINTER(ctypes.c_float)
F64P = ctypes.POINTER(ctypes.c_double)
ALLREDUCE = ctypes.CFUNCTYPE(None, F64P, ctypes.c_int64)


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        path = os.path.join(_HERE, 'lib', 'libhetu_ml.so')
        if not os.path.exists(path):
            raise RuntimeError('libhetu_ml.so is not built; run `python -c "import __graft_entry__ as g; g.build()"`')
        _lib = ctypes.CDLL(path)
        _lib.hml_last_error.restype = ctypes.c_char_p
        _lib.hml_gbdt_create.restype = ctypes.c_void_p
        _lib.hml_gbdt_max_dim.restype = ctypes.c_int64
        for name in ('hml_gbdt_free', 'hml_gbdt_fit', 'hml_gbdt_score_dim', 'hml_gbdt_num_trees', 'hml_gbdt_max_dim',
                     'hml_gbdt_round_metric', 'hml_gbdt_get_params', 'hml_gbdt_predict', 'hml_gbdt_save', 'hml_gbdt_load'):
            getattr(_lib, name).argtypes = None
    return _lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError('%s failed: %s' % (what, lib().hml_last_error().decode()))
