"""Factorization machine (second-order, logistic loss).

reference: third_party/HetuML/hetuml/python/hetuml/factorization_machines/fm.py:6-45 and
hetuml/cc/ml/include/model/fm/{fm,parallel_fm}.h."""
from __future__ import annotations

import ctypes
from typing import Optional

import numpy as np

from . import _core as C
from .data import CSRMatrix, Dataset
from .model_base import SupervisedMLBase, join_args


class FM(SupervisedMLBase):
    ps_data_type = 'float'
    _NAME = 'FactorizationMachine'

    def __init__(self, embedding_dim: int = 16, num_epoch: int = 10, batch_size: int = 1000, learning_rate: float = 0.1,
                 l1_reg: float = 0.0, l2_reg: float = 0.0, loss: str = 'logistic', metrics: str = 'log-loss',
                 parallel=False, aux_args=None):
        assert embedding_dim > 0, 'Embedding dimension should be positive'
        assert num_epoch > 0, 'Number of epochs should be positive'
        assert batch_size > 0, 'Batch size should be positive'
        assert learning_rate > 0, 'Learning rate should be positive'
        assert l1_reg >= 0 and l2_reg >= 0, 'regularization terms should be non-negative'
        assert loss in ('logistic',), 'Loss %s is not allowed for %s' % (loss, type(self))
        super().__init__(join_args(aux_args, embedding_dim=embedding_dim, num_epoch=num_epoch, batch_size=batch_size,
                                   learning_rate=learning_rate, l1_reg=l1_reg, l2_reg=l2_reg, loss=loss, metrics=metrics,
                                   parallel=parallel, use_neg_y=True))

    def use_neg_y(self):
        return True

    def _fit(self, train: Dataset, valid: Optional[Dataset]) -> None:
        a = self._args
        comm = self._comm()
        dim, k = train.get_max_dim(), int(a['embedding_dim'])
        if comm is not None:
            dim = int(comm.allreduce_max(np.array([dim], dtype=np.float64))[0])
        rng = np.random.RandomState(int(a.get('seed', 0)))
        w = np.zeros(dim, dtype=np.float32)
        v = (rng.standard_normal((dim, k)) * 0.01).astype(np.float32)     # same seed -> same start on every worker
        y = train.labels_pm1()
        ip, idx, val = train.X.ptrs()
        step = 1 if (comm is not None or a.get('verbose', False)) else a['num_epoch']
        for ep in range(0, a['num_epoch'], step):
            C.check(C.lib().hml_fm_fit(ip, idx, val, y.ctypes.data_as(C.F32P), ctypes.c_int64(train.get_num_instances()),
                                       ctypes.c_int64(dim), k, w.ctypes.data_as(C.F32P), v.ctypes.data_as(C.F32P), step,
                                       int(a['batch_size']), ctypes.c_float(a['learning_rate']),
                                       ctypes.c_float(a['l1_reg']), ctypes.c_float(a['l2_reg'])), 'FM fit')
            if comm is not None:
                w = comm.average(w)
                v = comm.average(v.reshape(-1)).reshape(dim, k)
            self._arrays.update(w=w, v=v)
            self._log_metrics('epoch %d' % (ep + step), train, valid)
        self._arrays.update(w=w, v=v)

    def _scores(self, X: CSRMatrix) -> np.ndarray:
        w, v = self._arrays['w'], np.ascontiguousarray(self._arrays['v'])
        out = np.empty(X.get_num_rows(), dtype=np.float32)
        ip, idx, val = X.ptrs()
        C.check(C.lib().hml_fm_predict(ip, idx, val, ctypes.c_int64(X.get_num_rows()), ctypes.c_int64(w.size),
                                       int(v.shape[1]), w.ctypes.data_as(C.F32P), v.ctypes.data_as(C.F32P),
                                       out.ctypes.data_as(C.F32P)), 'FM predict')
        return out
