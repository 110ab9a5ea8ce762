"""Linear models: logistic regression, linear SVM, linear (least-squares) regression.

reference: third_party/HetuML/hetuml/python/hetuml/linear/{log_reg,svm,linear_reg}.py and
hetuml/cc/ml/include/model/linear/{linear,parallel_linear}.h.  ``parallel=True`` trains
data-parallel: each worker runs an epoch on its shard and the weights are averaged through the
parameter server."""
from __future__ import annotations

import ctypes
from typing import Optional

import numpy as np

from . import _core as C
from .data import CSRMatrix, Dataset
from .model_base import SupervisedMLBase, join_args


class _LinearModel(SupervisedMLBase):
    _LOSS = 0
    _LOSSES = ()

    def __init__(self, num_epoch, batch_size, learning_rate, l1_reg, l2_reg, loss, metrics, parallel, aux_args, **extra):
        assert num_epoch > 0, 'Number of epochs should be positive'
        assert batch_size > 0, 'Batch size should be positive'
        assert learning_rate > 0, 'Learning rate should be positive'
        assert l1_reg >= 0, 'L1 regularization term should be non-negative'
        assert l2_reg >= 0, 'L2 regularization term should be non-negative'
        assert loss in self._LOSSES, 'Loss %s is not allowed for %s' % (loss, type(self))
        super().__init__(join_args(aux_args, num_epoch=num_epoch, batch_size=batch_size, learning_rate=learning_rate,
                                   l1_reg=l1_reg, l2_reg=l2_reg, loss=loss, metrics=metrics, parallel=parallel, **extra))

    def _targets(self, data: Dataset) -> np.ndarray:
        return data.labels_pm1() if self.use_neg_y() else data.y

    def _fit(self, train: Dataset, valid: Optional[Dataset]) -> None:
        a = self._args
        comm = self._comm()
        dim = train.get_max_dim()
        if comm is not None:
            dim = int(comm.allreduce_max(np.array([dim], dtype=np.float64))[0])
        w = np.zeros(dim, dtype=np.float32)
        y = self._targets(train)
        ip, idx, val = train.X.ptrs()
        epochs_per_call = 1 if (comm is not None or a.get('verbose', False)) else a['num_epoch']
        for ep in range(0, a['num_epoch'], epochs_per_call):
            C.check(C.lib().hml_linear_fit(self._LOSS, ip, idx, val, y.ctypes.data_as(C.F32P),
                                           ctypes.c_int64(train.get_num_instances()), ctypes.c_int64(dim),
                                           w.ctypes.data_as(C.F32P), epochs_per_call, int(a['batch_size']),
                                           ctypes.c_float(a['learning_rate']), ctypes.c_float(a['l1_reg']),
                                           ctypes.c_float(a['l2_reg'])), 'linear fit')
            if comm is not None:
                w = comm.average(w)
            self._arrays['w'] = w
            self._log_metrics('epoch %d' % (ep + epochs_per_call), train, valid)
        self._arrays['w'] = w

    def _scores(self, X: CSRMatrix) -> np.ndarray:
        w = self._arrays['w']
        out = np.empty(X.get_num_rows(), dtype=np.float32)
        ip, idx, val = X.ptrs()
        C.check(C.lib().hml_linear_predict(ip, idx, val, ctypes.c_int64(X.get_num_rows()), ctypes.c_int64(w.size),
                                           w.ctypes.data_as(C.F32P), out.ctypes.data_as(C.F32P)), 'linear predict')
        return out

    @property
    def weights(self) -> np.ndarray:
        return self._arrays['w']


class LogisticRegression(_LinearModel):
    ps_data_type = 'float'
    _NAME, _LOSS, _LOSSES = 'LogisticRegression', 0, ('logistic',)

    def __init__(self, num_epoch: int = 10, batch_size: int = 1000, learning_rate: float = 0.1, l1_reg: float = 0.0,
                 l2_reg: float = 0.0, loss: str = 'logistic', metrics: str = 'log-loss', parallel: bool = False,
                 aux_args=None):
        super().__init__(num_epoch, batch_size, learning_rate, l1_reg, l2_reg, loss, metrics, parallel, aux_args,
                         use_neg_y=True)

    def use_neg_y(self):
        return True


class SVM(_LinearModel):
    ps_data_type = 'float'
    _NAME, _LOSS, _LOSSES = 'SupportVectorMachine', 1, ('hinge',)

    def __init__(self, num_epoch: int = 10, batch_size: int = 1000, learning_rate: float = 0.1, l1_reg: float = 0.0,
                 l2_reg: float = 0.0, loss: str = 'hinge', metrics: str = 'hinge-loss', parallel: bool = False,
                 aux_args=None):
        super().__init__(num_epoch, batch_size, learning_rate, l1_reg, l2_reg, loss, metrics, parallel, aux_args,
                         use_neg_y=True)

    def use_neg_y(self):
        return True


class LinearRegression(_LinearModel):
    ps_data_type = 'float'
    _NAME, _LOSS, _LOSSES = 'LinearRegression', 2, ('square',)

    def __init__(self, num_epoch: int = 10, batch_size: int = 1000, learning_rate: float = 0.1, l1_reg: float = 0.0,
                 l2_reg: float = 0.0, loss: str = 'square', metrics: str = 'mse', parallel: bool = False, aux_args=None):
        super().__init__(num_epoch, batch_size, learning_rate, l1_reg, l2_reg, loss, metrics, parallel, aux_args,
                         is_regression=True)
