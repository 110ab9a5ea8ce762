"""Gaussian naive Bayes over sparse features.

reference: third_party/HetuML/hetuml/python/hetuml/naive_bayes/naive_bayes.py:6-24 and
hetuml/cc/ml/include/model/naive_bayes/{naive_bayes,parallel_naive_bayes}.h.  Scores are log joint
probabilities, so ``predict_proba`` (softmax of the scores) is the class posterior."""
from __future__ import annotations

import ctypes
from typing import Optional

import numpy as np

from . import _core as C
from .data import CSRMatrix, Dataset
from .model_base import SupervisedMLBase, join_args


class NaiveBayes(SupervisedMLBase):
    ps_data_type = 'float'
    _NAME = 'NaiveBayes'

    def __init__(self, num_label: int, metrics: str = 'cross-entropy', parallel=False, aux_args=None):
        assert num_label > 1, 'Number of labels should be greater than 1'
        super().__init__(join_args(aux_args, num_label=num_label, metrics=metrics, parallel=parallel))

    def predict_score(self, pred_data):
        return super().predict_score(pred_data).reshape(-1, self.get_num_label())

    def _fit(self, train: Dataset, valid: Optional[Dataset]) -> None:
        comm = self._comm()
        K, dim = self.get_num_label(), train.get_max_dim()
        if comm is not None:
            dim = int(comm.allreduce_max(np.array([dim], dtype=np.float64))[0])
        if train.y.size and (train.y.min() < 0 or train.y.max() >= K):
            raise ValueError('labels must lie in [0, %d)' % K)
        stats = np.zeros(3 * dim * K + K, dtype=np.float64)
        ip, idx, val = train.X.ptrs()
        C.check(C.lib().hml_nb_stats(ip, idx, val, train.y.ctypes.data_as(C.F32P), ctypes.c_int64(train.get_num_instances()),
                                     ctypes.c_int64(dim), K, stats.ctypes.data_as(C.F64P)), 'NaiveBayes statistics')
        if comm is not None:
            comm.allreduce(stats)
        model = np.zeros(K + 2 * dim * K, dtype=np.float32)
        C.check(C.lib().hml_nb_finalize(stats.ctypes.data_as(C.F64P), ctypes.c_int64(dim), K,
                                        model.ctypes.data_as(C.F32P)), 'NaiveBayes finalize')
        self._arrays['model'] = model
        self._args['max_dim'] = dim
        self._log_metrics('fit', train, valid)

    def _scores(self, X: CSRMatrix) -> np.ndarray:
        K, dim = self.get_num_label(), int(self._args['max_dim'])
        model = self._arrays['model']
        out = np.empty(X.get_num_rows() * K, dtype=np.float32)
        ip, idx, val = X.ptrs()
        C.check(C.lib().hml_nb_predict(ip, idx, val, ctypes.c_int64(X.get_num_rows()), ctypes.c_int64(dim), K,
                                       model.ctypes.data_as(C.F32P), out.ctypes.data_as(C.F32P)), 'NaiveBayes predict')
        return out
