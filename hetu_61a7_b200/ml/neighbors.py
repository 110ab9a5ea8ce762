"""k-nearest-neighbour classifier (brute force over sparse rows).

reference: third_party/HetuML/hetuml/python/hetuml/neighbors/knn.py:10-70 and
hetuml/cc/ml/include/model/knn/knn.h."""
from __future__ import annotations

import ctypes
from typing import Iterable, Union

import numpy as np

from . import _core as C
from .data import Dataset
from .model_base import MLBase, _as_features, evaluate_metrics, join_args


class KNN(MLBase):
    _NAME = 'KNN'

    def __init__(self, num_label: int, num_neighbor: int = 5, aux_args=None):
        assert num_label > 1, 'Number of labels should be greater than 1'
        assert num_neighbor > 0, 'Number of neighbors should be positive'
        super().__init__(join_args(aux_args, num_label=num_label, num_neighbor=num_neighbor, parallel=False))

    def predict_score(self, labeled_data, pred_data) -> np.ndarray:
        """[n, num_label] neighbour votes per label."""
        lab = Dataset.from_data(labeled_data)
        Q = _as_features(pred_data)
        k, nl = int(self._args['num_neighbor']), int(self._args['num_label'])
        dim = max(lab.get_max_dim(), Q.get_num_cols())
        out = np.empty((Q.get_num_rows(), nl), dtype=np.float32)
        lp, li, lv = lab.X.ptrs()
        qp, qi, qv = Q.ptrs()
        C.check(C.lib().hml_knn_predict(lp, li, lv, lab.y.ctypes.data_as(C.F32P), ctypes.c_int64(lab.get_num_instances()),
                                        qp, qi, qv, ctypes.c_int64(Q.get_num_rows()), ctypes.c_int64(dim), k, nl,
                                        out.ctypes.data_as(C.F32P)), 'KNN predict')
        return out

    def predict_proba(self, labeled_data, pred_data) -> np.ndarray:
        scores = self.predict_score(labeled_data, pred_data)
        tmp = np.exp(scores - np.max(scores, axis=1).reshape(-1, 1))
        return tmp / tmp.sum(axis=1).reshape(-1, 1)

    def predict(self, labeled_data, pred_data) -> np.ndarray:
        return np.argmax(self.predict_score(labeled_data, pred_data), axis=1)

    def evaluate(self, labeled_data, eval_data, metrics: Union[str, Iterable[str]]):
        ev = Dataset.from_data(eval_data)
        names = metrics.split(',') if isinstance(metrics, str) else list(metrics)
        return evaluate_metrics(self.predict_score(labeled_data, ev), ev.y, names, int(self._args['num_label']))
