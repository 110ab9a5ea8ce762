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

    def __init__(self, num_label: int, num_neighbor: int = 5, device: str = 'auto', aux_args=None):
        """``device``: 'cpu' (sparse brute force in the native core), 'gpu' (dense distance matrix
        ``|q|^2 + |l|^2 - 2 Q L^T`` on the framework's tcgen05 GEMM, top-k on the device) or 'auto'
        (GPU when one is present and the dense matrices fit comfortably)."""
        assert num_label > 1, 'Number of labels should be greater than 1'
        assert num_neighbor > 0, 'Number of neighbors should be positive'
        assert device in ('auto', 'cpu', 'gpu')
        super().__init__(join_args(aux_args, num_label=num_label, num_neighbor=num_neighbor, parallel=False, device=device))

    def _use_gpu(self, n_lab, n_query, dim):
        import torch
        mode = self._args.get('device', 'auto')
        if mode == 'cpu' or not torch.cuda.is_available():
            if mode == 'gpu':
                raise RuntimeError("KNN(device='gpu') needs a CUDA device")
            return False
        dense_bytes = 4 * (n_lab + n_query) * dim + 4 * min(n_query, 8192) * n_lab
        return mode == 'gpu' or (dense_bytes < (8 << 30) and n_lab * n_query * dim > (1 << 24))

    def _votes_gpu(self, lab, Q, dim, k, nl):
        import torch
        from ..ops.linalg import gemm, get_matmul_precision, set_matmul_precision
        dev = torch.device('cuda')

        def dense(X):
            rows = np.repeat(np.arange(X.get_num_rows()), np.diff(X.indptr))
            out = torch.zeros((X.get_num_rows(), dim), dtype=torch.float32, device=dev)
            out[torch.from_numpy(rows).to(dev), torch.from_numpy(X.indices.astype(np.int64)).to(dev)] = \
                torch.from_numpy(X.data).to(dev)
            return out
        L, labels = dense(lab.X), torch.from_numpy(lab.y.astype(np.int64)).to(dev)
        ln = (L * L).sum(1)
        votes = torch.zeros((Q.get_num_rows(), nl), dtype=torch.float32, device=dev)
        kk = min(k, L.shape[0])
        prev = get_matmul_precision()
        set_matmul_precision('fp32' if dim <= 64 else 'tf32')       # tiny K: exact FFMA path costs nothing
        try:
            for b0 in range(0, Q.get_num_rows(), 8192):                # bound the [chunk, n_labeled] distance block
                q = dense(Q.rows(b0, min(b0 + 8192, Q.get_num_rows())))
                d = (q * q).sum(1, keepdim=True) + ln.unsqueeze(0) - 2.0 * gemm(q, L, trans_b=True)
                idx = torch.topk(d, kk, dim=1, largest=False).indices
                votes[b0:b0 + q.shape[0]].scatter_add_(1, labels[idx].clamp_(0, nl - 1), torch.ones_like(idx, dtype=torch.float32))
        finally:
            set_matmul_precision(prev)
        return votes.cpu().numpy()

    def predict_score(self, labeled_data, pred_data) -> np.ndarray:
        """[n, num_label] neighbour votes per label."""
        lab = Dataset.from_data(labeled_data)
        Q = _as_features(pred_data)
        k, nl = int(self._args['num_neighbor']), int(self._args['num_label'])
        dim = max(lab.get_max_dim(), Q.get_num_cols())
        if self._use_gpu(lab.get_num_instances(), Q.get_num_rows(), dim):
            return self._votes_gpu(lab, Q, dim, k, nl)
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
