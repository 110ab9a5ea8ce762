"""Gradient boosted decision trees and random forest (histogram-based trees over sparse features).

reference: third_party/HetuML/hetuml/python/hetuml/ensemble/{gbdt,random_boosted_forest}.py and
hetuml/cc/ml/include/model/{gbdt,rf,tree}/**.  ``parallel=True`` is data parallel: workers hold
row shards, agree on feature ranges, and sum node histograms through the parameter server, so all
of them grow the same trees."""
from __future__ import annotations

import ctypes
from typing import Optional

import numpy as np

from . import _core as C
from .data import CSRMatrix, Dataset
from .model_base import SupervisedMLBase, join_args

_ORDER = ('max_depth', 'max_node_num', 'leafwise', 'num_split', 'ins_sp_ratio', 'feat_sp_ratio', 'is_regression',
          'num_label', 'num_round', 'learning_rate', 'min_split_gain', 'min_node_ins', 'min_child_weight', 'reg_alpha',
          'reg_lambda', 'max_leaf_weight', 'multi_tree', 'majority_voting', 'seed')


class GBDT(SupervisedMLBase):
    ps_data_type = 'double'
    _NAME = 'GBDT'
    _MAJORITY = False

    def __init__(self, max_depth: int = 4, max_node_num: int = -1, leafwise: bool = False, num_split: int = 10,
                 ins_sp_ratio: float = 1.0, feat_sp_ratio: float = 1.0, is_regression: bool = False, num_label: int = 2,
                 num_round: int = 10, learning_rate: float = 0.1, min_split_gain: float = 0.0, min_node_ins: int = 1024,
                 min_child_weight: float = 0, reg_alpha: float = 0, reg_lambda: float = 0, max_leaf_weight: float = 0,
                 multi_tree: bool = False, loss: str = 'auto', metrics: str = 'auto', parallel=False, aux_args=None,
                 device: str = 'auto'):
        """``device``: 'cpu' (native OpenMP trainer), 'gpu' (level-at-a-time trainer on the CUDA kernels of
        csrc/gbdt.cu; level-wise scalar-leaf trees) or 'auto' (GPU for large inputs when one is present)."""
        assert device in ('auto', 'cpu', 'gpu')
        assert max_depth > 0, 'Maximum tree depth should be positive'
        if max_node_num == -1:
            max_node_num = (2 ** (max_depth + 1)) - 1
        assert max_node_num > 0, 'Maximum number of tree nodes should be positive'
        assert 0 < num_split <= 254, 'Invalid number of splits: ' + str(num_split)
        assert 0 < ins_sp_ratio <= 1, 'Invalid sampling ratio for instances: ' + str(ins_sp_ratio)
        assert 0 < feat_sp_ratio <= 1, 'Invalid sampling ratio for features: ' + str(feat_sp_ratio)
        assert num_label > 1, 'Number of labels should be greater than 1'
        assert num_round > 0, 'Number of rounds should be positive'
        assert learning_rate > 0, 'Learning rate should be positive'
        assert min_split_gain >= 0 and min_node_ins >= 0 and min_child_weight >= 0
        assert reg_alpha >= 0 and reg_lambda >= 0 and max_leaf_weight >= 0
        if loss == 'auto':
            loss = 'square' if is_regression else 'logistic'
        if metrics == 'auto':
            metrics = 'mse' if is_regression else ('log-loss' if num_label == 2 else 'cross-entropy')
        args = join_args(aux_args, max_depth=max_depth, max_node_num=max_node_num, leafwise=leafwise, num_split=num_split,
                         ins_sp_ratio=ins_sp_ratio, feat_sp_ratio=feat_sp_ratio, is_regression=is_regression,
                         num_label=num_label, num_round=num_round, learning_rate=learning_rate,
                         min_split_gain=min_split_gain, min_node_ins=min_node_ins, min_child_weight=min_child_weight,
                         reg_alpha=reg_alpha, reg_lambda=reg_lambda, max_leaf_weight=max_leaf_weight, multi_tree=multi_tree,
                         loss=loss, metrics=metrics, parallel=parallel, majority_voting=self._MAJORITY)
        args.setdefault('seed', 42)
        args['device'] = device
        super().__init__(args)
        self._handle = None

    def __del__(self):
        try:
            self._free()
        except Exception:                    # interpreter shutdown: module globals may already be gone
            pass

    def _free(self):
        if getattr(self, '_handle', None):
            C.lib().hml_gbdt_free(ctypes.c_void_p(self._handle))
            self._handle = None

    def _new_handle(self):
        self._free()
        vals = (ctypes.c_double * len(_ORDER))(*[float(self._args[k]) for k in _ORDER])
        self._handle = C.lib().hml_gbdt_create(vals, len(_ORDER))

    def _on_gpu(self, train: Dataset) -> bool:
        mode = self._args.get('device', 'auto')
        if mode == 'cpu' or self._args.get('parallel', False):
            return False
        import torch
        from . import _gbdt_gpu
        if not torch.cuda.is_available() or not _gbdt_gpu.supported(self._args):
            if mode == 'gpu':
                raise RuntimeError("GBDT(device='gpu') needs a CUDA device and level-wise scalar-leaf trees "
                                   "(leafwise=False; multi_tree=True for more than two labels)")
            return False
        return mode == 'gpu' or train.X.indices.size >= (1 << 22)

    def _fit(self, train: Dataset, valid: Optional[Dataset]) -> None:
        self._gpu_curve = None
        if self._on_gpu(train):
            from . import _gbdt_gpu
            _gbdt_gpu.fit(self, train, train.get_max_dim())
            self._log_metrics('fit', train, valid)
            return
        comm = self._comm()
        self._new_handle()
        dim = train.get_max_dim()
        y = train.y if (self.is_regression() or self.get_num_label() > 2) else train.labels01()
        ip, idx, val = train.X.ptrs()
        cb, rng_ptr, keep = None, None, None
        if comm is not None:
            dim = int(comm.allreduce_max(np.array([dim], dtype=np.float64))[0])
            lo, hi = np.zeros(dim), np.zeros(dim)
            if train.X.indices.size:
                np.minimum.at(lo, train.X.indices, train.X.data)
                np.maximum.at(hi, train.X.indices, train.X.data)
            hi = comm.allreduce_max(hi)
            lo = -comm.allreduce_max(-lo)
            keep = np.ascontiguousarray(np.stack([lo, hi], axis=1).reshape(-1), dtype=np.float64)
            rng_ptr = keep.ctypes.data_as(C.F64P)

            def _ar(buf, n):
                arr = np.ctypeslib.as_array(buf, shape=(n,))
                comm.allreduce(arr)
            cb = C.ALLREDUCE(_ar)
        C.check(C.lib().hml_gbdt_fit(ctypes.c_void_p(self._handle), ip, idx, val, y.ctypes.data_as(C.F32P),
                                     ctypes.c_int64(train.get_num_instances()), ctypes.c_int64(dim), cb, rng_ptr), 'GBDT fit')
        self._log_metrics('fit', train, valid)

    def train_loss_curve(self) -> np.ndarray:
        """Mean training loss after every boosting round."""
        if getattr(self, '_gpu_curve', None) is not None:
            return self._gpu_curve
        out = (ctypes.c_double * int(self._args['num_round']))()
        n = C.lib().hml_gbdt_round_metric(ctypes.c_void_p(self._handle), out, len(out))
        return np.array(out[:n])

    def num_trees(self) -> int:
        return int(C.lib().hml_gbdt_num_trees(ctypes.c_void_p(self._handle))) if self._handle else 0

    def _scores(self, X: CSRMatrix) -> np.ndarray:
        if not self._handle:
            raise RuntimeError('Model is empty')
        k = int(C.lib().hml_gbdt_score_dim(ctypes.c_void_p(self._handle)))
        out = np.empty(X.get_num_rows() * k, dtype=np.float32)
        ip, idx, val = X.ptrs()
        C.check(C.lib().hml_gbdt_predict(ctypes.c_void_p(self._handle), ip, idx, val, ctypes.c_int64(X.get_num_rows()),
                                         out.ctypes.data_as(C.F32P)), 'GBDT predict')
        return out

    # trees are written by the native core as text (one line per node)
    def save_model(self, path: str) -> None:
        C.check(C.lib().hml_gbdt_save(ctypes.c_void_p(self._handle), str(path).encode()), 'GBDT save')

    def load_model(self, path: str) -> None:
        self._new_handle()
        C.check(C.lib().hml_gbdt_load(ctypes.c_void_p(self._handle), str(path).encode()), 'GBDT load')
        vals = (ctypes.c_double * len(_ORDER))()
        C.lib().hml_gbdt_get_params(ctypes.c_void_p(self._handle), vals)
        for k, v in zip(_ORDER, vals):
            old = self._args.get(k)
            self._args[k] = bool(v) if isinstance(old, bool) else (int(v) if isinstance(old, int) else float(v))


class RandomForest(GBDT):
    """Bagged trees: every tree is fitted independently on an instance / feature sample and the
    forest averages them (reference rf.h: GBDT trainer with IS_MAJORITY_VOTING)."""
    _NAME = 'RandomForest'
    _MAJORITY = True

    def __init__(self, max_depth: int = 4, max_node_num: int = -1, leafwise: bool = False, num_split: int = 10,
                 ins_sp_ratio: float = 0.3, feat_sp_ratio: float = 0.3, is_regression: bool = False, num_label: int = 2,
                 num_round: int = 100, learning_rate: float = 0.1, min_split_gain: float = 0.0, min_node_ins: int = 1024,
                 min_child_weight: float = 0, reg_alpha: float = 0, reg_lambda: float = 0, max_leaf_weight: float = 0,
                 multi_tree: bool = False, loss: str = 'auto', metrics: str = 'auto', parallel=False, aux_args=None,
                 device: str = 'auto'):
        super().__init__(max_depth, max_node_num, leafwise, num_split, ins_sp_ratio, feat_sp_ratio, is_regression, num_label,
                         num_round, learning_rate, min_split_gain, min_node_ins, min_child_weight, reg_alpha, reg_lambda,
                         max_leaf_weight, multi_tree, loss, metrics, parallel, aux_args, device)


RanomBoostedForest = RandomForest          # the reference's (misspelt) class name
RandomBoostedForest = RandomForest
