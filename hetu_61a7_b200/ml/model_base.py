"""Model base classes, evaluation metrics and the model-file format of the ML library.

reference: third_party/HetuML/hetuml/python/hetuml/model_base.py:10-145 (MLBase /
SupervisedMLBase surface: fit / predict_score / predict_proba / evaluate / save_model /
load_model / name) and hetuml/cc/ml/include/objective/metric.h:100-360 (metric names and
definitions: mse, error, precision (= share of correct predictions), log-loss, cross-entropy,
hinge-loss)."""
from __future__ import annotations

import json
from copy import deepcopy
from typing import Dict, Iterable, Optional, Union

import numpy as np

from .data import CSRMatrix, Dataset

_EPS = 1e-15


def evaluate_metrics(scores: np.ndarray, labels: np.ndarray, metrics: Iterable[str], num_label: int = 2,
                     is_regression: bool = False) -> Dict[str, float]:
    """``scores`` are raw model outputs (margins / logits / regression values)."""
    out = {}
    scores = np.asarray(scores, dtype=np.float64)
    labels = np.asarray(labels, dtype=np.float64).reshape(-1)
    multi = scores.ndim == 2 and scores.shape[1] > 1
    y01 = (labels > 0).astype(np.float64)
    for m in metrics:
        m = m.strip()
        if not m:
            continue
        if m == 'mse':
            v = np.mean((scores.reshape(labels.shape) - labels) ** 2)
        elif m == 'rmse':
            v = np.sqrt(np.mean((scores.reshape(labels.shape) - labels) ** 2))
        elif m in ('error', 'precision'):
            if multi:
                correct = np.argmax(scores, axis=1) == labels.astype(np.int64)
            else:
                correct = (scores.reshape(-1) >= 0) == (y01 > 0.5)
            v = np.mean(correct) if m == 'precision' else 1.0 - np.mean(correct)
        elif m == 'log-loss':
            if multi:
                raise ValueError('log-loss is a binary metric; use cross-entropy')
            p = np.clip(1.0 / (1.0 + np.exp(-scores.reshape(-1))), _EPS, 1 - _EPS)
            v = np.mean(-(y01 * np.log(p) + (1 - y01) * np.log(1 - p)))
        elif m == 'cross-entropy':
            if not multi:
                s = np.stack([np.zeros_like(scores.reshape(-1)), scores.reshape(-1)], axis=1)
                lab = y01.astype(np.int64)
            else:
                s, lab = scores, labels.astype(np.int64)
            s = s - s.max(axis=1, keepdims=True)
            logp = s - np.log(np.exp(s).sum(axis=1, keepdims=True))
            v = np.mean(-logp[np.arange(lab.size), lab])
        elif m == 'hinge-loss':
            ypm = np.where(labels > 0, 1.0, -1.0)
            v = np.mean(np.maximum(0.0, 1.0 - ypm * scores.reshape(-1)))
        else:
            raise ValueError('No such metric: ' + m)
        out[m] = float(v)
    return out


def join_args(aux_args=None, **kwargs):
    res = deepcopy(aux_args) if aux_args else {}
    res.update(kwargs)
    return res


def args_dict_to_string(args: dict) -> str:
    return ' '.join('%s=%s' % (k.upper(), str(v).lower() if isinstance(v, bool) else v) for k, v in args.items())


class MLBase(object):
    _NAME = 'MLBase'

    def __init__(self, args: dict):
        self._args = args
        self._arrays: Dict[str, np.ndarray] = {}

    # -- persistence: one .npz-compatible archive (arrays) with the hyper-parameters as JSON -------------
    def save_model(self, path: str) -> None:
        with open(path, 'wb') as f:
            np.savez(f, __args__=np.frombuffer(json.dumps(self._args).encode(), dtype=np.uint8),
                     __name__=np.frombuffer(self.name().encode(), dtype=np.uint8), **self._arrays)

    def load_model(self, path: str) -> None:
        with np.load(path, allow_pickle=False) as z:
            name = bytes(z['__name__']).decode()
            if name != self.name():
                raise ValueError('%s holds a %s model, not %s' % (path, name, self.name()))
            loaded = json.loads(bytes(z['__args__']).decode())
            self._arrays = {k: np.ascontiguousarray(z[k]) for k in z.files if not k.startswith('__')}
        self._args.update(loaded)
        self._after_load()

    def _after_load(self) -> None:
        pass

    def name(self) -> str:
        return self._NAME

    def get_params(self) -> dict:
        return dict(self._args)

    def __getitem__(self, name):
        if name in self._args:
            return self._args[name]
        if hasattr(self, name):
            return getattr(self, name)
        raise AttributeError('Type %s does not have attribute %s' % (type(self), name))

    # -- distributed helpers (parallel=True): sum over workers through the parameter server -------------
    def _comm(self):
        if not self._args.get('parallel', False):
            return None
        from .cluster import current_comm
        return current_comm()


def _as_features(pred_data) -> CSRMatrix:
    if isinstance(pred_data, Dataset):
        return pred_data.X
    if isinstance(pred_data, CSRMatrix):
        return pred_data
    if isinstance(pred_data, tuple):
        return Dataset.from_data(pred_data).X
    return CSRMatrix.from_data(pred_data)


class SupervisedMLBase(MLBase):

    def fit(self, train_data, valid_data=None) -> None:
        train = Dataset.from_data(train_data)
        valid = Dataset.from_data(valid_data) if valid_data is not None else None
        self._fit(train, valid)

    def _fit(self, train: Dataset, valid: Optional[Dataset]) -> None:
        raise NotImplementedError

    def _scores(self, X: CSRMatrix) -> np.ndarray:
        raise NotImplementedError

    def predict_score(self, pred_data) -> np.ndarray:
        scores = self._scores(_as_features(pred_data))
        if self.is_classification() and self.get_num_label() > 2:
            scores = scores.reshape(-1, self.get_num_label())
        return scores

    def predict_proba(self, pred_data) -> np.ndarray:
        scores = self.predict_score(pred_data)
        if self.is_regression():
            return scores
        if scores.ndim == 1:
            return 1 / (1 + np.exp(-scores))
        tmp = np.exp(scores - np.max(scores, axis=1).reshape(-1, 1))
        return tmp / tmp.sum(axis=1).reshape(-1, 1)

    def predict(self, pred_data) -> np.ndarray:
        scores = self.predict_score(pred_data)
        if self.is_regression():
            return scores
        if scores.ndim == 1:
            pos = scores >= 0
            return np.where(pos, 1, -1 if self.use_neg_y() else 0)
        return np.argmax(scores, axis=1)

    def evaluate(self, eval_data, metrics: Union[str, Iterable[str]]) -> Dict[str, float]:
        data = Dataset.from_data(eval_data)
        names = [m for m in (metrics.split(',') if isinstance(metrics, str) else list(metrics)) if m.strip()]
        res = evaluate_metrics(self.predict_score(data), data.y, names, self.get_num_label(), self.is_regression())
        comm = self._comm()
        if comm is not None:                 # instance-weighted mean over the workers' shards
            n = float(data.get_num_instances())
            buf = np.array([res[m] * n for m in names] + [n], dtype=np.float64)
            comm.allreduce(buf)
            res = {m: buf[i] / max(buf[-1], 1.0) for i, m in enumerate(names)}
        return res

    def _log_metrics(self, tag: str, train: Dataset, valid: Optional[Dataset]) -> None:
        names = self._args.get('metrics', '')
        if not names or not self._args.get('verbose', False):
            return
        msg = '%s %s train %s' % (self.name(), tag, self.evaluate(train, names))
        if valid is not None and valid.get_num_instances():
            msg += ' valid %s' % self.evaluate(valid, names)
        print(msg, flush=True)

    def use_neg_y(self) -> bool:
        return self._args.get('use_neg_y', False)

    def is_regression(self) -> bool:
        return self._args.get('is_regression', False)

    def is_classification(self) -> bool:
        return not self.is_regression()

    def get_num_label(self) -> int:
        return self._args.get('num_label', 2)
