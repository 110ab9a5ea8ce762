"""numpy metrics.  reference: python/hetu/metrics.py:4-359 (TF-style AUC with 200
thresholds, accuracy, confusion matrix, precision / recall / F-score)."""
from __future__ import annotations

import numpy as np


def _np(x):
    return x.asnumpy() if hasattr(x, 'asnumpy') else np.asarray(x)


def softmax_func(y):
    b = y - np.max(y, axis=-1, keepdims=True)
    e = np.exp(b)
    return e / np.sum(e, axis=-1, keepdims=True)


def accuracy(y_pred, y_true):
    y_pred, y_true = _np(y_pred), _np(y_true)
    if y_pred.ndim > 1 and y_pred.shape[-1] > 1:
        p = y_pred.argmax(-1)
        t = y_true.argmax(-1) if y_true.ndim == y_pred.ndim else y_true.astype(np.int64)
    else:
        p = (y_pred.reshape(-1) > 0.5).astype(np.int64)
        t = y_true.reshape(-1).astype(np.int64)
    return float((p == t).mean())


def confusion_matrix(y_pred, y_true, num_classes=None):
    p = _np(y_pred).astype(np.int64).reshape(-1)
    t = _np(y_true).astype(np.int64).reshape(-1)
    n = num_classes or int(max(p.max(), t.max())) + 1
    cm = np.zeros((n, n), dtype=np.int64)
    np.add.at(cm, (t, p), 1)
    return cm


def precision(y_pred, y_true, average='binary', num_classes=None):
    cm = confusion_matrix(y_pred, y_true, num_classes)
    tp = np.diag(cm).astype(np.float64)
    denom = cm.sum(axis=0).astype(np.float64)
    per = np.divide(tp, denom, out=np.zeros_like(tp), where=denom > 0)
    if average == 'binary':
        return float(per[1]) if len(per) > 1 else float(per[0])
    if average == 'micro':
        return float(tp.sum() / max(cm.sum(), 1))
    return float(per.mean())


def recall(y_pred, y_true, average='binary', num_classes=None):
    cm = confusion_matrix(y_pred, y_true, num_classes)
    tp = np.diag(cm).astype(np.float64)
    denom = cm.sum(axis=1).astype(np.float64)
    per = np.divide(tp, denom, out=np.zeros_like(tp), where=denom > 0)
    if average == 'binary':
        return float(per[1]) if len(per) > 1 else float(per[0])
    if average == 'micro':
        return float(tp.sum() / max(cm.sum(), 1))
    return float(per.mean())


def f_score(y_pred, y_true, beta=1.0, average='binary', num_classes=None):
    p = precision(y_pred, y_true, average, num_classes)
    r = recall(y_pred, y_true, average, num_classes)
    if p + r == 0:
        return 0.0
    b2 = beta * beta
    return float((1 + b2) * p * r / (b2 * p + r))


def auc(y_pred, y_true, num_thresholds=200):
    """ROC AUC by the trapezoid rule over evenly spaced thresholds (TF1 ``tf.metrics.auc``)."""
    y_pred, y_true = _np(y_pred).reshape(-1), _np(y_true).reshape(-1) > 0.5
    eps = 1e-7
    th = np.concatenate([[0.0 - eps], (np.arange(1, num_thresholds - 1) / (num_thresholds - 1)),
                         [1.0 + eps]])
    pos = y_pred[None, :] > th[:, None]
    tp = (pos & y_true[None, :]).sum(1).astype(np.float64)
    fp = (pos & ~y_true[None, :]).sum(1).astype(np.float64)
    fn = ((~pos) & y_true[None, :]).sum(1).astype(np.float64)
    tn = ((~pos) & ~y_true[None, :]).sum(1).astype(np.float64)
    tpr = (tp + eps) / (tp + fn + eps)
    fpr = fp / (fp + tn + eps)
    return float(np.sum((fpr[:-1] - fpr[1:]) * (tpr[:-1] + tpr[1:]) / 2.0))


def roc_auc_exact(y_pred, y_true):
    """Rank-based (Mann-Whitney) AUC; equals sklearn.metrics.roc_auc_score."""
    from scipy.stats import rankdata
    y_pred, y_true = _np(y_pred).reshape(-1), _np(y_true).reshape(-1) > 0.5
    n_pos, n_neg = int(y_true.sum()), int((~y_true).sum())
    if n_pos == 0 or n_neg == 0:
        return float('nan')
    ranks = rankdata(y_pred)
    return float((ranks[y_true].sum() - n_pos * (n_pos + 1) / 2.0) / (n_pos * n_neg))
