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


# ---------------------------------------------------------------------------------------------
# threshold / one-hot variants (reference metrics.py:17-359 spellings)
def confusion_matrix_at_thresholds(labels, predictions, thresholds, includes=None):
    """{'tp','fn','tn','fp'} -> arrays of shape [len(thresholds)] for binary ``labels`` and scores in [0, 1]."""
    keys = ('tp', 'fn', 'tn', 'fp')
    includes = keys if includes is None else tuple(includes)
    for k in includes:
        if k not in keys:
            raise ValueError('Invalid key: %s.' % k)
    y = _np(labels).reshape(-1).astype(bool)
    p = _np(predictions).reshape(-1).astype(np.float64)
    th = np.asarray(list(thresholds), dtype=np.float64).reshape(-1, 1)
    pred_pos = p[None, :] > th                                   # [T, n]
    out = {}
    if 'tp' in includes:
        out['tp'] = (pred_pos & y[None, :]).sum(1).astype(np.float64)
    if 'fn' in includes:
        out['fn'] = (~pred_pos & y[None, :]).sum(1).astype(np.float64)
    if 'tn' in includes:
        out['tn'] = (~pred_pos & ~y[None, :]).sum(1).astype(np.float64)
    if 'fp' in includes:
        out['fp'] = (pred_pos & ~y[None, :]).sum(1).astype(np.float64)
    return out


def roc_pr_curve(values, curve='ROC'):
    """(x, y) lists of the ROC (fpr, tpr) or PR (recall, precision) curve from confusion counts."""
    for k in ('tp', 'fp', 'fn', 'tn'):
        if k not in values:
            raise ValueError('values must have the key %s' % k)
    eps = 1e-6
    tp, fp, fn, tn = (np.asarray(values[k], dtype=np.float64) for k in ('tp', 'fp', 'fn', 'tn'))
    tpr = (tp + eps) / (tp + fn + eps)
    if curve == 'ROC':
        return list(fp / (fp + tn + eps)), list(tpr)
    if curve == 'PR':
        return list(tpr), list((tp + eps) / (tp + fp + eps))
    raise ValueError("curve must be 'ROC' or 'PR'")


def confusion_matrix_one_hot(labels, predictions):
    """Per-class tp / tn / fp / fn for one-hot ``labels`` and score matrix ``predictions`` (argmax decides)."""
    y = _np(labels).astype(bool)
    p = _np(predictions)
    hit = np.zeros_like(y)
    hit[np.arange(p.shape[0]), p.argmax(1)] = True
    return {'tp': (hit & y).sum(0).astype(np.float64), 'tn': (~hit & ~y).sum(0).astype(np.float64),
            'fp': (hit & ~y).sum(0).astype(np.float64), 'fn': (~hit & y).sum(0).astype(np.float64)}


def _avg_score(num, den, average):
    if average is None:
        return num / np.maximum(den, 1e-12)
    if average == 'micro':
        return float(num.sum() / max(den.sum(), 1e-12))
    if average == 'macro':
        return float(np.mean(num / np.maximum(den, 1e-12)))
    raise ValueError("average must be None, 'micro' or 'macro'")


def precision_score_one_hot(labels, predictions, average=None):
    v = confusion_matrix_one_hot(labels, predictions)
    return _avg_score(v['tp'], v['tp'] + v['fp'], average)


def recall_score_one_hot(labels, predictions, average=None):
    v = confusion_matrix_one_hot(labels, predictions)
    return _avg_score(v['tp'], v['tp'] + v['fn'], average)


def f_score_one_hot(labels, predictions, beta=1.0, average=None):
    p = precision_score_one_hot(labels, predictions, average)
    r = recall_score_one_hot(labels, predictions, average)
    b2 = beta * beta
    return (1 + b2) * p * r / np.maximum(b2 * p + r, 1e-12)
