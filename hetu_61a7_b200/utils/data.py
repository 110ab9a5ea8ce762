"""Dataset helpers.  reference: python/hetu/data.py:5-338 (MNIST / CIFAR loaders and
augmentation).  There is no network on the build boxes, so each loader falls back
to a deterministic synthetic dataset of the same shape/dtype when the files are
not on disk (``synthetic=True`` forces it)."""
from __future__ import annotations

import gzip
import os
import pickle

import numpy as np


def one_hot(labels, num_classes):
    out = np.zeros((len(labels), num_classes), dtype=np.float32)
    out[np.arange(len(labels)), np.asarray(labels, dtype=np.int64)] = 1.0
    return out


def _synthetic(n, shape, num_classes, seed):
    rng = np.random.RandomState(seed)
    y = rng.randint(0, num_classes, size=n)
    centers = rng.normal(0, 1, size=(num_classes,) + tuple(shape)).astype(np.float32)
    x = centers[y] * 0.5 + rng.normal(0, 1, size=(n,) + tuple(shape)).astype(np.float32)
    return x.astype(np.float32), y


def mnist(dataset='mnist.pkl.gz', onehot=True, synthetic=False, n_train=50000, n_valid=10000):
    if not synthetic and os.path.exists(dataset):
        with gzip.open(dataset, 'rb') as f:
            train, valid, test = pickle.load(f, encoding='latin1')
        sets = [train, valid, test]
    else:
        sets = [_synthetic(n_train, (784,), 10, 1), _synthetic(n_valid, (784,), 10, 2),
                _synthetic(n_valid, (784,), 10, 3)]
    out = []
    for x, y in sets:
        out.append((np.asarray(x, dtype=np.float32), one_hot(y, 10) if onehot else np.asarray(y)))
    return out


def _cifar_dir(directory, names, num_classes, label_key, onehot):
    xs, ys = [], []
    for n in names:
        with open(os.path.join(directory, n), 'rb') as f:
            d = pickle.load(f, encoding='latin1')
        xs.append(np.asarray(d['data'], dtype=np.float32).reshape(-1, 3, 32, 32))
        ys.append(np.asarray(d[label_key]))
    x, y = np.concatenate(xs), np.concatenate(ys)
    return x, (one_hot(y, num_classes) if onehot else y)


def normalize_cifar(num_class=10, directory=None, onehot=True, synthetic=False, n_train=50000,
                    n_test=10000):
    """Returns (train_x, train_y, test_x, test_y), images NCHW float32 normalised per channel."""
    directory = directory or ('datasets/cifar-10-batches-py' if num_class == 10 else 'datasets/cifar-100-python')
    if not synthetic and os.path.isdir(directory):
        if num_class == 10:
            tx, ty = _cifar_dir(directory, ['data_batch_%d' % i for i in range(1, 6)], 10, 'labels', onehot)
            vx, vy = _cifar_dir(directory, ['test_batch'], 10, 'labels', onehot)
        else:
            tx, ty = _cifar_dir(directory, ['train'], 100, 'fine_labels', onehot)
            vx, vy = _cifar_dir(directory, ['test'], 100, 'fine_labels', onehot)
    else:
        tx, ty = _synthetic(n_train, (3, 32, 32), num_class, 11)
        vx, vy = _synthetic(n_test, (3, 32, 32), num_class, 12)
        if onehot:
            ty, vy = one_hot(ty, num_class), one_hot(vy, num_class)
    mean = tx.mean(axis=(0, 2, 3), keepdims=True)
    std = tx.std(axis=(0, 2, 3), keepdims=True) + 1e-7
    return (tx - mean) / std, ty, (vx - mean) / std, vy


def cifar10(directory=None, onehot=True, synthetic=False):
    return normalize_cifar(10, directory, onehot, synthetic)


def cifar100(directory=None, onehot=True, synthetic=False):
    return normalize_cifar(100, directory, onehot, synthetic)


def random_crop(x, padding=4, rng=None):
    rng = rng or np.random
    n, c, h, w = x.shape
    padded = np.pad(x, ((0, 0), (0, 0), (padding, padding), (padding, padding)), mode='constant')
    out = np.empty_like(x)
    ox = rng.randint(0, 2 * padding + 1, size=n)
    oy = rng.randint(0, 2 * padding + 1, size=n)
    for i in range(n):
        out[i] = padded[i, :, oy[i]:oy[i] + h, ox[i]:ox[i] + w]
    return out


def random_flip(x, rng=None):
    rng = rng or np.random
    flip = rng.rand(x.shape[0]) < 0.5
    out = x.copy()
    out[flip] = out[flip][..., ::-1]
    return out


def augment(x, rng=None):
    return random_flip(random_crop(x, 4, rng), rng)


def criteo_synthetic(n, num_dense=13, num_sparse=26, vocab=33762577, seed=0, zipf_a=1.05):
    """Criteo-shaped synthetic batch: dense [n,13] float, sparse ids [n,26] (zipf within
    per-field vocab slices that tile ``vocab``), labels [n,1] in {0,1}."""
    rng = np.random.RandomState(seed)
    dense = rng.rand(n, num_dense).astype(np.float32)
    field = vocab // num_sparse
    ranks = rng.zipf(zipf_a, size=(n, num_sparse)).astype(np.int64)
    ids = (ranks - 1) % field + np.arange(num_sparse, dtype=np.int64)[None, :] * field
    labels = (rng.rand(n, 1) < 0.25).astype(np.float32)
    return dense, ids, labels


# ---------------------------------------------------------------------------------------------
def convert_to_one_hot(vals, max_val=0):
    """Label vector -> one-hot matrix (reference data.py:250-256)."""
    vals = np.asarray(vals).astype(np.int64).reshape(-1)
    if max_val == 0:
        max_val = int(vals.max()) + 1
    out = np.zeros((vals.size, max_val), dtype=np.float32)
    out[np.arange(vals.size), vals] = 1
    return out


def tf_normalize_cifar(num_class=10, onehot=True, directory=None, synthetic=False):
    """CIFAR in NHWC layout with per-channel mean/std normalisation (reference data.py:221-247)."""
    tx, ty, vx, vy = normalize_cifar(num_class=num_class, directory=directory, onehot=onehot, synthetic=synthetic)
    tx = tx.reshape((-1, 3, 32, 32)).transpose(0, 2, 3, 1)
    vx = vx.reshape((-1, 3, 32, 32)).transpose(0, 2, 3, 1)
    return np.ascontiguousarray(tx), ty, np.ascontiguousarray(vx), vy


def data_augmentation(images, mode='train', flip=False, crop=False, crop_shape=(24, 24, 3), whiten=False,
                      noise=False, noise_mean=0, noise_std=0.01, rng=None):
    """NHWC batch augmentation: random (train) / centre (test) crop, horizontal flip, per-image
    whitening, additive Gaussian noise (reference data.py:263-338)."""
    rng = rng if rng is not None else np.random
    x = np.asarray(images, dtype=np.float32)
    if crop:
        n, h, w, _ = x.shape
        ch, cw = crop_shape[0], crop_shape[1]
        out = np.empty((n, ch, cw, x.shape[3]), dtype=np.float32)
        for i in range(n):
            top, left = ((rng.randint(0, h - ch + 1), rng.randint(0, w - cw + 1)) if mode == 'train'
                         else ((h - ch) // 2, (w - cw) // 2))
            out[i] = x[i, top:top + ch, left:left + cw]
        x = out
    if flip:
        mask = rng.rand(x.shape[0]) < 0.5
        x[mask] = x[mask, :, ::-1]
    if whiten:
        mean = x.mean(axis=(1, 2, 3), keepdims=True)
        std = np.maximum(x.std(axis=(1, 2, 3), keepdims=True), 1.0 / np.sqrt(np.prod(x.shape[1:])))
        x = (x - mean) / std
    if noise:
        x = x + rng.normal(noise_mean, noise_std, size=x.shape).astype(np.float32)
    return x
