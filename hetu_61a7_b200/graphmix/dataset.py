"""Graph datasets for the sampling library.

reference: third_party/GraphMix/python/graphmix/dataset/{dataset,utils}.py (Cora / PubMed / Reddit /
ogbn-* downloaders returning x, y, edge_index, train mask).  There is no network here, so the
named datasets are generated: a planted-partition graph with the real dataset's node / feature /
class counts whose features carry the class signal.  ``load_dataset(path)`` reads an ``.npz`` with
``x, y, edge_index[, train_mask]`` for real data."""
from __future__ import annotations

import os

import numpy as np

_SHAPES = {  # name: (nodes, feature dim, classes, average degree)
    'Cora': (2708, 1433, 7, 4), 'PubMed': (19717, 500, 3, 5), 'Citeseer': (3327, 3703, 6, 3),
    'Flickr': (89250, 500, 7, 10), 'Reddit': (232965, 602, 41, 50), 'ogbn-arxiv': (169343, 128, 40, 14),
    'Yelp': (716847, 300, 100, 20), 'ogbn-products': (2449029, 100, 47, 50),
}


class _Graph(object):
    def __init__(self, edge_index, num_nodes):
        self.edge_index = edge_index
        self.num_nodes = num_nodes
        self.num_edges = int(edge_index.shape[1])


class GraphDataset(object):
    """x [N, F] float32, y [N] int, graph.edge_index [2, E] (both directions stored), train_mask [N]."""

    def __init__(self, name, x, y, edge_index, train_mask=None, num_classes=None):
        self.name = name
        self.x = np.ascontiguousarray(x, dtype=np.float32)
        self.y = np.ascontiguousarray(y, dtype=np.int32)
        self.graph = _Graph(np.ascontiguousarray(edge_index, dtype=np.int64), self.x.shape[0])
        self.train_mask = np.ascontiguousarray(train_mask if train_mask is not None else np.ones(self.x.shape[0]), dtype=np.int32)
        self.num_classes = int(num_classes if num_classes is not None else self.y.max() + 1)


def synthetic_dataset(name='synthetic', num_nodes=2000, feat_dim=32, num_classes=4, avg_degree=6, p_in=0.85, seed=0,
                      scale=1.0) -> GraphDataset:
    rng = np.random.RandomState(seed)
    n = max(int(num_nodes * scale), num_classes * 4)
    y = rng.randint(0, num_classes, n)
    m = n * avg_degree // 2
    src = rng.randint(0, n, m)
    same = rng.rand(m) < p_in
    order = np.argsort(y, kind='stable')
    starts = np.searchsorted(y[order], np.arange(num_classes))
    counts = np.bincount(y, minlength=num_classes)
    pick = starts[y[src]] + (rng.rand(m) * counts[y[src]]).astype(np.int64)
    dst = np.where(same, order[np.minimum(pick, n - 1)], rng.randint(0, n, m))
    keep = src != dst
    src, dst = src[keep], dst[keep]
    e = np.unique(np.stack([np.concatenate([src, dst]), np.concatenate([dst, src])], axis=1), axis=0).T
    centers = rng.standard_normal((num_classes, feat_dim)).astype(np.float32)
    x = centers[y] + 1.5 * rng.standard_normal((n, feat_dim)).astype(np.float32)
    train_mask = (rng.rand(n) < 0.6).astype(np.int32)
    return GraphDataset(name, x, y, e, train_mask, num_classes)


def load_dataset(name: str, scale: float = 1.0) -> GraphDataset:
    if os.path.exists(name):
        z = np.load(name)
        return GraphDataset(os.path.basename(name), z['x'], z['y'], z['edge_index'],
                            z['train_mask'] if 'train_mask' in z.files else None)
    if name not in _SHAPES:
        raise ValueError('unknown dataset %s (known: %s, or a path to an .npz)' % (name, ', '.join(_SHAPES)))
    n, f, c, deg = _SHAPES[name]
    return synthetic_dataset(name, n, f, c, deg, scale=scale)
