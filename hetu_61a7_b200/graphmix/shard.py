"""One partition of a partitioned graph on disk (reference python/graphmix/shard.py:6-23)."""
from __future__ import annotations

import os

import numpy as np
import yaml


class Shard(object):
    def __init__(self, path):
        self.path = os.path.expanduser(path)
        with open(os.path.join(self.path, 'meta.yml'), 'rb') as f:
            self.meta = yaml.safe_load(f.read())

    def load_graph_shard(self, shard_idx):
        assert 0 <= shard_idx < self.meta['num_part']
        d = os.path.join(self.path, 'part%d' % shard_idx)
        self.index = shard_idx
        self.edges = np.load(os.path.join(d, 'graph.npy'))
        self.f_feat = np.load(os.path.join(d, 'float_feature.npy'))
        self.i_feat = np.load(os.path.join(d, 'int_feature.npy'))
        return self

    def csr(self):
        """(indptr, neighbours) of the nodes this shard owns; neighbours are global ids."""
        off = self.meta['partition']['offset']
        lo, n_local = off[self.index], off[self.index + 1] - off[self.index]
        src = self.edges[0] - lo
        order = np.argsort(src, kind='stable')
        indptr = np.zeros(n_local + 1, dtype=np.int64)
        np.add.at(indptr, src + 1, 1)
        return np.cumsum(indptr).astype(np.int64), np.ascontiguousarray(self.edges[1][order], dtype=np.int64)
