"""Graph partitioning: writes ``<path>/meta.yml`` and ``part<i>/{graph,float_feature,int_feature}.npy``.

reference: third_party/GraphMix/python/graphmix/partition.py:1-94 (METIS k-way partition, nodes
renumbered so that every part owns a contiguous id range, per-part edge / feature files, meta.yml).
METIS is not available here; the partitioner below grows ``nparts`` balanced regions breadth-first
from spread-out seeds and then runs a few label-propagation sweeps that move nodes to the part
holding most of their neighbours under a balance cap - the same objective (few cut edges, equal
sizes).

    python -m hetu_61a7_b200.graphmix.partition -d Cora -n 4 -p ~/mydata
"""
from __future__ import annotations

import argparse
import os
from collections import deque

import numpy as np
import yaml

from .dataset import GraphDataset, load_dataset


def _csr(n, edge_index):
    order = np.argsort(edge_index[0], kind='stable')
    src, dst = edge_index[0][order], edge_index[1][order]
    indptr = np.zeros(n + 1, dtype=np.int64)
    np.add.at(indptr, src + 1, 1)
    return np.cumsum(indptr), dst


def partition_nodes(n: int, edge_index: np.ndarray, nparts: int, seed: int = 0, sweeps: int = 4) -> np.ndarray:
    """part id per node."""
    if nparts <= 1:
        return np.zeros(n, dtype=np.int64)
    rng = np.random.RandomState(seed)
    indptr, nbr = _csr(n, edge_index)
    part = np.full(n, -1, dtype=np.int64)
    cap = -(-n // nparts)
    sizes = np.zeros(nparts, dtype=np.int64)
    queues = [deque() for _ in range(nparts)]
    for p, s in enumerate(rng.choice(n, nparts, replace=False)):
        part[s] = p
        sizes[p] += 1
        queues[p].append(int(s))
    unassigned = n - nparts
    while unassigned > 0:
        progressed = False
        for p in np.argsort(sizes):
            q = queues[p]
            while q and sizes[p] < cap:
                u = q.popleft()
                grew = False
                for v in nbr[indptr[u]:indptr[u + 1]]:
                    if part[v] < 0 and sizes[p] < cap:
                        part[v] = p
                        sizes[p] += 1
                        unassigned -= 1
                        q.append(int(v))
                        grew = progressed = True
                if grew:
                    break
        if not progressed:                                   # disconnected remainder: hand out to the smallest parts
            rest = np.nonzero(part < 0)[0]
            for v in rest:
                p = int(np.argmin(sizes))
                part[v] = p
                sizes[p] += 1
                queues[p].append(int(v))
            unassigned = 0
    limit = int(cap * 1.05) + 1
    for _ in range(sweeps):
        moved = 0
        for u in rng.permutation(n):
            nb = nbr[indptr[u]:indptr[u + 1]]
            if nb.size == 0:
                continue
            votes = np.bincount(part[nb], minlength=nparts)
            best = int(np.argmax(votes))
            cur = int(part[u])
            if best != cur and votes[best] > votes[cur] and sizes[best] < limit and sizes[cur] > 1:
                part[u] = best
                sizes[best] += 1
                sizes[cur] -= 1
                moved += 1
        if moved == 0:
            break
    return part


def part_graph(dataset, nparts: int, path: str, seed: int = 0) -> dict:
    """Partition ``dataset`` (a name, an .npz path or a GraphDataset) into ``nparts`` shards under
    ``path/<name>`` and return the meta dict."""
    ds = dataset if isinstance(dataset, GraphDataset) else load_dataset(dataset)
    n = ds.x.shape[0]
    edge_index = ds.graph.edge_index
    part = partition_nodes(n, edge_index, nparts, seed)
    order = np.argsort(part, kind='stable')                  # new id -> old id ; parts become contiguous ranges
    new_id = np.empty(n, dtype=np.int64)
    new_id[order] = np.arange(n)
    counts = np.bincount(part, minlength=nparts)
    offset = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    src, dst = new_id[edge_index[0]], new_id[edge_index[1]]
    root = os.path.join(os.path.expanduser(path), ds.name)
    os.makedirs(root, exist_ok=True)
    owner = part[edge_index[0]]
    edges_per_part = []
    for p in range(nparts):
        d = os.path.join(root, 'part%d' % p)
        os.makedirs(d, exist_ok=True)
        sel = owner == p
        np.save(os.path.join(d, 'graph.npy'), np.stack([src[sel], dst[sel]]).astype(np.int64))
        old = order[offset[p]:offset[p + 1]]
        np.save(os.path.join(d, 'float_feature.npy'), ds.x[old])
        # int features: [label, train_mask, original node id]
        np.save(os.path.join(d, 'int_feature.npy'), np.stack([ds.y[old], ds.train_mask[old], old.astype(np.int32)], axis=1).astype(np.int32))
        edges_per_part.append(int(sel.sum()))
    cut = int((part[edge_index[0]] != part[edge_index[1]]).sum())
    meta = {'name': ds.name, 'num_part': int(nparts), 'node': int(n), 'edge': int(edge_index.shape[1]),
            'float_feature': int(ds.x.shape[1]), 'int_feature': 3, 'class': int(ds.num_classes),
            'train_node': int(ds.train_mask.sum()), 'cut_edge': cut,
            'partition': {'nodes': [int(c) for c in counts], 'edges': edges_per_part, 'offset': [int(o) for o in offset]}}
    with open(os.path.join(root, 'meta.yml'), 'w') as f:
        yaml.safe_dump(meta, f)
    return meta


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--dataset', '-d', required=True)
    ap.add_argument('--nparts', '-n', type=int, required=True)
    ap.add_argument('--path', '-p', required=True)
    ap.add_argument('--scale', type=float, default=1.0, help='shrink a generated dataset')
    args = ap.parse_args()
    meta = part_graph(load_dataset(args.dataset, args.scale), args.nparts, args.path)
    print(yaml.safe_dump(meta))


if __name__ == '__main__':
    main()
