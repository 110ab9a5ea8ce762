#!/usr/bin/env python
"""GCN / GraphSAGE node classification on one device (reference examples/gnn/run_single.py).
A random stochastic-block-model graph stands in for the OGB datasets (no network in this image)."""
import argparse
import os
import sys

import numpy as np
import scipy.sparse as sp

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..'))
import hetu as ht                                              # noqa: E402
from hetu_61a7_b200.models import gnn                          # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--model', default='gcn', choices=['gcn', 'sage'])
    ap.add_argument('--nodes', type=int, default=2000)
    ap.add_argument('--classes', type=int, default=5)
    ap.add_argument('--features', type=int, default=64)
    ap.add_argument('--hidden', type=int, default=64)
    ap.add_argument('--epochs', type=int, default=50)
    ap.add_argument('--gpu', type=int, default=0)
    args = ap.parse_args()
    rs = np.random.RandomState(0)
    lab = rs.randint(0, args.classes, args.nodes)
    same = lab[:, None] == lab[None, :]
    adj = sp.coo_matrix(((rs.rand(args.nodes, args.nodes) < np.where(same, 0.02, 0.002)) | np.eye(args.nodes, dtype=bool))
                        .astype(np.float32))
    adj = ((adj + adj.T) > 0).astype(np.float32)
    deg = np.asarray(adj.sum(1)).ravel()
    norm = (sp.diags(deg ** -0.5) @ adj @ sp.diags(deg ** -0.5)).tocoo()
    feat_v = (np.eye(args.classes)[lab] @ rs.randn(args.classes, args.features) + rs.randn(args.nodes, args.features)).astype(np.float32)
    ctx = ht.cpu(0) if args.gpu < 0 else ht.gpu(args.gpu)
    (feat, y_, mp, mask), (loss, y, train) = gnn.dense_model(args.features, args.hidden, args.classes, lr=0.5,
                                                             layer=gnn.GCN if args.model == 'gcn' else gnn.SageConv)
    ex = ht.Executor([loss, y, train], ctx=ctx)
    mp_v = ht.sparse_array(norm.data.astype(np.float32), (norm.row, norm.col), shape=norm.shape, ctx=ctx)
    train_mask = (rs.rand(args.nodes) < 0.5).astype(np.float32)
    onehot = np.eye(args.classes, dtype=np.float32)[lab]
    for ep in range(args.epochs):
        l, p, _ = ex.run(feed_dict={feat: feat_v, y_: onehot, mp: mp_v, mask: train_mask}, convert_to_numpy_ret_vals=True)
        if ep % 10 == 0 or ep == args.epochs - 1:
            acc = ((p.argmax(1) == lab) * (1 - train_mask)).sum() / (1 - train_mask).sum()
            print('epoch %d loss %.4f test acc %.4f' % (ep, float(l), acc), flush=True)


if __name__ == '__main__':
    main()
