#!/usr/bin/env python
"""Tensor (model) parallelism with ``ht.dispatch`` inside a tuple context
(reference examples/runner/parallel/test_mlp_mp.py): the middle layer is split over two GPUs.

    heturun -w 2 python examples/runner/parallel/test_mlp_mp.py --split middle

split: left   = activation rows split (2,1), weight replicated
       right  = weight columns split (1,2)              (column parallel)
       middle = activation (1,2) x weight (2,1)          (row parallel, partial sums all-reduced)
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..', '..'))
import hetu as ht                                              # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--split', default='middle', choices=['left', 'right', 'middle'])
    ap.add_argument('--hidden', type=int, default=1024)
    ap.add_argument('--steps', type=int, default=10)
    args = ap.parse_args()
    comm = ht.wrapped_mpi_nccl_init()
    H = args.hidden
    with ht.context(ht.gpu(0)):
        x = ht.Variable('x')
        a = ht.relu_op(ht.matmul_op(x, ht.init.xavier_normal((784, H), name='w1')))
    with ht.context((ht.gpu(0), ht.gpu(1))):
        w2 = ht.init.xavier_normal((H, H), name='w2')
        act, wt = {'left': ((2, 1), (1, 1)), 'right': ((1, 1), (1, 2)), 'middle': ((1, 2), (2, 1))}[args.split]
        a = ht.relu_op(ht.matmul_op(ht.dispatch(a, act), ht.dispatch(w2, wt)))
    with ht.context(ht.gpu(1)):
        a = ht.dispatch(a, (1, 1))
        y_ = ht.Variable('y_')
        logits = ht.matmul_op(a, ht.init.xavier_normal((H, 10), name='w3'))
        loss = ht.reduce_mean_op(ht.softmaxcrossentropy_op(logits, y_), [0])
        train = ht.optim.SGDOptimizer(0.05).minimize(loss)
        ex = ht.Executor([loss, train], seed=0)
    rs = np.random.RandomState(0)
    for step in range(args.steps):
        xv = rs.randn(128, 784).astype(np.float32)
        yv = np.eye(10, dtype=np.float32)[rs.randint(0, 10, 128)]
        out = ex.run(feed_dict={x: xv, y_: yv})
        if out[0] is not None:
            print('[rank %d] step %d loss %.4f' % (comm.rank, step, float(out[0].asnumpy())), flush=True)


if __name__ == '__main__':
    main()
