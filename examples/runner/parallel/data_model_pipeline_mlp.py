#!/usr/bin/env python
"""Data + model + pipeline parallelism in one graph (reference examples/runner/parallel/data_model_pipeline_mlp.py):
stage 0 is data parallel over two ranks, stage 1 is ONE tensor-parallel worker of two ranks whose middle layer is split
with ``ht.dispatch`` (activation columns x weight rows -> partial sums reduced by the lowering,
parallel/mp_lowering.py), stage 2 is a single rank.

    heturun -w 5 python examples/runner/parallel/data_model_pipeline_mlp.py
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..', '..'))
import hetu as ht                                              # noqa: E402
from gpipe_multi import dense                                  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--batch-size', type=int, default=64)
    ap.add_argument('--hidden', type=int, default=128)
    ap.add_argument('--steps', type=int, default=5)
    args = ap.parse_args()
    comm = ht.wrapped_mpi_nccl_init()
    assert comm.nrank == 5, 'launch 5 ranks: 2 (data parallel) + 2 (one tensor-parallel worker) + 1'
    h = args.hidden
    with ht.context([ht.gpu(0), ht.gpu(1)]):                    # stage 0: two replicas
        x = ht.Variable('x', trainable=False)
        act = dense(x, (784, h), 'fc0')
    with ht.context([(ht.gpu(2), ht.gpu(3))]):                  # stage 1: one worker made of two GPUs
        w = ht.init.random_normal((h, h), stddev=0.04, name='fc1_weight')
        act = ht.matmul_op(ht.dispatch(act, (1, 2)), ht.dispatch(w, (2, 1)))       # split the contraction dimension
        act = ht.relu_op(ht.dispatch(act, (1, 1)))
    with ht.context(ht.gpu(4)):                                 # stage 2
        y_ = ht.Variable('y_', trainable=False)
        loss = ht.reduce_mean_op(ht.softmaxcrossentropy_op(dense(act, (h, 10), 'head', relu=False), y_), [0])
        train = ht.optim.SGDOptimizer(0.05).minimize(loss)
        ex = ht.Executor({'train': [loss, train]}, pipeline='gpipe', seed=0)
    rs = np.random.RandomState(0)
    M = 2                                                       # one micro-batch per stage-0 replica and tick
    for step in range(args.steps):
        xs = [rs.randn(args.batch_size, 784).astype(np.float32) for _ in range(M)]
        ys = [np.eye(10, dtype=np.float32)[x_[:, :10].argmax(1)] for x_ in xs]
        out = ex.run('train', feed_dict={x: xs, y_: ys}, batch_num=M)
        if out[0] is not None:
            print('[rank %d] step %d loss %.4f' % (comm.rank, step, float(np.mean([o.asnumpy() for o in out[0]]))), flush=True)
    comm.barrier()
    print('DATA_MODEL_PIPELINE rank %d OK' % comm.rank, flush=True)


if __name__ == '__main__':
    main()
