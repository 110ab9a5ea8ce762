#!/usr/bin/env python
"""Pipeline parallelism whose parameters live on the parameter server: every stage is declared on
``[ht.cpu(0), ht.gpu(i)]`` - the CPU entry makes the stage's parameters PS-resident (workers push ``-lr * grad``, the
servers add it and return the fresh value), the GPU entry is the worker that runs the stage; activations travel between
the stages over p2p (reference examples/runner/parallel/ps_pipeline.py, PipeDream schedule).

    heturun -s 1 -w 3 python examples/runner/parallel/ps_pipeline.py
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
    ap.add_argument('--micro-batches', type=int, default=4)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--learning-rate', type=float, default=0.05)
    args = ap.parse_args()
    comm = ht.wrapped_mpi_nccl_init()
    S = comm.nrank
    h = args.hidden
    with ht.context([ht.cpu(0), ht.gpu(0)]):
        x = ht.Variable('x', trainable=False)
        act = dense(x, (784, h), 'mlp_fc1')
    for i in range(1, S - 1):
        with ht.context([ht.cpu(0), ht.gpu(i)]):
            act = dense(act, (h, h), 'mlp_fc%d' % (i + 1))
    with ht.context([ht.cpu(0), ht.gpu(S - 1)]):
        y_pred = dense(act, (h, 10), 'mlp_fc_out', relu=False)
        y_ = ht.Variable('y_', trainable=False)
        loss = ht.reduce_mean_op(ht.softmaxcrossentropy_op(y_pred, y_), [0])
        train_op = ht.optim.SGDOptimizer(learning_rate=args.learning_rate).minimize(loss)
        executor = ht.Executor({'train': [loss, train_op]}, seed=0, pipeline='pipedream')
    rs = np.random.RandomState(0)
    M = args.micro_batches
    seen = []
    for step in range(args.steps):
        xs = [rs.randn(args.batch_size, 784).astype(np.float32) for _ in range(M)]
        ys = [np.eye(10, dtype=np.float32)[x_[:, :10].argmax(1)] for x_ in xs]
        out = executor.run('train', feed_dict={x: xs, y_: ys}, batch_num=M)
        if out[0] is not None:
            seen.append(float(np.mean([o.asnumpy() for o in out[0]])))
            print('[rank %d] step %d loss %.4f' % (comm.rank, step, seen[-1]), flush=True)
    comm.barrier()
    print('PS_PIPELINE rank %d OK' % comm.rank, flush=True)
    ht.worker_finish()


if __name__ == '__main__':
    main()
