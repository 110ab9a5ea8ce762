#!/usr/bin/env python
"""Pipeline parallelism over context blocks: GPipe, PipeDream (1F1B + weight stashing) or HetPipe
(reference examples/runner/parallel/{gpipe,pipedream,complex_pipeline_mlp}.py).

    heturun -w 4 python examples/runner/parallel/pipeline_mlp.py --pipeline gpipe --stages 4 --micro-batches 8
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..', '..'))
import hetu as ht                                              # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--pipeline', default='gpipe', choices=['gpipe', 'pipedream', 'hetpipe'])
    ap.add_argument('--stages', type=int, default=2)
    ap.add_argument('--micro-batches', type=int, default=4)
    ap.add_argument('--batch-size', type=int, default=128)
    ap.add_argument('--hidden', type=int, default=1024)
    ap.add_argument('--steps', type=int, default=10)
    args = ap.parse_args()
    comm = ht.wrapped_mpi_nccl_init()
    assert comm.nrank == args.stages, 'launch one rank per stage (heturun -w %d)' % args.stages
    dims = [784] + [args.hidden] * args.stages + [10]
    with ht.context(ht.gpu(0)):
        x = ht.Variable('x')
        act = x
    for s in range(args.stages):
        with ht.context(ht.gpu(s)):
            w = ht.init.xavier_normal((dims[s], dims[s + 1]), name='w%d' % s)
            act = ht.relu_op(ht.matmul_op(act, w))
    with ht.context(ht.gpu(args.stages - 1)):
        w = ht.init.xavier_normal((dims[-2], dims[-1]), name='w_out')
        y_ = ht.Variable('y_')
        loss = ht.reduce_mean_op(ht.softmaxcrossentropy_op(ht.matmul_op(act, w), y_), [0])
        train = ht.optim.SGDOptimizer(0.05).minimize(loss)
        ex = ht.Executor({'train': [loss, train]}, pipeline=args.pipeline, seed=0)
    rs = np.random.RandomState(0)
    M = args.micro_batches
    for step in range(args.steps):
        xs = [rs.randn(args.batch_size, 784).astype(np.float32) for _ in range(M)]
        ys = [np.eye(10, dtype=np.float32)[rs.randint(0, 10, args.batch_size)] for _ in range(M)]
        out = ex.run('train', feed_dict={x: xs, y_: ys}, batch_num=M)
        if out[0] is not None:
            print('step %d loss %.4f' % (step, float(np.mean([o.asnumpy() for o in out[0]]))), flush=True)


if __name__ == '__main__':
    main()
