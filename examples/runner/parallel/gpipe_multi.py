#!/usr/bin/env python
"""GPipe over stages that are themselves data-parallel: S stages x R replicas, micro-batches flow through the stages,
each stage's replicas all-reduce their gradients at the end of the batch (reference
examples/runner/parallel/gpipe_multi.py: two pipelines side by side on 8 GPUs).

    heturun -w 4 python examples/runner/parallel/gpipe_multi.py --stages 2 --replicas 2
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..', '..'))
import hetu as ht                                              # noqa: E402


def dense(x, shape, name, relu=True):
    w = ht.init.random_normal(shape, stddev=0.04, name=name + '_weight')
    b = ht.init.random_normal(shape[-1:], stddev=0.04, name=name + '_bias')
    x = ht.matmul_op(x, w)
    x = x + ht.broadcastto_op(b, x)
    return ht.relu_op(x) if relu else x


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--stages', type=int, default=2)
    ap.add_argument('--replicas', type=int, default=2)
    ap.add_argument('--micro-batches', type=int, default=4)
    ap.add_argument('--batch-size', type=int, default=64)
    ap.add_argument('--hidden', type=int, default=256)
    ap.add_argument('--steps', type=int, default=5)
    args = ap.parse_args()
    comm = ht.wrapped_mpi_nccl_init()
    S, R = args.stages, args.replicas
    assert comm.nrank == S * R, 'launch %d ranks (heturun -w %d)' % (S * R, S * R)
    group = lambda s: [ht.gpu(s * R + r) for r in range(R)]    # noqa: E731  stage s lives on R consecutive ranks
    dims = [784] + [args.hidden] * S
    with ht.context(group(0)):
        x = ht.Variable('x', trainable=False)
        act = x
    for s in range(S):
        with ht.context(group(s)):
            act = dense(act, (dims[s], dims[s + 1]), 'fc%d' % s)
    with ht.context(group(S - 1)):
        y_ = ht.Variable('y_', trainable=False)
        loss = ht.reduce_mean_op(ht.softmaxcrossentropy_op(dense(act, (dims[-1], 10), 'head', relu=False), y_), [0])
        train = ht.optim.SGDOptimizer(0.05).minimize(loss)
        ex = ht.Executor({'train': [loss, train]}, pipeline='gpipe', seed=0)
    replica = comm.rank % R
    rs = np.random.RandomState(replica)                         # every replica column trains on its own shard
    M = args.micro_batches
    first = None
    for step in range(args.steps):
        xs = [rs.randn(args.batch_size, 784).astype(np.float32) for _ in range(M)]
        ys = [np.eye(10, dtype=np.float32)[(x_[:, :10].argmax(1))] for x_ in xs]      # learnable labels
        out = ex.run('train', feed_dict={x: xs, y_: ys}, batch_num=M)
        if out[0] is not None:
            l = float(np.mean([o.asnumpy() for o in out[0]]))
            first = l if first is None else first
            print('[rank %d] step %d loss %.4f' % (comm.rank, step, l), flush=True)
    comm.barrier()
    print('GPIPE_MULTI rank %d OK' % comm.rank, flush=True)


if __name__ == '__main__':
    main()
