#!/usr/bin/env python
"""A 4-stage MLP pipeline in which every stage is replicated (data parallel) - the layout of the reference's
examples/runner/parallel/dist_data_pipeline_mlp.py (stages on ``ht.rgpu(host, i)`` pairs of two machines).  Here the
device groups are built from ``--hosts``; with one host the remote contexts collapse to local GPUs.

    heturun -w 8 python examples/runner/parallel/dist_data_pipeline_mlp.py                   # one node
    heturun -c cluster.yml python examples/runner/parallel/dist_data_pipeline_mlp.py --hosts nodeA nodeB
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
    ap.add_argument('--hosts', nargs='*', default=[])
    ap.add_argument('--replicas', type=int, default=2)
    ap.add_argument('--batch-size', type=int, default=128)
    ap.add_argument('--hidden', type=int, default=256)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--learning-rate', type=float, default=0.01)
    args = ap.parse_args()
    comm = ht.wrapped_mpi_nccl_init()
    R, S = args.replicas, 4
    assert comm.nrank == S * R, 'launch %d ranks' % (S * R)
    per_host = S * R // max(len(args.hosts), 1)

    def device(rank):
        if not args.hosts:
            return ht.gpu(rank)
        return ht.rgpu(args.hosts[rank // per_host], rank % per_host)

    group = lambda s: [device(s * R + r) for r in range(R)]    # noqa: E731
    h = args.hidden
    with ht.context(group(0)):
        x = ht.Variable(name='dataloader_x', trainable=False)
        act = dense(x, (784, h), 'mlp_fc0')
    with ht.context(group(1)):
        act = dense(dense(act, (h, h), 'mlp_fc1'), (h, h), 'mlp_fc11')
    with ht.context(group(2)):
        act = dense(dense(act, (h, h), 'mlp_fc2'), (h, h), 'mlp_fc22')
    with ht.context(group(3)):
        y_pred = dense(act, (h, 10), 'mlp_fc3', relu=False)
        y_ = ht.Variable(name='dataloader_y', trainable=False)
        loss = ht.reduce_mean_op(ht.softmaxcrossentropy_op(y_pred, y_), [0])
        train_op = ht.optim.SGDOptimizer(learning_rate=args.learning_rate).minimize(loss)
        executor = ht.Executor([loss, train_op], seed=0)
    rs = np.random.RandomState(comm.rank % R)
    for step in range(args.steps):
        xv = rs.randn(args.batch_size, 784).astype(np.float32)
        yv = np.eye(10, dtype=np.float32)[xv[:, :10].argmax(1)]
        out = executor.run(feed_dict={x: xv, y_: yv}, convert_to_numpy_ret_vals=True)
        if out[0] is not None:
            print('[rank %d] step %d loss %.4f' % (comm.rank, step, float(np.asarray(out[0]).reshape(-1)[0])), flush=True)
    comm.barrier()
    print('DIST_DATA_PIPELINE rank %d OK' % comm.rank, flush=True)


if __name__ == '__main__':
    main()
