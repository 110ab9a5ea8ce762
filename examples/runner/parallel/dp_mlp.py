#!/usr/bin/env python
"""Data parallelism with a context list (reference examples/runner/parallel/data_mlp.py):
``with ht.context([ht.gpu(0), ..., ht.gpu(n-1)])`` -> n replicas, gradients all-reduced and the
optimizer applied inside ONE fused NVLink kernel (parallel/symm.py FusedDPUpdate).

    heturun -w 8 python examples/runner/parallel/dp_mlp.py
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..', '..'))
import hetu as ht                                              # noqa: E402

comm = ht.wrapped_mpi_nccl_init()
with ht.context([ht.gpu(i) for i in range(comm.nrank)]):
    x, y_ = ht.Variable('x'), ht.Variable('y_')
    h = ht.relu_op(ht.matmul_op(x, ht.init.xavier_normal((784, 1024), name='w1')))
    h = ht.relu_op(ht.matmul_op(h, ht.init.xavier_normal((1024, 1024), name='w2')))
    loss = ht.reduce_mean_op(ht.softmaxcrossentropy_op(ht.matmul_op(h, ht.init.xavier_normal((1024, 10), name='w3')), y_), [0])
    train = ht.optim.AdamOptimizer(1e-3).minimize(loss)
    ex = ht.Executor([loss, train], seed=0)
rs = np.random.RandomState(comm.rank)
for step in range(20):
    xv = rs.randn(128, 784).astype(np.float32)
    yv = np.eye(10, dtype=np.float32)[rs.randint(0, 10, 128)]
    l = float(ex.run(feed_dict={x: xv, y_: yv})[0].asnumpy())
    if comm.rank == 0:
        print('step %d loss %.4f' % (step, l), flush=True)
