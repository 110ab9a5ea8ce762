#!/usr/bin/env python
"""Neural collaborative filtering (reference examples/rec/run_hetu.py, MovieLens-shaped synthetic data)."""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..'))
import hetu as ht                                              # noqa: E402
from hetu_61a7_b200.models.rec import neural_mf                # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--num-users', type=int, default=138493)
    ap.add_argument('--num-items', type=int, default=26744)
    ap.add_argument('--batch-size', type=int, default=1024)
    ap.add_argument('--steps', type=int, default=100)
    ap.add_argument('--comm', default=None, choices=[None, 'PS', 'Hybrid', 'AllReduce'])
    ap.add_argument('--gpu', type=int, default=0)
    args = ap.parse_args()
    ctx = ht.cpu(0) if args.gpu < 0 else ht.gpu(args.gpu)
    u, i, y_ = ht.Variable('user'), ht.Variable('item'), ht.Variable('y_')
    loss, y, train = neural_mf(u, i, y_, args.num_users, args.num_items)
    ex = ht.Executor([loss, y, train], ctx=ctx, comm_mode=args.comm)
    rs = np.random.RandomState(0)
    t0 = time.time()
    for step in range(args.steps):
        uu = rs.randint(0, args.num_users, args.batch_size).astype(np.float32)
        ii = rs.randint(0, args.num_items, args.batch_size).astype(np.float32)
        yy = ((uu.astype(np.int64) + ii.astype(np.int64)) % 5 == 0).astype(np.float32).reshape(-1, 1)
        l = ex.run(feed_dict={u: uu, i: ii, y_: yy})[0].asnumpy()
        if step % 20 == 0:
            print('step %d loss %.4f' % (step, float(np.mean(l))), flush=True)
    print('%.0f samples/s' % (args.steps * args.batch_size / (time.time() - t0)))


if __name__ == '__main__':
    main()
