#!/usr/bin/env python
"""Wide&Deep on Adult declared with context lists (reference examples/runner/run_wdl.py): the
embedding tables can sit on a parameter server while the dense part is all-reduced.

    python run_wdl.py --config local
    heturun -c local_ps.yml        python run_wdl.py --config lps     # everything through the PS
    heturun -c local_ps.yml        python run_wdl.py --config lhy     # Hybrid: PS embeddings + all-reduced dense part
    heturun -c local_allreduce.yml python run_wdl.py --config lar

Adult is generated with the real shapes (8 categorical + 4 continuous deep inputs, 809 wide
columns) because the image has no network."""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..'))
import hetu as ht                                              # noqa: E402
from hetu_61a7_b200.models import ctr                          # noqa: E402


from hetu_61a7_b200.utils.ctr_data import adult_synthetic as synthetic_adult   # noqa: E402


if __name__ == '__main__':
    parser = argparse.ArgumentParser()
    parser.add_argument('--config', default='local', help='[local, lps, lhy(hybrid), lar, rps, rhy, rar]')
    parser.add_argument('--batch-size', type=int, default=128)
    parser.add_argument('--num-epochs', type=int, default=3)
    parser.add_argument('--samples', type=int, default=32561)
    parser.add_argument('--cpu', action='store_true')
    parser.add_argument('--validate', action='store_true')
    args = parser.parse_args()
    workers = {'local': ['gpu:0'], 'lps': ['gpu:0', 'gpu:1'], 'lhy': ['gpu:0', 'gpu:1'], 'lar': ['gpu:0', 'gpu:1', 'gpu:2', 'gpu:3'],
               'rps': ['node1:gpu:0', 'node1:gpu:1', 'node2:gpu:0', 'node2:gpu:1'],
               'rhy': ['node1:gpu:0', 'node1:gpu:1', 'node2:gpu:0', 'node2:gpu:1'],
               'rar': ['node1:gpu:0', 'node1:gpu:1', 'node2:gpu:0', 'node2:gpu:1']}[args.config]
    with_ps = args.config in ('lps', 'rps')
    hybrid = args.config in ('lhy', 'rhy')
    ctx = ht.cpu(0) if args.cpu else ((['cpu:0'] if (with_ps or hybrid) else []) + workers)
    deep, wide, y = synthetic_adult(args.samples)
    ntrain = int(args.samples * 0.8) // args.batch_size * args.batch_size
    with ht.context(ctx):
        def loader(a):
            return ht.dataloader_op([ht.Dataloader(a[:ntrain], args.batch_size, 'train'),
                                     ht.Dataloader(a[ntrain:], args.batch_size, 'validate')])
        X_deep = [loader(a) for a in deep]
        X_wide, y_ = loader(wide), loader(y)
        loss, pred, label, train_op = ctr.wdl_adult(X_deep, X_wide, y_)
        executor = ht.Executor({'train': [loss, pred, label, train_op], 'validate': [loss, pred, label]},
                               comm_mode='Hybrid' if hybrid else None)
    nb, nv = executor.get_batch_num('train'), executor.get_batch_num('validate')
    for ep in range(args.num_epochs):
        t0 = time.time()
        tot, acc = 0.0, []
        for _ in range(nb):
            l, p, lab, _ = executor.run('train', convert_to_numpy_ret_vals=True)
            tot += float(l)
            acc.extend(np.argmax(p, 1) == np.argmax(lab, 1))
        msg = 'epoch %d loss %.4f train acc %.4f time %.2fs' % (ep, tot / nb, np.mean(acc), time.time() - t0)
        if args.validate:
            vacc = []
            for _ in range(nv):
                l, p, lab = executor.run('validate', convert_to_numpy_ret_vals=True)
                vacc.extend(np.argmax(p, 1) == np.argmax(lab, 1))
            msg += ' valid acc %.4f' % np.mean(vacc)
        print(msg, flush=True)
