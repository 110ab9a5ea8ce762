#!/usr/bin/env python
"""Train the vision / sequence zoo (reference examples/cnn/main.py).

    python examples/cnn/main.py --model resnet18 --dataset CIFAR10 --batch-size 128 --gpu 0 --timing
    heturun -w 8 python examples/cnn/main.py --model resnet18 --dataset CIFAR10 --comm-mode AllReduce

Without the dataset files on disk (this image has no network) synthetic data of the same shape is used.
"""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..'))
import hetu as ht                                              # noqa: E402
from hetu_61a7_b200.models import cnn as models                # noqa: E402


def load(dataset, n_train, n_val):
    shape, ncls = ((784,), 10) if dataset == 'MNIST' else ((3072,), 100 if dataset == 'CIFAR100' else 10)
    try:
        if dataset == 'MNIST':
            (tx, ty), (vx, vy), _ = ht.data.mnist()
        else:
            tx, ty, vx, vy = ht.data.normalize_cifar(num_class=ncls)
            tx, vx = tx.reshape(len(tx), -1), vx.reshape(len(vx), -1)
        return tx, ty, vx, vy, ncls
    except Exception:
        rs = np.random.RandomState(0)
        tx = rs.randn(n_train, *shape).astype(np.float32)
        vx = rs.randn(n_val, *shape).astype(np.float32)
        ty = np.eye(ncls, dtype=np.float32)[rs.randint(0, ncls, n_train)]
        vy = np.eye(ncls, dtype=np.float32)[rs.randint(0, ncls, n_val)]
        print('dataset %s not found on disk: using synthetic data of the same shape' % dataset)
        return tx, ty, vx, vy, ncls


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--model', required=True, choices=sorted(models.MODELS))
    ap.add_argument('--dataset', default='CIFAR10', choices=['MNIST', 'CIFAR10', 'CIFAR100'])
    ap.add_argument('--batch-size', type=int, default=128)
    ap.add_argument('--learning-rate', type=float, default=0.1)
    ap.add_argument('--opt', default='sgd', choices=['sgd', 'momentum', 'nesterov', 'adagrad', 'adam'])
    ap.add_argument('--num-epochs', type=int, default=2)
    ap.add_argument('--gpu', type=int, default=0, help='-1 = CPU executor')
    ap.add_argument('--validate', action='store_true')
    ap.add_argument('--timing', action='store_true')
    ap.add_argument('--comm-mode', default=None, choices=[None, 'AllReduce', 'PS', 'Hybrid'])
    ap.add_argument('--synthetic-samples', type=int, default=2048)
    args = ap.parse_args()

    tx, ty, vx, vy, ncls = load(args.dataset, args.synthetic_samples, args.synthetic_samples // 4)
    ctx = ht.cpu(0) if args.gpu < 0 else ht.gpu(args.gpu)
    opt = {'sgd': lambda: ht.optim.SGDOptimizer(args.learning_rate),
           'momentum': lambda: ht.optim.MomentumOptimizer(args.learning_rate),
           'nesterov': lambda: ht.optim.MomentumOptimizer(args.learning_rate, nesterov=True),
           'adagrad': lambda: ht.optim.AdaGradOptimizer(args.learning_rate, initial_accumulator_value=0.1),
           'adam': lambda: ht.optim.AdamOptimizer(args.learning_rate)}[args.opt]()
    x = ht.dataloader_op([ht.Dataloader(tx, args.batch_size, 'train'), ht.Dataloader(vx, args.batch_size, 'validate')])
    y_ = ht.dataloader_op([ht.Dataloader(ty, args.batch_size, 'train'), ht.Dataloader(vy, args.batch_size, 'validate')])
    kw = {'num_class': ncls} if args.model not in ('rnn', 'lstm') else {}
    loss, y = models.MODELS[args.model](x, y_, **kw)
    train_op = opt.minimize(loss)
    ex = ht.Executor({'train': [loss, y, y_, train_op], 'validate': [loss, y, y_]}, ctx=ctx, comm_mode=args.comm_mode)
    n_train, n_val = ex.get_batch_num('train'), ex.get_batch_num('validate')
    for ep in range(args.num_epochs):
        t0, losses, accs = time.time(), [], []
        for _ in range(n_train):
            l, p, lab, _ = ex.run('train', convert_to_numpy_ret_vals=True)
            losses.append(float(l))
            accs.append(float((p.argmax(1) == lab.argmax(1)).mean()))
        dt = time.time() - t0
        msg = 'epoch %d  loss %.4f  acc %.4f' % (ep, np.mean(losses), np.mean(accs))
        if args.timing:
            msg += '  %.3f s/epoch  %.1f samples/s' % (dt, n_train * args.batch_size / dt)
        if args.validate:
            vl, va = [], []
            for _ in range(n_val):
                l, p, lab = ex.run('validate', convert_to_numpy_ret_vals=True)
                vl.append(float(l))
                va.append(float((p.argmax(1) == lab.argmax(1)).mean()))
            msg += '  val loss %.4f acc %.4f' % (np.mean(vl), np.mean(va))
        print(msg, flush=True)


if __name__ == '__main__':
    main()
