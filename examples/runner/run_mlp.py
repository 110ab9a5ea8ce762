#!/usr/bin/env python
"""MLP on MNIST declared with *context lists* (reference examples/runner/run_mlp.py:65-72): GPU
entries of the list are data-parallel workers, CPU entries are parameter servers, so the same
script is single-device, PS or AllReduce depending on ``--config``:

    python run_mlp.py --config local
    heturun -c local_ps.yml        python run_mlp.py --config lps
    heturun -c local_allreduce.yml python run_mlp.py --config lar
    heturun -c remote_ps.yml       python run_mlp.py --config rps     # hosts of remote_ps.yml

MNIST is generated with the real shapes when mnist.pkl.gz is absent (no network here)."""
import argparse
import os
import sys
from time import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..'))
import hetu as ht                                              # noqa: E402
from hetu_61a7_b200.models import cnn as zoo                   # noqa: E402


def make_optimizer(name, lr):
    assert name in ['sgd', 'momentum', 'nesterov', 'adagrad', 'adam'], 'Optimizer not supported!'
    if name == 'sgd':
        return ht.optim.SGDOptimizer(learning_rate=lr)
    if name == 'momentum':
        return ht.optim.MomentumOptimizer(learning_rate=lr)
    if name == 'nesterov':
        return ht.optim.MomentumOptimizer(learning_rate=lr, nesterov=True)
    if name == 'adagrad':
        return ht.optim.AdaGradOptimizer(learning_rate=lr, initial_accumulator_value=0.1)
    return ht.optim.AdamOptimizer(learning_rate=lr)


def contexts(config, cpu=False):
    if cpu:
        return ht.cpu(0)
    return {
        'local': ht.gpu(0),
        'lps': [ht.cpu(0), ht.gpu(0), ht.gpu(1)],
        'lar': [ht.gpu(0), ht.gpu(1), ht.gpu(2), ht.gpu(3)],
        'rps': ['cpu:0', 'node1:gpu:0', 'node1:gpu:1', 'node1:gpu:2', 'node1:gpu:3', 'node2:gpu:0', 'node2:gpu:1',
                'node2:gpu:2', 'node2:gpu:3'],
        'rar': ['node1:gpu:%d' % i for i in range(8)] + ['node2:gpu:%d' % i for i in range(8)],
    }[config]


if __name__ == '__main__':
    parser = argparse.ArgumentParser()
    parser.add_argument('--config', type=str, default='local', help='[local, lps(localps), lar(localallreduce), rps(remoteps), rar]')
    parser.add_argument('--batch-size', type=int, default=128)
    parser.add_argument('--learning-rate', type=float, default=0.1)
    parser.add_argument('--opt', type=str, default='sgd', help='sgd / momentum / nesterov / adagrad / adam')
    parser.add_argument('--num-epochs', type=int, default=10)
    parser.add_argument('--validate', action='store_true')
    parser.add_argument('--timing', action='store_true')
    parser.add_argument('--cpu', action='store_true', help='single CPU executor (no GPU needed)')
    parser.add_argument('--samples', type=int, default=50000, help='training samples when MNIST is generated')
    args = parser.parse_args()
    opt = make_optimizer(args.opt, args.learning_rate)
    print('Loading MNIST data...')
    have = os.path.exists('mnist.pkl.gz')
    datasets = ht.data.mnist(synthetic=not have, n_train=args.samples, n_valid=max(args.samples // 5, args.batch_size))
    (train_set_x, train_set_y), (valid_set_x, valid_set_y) = datasets[0], datasets[1]
    with ht.context(contexts(args.config, args.cpu)):
        print('Building model...')
        x = ht.dataloader_op([ht.Dataloader(train_set_x, args.batch_size, 'train'),
                              ht.Dataloader(valid_set_x, args.batch_size, 'validate')])
        y_ = ht.dataloader_op([ht.Dataloader(train_set_y, args.batch_size, 'train'),
                               ht.Dataloader(valid_set_y, args.batch_size, 'validate')])
        loss, y = zoo.mlp(x, y_, in_dim=784)
        train_op = opt.minimize(loss)
        executor = ht.Executor({'train': [loss, y, train_op], 'validate': [loss, y, y_]})
        n_train_batches = executor.get_batch_num('train')
        n_valid_batches = executor.get_batch_num('validate')
    print('Start training loop...')
    for i in range(args.num_epochs):
        print('Epoch %d' % i)
        loss_all = 0
        start = time()
        for _ in range(n_train_batches):
            loss_val, predict_y, _ = executor.run('train')
            loss_all += float(loss_val.asnumpy()) * x.dataloaders['train'].last_batch_size
        print('Loss = %f' % (loss_all / (n_train_batches * args.batch_size)))
        if args.timing:
            print('Time = %f' % (time() - start))
        if args.validate:
            correct = []
            for _ in range(n_valid_batches):
                loss_val, pred, y_val = executor.run('validate', convert_to_numpy_ret_vals=True)
                correct.extend(np.equal(np.argmax(y_val, 1), np.argmax(pred, 1)).astype(np.float32))
            print('Validation accuracy = %f' % np.mean(correct))
