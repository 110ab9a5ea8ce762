"""Shared pieces of the parallel-equivalence suite (reference examples/runner/parallel/
test_mlp_{base,pp,mp,mp_pp}.py + validate_results.py): a 5-layer MLP with FIXED weights and
FIXED feeds, built under different device placements; every variant writes its loss curve to
``results/<name>.npy`` and ``validate_results.py`` asserts they all equal the single-device run.

Weights come from a seeded generator instead of the reference's ``std/*.npy`` files."""
import os
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..', '..')
sys.path.insert(0, ROOT)
if os.environ.get('HETU_CPU', '0') == '1':                   # HETU_CPU=1: gloo world on CPU, same scripts
    os.environ.setdefault('HETU_COMM_BACKEND', 'gloo')
import hetu as ht                                             # noqa: E402

DIMS = [784, 256, 256, 256, 256, 10]
BATCH = 128
HERE = os.path.dirname(os.path.abspath(__file__))


def fixed_weights():
    rs = np.random.RandomState(123)
    return [(rs.randn(DIMS[i], DIMS[i + 1]) * np.sqrt(2.0 / DIMS[i])).astype(np.float32) for i in range(len(DIMS) - 1)]


def fixed_feeds(steps):
    rs = np.random.RandomState(7)
    xs = rs.randn(steps, BATCH, DIMS[0]).astype(np.float32)
    ys = np.eye(DIMS[-1], dtype=np.float32)[rs.randint(0, DIMS[-1], (steps, BATCH))]
    return xs, ys


def build(placement, lr=0.05):
    """``placement[i]`` = context of layer i: a device (ht.gpu(k)) or a tuple of devices with the
    split kind, e.g. ((ht.gpu(1), ht.gpu(2)), 'middle').  Returns x, y_, loss, train_op."""
    W = fixed_weights()
    kinds = {'left': ((2, 1), (1, 1)), 'right': ((1, 1), (1, 2)), 'middle': ((1, 2), (2, 1))}
    x = y_ = None
    act = None
    n = len(W)
    prev_split = False
    for i in range(n):
        ctx = placement[i]
        split = None
        if isinstance(ctx, tuple) and len(ctx) == 2 and isinstance(ctx[1], str):
            ctx, split = ctx
        with ht.context(ctx):
            if i == 0:
                x = ht.Variable('x')
                act = x
            if prev_split and split is None:
                act = ht.dispatch(act, (1, 1))
            w = ht.Variable('mlp_fc%d_weight' % i, value=W[i].copy())
            if split is not None:
                a_part, w_part = kinds[split]
                act = ht.matmul_op(ht.dispatch(act, a_part), ht.dispatch(w, w_part))
            else:
                act = ht.matmul_op(act, w)
            if i < n - 1:
                act = ht.relu_op(act)
            prev_split = split is not None
            if i == n - 1:
                if prev_split:
                    act = ht.dispatch(act, (1, 1))
                y_ = ht.Variable('y_')
                loss = ht.reduce_mean_op(ht.softmaxcrossentropy_op(act, y_), [0])
                train = ht.optim.SGDOptimizer(lr).minimize(loss)
    return x, y_, loss, train


def run_and_save(name, x, y_, loss, train, steps=20, pipeline=None, micro=4):
    comm = ht.wrapped_mpi_nccl_init()
    xs, ys = fixed_feeds(steps)
    losses = []
    if pipeline is None:
        ex = ht.Executor({'train': [loss, train]}, seed=0)
        for s in range(steps):
            out = ex.run('train', feed_dict={x: xs[s], y_: ys[s]})
            if out[0] is not None:
                losses.append(float(out[0].asnumpy()))
    else:
        ex = ht.Executor({'train': [loss, train]}, pipeline=pipeline, seed=0)
        for s in range(0, steps - steps % micro, micro):
            out = ex.run('train', feed_dict={x: [xs[s + m] for m in range(micro)], y_: [ys[s + m] for m in range(micro)]},
                         batch_num=micro)
            if out[0] is not None:
                losses += [float(o.asnumpy()) for o in out[0]]
    if losses:                                                # only the rank holding the loss writes
        os.makedirs(os.path.join(HERE, 'results'), exist_ok=True)
        np.save(os.path.join(HERE, 'results', name + '.npy'), np.asarray(losses, dtype=np.float64))
        print('[%s] rank %d: %d steps, loss %.6f -> %.6f' % (name, comm.rank, len(losses), losses[0], losses[-1]), flush=True)
    comm.barrier()
