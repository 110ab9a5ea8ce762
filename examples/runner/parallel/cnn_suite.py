"""CNN counterpart of mlp_suite.py (reference examples/runner/parallel/test_model_cnn*.py + all_cnn_tests.sh):
a small conv net with FIXED weights / feeds under different stage placements; loss curves go to
``results/cnn_<name>.npy`` and validate_results.py compares them with ``cnn_base``."""
import os

import numpy as np

from mlp_suite import HERE, ht

BATCH = 32


def fixed_weights():
    rs = np.random.RandomState(321)
    conv = lambda o, i, k: (rs.randn(o, i, k, k) * np.sqrt(2.0 / (i * k * k))).astype(np.float32)     # noqa: E731
    return dict(c1=conv(8, 3, 3), c2=conv(16, 8, 3), c3=conv(16, 16, 3),
                fc1=(rs.randn(16 * 4 * 4, 64) * np.sqrt(2.0 / 256)).astype(np.float32),
                fc2=(rs.randn(64, 10) * np.sqrt(2.0 / 64)).astype(np.float32))


def fixed_feeds(steps):
    rs = np.random.RandomState(9)
    xs = rs.randn(steps, BATCH, 3, 16, 16).astype(np.float32)
    ys = np.eye(10, dtype=np.float32)[rs.randint(0, 10, (steps, BATCH))]
    return xs, ys


def build(stage_ctx, lr=0.05):
    """``stage_ctx`` = contexts of the 5 blocks (conv1+pool, conv2+pool, conv3, fc1, fc2+loss)."""
    W = fixed_weights()
    with ht.context(stage_ctx[0]):
        x = ht.Variable('x')
        a = ht.relu_op(ht.conv2d_op(x, ht.Variable('cnn_c1', value=W['c1'].copy()), padding=1, stride=1))
        a = ht.max_pool2d_op(a, 2, 2, 0, 2)
    with ht.context(stage_ctx[1]):
        a = ht.relu_op(ht.conv2d_op(a, ht.Variable('cnn_c2', value=W['c2'].copy()), padding=1, stride=1))
        a = ht.max_pool2d_op(a, 2, 2, 0, 2)
    with ht.context(stage_ctx[2]):
        a = ht.relu_op(ht.conv2d_op(a, ht.Variable('cnn_c3', value=W['c3'].copy()), padding=1, stride=1))
        a = ht.array_reshape_op(a, (-1, 16 * 4 * 4))
    with ht.context(stage_ctx[3]):
        a = ht.relu_op(ht.matmul_op(a, ht.Variable('cnn_fc1', value=W['fc1'].copy())))
    with ht.context(stage_ctx[4]):
        y_ = ht.Variable('y_')
        logits = ht.matmul_op(a, ht.Variable('cnn_fc2', value=W['fc2'].copy()))
        loss = ht.reduce_mean_op(ht.softmaxcrossentropy_op(logits, y_), [0])
        train = ht.optim.SGDOptimizer(lr).minimize(loss)
    return x, y_, loss, train


def run_and_save(name, x, y_, loss, train, steps=12, pipeline=None, micro=4):
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
    if losses:
        os.makedirs(os.path.join(HERE, 'results'), exist_ok=True)
        np.save(os.path.join(HERE, 'results', 'cnn_' + name + '.npy'), np.asarray(losses, dtype=np.float64))
        print('[cnn_%s] rank %d: %d steps, loss %.6f -> %.6f' % (name, comm.rank, len(losses), losses[0], losses[-1]), flush=True)
    comm.barrier()
