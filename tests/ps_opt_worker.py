"""Worker for the server-side optimizer test: one step sequence per optimizer against numpy."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from hetu_61a7_b200.parallel.ps import PSWorker          # noqa: E402

w = PSWorker.from_env()
rng = np.random.RandomState(0)
grads = [rng.standard_normal(64).astype(np.float32) for _ in range(4)]
lr, mom, b1, b2, eps = 0.1, 0.9, 0.9, 0.999, 1e-7


def oracle(kind):
    x = np.full(64, 1.0)
    v = np.zeros(64); m = np.zeros(64); acc = np.zeros(64)
    for t, g in enumerate(grads, 1):
        g = g.astype(np.float64)
        if kind == 'sgd':
            x -= lr * g
        elif kind == 'momentum':
            v = mom * v - lr * g; x += v
        elif kind == 'nesterov':
            v = mom * v - lr * g; x += mom * v - lr * g
        elif kind == 'adagrad':
            acc += g * g; x -= lr * g / (np.sqrt(acc) + eps)
        elif kind == 'adam':
            m = b1 * m + (1 - b1) * g; v = b2 * v + (1 - b2) * g * g
            x -= lr * (m / (1 - b1 ** t)) / (np.sqrt(v / (1 - b2 ** t)) + eps)
    return x


for key, (kind, args) in enumerate([('sgd', [lr]), ('momentum', [lr, mom]), ('nesterov', [lr, mom]),
                                    ('adagrad', [lr, 0, 0, eps]), ('adam', [lr, b1, b2, eps])], start=20):
    w.InitTensor(key, 0, 64, 1, 0, 1.0, 1.0, 0, opt=(kind, args))
    for g in grads:
        w.Push(key, g)
    np.testing.assert_allclose(w.Pull(key), oracle(kind), rtol=2e-5, atol=2e-5, err_msg=kind)
# sparse rows: Adam keeps a step count per row
w.InitTensor(40, 1, 10, 4, 0, 0.0, 0.0, 0, opt=('adam', [lr, b1, b2, eps]))
w.SparsePush(40, np.array([2, 7]), np.ones((2, 4), np.float32))
w.SparsePush(40, np.array([2]), np.ones((1, 4), np.float32))
rows = w.SparsePull(40, np.array([2, 7, 3]))
assert np.allclose(rows[0], -2 * lr, atol=1e-5) and np.allclose(rows[1], -lr, atol=1e-5) and np.allclose(rows[2], 0)
# async pool: requests of one key keep their order (push, push, pull sees both), priorities reorder across keys
import time
w.InitTensor(60, 0, 1 << 16, 1, 0, 0.0, 0.0, 0)
w.InitTensor(61, 0, 8, 1, 0, 0.0, 0.0, 0)
one = np.ones(1 << 16, np.float32)
out = np.empty(1 << 16, np.float32)
for rep in range(20):
    w.PushAsync(60, one)
    w.PushAsync(60, one)
    w.PullAsync(60, out)
    w.Wait(60)
    assert np.all(out == 2.0 * (rep + 1)), (rep, out[:3])
w.SetPriority(61, 10)
small = np.empty(8, np.float32)
for _ in range(30):
    w.PushAsync(60, one)
w.PushAsync(61, np.ones(8, np.float32))
w.PullAsync(61, small)
w.Wait(61)
assert np.all(small == 1.0)
w.Wait(60)
assert np.allclose(w.Pull(60), 2.0 * 20 + 30)
print('OK')
w.finalize()
