"""BASELINE config 1: 3-layer MLP (3072-256-256-10), batch 128, CPU executor, synthetic CIFAR10 shape.
usage: cpu_mlp_bench.py [steps]   -> samples/s over the timed steps (host wall clock; there is no device)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import hetu as ht
from hetu_61a7_b200.models import cnn
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 300
rs = np.random.RandomState(0)
x, y_ = ht.Variable('x'), ht.Variable('y_')
loss, y = cnn.mlp(x, y_)
train = ht.optim.SGDOptimizer(0.01).minimize(loss)
ex = ht.Executor([loss, train], ctx=ht.cpu(0))
xs = [rs.randn(128, 3072).astype(np.float32) for _ in range(8)]
ys = [np.eye(10, dtype=np.float32)[rs.randint(0, 10, 128)] for _ in range(8)]
for i in range(20):
    ex.run(feed_dict={x: xs[i % 8], y_: ys[i % 8]})
t0 = time.time()
for i in range(steps):
    out = ex.run(feed_dict={x: xs[i % 8], y_: ys[i % 8]})
l = float(out[0].asnumpy())
dt = time.time() - t0
print('cpu mlp: %.1f us/step, %.0f samples/s, loss %.4f' % (dt / steps * 1e6, 128 * steps / dt, l))
