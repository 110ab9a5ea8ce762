import sys, numpy as np, torch
sys.path.insert(0, '/root/repo')
import hetu as ht
rs = np.random.RandomState(0)
W1, W2, W3 = (rs.randn(12, 16) * 0.3).astype(np.float32), (rs.randn(16, 16) * 0.3).astype(np.float32), (rs.randn(16, 4) * 0.3).astype(np.float32)
X = rs.randn(4, 8, 12).astype(np.float32)
Y = np.eye(4, dtype=np.float32)[rs.randint(0, 4, (4, 8))]
with ht.context(ht.gpu(0)):
    x = ht.Variable('x'); y_ = ht.Variable('y_')
    w1 = ht.Variable('w1', value=W1.copy()); w2 = ht.Variable('w2', value=W2.copy()); w3 = ht.Variable('w3', value=W3.copy())
    a = ht.relu_op(ht.matmul_op(x, w1)); a = ht.relu_op(ht.matmul_op(a, w2))
    loss = ht.reduce_mean_op(ht.softmaxcrossentropy_op(ht.matmul_op(a, w3), y_), [0])
    train = ht.optim.SGDOptimizer(0.1).minimize(loss)
ex = ht.Executor({'train': [loss, train]}, seed=0)
for step in range(6):
    out = ex.run('train', feed_dict={x: X[step % 4], y_: Y[step % 4]})
    print(step, float(out[0].asnumpy()), flush=True)
torch.cuda.synchronize()
print('TINY OK')
