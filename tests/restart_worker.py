"""Worker for the heturun restart test: rank 1 dies once at step 3; after the relaunch both ranks resume from
the checkpoint that rank 0 wrote and finish."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ['HETU_COMM_BACKEND'] = 'gloo'
import hetu as ht                                           # noqa: E402

out_dir = sys.argv[1]
comm = ht.wrapped_mpi_nccl_init()
rank, incarnation = comm.rank, int(os.environ.get('HETU_RESTART_COUNT', '0'))
rs = np.random.RandomState(0)
x, y_ = ht.Variable('x'), ht.Variable('y_')
w = ht.Variable('w', value=(rs.randn(8, 3) * 0.1).astype(np.float32))
loss = ht.reduce_mean_op(ht.softmaxcrossentropy_op(ht.matmul_op(x, w), y_), [0])
train = ht.optim.SGDOptimizer(0.1).minimize(loss)
ex = ht.Executor([loss, train], ctx=ht.cpu(0), comm_mode='AllReduce')
start = 0
marker = os.path.join(out_dir, 'step.txt')
if os.path.exists(marker):
    start = int(open(marker).read())
    ex.load(out_dir, 'ckpt.pkl')
X = rs.randn(10, 16, 8).astype(np.float32)
Y = np.eye(3, dtype=np.float32)[rs.randint(0, 3, (10, 16))]
for step in range(start, 10):
    if step == 3 and rank == 1 and incarnation == 0:
        os._exit(17)                                        # simulated crash
    l = float(ex.run(feed_dict={x: X[step], y_: Y[step]})[0].asnumpy())
    comm.barrier()
    if rank == 0:
        ex.save(out_dir, 'ckpt.pkl')
        with open(marker, 'w') as f:
            f.write(str(step + 1))
    comm.barrier()
print('rank %d incarnation %d finished from step %d, loss %.4f' % (rank, incarnation, start, l), flush=True)
