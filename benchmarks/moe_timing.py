"""Per-op GPU timing of the MoE step at N ranks (TimerSubExecutor, CUDA events per node, rank 0 prints)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import hetu as ht
from hetu_61a7_b200.models import moe

comm = ht.wrapped_mpi_nccl_init()
world, rank = comm.nrank, comm.rank
ctx = ht.gpu(comm.local_rank)
x, t = ht.Variable(name='x'), ht.Variable(name='t')
loss, _ = moe.moe_model(x, t, world=world, rank=rank, top=2)
train = ht.optim.SGDOptimizer(learning_rate=1e-4).minimize(loss)
ex = ht.Executor({'train': [loss, train]}, ctx=ctx, comm_mode='AllReduce' if world > 1 else None, timing='gpu')
xv = ht.array(np.random.RandomState(rank).randn(16, 1024, 2048).astype(np.float32), ctx)
tv = ht.array(np.zeros((16,), np.float32), ctx)
for i in range(3):
    ex.run('train', feed_dict={x: xv, t: tv})
ex.subexecutor['train'].clearTimer()
for i in range(5):
    ex.run('train', feed_dict={x: xv, t: tv})
path = 'gpurun_out/moe_timing_n%d_r%d.txt' % (world, rank)
os.makedirs('gpurun_out', exist_ok=True)
ex.subexecutor['train'].logOut(path, log_level='node')
if rank == 0:
    rows = [l.split() for l in open(path)]
    rows = [(r[0], r[1], float(r[2])) for r in rows]
    print('total %.3f ms' % sum(r[2] for r in rows))
    for r in sorted(rows, key=lambda r: -r[2])[:28]:
        print('%8.3f ms  %-28s %s' % (r[2], r[1], r[0]))
