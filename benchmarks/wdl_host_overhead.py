"""Where a Wide&Deep step's time goes on the host side: bare graph replays vs Executor.run vs the bench loops.
Run on a GPU box: python benchmarks/wdl_host_overhead.py"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import hetu_61a7_b200 as ht  # noqa: E402
from hetu_61a7_b200.models import ctr  # noqa: E402

dev = torch.device('cuda', 0)
ctx = ht.gpu(0)
B, vocab = 128, int(os.environ.get('VOCAB', 33762577))
pool = [ht.data.criteo_synthetic(B, vocab=vocab, seed=i) for i in range(32)]
dense = ht.dataloader_op([ht.Dataloader(np.concatenate([p[0] for p in pool]), B, 'train')])
sparse = ht.dataloader_op([ht.Dataloader(np.concatenate([p[1] for p in pool]), B, 'train', dtype=np.int64)])
y_ = ht.dataloader_op([ht.Dataloader(np.concatenate([p[2] for p in pool]), B, 'train')])
loss, pred, _, train_op = ctr.wdl_criteo(dense, sparse, y_, feature_dimension=vocab, ctx=ctx)
ex = ht.Executor({'train': [loss, pred, train_op]}, ctx=ctx, seed=123)
d = ht.array(pool[0][0], ctx)
s = ht.NDArray(torch.from_numpy(pool[0][1]).to(dev))
l = ht.array(pool[0][2], ctx)
fd = {dense: d, sparse: s, y_: l}
for _ in range(5):
    ex.run('train', feed_dict=fd)
sub = ex.subexecutor['train']


def timeit(fn, n=300):
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    a.record()
    for i in range(n):
        fn(i)
    b.record()
    t_host = (time.perf_counter() - t0) / n * 1e6
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3, t_host


print('launches per step', sub.graph_launches)
print('graph.replay only        : %.1f us/step device, %.1f us host' % timeit(lambda i: sub._graph.replay()))
print('Executor.run, device feed: %.1f us/step device, %.1f us host' % timeit(lambda i: ex.run('train', feed_dict=fd)))
for _ in range(5):
    ex.run('train')
print('Executor.run, dataloaders: %.1f us/step device, %.1f us host' % timeit(lambda i: ex.run('train')))
