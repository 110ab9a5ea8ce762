"""cProfile of the host side of the Wide&Deep end-to-end step (dataloader -> H2D -> graph replay -> loss D2H)."""
import cProfile, os, pstats, sys, io
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import hetu as ht
from hetu_61a7_b200.models import ctr
ctx = ht.gpu(0)
B, npool, vocab = 128, 32, 33762577
pool = [ht.data.criteo_synthetic(B, vocab=vocab, seed=i) for i in range(npool)]
dense = ht.dataloader_op([ht.Dataloader(np.concatenate([p[0] for p in pool]), B, 'train')])
sparse = ht.dataloader_op([ht.Dataloader(np.concatenate([p[1] for p in pool]).astype(np.int64), B, 'train', dtype=np.int64)])
y_ = ht.dataloader_op([ht.Dataloader(np.concatenate([p[2] for p in pool]), B, 'train')])
loss, pred, _, train_op = ctr.wdl_criteo(dense, sparse, y_, feature_dimension=vocab, ctx=ctx)
ex = ht.Executor({'train': [loss, pred, train_op]}, ctx=ctx, seed=1)
for _ in range(20):
    ex.run('train')[0].asnumpy()
torch.cuda.synchronize()
import time
t0 = time.perf_counter()
for _ in range(500):
    ex.run('train')[0].asnumpy()
dt = (time.perf_counter() - t0) / 500
print('e2e %.1f us/step' % (dt * 1e6))
t0 = time.perf_counter()
for _ in range(500):
    ex.run('train')
torch.cuda.synchronize()
print('no-readback %.1f us/step (host launch rate)' % ((time.perf_counter() - t0) / 500 * 1e6))
pending = None
t0 = time.perf_counter()
for _ in range(500):
    fut = ex.run('train')[0].asnumpy_async()
    if pending is not None:
        pending.get()
    pending = fut
torch.cuda.synchronize()
print('async read-back %.1f us/step' % ((time.perf_counter() - t0) / 500 * 1e6))
pr = cProfile.Profile()
pr.enable()
for _ in range(2000):
    fut = ex.run('train')[0].asnumpy_async()
    if pending is not None:
        pending.get()
    pending = fut
pr.disable()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats('tottime').print_stats(32)
print(s.getvalue()[:9000])
