"""Worker launched by heturun in test_launch.py: WDL-like model in PS / Hybrid mode on CPU."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault('HETU_COMM_BACKEND', 'gloo')
import hetu as ht                                          # noqa: E402

mode, policy = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 and sys.argv[2] != 'none' else None)
from hetu_61a7_b200.models.ctr import wdl_criteo           # noqa: E402
prefetch = len(sys.argv) > 4 and sys.argv[4] == 'prefetch'
rank0 = int(os.environ.get('DMLC_WORKER_ID', os.environ.get('RANK', '0')) or 0)
d, s, l = ht.data.criteo_synthetic(64, vocab=2000, seed=rank0)
l = (d[:, :1] > 0.5).astype(np.float32)                    # learnable labels
dense, y_ = ht.Variable('dense'), ht.Variable('y_')
if prefetch:      # ids come from a dataloader: the executor can peek the next batch and pull its rows early (ASP)
    sparse = ht.dataloader_op([ht.Dataloader(np.tile(s.astype(np.float32), (4, 1)), 64, 'train')])
else:
    sparse = ht.Variable('sparse')
loss, pred, _, train = wdl_criteo(dense, sparse, y_, feature_dimension=2000, embedding_size=8, learning_rate=0.2)
ex = ht.Executor({'train': [loss, pred, train]}, ctx=ht.cpu(0), comm_mode=mode, cstable_policy=policy,
                 cache_bound=1, bsp=-1 if prefetch else 0, seed=3, shard_dataloaders=False)
rank = ex.rank
losses = []
for it in range(60):
    fd = {dense: d, y_: l} if prefetch else {dense: d, sparse: s.astype(np.float32), y_: l}
    out = ex.run('train', feed_dict=fd)
    losses.append(float(out[0].asnumpy().reshape(-1)[0]))
if prefetch:
    from hetu_61a7_b200.ops.embedding import EmbeddingLookUp
    hits = sum(getattr(n, 'prefetch_hits', 0) for n in ex.topo_order if isinstance(n, EmbeddingLookUp))
    assert hits >= 50, 'embedding rows were not prefetched (%d hits)' % hits
assert losses[-1] < losses[0] - 0.0003, (losses[0], losses[-1])
ex.save(sys.argv[3], 'ckpt_%s.bin' % mode)
print('TRAIN %s rank %d OK %.4f -> %.4f' % (mode, rank, losses[0], losses[-1]), flush=True)
if ex.config.ps_comm is not None:
    ex.config.ps_comm.BarrierWorker()
