"""Launched by heturun (test_launch.py): checkpoint round trip in PS / Hybrid mode.

train 5 steps -> save -> 1 step (loss A) -> 3 more steps (parameters drift) -> load -> 1 step (loss B).
A == B only if load() restored the SERVERS' copy of every parameter (dense ones included), refreshed the device
replicas and dropped the HET cache contents.  Fixed batch, BSP, so the loss depends on the parameters alone.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault('HETU_COMM_BACKEND', 'gloo')
import hetu as ht                                          # noqa: E402
from hetu_61a7_b200.models.ctr import wdl_criteo           # noqa: E402

mode, policy, out_dir = sys.argv[1], (sys.argv[2] if sys.argv[2] != 'none' else None), sys.argv[3]
rank0 = int(os.environ.get('DMLC_WORKER_ID', os.environ.get('RANK', '0')) or 0)
d, s, l = ht.data.criteo_synthetic(64, vocab=2000, seed=rank0)
l = (d[:, :1] > 0.5).astype(np.float32)
dense, sparse, y_ = ht.Variable('dense'), ht.Variable('sparse'), ht.Variable('y_')
loss, pred, _, train = wdl_criteo(dense, sparse, y_, feature_dimension=2000, embedding_size=8, learning_rate=0.2)
ex = ht.Executor({'train': [loss, pred, train]}, ctx=ht.cpu(0), comm_mode=mode, cstable_policy=policy,
                 cache_bound=1, bsp=0, seed=3, shard_dataloaders=False)
fd = {dense: d, sparse: s.astype(np.float32), y_: l}


def step():
    return float(ex.run('train', feed_dict=fd)[0].asnumpy().reshape(-1)[0])


for _ in range(5):
    step()
ex.save(out_dir, 'ckpt.bin')
a = step()
for _ in range(3):
    step()
drifted = step()
ex.load(out_dir, 'ckpt.bin')
b = step()
assert abs(a - b) < 2e-6, 'restored run diverges: %.7f vs %.7f (drifted %.7f)' % (a, b, drifted)
assert abs(a - drifted) > 1e-5, 'the test needs the parameters to move: %.7f %.7f' % (a, drifted)
print('CKPT %s rank %d OK %.6f == %.6f (drifted %.6f)' % (mode, ex.rank, a, b, drifted), flush=True)
if ex.config.ps_comm is not None:
    ex.config.ps_comm.BarrierWorker()
