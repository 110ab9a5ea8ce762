"""Worker process for the PS fault-injection test: pushes through a lossy server and checks that
every update is applied exactly once."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from hetu_61a7_b200.parallel.ps import PSWorker          # noqa: E402

w = PSWorker.from_env()
w.InitTensor(5, 0, 256, 1, 0, 0.0, 0.0, 3)
w.InitTensor(6, 1, 64, 4, 0, 0.0, 0.0, 3)
n = 60
for i in range(n):
    w.Push(5, np.full(256, 1.0, np.float32))
    w.SparsePush(6, np.array([1, 7, 63], dtype=np.int64), np.ones((3, 4), np.float32))
v = w.Pull(5)
assert np.allclose(v, float(n)), (v[:4], n)
r = w.SparsePull(6, np.array([1, 7, 63, 2], dtype=np.int64))
assert np.allclose(r[:3], float(n)) and np.allclose(r[3], 0.0), r
assert w.num_retransmits() > 0, 'PS_DROP_MSG did not lose anything'
print('OK retransmits=%d' % w.num_retransmits())
w.finalize()
