"""One rank of the CPU ring all-reduce test."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from hetu_61a7_b200.parallel.ring_comm import RingCommunicator    # noqa: E402

rank, n, base = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
ring = RingCommunicator(rank, [('127.0.0.1', base + r) for r in range(n)])
for length in (1, 7, 1000, 300003):
    rs = [np.random.RandomState(100 * r + length % 97) for r in range(n)]
    parts = [g.standard_normal(length).astype(np.float32) for g in rs]
    mine = parts[rank].copy()
    ring.allreduce(mine, average=False)
    np.testing.assert_allclose(mine, np.sum(parts, axis=0), rtol=1e-5, atol=1e-5)
    mine = parts[rank].copy()
    ring.allreduce(mine)
    np.testing.assert_allclose(mine, np.mean(parts, axis=0), rtol=1e-5, atol=1e-5)
    full = np.zeros(length, np.float32)
    b, e = ring.segment(length)
    full[b:e] = rank + 1
    ring.allgather(full)
    for r in range(n):
        b, e = ring.segment(length, r)
        assert np.all(full[b:e] == r + 1)
ring.close()
print('OK')
