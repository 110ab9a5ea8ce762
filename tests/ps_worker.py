"""PS worker process for test_ps.py (pattern: reference tests/pstests/test_apis.py)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from hetu_61a7_b200.parallel.ps import PSWorker          # noqa: E402
from hetu_61a7_b200.parallel.cstable import CacheSparseTable   # noqa: E402

w = PSWorker.from_env()
rank, nrank = w.rank(), w.nrank()
out_dir = sys.argv[1]
# dense: constant init 1.0, every worker pushes -0.5 -> 1 - 0.5 * nrank
w.InitTensor(1, 0, 1000, 1, 0, 1.0, 1.0, 7)
w.BarrierWorker()
w.Push(1, np.full(1000, -0.5, np.float32))
w.BarrierWorker()
v = w.Pull(1)
assert np.allclose(v, 1.0 - 0.5 * nrank), v[:4]
# DDPushPull returns the post-push value
w.BarrierWorker()
v2 = w.DDPushPull(1, np.full(1000, 0.25, np.float32))
w.BarrierWorker()
assert np.allclose(w.Pull(1), 1.0 - 0.5 * nrank + 0.25 * nrank)
# sparse 2-D table: normal init, SparsePush on rows then SparsePull
w.InitTensor(2, 1, 500, 8, 2, 0.0, 0.01, 11)
w.BarrierWorker()
rows = np.array([3, 499, 250, 3], dtype=np.int64)
before = w.SparsePull(2, rows)
assert np.allclose(before[0], before[3]) and np.abs(before).max() < 0.1
w.BarrierWorker()
w.SparsePush(2, rows, np.ones((4, 8), np.float32))
w.BarrierWorker()
after = w.SparsePull(2, rows)
assert np.allclose(after[1] - before[1], nrank * 1.0, atol=1e-5)
assert np.allclose(after[0] - before[0], nrank * 2.0, atol=1e-5)       # row 3 pushed twice per worker
w.BarrierWorker()
# SSPushPull / SDPushPull
got = w.SSPushPull(2, rows[:2], np.zeros((2, 8), np.float32), np.array([250], dtype=np.int64))
assert np.allclose(got[0], after[2], atol=1e-5)
full = w.SDPushPull(2, rows[:1], np.zeros((1, 8), np.float32))
assert full.shape == (500 * 8,) and np.allclose(full.reshape(500, 8)[499], after[1], atol=1e-5)
# save / load on the servers
w.BarrierWorker()
if rank == 0:
    w.SaveParam(2, out_dir)
w.BarrierWorker()
w.SparsePush(2, rows[:1], np.ones((1, 8), np.float32))
w.BarrierWorker()
if rank == 0:
    w.LoadParam(2, out_dir)
w.BarrierWorker()
assert np.allclose(w.SparsePull(2, rows[:1])[0], after[0], atol=1e-5)
# a push whose data does not hold rows x width floats is refused before it reaches the wire (the server checks the same
# on its side), a checkpoint directory that climbs out with '..' is refused by the server, which keeps serving
w.BarrierWorker()
try:
    w.SparsePush(2, rows[:2], np.ones((2, 5), np.float32))
    raise AssertionError('a sparse push with the wrong width must be refused')
except ValueError as exc:
    assert 'rows of width' in str(exc), exc
try:
    w.SaveParam(2, out_dir + '/../escape')
    raise AssertionError("a checkpoint path with '..' must be refused")
except RuntimeError as exc:
    assert '..' in str(exc), exc
assert np.allclose(w.SparsePull(2, rows[:1])[0], after[0], atol=1e-5)
w.BarrierWorker()
# SSP: nobody may run more than 1 clock ahead of the slowest
w.ssp_init(rank, nrank, 1)
import time
for clock in range(1, 5):
    if rank == 0:
        time.sleep(0.05)
    w.ssp_sync(rank, clock)
# partial reduce matchmaking: all workers arrive -> one group of everybody
grp = w.preduce_get_partner(0, rank, nrank, 2000.0)
assert sorted(grp) == list(range(nrank)), grp
# HET cache on a versioned table
w.InitTensor(3, 2, 200, 4, 0, 0.0, 1.0, 5)
w.BarrierWorker()
cache = CacheSparseTable(limit=50, length=200, width=4, node_id=3, policy=['LRU', 'LFU', 'LFUOpt'][rank % 3], bound=2)
rng = np.random.RandomState(rank)
ref = np.zeros((200, 4), np.float32)
for it in range(30):
    keys = rng.randint(0, 200, 16)
    emb = cache.embedding_lookup(keys, sync=True)
    assert emb.shape == (16, 4)
    g = np.full((16, 4), -0.01, np.float32)
    cache.embedding_update(keys, g, sync=True)
    np.add.at(ref, keys, g)
p = cache.perf
assert p['num_lookup'] > 0 and p['num_miss'] > 0 and p['num_evict'] >= 0
# flush the un-pushed updates, then compare server state with the sum over workers
assert cache.flush() >= 0
assert cache.flush() == 0                      # nothing left
w.BarrierWorker()
np.save(os.path.join(out_dir, 'ref_%d.npy' % rank), ref)
w.BarrierWorker()
total = sum(np.load(os.path.join(out_dir, 'ref_%d.npy' % r)) for r in range(nrank))
srv = w.SparsePull(3, np.arange(200))
assert np.allclose(srv, total, atol=1e-4), np.abs(srv - total).max()
fut = cache.embedding_push_pull(np.arange(8), None, np.arange(4), np.zeros((4, 4), np.float32))
fut.wait()
w.startRecord(out_dir)
w.Pull(1)
s, r = w.getLoads()
assert r > 4000
by = w.getLoadsByFunction()
assert set(by) == {'DensePull'} and by['DensePull'][1] == r and by['DensePull'][2] == 2          # one request per server
assert 'DensePull:' in open(os.path.join(out_dir, 'loads_%d.txt' % rank)).read()
w.BarrierWorker()
print('PS WORKER %d OK' % rank, flush=True)
w.finalize()
