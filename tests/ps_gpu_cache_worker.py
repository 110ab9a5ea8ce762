"""Worker for test_zt_het_gpu_cache.py: the GPU-resident HET cache against a numpy model of the table.

One worker, so the cache's own view (server row + kept local gradient) must equal the true table at every lookup, for any
staleness bound and also when the cache is far smaller than the working set (evictions push their gradients first).  A
second worker id ("rival") writes to the server behind the cache's back to check the pull bound: a cached row may lag by
at most `bound` versions."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from hetu_61a7_b200.parallel.ps import PSWorker            # noqa: E402
from hetu_61a7_b200.parallel.gpu_cache import GPUHetCache  # noqa: E402

w = PSWorker.from_env()
V, D = 4000, 32
dev = torch.device('cuda', 0)
rs = np.random.RandomState(0)
for case, (policy, limit, bound) in enumerate([('LRU', 4000, 0), ('LFU', 256, 3), ('LFUOpt', 512, 5)]):
    node = 100 + case
    w.InitTensor(node, 2, V, D, 2, 0.0, 0.01, 17 + case)     # cached embedding table, normal init on the server
    w.BarrierWorker()
    truth = w.SparsePull(node, np.arange(V)).copy()
    cache = GPUHetCache(limit, V, D, node, policy, bound, comm=w, device=dev)
    for step in range(30):
        ids = (rs.zipf(1.3, size=(16, 6)) - 1) % V
        rows = cache.device_lookup(torch.from_numpy(ids).to(dev))
        np.testing.assert_allclose(rows.cpu().numpy(), truth[ids], rtol=1e-5, atol=1e-6, err_msg='%s step %d' % (policy, step))
        delta = (rs.randn(ids.size, D) * 0.01).astype(np.float32)
        cache.device_update(torch.from_numpy(ids).to(dev), torch.from_numpy(delta).to(dev))
        np.add.at(truth, ids.reshape(-1), delta)
    perf = cache.perf()
    assert perf['num_lookup'] == 30 * 96 and 0 < perf['num_miss'] < perf['num_lookup'], perf
    if limit < V:
        assert cache.size() <= limit and (policy != 'LFU' or perf['num_evict'] > 0), (cache.size(), perf)
    cache.flush()
    np.testing.assert_allclose(w.SparsePull(node, np.arange(V)), truth, rtol=1e-5, atol=1e-5, err_msg='%s after flush' % policy)
    print('case %s: miss rate %.3f, pulled %d rows, pushed %d rows, evictions %d' % (
        policy, cache.overall_miss_rate(), perf['num_transfered'], perf['num_pushed'], perf['num_evict']), flush=True)

# pull bound: somebody else updates row 5 on the server; the cached copy is refreshed once it is more than `bound` behind
node = 200
w.InitTensor(node, 2, 64, 8, 0, 1.0, 1.0, 3)
w.BarrierWorker()
cache = GPUHetCache(64, 64, 8, node, 'LRU', 2, comm=w, device=dev)
five = torch.tensor([5], device=dev)
assert np.allclose(cache.device_lookup(five).cpu().numpy(), 1.0)
for k in range(1, 5):
    w.push_embedding(node, np.array([5]), np.full((1, 8), 0.5, np.float32), np.array([1]))      # the rival's pushed update
    got = float(cache.device_lookup(five).cpu().numpy()[0, 0])
    want = {1: 1.0, 2: 1.0, 3: 2.5, 4: 2.5}[k]                   # served stale while <= 2 versions behind; the third push
                                                                 # triggers a refresh (version 3), the fourth is 1 behind again
    assert abs(got - want) < 1e-6, (k, got, want)
print('GPU HET CACHE OK', flush=True)
w.BarrierWorker()
