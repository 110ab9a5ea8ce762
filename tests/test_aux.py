"""Auxiliary subsystems: per-node timer executor, op micro-profiler, collective profiler (loop-back
world), distribution strategies, GNN ops.  Pattern: reference tests/test_profiler.py (smoke) and
examples/gnn scripts."""
import numpy as np
import pytest
import scipy.sparse as sp
import torch

import hetu as ht


def _mlp():
    x, y_ = ht.Variable('x'), ht.Variable('y_')
    h = ht.relu_op(ht.matmul_op(x, ht.init.xavier_normal((20, 16), name='w1')))
    logits = ht.matmul_op(h, ht.init.xavier_normal((16, 4), name='w2'))
    loss = ht.reduce_mean_op(ht.softmaxcrossentropy_op(logits, y_), [0])
    return x, y_, loss, ht.optim.SGDOptimizer(0.1).minimize(loss)


def test_timer_executor_and_profiler(tmp_path):
    rs = np.random.RandomState(0)
    x, y_, loss, train = _mlp()
    ex = ht.Executor([loss, train], ctx=ht.cpu(0), timing='cpu')
    X, Y = rs.randn(8, 20).astype(np.float32), np.eye(4, dtype=np.float32)[rs.randint(0, 4, 8)]
    for _ in range(3):
        ex.run(feed_dict={x: X, y_: Y})
    import json
    trace_path = str(tmp_path / 'trace.json')
    n_events = ex.export_chrome_trace(trace_path)
    ev = json.load(open(trace_path))['traceEvents']
    runs = [e for e in ev if e['ph'] == 'X']
    assert n_events == len(ev) and len(runs) >= 3 * 5 and all(e['dur'] >= 0 for e in runs)
    assert any(e['cat'] == 'MatMulOp' for e in runs) and any(e['ph'] == 'M' for e in ev)
    assert [e['ts'] for e in runs] == sorted(e['ts'] for e in runs)
    for level in ('node', 'type'):
        path = str(tmp_path / ('timer_%s.txt' % level))
        ex.logOut(path, log_level=level)
        txt = open(path).read()
        assert 'MatMulOp' in txt and 'ms' in txt
    # op micro-profiler (Executor.profile -> HetuProfiler.profile_n_log)
    x, y_, loss, train = _mlp()
    ex = ht.Executor([loss, train], ctx=ht.cpu(0))
    ex.run(feed_dict={x: X, y_: Y})
    path = str(tmp_path / 'profile.txt')
    ex.profile({x: (8, 20), y_: (8, 4)}, path, profiler='cpu')
    lines = open(path).read().strip().splitlines()
    assert len(lines) >= 3 and all(l.endswith('ms') for l in lines)


def test_nccl_profiler_loopback():
    prof = ht.NCCLProfiler()
    for prim in ('allreduce', 'allgather', 'reducescatter', 'broadcast', 'reduce', 'alltoall'):
        res = prof.profile(prim, [1024, 4096])
        assert set(res) == {1024, 4096} and all(v >= 0 for v in res.values())


def test_strategies_assign_contexts(monkeypatch):
    from hetu_61a7_b200.distributed_strategies import hybrid, simple
    x, y_, loss, train = _mlp()
    monkeypatch.setattr(simple.DataParallel, 'worker_devices', lambda self: ['gpu:0', 'gpu:1', 'gpu:2', 'gpu:3'], raising=False)
    monkeypatch.setattr(hybrid.ModelParallel, 'worker_devices', lambda self: ['gpu:0', 'gpu:1', 'gpu:2', 'gpu:3'], raising=False)
    mp = hybrid.ModelParallel(tp=2)
    ctx = mp.set_raw_ctxs_n_states([loss, train])
    assert ctx.is_mp and ctx.worker_num == 2 and all(isinstance(w, tuple) and len(w) == 2 for w in ctx.workers)
    assert loss.raw_ctx is ctx
    plan = hybrid.LayerPlan(dp=2, tp=2, pp_stage=1)
    assert 'tp=2' in repr(plan)
    g = hybrid._stage_group(['gpu:0', 'gpu:1', 'gpu:2', 'gpu:3'], 2, 2)
    assert g.worker_num == 2 and g.mp_dev_num == 2


def test_gnn_ops_match_dense():
    rs = np.random.RandomState(0)
    N, F, O = 30, 8, 5
    A = sp.random(N, N, 0.2, random_state=rs, format='coo', dtype=np.float32)
    H, W = rs.randn(N, F).astype(np.float32), rs.randn(F, O).astype(np.float32)
    adj = ht.Variable('adj', trainable=False)
    h, w = ht.Variable('h', value=H), ht.Variable('w', value=W)
    out = ht.csrmm_op(adj, ht.matmul_op(h, w))
    loss = ht.reduce_sum_op(ht.mul_op(out, out), [0, 1])
    gh, gw = ht.gradients(loss, [h, w])
    ex = ht.Executor([out, gh, gw], ctx=ht.cpu(0))
    mp = ht.sparse_array(A.data, (A.row, A.col), shape=(N, N), ctx=ht.cpu(0))
    o, g1, g2 = [r.asnumpy() for r in ex.run(feed_dict={adj: mp})]
    Ad = torch.tensor(A.toarray())
    th, tw = torch.tensor(H, requires_grad=True), torch.tensor(W, requires_grad=True)
    ref = Ad @ (th @ tw)
    (ref * ref).sum().backward()
    np.testing.assert_allclose(o, ref.detach().numpy(), rtol=1e-4, atol=1e-4)
    np.testing.assert_allclose(g1, th.grad.numpy(), rtol=1e-3, atol=1e-3)
    np.testing.assert_allclose(g2, tw.grad.numpy(), rtol=1e-3, atol=1e-3)
    # 1.5-D distributed GCN layer on a single-rank world == plain A (H W)
    comm = ht.wrapped_mpi_nccl_init()
    z = ht.distgcn_15d_op(adj, h, w, N, N, 1, 1, 0, comm, [None, None], True)
    ex = ht.Executor([z], ctx=ht.cpu(0))
    np.testing.assert_allclose(ex.run(feed_dict={adj: mp})[0].asnumpy(), (Ad @ torch.tensor(H)) @ torch.tensor(W), rtol=1e-4, atol=1e-4)


def test_bfc_allocator_and_memory_plan():
    from hetu_61a7_b200.memory_pool import BFCAllocator, HetuMemoryPool
    a = BFCAllocator(1 << 20, alignment=256)
    o1, o2, o3 = a.alloc(1000), a.alloc(5000), a.alloc(300)
    assert o1 == 0 and o2 == 1024 and o3 == 1024 + 5120
    a.free(o2)
    assert a.alloc(4096) == o2                          # best fit reuses the hole
    assert a.alloc(2048) == o3 + 512                    # the hole's 1 KB remainder is too small
    st = a.stats()
    assert st['live_blocks'] == 4 and st['in_use'] == 1024 + 4096 + 512 + 2048
    for o in (o1, o2, o3, o3 + 512):
        a.free(o)
    st = a.stats()
    assert st['in_use'] == 0 and st['free_fragments'] == 1 and st['largest_free'] == 1 << 20     # fully coalesced
    assert BFCAllocator(4096).alloc(8192) == -1
    with pytest.raises(ValueError):
        a.free(12345)
    # planner: peak of a 3-layer MLP step is far below the sum of all buffers
    x, y_, loss, train = _mlp()
    ex = ht.Executor([loss, train], ctx=ht.cpu(0))
    rep = ex.subexecutor['default'].memory_plan({x: (4096, 20), y_: (4096, 4)})
    pool = HetuMemoryPool()
    shapes = ex.subexecutor['default'].infer_shape({x: (4096, 20), y_: (4096, 4)})
    plan = pool.memory_plan(ex.subexecutor['default'].topo_order, shapes, eval_nodes=[loss])
    total = sum(sz for _, sz in plan.values())
    assert rep['parameter_bytes'] == (20 * 16 + 16 * 4) * 4
    assert 0 < rep['activation_peak_bytes'] < total and rep['buffers'] == len(plan)


def test_metric_and_data_helpers():
    from hetu_61a7_b200.utils import metrics as M, data as D
    y = np.eye(3)[[0, 1, 2, 1]]
    p = np.array([[9.1, 5, 7.8], [.3, .7, 1.4], [4.3, 1.3, 5.3], [0, 3, 1]])
    cm = M.confusion_matrix_one_hot(y, p)
    assert cm['tp'].tolist() == [1, 1, 1] and cm['fp'].tolist() == [0, 0, 1] and cm['fn'].tolist() == [0, 1, 0]
    assert abs(M.precision_score_one_hot(y, p, 'macro') - (1 + 1 + 0.5) / 3) < 1e-9
    assert abs(M.recall_score_one_hot(y, p, 'micro') - 0.75) < 1e-9
    assert abs(M.f_score_one_hot(y, p, average='micro') - 0.75) < 1e-9
    v = M.confusion_matrix_at_thresholds([1, 0, 1, 0], [.9, .4, .6, .1], [0.0, 0.5, 1.0])
    assert v['tp'].tolist() == [2, 2, 0] and v['fp'].tolist() == [2, 0, 0]
    fpr, tpr = M.roc_pr_curve(v)
    assert abs(fpr[0] - 1) < 1e-5 and abs(tpr[1] - 1) < 1e-5 and tpr[2] < 1e-5
    assert D.convert_to_one_hot([0, 2, 1]).tolist() == [[1, 0, 0], [0, 0, 1], [0, 1, 0]]
    aug = D.data_augmentation(np.random.rand(4, 32, 32, 3), flip=True, crop=True, whiten=True, noise=True)
    assert aug.shape == (4, 24, 24, 3) and abs(float(aug.mean())) < 0.1
    from hetu_61a7_b200.graph.executor import topo_sort_dfs
    x = ht.Variable('x')
    order = []
    topo_sort_dfs(ht.relu_op(x), set(), order)
    assert order[0] is x and len(order) == 2


def test_step_monitor_jsonl_and_prometheus(tmp_path):
    """utils/monitor.py: step metrics as JSON lines and Prometheus gauges (no counterpart in the reference)."""
    import json
    import socket
    import urllib.request
    from hetu_61a7_b200.utils.monitor import StepMonitor
    from conftest import free_port_block
    port = free_port_block(2)
    path = str(tmp_path / 'metrics.jsonl')
    mon = StepMonitor(samples_per_step=128, log_path=path, prometheus_port=port, rank=0)
    for i in range(3):
        mon.observe(0.01 * (i + 1), loss=1.0 / (i + 1), lr=0.1)
    mon.close()
    rows = [json.loads(l) for l in open(path)]
    assert [r['step'] for r in rows] == [1, 2, 3] and rows[-1]['lr'] == 0.1
    assert abs(rows[-1]['samples_per_second'] - 128 / 0.02) < 1e-6
    txt = urllib.request.urlopen('http://127.0.0.1:%d/metrics' % port).read().decode()
    assert 'hetu_loss{rank="0"} 0.333' in txt and 'hetu_step{rank="0"} 3.0' in txt


def test_unequal_dp_schedule_matches_the_reference_rule():
    """Micro-batch turns of the replicas of a stage that is wider than the narrowest stage
    (reference pipeline_subexecutor.py:83-106)."""
    from hetu_61a7_b200.parallel.pipeline import schedule_for_different_dp as sched
    assert sched(0, 2, 2) == [True] and sched(1, 4, 4) == [True]
    assert sched(0, 2, 1) == [True, False] and sched(1, 2, 1) == [False, True]
    # 4 replicas fed by 2: two of them work per tick, every replica works every second tick
    s4 = [sched(r, 4, 2) for r in range(4)]
    assert s4 == [[True, False], [True, False], [False, True], [False, True]]
    # 3 replicas fed by 2: cycle of 3 ticks, two active replicas per tick, every replica active twice per cycle
    s3 = [sched(r, 3, 2) for r in range(3)]
    assert all(len(s) == 3 for s in s3)
    for t in range(3):
        assert sum(s[t] for s in s3) == 2
    assert all(sum(s) == 2 for s in s3)
