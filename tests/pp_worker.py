"""2-rank pipeline / model-parallel equivalence worker (gloo on CPU or NCCL on GPUs).
Pattern: reference examples/runner/parallel/all_mlp_tests.sh + validate_results.py - same
weights, same feeds, losses must equal the single-process run."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
mode, case = sys.argv[1], sys.argv[2]
if mode == 'cpu':
    os.environ['HETU_COMM_BACKEND'] = 'gloo'
import hetu as ht                                           # noqa: E402

comm = ht.wrapped_mpi_nccl_init()
rank = comm.rank
rs = np.random.RandomState(0)
W1, W2, W3 = (rs.randn(12, 16) * 0.3).astype(np.float32), (rs.randn(16, 16) * 0.3).astype(np.float32), (rs.randn(16, 4) * 0.3).astype(np.float32)
X = rs.randn(4, 8, 12).astype(np.float32)                    # 4 micro-batches of 8
Y = np.eye(4, dtype=np.float32)[rs.randint(0, 4, (4, 8))]
dev = (lambda i: ht.gpu(i)) if mode == 'gpu' else (lambda i: ht.gpu(i))


def oracle(steps, micro, per_micro_update):
    w = [torch.tensor(a.copy(), requires_grad=True) for a in (W1, W2, W3)]
    opt = torch.optim.SGD(w, 0.1)
    losses = []
    for _ in range(steps):
        if not per_micro_update:
            opt.zero_grad()
        for m in range(micro):
            if per_micro_update:
                opt.zero_grad()
            h = torch.relu(torch.relu(torch.tensor(X[m]) @ w[0]) @ w[1]) @ w[2]
            l = torch.nn.functional.cross_entropy(h, torch.tensor(Y[m]).argmax(1))
            l.backward()
            losses.append(float(l))
            if per_micro_update:
                opt.step()
        if not per_micro_update:
            opt.step()
    return losses, [t.detach().numpy() for t in w]


def build(split=None):
    with ht.context(dev(0)):
        x = ht.Variable('x')
        w1 = ht.Variable('w1', value=W1.copy())
        a = ht.relu_op(ht.matmul_op(x, w1))
    if split is None:
        with ht.context(dev(1)):
            w2 = ht.Variable('w2', value=W2.copy())
            a = ht.relu_op(ht.matmul_op(a, w2))
    else:
        with ht.context((dev(0), dev(1))):
            w2 = ht.Variable('w2', value=W2.copy())
            act, wt = {'left': ((2, 1), (1, 1)), 'right': ((1, 1), (1, 2)), 'middle': ((1, 2), (2, 1))}[split]
            a = ht.matmul_op(ht.dispatch(a, act), ht.dispatch(w2, wt))
            a = ht.relu_op(a)
    with ht.context(dev(1)):
        if split is not None:
            a = ht.dispatch(a, (1, 1))
        w3 = ht.Variable('w3', value=W3.copy())
        y_ = ht.Variable('y_')
        loss = ht.reduce_mean_op(ht.softmaxcrossentropy_op(ht.matmul_op(a, w3), y_), [0])
        train = ht.optim.SGDOptimizer(0.1).minimize(loss)
    return x, y_, loss, train, (w1, w2, w3)


def mine(ex, p):
    t = ex.config.placeholder_to_arr_map.get(p)
    return None if t is None else t.detach().cpu().numpy()


if case == 'sp':
    # sequence parallel + tensor parallel block: rows-split activation -> (all-gather -> GEMM) column-parallel
    # -> row-parallel (GEMM -> reduce-scatter) -> rows-split activation.  On GPUs both pairs must run as ONE
    # fused kernel each (hetu_gemm_ag / hetu_gemm_rs over the symmetric heap).
    from hetu_61a7_b200.parallel.mp_lowering import MPAllGatherMatMulOp, MPMatMulReduceScatterOp
    from hetu_61a7_b200.graph.autodiff import find_topo_sort
    Bn, Hd = (512, 256) if mode == 'gpu' else (8, 16)
    rs2 = np.random.RandomState(1)
    V1, V2, V3, V4 = [(rs2.randn(*shp) * sc).astype(np.float32) for shp, sc in
                      (((12, Hd), 0.3), ((Hd, Hd), 0.08 if mode == 'gpu' else 0.3), ((Hd, Hd), 0.08 if mode == 'gpu' else 0.3), ((Hd, 4), 0.1))]
    Xs = rs2.randn(4, Bn, 12).astype(np.float32)
    Ys = np.eye(4, dtype=np.float32)[rs2.randint(0, 4, (4, Bn))]
    with ht.context(dev(0)):
        x = ht.Variable('x')
        w1 = ht.Variable('w1', value=V1.copy())
        a = ht.relu_op(ht.matmul_op(x, w1))
    with ht.context((dev(0), dev(1))):
        w2 = ht.Variable('w2', value=V2.copy())
        w3 = ht.Variable('w3', value=V3.copy())
        a = ht.dispatch(a, (2, 1))
        h = ht.relu_op(ht.matmul_op(a, ht.dispatch(w2, (1, 2))))
        o = ht.matmul_op(h, ht.dispatch(w3, (2, 1)))
        o = ht.relu_op(ht.dispatch(o, (2, 1)))
    with ht.context(dev(1)):
        o = ht.dispatch(o, (1, 1))
        w4 = ht.Variable('w4', value=V4.copy())
        y_ = ht.Variable('y_')
        loss = ht.reduce_mean_op(ht.softmaxcrossentropy_op(ht.matmul_op(o, w4), y_), [0])
        train = ht.optim.SGDOptimizer(0.1).minimize(loss)
    topo = find_topo_sort([loss, train])
    fused = [n for n in topo if isinstance(n, (MPAllGatherMatMulOp, MPMatMulReduceScatterOp))]
    assert len([n for n in fused if isinstance(n, MPAllGatherMatMulOp)]) >= 2, [n.name for n in fused]
    assert len([n for n in fused if isinstance(n, MPMatMulReduceScatterOp)]) >= 2, [n.name for n in fused]
    ex = ht.Executor([loss, train], seed=0)
    got = []
    for step in range(8):
        out = ex.run(feed_dict={x: Xs[step % 4], y_: Ys[step % 4]})
        if out[0] is not None:
            got.append(float(out[0].asnumpy()))
    tw = [torch.tensor(a.copy(), requires_grad=True) for a in (V1, V2, V3, V4)]
    opt = torch.optim.SGD(tw, 0.1)
    ref = []
    for step in range(8):
        opt.zero_grad()
        hh = torch.relu(torch.relu(torch.relu(torch.tensor(Xs[step % 4]) @ tw[0]) @ tw[1]) @ tw[2])
        l = torch.nn.functional.cross_entropy(torch.relu(hh) @ tw[3] if False else hh @ tw[3], torch.tensor(Ys[step % 4]).argmax(1))
        l.backward()
        opt.step()
        ref.append(float(l))
    tol = 2e-2 if mode == 'gpu' else 2e-4                       # tf32 tensor cores on the GPU path
    if got:
        assert np.allclose(got, ref, rtol=tol, atol=1e-4), (got, ref)
    for p, r in zip((w1, w2, w3, w4), tw):
        m = mine(ex, p)
        if m is not None:
            assert np.allclose(m, r.detach().numpy(), rtol=5 * tol, atol=5e-3 if mode == 'gpu' else 1e-5), p.name
    if mode == 'gpu':
        launched = sum(n.fused_launches for n in fused)
        assert launched > 0, 'fused TP kernels did not run'
        print('fused TP kernel launches on rank %d: %d' % (rank, launched), flush=True)
    comm.barrier()
    print('PP %s rank %d OK' % (case, rank), flush=True)
    sys.exit(0)
if case == 'unequal':
    # stage 0 on one rank, stage 1 replicated on two: the replicas of the wide stage take the micro-batches in turns
    # (schedule_for_different_dp), the narrow stage hands them out round-robin; result = plain GPipe
    with ht.context(dev(0)):
        x = ht.Variable('x')
        w1 = ht.Variable('w1', value=W1.copy())
        a = ht.relu_op(ht.matmul_op(x, w1))
    with ht.context([dev(1), dev(2)]):
        w2 = ht.Variable('w2', value=W2.copy())
        a = ht.relu_op(ht.matmul_op(a, w2))
        w3 = ht.Variable('w3', value=W3.copy())
        y_ = ht.Variable('y_')
        loss = ht.reduce_mean_op(ht.softmaxcrossentropy_op(ht.matmul_op(a, w3), y_), [0])
        train = ht.optim.SGDOptimizer(0.1).minimize(loss)
    ws = (w1, w2, w3)
    ex = ht.Executor({'train': [loss, train]}, pipeline='gpipe', seed=0)
    assert ex.config.min_dp_nrank == 1 and ex.config.dp_nrank == (1 if rank == 0 else 2), (ex.config.min_dp_nrank, ex.config.dp_nrank)
    got = []
    for step in range(3):
        out = ex.run('train', feed_dict={x: [X[m] for m in range(4)], y_: [Y[m] for m in range(4)]}, batch_num=4)
        if out[0] is not None:
            got += [float(o.asnumpy()) for o in out[0]]
    ref_l, ref_w = oracle(3, 4, per_micro_update=False)
    assert (mine(ex, ws[0]) is not None) == (rank == 0)
    assert (mine(ex, ws[2]) is not None) == (rank in (1, 2))
    if rank in (1, 2):
        mine_l = [l for s_ in range(3) for m, l in enumerate(ref_l[4 * s_:4 * s_ + 4]) if m % 2 == rank - 1]
        assert len(got) == 6 and np.allclose(got, mine_l, rtol=2e-4, atol=1e-5), (got, mine_l)
    for p_, r_ in zip(ws, ref_w):
        m_ = mine(ex, p_)
        if m_ is not None:
            assert np.allclose(m_, r_, rtol=2e-3, atol=1e-5), (p_.name, np.abs(m_ - r_).max())
    comm.barrier()
    print('PP %s rank %d OK' % (case, rank), flush=True)
    sys.exit(0)
if case == 'gpipe':
    x, y_, loss, train, ws = build()
    ex = ht.Executor({'train': [loss, train]}, pipeline='gpipe', seed=0)
    got = []
    for step in range(3):
        out = ex.run('train', feed_dict={x: [X[m] for m in range(4)], y_: [Y[m] for m in range(4)]}, batch_num=4)
        if out[0] is not None:
            got += [float(o.asnumpy()) for o in out[0]]
    # GPipe sums the micro-batch gradients before one update
    ref_l, ref_w = oracle(3, 4, per_micro_update=False)
elif case == 'pipedream':
    x, y_, loss, train, ws = build()
    ex = ht.Executor({'train': [loss, train]}, pipeline='pipedream', seed=0)
    got = []
    for step in range(3):
        out = ex.run('train', feed_dict={x: [X[m] for m in range(4)], y_: [Y[m] for m in range(4)]}, batch_num=4)
        if out[0] is not None:
            got += [float(o.asnumpy()) for o in out[0]]
    ref_l, ref_w = None, None
else:
    x, y_, loss, train, ws = build(split=case)
    ex = ht.Executor({'train': [loss, train]}, seed=0)
    got = []
    for step in range(12):
        out = ex.run('train', feed_dict={x: X[step % 4], y_: Y[step % 4]})
        if out[0] is not None:
            got.append(float(out[0].asnumpy()))
    ref_l, ref_w = oracle(3, 4, per_micro_update=True)

# the graph must really be cut: stage-0 weights live on rank 0 only, stage-2 weights on rank 1 only
assert (mine(ex, ws[0]) is not None) == (rank == 0), 'placement did not prune w1 on rank %d' % rank
assert (mine(ex, ws[2]) is not None) == (rank == 1), 'placement did not prune w3 on rank %d' % rank
if ref_l is not None:
    if got:
        assert np.allclose(got, ref_l, rtol=2e-4, atol=1e-5), (got[:4], ref_l[:4])
    for p, r in zip(ws, ref_w):
        m = mine(ex, p)
        if m is not None:
            assert np.allclose(m, r, rtol=2e-3, atol=1e-5), (p.name, np.abs(m - r).max())
else:
    if got:
        assert got[-1] < got[0], got
# checkpoint round trip: every rank holds other parameters (pipeline stages) or other slices of the same parameter
# (tensor parallel): save must not let the ranks overwrite each other, load must give every rank its own values back
ck = os.environ.get('HETU_TEST_CKPT_DIR')
if ck and case in ('gpipe', 'left'):
    before = {p.name: mine(ex, p).copy() for p in ws if mine(ex, p) is not None}
    ex.save(ck, 'pp.bin', save_optimizer=True)
    comm.barrier()
    assert os.path.exists(os.path.join(ck, 'pp.bin.rank%d' % rank)), os.listdir(ck)
    for p in ws:
        t = ex.config.placeholder_to_arr_map.get(p)
        if t is not None:
            t.zero_()
    ex.load(ck, 'pp.bin')
    for p in ws:
        if p.name in before:
            assert np.array_equal(mine(ex, p), before[p.name]), 'checkpoint did not restore %s on rank %d' % (p.name, rank)
    print('PP %s rank %d checkpoint OK' % (case, rank), flush=True)
comm.barrier()
print('PP %s rank %d OK' % (case, rank), flush=True)
