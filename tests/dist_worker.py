"""Multi-process worker used by test_distributed.py (launched with torch.distributed.run).
mode 'cpu'  : gloo, host-side logic (comm ops, DP all-reduce equivalence, PS-free paths)
mode 'gpu'  : nccl + symmetric heap: in-kernel collectives vs NCCL, fused DP update, sharded
              embedding, fused MoE dispatch/combine."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import hetu as ht                                                     # noqa: E402
from hetu_61a7_b200.parallel.comm import world                       # noqa: E402

mode = sys.argv[1]
if mode == 'cpu':
    os.environ['HETU_COMM_BACKEND'] = 'gloo'
comm = ht.wrapped_mpi_nccl_init()
rank, nrank = comm.rank, comm.nrank
gpu = mode == 'gpu'
dev = torch.device('cuda', torch.cuda.current_device()) if gpu else torch.device('cpu')
ctx = ht.gpu(dev.index) if gpu else ht.cpu(0)


def log(*a):
    if rank == 0:
        print(*a, flush=True)


def check(name, a, b, rtol=1e-5, atol=1e-6):
    torch.testing.assert_close(a, b, rtol=rtol, atol=atol, msg=lambda m: '%s: %s' % (name, m))
    log('ok', name)


# ---------------------------------------------------------------- raw collectives
g = torch.Generator().manual_seed(rank)
x = torch.randn(nrank * 64, 32, generator=g).to(dev)
ref = x.clone(); comm.allreduce(ref)
if gpu:
    from hetu_61a7_b200.parallel.symm import SymmEngine
    eng = SymmEngine.get(comm)
    assert eng is not None, 'symmetric heap must come up on a multi-GPU box'
    for it in range(3):
        check('symm allreduce %d' % it, eng.allreduce(x, key='t_ar'), ref, rtol=1e-5, atol=1e-5)
    check('symm allgather', eng.allgather(x, key='t_ag').clone(), comm.allgather(x))
    check('symm reducescatter', eng.reducescatter(x, key='t_rs'), comm.reducescatter(x), rtol=1e-5, atol=1e-5)
    check('symm alltoall', eng.alltoall(x, key='t_a2a').clone(), comm.alltoall(x))
    big = torch.randn(nrank * 4096, 512, generator=g).to(dev)
    check('symm alltoall big', eng.alltoall(big, key='t_a2ab').clone(), comm.alltoall(big))
    refb = big.clone(); comm.allreduce(refb)
    check('symm allreduce big', eng.allreduce(big, key='t_arb'), refb, rtol=1e-4, atol=1e-4)
    # NVLS: in-switch reduction / multicast (falls back to None where the fabric has no multicast)
    if eng.mc_buffer('probe', 4096) is not None:
        for it in range(2):
            check('nvls allreduce %d' % it, eng._nvls(big, 't_nvls_ar', 0), refb, rtol=1e-4, atol=1e-4)
        bsrc = big * (rank + 1)
        want = bsrc.clone(); comm.broadcast(want, 1 % nrank)
        check('nvls broadcast', eng.broadcast(bsrc, 1 % nrank, key='t_nvls_b'), want)
        red = eng.reduce(big, 0, key='t_nvls_r')
        if rank == 0:
            check('nvls reduce', red, refb, rtol=1e-4, atol=1e-4)
        os.environ['HETU_NVLS_ALLREDUCE'] = '0'
        check('peer-load allreduce (NVLS off)', eng.allreduce(big, key='t_arb2'), refb, rtol=1e-4, atol=1e-4)
        os.environ['HETU_NVLS_ALLREDUCE'] = '1'
        log('ok NVLS collectives')
    else:
        log('NVLS multicast not available on this fabric: skipped')


# ---------------------------------------------- data-parallel equivalence (graph level)
def build_mlp(seed_data):
    rng = np.random.RandomState(0)
    W1 = ht.Variable('W1', value=(rng.randn(20, 32) * 0.1).astype(np.float32))
    b1 = ht.Variable('b1', value=np.zeros(32, np.float32))
    W2 = ht.Variable('W2', value=(rng.randn(32, 4) * 0.1).astype(np.float32))
    x, y_ = ht.Variable('x'), ht.Variable('y_')
    h = ht.relu_op(ht.linear_op(x, W1, b1))
    loss = ht.reduce_mean_op(ht.softmaxcrossentropy_op(ht.matmul_op(h, W2), y_), [0])
    return x, y_, loss, [W1, b1, W2]


data_rng = np.random.RandomState(42)
X = data_rng.randn(nrank * 16, 20).astype(np.float32)
Y = np.eye(4, dtype=np.float32)[data_rng.randint(0, 4, nrank * 16)]
for opt_name, mk in (('sgd', lambda: ht.optim.SGDOptimizer(0.1 / nrank)), ('adam', lambda: ht.optim.AdamOptimizer(0.01))):
    for engine in (['nccl', 'symm', 'symm2shot'] if gpu else ['nccl']):
        if gpu:          # 'symm2shot': force the large-model path (reduce-scatter + all-gather, then the optimizer)
            from hetu_61a7_b200.parallel.symm import SymmEngine
            SymmEngine.MAX_FUSED_BYTES = 0 if engine == 'symm2shot' else (64 << 20)
        x, y_, loss, params = build_mlp(0)
        # sum-all-reduce of per-shard mean losses == nrank * full-batch mean gradient
        train = mk().minimize(loss)
        ex = ht.Executor({'train': [loss, train]}, ctx=ctx, comm_mode='AllReduce', seed=1,
                         use_nccl_collectives=(engine == 'nccl'))
        sl = slice(rank * 16, (rank + 1) * 16)
        for _ in range(5):
            ex.run('train', feed_dict={x: X[sl], y_: Y[sl]})
        got = [ex.config.placeholder_to_arr_map[p].detach().cpu() for p in params]
        # oracle: single process, full batch, gradient scaled by nrank (sum of shard means)
        tW1 = torch.tensor((np.random.RandomState(0).randn(20, 32) * 0.1).astype(np.float32), requires_grad=True)
        rs = np.random.RandomState(0); w1 = (rs.randn(20, 32) * 0.1).astype(np.float32); w2 = (rs.randn(32, 4) * 0.1).astype(np.float32)
        tW1 = torch.tensor(w1, requires_grad=True); tb1 = torch.zeros(32, requires_grad=True); tW2 = torch.tensor(w2, requires_grad=True)
        topt = torch.optim.SGD([tW1, tb1, tW2], 0.1 / nrank) if opt_name == 'sgd' else torch.optim.Adam([tW1, tb1, tW2], 0.01, eps=1e-7)
        for _ in range(5):
            topt.zero_grad()
            l = sum(torch.nn.functional.cross_entropy(torch.relu(torch.tensor(X[r * 16:(r + 1) * 16]) @ tW1 + tb1) @ tW2,
                                                      torch.tensor(Y[r * 16:(r + 1) * 16]).argmax(1)) for r in range(nrank))
            l.backward()
            topt.step()
        for a, b, n in zip(got, [tW1, tb1, tW2], ['W1', 'b1', 'W2']):
            check('dp %s %s %s' % (opt_name, engine, n), a, b.detach(), rtol=2e-3, atol=2e-5)

# ---------------------------------------- 1.5-D distributed GCN layer: forward and backward against the dense product
if True:
    rs = np.random.RandomState(11)
    Nn, Fi, Fo = 6 * nrank + 3, 7, 5                          # the last block-row is shorter than the others
    Ad = rs.rand(Nn, Nn).astype(np.float32) * (rs.rand(Nn, Nn) < 0.4)
    Ad = ((Ad + Ad.T) / 2).astype(np.float32)                 # symmetric, as a normalised adjacency is
    Hd, Wd = rs.randn(Nn, Fi).astype(np.float32), rs.randn(Fi, Fo).astype(np.float32)
    per = -(-Nn // nrank)
    lo, hi = rank * per, min(Nn, (rank + 1) * per)
    adj_n = ht.Variable('adj_blk', trainable=False)
    h_n, w_n = ht.Variable('h_blk', value=Hd[lo:hi].copy(), ctx=ctx), ht.Variable('w_gcn', value=Wd.copy(), ctx=ctx)
    z_n = ht.distgcn_15d_op(adj_n, h_n, w_n, hi - lo, Nn, nrank, 1, rank, comm, [None, None], True)
    loss_n = ht.reduce_sum_op(ht.mul_op(z_n, z_n), [0, 1])
    gh_n, gw_n = ht.gradients(loss_n, [h_n, w_n])
    exg = ht.Executor([z_n, gh_n, gw_n], ctx=ctx)
    zz, ghh, gww = [r.asnumpy() for r in exg.run(feed_dict={adj_n: torch.tensor(Ad[lo:hi]).to(dev)})]
    tH, tW = torch.tensor(Hd, requires_grad=True), torch.tensor(Wd, requires_grad=True)
    tz = torch.tensor(Ad) @ tH @ tW
    (tz * tz).sum().backward()
    check('distgcn forward', torch.tensor(zz), tz.detach()[lo:hi], rtol=1e-4, atol=1e-4)
    check('distgcn grad H', torch.tensor(ghh), tH.grad[lo:hi], rtol=1e-3, atol=1e-3)
    check('distgcn grad W (all-reduced)', torch.tensor(gww), tW.grad, rtol=1e-3, atol=1e-3)

# -------------------------------------------------------- sharded embedding (hybrid)
if gpu:
    V, D = 1000, 16
    table0 = np.random.RandomState(3).randn(V, D).astype(np.float32)
    ids_all = np.random.RandomState(4).randint(0, V, (nrank, 8, 4))
    emb = ht.Variable('emb', value=table0.copy())
    idn = ht.Variable('ids')
    w = ht.Variable('w_emb', value=np.ones((4 * D, 1), np.float32))
    e = ht.array_reshape_op(ht.embedding_lookup_op(emb, idn), (-1, 4 * D))
    loss = ht.reduce_mean_op(ht.reduce_sum_op(ht.mul_op(ht.matmul_op(e, w), ht.matmul_op(e, w)), [1]), [0])
    train = ht.optim.SGDOptimizer(0.01).minimize(loss, var_list=[emb])
    ex = ht.Executor({'train': [loss, train]}, ctx=ctx, comm_mode='Hybrid', seed=1, bsp=0)
    tab = ex.config.placeholder_to_arr_map[emb]
    assert getattr(tab, 'sharded', False), 'embedding must be sharded over the symmetric heap'
    check('sharded lookup', tab.lookup(torch.arange(V, device=dev)).cpu(), torch.tensor(table0))
    for step in range(3):
        ex.run('train', feed_dict={idn: ids_all[rank].astype(np.float32)})
    te = torch.tensor(table0.copy(), requires_grad=True)
    for step in range(3):
        te.grad = None
        tot = 0
        for r in range(nrank):
            ee = te[torch.tensor(ids_all[r])].reshape(-1, 4 * D)
            tot = tot + (ee.sum(1) ** 2).mean()
        tot.backward()
        with torch.no_grad():
            te -= 0.01 * te.grad
    check('sharded sgd (bsp)', tab.export_dense(), te.detach(), rtol=1e-3, atol=1e-4)

# ------------------------------ sharded embedding, optimizers with state applied at the owner (Momentum, Adam)
if gpu:
    for opt_name in ('momentum', 'adam'):
        V, D = 600, 16
        table0 = np.random.RandomState(13).randn(V, D).astype(np.float32)
        ids_all = np.random.RandomState(14).randint(0, V, (3, nrank, 8, 4))
        ids_all[:, :, 0, 0] = 7                                   # a row every rank touches every step, twice on rank 0
        ids_all[:, 0, 1, 1] = 7
        emb = ht.Variable('emb_' + opt_name, value=table0.copy())
        idn = ht.Variable('ids_' + opt_name)
        w = ht.Variable('w_emb_' + opt_name, value=np.ones((4 * D, 1), np.float32))
        e = ht.array_reshape_op(ht.embedding_lookup_op(emb, idn), (-1, 4 * D))
        loss = ht.reduce_mean_op(ht.reduce_sum_op(ht.mul_op(ht.matmul_op(e, w), ht.matmul_op(e, w)), [1]), [0])
        opt = ht.optim.MomentumOptimizer(0.01, 0.9) if opt_name == 'momentum' else ht.optim.AdamOptimizer(0.01)
        train = opt.minimize(loss, var_list=[emb])
        ex = ht.Executor({'train': [loss, train]}, ctx=ctx, comm_mode='Hybrid', seed=1, bsp=0)
        tab = ex.config.placeholder_to_arr_map[emb]
        assert getattr(tab, 'sharded', False)
        for step in range(3):
            ex.run('train', feed_dict={idn: ids_all[step, rank].astype(np.float32)})
        # oracle: gradients of all ranks summed per row, the optimizer applied once to every touched row
        P = torch.tensor(table0.copy(), dtype=torch.float64)
        s1, s2 = torch.zeros_like(P), torch.zeros_like(P)
        for step in range(3):
            te = P.clone().float().requires_grad_(True)
            tot = 0
            for r in range(nrank):
                ee = te[torch.tensor(ids_all[step, r])].reshape(-1, 4 * D)
                tot = tot + (ee.sum(1) ** 2).mean()
            tot.backward()
            g = te.grad.double()
            rows = torch.unique(torch.tensor(ids_all[step]).reshape(-1))
            if opt_name == 'momentum':
                v = 0.9 * s1[rows] - 0.01 * g[rows]
                s1[rows] = v
                P[rows] += v
            else:
                t_ = step + 1
                m = 0.9 * s1[rows] + 0.1 * g[rows]
                v = 0.999 * s2[rows] + 0.001 * g[rows] ** 2
                s1[rows], s2[rows] = m, v
                P[rows] -= 0.01 * (m / (1 - 0.9 ** t_)) / (torch.sqrt(v / (1 - 0.999 ** t_)) + opt.hyper()[2])
        check('sharded %s at the owner' % opt_name, tab.export_dense().double(), P, rtol=2e-3, atol=2e-4)

# ------------------------------------------------------------------ MoE fused vs NCCL
# (on the gloo world the one-op MoE layer runs its stage-by-stage path with all-to-alls, against the reference op chain)
if True:
    from hetu_61a7_b200.models.moe import moe_model
    for pre in (False, True):       # pre: a trainable projection in front, so the layer must also return d(input)
        losses = {}
        for fused in ('1', '0'):
            os.environ['HETU_MOE_FUSED'] = fused
            xx, tt = ht.Variable('x'), ht.Variable('t')
            xin = xx
            if pre:
                wp = ht.init.random_normal((64, 64), stddev=0.2, name='pre_proj')
                xin = ht.array_reshape_op(ht.matmul_op(ht.array_reshape_op(xx, [-1, 64]), wp), [2, 64, 64])
            ht_loss, _ = moe_model(xin, tt, batch_size=2, num_tokens=64, model_dim=64, hidden_size=64,
                                   num_local_experts=2, world=nrank, rank=rank, top=2)
            tr = ht.optim.SGDOptimizer(0.05).minimize(ht_loss)
            ex = ht.Executor({'train': [ht_loss, tr]}, ctx=ctx, comm_mode='AllReduce', seed=5,
                             use_nccl_collectives=(fused == '0'))
            # identical initial weights for both engines (node ids, hence init seeds, differ per build)
            named = ex._named_params()
            if fused == '1':
                snapshot = {n: ex.config.placeholder_to_arr_map[p].detach().cpu().numpy().copy()
                            for n, p in named.items()}
            else:
                ex.load_dict(snapshot)
            xv = np.random.RandomState(10 + rank).randn(2, 64, 64).astype(np.float32)
            ls = []
            for _ in range(4):
                ls.append(float(ex.run('train', feed_dict={xx: xv, tt: np.zeros(2, np.float32)})[0].asnumpy().reshape(-1)[0]))
            losses[fused] = ls
        log('moe losses (pre-projection %s) fused' % pre, losses['1'], 'nccl', losses['0'])
        assert np.allclose(losses['1'], losses['0'], rtol=2e-3, atol=1e-4), losses

# ------------------------------------- Wide&Deep: fused step (in-kernel push all-reduce + SGD) vs node-by-node
if gpu:
    from hetu_61a7_b200.models import ctr
    res = {}
    for fuse in ('1', '0'):
        os.environ['HETU_FUSE_WDL'] = fuse
        B, vocab = 128, 40000
        pool = [ht.data.criteo_synthetic(B, vocab=vocab, seed=100 * rank + i) for i in range(4)]
        dense = ht.dataloader_op([ht.Dataloader(np.concatenate([p_[0] for p_ in pool]), B, 'train')])
        sparse = ht.dataloader_op([ht.Dataloader(np.concatenate([p_[1] for p_ in pool]), B, 'train', dtype=np.int64)])
        y_ = ht.dataloader_op([ht.Dataloader(np.concatenate([p_[2] for p_ in pool]), B, 'train')])
        wl, wp, _, wt = ctr.wdl_criteo(dense, sparse, y_, feature_dimension=vocab, learning_rate=0.5, ctx=ctx)
        ex = ht.Executor({'train': [wl, wp, wt]}, ctx=ctx, comm_mode='Hybrid', seed=123, bsp=0, shard_dataloaders=False)
        assert ex.fused_wdl == int(fuse), (ex.fused_wdl, fuse)
        named = ex._named_params()
        if fuse == '1':
            snap = {n: (ex.config.placeholder_to_arr_map[p_].export_dense().numpy().copy()
                        if hasattr(ex.config.placeholder_to_arr_map[p_], 'export_dense')
                        else ex.config.placeholder_to_arr_map[p_].detach().cpu().numpy().copy()) for n, p_ in named.items()}
        else:
            ex.load_dict(snap)
        ls = [float(ex.run('train')[0].asnumpy().reshape(-1)[0]) for _ in range(8)]
        torch.cuda.synchronize()
        comm.barrier()
        dense_params = {n: ex.config.placeholder_to_arr_map[p_].detach().cpu().clone() for n, p_ in named.items()
                        if not hasattr(ex.config.placeholder_to_arr_map[p_], 'export_dense')}
        res[fuse] = (ls, dense_params)
    log('wdl losses fused', res['1'][0], 'unfused', res['0'][0])
    assert np.allclose(res['1'][0], res['0'][0], rtol=3e-3), res
    for n in res['0'][1]:
        a, b = res['1'][1][n], res['0'][1][n]
        assert (a - b).abs().max().item() <= 3e-2 * b.abs().max().item() + 1e-5, n
        r0 = a.to(dev).clone()
        comm.broadcast(r0, 0)                       # every rank sums the pushed blocks in rank order: replicas stay bit-identical
        assert torch.equal(r0.cpu(), a), 'replicas of %s diverged' % n
    log('ok wdl fused step == node-by-node step on %d GPUs' % nrank)
    os.environ['HETU_FUSE_WDL'] = '1'

comm.barrier()
log('ALL DIST CHECKS PASSED (mode %s, %d ranks)' % (mode, nrank))
