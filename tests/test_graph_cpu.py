"""Graph core on CPU: ops + symbolic autodiff against torch autograd, optimizers against
closed forms, schedulers, dataloader, executor save/load (pattern: reference
tests/test_transformer_ops.py - second framework as oracle; test_optimizer.py)."""
import os
import tempfile

import numpy as np
import pytest
import torch

import hetu as ht


def run_graph(build, feeds, wrt_names, seed=0):
    """build(vars) -> scalar loss node; returns (loss, grads) evaluated on CPU."""
    vars_ = {k: ht.Variable(name=k, value=v.copy()) for k, v in feeds.items()}
    loss = build(vars_)
    grads = ht.gradients(loss, [vars_[k] for k in wrt_names])
    # a zero-lr optimizer makes this a *training* sub-executor (BN / dropout in train mode)
    train = ht.optim.SGDOptimizer(0.0).minimize(loss)
    ex = ht.Executor([loss] + grads + [train], ctx=ht.cpu(0), seed=seed)
    out = ex.run(convert_to_numpy_ret_vals=True)
    return out[0], out[1:-1]


def check(build_ht, build_torch, feeds, wrt=None, rtol=1e-4, atol=1e-5):
    wrt = wrt or list(feeds.keys())
    loss, grads = run_graph(build_ht, feeds, wrt)
    tvars = {k: torch.tensor(v, requires_grad=(k in wrt)) for k, v in feeds.items()}
    tl = build_torch(tvars)
    tl.backward()
    np.testing.assert_allclose(np.asarray(loss).reshape(-1), tl.detach().numpy().reshape(-1), rtol=rtol, atol=atol)
    for k, g in zip(wrt, grads):
        np.testing.assert_allclose(g, tvars[k].grad.numpy(), rtol=rtol, atol=atol, err_msg=k)


R = np.random.RandomState(0)


def f32(*s):
    return R.randn(*s).astype(np.float32)


def test_mlp_softmax_ce_grads():
    feeds = {'x': f32(8, 12), 'w1': f32(12, 16), 'b1': f32(16), 'w2': f32(16, 5)}
    y = np.eye(5, dtype=np.float32)[R.randint(0, 5, 8)]

    def b_ht(v):
        h = ht.relu_op(ht.linear_op(v['x'], v['w1'], v['b1']))
        logits = ht.matmul_op(h, v['w2'])
        return ht.reduce_mean_op(ht.softmaxcrossentropy_op(logits, ht.Variable('y', value=y)), [0])

    def b_t(v):
        h = torch.relu(v['x'] @ v['w1'] + v['b1'])
        return torch.nn.functional.cross_entropy(h @ v['w2'], torch.tensor(y).argmax(1))

    check(b_ht, b_t, feeds)


def test_elementwise_grads():
    feeds = {'a': f32(4, 6), 'b': np.abs(f32(4, 6)) + 0.5}

    def b_ht(v):
        a, b = v['a'], v['b']
        t = ht.tanh_op(a) * ht.sigmoid_op(b) + ht.exp_op(a * 0.1) + ht.log_op(b) + ht.sqrt_op(b)
        t = t + ht.gelu_op(a) + ht.leaky_relu_op(a, 0.1) + ht.div_op(a, b) + ht.pow_op(b, 1.5)
        t = t + ht.sin_op(a) + ht.cos_op(a) + ht.abs_op(a) + ht.rsqrt_op(b) + ht.div_const_op(2.0, b)
        t = t + ht.minus_op(a, b) + ht.minus_byconst_op(a, 3.0) + ht.opposite_op(a) + ht.const_pow_op(a, 2.0)
        return ht.reduce_sum_op(ht.reduce_sum_op(t, [1]), [0])

    def b_t(v):
        a, b = v['a'], v['b']
        t = torch.tanh(a) * torch.sigmoid(b) + torch.exp(a * 0.1) + torch.log(b) + torch.sqrt(b)
        t = t + torch.nn.functional.gelu(a) + torch.nn.functional.leaky_relu(a, 0.1) + a / b + b ** 1.5
        t = t + torch.sin(a) + torch.cos(a) + a.abs() + torch.rsqrt(b) + 2.0 / b
        t = t + (a - b) + (3.0 - a) + (-a) + torch.pow(torch.tensor(2.0), a)
        return t.sum()

    check(b_ht, b_t, feeds, rtol=1e-3, atol=1e-4)


def test_broadcast_reduce_shape_grads():
    feeds = {'a': f32(3, 4, 5), 'b': f32(5), 'c': f32(3, 1, 5)}

    def b_ht(v):
        a, b, c = v['a'], v['b'], v['c']
        t = a + ht.broadcastto_op(b, a)
        t = ht.mul_op(t, ht.broadcastto_op(c, a))
        t = ht.transpose_op(t, [2, 0, 1])
        t = ht.array_reshape_op(t, [5, 12])
        s = ht.slice_op(t, [1, 2], [3, 6])
        s2 = ht.split_op(t, [1], [1], [3])
        cat = ht.concat_op(s2, ht.pad_op(s, [[0, 2], [0, 0]]), axis=1)
        cat = ht.concatenate_op([cat, cat * 2.0], axis=0)
        r = ht.reduce_mean_op(cat, [0]) * ht.reduce_sum_op(ht.repeat_op(ht.reduce_sum_op(a, [0, 1]), [2]), [0])
        return ht.reduce_sum_op(r, [0])

    def b_t(v):
        a, b, c = v['a'], v['b'], v['c']
        t = ((a + b) * c).permute(2, 0, 1).reshape(5, 12)
        s = t[1:4, 2:8]
        s2 = t[:, 4:8]
        cat = torch.cat([s2, torch.nn.functional.pad(s, (0, 0, 0, 2))], dim=1)
        cat = torch.cat([cat, cat * 2.0], dim=0)
        r = cat.mean(0) * a.sum((0, 1)).repeat(2).sum()
        return r.sum()

    check(b_ht, b_t, feeds)


def test_norm_layers_and_losses():
    feeds = {'x': f32(6, 10), 'g': f32(10), 'b': f32(10), 'w': f32(10, 3)}
    lab = R.randint(0, 3, 6).astype(np.float32)

    def b_ht(v):
        h = ht.layer_normalization_op(v['x'], v['g'], v['b'], eps=1e-5)
        logits = ht.matmul_op(h, v['w'])
        l1 = ht.reduce_mean_op(ht.softmaxcrossentropy_sparse_op(logits, ht.Variable('lab', value=lab)), [0])
        p = ht.softmax_op(logits)
        l2 = ht.reduce_mean_op(ht.crossentropy_sparse_op(p, ht.Variable('lab2', value=lab)), [0])
        return l1 + l2

    def b_t(v):
        h = torch.nn.functional.layer_norm(v['x'], (10,), v['g'], v['b'], 1e-5)
        logits = h @ v['w']
        l = torch.nn.functional.cross_entropy(logits, torch.tensor(lab).long())
        return l + l

    check(b_ht, b_t, feeds, rtol=1e-3, atol=1e-5)


def test_conv_bn_pool_grads():
    feeds = {'x': f32(2, 3, 8, 8), 'w': f32(4, 3, 3, 3) * 0.3, 'bias': f32(4), 'g': np.abs(f32(4)) + 0.5, 'b': f32(4)}

    def b_ht(v):
        y = ht.conv2d_add_bias_op(v['x'], v['w'], v['bias'], padding=1, stride=1)
        y = ht.batch_normalization_op(y, v['g'], v['b'], eps=1e-5)
        y = ht.relu_op(y)
        y = ht.max_pool2d_op(y, 2, 2, 0, 2) + ht.avg_pool2d_op(y, 2, 2, 0, 2)
        y = ht.conv2d_op(y, ht.Variable('w2', value=np.ones((2, 4, 1, 1), np.float32)), padding=0, stride=2)
        return ht.reduce_sum_op(ht.array_reshape_op(y, [-1]), [0])

    def b_t(v):
        F = torch.nn.functional
        y = F.conv2d(v['x'], v['w'], v['bias'], padding=1)
        y = F.batch_norm(y, None, None, v['g'], v['b'], True, 0.1, 1e-5)
        y = torch.relu(y)
        y = F.max_pool2d(y, 2, 2) + F.avg_pool2d(y, 2, 2)
        y = F.conv2d(y, torch.ones(2, 4, 1, 1), stride=2)
        return y.sum()

    check(b_ht, b_t, feeds, rtol=2e-3, atol=1e-4)


def test_embedding_sparse_grad_and_optimizers():
    table = f32(20, 4)
    ids = R.randint(0, 20, (6, 3)).astype(np.float32)
    for make, torch_make in [
        (lambda: ht.optim.SGDOptimizer(0.1), lambda p: torch.optim.SGD(p, 0.1)),
        (lambda: ht.optim.AdaGradOptimizer(0.1, eps=1e-10), lambda p: torch.optim.Adagrad(p, 0.1, eps=1e-10)),
        (lambda: ht.optim.AdamOptimizer(0.01, epsilon=1e-8), lambda p: torch.optim.Adam(p, 0.01, eps=1e-8)),
    ]:
        emb = ht.Variable('emb', value=table.copy())
        w = ht.Variable('w', value=np.ones((12, 1), np.float32))
        idn = ht.Variable('ids', value=ids, trainable=False)
        e = ht.array_reshape_op(ht.embedding_lookup_op(emb, idn), (-1, 12))
        loss = ht.reduce_mean_op(ht.reduce_sum_op(ht.mul_op(ht.matmul_op(e, w), ht.matmul_op(e, w)), [1]), [0])
        train = make().minimize(loss)
        ex = ht.Executor([loss, train], ctx=ht.cpu(0))
        for _ in range(3):
            ex.run()
        got = ex.config.placeholder_to_arr_map[emb].numpy()
        te = torch.tensor(table.copy(), requires_grad=True)
        tw = torch.ones(12, 1, requires_grad=True)
        opt = torch_make([te, tw])
        for _ in range(3):
            opt.zero_grad()
            ee = te[torch.tensor(ids).long()].reshape(-1, 12)
            l = ((ee @ tw) ** 2).sum(1).mean()
            l.backward()
            opt.step()
        # AdaGrad/Adam sparse semantics: only touched rows move (matches dense for SGD; for the
        # adaptive ones untouched rows have zero grad so the dense torch step leaves them too,
        # except Adam's moment decay which has no effect on zero-initialised moments)
        np.testing.assert_allclose(got, te.detach().numpy(), rtol=2e-3, atol=2e-4)


def test_dense_optimizers_closed_form():
    p0, g = f32(5, 3), f32(5, 3)
    cases = {
        'momentum': (ht.optim.MomentumOptimizer(0.1, 0.9), lambda p, v: (p + (0.9 * v - 0.1 * g), 0.9 * v - 0.1 * g)),
        'nesterov': (ht.optim.MomentumOptimizer(0.1, 0.9, nesterov=True),
                     lambda p, v: (p + 0.9 * (v - 0.1 * g) - 0.1 * g, 0.9 * (v - 0.1 * g))),
    }
    for name, (opt, step) in cases.items():
        w = ht.Variable('w_' + name, value=p0.copy())
        loss = ht.reduce_sum_op(ht.reduce_sum_op(ht.mul_op(w, ht.Variable('g', value=g, trainable=False)), [1]), [0])
        ex = ht.Executor([loss, opt.minimize(loss)], ctx=ht.cpu(0))
        p, v = p0.copy(), np.zeros_like(p0)
        for _ in range(3):
            ex.run()
            p, v = step(p, v)
        np.testing.assert_allclose(ex.config.placeholder_to_arr_map[w].numpy(), p, rtol=1e-5, atol=1e-6)
    # AdamW / Lamb one step vs formula (reference tests/test_optimizer.py)
    for cls, lamb in ((ht.optim.AdamWOptimizer, False), (ht.optim.LambOptimizer, True)):
        w = ht.Variable('w', value=p0.copy())
        loss = ht.reduce_sum_op(ht.reduce_sum_op(ht.mul_op(w, ht.Variable('g', value=g, trainable=False)), [1]), [0])
        ex = ht.Executor([loss, cls(0.01, 0.9, 0.999, 1e-7, 0.1).minimize(loss)], ctx=ht.cpu(0))
        ex.run()
        m, v = 0.1 * g, 0.001 * g * g
        upd = (m / 0.1) / (np.sqrt(v / 0.001) + 1e-7)
        if lamb:
            ratio = np.linalg.norm(p0) / np.linalg.norm(upd)
            exp = p0 - 0.01 * ratio * (upd + 0.1 * p0)
        else:
            exp = p0 - 0.01 * (upd + 0.1 * p0)
        np.testing.assert_allclose(ex.config.placeholder_to_arr_map[w].numpy(), exp, rtol=1e-4, atol=1e-6)


def test_lr_schedulers():
    s = ht.lr.StepScheduler(1.0, 2, 0.5)
    assert [s.step() for _ in range(5)] == [1.0, 0.5, 0.5, 0.25, 0.25]
    m = ht.lr.MultiStepScheduler(1.0, [2, 4], 0.1)
    vals = [m.step() for _ in range(5)]
    assert vals[0] == 1.0 and abs(vals[1] - 0.1) < 1e-12 and abs(vals[3] - 0.01) < 1e-12
    e = ht.lr.ExponentialScheduler(1.0, 0.5, ending=0.2)
    assert [e.step() for _ in range(4)] == [0.5, 0.25, 0.2, 0.2]
    r = ht.lr.ReduceOnPlateauScheduler(1.0, patience=1, factor=0.1)
    out = [r.step(v) for v in [1.0, 1.0, 1.0, 1.0]]
    assert out[-1] < 1.0
    # scheduler drives the optimizer's device-side lr
    w = ht.Variable('w', value=np.ones((2, 2), np.float32))
    loss = ht.reduce_sum_op(ht.reduce_sum_op(w, [1]), [0])
    sched = ht.lr.StepScheduler(1.0, 1, 0.5)
    opt = ht.optim.SGDOptimizer(sched)
    ex = ht.Executor([loss, opt.minimize(loss)], ctx=ht.cpu(0))
    ex.run()
    sched.step()
    ex.run()
    np.testing.assert_allclose(ex.config.placeholder_to_arr_map[w].numpy(), np.ones((2, 2)) - 1.0 - 0.5)


def test_dataloader_and_subexecutors():
    x = np.arange(40, dtype=np.float32).reshape(20, 2)
    dl = ht.dataloader_op([ht.Dataloader(x, 4, 'train'), ht.Dataloader(x[:8], 4, 'validate')])
    w = ht.Variable('w', value=np.ones((2, 1), np.float32))
    y = ht.matmul_op(dl, w)
    loss = ht.reduce_mean_op(ht.reduce_sum_op(y, [1]), [0])
    train = ht.optim.SGDOptimizer(0.0).minimize(loss)
    ex = ht.Executor({'train': [loss, train], 'validate': [loss]}, ctx=ht.cpu(0))
    assert ex.get_batch_num('train') == 5 and ex.get_batch_num('validate') == 2
    seen = [float(ex.run('train')[0].asnumpy()) for _ in range(7)]
    expect = [x[i * 4:(i + 1) * 4].sum(1).mean() for i in range(5)]
    np.testing.assert_allclose(seen, expect + expect[:2], rtol=1e-6)
    v = [float(ex.run('validate')[0].asnumpy()) for _ in range(3)]
    np.testing.assert_allclose(v, [expect[0], expect[1], expect[0]], rtol=1e-6)


def test_save_load_roundtrip():
    w = ht.init.random_normal((4, 3), name='W_save')
    b = ht.init.zeros((3,), name='b_save')
    x = ht.Variable('x')
    loss = ht.reduce_sum_op(ht.reduce_sum_op(ht.linear_op(x, w, b), [1]), [0])
    ex = ht.Executor([loss, ht.optim.AdamOptimizer(0.1).minimize(loss)], ctx=ht.cpu(0), seed=3)
    xv = f32(5, 4)
    ex.run(feed_dict={x: xv})
    with tempfile.TemporaryDirectory() as d:
        ex.save(d, 'ckpt.bin', save_optimizer=True)
        import pickle
        state = pickle.load(open(os.path.join(d, 'ckpt.bin'), 'rb'))
        assert set(state.keys()) == {'W_save', 'b_save'} and state['W_save'].dtype == np.float32
        saved = {k: v.copy() for k, v in state.items()}
        ex.run(feed_dict={x: xv})
        after1 = ex.config.placeholder_to_arr_map[w].numpy().copy()
        ex.load(d, 'ckpt.bin')
        np.testing.assert_allclose(ex.config.placeholder_to_arr_map[w].numpy(), saved['W_save'])
        ex.run(feed_dict={x: xv})      # optimizer slots were restored too -> same step as before
        np.testing.assert_allclose(ex.config.placeholder_to_arr_map[w].numpy(), after1, rtol=1e-6)


def test_moe_layer_cpu_forward_backward():
    from hetu_61a7_b200.models.moe import moe_model
    x, t = ht.Variable('x'), ht.Variable('t')
    for gate in ('top', 'ktop1', 'sam'):
        loss, y = moe_model(x, t, batch_size=2, num_tokens=32, model_dim=16, hidden_size=24,
                            num_local_experts=4, world=1, top=2, gate=gate, num_local_gpus=2)
        train = ht.optim.SGDOptimizer(0.01).minimize(loss)
        ex = ht.Executor([loss, y, train], ctx=ht.cpu(0), seed=0)
        out = ex.run(feed_dict={x: f32(2, 32, 16), t: np.zeros(2, np.float32)})
        assert np.isfinite(out[0].asnumpy()).all() and out[1].shape == (2, 32)


def test_moe_ffn_op_matches_op_chain(monkeypatch):
    """ops/moe_ffn.py (dispatch + expert FFN + combine and their gradients in one op) trains exactly like the
    dispatch -> split -> matmul/relu/matmul -> concatenate -> combine chain it replaces."""
    from hetu_61a7_b200.models.moe import moe_model
    from hetu_61a7_b200.ops.moe_ffn import FusedMoEFFNOp
    from hetu_61a7_b200.graph.autodiff import find_topo_sort
    for gate, pre in (('top', False), ('top', True), ('hash', False), ('hash', True), ('ktop1', True), ('sam', True)):
        losses, snapshot = {}, None
        for overlap in ('1', '0'):
            monkeypatch.setenv('HETU_MOE_OVERLAP', overlap)
            x, t = ht.Variable('x'), ht.Variable('t')
            hi = ht.Variable('hash_idx') if gate == 'hash' else None
            xin = x
            if pre:         # a trainable layer in front: the MoE layer has to return the gradient of its input too
                wp = ht.init.random_normal((16, 16), stddev=0.3, name='pre_proj')
                xin = ht.array_reshape_op(ht.matmul_op(ht.array_reshape_op(x, [-1, 16]), wp), [2, 32, 16])
            loss, y = moe_model(xin, t, batch_size=2, num_tokens=32, model_dim=16, hidden_size=24, num_local_experts=4,
                                world=1, top=2, gate=gate, num_local_gpus=2, hash_indices=hi)
            assert any(isinstance(n, FusedMoEFFNOp) for n in find_topo_sort([loss])) == (overlap == '1')
            train = ht.optim.SGDOptimizer(0.05).minimize(loss)
            ex = ht.Executor([loss, train], ctx=ht.cpu(0), seed=0)
            if snapshot is None:
                snapshot = {n: ex.config.placeholder_to_arr_map[p_].detach().numpy().copy()
                            for n, p_ in ex._named_params().items()}
            else:
                ex.load_dict(snapshot)
            feed = {x: np.random.RandomState(3).randn(2, 32, 16).astype(np.float32), t: np.zeros(2, np.float32)}
            if hi is not None:
                feed[hi] = np.random.RandomState(4).randint(0, 4, 64).astype(np.float32)
            losses[overlap] = [float(ex.run(feed_dict=feed)[0].asnumpy().reshape(-1)[0]) for _ in range(4)]
        np.testing.assert_allclose(losses['1'], losses['0'], rtol=1e-5, atol=1e-6, err_msg=gate)


def test_moe_dispatch_combine_identity():
    """combine(dispatch(x)) with gate 1 and enough capacity returns x for kept tokens."""
    T, d, E = 64, 8, 4
    xv = f32(T, d)
    idx = R.randint(0, E, T).astype(np.float32)
    onehot = np.eye(E, dtype=np.float32)[idx.astype(int)]
    loc = ((np.cumsum(onehot, 0) - 1) * onehot).sum(1).astype(np.float32)
    x = ht.Variable('x', value=xv)
    i, l = ht.Variable('i', value=idx, trainable=False), ht.Variable('l', value=loc, trainable=False)
    g = ht.Variable('g', value=np.ones(T, np.float32), trainable=False)
    disp = ht.layout_transform_op(x, [i], [l], T, E)
    back = ht.reverse_layout_transform_op(disp, [i], [l], [g], T, E)
    ex = ht.Executor([back], ctx=ht.cpu(0))
    np.testing.assert_allclose(ex.run()[0].asnumpy(), xv)


def test_wdl_trains_on_cpu():
    from hetu_61a7_b200.models.ctr import wdl_criteo
    dense, sparse, y_ = ht.Variable('dense'), ht.Variable('sparse'), ht.Variable('y_')
    loss, pred, _, train = wdl_criteo(dense, sparse, y_, feature_dimension=1000, learning_rate=0.5)
    ex = ht.Executor({'train': [loss, pred, train]}, ctx=ht.cpu(0), seed=1)
    d, s, l = ht.data.criteo_synthetic(64, vocab=1000, seed=1)
    losses = [float(ex.run('train', feed_dict={dense: d, sparse: s.astype(np.float32), y_: l})[0].asnumpy())
              for _ in range(30)]
    assert losses[-1] < losses[0]


def test_context_and_device_group():
    g = ht.DeviceGroup(['cpu:0', 'gpu:0', 'gpu:1'])
    assert g.worker_num == 2 and g.server_num == 1 and not g.is_mp
    g2 = ht.DeviceGroup([(ht.gpu(0), ht.gpu(1)), (ht.gpu(2), ht.gpu(3))])
    assert g2.worker_num == 2 and g2.is_mp and g2.mp_dev_num == 2
    with ht.context([ht.gpu(0), ht.gpu(1)]):
        n = ht.Variable('ctxvar', value=np.zeros(1, np.float32))
    assert n.raw_ctx.worker_num == 2
    r = ht.DeviceGroup('node1:gpu:3')
    assert not r.workers[0].local and r.workers[0].device_id == 3


def test_metrics_tokenizer():
    y = (R.rand(500) > 0.5).astype(np.float32)
    p = np.clip(y * 0.6 + R.rand(500) * 0.5, 0, 1)
    a1, a2 = ht.metrics.auc(p, y), ht.metrics.roc_auc_exact(p, y)
    assert abs(a1 - a2) < 0.02 and a2 > 0.8
    vocab = ['[PAD]', '[UNK]', '[CLS]', '[SEP]', 'hello', 'world', '##s', ',']
    tok = ht.BertTokenizer(vocab=vocab)
    assert tok.tokenize('Hello, worlds foo') == ['hello', ',', 'world', '##s', '[UNK]']
    ids, types, mask = tok.encode('hello world', 'hello', max_seq_length=8)
    assert ids[:6] == [2, 4, 5, 3, 4, 3] and sum(mask) == 6 and types[4] == 1


def test_fused_attention_matches_autograd():
    rs = np.random.RandomState(0)
    B, H, S, D = 2, 3, 16, 8
    qv, kv, vv = [rs.randn(B, H, S, D).astype(np.float32) for _ in range(3)]
    mv = (rs.rand(B, 1, 1, S) > 0.8).astype(np.float32) * -1e4
    q, k, v, m = ht.Variable('q', value=qv), ht.Variable('k', value=kv), ht.Variable('v', value=vv), ht.Variable('m')
    out = ht.fused_attention_op(q, k, v, m)
    loss = ht.reduce_sum_op(ht.mul_op(out, out), [0, 1, 2, 3])
    grads = ht.gradients(loss, [q, k, v])
    ex = ht.Executor([out] + grads, ctx=ht.cpu(0))
    res = [r.asnumpy() for r in ex.run(feed_dict={m: mv})]
    tq, tk, tv = [torch.tensor(a, requires_grad=True) for a in (qv, kv, vv)]
    s = tq @ tk.transpose(-1, -2) / np.sqrt(D) + torch.tensor(mv)
    o = torch.softmax(s, -1) @ tv
    (o * o).sum().backward()
    for got, ref in zip(res, [o.detach(), tq.grad, tk.grad, tv.grad]):
        np.testing.assert_allclose(got, ref.numpy(), rtol=1e-4, atol=1e-5)


def test_matmul_relu_epilogue_fusion_matches_oracle():
    """graph/fusion.py moves relu into the GEMM epilogue; forward and backward must be unchanged."""
    rs = np.random.RandomState(0)
    W1, W2, W3 = [(rs.randn(*s) * 0.3).astype(np.float32) for s in ((12, 16), (16, 16), (16, 4))]
    X = rs.randn(8, 12).astype(np.float32)
    Y = np.eye(4, dtype=np.float32)[rs.randint(0, 4, 8)]
    x, y_ = ht.Variable('x'), ht.Variable('y_')
    w1, w2, w3 = ht.Variable('w1', value=W1.copy()), ht.Variable('w2', value=W2.copy()), ht.Variable('w3', value=W3.copy())
    h = ht.relu_op(ht.matmul_op(ht.relu_op(ht.matmul_op(x, w1)), w2))
    loss = ht.reduce_mean_op(ht.softmaxcrossentropy_op(ht.matmul_op(h, w3), y_), [0])
    train = ht.optim.SGDOptimizer(0.1).minimize(loss)
    ex = ht.Executor([loss, train], ctx=ht.cpu(0))
    assert ex.fused_epilogues == 2
    got = [float(ex.run(feed_dict={x: X, y_: Y})[0].asnumpy()) for _ in range(5)]
    tw = [torch.tensor(a.copy(), requires_grad=True) for a in (W1, W2, W3)]
    opt = torch.optim.SGD(tw, 0.1)
    ref = []
    for _ in range(5):
        opt.zero_grad()
        l = torch.nn.functional.cross_entropy(torch.relu(torch.relu(torch.tensor(X) @ tw[0]) @ tw[1]) @ tw[2], torch.tensor(Y).argmax(1))
        l.backward()
        opt.step()
        ref.append(float(l))
    np.testing.assert_allclose(got, ref, rtol=1e-4, atol=1e-6)


def _fused_gate_vs_composed(ctx, tol):
    """ops/moe_gate.py: one op for softmax + top-k + slots + balance loss equals the ~30-op composed gate
    (values, slots and the gradient w.r.t. the logits)."""
    import os
    from hetu_61a7_b200.layers import moe as M
    rs = np.random.RandomState(0)
    res = {}
    for T, E, k in ((300, 8, 2), (1000, 16, 1), (517, 12, 3)):
        lv = rs.randn(T, E).astype(np.float32)
        for flag in ('0', '1'):
            os.environ['HETU_MOE_FUSED_GATE'] = flag
            try:
                logits = ht.Variable('gate_logits_%s_%d' % (flag, T), value=lv.copy())
                l_aux, idx, loc, gates, cap = M.topkgating(logits, k, 1.0, T, E, 16)
                obj = l_aux
                for i in range(k):
                    r = ht.Variable('gate_r%d_%s_%d' % (i, flag, T), value=np.random.RandomState(1 + i).randn(T).astype(np.float32),
                                    trainable=False)
                    obj = obj + ht.reduce_sum_op(ht.mul_op(gates[i], r), axes=0)
                g = ht.gradients(obj, [logits])[0]
                ex = ht.Executor([l_aux] + idx + loc + gates + [g], ctx=ctx)
                res[flag] = ex.run(convert_to_numpy_ret_vals=True)
            finally:
                os.environ.pop('HETU_MOE_FUSED_GATE', None)
        for a, b in zip(res['0'], res['1']):
            np.testing.assert_allclose(np.ravel(a), np.ravel(b), rtol=tol, atol=tol)
        loc = np.stack(res['1'][1 + k:1 + 2 * k]).astype(np.int64)
        idx = np.stack(res['1'][1:1 + k]).astype(np.int64)
        pairs = set(zip(idx.reshape(-1).tolist(), loc.reshape(-1).tolist()))
        assert len(pairs) == T * k                              # every (expert, slot) is used once


def test_fused_moe_gate_matches_composed_gate():
    _fused_gate_vs_composed(ht.cpu(0), 1e-6)


def test_matmul_bias_fusion_keeps_training_identical(tmp_path):
    """graph/fusion.py::fuse_matmul_bias: matmul + broadcast bias (+ relu) folded into one GEMM on host executors -
    same losses step by step, same outputs, and the ONNX export of the fused graph still computes the same function."""
    import os
    from hetu.onnx import hetu2onnx, onnx2hetu
    from hetu_61a7_b200.models import cnn

    def run(flag):
        os.environ['HETU_FUSE_BIAS'] = flag
        try:
            np.random.seed(0)
            rs = np.random.RandomState(0)
            x, y_ = ht.Variable('x'), ht.Variable('y_')
            loss, y = cnn.mlp(x, y_, in_dim=48)
            train = ht.optim.SGDOptimizer(0.05).minimize(loss)
            ex = ht.Executor([loss, train], ctx=ht.cpu(0), seed=0)
            out = []
            for _ in range(6):
                xs = rs.randn(16, 48).astype(np.float32)
                ys = np.eye(10, dtype=np.float32)[rs.randint(0, 10, 16)]
                out.append(float(ex.run(feed_dict={x: xs, y_: ys}, convert_to_numpy_ret_vals=True)[0]))
            return out, ex.fused_biases, ex.fused_epilogues
        finally:
            os.environ.pop('HETU_FUSE_BIAS', None)
    plain, nb0, _ = run('0')
    fused, nb1, ne1 = run('all')
    assert nb0 == 0 and nb1 == 3 and ne1 == 2
    np.testing.assert_allclose(fused, plain, rtol=1e-6)
    os.environ['HETU_FUSE_BIAS'] = 'all'
    try:
        x, y_ = ht.Variable('x'), ht.Variable('y_')
        _, y = cnn.mlp(x, y_, in_dim=48)
        h = ht.relu_op(y)
        ex = ht.Executor([h], ctx=ht.cpu(0))
        xv = np.random.RandomState(1).randn(4, 48).astype(np.float32)
        ref = ex.run(feed_dict={x: xv}, convert_to_numpy_ret_vals=True)[0]
        assert ex.fused_biases == 3
        path = str(tmp_path / 'mlp.onnx')
        hetu2onnx.export(ex, [x], [h], path)
        ins, outs = onnx2hetu.load_onnx(path)
        got = ht.Executor(outs, ctx=ht.cpu(0)).run(feed_dict={ins[0]: xv}, convert_to_numpy_ret_vals=True)[0]
        np.testing.assert_allclose(got, ref, rtol=1e-5, atol=1e-6)
    finally:
        os.environ.pop('HETU_FUSE_BIAS', None)


def test_checkpoint_resume_is_exact_with_run_state(tmp_path):
    """save(..., save_optimizer=True) also records dataloader positions and step counters, so a fresh executor that
    loads the checkpoint continues on the same batches with the same optimizer state (the reference drops all of this)."""
    rs = np.random.RandomState(0)
    X = rs.randn(80, 6).astype(np.float32)
    Y = np.eye(3, dtype=np.float32)[rs.randint(0, 3, 80)]

    def build():
        x = ht.dataloader_op([ht.Dataloader(X, 8, 'train')])
        y_ = ht.dataloader_op([ht.Dataloader(Y, 8, 'train')])
        w = ht.Variable('w_resume', value=np.zeros((6, 3), np.float32))
        loss = ht.reduce_mean_op(ht.softmaxcrossentropy_op(ht.matmul_op(x, w), y_), [0])
        return ht.Executor({'train': [loss, ht.optim.AdamOptimizer(0.05).minimize(loss)]}, ctx=ht.cpu(0))
    ex = build()
    ref = [float(ex.run('train')[0].asnumpy()) for _ in range(13)]
    ex = build()
    first = [float(ex.run('train')[0].asnumpy()) for _ in range(7)]
    ex.save(str(tmp_path), 'ckpt.pkl', save_optimizer=True)
    ex2 = build()
    ex2.load(str(tmp_path), 'ckpt.pkl')
    rest = [float(ex2.run('train')[0].asnumpy()) for _ in range(6)]
    np.testing.assert_allclose(first + rest, ref, rtol=1e-6)


def test_fused_gate_partial_uses_match_composed_gate(monkeypatch):
    """The fused gate's backward takes the gradients of its pieces directly: a loss that leaves the balance loss out,
    uses one gate value twice, or uses the balance loss alone must still give the composed gate's gradient."""
    from hetu_61a7_b200.layers import moe as M
    T, E, k = 96, 8, 2
    lv = np.random.RandomState(0).randn(T, E).astype(np.float32)

    def grad(case, flag):
        monkeypatch.setenv('HETU_MOE_FUSED_GATE', flag)
        logits = ht.Variable('lg_%s_%s' % (case, flag), value=lv.copy())
        l_aux, idx, loc, gates, cap = M.topkgating(logits, k, 1.0, T, E, 16)
        if case == 'no_aux':
            obj = ht.reduce_sum_op(gates[0], axes=0) + ht.reduce_sum_op(ht.mul_byconst_op(gates[1], 2.0), axes=0)
        elif case == 'gate_twice':
            obj = ht.reduce_sum_op(ht.mul_op(gates[0], gates[0]), axes=0) + l_aux
        else:
            obj = l_aux
        g = ht.gradients(obj, [logits])[0]
        return ht.Executor([obj, g], ctx=ht.cpu(0)).run(convert_to_numpy_ret_vals=True)

    for case in ('no_aux', 'gate_twice', 'aux_only'):
        a, b = grad(case, '1'), grad(case, '0')
        np.testing.assert_allclose(np.ravel(a[0]), np.ravel(b[0]), rtol=1e-5, atol=1e-6, err_msg=case)
        np.testing.assert_allclose(a[1], b[1], rtol=1e-4, atol=1e-6, err_msg=case)
