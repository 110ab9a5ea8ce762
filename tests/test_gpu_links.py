"""``ht.gpu_links`` kernel API against numpy / torch-autograd oracles (pattern: reference
tests/test_gpu_op.py - call the link on ``ht.array`` inputs, compare the output array with the
numpy expression).  The same checks run on host arrays here and, marked ``gpu``, on a B200 where
the calls go through the sm_100a kernels."""
import numpy as np
import pytest
import torch

import hetu as ht
from hetu import gpu_links as gl

RS = np.random.RandomState(0)


def _run_all(ctx, tol):
    def arr(a):
        return ht.array(np.asarray(a, dtype=np.float32), ctx)

    def out(*shape):
        return ht.empty(shape, ctx=ctx)

    def close(o, ref, t=None, msg=''):
        np.testing.assert_allclose(o.asnumpy(), ref, rtol=t or tol, atol=t or tol, err_msg=msg)

    x = RS.randn(6, 10).astype(np.float32)
    y = RS.randn(6, 10).astype(np.float32)
    pos = np.abs(x) + 0.5
    # --- element-wise ---------------------------------------------------------------------
    unary = [(gl.abs_val, np.abs(x)), (gl.exp, np.exp(x)), (gl.floor, np.floor(x)), (gl.sin, np.sin(x)), (gl.cos, np.cos(x)),
             (gl.sigmoid, 1 / (1 + np.exp(-x))), (gl.tanh, np.tanh(x)), (gl.relu, np.maximum(x, 0)), (gl.matrix_opposite, -x),
             (gl.bool, (x > 0).astype(np.float32))]
    for fn, ref in unary:
        o = out(6, 10)
        fn(arr(x), o)
        close(o, ref, msg=fn.__name__)
    for fn, ref in [(gl.matrix_sqrt, np.sqrt(pos)), (gl.matrix_rsqrt, 1 / np.sqrt(pos)), (gl.log_link, np.log(pos))]:
        o = out(6, 10)
        fn(arr(pos), o)
        close(o, ref, msg=fn.__name__)
    o = out(6, 10)
    gl.gelu(arr(x), o)
    close(o, torch.nn.functional.gelu(torch.tensor(x)).numpy(), 2e-3, 'gelu')
    for fn, ref in [(gl.matrix_elementwise_add, x + y), (gl.matrix_elementwise_minus, x - y), (gl.matrix_elementwise_multiply, x * y),
                    (gl.max_mat, np.maximum(x, y)), (gl.min_mat, np.minimum(x, y))]:
        o = out(6, 10)
        fn(arr(x), arr(y), o)
        close(o, ref, msg=fn.__name__)
    o = out(6, 10)
    gl.matrix_elementwise_divide(arr(x), arr(pos), o)
    close(o, x / pos)
    o = out(6, 10)
    gl.matrix_elementwise_add_by_const(arr(x), 2.5, o)
    close(o, x + 2.5)
    gl.matrix_elementwise_multiply_by_const(arr(x), -3.0, o)
    close(o, x * -3.0)
    gl.minus_by_const(arr(x), o, 4.0)
    close(o, 4.0 - x)
    gl.matrix_elementwise_divide_const(2.0, arr(pos), o)
    close(o, 2.0 / pos)
    gl.leaky_relu(arr(x), 0.1, o)
    close(o, np.where(x > 0, x, 0.1 * x))
    gl.pow_matrix(arr(pos), o, 1.7)
    close(o, pos ** 1.7, 1e-4)
    gl.const_pow(arr(x), o, 2.0)
    close(o, 2.0 ** x, 1e-4)
    gl.clamp(arr(x), o, mmin=-0.3, mmax=0.4)
    close(o, np.clip(x, -0.3, 0.4))
    gl.bool_val(arr(x), o, 0.2, 2)
    close(o, (x > 0.2).astype(np.float32))
    gl.bool_matrix(arr(x), arr(y), o, 1)
    close(o, (x < y).astype(np.float32))
    mask = (RS.rand(6, 10) < 0.3).astype(np.float32)
    gl.masked_fill(arr(x), arr(mask), o, -9.0)
    close(o, np.where(mask > 0, -9.0, x))
    gl.where(arr(mask), arr(x), arr(y), o)
    close(o, np.where(mask > 0, x, y))
    gl.where_const(arr(mask), arr(x), 7.0, o)
    close(o, np.where(mask > 0, x, 7.0))
    gl.array_set(o, 3.0)
    close(o, np.full((6, 10), 3.0))
    o1 = out(8)
    gl.arange(2, 10, 1, o1)
    close(o1, np.arange(2, 10))
    # gradients of the element-wise ops against autograd
    g = RS.randn(6, 10).astype(np.float32)
    for fn, tfn, inp in [(gl.relu_gradient, torch.relu, x), (gl.gelu_gradient, torch.nn.functional.gelu, x),
                         (gl.abs_gradient, torch.abs, x), (lambda a, b, c: gl.leaky_relu_gradient(a, b, 0.1, c),
                                                            lambda t: torch.nn.functional.leaky_relu(t, 0.1), x),
                         (lambda a, b, c: gl.pow_gradient(a, b, c, 1.7), lambda t: t ** 1.7, pos),
                         (lambda a, b, c: gl.log_grad_link(b, a, c), torch.log, pos)]:
        t = torch.tensor(inp, requires_grad=True)
        tfn(t).backward(torch.tensor(g))
        o = out(6, 10)
        fn(arr(inp), arr(g), o)
        close(o, t.grad.numpy(), 2e-3, getattr(fn, '__name__', 'grad'))
    o = out(6, 10)
    gl.const_pow_gradient(arr(2.0 ** x), arr(g), o, 2.0)            # takes the forward OUTPUT, like the reference kernel
    close(o, (2.0 ** x) * np.log(2.0) * g, 1e-4)
    gl.tanh_gradient(arr(np.tanh(x)), arr(g), o)
    close(o, g * (1 - np.tanh(x) ** 2), 1e-4)
    # --- linear algebra -------------------------------------------------------------------
    A, B = RS.randn(16, 24).astype(np.float32), RS.randn(24, 8).astype(np.float32)
    mm_tol = 5e-2 if ctx.device_type != 'cpu' else 1e-4          # tf32 tensor cores on the GPU
    o = out(16, 8)
    gl.matrix_multiply(arr(A), False, arr(B), False, o)
    close(o, A @ B, mm_tol)
    gl.matrix_multiply(arr(A.T.copy()), True, arr(B.T.copy()), True, o)
    close(o, A @ B, mm_tol)
    bias = RS.randn(8).astype(np.float32)
    gl.matmul_with_bias(arr(A), False, arr(B), False, arr(bias), o)
    close(o, A @ B + bias, mm_tol)
    C = RS.randn(16, 8).astype(np.float32)
    gl.addmm(arr(C), arr(A), arr(B), o, 0.5, 2.0)
    close(o, 2.0 * C + 0.5 * (A @ B), mm_tol)
    bA, bB = RS.randn(3, 5, 7).astype(np.float32), RS.randn(3, 7, 4).astype(np.float32)
    ob = out(3, 5, 4)
    gl.batch_matrix_multiply(arr(bA), False, arr(bB), False, ob)
    close(ob, bA @ bB, mm_tol)
    bC = RS.randn(3, 5, 4).astype(np.float32)
    gl.baddbmm(arr(bC), arr(bA), arr(bB), ob, 0.5, 2.0)
    close(ob, 2.0 * bC + 0.5 * (bA @ bB), mm_tol)
    # --- reductions / arg ops -------------------------------------------------------------
    o = out(10)
    gl.reduce_sum(arr(x), o, [0])
    close(o, x.sum(0), 1e-4)
    gl.reduce_sum_axis_zero(arr(x), o, None)
    close(o, x.sum(0), 1e-4)
    o6 = out(6)
    gl.reduce_mean(arr(x), o6, [1])
    close(o6, x.mean(1), 1e-4)
    gl.max(arr(x), o6, 1)
    close(o6, x.max(1))
    gl.min(arr(x), o6, 1)
    close(o6, x.min(1))
    gl.argmax(arr(x), o6, 1)
    close(o6, x.argmax(1))
    oi, ov = out(6, 3), out(6, 3)
    gl.topk_idx(arr(x), oi, 3)
    close(oi, np.argsort(-x, 1)[:, :3])
    gl.topk_val(arr(x), oi, ov, 3)
    close(ov, -np.sort(-x, 1)[:, :3])
    oc = out(6, 10)
    gl.cumsum_with_bias(arr(x), oc, -1.0, 0)
    close(oc, np.cumsum(x, 0) - 1.0, 1e-4)
    so, si = out(6, 10), out(6, 10)
    gl.argsort(arr(x), so, out(6, 10), si, 1, False)
    close(so, np.sort(x, 1))
    close(si, np.argsort(x, 1, kind='stable'))
    # --- shape / indexing -----------------------------------------------------------------
    o = out(10, 6)
    gl.matrix_transpose(arr(x), o, [1, 0])
    close(o, x.T)
    o = out(4, 6, 10)
    gl.broadcast_to(arr(x), o)
    close(o, np.broadcast_to(x, (4, 6, 10)))
    o = out(12, 10)
    gl.concat(arr(x), arr(y), o, 0)
    close(o, np.concatenate([x, y], 0))
    o = out(6, 30)
    gl.concatenate([arr(x), arr(y), arr(x)], o, 1)
    close(o, np.concatenate([x, y, x], 1))
    o = out(3, 4)
    gl.matrix_slice(arr(x), o, [2, 5])
    close(o, x[2:5, 5:9])
    full = out(6, 10)
    gl.matrix_slice_gradient(o, full, [2, 5])
    ref = np.zeros_like(x)
    ref[2:5, 5:9] = x[2:5, 5:9]
    close(full, ref)
    idx = RS.randint(0, 10, (6, 4))
    o = out(6, 4)
    gl.gather(arr(x), arr(idx), o, 1)
    close(o, np.take_along_axis(x, idx, 1))
    rows = RS.randint(0, 6, 5)
    o = out(5, 10)
    gl.indexing(arr(x), arr(rows), o)
    close(o, x[rows])
    o = out(12, 20)
    gl.repeat(arr(x), o)
    close(o, np.tile(x, (2, 2)))
    o = out(6, 10)
    gl.roll(arr(x), o, 3, 1)
    close(o, np.roll(x, 3, 1))
    lab = RS.randint(0, 7, 9)
    o = out(9, 7)
    gl.one_hot(arr(lab), o)
    close(o, np.eye(7)[lab])
    img = RS.randn(2, 3, 5, 5).astype(np.float32)
    o = out(2, 3, 7, 9)
    gl.pad(arr(img), o, [[0, 0], [0, 0], [1, 1], [2, 2]])
    close(o, np.pad(img, [[0, 0], [0, 0], [1, 1], [2, 2]]))
    # --- nn -------------------------------------------------------------------------------
    f = RS.randn(4, 3, 3, 3).astype(np.float32)
    o = out(2, 4, 5, 5)
    gl.CuDNN_conv2d(arr(img), arr(f), o, padding=(1, 1), stride=(1, 1))
    ref = torch.nn.functional.conv2d(torch.tensor(img), torch.tensor(f), padding=1).numpy()
    close(o, ref, mm_tol)
    cb = RS.randn(4).astype(np.float32)
    gl.CuDNN_conv2d_with_bias(arr(img), arr(f), arr(cb), o, padding=(1, 1), stride=(1, 1))
    close(o, ref + cb.reshape(1, 4, 1, 1), mm_tol)
    gy = RS.randn(2, 4, 5, 5).astype(np.float32)
    ti, tf_ = torch.tensor(img, requires_grad=True), torch.tensor(f, requires_grad=True)
    torch.nn.functional.conv2d(ti, tf_, padding=1).backward(torch.tensor(gy))
    gx, gf = out(2, 3, 5, 5), out(4, 3, 3, 3)
    gl.CuDNN_conv2d_gradient_of_data(arr(f), arr(gy), gx, padding=(1, 1), stride=(1, 1))
    gl.CuDNN_conv2d_gradient_of_filter(arr(img), arr(gy), gf, padding=(1, 1), stride=(1, 1))
    close(gx, ti.grad.numpy(), mm_tol)
    close(gf, tf_.grad.numpy(), mm_tol)
    pimg = RS.randn(2, 3, 6, 6).astype(np.float32)
    o = out(2, 3, 3, 3)
    gl.CuDNN_max_pooling2d(arr(pimg), 2, 2, o, padding=0, stride=2)
    close(o, torch.nn.functional.max_pool2d(torch.tensor(pimg), 2, 2).numpy())
    gl.CuDNN_average_pooling2d(arr(pimg), 2, 2, o, padding=0, stride=2)
    close(o, torch.nn.functional.avg_pool2d(torch.tensor(pimg), 2, 2).numpy(), 1e-4)
    o = out(6, 10)
    gl.softmax(arr(x), o)
    close(o, torch.softmax(torch.tensor(x), -1).numpy(), 1e-4)
    onehot = np.eye(10, dtype=np.float32)[RS.randint(0, 10, 6)]
    gl.softmax_cross_entropy(arr(x), arr(onehot), o6)
    close(o6, -(onehot * torch.log_softmax(torch.tensor(x), -1).numpy()).sum(1), 1e-4)
    o = out(6, 10)
    gl.softmax_cross_entropy_gradient(arr(x), arr(onehot), arr(np.ones(6)), o)
    close(o, torch.softmax(torch.tensor(x), -1).numpy() - onehot, 1e-4)
    sl = np.array([1, 3, -1, 0, 9, 2])
    gl.softmax_cross_entropy_sparse(arr(x), arr(sl), -1, o6)
    ls = -torch.log_softmax(torch.tensor(x), -1).numpy()
    close(o6, np.array([ls[i, sl[i]] if sl[i] >= 0 else 0.0 for i in range(6)]), 1e-4)
    scale, lbias = RS.rand(10).astype(np.float32) + 0.5, RS.randn(10).astype(np.float32)
    mean, var, o = out(6, 1), out(6, 1), out(6, 10)
    gl.layer_normalization(arr(x), arr(scale), arr(lbias), mean, var, o, 1e-5)
    close(o, torch.nn.functional.layer_norm(torch.tensor(x), (10,), torch.tensor(scale), torch.tensor(lbias), 1e-5).numpy(), 1e-3)
    tx = torch.tensor(x, requires_grad=True)
    ts, tb = torch.tensor(scale, requires_grad=True), torch.tensor(lbias, requires_grad=True)
    torch.nn.functional.layer_norm(tx, (10,), ts, tb, 1e-5).backward(torch.tensor(g))
    dx, ds, db = out(6, 10), out(10), out(10)
    gl.layer_normalization_gradient(arr(g), arr(x), arr(scale), dx, ds, db, mean, var, 1e-5)
    close(dx, tx.grad.numpy(), 1e-3)
    close(ds, ts.grad.numpy(), 1e-3)
    close(db, tb.grad.numpy(), 1e-3)
    bscale, bbias = RS.rand(3).astype(np.float32) + 0.5, RS.randn(3).astype(np.float32)
    rm, rv, sm, sv, o = arr(np.zeros(3)), arr(np.ones(3)), out(3), out(3), out(2, 3, 6, 6)
    gl.CuDNN_Batch_Normalization(arr(pimg), arr(bscale), arr(bbias), o, rm, rv, sm, sv, 0.1, 1e-5)
    tbn = torch.nn.BatchNorm2d(3, momentum=0.1, eps=1e-5)
    with torch.no_grad():
        tbn.weight.copy_(torch.tensor(bscale))
        tbn.bias.copy_(torch.tensor(bbias))
    tp = torch.tensor(pimg, requires_grad=True)
    ty = tbn(tp)
    close(o, ty.detach().numpy(), 1e-3)
    close(rm, tbn.running_mean.numpy(), 1e-4)
    close(rv, tbn.running_var.numpy(), 1e-4)
    gbn = RS.randn(2, 3, 6, 6).astype(np.float32)
    ty.backward(torch.tensor(gbn))
    dx, ds, db = out(2, 3, 6, 6), out(3), out(3)
    gl.CuDNN_Batch_Normalization_gradient(arr(gbn), arr(pimg), arr(bscale), dx, ds, db, sm, sv, 1e-5)
    close(dx, tp.grad.numpy(), 1e-3)
    close(ds, tbn.weight.grad.numpy(), 1e-3)
    close(db, tbn.bias.grad.numpy(), 1e-3)
    table = RS.randn(20, 5).astype(np.float32)
    ids = RS.randint(0, 20, (4, 3))
    o = out(4, 3, 5)
    gl.embedding_lookup(arr(table), arr(ids), o)
    close(o, table[ids])
    gt = out(20, 5)
    ge = RS.randn(4, 3, 5).astype(np.float32)
    gl.embedding_lookup_gradient(arr(ge), arr(ids), gt)
    ref = np.zeros_like(table)
    np.add.at(ref, ids.reshape(-1), ge.reshape(-1, 5))
    close(gt, ref, 1e-4)
    d1, d2 = out(6, 10), out(6, 10)
    gl.dropout(arr(np.ones((6, 10))), 0.5, d1, 123)
    gl.dropout_gradient_recompute(arr(np.ones((6, 10))), 0.5, d2, 123)
    close(d1, d2.asnumpy())                                         # same seed -> same mask
    assert set(np.unique(d1.asnumpy())) <= {0.0, 2.0}
    # --- optimizers (reference tests/test_optimizer.py: one step vs the numpy formula) ---------
    p0, gr = RS.randn(7, 3).astype(np.float32), RS.randn(7, 3).astype(np.float32)
    p = arr(p0)
    gl.sgd_update(p, arr(gr), 0.1, 0.0)
    close(p, p0 - 0.1 * gr, 1e-5)
    p, v = arr(p0), arr(np.zeros((7, 3)))
    gl.momentum_update(p, arr(gr), v, 0.1, 0.9, False, 0.0)
    close(p, p0 - 0.1 * gr, 1e-5)
    p, m, vv = arr(p0), arr(np.zeros((7, 3))), arr(np.zeros((7, 3)))
    gl.adam_update(p, arr(gr), m, vv, 0.01, 0.9, 0.999, 0.9, 0.999, 1e-7, 0.0)
    mh, vh = 0.1 * gr / 0.1, 0.001 * gr * gr / 0.001
    close(p, p0 - 0.01 * mh / (np.sqrt(vh) + 1e-7), 1e-4)
    p, m, vv = arr(p0), arr(np.zeros((7, 3))), arr(np.zeros((7, 3)))
    gl.adamw_update(p, arr(gr), m, vv, 0.01, 0.9, 0.999, 0.9, 0.999, 1e-7, 0.1)
    close(p, p0 - 0.01 * (mh / (np.sqrt(vh) + 1e-7) + 0.1 * p0), 1e-4)
    p, m, vv = arr(p0), arr(np.zeros((7, 3))), arr(np.zeros((7, 3)))
    gl.lamb_update(p, arr(gr), m, vv, 0.01, 0.9, 0.999, 0.9, 0.999, 1e-7, 0.1)
    upd = mh / (np.sqrt(vh) + 1e-7) + 0.1 * p0
    close(p, p0 - 0.01 * np.linalg.norm(p0) / np.linalg.norm(upd) * upd, 1e-4)
    p, acc = arr(p0), arr(np.zeros((7, 3)))
    gl.adagrad_update(p, arr(gr), acc, 0.1, 1e-7, 0.0)
    close(p, p0 - 0.1 * gr / (np.abs(gr) + 1e-7), 1e-4)
    # --- MoE layout transforms --------------------------------------------------------------
    T, E, Cap, D = 12, 3, 5, 4
    tok = RS.randn(T, D).astype(np.float32)
    eidx = RS.randint(0, E, T)
    loc = np.zeros(T, dtype=np.int64)
    cnt = [0] * E
    for t in range(T):
        loc[t] = cnt[eidx[t]]
        cnt[eidx[t]] += 1
    disp = out(E * Cap, D)
    gl.layout_transform_top1(arr(tok), arr(eidx), arr(loc), disp, Cap)
    ref = np.zeros((E * Cap, D), np.float32)
    for t in range(T):
        if loc[t] < Cap:
            ref[eidx[t] * Cap + loc[t]] = tok[t]
    close(disp, ref)
    gates = RS.rand(T).astype(np.float32)
    comb = out(T, D)
    gl.reverse_layout_transform_top1(disp, arr(eidx), arr(loc), arr(gates), comb, Cap)
    ref2 = np.stack([gates[t] * tok[t] if loc[t] < Cap else np.zeros(D) for t in range(T)])
    close(comb, ref2, 1e-5)
    gg = out(T)
    gl.reverse_layout_transform_top1_gradient_gate(arr(tok), disp, arr(eidx), arr(loc), gg, Cap)
    close(gg, np.array([(tok[t] * tok[t]).sum() if loc[t] < Cap else 0.0 for t in range(T)]), 1e-4)
    h = RS.randn(2 * 3 * 2 * 2, 4).astype(np.float32)          # G=2, N=3: [gpu][node][gpu'][rows=2]
    ho, hb = out(*h.shape), out(*h.shape)
    gl.ha2a_layout_transform(arr(h), ho, 3, 2)
    gl.ha2a_reverse_layout_transform(ho, hb, 3, 2)
    close(hb, h)
    close(ho, h.reshape(2, 3, 2, 2, 4).transpose(1, 2, 0, 3, 4).reshape(h.shape))
    # --- initialisers (reference tests/test_gpu_initializers.py: moments) ---------------------
    big = out(200, 200)
    gl.normal_init(big, 1.0, 2.0, 5)
    v = big.asnumpy()
    assert abs(v.mean() - 1.0) < 0.05 and abs(v.std() - 2.0) < 0.05
    gl.uniform_init(big, -1.0, 3.0, 5)
    v = big.asnumpy()
    assert v.min() >= -1.0 and v.max() <= 3.0 and abs(v.mean() - 1.0) < 0.05
    gl.truncated_normal_init(big, 0.0, 1.0, 5)
    v = big.asnumpy()
    assert np.abs(v).max() <= 2.0 + 1e-5 and abs(v.mean()) < 0.05


def test_gpu_links_surface_matches_reference_names():
    names = ('abs_val matrix_elementwise_add addmm arange argmax argsort array_set average_pooling2d baddbmm batch_matrix_multiply '
             'binary_cross_entropy bool bool_val bool_matrix broadcast_to broadcast_shape clamp clone concat concatenate const_pow '
             'conv2d_broadcast_to conv2d conv2d_reduce_sum cross_entropy cross_entropy_sparse CuSparse_Csrmv CuSparse_Csrmm '
             'CuDNN_average_pooling2d CuDNN_Batch_Normalization CuDNN_conv2d CuDNN_conv2d_with_bias CuDNN_Dropout CuDNN_max_pooling2d '
             'CuDNN_softmax_cross_entropy CuDNN_softmax cumsum_with_bias matrix_dot dropout2d dropout embedding_lookup exp floor gather '
             'gelu group_topk_idx ha2a_layout_transform indexedslice_oneside_add indexing normal_init uniform_init truncated_normal_init '
             'instance_normalization2d bicubic_interpolate layer_normalization layout_transform_top1 layout_transform_top2 leaky_relu '
             'matmul_with_bias log_link masked_fill matrix_elementwise_divide_const matrix_elementwise_divide matrix_multiply matrix_rsqrt '
             'matrix_sqrt matrix_transpose max max_mat max_pooling2d min min_mat minus_by_const matrix_elementwise_minus '
             'matrix_elementwise_multiply_by_const matrix_elementwise_multiply nll_loss_link norm one_hot matrix_opposite '
             'add_l2_regularization sgd_update momentum_update adagrad_update adam_update adamw_update lamb_update pad pow_matrix '
             'reduce_mean reduce_sum_axis_zero reduce_sum relu repeat array_reshape reverse_layout_transform_top1 '
             'reverse_layout_transform_top2 reverse_layout_transform_no_gate roll sam_group_sum_link sammax_link scatter1d scatter '
             'sigmoid sin cos slice_assign slice_assign_matrix slice_by_matrix matrix_slice softmax_cross_entropy '
             'softmax_cross_entropy_sparse softmax tanh topk_idx topk_val where where_const').split()
    missing = [n for n in names if not hasattr(gl, n)]
    assert not missing, missing
    for n in names:                                              # every one also has its *_gradient twin where the reference has
        assert callable(getattr(gl, n))


def test_gpu_links_on_host_arrays():
    _run_all(ht.cpu(0), 1e-5)


@pytest.mark.gpu
def test_gpu_links_on_b200():
    _run_all(ht.gpu(0), 2e-4)
