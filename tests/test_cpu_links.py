"""CPU op library vs numpy / torch oracles.  Pattern: reference tests/test_dnnl_op.py (every
``cpu_links`` entry point checked against numpy)."""
import numpy as np
import pytest
import torch

import hetu as ht
from hetu_61a7_b200 import cpu_links as cl

pytestmark = pytest.mark.skipif(not cl.available(), reason='libhetu_cpu.so not built')
rs = np.random.RandomState(3)


def T(*shape):
    return torch.from_numpy(rs.randn(*shape).astype(np.float32))


@pytest.mark.parametrize('ta,tb', [(False, False), (True, False), (False, True), (True, True)])
def test_matmul(ta, tb):
    M, N, K = 67, 45, 300
    a = T(K, M) if ta else T(M, K)
    b = T(N, K) if tb else T(K, N)
    c = torch.empty(M, N)
    cl.matrix_multiply(a, ta, b, tb, c)
    ref = (a.t() if ta else a).double() @ (b.t() if tb else b).double()
    assert torch.allclose(c.double(), ref, atol=1e-4, rtol=1e-5)


def test_elementwise_and_activations():
    x, y = T(33, 17), T(33, 17)
    out = torch.empty_like(x)
    for fn, ref in [(cl.relu, x.relu()), (cl.sigmoid, x.sigmoid()), (cl.tanh, x.tanh()), (cl.opposite, -x),
                    (cl.gelu, torch.nn.functional.gelu(x))]:
        fn(x, out)
        assert torch.allclose(out, ref, atol=1e-6), fn.__name__
    cl.sqrt(x.abs(), out)
    assert torch.allclose(out, x.abs().sqrt(), atol=1e-6)
    cl.rsqrt(x.abs() + 1, out)
    assert torch.allclose(out, (x.abs() + 1).rsqrt(), atol=1e-6)
    cl.matrix_elementwise_add(x, y, out)
    assert torch.equal(out, x + y)
    cl.matrix_elementwise_multiply(x, y, out)
    assert torch.equal(out, x * y)
    cl.matrix_elementwise_divide(x, y, out)
    assert torch.allclose(out, x / y, rtol=1e-6)
    cl.matrix_elementwise_add_by_const(x, 2.5, out)
    assert torch.equal(out, x + 2.5)
    cl.matrix_elementwise_multiply_by_const(x, -3.0, out)
    assert torch.equal(out, x * -3.0)
    cl.relu_gradient(x, y, out)
    assert torch.equal(out, y * (x > 0))
    xg = x.clone().requires_grad_()
    torch.nn.functional.gelu(xg).backward(y)
    cl.gelu_gradient(x, y, out)
    assert torch.allclose(out, xg.grad, atol=1e-5)


def test_softmax_and_ce():
    x = T(29, 10)
    lab = torch.eye(10)[torch.from_numpy(rs.randint(0, 10, 29))]
    out = torch.empty_like(x)
    cl.softmax(x, out)
    assert torch.allclose(out, x.softmax(-1), atol=1e-6)
    loss = torch.empty(29)
    cl.softmax_crossentropy(x, lab, loss)
    assert torch.allclose(loss, -(lab * x.log_softmax(-1)).sum(-1), atol=1e-5)
    g = T(29)
    gx = torch.empty_like(x)
    cl.softmax_crossentropy_gradient(x, lab, g, gx)
    assert torch.allclose(gx, (x.softmax(-1) - lab) * g[:, None], atol=1e-5)


def test_conv_pool_bn():
    x, w = T(3, 5, 12, 12), T(7, 5, 3, 3)
    for pad, stride in [(0, 1), (1, 2)]:
        ref = torch.nn.functional.conv2d(x, w, padding=pad, stride=stride)
        y = torch.empty_like(ref)
        cl.conv2d(x, w, y, pad, stride)
        assert torch.allclose(y, ref, atol=1e-4)
        gy = torch.randn_like(ref)
        xg, wg = x.clone().requires_grad_(), w.clone().requires_grad_()
        torch.nn.functional.conv2d(xg, wg, padding=pad, stride=stride).backward(gy)
        gx, gw = torch.empty_like(x), torch.empty_like(w)
        cl.conv2d_gradient_of_data(w, gy, gx, pad, stride)
        cl.conv2d_gradient_of_filter(x, gy, gw, pad, stride)
        assert torch.allclose(gx, xg.grad, atol=1e-4)
        assert torch.allclose(gw, wg.grad, atol=1e-3)
    ref = torch.nn.functional.max_pool2d(x, 2, 2)
    y = torch.empty_like(ref)
    cl.max_pool(x, 2, 2, y, 0, 2)
    assert torch.equal(y, ref)
    xg = x.clone().requires_grad_()
    gy = torch.randn_like(ref)
    torch.nn.functional.max_pool2d(xg, 2, 2).backward(gy)
    gx = torch.empty_like(x)
    cl.max_pool_gradient(x, gy, 2, 2, gx, 0, 2)
    assert torch.allclose(gx, xg.grad)
    ref = torch.nn.functional.avg_pool2d(x, 2, 2)
    cl.avg_pool(x, 2, 2, y, 0, 2)
    assert torch.allclose(y, ref, atol=1e-6)
    xg = x.clone().requires_grad_()
    torch.nn.functional.avg_pool2d(xg, 2, 2).backward(gy)
    cl.avg_pool_gradient(gy, 2, 2, gx, 0, 2)
    assert torch.allclose(gx, xg.grad, atol=1e-6)
    # batch norm
    scale, bias = T(5), T(5)
    rm, rv, sm, sv = torch.zeros(5), torch.ones(5), torch.empty(5), torch.empty(5)
    y = torch.empty_like(x)
    cl.batch_norm(x, scale, bias, y, rm, rv, sm, sv, momentum=0.1, eps=1e-5)
    xg, sg, bg = x.clone().requires_grad_(), scale.clone().requires_grad_(), bias.clone().requires_grad_()
    ref = torch.nn.functional.batch_norm(xg, None, None, sg, bg, training=True, eps=1e-5)
    assert torch.allclose(y, ref, atol=1e-4)
    gy = torch.randn_like(x)
    ref.backward(gy)
    gx, gs, gb = torch.empty_like(x), torch.empty(5), torch.empty(5)
    cl.batch_norm_gradient(gy, x, scale, gx, gs, gb, sm, sv)
    assert torch.allclose(gx, xg.grad, atol=1e-4)
    assert torch.allclose(gs, sg.grad, atol=1e-3) and torch.allclose(gb, bg.grad, atol=1e-3)


def test_shape_ops():
    x = T(4, 5, 6)
    y = torch.empty(6, 4, 5)
    cl.transpose(x, y, (2, 0, 1))
    assert torch.equal(y, x.permute(2, 0, 1))
    a, b = T(3, 4), T(3, 6)
    out = torch.empty(3, 10)
    cl.concat(a, b, out, axis=1)
    assert torch.equal(out, torch.cat([a, b], 1))
    ga, gb = torch.empty_like(a), torch.empty_like(b)
    cl.concat_gradient(out, ga, axis=1, id=0)
    cl.concat_gradient(out, gb, axis=1, id=1)
    assert torch.equal(ga, a) and torch.equal(gb, b)
    x4 = T(2, 3, 4, 5)
    y4 = torch.empty(2, 3, 7, 9)
    pads = [(0, 0), (0, 0), (1, 2), (3, 1)]
    cl.pad(x4, y4, pads)
    assert torch.equal(y4, torch.nn.functional.pad(x4, (3, 1, 1, 2)))
    g4 = torch.empty_like(x4)
    cl.pad_gradient(y4, g4, pads)
    assert torch.equal(g4, x4)
    bias = T(6)
    out = torch.empty(4, 6)
    cl.broadcast_to(bias, out)
    assert torch.equal(out, bias.expand(4, 6))
    red = torch.empty(5, 6)
    cl.reduce_sum_axis_zero(x, red)
    assert torch.allclose(red, x.sum(0), atol=1e-6)
    cl.array_set(red, 1.5)
    assert bool((red == 1.5).all())
    tab = T(50, 8)
    ids = torch.from_numpy(rs.randint(0, 50, (7, 3)).astype(np.float32))
    out = torch.empty(7, 3, 8)
    cl.embedding_lookup(tab, ids, out)
    assert torch.equal(out, tab[ids.long()])
    keep = 0.7
    big = torch.ones(100000)
    dr = torch.empty_like(big)
    cl.dropout(big, keep, dr)
    assert abs(float((dr > 0).float().mean()) - keep) < 0.01
    assert torch.allclose(dr[dr > 0], torch.tensor(1 / keep))


def test_optimizers_and_init():
    p, g = T(1000), T(1000)
    ref = p - 0.1 * g
    cl.sgd_update(p, g, 0.1)
    assert torch.allclose(p, ref, atol=1e-6)
    tab, val = T(20, 4), T(3, 4)
    ids = torch.tensor([1, 5, 1])
    ref = tab.clone()
    ref.index_add_(0, ids, -0.5 * val)
    cl.sgd_update_sparse(tab, ids, val, 0.5)
    assert torch.allclose(tab, ref, atol=1e-6)
    p, g, acc = T(100), T(100), torch.zeros(100)
    ref = p - 0.1 * g / (g.abs() + 1e-7)
    cl.adagrad_update(p, g, acc, 0.1, 1e-7)
    assert torch.allclose(p, ref, atol=1e-5)
    p0 = T(100)
    p, g = p0.clone(), T(100)
    m, v = torch.zeros(100), torch.zeros(100)
    tp = p0.clone().requires_grad_()
    opt = torch.optim.Adam([tp], lr=0.01, betas=(0.9, 0.999), eps=1e-7)
    for step in range(1, 4):
        tp.grad = g.clone()
        opt.step()
        cl.adam_update(p, g, m, v, 0.01, 0.9, 0.999, 0.9 ** step, 0.999 ** step, 1e-7, step=step)
    assert torch.allclose(p, tp.detach(), atol=1e-5)
    p, g, vel = T(100), T(100), torch.zeros(100)
    ref = p - 0.1 * g
    cl.momentum_update(p, g, vel, 0.1, 0.9, False)
    assert torch.allclose(p, ref, atol=1e-6)
    big = torch.empty(200000)
    cl.uniform_init(big, -2.0, 2.0, 1)
    assert -2.0 <= float(big.min()) and float(big.max()) <= 2.0 and abs(float(big.mean())) < 0.02
    cl.normal_init(big, 1.0, 0.5, 2)
    assert abs(float(big.mean()) - 1.0) < 0.01 and abs(float(big.std()) - 0.5) < 0.01
    cl.truncated_normal_init(big, 0.0, 1.0, 3)
    assert float(big.abs().max()) <= 2.0


def test_cpu_executor_uses_native_library():
    """Config 1 (BASELINE.json): 3-layer MLP, CIFAR10-shaped synthetic data, CPU executor."""
    from hetu_61a7_b200 import _lib
    x = ht.Variable('x')
    y_ = ht.Variable('y_')
    h = x
    dims = [3072, 256, 256, 10]
    for i in range(3):
        w = ht.init.random_normal((dims[i], dims[i + 1]), stddev=0.05, name='w%d' % i)
        h = ht.matmul_op(h, w)
        if i < 2:
            h = ht.relu_op(h)
    loss = ht.reduce_mean_op(ht.softmaxcrossentropy_op(h, y_), [0])
    train = ht.optim.SGDOptimizer(0.05).minimize(loss)
    ex = ht.Executor([loss, train], ctx=ht.cpu(0))
    X = rs.randn(64, 3072).astype(np.float32)
    Y = np.eye(10, dtype=np.float32)[rs.randint(0, 10, 64)]
    assert _lib.cpu_lib() is not None
    losses = [float(ex.run(feed_dict={x: X, y_: Y})[0].asnumpy()) for _ in range(15)]
    assert losses[-1] < losses[0] * 0.7, losses
