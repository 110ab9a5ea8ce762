"""Numerics of every hand-written sm_100a kernel against a plain PyTorch fp32 reference of
the same op (pattern: reference tests/test_gpu_op.py - kernel vs numpy oracle)."""
import numpy as np
import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]


@pytest.fixture(scope='module')
def L():
    from hetu_61a7_b200 import _lib
    assert _lib.cuda_lib() is not None, 'libhetu_b200.so must be built (python __graft_entry__.py)'
    return _lib


def dev():
    return torch.device('cuda', 0)


def rnd(*shape, seed=0, scale=1.0):
    g = torch.Generator(device='cpu').manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(dev())


# ------------------------------------------------------------------ elementwise
def test_unary_binary(L):
    from hetu_61a7_b200.ops import elementwise as E
    x = rnd(1000, 37)
    y = rnd(1000, 37, seed=1)
    cases = {
        'RELU': torch.relu(x), 'EXP': torch.exp(x), 'TANH': torch.tanh(x), 'SIGMOID': torch.sigmoid(x),
        'GELU': torch.nn.functional.gelu(x), 'ABS': x.abs(), 'OPPOSITE': -x, 'SIN': x.sin(), 'COS': x.cos(),
        'FLOOR': x.floor(), 'SIGN': x.sign(),
    }
    for name, ref in cases.items():
        out = E._unary_kernel(E.U[name], x)
        torch.testing.assert_close(out, ref, rtol=2e-5, atol=2e-6, msg=name)
    torch.testing.assert_close(E._unary_kernel(E.U['ADD_CONST'], x, 3.0), x + 3)
    torch.testing.assert_close(E._unary_kernel(E.U['MUL_CONST'], x, -2.5), x * -2.5)
    torch.testing.assert_close(E._unary_kernel(E.U['SQRT'], x.abs()), x.abs().sqrt())
    torch.testing.assert_close(E._unary_kernel(E.U['CLAMP'], x, -0.5, 0.5), x.clamp(-0.5, 0.5))
    torch.testing.assert_close(E._unary_kernel(E.U['LEAKY_RELU'], x, 0.1), torch.nn.functional.leaky_relu(x, 0.1))
    for name, ref in {'ADD': x + y, 'SUB': x - y, 'MUL': x * y, 'DIV': x / y,
                      'RELU_GRAD': y * (x > 0), 'MAX': torch.maximum(x, y)}.items():
        torch.testing.assert_close(E._binary_kernel(E.B[name], x, y), ref, rtol=2e-5, atol=2e-6, msg=name)
    b = rnd(37, seed=2)
    torch.testing.assert_close(E._bcast_kernel(E.B['ADD'], x, b), x + b)
    c = rnd(1000, 1, seed=3)
    torch.testing.assert_close(E._bcast_kernel(E.B['MUL'], x, c), x * c)
    # odd length / misaligned view
    z = rnd(4099)[1:]
    torch.testing.assert_close(E._unary_kernel(E.U['RELU'], z), z.relu())


def test_strided_copy_and_concat(L):
    from hetu_61a7_b200.ops import shape as S
    x = rnd(6, 5, 8, 12)
    torch.testing.assert_close(S.materialize(x.permute(0, 2, 1, 3)), x.permute(0, 2, 1, 3).contiguous())
    torch.testing.assert_close(S.materialize(x.permute(3, 1, 0, 2)), x.permute(3, 1, 0, 2).contiguous())
    torch.testing.assert_close(S.materialize(x[:, 1:4, ::2]), x[:, 1:4, ::2].contiguous())
    for shp in ((5, 37, 9, 11), (3, 64, 32, 32), (1, 130, 7, 5)):      # NCHW <-> NHWC: batched tile transpose kernel
        t = rnd(*shp, seed=7)
        torch.testing.assert_close(S.materialize(t.permute(0, 2, 3, 1)), t.permute(0, 2, 3, 1).contiguous())
        torch.testing.assert_close(S.materialize(t.permute(0, 3, 1, 2)), t.permute(0, 3, 1, 2).contiguous())
    m2 = rnd(300, 70, seed=8)
    torch.testing.assert_close(S.materialize(m2.t()), m2.t().contiguous())
    g = rnd(16, 40)                                       # gradient of a row reduction: innermost dim broadcast
    torch.testing.assert_close(S.materialize(g.unsqueeze(-1).expand(16, 40, 64)), g.unsqueeze(-1).expand(16, 40, 64).contiguous())
    torch.testing.assert_close(S.materialize(g.unsqueeze(-1).expand(16, 40, 6)), g.unsqueeze(-1).expand(16, 40, 6).contiguous())
    a, b = rnd(128, 3328), rnd(128, 256, seed=1)
    torch.testing.assert_close(S._cat([a, b], 1), torch.cat([a, b], 1))
    torch.testing.assert_close(S._cat([a, a], 0), torch.cat([a, a], 0))


def test_reduce_topk_cumsum(L):
    from hetu_61a7_b200.ops import reduce as R
    x = rnd(33, 17, 29)
    for axes in ([0], [1], [2], [0, 1], [1, 2], [0, 1, 2]):
        torch.testing.assert_close(R._reduce(x, axes, False, R.R_SUM), x.sum(dim=axes), rtol=1e-4, atol=1e-4)
        torch.testing.assert_close(R._reduce(x, axes, True, R.R_MEAN), x.mean(dim=axes, keepdim=True), rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(R._reduce(x, [2], False, R.R_MAX), x.amax(dim=2))
    torch.testing.assert_close(R._reduce(x, [0], False, R.R_MIN), x.amin(dim=0))
    big = rnd(16384, 16)
    torch.testing.assert_close(R._reduce(big, [0], False, R.R_SUM), big.sum(0), rtol=1e-4, atol=1e-3)
    wide = rnd(7, 5000)
    torch.testing.assert_close(R._reduce(wide, [1], False, R.R_SUM), wide.sum(1), rtol=1e-4, atol=1e-3)
    # topk / argmax / cumsum through the ops
    g = torch.softmax(rnd(4096, 16), dim=-1)
    op = R.TopKIdxOp.__new__(R.TopKIdxOp)
    op.k = 2
    idx = R.TopKIdxOp.compute(op, [g])
    assert torch.equal(idx.long(), torch.topk(g, 2, dim=-1).indices)
    am = R.ArgmaxOp.__new__(R.ArgmaxOp)
    am.dim = 1
    assert torch.equal(R.ArgmaxOp.compute(am, [g]).long(), g.argmax(1))
    mask = torch.nn.functional.one_hot(g.argmax(1), 16).float()
    cs = R.CumsumOp.__new__(R.CumsumOp)
    cs.bias, cs.dim = -1, 0
    torch.testing.assert_close(R.CumsumOp.compute(cs, [mask]), torch.cumsum(mask, 0) - 1)
    for T, E in [(33001, 2), (1500, 100)]:                      # few-column and many-column kernels
        m2 = (torch.rand(T, E, device='cuda') < 0.3).float()
        torch.testing.assert_close(R.CumsumOp.compute(cs, [m2]), torch.cumsum(m2, 0) - 1)


# ------------------------------------------------------------------------ GEMM
GEMM_SHAPES = [(128, 128, 128), (256, 384, 512), (1000, 520, 264), (2048, 2048, 2048), (128, 256, 256),
               (4096, 1024, 512), (1160, 776, 2048), (16384, 2, 2048), (2048, 3, 16384), (5000, 7, 1001),
               (96, 40, 72), (128, 1, 3584), (128, 256, 13), (4096, 64, 1024), (300, 32, 2048)]


@pytest.mark.parametrize('tA', [False, True])
@pytest.mark.parametrize('tB', [False, True])
def test_gemm_all_layouts(L, tA, tB):
    from hetu_61a7_b200.ops.linalg import gemm, set_matmul_precision
    set_matmul_precision('tf32')
    for (m, n, k) in GEMM_SHAPES:
        a = rnd(*((k, m) if tA else (m, k)), seed=m + k)
        b = rnd(*((n, k) if tB else (k, n)), seed=n + k + 1)
        ref = (a.t() if tA else a).double() @ (b.t() if tB else b).double()
        out = gemm(a, b, tA, tB)
        err = (out.double() - ref).abs().max().item()
        scale = ref.abs().max().item() + 1e-6
        assert err / scale < 2e-3, 'gemm %s tA=%s tB=%s rel err %.3e' % ((m, n, k), tA, tB, err / scale)


def test_gemm_fp32_mode_and_epilogue(L):
    from hetu_61a7_b200.ops.linalg import gemm, set_matmul_precision, ACT_RELU
    a, b, bias, c = rnd(333, 200), rnd(200, 150, seed=1), rnd(150, seed=2), rnd(333, 150, seed=3)
    set_matmul_precision('fp32')
    try:
        out = gemm(a, b)
        torch.testing.assert_close(out, a @ b, rtol=1e-5, atol=1e-4)
        out = gemm(a, b, bias=bias, c_in=c, alpha=0.5, beta=2.0, act=ACT_RELU)
        torch.testing.assert_close(out, torch.relu(0.5 * (a @ b) + 2.0 * c + bias), rtol=1e-5, atol=1e-4)
    finally:
        set_matmul_precision('tf32')
    a, b, bias = rnd(512, 256), rnd(256, 384, seed=1), rnd(384, seed=2)
    out = gemm(a, b, bias=bias, act=ACT_RELU)
    ref = torch.relu(a @ b + bias)
    assert (out - ref).abs().max().item() / ref.abs().max().item() < 2e-3      # tf32 operands
    # persistent 2-SM (cta_group::2) kernel: fused epilogue, odd edges, batched operands
    a, b, bias, c = rnd(1100, 4096), rnd(4096, 1032, seed=1), rnd(1032, seed=2), rnd(1100, 1032, seed=3)
    out = gemm(a, b, bias=bias, c_in=c, alpha=0.5, beta=2.0, act=ACT_RELU)
    ref = torch.relu(0.5 * (a.double() @ b.double()) + 2.0 * c.double() + bias.double())
    assert (out.double() - ref).abs().max().item() / ref.abs().max().item() < 2e-3
    a, b = rnd(3, 1024, 2048), rnd(3, 2048, 512, seed=1)
    ref = torch.bmm(a.double(), b.double())
    assert (gemm(a, b).double() - ref).abs().max().item() / ref.abs().max().item() < 2e-3
    w = rnd(2048, 512, seed=4)                                                  # shared (un-batched) B
    ref = a.double() @ w.double()
    assert (gemm(a, w).double() - ref).abs().max().item() / ref.abs().max().item() < 2e-3


def test_gemm_batched(L):
    from hetu_61a7_b200.ops.linalg import gemm
    a, b = rnd(6, 128, 64), rnd(6, 64, 128, seed=1)
    ref = torch.bmm(a, b)
    out = gemm(a, b)
    assert (out - ref).abs().max().item() / ref.abs().max().item() < 2e-3
    out = gemm(a, a, False, True)
    ref = torch.bmm(a, a.transpose(1, 2))
    assert (out - ref).abs().max().item() / ref.abs().max().item() < 2e-3


# --------------------------------------------------------------------------- nn
def test_softmax_layernorm_xent(L):
    from hetu_61a7_b200.ops import nn as N, loss as LS
    for cols in (10, 16, 100, 1000, 3000):
        x = rnd(257, cols)
        torch.testing.assert_close(N.softmax_rows(x), torch.softmax(x, -1), rtol=1e-4, atol=1e-6)
    x = rnd(300, 1024)
    gamma, beta = rnd(1024, seed=1), rnd(1024, seed=2)
    ln = N.LayerNormOp.__new__(N.LayerNormOp)
    ln.eps = 1e-5
    y = N.LayerNormOp.compute(ln, [x, gamma, beta])
    torch.testing.assert_close(y, torch.nn.functional.layer_norm(x, (1024,), gamma, beta, 1e-5), rtol=1e-4, atol=1e-4)
    dy = rnd(300, 1024, seed=3)
    g = N.LayerNormGradientOp.__new__(N.LayerNormGradientOp)
    g.forward_node, g.eps = ln, 1e-5
    N.LayerNormGradientOp.compute(g, [dy, x, gamma])
    xr = x.clone().requires_grad_(True)
    gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    torch.nn.functional.layer_norm(xr, (1024,), gr, br, 1e-5).backward(dy)
    torch.testing.assert_close(g.results[0], xr.grad, rtol=1e-3, atol=1e-4)
    torch.testing.assert_close(g.results[1], gr.grad, rtol=1e-3, atol=1e-3)
    torch.testing.assert_close(g.results[2], br.grad, rtol=1e-3, atol=1e-3)
    # softmax cross entropy (dense + sparse labels)
    logits = rnd(513, 10)
    lab = torch.randint(0, 10, (513,), device=dev())
    onehot = torch.nn.functional.one_hot(lab, 10).float()
    op = LS.SoftmaxCrossEntropyOp.__new__(LS.SoftmaxCrossEntropyOp)
    loss = LS.SoftmaxCrossEntropyOp.compute(op, [logits, onehot])
    ref = torch.nn.functional.cross_entropy(logits, lab, reduction='none')
    torch.testing.assert_close(loss, ref, rtol=1e-4, atol=1e-5)
    gop = LS.SoftmaxCrossEntropyGradientOp.__new__(LS.SoftmaxCrossEntropyGradientOp)
    gg = rnd(513, seed=5)
    dx = LS.SoftmaxCrossEntropyGradientOp.compute(gop, [logits, onehot, gg])
    torch.testing.assert_close(dx, (torch.softmax(logits, -1) - onehot) * gg[:, None], rtol=1e-4, atol=1e-6)
    sop = LS.SoftmaxCrossEntropySparseOp.__new__(LS.SoftmaxCrossEntropySparseOp)
    sop.ignored_index = 3
    sl = LS.SoftmaxCrossEntropySparseOp.compute(sop, [logits, lab.float()])
    torch.testing.assert_close(sl, torch.where(lab == 3, torch.zeros_like(ref), ref), rtol=1e-4, atol=1e-5)


def test_conv_pool_bn(L):
    from hetu_61a7_b200.ops import nn as N
    import torch.nn.functional as F
    x, w, b = rnd(8, 16, 32, 32), rnd(32, 16, 3, 3, seed=1, scale=0.1), rnd(32, seed=2)
    for pad, stride in ((1, 1), (0, 2), (1, 2)):
        y = N.conv2d_forward(x, w, b, pad, stride)
        ref = F.conv2d(x, w, b, stride=stride, padding=pad)
        assert (y - ref).abs().max().item() < 2e-2
        dy = rnd(*ref.shape, seed=3)
        dx = N.conv2d_grad_data(dy, w, tuple(x.shape), pad, stride)
        dw = N.conv2d_grad_filter(x, dy, tuple(w.shape), pad, stride)
        assert (dx - torch.nn.grad.conv2d_input(x.shape, w, dy, stride=stride, padding=pad)).abs().max().item() < 2e-2
        rw = torch.nn.grad.conv2d_weight(x, w.shape, dy, stride=stride, padding=pad)
        assert (dw - rw).abs().max().item() / rw.abs().max().item() < 3e-3
    for is_max in (True, False):
        op = (N.MaxPool2dOp if is_max else N.AvgPool2dOp).__new__(N.MaxPool2dOp if is_max else N.AvgPool2dOp)
        op.kernel_H = op.kernel_W = 2
        op.padding, op.stride = 0, 2
        y = type(op).compute(op, [x])
        ref = F.max_pool2d(x, 2, 2) if is_max else F.avg_pool2d(x, 2, 2)
        torch.testing.assert_close(y, ref)
    bn = N.BatchNormOp.__new__(N.BatchNormOp)
    bn.momentum, bn.eps = 0.1, 1e-5
    bn.running_mean = bn.running_var = bn.save_mean = bn.save_rstd = None
    g, bb = rnd(16, seed=4), rnd(16, seed=5)
    y = N.BatchNormOp.compute(bn, [x, g, bb])
    ref = F.batch_norm(x, None, None, g, bb, True, 0.1, 1e-5)
    torch.testing.assert_close(y, ref, rtol=1e-3, atol=1e-4)


def test_dropout_recompute(L):
    from hetu_61a7_b200.ops import nn as N
    op = N.DropoutOp.__new__(N.DropoutOp)
    op.keep_prob, op.seed, op.mask, op.id = 0.7, 12345, None, 1
    x = torch.ones(1 << 20, device=dev())
    y1 = op._apply(x)
    y2 = op._apply(x)
    assert torch.equal(y1, y2)                       # same (seed, step) -> same mask
    keep = (y1 != 0).float().mean().item()
    assert abs(keep - 0.7) < 0.01
    torch.testing.assert_close(y1[y1 != 0], torch.full_like(y1[y1 != 0], 1 / 0.7))
    L.call('increment_i64', L.p(N.step_counter(x.device)), L.stream_ptr(x.device))
    assert not torch.equal(op._apply(x), y1)


# -------------------------------------------------------------- embedding / optimizers
def test_embedding_and_sparse_opt(L):
    from hetu_61a7_b200.ops.embedding import gather_rows_kernel
    table = rnd(5000, 128)
    for dt in (torch.float32, torch.int32, torch.int64):
        idx = torch.randint(0, 5000, (128, 26), device=dev()).to(dt)
        torch.testing.assert_close(gather_rows_kernel(table, idx), table[idx.long()])
    import hetu_61a7_b200 as ht
    from hetu_61a7_b200.ndarray import IndexedSlices
    import itertools
    makers = (lambda: ht.optim.SGDOptimizer(0.1), lambda: ht.optim.AdaGradOptimizer(0.1),
              lambda: ht.optim.AdamOptimizer(0.01), lambda: ht.optim.MomentumOptimizer(0.1))
    # (ids, vocab, width): short runs (head sums its run) and very long runs (atomic accumulation)
    for make, (nids, vocab, width) in itertools.product(makers, ((400, 50, 16), (4096, 3, 96))):
        idx = torch.randint(0, vocab, (nids,), device=dev())
        vals = rnd(nids, width, seed=7)
        res = []
        for use_kernel in (True, False):
            opt = make()
            p = rnd(50, width, seed=9).to(dev() if use_kernel else 'cpu')
            opt.tensors, opt.params = [p], [None]
            opt.states = [[opt._new_state(p, i) for i in range(opt.n_states)]]
            opt.step_t = torch.zeros(1, dtype=torch.int64, device=p.device)
            opt.lr_t = torch.full((1,), float(opt.learning_rate), device=p.device)
            for _ in range(2):
                opt.update([IndexedSlices(idx.to(p.device), vals.to(p.device), (50, width))])
            res.append(p.cpu())
        torch.testing.assert_close(res[0], res[1], rtol=2e-4, atol=2e-5, msg=type(opt).__name__)


def test_dense_optimizers_multi_tensor(L):
    import hetu_61a7_b200 as ht
    makers = [lambda: ht.optim.SGDOptimizer(0.1, l2reg=0.01), lambda: ht.optim.MomentumOptimizer(0.1),
              lambda: ht.optim.MomentumOptimizer(0.1, nesterov=True), lambda: ht.optim.AdaGradOptimizer(0.1),
              lambda: ht.optim.AdamOptimizer(0.01), lambda: ht.optim.AdamWOptimizer(0.01, weight_decay=0.1),
              lambda: ht.optim.LambOptimizer(0.01, weight_decay=0.1)]
    shapes = [(13, 256), (256, 256), (3584, 1), (7,), (1000, 33)]
    for make in makers:
        res = []
        for d in (dev(), torch.device('cpu')):
            opt = make()
            ps = [rnd(*s, seed=i).to(d) for i, s in enumerate(shapes)]
            opt.tensors, opt.params = ps, [None] * len(ps)
            opt.states = [[opt._new_state(p, i) for i in range(opt.n_states)] for p in ps]
            opt.step_t = torch.zeros(1, dtype=torch.int64, device=d)
            opt.lr_t = torch.full((1,), float(opt.learning_rate), device=d)
            for step in range(3):
                opt.update([rnd(*s, seed=100 + i + step).to(d) for i, s in enumerate(shapes)])
            res.append([p.cpu() for p in ps])
        for a, b in zip(*res):
            torch.testing.assert_close(a, b, rtol=2e-4, atol=1e-5, msg=type(opt).__name__)


# ------------------------------------------------------------------------- MoE
def test_moe_dispatch_combine(L):
    from hetu_61a7_b200.ops import moe as M
    T, d, E, k = 4096, 256, 8, 2
    cap = k * (T // E)
    x = rnd(T, d)
    logits = rnd(T, E, seed=1)
    gates = torch.softmax(logits, -1)
    top = torch.topk(gates, k, dim=-1)
    idxs = [top.indices[:, j].float().contiguous() for j in range(k)]
    locs = []
    base = torch.zeros(1, E, device=dev())
    for j in range(k):
        m = torch.nn.functional.one_hot(top.indices[:, j], E).float()
        loc = torch.cumsum(m, 0) - 1 + base
        locs.append((loc * m).sum(1).contiguous())
        base = base + m.sum(0, keepdim=True)
    gs = [top.values[:, j].contiguous() for j in range(k)]
    import os
    out = M.moe_dispatch(x, idxs, locs, cap, E)
    os.environ['HETU_FORCE_TORCH'] = '1'
    from hetu_61a7_b200 import _lib
    _lib._FORCE_TORCH = True
    try:
        ref = M.moe_dispatch(x, idxs, locs, cap, E)
        y = rnd(E * cap, d, seed=3)
        ref_c = M.moe_combine(y, idxs, locs, cap, gates=gs)
        ref_g = M.moe_gate_grad(x, y, idxs[0], locs[0], cap)
    finally:
        _lib._FORCE_TORCH = False
        os.environ.pop('HETU_FORCE_TORCH')
    torch.testing.assert_close(out, ref)
    torch.testing.assert_close(M.moe_combine(y, idxs, locs, cap, gates=gs), ref_c, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(M.moe_gate_grad(x, y, idxs[0], locs[0], cap), ref_g, rtol=1e-4, atol=1e-3)


def test_init_random(L):
    import hetu_61a7_b200 as ht
    from hetu_61a7_b200 import initializers as I

    class N:
        id = 3
        ctx = ht.gpu(0)
    for init, mean, std in ((I.NormalInit(0.5, 2.0, (1000, 1000)), 0.5, 2.0),
                            (I.UniformInit(-1, 3, (1000, 1000)), 1.0, 4 / np.sqrt(12)),
                            (I.TruncatedNormalInit(0.0, 1.0, (1000, 1000)), 0.0, 0.8796)):
        n = N()
        init(n, 1)
        t = n.tensor_value.tensor
        assert abs(t.mean().item() - mean) < 0.02 and abs(t.std().item() - std) < 0.02
    assert I.TruncatedNormalInit(0.0, 1.0, (10,)).kind == 3


def test_dataloader_prefetch_matches_feed_dict(L):
    """Dataloader node (pinned dataset views + H2D prefetch on the copy stream, CUDA-graph replay)
    must give the same losses as feeding the same batches through feed_dict."""
    import hetu as ht
    rs = np.random.RandomState(5)
    data = rs.randn(6 * 32, 64).astype(np.float32)
    labels = np.eye(10, dtype=np.float32)[rs.randint(0, 10, 6 * 32)]
    w0 = (rs.randn(64, 10) * 0.1).astype(np.float32)

    def build(use_loader):
        if use_loader:
            x = ht.dataloader_op([ht.Dataloader(data, 32, 'train')])
            y_ = ht.dataloader_op([ht.Dataloader(labels, 32, 'train')])
        else:
            x, y_ = ht.Variable('x'), ht.Variable('y_')
        w = ht.Variable('w', value=w0.copy())
        loss = ht.reduce_mean_op(ht.softmaxcrossentropy_op(ht.matmul_op(x, w), y_), [0])
        train = ht.optim.SGDOptimizer(0.1).minimize(loss)
        return x, y_, loss, ht.Executor({'train': [loss, train]}, ctx=ht.gpu(0), seed=0)

    _, _, _, ex1 = build(True)
    x, y_, _, ex2 = build(False)
    a, b = [], []
    for i in range(14):                       # > 2 epochs: eager steps, graph capture, ring wrap
        a.append(float(ex1.run('train')[0].asnumpy()))
        k = i % 6
        b.append(float(ex2.run('train', feed_dict={x: data[k * 32:(k + 1) * 32], y_: labels[k * 32:(k + 1) * 32]})[0].asnumpy()))
    np.testing.assert_allclose(a, b, rtol=1e-5, atol=1e-6)
    sub = ex1.subexecutor['train']
    assert sub._graph is not None
    # steady state runs through the native step plan (csrc/step_plan.cu): batch slices -> pinned slot -> one H2D ->
    # cudaGraphLaunch in one call, loss copied back behind every step
    assert sub._fast is not None
    c, pending = [], None
    for i in range(14, 30):
        fut = ex1.run('train')[0].asnumpy_async()
        if pending is not None:
            c.append(float(pending.get()))
            assert pending.ready()
        pending = fut
        k = i % 6
        b.append(float(ex2.run('train', feed_dict={x: data[k * 32:(k + 1) * 32], y_: labels[k * 32:(k + 1) * 32]})[0].asnumpy()))
    c.append(float(pending.get()))
    np.testing.assert_allclose(c, b[14:], rtol=1e-5, atol=1e-6)
    # a differently shaped call drops the plan and gives the same values through the general path
    k = 30 % 6
    v = float(ex1.run('train', convert_to_numpy_ret_vals=True)[0])
    w = float(ex2.run('train', feed_dict={x: data[k * 32:(k + 1) * 32], y_: labels[k * 32:(k + 1) * 32]})[0].asnumpy())
    np.testing.assert_allclose(v, w, rtol=1e-5, atol=1e-6)


def test_fused_attention_gpu(L):
    """Attention op on the GPU path (batched tcgen05 GEMMs + in-place softmax kernels) vs fp32 torch."""
    import hetu as ht
    rs = np.random.RandomState(0)
    B, H, S, D = 4, 8, 128, 64
    qv, kv, vv = [rs.randn(B, H, S, D).astype(np.float32) * 0.5 for _ in range(3)]
    mv = (rs.rand(B, 1, 1, S) > 0.8).astype(np.float32) * -1e4
    q, k, v, m = ht.Variable('q', value=qv), ht.Variable('k', value=kv), ht.Variable('v', value=vv), ht.Variable('m')
    out = ht.fused_attention_op(q, k, v, m)
    loss = ht.reduce_sum_op(ht.mul_op(out, out), [0, 1, 2, 3])
    grads = ht.gradients(loss, [q, k, v])
    ex = ht.Executor([out] + grads, ctx=ht.gpu(0))
    for _ in range(4):                                   # eager steps, then CUDA-graph replays
        res = [r.asnumpy() for r in ex.run(feed_dict={m: mv})]
    tq, tk, tv = [torch.tensor(a, dtype=torch.float64, requires_grad=True) for a in (qv, kv, vv)]
    s = tq @ tk.transpose(-1, -2) / np.sqrt(D) + torch.tensor(mv, dtype=torch.float64)
    o = torch.softmax(s, -1) @ tv
    (o * o).sum().backward()
    for got, ref in zip(res, [o.detach(), tq.grad, tk.grad, tv.grad]):
        ref = ref.numpy()
        assert np.abs(got - ref).max() / (np.abs(ref).max() + 1e-9) < 5e-3    # tf32 operands


@pytest.mark.gpu
@pytest.mark.parametrize('mode,tol', [('bf16', 2e-2), ('fp16', 4e-3)])
@pytest.mark.parametrize('m,n,k,tA,tB', [(256, 256, 512, False, False), (384, 200, 1000, True, False),
                                          (512, 96, 136, False, True), (130, 520, 264, True, True)])
def test_gemm_16bit_operands_match_fp32_reference(mode, tol, m, n, k, tA, tB):
    """hetu_gemm_f16 (tcgen05 kind::f16, csrc/gemm_f16_sm100.cu) against the plain fp32 product; reference op:
    src/ops/MatrixMult.cu:22-26."""
    import torch
    from hetu_61a7_b200.ops import linalg
    dev = torch.device('cuda', 0)
    g = torch.Generator().manual_seed(m + n + k)
    a = torch.randn((k, m) if tA else (m, k), generator=g).to(dev)
    b = torch.randn((n, k) if tB else (k, n), generator=g).to(dev)
    bias = torch.randn(n, generator=g).to(dev)
    ref = torch.relu((a.t() if tA else a).double() @ (b.t() if tB else b).double() + bias.double())
    linalg.set_matmul_precision(mode)
    try:
        out = linalg.gemm(a, b, tA, tB, bias=bias, act=linalg.ACT_RELU)
    finally:
        linalg.set_matmul_precision('tf32')
    scale = ref.abs().max().item()
    assert (out.double() - ref).abs().max().item() <= tol * scale


@pytest.mark.gpu
def test_gemm_16bit_batched():
    import torch
    from hetu_61a7_b200.ops import linalg
    dev = torch.device('cuda', 0)
    g = torch.Generator().manual_seed(5)
    a = torch.randn(4, 256, 320, generator=g).to(dev)
    b = torch.randn(4, 320, 192, generator=g).to(dev)
    linalg.set_matmul_precision('bf16')
    try:
        out = linalg.gemm(a, b)
    finally:
        linalg.set_matmul_precision('tf32')
    ref = a.double() @ b.double()
    assert (out.double() - ref).abs().max().item() <= 2e-2 * ref.abs().max().item()


@pytest.mark.gpu
@pytest.mark.parametrize('m,n,k', [(1024, 512, 256), (2048, 2048, 2048), (700, 300, 260)])
def test_gemm_relu_mask_epilogue(m, n, k):
    """ACT_RELU_MASK: out holds the forward activations h on entry; out = (h > 0) ? A @ B^T : 0 (relu backward folded
    into the GEMM that produces dH) - one-shot and persistent 2-SM tensor-core kernels."""
    from hetu_61a7_b200.ops.linalg import gemm, ACT_RELU_MASK
    dy, w2, h = rnd(m, k), rnd(n, k, seed=1), torch.relu(rnd(m, n, seed=2))
    ref = (dy.double() @ w2.double().t()) * (h > 0).double()
    out = h.clone()
    res = gemm(dy, w2, False, True, act=ACT_RELU_MASK, out=out)
    assert res.data_ptr() == out.data_ptr()
    assert (out.double() - ref).abs().max().item() <= 2e-3 * ref.abs().max().item()
    assert (out[h <= 0] == 0).all()


@pytest.mark.gpu
@pytest.mark.parametrize('m,n,k,tB', [(2048, 2, 16384, False), (512, 5, 4099, False), (260, 8, 1000, True), (130, 3, 700, False)])
def test_gemm_skinny_weight_gradient(m, n, k, tB):
    """A^T B with a handful of output columns (MoE gate / logistic-regression weight gradient): 128-bit-load kernel
    (M % 4 == 0) and the scalar one, fp32 accumulation."""
    from hetu_61a7_b200.ops.linalg import gemm
    a = rnd(k, m)
    b = rnd(n, k, seed=1) if tB else rnd(k, n, seed=1)
    ref = a.double().t() @ (b.double().t() if tB else b.double())
    out = gemm(a, b, True, tB)
    assert (out.double() - ref).abs().max().item() <= 1e-4 * ref.abs().max().item() + 1e-4


@pytest.mark.gpu
@pytest.mark.parametrize('B,H,S,Sk,D,masked', [(2, 3, 128, 128, 64, False), (2, 2, 200, 333, 64, True), (3, 2, 77, 50, 32, True),
                                               (1, 2, 512, 512, 64, False), (1, 1, 130, 1000, 64, True)])
def test_flash_attention_forward_kernel(B, H, S, Sk, D, masked):
    """csrc/attention_sm100.cu (one tcgen05 kernel: QK^T -> online softmax -> PV) against fp64 torch: output and
    log-sum-exp, ragged tile edges, additive key mask, several key blocks."""
    from hetu_61a7_b200 import _lib as L
    q, k, v = rnd(B, H, S, D, scale=0.7), rnd(B, H, Sk, D, seed=1, scale=0.7), rnd(B, H, Sk, D, seed=2)
    mask = None
    if masked:
        mask = ((torch.rand(B, Sk, generator=torch.Generator().manual_seed(3)) > 0.75).float() * -1e4).to(dev())
    out = torch.empty_like(q)
    lse = torch.empty(B, H, S, device=dev())
    scale = 1.0 / np.sqrt(D)
    L.call('flash_attn_fwd', L.p(q), L.p(k), L.p(v), L.p(mask), L.p(out), L.p(lse), L.c_int(B), L.c_int(H), L.c_int(S),
           L.c_int(Sk), L.c_int(D), L.c_f32(scale), L.stream_ptr(dev()))
    s = q.double() @ k.double().transpose(-1, -2) * scale
    if mask is not None:
        s = s + mask.double().view(B, 1, 1, Sk)
    ref = torch.softmax(s, -1) @ v.double()
    assert (out.double() - ref).abs().max().item() <= 4e-3 * ref.abs().max().item()
    torch.testing.assert_close(lse.double(), torch.logsumexp(s, -1), rtol=2e-3, atol=2e-3)


@pytest.mark.gpu
def test_fused_attention_without_flash_kernel(L, monkeypatch):
    monkeypatch.setenv('HETU_FLASH_ATTN', '0')
    test_fused_attention_gpu(L)


@pytest.mark.gpu
@pytest.mark.parametrize('m,n,k,tA,tB', [(1024, 768, 2048, False, False), (900, 516, 1000, True, False), (512, 512, 4096, False, True)])
def test_gemm_3xtf32_is_fp32_faithful(m, n, k, tA, tB):
    """set_matmul_precision('3xtf32'): three tensor-core products of split operands reproduce the fp32 product to
    ~1e-6 of its magnitude (plain tf32: ~1e-3), with the fused epilogue and the accumulate path."""
    from hetu_61a7_b200.ops import linalg
    a = rnd(k, m) if tA else rnd(m, k)
    b = rnd(n, k, seed=1) if tB else rnd(k, n, seed=1)
    bias, c0 = rnd(n, seed=2), rnd(m, n, seed=3)
    A = a.double().t() if tA else a.double()
    Bm = b.double().t() if tB else b.double()
    ref = torch.relu(0.5 * (A @ Bm) + 2.0 * c0.double() + bias.double())
    linalg.set_matmul_precision('3xtf32')
    try:
        out = linalg.gemm(a, b, tA, tB, bias=bias, c_in=c0, alpha=0.5, beta=2.0, act=linalg.ACT_RELU)
        plain = linalg.gemm(a, b, tA, tB)
    finally:
        linalg.set_matmul_precision('tf32')
    scale = (A.abs() @ Bm.abs()).max().item()
    assert (out.double() - ref).abs().max().item() <= 3e-6 * scale
    assert (plain.double() - A @ Bm).abs().max().item() <= 3e-6 * scale
    tf = linalg.gemm(a, b, tA, tB)
    assert (tf.double() - A @ Bm).abs().max().item() > 2e-5 * scale          # the plain tf32 product is visibly coarser


@pytest.mark.gpu
@pytest.mark.parametrize('m,n,k,tA,tB', [(1024, 1024, 4096, True, False), (768, 3072, 4096, True, False), (768, 768, 2048, True, False),
                                         (520, 1000, 2100, False, True), (3072, 768, 4096, True, False)])
def test_gemm_stream_k_transformer_weight_gradients(m, n, k, tA, tB):
    """Few output tiles, K of a few thousand (weight gradients of transformer layers): the persistent 2-SM kernel cuts
    the k-blocks of all tiles into equal ranges per CTA pair and adds partial tiles with red.global."""
    from hetu_61a7_b200.ops.linalg import gemm
    a = rnd(k, m) if tA else rnd(m, k)
    b = rnd(n, k, seed=1) if tB else rnd(k, n, seed=1)
    bias = rnd(n, seed=2)
    A = a.double().t() if tA else a.double()
    Bm = b.double().t() if tB else b.double()
    ref = A @ Bm + bias.double()
    out = gemm(a, b, tA, tB, bias=bias)
    assert (out.double() - ref).abs().max().item() <= 2e-3 * ref.abs().max().item()


@pytest.mark.gpu
@pytest.mark.parametrize('m,n,k,tB', [(5000, 2048, 5, True), (70000, 256, 8, False), (3000, 1024, 16, True), (16384, 2048, 8, True)])
def test_gemm_tiny_k_rows_and_wide_skinny(m, n, k, tB):
    """[tokens, E] x [E, d] (input gradient of the MoE gate: K = E <= 16) and [tokens, d] x [d, E] with an E x d table
    above 48 KB of shared memory (8 experts, d = 2048)."""
    from hetu_61a7_b200.ops.linalg import gemm
    a, bias = rnd(m, k), rnd(n, seed=2)
    b = rnd(n, k, seed=1) if tB else rnd(k, n, seed=1)
    ref = a.double() @ (b.double().t() if tB else b.double()) + bias.double()
    out = gemm(a, b, False, tB, bias=bias)
    assert (out.double() - ref).abs().max().item() <= 1e-5 * ref.abs().max().item() + 1e-5
    x, w = rnd(4096, 2048, seed=3), rnd(2048, 8, seed=4, scale=0.05)       # skinny rows kernel, 64 KB table
    ref2 = x.double() @ w.double()
    out2 = gemm(x, w)
    assert (out2.double() - ref2).abs().max().item() <= 1e-4 * ref2.abs().max().item() + 1e-5


@pytest.mark.gpu
def test_graph_capture_with_release_matches(L):
    """Executor(dynamic_memory=True): values dropped at their last use during capture (the graph pool reuses their blocks)
    - same losses as the default plan, lower peak memory."""
    import gc
    import hetu as ht
    from hetu_61a7_b200.models import cnn
    rs = np.random.RandomState(3)
    xs = rs.randn(64, 3072).astype(np.float32)
    ys = np.eye(10, dtype=np.float32)[rs.randint(0, 10, 64)]
    out, snapshot = {}, None
    for dyn in (True, False):               # the planned run first: leftovers of it can only raise the other run's peak
        x, y_ = ht.Variable('x'), ht.Variable('y_')
        loss, _ = cnn.resnet18(x, y_)
        train = ht.optim.SGDOptimizer(0.01).minimize(loss)
        ex = ht.Executor([loss, train], ctx=ht.gpu(0), seed=4, dynamic_memory=dyn)
        if snapshot is None:                # same initial weights for both (node ids, hence init seeds, differ per build)
            snapshot = {n: ex.config.placeholder_to_arr_map[p].detach().cpu().numpy().copy() for n, p in ex._named_params().items()}
        else:
            ex.load_dict(snapshot)
        torch.cuda.synchronize()
        torch.cuda.reset_peak_memory_stats()
        base = torch.cuda.memory_allocated()
        losses = [float(ex.run(feed_dict={x: xs, y_: ys})[0].asnumpy()) for _ in range(6)]
        out[dyn] = (losses, torch.cuda.max_memory_allocated() - base)
        assert next(iter(ex.subexecutor.values()))._graph is not None
        del ex
        gc.collect()
    np.testing.assert_allclose(out[True][0], out[False][0], rtol=5e-3, atol=1e-4)
    assert out[True][1] < out[False][1], out
