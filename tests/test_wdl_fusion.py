"""Whole-step fusion of the Wide&Deep graph (graph/fusion_wdl.py, csrc/wdl_tower.cu).

CPU: the graph rewrite itself - the fused node's torch restatement must train exactly like the node-by-node
graph.  GPU: the cluster kernel against the fp32 torch restatement of the same step (tf32 tolerance), and the
fused executor against the unfused one over several steps.
Reference model: examples/ctr/models/wdl_criteo.py:8-42.
"""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

_WORKER = r'''
import os, sys, json
sys.path.insert(0, %(root)r)
import numpy as np
import hetu_61a7_b200 as ht
from hetu_61a7_b200.models import ctr
dev = %(dev)r
ctx = ht.gpu(0) if dev == 'gpu' else ht.cpu(0)
B, vocab = 128, 20000
pool = [ht.data.criteo_synthetic(B, vocab=vocab, seed=i) for i in range(4)]
dense = ht.dataloader_op([ht.Dataloader(np.concatenate([p[0] for p in pool]), B, 'train')])
sparse = ht.dataloader_op([ht.Dataloader(np.concatenate([p[1] for p in pool]), B, 'train', dtype=np.int64)])
y_ = ht.dataloader_op([ht.Dataloader(np.concatenate([p[2] for p in pool]), B, 'train')])
loss, pred, _, train_op = ctr.wdl_criteo(dense, sparse, y_, feature_dimension=vocab, learning_rate=0.5, ctx=ctx)
ex = ht.Executor({'train': [loss, pred, train_op]}, ctx=ctx, seed=123)
losses, preds = [], []
for i in range(8):
    r = ex.run('train')
    losses.append(float(r[0].asnumpy().reshape(-1)[0]))
    preds.append(r[1].asnumpy().reshape(-1).tolist())
params = {n.name: np.asarray(ex.config.placeholder_to_arr_map[n].detach().cpu()).reshape(-1)[:4096].tolist()
          for n in ex.param_nodes}
print('RESULT' + json.dumps({'fused': ex.fused_wdl, 'losses': losses, 'preds': preds, 'params': params}))
'''


def _run(dev, fuse):
    import json
    env = dict(os.environ, HETU_FUSE_WDL='1' if fuse else '0')
    out = subprocess.run([sys.executable, '-c', _WORKER % {'root': ROOT, 'dev': dev}], env=env, capture_output=True,
                         text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-3000:]
    line = [l for l in out.stdout.splitlines() if l.startswith('RESULT')][-1]
    return json.loads(line[len('RESULT'):])


def test_rewrite_trains_like_the_unfused_graph_cpu():
    a, b = _run('cpu', True), _run('cpu', False)
    assert a['fused'] == 1 and b['fused'] == 0
    np.testing.assert_allclose(a['losses'], b['losses'], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(a['preds'], b['preds'], rtol=1e-5, atol=1e-6)
    for k in b['params']:
        np.testing.assert_allclose(a['params'][k], b['params'][k], rtol=1e-5, atol=1e-7)


def test_other_graphs_are_left_alone_cpu():
    """Adam instead of SGD, or an extra consumer of an intermediate: no rewrite."""
    sys.path.insert(0, ROOT)
    import hetu_61a7_b200 as ht
    from hetu_61a7_b200 import initializers as init
    B, vocab = 16, 100
    d, s, l = ht.data.criteo_synthetic(B, vocab=vocab, seed=0)
    x = ht.dataloader_op([ht.Dataloader(d, B, 'train')])
    ids = ht.dataloader_op([ht.Dataloader(s, B, 'train', dtype=np.int64)])
    y_ = ht.dataloader_op([ht.Dataloader(l, B, 'train')])
    E = init.random_normal([vocab, 128], stddev=0.01, name='E', ctx=ht.cpu(0))
    sp = ht.array_reshape_op(ht.embedding_lookup_op(E, ids, ctx=ht.cpu(0)), (-1, 26 * 128))
    W1 = init.random_normal([13, 256], stddev=0.01, name='W1')
    W2 = init.random_normal([256, 256], stddev=0.01, name='W2')
    W3 = init.random_normal([256, 256], stddev=0.01, name='W3')
    W4 = init.random_normal([256 + 26 * 128, 1], stddev=0.01, name='W4')
    y3 = ht.matmul_op(ht.relu_op(ht.matmul_op(ht.relu_op(ht.matmul_op(x, W1)), W2)), W3)
    y = ht.sigmoid_op(ht.matmul_op(ht.concat_op(sp, y3, axis=1), W4))
    loss = ht.reduce_mean_op(ht.binarycrossentropy_op(y, y_), [0])
    train = ht.optim.AdamOptimizer(learning_rate=0.01).minimize(loss)
    ex = ht.Executor({'train': [loss, train]}, ctx=ht.cpu(0), seed=1)
    assert ex.fused_wdl == 0
    ex.run('train')


@pytest.mark.gpu
def test_tower_kernel_matches_fp32_torch_step():
    sys.path.insert(0, ROOT)
    from hetu_61a7_b200.graph import fusion_wdl as fw
    dev = torch.device('cuda', 0)
    g = torch.Generator().manual_seed(0)
    B, nd, hid, fields, width, vocab = 128, 13, 256, 26, 128, 50000
    m = dict(nd=nd, width=width, emb_cols=fields * width, fields=fields, x=None, ids=None, labels=None, E=None,
             W1=None, W2=None, W3=None, W4=None)
    cls = fw._make_op_class()
    op = cls.__new__(cls)
    op.m, op._bufs, op._dp, op.last, op.fused_launches = m, None, None, None, 0
    x = torch.rand(B, nd, generator=g)
    ids = torch.randint(0, vocab, (B, fields), generator=g)
    ids[5] = ids[4]                                    # duplicate rows
    labels = (torch.rand(B, 1, generator=g) < 0.3).float()
    E = torch.randn(vocab, width, generator=g) * 0.05
    Ws = [torch.randn(nd, hid, generator=g) * 0.3, torch.randn(hid, hid, generator=g) * 0.08,
          torch.randn(hid, hid, generator=g) * 0.08, torch.randn(fields * width + hid, 1, generator=g) * 0.05]
    loss_r, pred_r = op._torch_step(x.double(), ids, labels.double(), E.double(), *[w.double() for w in Ws])
    ref = op.last
    to = lambda t: t.to(dev)                           # noqa: E731
    loss_k, pred_k = op.compute([to(x), to(ids), to(labels), to(E), *[to(w) for w in Ws]])
    torch.cuda.synchronize()
    got = op.last
    assert not got['torch_path']
    assert abs(loss_k.item() - loss_r.item()) < 2e-3 * abs(loss_r.item())
    torch.testing.assert_close(pred_k.cpu().double(), pred_r, rtol=5e-3, atol=5e-4)
    torch.testing.assert_close(got['dz'].cpu().double(), ref['dz'], rtol=5e-3, atol=1e-6)
    for name, a, b in zip(('dW1', 'dW2', 'dW3', 'dW4'), got['grads'], ref['grads']):
        # tf32 keeps 10 mantissa bits of each operand: dW2 = h1^T dh2 is a sum of 128 products of mixed sign whose
        # magnitude is far above the result's, so its tolerance is set against the largest entry, not per element
        scale = b.abs().max().item()
        err = (a.cpu().double() - b).abs().max().item()
        assert err <= (1.5e-2 if name == 'dW2' else 4e-3) * scale, (name, err, scale)


@pytest.mark.gpu
def test_fused_executor_trains_like_the_unfused_one_gpu():
    a, b = _run('gpu', True), _run('gpu', False)
    assert a['fused'] == 1 and b['fused'] == 0
    np.testing.assert_allclose(a['losses'], b['losses'], rtol=2e-3)
    np.testing.assert_allclose(a['preds'], b['preds'], rtol=1e-2, atol=2e-3)
    for k in b['params']:
        pa, pb = np.asarray(a['params'][k]), np.asarray(b['params'][k])
        assert np.abs(pa - pb).max() <= 2e-2 * np.abs(pb).max() + 1e-5, k
