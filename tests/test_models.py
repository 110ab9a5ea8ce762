"""Model zoo smoke / convergence tests on the CPU executor (reference: examples/*/ scripts are the
upstream "tests" for these models; here each model must build, differentiate and reduce its loss)."""
import numpy as np
import pytest

import hetu as ht
from hetu_61a7_b200.models import cnn, gnn, nlp, rec

rs = np.random.RandomState(0)


@pytest.mark.parametrize('name', ['logreg', 'mlp', 'cnn_3_layers', 'lenet', 'resnet18', 'rnn', 'lstm'])
def test_cnn_zoo(name):
    d = 3072 if name in ('mlp', 'resnet18') else 784
    x, y_ = ht.Variable('x'), ht.Variable('y_')
    loss, y = cnn.MODELS[name](x, y_)
    train = ht.optim.SGDOptimizer(0.01).minimize(loss)
    ex = ht.Executor([loss, y, train], ctx=ht.cpu(0), seed=3)
    X = rs.randn(4, d).astype(np.float32)
    Y = np.eye(10, dtype=np.float32)[rs.randint(0, 10, 4)]
    out = [ex.run(feed_dict={x: X, y_: Y}) for _ in range(3)]
    assert out[0][1].asnumpy().shape == (4, 10)
    losses = [float(o[0].asnumpy()) for o in out]
    assert np.isfinite(losses).all()
    if name in ('logreg', 'lenet', 'resnet18'):
        assert losses[-1] < losses[0]


@pytest.mark.parametrize('fused', [True, False])
def test_bert_pretraining_tiny(fused):
    cfg = nlp.BertConfig(vocab_size=100, hidden_size=32, num_hidden_layers=2, num_attention_heads=4,
                         intermediate_size=64, max_position_embeddings=16, batch_size=2, seq_len=8,
                         hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, fused_attention=fused)
    ids, tt, am, mlm, nsp = [ht.Variable(n) for n in ('ids', 'tt', 'am', 'mlm', 'nsp')]
    _, _, l1, l2 = nlp.BertForPreTraining(cfg)(ids, tt, am, mlm, nsp)
    loss = l1 + l2
    train = ht.optim.AdamOptimizer(1e-3).minimize(loss)
    ex = ht.Executor([loss, train], ctx=ht.cpu(0), seed=1)
    am_v = np.ones((2, 8), np.float32)
    am_v[1, 6:] = 0
    fd = {ids: rs.randint(0, 100, (2, 8)).astype(np.float32), tt: np.zeros((2, 8), np.float32), am: am_v,
          mlm: rs.randint(0, 100, (2, 8)).astype(np.float32), nsp: np.array([0, 1], np.float32)}
    losses = [float(ex.run(feed_dict=fd)[0].asnumpy()) for _ in range(6)]
    assert losses[-1] < losses[0] - 0.1, losses


def test_transformer_tiny():
    hp = nlp.TransformerHParams(vocab_size=50, d_model=32, d_ff=64, num_blocks=1, num_heads=4, maxlen1=6, maxlen2=7,
                                dropout_rate=0.0, batch_size=2)
    xs, yi, yo = ht.Variable('xs'), ht.Variable('yi'), ht.Variable('yo')
    loss, logits = nlp.Transformer(hp)(xs, yi, yo)
    train = ht.optim.AdamOptimizer(1e-3).minimize(loss)
    ex = ht.Executor([loss, train], ctx=ht.cpu(0))
    fd = {xs: rs.randint(1, 50, (2, 6)).astype(np.float32), yi: rs.randint(1, 50, (2, 7)).astype(np.float32),
          yo: rs.randint(1, 50, (2, 7)).astype(np.float32)}
    losses = [float(ex.run(feed_dict=fd)[0].asnumpy()) for _ in range(6)]
    assert losses[-1] < losses[0] - 0.1, losses


def test_ncf_and_gnn():
    import scipy.sparse as sp
    u, i, y_ = ht.Variable('u'), ht.Variable('i'), ht.Variable('y_')
    loss, y, train = rec.neural_mf(u, i, y_, 50, 40)
    ex = ht.Executor([loss, y, train], ctx=ht.cpu(0))
    out = ex.run(feed_dict={u: rs.randint(0, 50, 64).astype(np.float32), i: rs.randint(0, 40, 64).astype(np.float32),
                            y_: rs.randint(0, 2, (64, 1)).astype(np.float32)})
    assert out[1].asnumpy().shape == (64, 1)
    N, F, C = 80, 16, 4
    lab = rs.randint(0, C, N)
    A = sp.coo_matrix(((rs.rand(N, N) < np.where(lab[:, None] == lab[None, :], 0.2, 0.01)) | np.eye(N, dtype=bool)).astype(np.float32))
    A = ((A + A.T) > 0).astype(np.float32)
    d = np.asarray(A.sum(1)).ravel()
    An = (sp.diags(d ** -0.5) @ A @ sp.diags(d ** -0.5)).tocoo()
    X = (np.eye(C)[lab] @ rs.randn(C, F) + 0.5 * rs.randn(N, F)).astype(np.float32)
    for layer in (gnn.GCN, gnn.SageConv):
        (feat, yy, adj, mask), (loss, y, train) = gnn.dense_model(F, 8, C, lr=0.5, layer=layer)
        ex = ht.Executor([loss, y, train], ctx=ht.cpu(0))
        mp = ht.sparse_array(An.data.astype(np.float32), (An.row, An.col), shape=(N, N), ctx=ht.cpu(0))
        fd = {feat: X, yy: np.eye(C, dtype=np.float32)[lab], adj: mp, mask: np.ones(N, np.float32)}
        losses = [float(ex.run(feed_dict=fd)[0].asnumpy()) for _ in range(25)]
        assert losses[-1] < losses[0] * 0.8, (layer.__name__, losses[0], losses[-1])
