"""Classical-ML library (hetuml): every model family against a numpy / scikit-learn oracle on
synthetic data (pattern: reference third_party/HetuML/examples/standalone_ml/*.py - fit, save &
load, predict, evaluate)."""
import os

import numpy as np
import pytest
import scipy.sparse as sp

from hetuml.data import COOMatrix, Corpus, CSRMatrix, Dataset
from hetuml.decomposition import LDA, NMF
from hetuml.ensemble import GBDT, RandomForest, RanomBoostedForest
from hetuml.factorization_machines import FM
from hetuml.linear import SVM, LinearRegression, LogisticRegression
from hetuml.naive_bayes import NaiveBayes
from hetuml.neighbors import KNN


def _binary(n=4000, d=40, seed=0, density=0.3):
    rng = np.random.RandomState(seed)
    X = sp.random(n, d, density=density, random_state=rng, format='csr', dtype=np.float32)
    w = rng.standard_normal(d)
    y = (X @ w + 0.1 * rng.standard_normal(n) > 0).astype(np.float32)
    return X, y, w


def test_dataset_libsvm_roundtrip(tmp_path):
    X, y, _ = _binary(50, 7)
    p = tmp_path / 'a.libsvm'
    with open(p, 'w') as f:
        for i in range(X.shape[0]):
            r = X.getrow(i)
            f.write('%d %s\n' % (1 if y[i] > 0 else -1, ' '.join('%d:%.6f' % (c, v) for c, v in zip(r.indices, r.data))))
    ds = Dataset.from_file(str(p), neg_y=True)
    assert ds.get_num_instances() == 50 and ds.get_max_dim() <= 7
    assert set(np.unique(ds.y)) <= {-1.0, 1.0}
    np.testing.assert_allclose(ds.X.data, X.data, atol=1e-6)
    half = Dataset.from_file(str(p), rank=1, total_ranks=2)
    assert half.get_num_instances() == 25
    # unsorted rows are canonicalised
    m = CSRMatrix([0, 2], [3, 1], [1.0, 2.0])
    assert list(m.indices) == [1, 3] and list(m.data) == [2.0, 1.0]


def test_logistic_regression_matches_numpy_sgd(tmp_path):
    X, y, _ = _binary(2000, 20)
    ypm = np.where(y > 0, 1.0, -1.0).astype(np.float32)
    lr = LogisticRegression(num_epoch=3, batch_size=100, learning_rate=0.5, l2_reg=0.01)
    lr.fit((X, ypm))
    # oracle: the same mini-batch schedule in numpy (regularisation on touched coordinates only)
    Xd = X.toarray().astype(np.float64)
    w = np.zeros(20)
    for _ in range(3):
        for b in range(0, 2000, 100):
            xb, yb = Xd[b:b + 100], ypm[b:b + 100]
            c = -yb / (1 + np.exp(yb * (xb @ w)))
            touched = (xb != 0).any(axis=0)
            g = xb.T @ c / len(yb) + 0.01 * w
            w[touched] -= 0.5 * g[touched]
    np.testing.assert_allclose(lr.weights, w, rtol=2e-3, atol=2e-4)
    m = lr.evaluate((X, ypm), 'log-loss,error,precision')
    assert m['error'] < 0.15 and abs(m['error'] + m['precision'] - 1) < 1e-9
    path = str(tmp_path / 'lr.model')
    lr.save_model(path)
    again = LogisticRegression()
    again.load_model(path)
    np.testing.assert_array_equal(again.predict_score(X), lr.predict_score(X))
    assert again['learning_rate'] == 0.5
    proba = again.predict_proba(X)
    assert proba.min() >= 0 and proba.max() <= 1


def test_svm_and_linear_regression():
    X, y, w = _binary(3000, 15)
    ypm = np.where(y > 0, 1.0, -1.0).astype(np.float32)
    svm = SVM(num_epoch=10, batch_size=50, learning_rate=0.2)
    svm.fit((X, ypm))
    assert svm.evaluate((X, ypm), ['error'])['error'] < 0.12
    target = (X @ w).astype(np.float32)
    reg = LinearRegression(num_epoch=60, batch_size=50, learning_rate=0.5)
    reg.fit((X, target))
    assert reg.evaluate((X, target), 'mse')['mse'] < 0.05 * float(np.var(target))


def test_fm_learns_pairwise_interaction():
    rng = np.random.RandomState(1)
    n, d = 6000, 12
    Xd = (rng.rand(n, d) < 0.4).astype(np.float32)
    y = np.where(Xd[:, 0] * Xd[:, 1] + Xd[:, 2] * Xd[:, 3] > 0.5, 1.0, -1.0).astype(np.float32)   # no linear separator
    fm = FM(embedding_dim=8, num_epoch=60, batch_size=100, learning_rate=0.5, aux_args={'seed': 3})
    fm.fit((sp.csr_matrix(Xd), y))
    lin = LogisticRegression(num_epoch=60, batch_size=100, learning_rate=0.5)
    lin.fit((sp.csr_matrix(Xd), y))
    e_fm = fm.evaluate((sp.csr_matrix(Xd), y), 'error')['error']
    e_lin = lin.evaluate((sp.csr_matrix(Xd), y), 'error')['error']
    assert e_fm < 0.05 and e_fm < e_lin - 0.03, (e_fm, e_lin)
    # the closed form of the score
    w, v = fm._arrays['w'], fm._arrays['v']
    x = Xd[:5].astype(np.float64)
    ref = x @ w + 0.5 * (((x @ v) ** 2).sum(1) - ((x ** 2) @ (v.astype(np.float64) ** 2)).sum(1))
    np.testing.assert_allclose(fm.predict_score(sp.csr_matrix(Xd[:5])), ref, rtol=1e-4, atol=1e-4)


def test_gbdt_binary_against_sklearn_and_save_load(tmp_path):
    from sklearn.ensemble import GradientBoostingClassifier
    rng = np.random.RandomState(0)
    n, d = 5000, 10
    Xd = rng.standard_normal((n, d)).astype(np.float32)
    Xd[rng.rand(n, d) < 0.5] = 0                                         # genuinely sparse
    y = ((Xd[:, 0] > 0.3) ^ (Xd[:, 1] < -0.2) | (Xd[:, 2] * Xd[:, 3] > 0.5)).astype(np.float32)
    X = sp.csr_matrix(Xd)
    gb = GBDT(max_depth=4, num_round=30, learning_rate=0.3, num_split=32, min_node_ins=20)
    gb.fit((X, y))
    curve = gb.train_loss_curve()
    assert len(curve) == 30 and np.all(np.diff(curve) < 1e-9), curve      # boosting decreases the training loss
    m = gb.evaluate((X, y), 'log-loss,error')
    assert abs(m['log-loss'] - curve[-1]) < 1e-4
    sk = GradientBoostingClassifier(max_depth=4, n_estimators=30, learning_rate=0.3).fit(Xd, y)
    sk_err = 1 - sk.score(Xd, y)
    assert m['error'] < sk_err + 0.03, (m, sk_err)
    path = str(tmp_path / 'gbdt.model')
    gb.save_model(path)
    again = GBDT()
    again.load_model(path)
    np.testing.assert_allclose(again.predict_score(X), gb.predict_score(X), rtol=1e-6, atol=1e-6)
    assert again['max_depth'] == 4 and again.num_trees() == 30
    # leaf-wise growth with a node budget
    lw = GBDT(max_depth=8, max_node_num=15, leafwise=True, num_round=10, learning_rate=0.3, num_split=32, min_node_ins=20)
    lw.fit((X, y))
    assert lw.evaluate((X, y), 'error')['error'] < 0.2


def test_gbdt_regression_multiclass_and_forest():
    rng = np.random.RandomState(2)
    n, d = 4000, 8
    Xd = rng.rand(n, d).astype(np.float32)
    X = sp.csr_matrix(Xd)
    target = (np.sin(4 * Xd[:, 0]) + Xd[:, 1] * Xd[:, 2]).astype(np.float32)
    reg = GBDT(is_regression=True, max_depth=4, num_round=40, learning_rate=0.3, num_split=32, min_node_ins=20)
    reg.fit((X, target))
    assert reg.evaluate((X, target), 'mse')['mse'] < 0.05 * float(np.var(target))
    lab = (Xd[:, 0] > 0.5).astype(np.int64) + (Xd[:, 1] > 0.5).astype(np.int64)     # 3 classes
    for multi_tree in (False, True):
        mc = GBDT(num_label=3, max_depth=3, num_round=15, learning_rate=0.5, num_split=16, min_node_ins=20,
                  multi_tree=multi_tree)
        mc.fit((X, lab.astype(np.float32)))
        s = mc.predict_score(X)
        assert s.shape == (n, 3)
        ev = mc.evaluate((X, lab.astype(np.float32)), 'cross-entropy,precision')
        assert ev['precision'] > 0.97, ev
        assert mc.num_trees() == (45 if multi_tree else 15)
        np.testing.assert_allclose(mc.predict_proba(X).sum(1), 1, rtol=1e-5)
    assert RanomBoostedForest is RandomForest
    rf = RandomForest(max_depth=5, num_round=30, num_split=16, min_node_ins=20, ins_sp_ratio=0.5, feat_sp_ratio=0.7)
    y = (Xd[:, 0] + Xd[:, 1] > 1).astype(np.float32)
    rf.fit((X, y))
    assert rf.evaluate((X, y), 'error')['error'] < 0.1


def test_knn_matches_sklearn():
    from sklearn.neighbors import KNeighborsClassifier
    rng = np.random.RandomState(3)
    Xl = sp.random(600, 30, density=0.3, random_state=rng, format='csr', dtype=np.float32)
    yl = rng.randint(0, 4, 600).astype(np.float32)
    Xq = sp.random(80, 30, density=0.3, random_state=rng, format='csr', dtype=np.float32)
    knn = KNN(num_label=4, num_neighbor=7)
    votes = knn.predict_score((Xl, yl), Xq)
    assert votes.shape == (80, 4) and np.all(votes.sum(1) == 7)
    sk = KNeighborsClassifier(n_neighbors=7, algorithm='brute').fit(Xl.toarray(), yl)
    ref = sk.predict_proba(Xq.toarray()) * 7
    assert np.mean(np.abs(votes - ref).sum(1) == 0) > 0.95              # ties in distance may differ
    assert set(knn.evaluate((Xl, yl), (Xl, yl), 'precision')) == {'precision'}


def test_naive_bayes_matches_formula(tmp_path):
    rng = np.random.RandomState(4)
    n, d, K = 3000, 6, 3
    y = rng.randint(0, K, n)
    means = rng.standard_normal((K, d)) * 2 + 4
    Xd = (means[y] + rng.standard_normal((n, d))).astype(np.float32)
    Xd[rng.rand(n, d) < 0.3] = 0
    X = sp.csr_matrix(Xd)
    nb = NaiveBayes(num_label=K)
    nb.fit((X, y.astype(np.float32)))
    s = nb.predict_score(X)
    # oracle: Gaussian over the stored entries of each (feature, class)
    ref = np.zeros((n, K))
    for k in range(K):
        ref[:, k] = np.log((y == k).mean())
        for f in range(d):
            col = Xd[:, f]
            sel = (y == k) & (col != 0)
            mu, var = col[sel].mean(), col[sel].var()
            lp = -0.5 * np.log(2 * np.pi * var) - (col - mu) ** 2 / (2 * var)
            ref[:, k] += np.where(col != 0, np.maximum(lp, np.log(1e-8)), 0.0)
    np.testing.assert_allclose(s, ref, rtol=2e-3, atol=2e-2)
    assert nb.evaluate((X, y.astype(np.float32)), 'precision')['precision'] > 0.9
    path = str(tmp_path / 'nb.model')
    nb.save_model(path)
    again = NaiveBayes(num_label=K)
    again.load_model(path)
    np.testing.assert_array_equal(again.predict_score(X), s)


def test_nmf_reconstructs_low_rank_matrix():
    rng = np.random.RandomState(5)
    P, Q = rng.rand(60, 4), rng.rand(50, 4)
    R = P @ Q.T
    mask = rng.rand(60, 50) < 0.5
    r, c = np.nonzero(mask)
    train = COOMatrix.from_data((r, c, R[r, c]))
    mf = NMF(embedding_dim=4, num_epoch=300, learning_rate=0.05, l2_reg=1e-4)
    mf.fit(train)
    assert mf.train_rmse[-1] < 0.05 and mf.train_rmse[-1] < mf.train_rmse[0]
    r2, c2 = np.nonzero(~mask)
    pred = mf.predict(COOMatrix.from_data((r2, c2, R[r2, c2])))
    assert np.sqrt(np.mean((pred - R[r2, c2]) ** 2)) < 0.12                # held-out entries
    assert mf._arrays['P'].min() >= 0 and mf._arrays['Q'].min() >= 0


def test_lda_recovers_planted_topics(tmp_path):
    rng = np.random.RandomState(6)
    V, K = 40, 4
    docs = []
    for d in range(200):
        k = d % K                                                         # each document draws from one block of 10 words
        docs.append(list(rng.randint(k * 10, (k + 1) * 10, size=60)))
    corpus = Corpus.from_data(V, docs)
    assert corpus.get_num_docs() == 200 and corpus.get_num_tokens() == 12000
    lda = LDA(num_words=V, num_topics=K, num_iters=60, alpha=0.1, beta=0.01)
    lda.fit(corpus)
    tw = lda.topic_word()                                                 # [K, V]
    blocks = tw.reshape(K, K, 10).sum(axis=2)                             # topic mass per word block
    assert np.all(blocks.max(axis=1) > 0.9), blocks
    assert sorted(blocks.argmax(axis=1)) == list(range(K))
    z = lda.predict(Corpus.from_data(V, docs[:8]))
    assert z.shape == (480,)
    first = z[:60]
    assert np.bincount(first, minlength=K).max() > 50                     # one topic explains a document
    p = tmp_path / 'corpus.txt'
    with open(p, 'w') as f:
        for i, d in enumerate(docs[:5]):
            f.write('%d %s\n' % (i, ' '.join(map(str, d))))
    assert Corpus.from_file(V, str(p)).get_num_docs() == 5


def test_parallel_models_over_parameter_server(tmp_path):
    """2 workers + 1 server + scheduler on loopback: data-parallel GBDT grows the same trees on both
    workers, naive Bayes equals single-process training on the union, LR / LDA agree across workers."""
    import socket
    import subprocess
    import sys
    import time
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    from conftest import free_port_block
    port = free_port_block(2)
    sched = '127.0.0.1:%d' % port
    env = dict(os.environ, PYTHONPATH=root + os.pathsep + os.environ.get('PYTHONPATH', ''), OMP_NUM_THREADS='2')
    role = ("import sys; sys.path.insert(0, %r); from hetuml.cluster import Cluster; c = Cluster(%r, 1, 2, %%r); "
            "c.start_server('double'); c.join()") % (root, sched)
    procs = []
    try:
        procs.append(subprocess.Popen([sys.executable, '-c', role % 'scheduler'], env=env))
        time.sleep(0.5)
        procs.append(subprocess.Popen([sys.executable, '-c', role % 'server'], env=env))
        workers = [subprocess.Popen([sys.executable, os.path.join(root, 'tests', 'ml_parallel_worker.py'), sched, str(tmp_path)],
                                    env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for _ in range(2)]
        outs = [w.communicate(timeout=600)[0] for w in workers]
        for w, o in zip(workers, outs):
            assert w.returncode == 0 and 'OK' in o, o[-3000:]
    finally:
        for p in procs:
            p.terminate()
    g0, g1 = np.load(tmp_path / 'gbdt_0.npy'), np.load(tmp_path / 'gbdt_1.npy')
    np.testing.assert_array_equal(g0, g1)
    np.testing.assert_array_equal(np.load(tmp_path / 'lr_0.npy'), np.load(tmp_path / 'lr_1.npy'))
    np.testing.assert_allclose(np.load(tmp_path / 'nb_0.npy'), np.load(tmp_path / 'nb_single.npy'), rtol=1e-4, atol=1e-3)
    np.testing.assert_array_equal(np.load(tmp_path / 'nb_0.npy'), np.load(tmp_path / 'nb_1.npy'))
    t0, t1 = np.load(tmp_path / 'lda_0.npy'), np.load(tmp_path / 'lda_1.npy')
    np.testing.assert_array_equal(t0, t1)
    blocks = t0.reshape(3, 3, 10).sum(axis=2)
    assert np.all(blocks.max(axis=1) > 0.85), blocks


@pytest.mark.gpu
def test_knn_gpu_distance_gemm_matches_cpu():
    """KNN(device='gpu'): distances from the tcgen05 GEMM, same votes as the sparse CPU core
    (rows whose k-th and (k+1)-th neighbour are within tf32 rounding of each other may differ)."""
    rng = np.random.RandomState(7)
    Xl = sp.random(3000, 200, density=0.2, random_state=rng, format='csr', dtype=np.float32)
    yl = rng.randint(0, 5, 3000).astype(np.float32)
    Xq = sp.random(500, 200, density=0.2, random_state=rng, format='csr', dtype=np.float32)
    cpu = KNN(num_label=5, num_neighbor=9, device='cpu').predict_score((Xl, yl), Xq)
    gpu = KNN(num_label=5, num_neighbor=9, device='gpu').predict_score((Xl, yl), Xq)
    assert gpu.shape == cpu.shape and np.all(gpu.sum(1) == 9)
    assert np.mean(np.abs(gpu - cpu).sum(1) == 0) > 0.97
    small_l = sp.random(400, 30, density=0.3, random_state=rng, format='csr', dtype=np.float32)
    small_q = sp.random(50, 30, density=0.3, random_state=rng, format='csr', dtype=np.float32)
    ys = rng.randint(0, 3, 400).astype(np.float32)
    a = KNN(num_label=3, num_neighbor=5, device='cpu').predict_score((small_l, ys), small_q)
    b = KNN(num_label=3, num_neighbor=5, device='gpu').predict_score((small_l, ys), small_q)      # exact fp32 path (K <= 64)
    assert np.mean(np.abs(a - b).sum(1) == 0) > 0.97


def _gbdt_level_trainer_matches_node_trainer(dev):
    from hetu_61a7_b200.ml import _gbdt_gpu
    rng = np.random.RandomState(11)
    n, d = 3000, 12
    Xd = rng.standard_normal((n, d)).astype(np.float32)
    Xd[rng.rand(n, d) < 0.5] = 0
    y = ((Xd[:, 0] > 0.3) ^ (Xd[:, 1] < -0.2) | (Xd[:, 2] * Xd[:, 3] > 0.5)).astype(np.float32)
    X = sp.csr_matrix(Xd)
    kw = dict(max_depth=4, num_round=8, learning_rate=0.3, num_split=16, min_node_ins=20, reg_lambda=1.0)
    ref = GBDT(device='cpu', **kw)
    ref.fit((X, y))
    lvl = GBDT(device='cpu', **kw)
    ds = Dataset.from_data((X, y))
    _gbdt_gpu.fit(lvl, ds, ds.get_max_dim(), dev=dev)
    # float32 histograms vs double: same trees unless two splits tie within rounding
    np.testing.assert_allclose(lvl.train_loss_curve(), ref.train_loss_curve(), rtol=2e-3, atol=2e-3)
    a, b = lvl.predict_score(X), ref.predict_score(X)
    # (on the GPU the float atomics also reorder the sums, so a near-tie may pick the other split in some tree)
    assert np.mean(np.abs(a - b) < 1e-3) > (0.98 if dev == 'cpu' else 0.85), np.abs(a - b).max()
    assert abs(lvl.evaluate((X, y), 'log-loss')['log-loss'] - ref.evaluate((X, y), 'log-loss')['log-loss']) < 5e-3
    assert lvl.num_trees() == 8 and lvl.evaluate((X, y), 'error')['error'] < 0.12
    # regression, sampling, multi-class with one tree per class
    target = (np.sin(2 * Xd[:, 0]) + Xd[:, 1]).astype(np.float32)
    reg = GBDT(device='cpu', is_regression=True, max_depth=4, num_round=20, learning_rate=0.3, num_split=16, min_node_ins=20,
               ins_sp_ratio=0.8, feat_sp_ratio=0.9)
    dr = Dataset.from_data((X, target))
    curve = _gbdt_gpu.fit(reg, dr, dr.get_max_dim(), dev=dev)
    assert curve[-1] < 0.3 * curve[0] and reg.evaluate((X, target), 'mse')['mse'] < 0.3 * float(np.var(target))
    lab = ((Xd[:, 0] > 0).astype(np.int64) + (Xd[:, 1] > 0).astype(np.int64)).astype(np.float32)
    mc = GBDT(device='cpu', num_label=3, multi_tree=True, max_depth=3, num_round=10, learning_rate=0.5, num_split=16, min_node_ins=20)
    dm = Dataset.from_data((X, lab))
    _gbdt_gpu.fit(mc, dm, dm.get_max_dim(), dev=dev)
    assert mc.num_trees() == 30 and mc.evaluate((X, lab), 'precision')['precision'] > 0.95


def test_gbdt_level_trainer_host_oracle():
    """The level-at-a-time trainer (device GBDT) with torch stand-ins for its two kernels agrees
    with the node-at-a-time native trainer."""
    _gbdt_level_trainer_matches_node_trainer('cpu')
