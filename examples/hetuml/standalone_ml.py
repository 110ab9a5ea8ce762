"""Standalone (single-process) runs of the classical-ML models: fit, save & load, predict, evaluate.

Mirrors reference third_party/HetuML/examples/standalone_ml/{lr,fm,gbdt,knn,mf,lda}.py.  Pass
``--train path.libsvm [--valid path.libsvm]`` (e.g. a9a) or run without data to use a synthetic set.

    python examples/hetuml/standalone_ml.py --model gbdt
"""
import argparse
import os
import sys
import tempfile

import numpy as np
import scipy.sparse as sp

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from hetuml.data import COOMatrix, Corpus, Dataset          # noqa: E402
from hetuml.decomposition import LDA, NMF                   # noqa: E402
from hetuml.ensemble import GBDT, RandomForest              # noqa: E402
from hetuml.factorization_machines import FM                # noqa: E402
from hetuml.linear import SVM, LinearRegression, LogisticRegression   # noqa: E402
from hetuml.naive_bayes import NaiveBayes                   # noqa: E402
from hetuml.neighbors import KNN                            # noqa: E402


def synthetic(n=20000, d=123, seed=0):
    rng = np.random.RandomState(seed)
    X = sp.random(n, d, density=0.11, random_state=rng, format='csr', dtype=np.float32)
    X.data[:] = 1.0                                          # a9a-like binary indicators
    w = rng.standard_normal(d)
    y = (X @ w + 0.3 * rng.standard_normal(n) > 0).astype(np.float32)
    return X, y


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--model', default='lr', choices=['lr', 'svm', 'linreg', 'fm', 'gbdt', 'rf', 'knn', 'nb', 'mf', 'lda'])
    ap.add_argument('--train')
    ap.add_argument('--valid')
    args = ap.parse_args()
    if args.model == 'mf':
        rng = np.random.RandomState(0)
        R = rng.rand(300, 6) @ rng.rand(200, 6).T
        r, c = np.nonzero(rng.rand(300, 200) < 0.3)
        mf = NMF(embedding_dim=6, num_epoch=100, learning_rate=0.05)
        mf.fit(COOMatrix.from_data((r, c, R[r, c])))
        print('MF train rmse %.4f' % mf.train_rmse[-1])
        return
    if args.model == 'lda':
        rng = np.random.RandomState(0)
        docs = [list(rng.randint((i % 5) * 20, (i % 5 + 1) * 20, size=80)) for i in range(500)]
        lda = LDA(num_words=100, num_topics=5, num_iters=50, alpha=0.1, beta=0.01)
        lda.fit(Corpus.from_data(100, docs))
        print('LDA loglikelihood %.1f' % lda.loglikelihood[-1])
        return
    if args.train:
        neg = args.model in ('lr', 'svm', 'fm')
        train = Dataset.from_file(args.train, neg_y=neg)
        valid = Dataset.from_file(args.valid or args.train, neg_y=neg)
    else:
        X, y = synthetic()
        train = Dataset.from_data((X[:16000], y[:16000]))
        valid = Dataset.from_data((X[16000:], y[16000:]))
    if args.model == 'knn':
        knn = KNN(num_label=2, num_neighbor=5)
        print(knn.evaluate(train, valid, 'precision'))
        return
    model = {'lr': lambda: LogisticRegression(learning_rate=0.5, metrics='log-loss,error,precision'),
             'svm': lambda: SVM(learning_rate=0.5), 'linreg': lambda: LinearRegression(),
             'fm': lambda: FM(learning_rate=0.1), 'gbdt': lambda: GBDT(max_depth=6, metrics='log-loss,error,precision'),
             'rf': lambda: RandomForest(max_depth=6, num_round=50), 'nb': lambda: NaiveBayes(num_label=2)}[args.model]()
    model.fit(train, valid)
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, model.name())
        model.save_model(path)
        loaded = type(model)() if args.model != 'nb' else NaiveBayes(num_label=2)
        loaded.load_model(path)
    names = 'mse' if args.model == 'linreg' else ('cross-entropy,precision' if args.model == 'nb' else 'log-loss,error,precision')
    print(model.name(), 'valid', loaded.evaluate(valid, names))


if __name__ == '__main__':
    main()
