"""Non-negative matrix factorization and latent Dirichlet allocation.

reference: third_party/HetuML/hetuml/python/hetuml/decomposition/{mf,lda}.py and
hetuml/cc/ml/include/model/{mf,lda}/**.  Distributed variants: NMF averages the factors over the
workers after every epoch (reference mf/common/model_average.h); LDA sums the workers' word-topic
count deltas after every sweep (the approximate-distributed Gibbs scheme of parallel_lda.h)."""
from __future__ import annotations

import ctypes
from typing import Optional

import numpy as np

from . import _core as C
from .data import COOMatrix, Corpus
from .model_base import MLBase, join_args


class NMF(MLBase):
    ps_data_type = 'float'
    _NAME = 'MF'

    def __init__(self, embedding_dim: int = 8, num_epoch: int = 10, learning_rate: float = 0.1, l1_reg: float = 0.0,
                 l2_reg: float = 0.0, parallel=False, aux_args=None):
        assert embedding_dim > 0, 'Embedding dimemsion should be positive'
        assert num_epoch > 0, 'Number of epochs should be positive'
        assert learning_rate > 0, 'Learning rate should be positive'
        assert l1_reg >= 0 and l2_reg >= 0, 'regularization terms should be non-negative'
        super().__init__(join_args(aux_args, embedding_dim=embedding_dim, num_epoch=num_epoch, learning_rate=learning_rate,
                                   l1_reg=l1_reg, l2_reg=l2_reg, parallel=parallel))
        self.train_rmse = np.zeros(0)

    def fit(self, train_data, valid_data=None) -> None:
        train = COOMatrix.from_data(train_data)
        a, comm = self._args, self._comm()
        k = int(a['embedding_dim'])
        nr, nc = train.shape
        if comm is not None:
            nr, nc = (int(v) for v in comm.allreduce_max(np.array([nr, nc], dtype=np.float64)))
        rng = np.random.RandomState(int(a.get('seed', 0)))
        scale = float(np.sqrt(max(float(train.data.mean()) if train.data.size else 1.0, 1e-6) / k))
        P = (rng.uniform(0, 1, (nr, k)) * scale).astype(np.float32)
        Q = (rng.uniform(0, 1, (nc, k)) * scale).astype(np.float32)
        step = 1 if comm is not None else int(a['num_epoch'])
        curve = []
        for ep in range(0, int(a['num_epoch']), step):
            rm = (ctypes.c_double * step)()
            C.check(C.lib().hml_mf_fit(train.row.ctypes.data_as(C.I32P), train.col.ctypes.data_as(C.I32P),
                                       train.data.ctypes.data_as(C.F32P), ctypes.c_int64(train.data.size), k,
                                       P.ctypes.data_as(C.F32P), Q.ctypes.data_as(C.F32P), step,
                                       ctypes.c_float(a['learning_rate']), ctypes.c_float(a['l1_reg']),
                                       ctypes.c_float(a['l2_reg']), ctypes.c_uint64(int(a.get('seed', 0)) + ep), rm), 'MF fit')
            curve.extend(rm[:])
            if comm is not None:
                P = comm.average(P.reshape(-1)).reshape(nr, k)
                Q = comm.average(Q.reshape(-1)).reshape(nc, k)
        self.train_rmse = np.array(curve)
        self._arrays.update(P=P, Q=Q)
        if valid_data is not None and a.get('verbose', False):
            v = COOMatrix.from_data(valid_data)
            print('%s valid rmse %.6f' % (self.name(), float(np.sqrt(np.mean((self.predict(v) - v.data) ** 2)))), flush=True)

    def predict(self, pred_data) -> np.ndarray:
        pred = COOMatrix.from_data(pred_data)
        P, Q = np.ascontiguousarray(self._arrays['P']), np.ascontiguousarray(self._arrays['Q'])
        if pred.row.size and (pred.row.max() >= P.shape[0] or pred.col.max() >= Q.shape[0]):
            raise ValueError('row / column id outside the fitted matrix')
        out = np.empty(pred.data.size, dtype=np.float32)
        C.check(C.lib().hml_mf_predict(pred.row.ctypes.data_as(C.I32P), pred.col.ctypes.data_as(C.I32P),
                                       ctypes.c_int64(pred.data.size), int(P.shape[1]), P.ctypes.data_as(C.F32P),
                                       Q.ctypes.data_as(C.F32P), out.ctypes.data_as(C.F32P)), 'MF predict')
        return out


MF = NMF


class LDA(MLBase):
    ps_data_type = 'int'
    _NAME = 'LDA'

    def __init__(self, num_words: int, num_topics: int = 100, num_iters: int = 100, alpha: float = 0.0,
                 beta: float = 0.001, long_doc_thres: int = 600, parallel=False, aux_args=None):
        assert num_words > 0, 'Number of words should be positive'
        assert num_topics > 0, 'Number of topics should be positive'
        assert num_iters > 0, 'Number of iterations should be positive'
        assert alpha >= 0 and beta >= 0, 'Alpha / beta should be non-negative'
        assert long_doc_thres >= 0, 'Threshold for long docs should be non-negative'
        if alpha == 0.0:
            alpha = 50.0 / num_topics                  # the customary default when unset
        super().__init__(join_args(aux_args, num_topics=num_topics, num_words=num_words, num_iters=num_iters, alpha=alpha,
                                   beta=beta, long_doc_thres=long_doc_thres, parallel=parallel))
        self.loglikelihood = []

    def _run(self, corpus: Corpus, z, nwk, nk, iters, train, seed):
        a = self._args
        llh = ctypes.c_double(0.0)
        C.check(C.lib().hml_lda_run(corpus.doc_ptr.ctypes.data_as(C.I64P), corpus.words.ctypes.data_as(C.I32P),
                                    ctypes.c_int64(corpus.get_num_docs()), int(a['num_words']), int(a['num_topics']),
                                    int(iters), ctypes.c_double(a['alpha']), ctypes.c_double(a['beta']),
                                    z.ctypes.data_as(C.I32P), nwk.ctypes.data_as(C.I32P), nk.ctypes.data_as(C.I32P),
                                    int(train), ctypes.c_uint64(seed), ctypes.byref(llh)), 'LDA')
        return llh.value

    def fit(self, train_data: Corpus) -> None:
        if not isinstance(train_data, Corpus):
            raise TypeError('Input for %s should be Corpus' % self.name())
        a, comm = self._args, self._comm()
        V, K = int(a['num_words']), int(a['num_topics'])
        seed = int(a.get('seed', 0)) + (comm.rank if comm is not None else 0)
        z = np.full(train_data.get_num_tokens(), -1, dtype=np.int32)
        nwk, nk = np.zeros((V, K), dtype=np.int32), np.zeros(K, dtype=np.int32)
        self.loglikelihood = []
        if comm is None:
            self.loglikelihood.append(self._run(train_data, z, nwk, nk, int(a['num_iters']), 1, seed))
        else:
            # every worker samples its documents against the global counts of the previous sweep
            glob = np.zeros(V * K, dtype=np.float64)
            for it in range(int(a['num_iters']) + 1):
                before = nwk.copy()
                if it == 0:
                    self._run(train_data, z, nwk, nk, 0, 1, seed)          # random initial assignment only
                else:
                    self.loglikelihood.append(self._run(train_data, z, nwk, nk, 1, 2, seed + 1000 * it))
                delta = (nwk - before).astype(np.float64).reshape(-1)
                comm.allreduce(delta)
                glob += delta
                nwk = np.ascontiguousarray(glob.reshape(V, K).astype(np.int32))
                nk = np.ascontiguousarray(nwk.sum(axis=0).astype(np.int32))
        self._arrays.update(nwk=nwk, nk=nk)
        self._z = z

    def predict(self, pred_data: Corpus) -> np.ndarray:
        """Topic of every token of ``pred_data`` (word-topic counts held fixed)."""
        if not isinstance(pred_data, Corpus):
            raise TypeError('Input for %s should be Corpus' % self.name())
        nwk = np.ascontiguousarray(self._arrays['nwk'], dtype=np.int32)
        nk = np.ascontiguousarray(self._arrays['nk'], dtype=np.int32)
        z = np.full(pred_data.get_num_tokens(), -1, dtype=np.int32)
        self.predict_loglikelihood = self._run(pred_data, z, nwk, nk, max(1, int(self._args['num_iters']) // 5), 0,
                                               int(self._args.get('seed', 0)) + 17)
        return z

    def topic_word(self) -> np.ndarray:
        """[num_topics, num_words] smoothed topic-word distributions."""
        nwk = self._arrays['nwk'].astype(np.float64) + float(self._args['beta'])
        return (nwk / nwk.sum(axis=0, keepdims=True)).T
