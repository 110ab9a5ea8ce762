"""Worker process of the data-parallel hetuml test (pattern: reference
third_party/HetuML/examples/parallel_ml/*.py: Cluster(...) then Model(parallel=True).fit)."""
import os
import sys

import numpy as np
import scipy.sparse as sp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from hetuml.cluster import Cluster                      # noqa: E402
from hetuml.ensemble import GBDT                        # noqa: E402
from hetuml.linear import LogisticRegression            # noqa: E402
from hetuml.naive_bayes import NaiveBayes               # noqa: E402
from hetuml.decomposition import LDA                    # noqa: E402
from hetuml.data import Corpus                          # noqa: E402

sched, out_dir = sys.argv[1], sys.argv[2]
cluster = Cluster(sched, 1, 2, 'worker')
rank = cluster.rank

rng = np.random.RandomState(0)                          # same full data everywhere, each worker keeps its half
n, d = 4000, 10
Xd = rng.standard_normal((n, d)).astype(np.float32)
Xd[rng.rand(n, d) < 0.4] = 0
y = ((Xd[:, 0] > 0.2) ^ (Xd[:, 1] < -0.1)).astype(np.float32)
X = sp.csr_matrix(Xd)
mine = slice(rank, n, 2)
Xs, ys = sp.csr_matrix(Xd[mine]), y[mine]

gb = GBDT(max_depth=4, num_round=12, learning_rate=0.3, num_split=24, min_node_ins=20, parallel=True)
gb.fit((Xs, ys))
np.save(os.path.join(out_dir, 'gbdt_%d.npy' % rank), gb.predict_score(X))
ev = gb.evaluate((Xs, ys), 'error')
assert ev['error'] < 0.1, ev

ypm = np.where(y > 0, 1.0, -1.0).astype(np.float32)
lr = LogisticRegression(num_epoch=5, batch_size=100, learning_rate=0.5, parallel=True)
lr.fit((Xs, ypm[mine]))
np.save(os.path.join(out_dir, 'lr_%d.npy' % rank), lr.weights)

lab = ((Xd[:, 0] > 0).astype(np.int64) + (Xd[:, 1] > 0).astype(np.int64)).astype(np.float32)
nb = NaiveBayes(num_label=3, parallel=True)
nb.fit((Xs, lab[mine]))
np.save(os.path.join(out_dir, 'nb_%d.npy' % rank), nb.predict_score(X))
if rank == 0:
    single = NaiveBayes(num_label=3)
    single.fit((X, lab))
    np.save(os.path.join(out_dir, 'nb_single.npy'), single.predict_score(X))

V, K = 30, 3
docs = [list(np.random.RandomState(i).randint((i % K) * 10, (i % K + 1) * 10, size=40)) for i in range(120)]
lda = LDA(num_words=V, num_topics=K, num_iters=40, alpha=0.1, beta=0.01, parallel=True)
lda.fit(Corpus.from_data(V, docs[rank::2]))
np.save(os.path.join(out_dir, 'lda_%d.npy' % rank), lda.topic_word())
print('OK', flush=True)
cluster.join()
