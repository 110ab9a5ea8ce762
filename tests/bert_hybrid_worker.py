"""Hybrid BERT worker: pipeline stages x tensor-parallel FFN (x data-parallel replicas) vs a single-process run
with the same weights.  argv: mode(cpu|gpu) pp tp dp"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
mode, pp, tp, dp = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
if mode == 'cpu':
    os.environ['HETU_COMM_BACKEND'] = 'gloo'
import hetu as ht                                           # noqa: E402
from hetu_61a7_b200.models.nlp import BertConfig, BertForPreTraining   # noqa: E402

comm = ht.wrapped_mpi_nccl_init()
rank = comm.rank
assert comm.nrank == pp * tp * dp
B, S = 4, 8
kw = dict(vocab_size=64, hidden_size=32, num_hidden_layers=2 * pp, num_attention_heads=4, intermediate_size=64,
          max_position_embeddings=16, batch_size=B, seq_len=S, hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0,
          tie_word_embeddings=False)
rs = np.random.RandomState(0)
feeds = dict(ids=rs.randint(0, 64, (B, S)).astype(np.float32), tt=np.zeros((B, S), np.float32), am=np.ones((B, S), np.float32),
             mlm=rs.randint(0, 64, (B, S)).astype(np.float32), nsp=rs.randint(0, 2, (B,)).astype(np.float32))


def stage_ctx(s):
    base = s * tp * dp
    workers = []
    for d in range(dp):
        devs = tuple(ht.gpu(base + d * tp + t) for t in range(tp))
        workers.append(devs if tp > 1 else devs[0])
    return workers if dp > 1 else workers[0]


def build(hybrid):
    cfg = BertConfig(tp=tp if hybrid else 1, **kw)
    names = ['ids', 'tt', 'am', 'mlm', 'nsp']
    if hybrid:
        stages = [stage_ctx(s) for s in range(pp)]
        with ht.context(stages[0]):
            v = {n: ht.Variable(n) for n in names[:3]}
        with ht.context(stages[-1]):
            v.update({n: ht.Variable(n) for n in names[3:]})
        model = BertForPreTraining(cfg, stage_contexts=stages)
    else:
        v = {n: ht.Variable(n) for n in names}
        model = BertForPreTraining(cfg)
    _, _, l1, l2 = model(v['ids'], v['tt'], v['am'], v['mlm'], v['nsp'])
    if hybrid:
        with ht.context(stages[-1]):
            loss = l1 + l2
            # replicas see the same batch and gradients are SUMMED over them (reference semantics: ncclSum,
            # scripts pre-divide the learning rate)
            train = ht.optim.SGDOptimizer(0.05 / dp).minimize(loss)
    else:
        loss = l1 + l2
        train = ht.optim.SGDOptimizer(0.05).minimize(loss)
    return v, loss, train


# --- single-process oracle on every rank (CPU executor, same seed => same initial weights by name) ---
v, loss, train = build(False)
ex0 = ht.Executor([loss, train], ctx=ht.cpu(0), seed=11)
init = {p.name: ex0.config.placeholder_to_arr_map[p].clone() for p in ex0.param_nodes}
ref = [float(ex0.run(feed_dict={v[n]: feeds[n] for n in v})[0].asnumpy()) for _ in range(4)]

v, loss, train = build(True)
pipeline = sys.argv[5] if len(sys.argv) > 5 else None
ex = ht.Executor({'train': [loss, train]}, seed=11, pipeline=pipeline)
for p in ex.param_nodes:                                   # same starting point as the oracle
    t = ex.config.placeholder_to_arr_map.get(p)
    if t is not None and p.name in init:
        t.copy_(init[p.name].to(t.device))
got = []
for _ in range(4):
    if pipeline:                                           # one micro-batch: same maths as the plain executor
        out = ex.run('train', feed_dict={v[n]: [feeds[n]] for n in v}, batch_num=1)
        if out[0] is not None:
            got.append(float(out[0][0].asnumpy()))
    else:
        out = ex.run('train', feed_dict={v[n]: feeds[n] for n in v})
        if out[0] is not None:
            got.append(float(out[0].asnumpy()))
if got:
    tol = 5e-2 if mode == 'gpu' else 2e-4
    assert np.allclose(got, ref, rtol=tol, atol=1e-4), (got, ref)
comm.barrier()
print('HYBRID pp%d tp%d dp%d rank %d OK %s' % (pp, tp, dp, rank, got[:2]), flush=True)
