#!/usr/bin/env python
"""BERT-large with a Galvatron-style hybrid layout: pipeline stages x tensor-parallel feed-forward blocks x
data-parallel replicas on one 8-GPU node (BASELINE.json config 5).

    heturun -w 8 python examples/nlp/bert/train_hetu_bert_hybrid.py --pp 2 --tp 2 --dp 2 --micro-batches 4

Placement is expressed with contexts only: stage s lives on ``[(gpu, gpu), (gpu, gpu)]`` = dp workers of
tp devices each; ``ht.dispatch`` inside the model marks the column-/row-parallel FFN weights; the executor
cuts the graph per rank, joins stages with NCCL p2p and runs GPipe micro-batching (--pipeline gpipe).
Prints device-timed tokens/s (max over ranks).
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..', '..'))
import hetu as ht                                              # noqa: E402
from hetu_61a7_b200.models.nlp import BertConfig, BertForPreTraining   # noqa: E402
from hetu_61a7_b200.parallel.comm import device_max_ms         # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--size', default='large', choices=['large', 'base', 'tiny'])
    ap.add_argument('--pp', type=int, default=2)
    ap.add_argument('--tp', type=int, default=2)
    ap.add_argument('--dp', type=int, default=2)
    ap.add_argument('--batch-size', type=int, default=32, help='per replica and micro-batch')
    ap.add_argument('--seq-len', type=int, default=128)
    ap.add_argument('--micro-batches', type=int, default=4)
    ap.add_argument('--pipeline', default='gpipe', choices=['gpipe', 'pipedream', 'none'])
    ap.add_argument('--steps', type=int, default=10)
    args = ap.parse_args()
    comm = ht.wrapped_mpi_nccl_init()
    assert comm.nrank == args.pp * args.tp * args.dp, 'launch pp*tp*dp ranks'
    kw = dict(batch_size=args.batch_size, seq_len=args.seq_len, tp=args.tp, tie_word_embeddings=args.pp == 1,
              attention_probs_dropout_prob=0.0, hidden_dropout_prob=0.0)
    cfg = {'large': BertConfig.large, 'base': BertConfig.base,
           'tiny': lambda **k: BertConfig(hidden_size=128, num_hidden_layers=4, num_attention_heads=2,
                                          intermediate_size=512, **k)}[args.size](**kw)

    def stage_ctx(s):
        base = s * args.tp * args.dp
        workers = []
        for d in range(args.dp):
            devs = tuple(ht.gpu(base + d * args.tp + t) for t in range(args.tp))
            workers.append(devs if args.tp > 1 else devs[0])
        return workers if args.dp > 1 else workers[0]

    stages = [stage_ctx(s) for s in range(args.pp)]
    with ht.context(stages[0]):
        ids, tt, am = ht.Variable('input_ids'), ht.Variable('token_type_ids'), ht.Variable('attention_mask')
    with ht.context(stages[-1]):
        mlm, nsp = ht.Variable('masked_lm_labels'), ht.Variable('next_sentence_label')
    _, _, l1, l2 = BertForPreTraining(cfg, stage_contexts=stages)(ids, tt, am, mlm, nsp)
    with ht.context(stages[-1]):
        loss = l1 + l2
        train = ht.optim.AdamOptimizer(learning_rate=1e-5 / args.dp).minimize(loss)
    pipeline = None if args.pipeline == 'none' else args.pipeline
    ex = ht.Executor({'train': [loss, train]}, pipeline=pipeline, seed=0)
    rs = np.random.RandomState(0)
    M = args.micro_batches if pipeline else 1

    def batch():
        lab = rs.randint(0, cfg.vocab_size, (args.batch_size, args.seq_len)).astype(np.float32)
        lab[rs.rand(args.batch_size, args.seq_len) > 0.15] = -1
        return {ids: rs.randint(0, cfg.vocab_size, (args.batch_size, args.seq_len)).astype(np.float32),
                tt: np.zeros((args.batch_size, args.seq_len), np.float32), am: np.ones((args.batch_size, args.seq_len), np.float32),
                mlm: lab, nsp: rs.randint(0, 2, (args.batch_size,)).astype(np.float32)}

    mbs = [batch() for _ in range(M)]
    feed = {k: [mb[k] for mb in mbs] for k in mbs[0]} if pipeline else mbs[0]

    def step():
        return ex.run('train', feed_dict=feed, batch_num=M) if pipeline else ex.run('train', feed_dict=feed)

    for _ in range(3):
        out = step()
    comm.barrier()
    if torch.cuda.is_available():
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(args.steps):
            out = step()
        e.record()
        torch.cuda.synchronize()
        ms = device_max_ms(s.elapsed_time(e) / args.steps)
    else:                                               # CPU / gloo dry run of the same layout
        import time
        t0 = time.perf_counter()
        for _ in range(args.steps):
            out = step()
        ms = device_max_ms((time.perf_counter() - t0) * 1e3 / args.steps)
    tokens = args.batch_size * args.seq_len * M * args.dp
    l = out[0]
    if l is not None:
        l = float(np.mean([o.asnumpy() for o in l])) if isinstance(l, list) else float(l.asnumpy())
        print(json.dumps({'config': 'bert_%s hybrid pp%d tp%d dp%d %s x%d micro-batches' % (args.size, args.pp, args.tp, args.dp, args.pipeline, M),
                          'n_gpus': comm.nrank, 'ms_per_step': ms, 'value': tokens / (ms / 1e3), 'unit': 'tokens/s', 'loss': l}), flush=True)


if __name__ == '__main__':
    main()
