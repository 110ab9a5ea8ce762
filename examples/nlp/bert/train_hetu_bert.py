#!/usr/bin/env python
"""BERT pre-training (MLM + NSP) on one GPU or data-parallel over all GPUs of the node
(reference examples/nlp/bert/train_hetu_bert.py, train_hetu_bert_dp.py, scripts/train_hetu_bert_large_dp.sh).

    python examples/nlp/bert/train_hetu_bert.py --size base --batch-size 32 --seq-len 128
    heturun -w 8 python examples/nlp/bert/train_hetu_bert.py --size large --dp

The wikipedia/bookcorpus shards are not in the image: synthetic token ids with 15 % masked
positions are generated when ``--data`` is not given (npz with input_ids, token_type_ids,
attention_mask, masked_lm_labels, next_sentence_label).
"""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..', '..'))
import hetu as ht                                              # noqa: E402
from hetu_61a7_b200.models.nlp import BertConfig, BertForPreTraining   # noqa: E402


def synthetic(n, cfg, seed):
    rs = np.random.RandomState(seed)
    ids = rs.randint(1000, cfg.vocab_size, (n, cfg.seq_len))
    labels = np.full((n, cfg.seq_len), -1, dtype=np.int64)
    m = rs.rand(n, cfg.seq_len) < 0.15
    labels[m] = ids[m]
    ids[m] = 103                                              # [MASK]
    tt = np.zeros((n, cfg.seq_len))
    tt[:, cfg.seq_len // 2:] = 1
    return dict(input_ids=ids, token_type_ids=tt, attention_mask=np.ones((n, cfg.seq_len)),
                masked_lm_labels=labels, next_sentence_label=rs.randint(0, 2, (n,)))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--size', default='base', choices=['base', 'large', 'tiny'])
    ap.add_argument('--batch-size', type=int, default=32)
    ap.add_argument('--seq-len', type=int, default=128)
    ap.add_argument('--lr', type=float, default=1e-5)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--dp', action='store_true', help='data parallel over WORLD_SIZE ranks (AllReduce)')
    ap.add_argument('--unfused-attention', action='store_true', help="reference's 5-op attention")
    ap.add_argument('--data', default=None)
    args = ap.parse_args()
    comm = ht.wrapped_mpi_nccl_init()
    kw = dict(batch_size=args.batch_size, seq_len=args.seq_len, fused_attention=not args.unfused_attention)
    cfg = {'base': BertConfig.base, 'large': BertConfig.large,
           'tiny': lambda **k: BertConfig(hidden_size=128, num_hidden_layers=2, num_attention_heads=2,
                                          intermediate_size=512, **k)}[args.size](**kw)
    if not args.unfused_attention:
        cfg.attention_probs_dropout_prob = 0.0
    data = dict(np.load(args.data)) if args.data else synthetic(args.batch_size * 8, cfg, comm.rank)
    names = ['input_ids', 'token_type_ids', 'attention_mask', 'masked_lm_labels', 'next_sentence_label']
    nodes = [ht.dataloader_op([ht.Dataloader(np.asarray(data[k], dtype=np.float32), args.batch_size, 'train')])
             for k in names]
    _, _, mlm_loss, nsp_loss = BertForPreTraining(cfg)(*nodes)
    loss = mlm_loss + nsp_loss
    train = ht.optim.AdamOptimizer(learning_rate=args.lr, l2reg=0.01).minimize(loss)
    ex = ht.Executor({'train': [loss, mlm_loss, nsp_loss, train]}, ctx=ht.gpu(comm.local_rank),
                     comm_mode='AllReduce' if args.dp and comm.nrank > 1 else None, shard_dataloaders=False)
    for step in range(args.steps):
        t0 = time.time()
        l, a, b, _ = ex.run('train', convert_to_numpy_ret_vals=True)
        if comm.rank == 0:
            print('step %d loss %.4f (mlm %.4f nsp %.4f) %.1f ms' % (step, float(l), float(a), float(b),
                                                                   (time.time() - t0) * 1e3), flush=True)


if __name__ == '__main__':
    main()
