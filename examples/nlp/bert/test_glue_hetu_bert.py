#!/usr/bin/env python
"""Fine-tune BERT on a GLUE classification task (reference examples/nlp/bert/test_glue_hetu_bert.py,
scripts/test_glue_hetu_bert_{base,large}.sh).

    python test_glue_hetu_bert.py --task sst-2 --data-dir glue_data/SST-2 --vocab vocab.txt --size base
    python test_glue_hetu_bert.py --task sst-2 --synthetic --size tiny --gpu -1      # no data needed

``--init-checkpoint`` loads weights saved by ``Executor.save`` (pre-training run) by name."""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..', '..'))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import hetu as ht                                              # noqa: E402
from glue_processor.glue import PROCESSORS, convert_examples_to_features   # noqa: E402
from hetu_61a7_b200.models.nlp import BertConfig, BertForSequenceClassification   # noqa: E402
from hetu_61a7_b200.utils.tokenizers import BertTokenizer      # noqa: E402


def synthetic_task(n, seq_len, vocab, num_labels, seed):
    """Label = which of ``num_labels`` marker tokens occurs in the sentence: learnable, like a keyword task."""
    rs = np.random.RandomState(seed)
    ids = rs.randint(1000, vocab, (n, seq_len))
    lab = rs.randint(0, num_labels, n)
    pos = rs.randint(1, seq_len - 1, n)
    ids[np.arange(n), pos] = 500 + lab
    ids[:, 0], ids[:, -1] = 101, 102
    return dict(input_ids=ids, token_type_ids=np.zeros((n, seq_len), np.int64), attention_mask=np.ones((n, seq_len), np.int64),
                labels=lab)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--task', default='sst-2', choices=sorted(PROCESSORS))
    ap.add_argument('--data-dir', default=None)
    ap.add_argument('--vocab', default=None)
    ap.add_argument('--synthetic', action='store_true')
    ap.add_argument('--size', default='base', choices=['tiny', 'base', 'large'])
    ap.add_argument('--batch-size', type=int, default=32)
    ap.add_argument('--seq-len', type=int, default=128)
    ap.add_argument('--lr', type=float, default=2e-5)
    ap.add_argument('--epochs', type=int, default=3)
    ap.add_argument('--init-checkpoint', default=None, help='directory written by Executor.save')
    ap.add_argument('--gpu', type=int, default=0, help='-1 = CPU executor')
    args = ap.parse_args()
    proc = PROCESSORS[args.task]()
    labels = proc.get_labels()
    kw = dict(batch_size=args.batch_size, seq_len=args.seq_len)
    cfg = {'base': BertConfig.base, 'large': BertConfig.large,
           'tiny': lambda **k: BertConfig(hidden_size=128, num_hidden_layers=2, num_attention_heads=2, intermediate_size=512,
                                          **k)}[args.size](**kw)
    cfg.attention_probs_dropout_prob = 0.0
    if args.synthetic or not args.data_dir:
        train = synthetic_task(args.batch_size * 40, args.seq_len, cfg.vocab_size, len(labels), 0)
        dev = synthetic_task(args.batch_size * 8, args.seq_len, cfg.vocab_size, len(labels), 1)
    else:
        tok = BertTokenizer(vocab_file=args.vocab, do_lower_case=True)
        train = convert_examples_to_features(proc.get_train_examples(args.data_dir), labels, args.seq_len, tok)
        dev = convert_examples_to_features(proc.get_dev_examples(args.data_dir), labels, args.seq_len, tok)
    names = ['input_ids', 'token_type_ids', 'attention_mask', 'labels']

    def loaders(k):
        return ht.dataloader_op([ht.Dataloader(np.asarray(train[k], np.float32), args.batch_size, 'train', drop_last=True),
                                 ht.Dataloader(np.asarray(dev[k], np.float32), args.batch_size, 'validate', drop_last=True)])
    nodes = [loaders(k) for k in names]
    logits, loss = BertForSequenceClassification(cfg, num_labels=len(labels))(*nodes)
    train_op = ht.optim.AdamOptimizer(learning_rate=args.lr, l2reg=0.01).minimize(loss)
    ctx = ht.cpu(0) if args.gpu < 0 else ht.gpu(args.gpu)
    ex = ht.Executor({'train': [loss, logits, nodes[3], train_op], 'validate': [loss, logits, nodes[3]]}, ctx=ctx)
    if args.init_checkpoint:
        ex.load(args.init_checkpoint, 'bert.pkl')
    for ep in range(args.epochs):
        t0 = time.time()
        tl, correct, seen = 0.0, 0, 0
        nb = ex.get_batch_num('train')
        for _ in range(nb):
            l, lg, lab, _ = ex.run('train', convert_to_numpy_ret_vals=True)
            tl += float(l)
            correct += int((lg.argmax(1) == lab.reshape(-1)).sum())
            seen += lab.size
        vc, vs = 0, 0
        for _ in range(ex.get_batch_num('validate')):
            l, lg, lab = ex.run('validate', convert_to_numpy_ret_vals=True)
            vc += int((lg.argmax(1) == lab.reshape(-1)).sum())
            vs += lab.size
        print('epoch %d loss %.4f train acc %.4f dev acc %.4f (%.1fs)' % (ep, tl / nb, correct / seen, vc / max(vs, 1),
                                                                         time.time() - t0), flush=True)


if __name__ == '__main__':
    main()
