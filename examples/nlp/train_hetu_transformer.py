#!/usr/bin/env python
"""Encoder-decoder Transformer (reference examples/nlp/train_hetu_transformer.py, hparams.py).
Synthetic copy task by default (no IWSLT data in the image): target = source sequence.  With
``--data-dir`` (output of prepare_data.py) it trains on the segmented parallel corpus instead:

    python prepare_data.py --src raw.de --tgt raw.en --out iwslt2016 --vocab_size 8000
    python train_hetu_transformer.py --data-dir iwslt2016 --maxlen 100 --epochs 20"""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..'))
import hetu as ht                                              # noqa: E402
from hetu_61a7_b200.models.nlp import Transformer, TransformerHParams   # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--batch-size', type=int, default=16)
    ap.add_argument('--d-model', type=int, default=512)
    ap.add_argument('--num-blocks', type=int, default=6)
    ap.add_argument('--maxlen', type=int, default=64)
    ap.add_argument('--vocab-size', type=int, default=8000)
    ap.add_argument('--steps', type=int, default=50)
    ap.add_argument('--lr', type=float, default=3e-4)
    ap.add_argument('--gpu', type=int, default=0)
    ap.add_argument('--data-dir', default=None)
    ap.add_argument('--epochs', type=int, default=1)
    args = ap.parse_args()
    data = None
    if args.data_dir:
        sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
        from data_load import DataLoader
        seg = os.path.join(args.data_dir, 'segmented')
        data = DataLoader(os.path.join(seg, 'train.de.bpe'), os.path.join(seg, 'train.en.bpe'), args.maxlen, args.maxlen,
                          os.path.join(seg, 'bpe.vocab'))
        args.vocab_size = max(args.vocab_size, len(data.token2idx))
    hp = TransformerHParams(vocab_size=args.vocab_size, d_model=args.d_model, d_ff=4 * args.d_model,
                            num_blocks=args.num_blocks, maxlen1=args.maxlen, maxlen2=args.maxlen,
                            dropout_rate=0.1, batch_size=args.batch_size)
    xs, ys_in, ys_out = ht.Variable('xs'), ht.Variable('ys_in'), ht.Variable('ys_out')
    loss, _ = Transformer(hp)(xs, ys_in, ys_out)
    train = ht.optim.AdamOptimizer(args.lr).minimize(loss)
    ex = ht.Executor([loss, train], ctx=ht.cpu(0) if args.gpu < 0 else ht.gpu(args.gpu))
    if data is not None:
        step = 0
        for ep in range(args.epochs):
            for src, tgt_in, tgt_out in data.batches(args.batch_size, seed=ep):
                t0 = time.time()
                l = float(ex.run(feed_dict={xs: src, ys_in: tgt_in, ys_out: tgt_out})[0].asnumpy())
                if step % 10 == 0:
                    print('epoch %d step %d loss %.4f %.1f ms' % (ep, step, l, (time.time() - t0) * 1e3), flush=True)
                step += 1
        return
    rs = np.random.RandomState(0)
    for step in range(args.steps):
        src = rs.randint(2, args.vocab_size, (args.batch_size, args.maxlen)).astype(np.float32)
        tgt_in = np.concatenate([np.ones((args.batch_size, 1), np.float32), src[:, :-1]], 1)
        t0 = time.time()
        l = float(ex.run(feed_dict={xs: src, ys_in: tgt_in, ys_out: src})[0].asnumpy())
        print('step %d loss %.4f %.1f ms' % (step, l, (time.time() - t0) * 1e3), flush=True)


if __name__ == '__main__':
    main()
