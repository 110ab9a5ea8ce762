"""Command-line hyper-parameters of the Transformer example (reference examples/nlp/hparams.py:
``Hparams.parser`` with the IWSLT'16 de-en defaults)."""
import argparse


class Hparams(object):
    parser = argparse.ArgumentParser()
    # prepare_data.py
    parser.add_argument('--vocab_size', default=32000, type=int)
    # data
    parser.add_argument('--train1', default='iwslt2016/segmented/train.de.bpe', help='source training text (one sentence per line)')
    parser.add_argument('--train2', default='iwslt2016/segmented/train.en.bpe', help='target training text')
    parser.add_argument('--eval1', default='iwslt2016/segmented/eval.de.bpe')
    parser.add_argument('--eval2', default='iwslt2016/segmented/eval.en.bpe')
    parser.add_argument('--vocab', default='iwslt2016/segmented/bpe.vocab', help='token per line (optionally "token<TAB>score")')
    # training
    parser.add_argument('--batch_size', default=16, type=int)
    parser.add_argument('--lr', default=0.0003, type=float)
    parser.add_argument('--warmup_steps', default=4000, type=int)
    parser.add_argument('--num_epochs', default=20, type=int)
    parser.add_argument('--logdir', default='log/1')
    # model
    parser.add_argument('--d_model', default=512, type=int)
    parser.add_argument('--d_ff', default=2048, type=int)
    parser.add_argument('--num_blocks', default=6, type=int)
    parser.add_argument('--num_heads', default=8, type=int)
    parser.add_argument('--maxlen1', default=100, type=int, help='maximum source length')
    parser.add_argument('--maxlen2', default=100, type=int, help='maximum target length')
    parser.add_argument('--dropout_rate', default=0.3, type=float)
    parser.add_argument('--smoothing', default=0.1, type=float)
    parser.add_argument('--gpu', default=0, type=int, help='-1 = CPU executor')
