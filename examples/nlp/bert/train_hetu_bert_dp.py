#!/usr/bin/env python
"""Data-parallel BERT pre-training (reference examples/nlp/bert/train_hetu_bert_dp.py): the same
trainer with gradients all-reduced over the launched ranks.
    heturun -w 8 python train_hetu_bert_dp.py --size large --batch-size 32"""
import sys

from train_hetu_bert import main

if '--dp' not in sys.argv:
    sys.argv.append('--dp')
main()
