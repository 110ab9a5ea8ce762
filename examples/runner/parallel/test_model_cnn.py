#!/usr/bin/env python
"""The conv net cut into pipeline stages (reference test_model_cnn.py / test_model_cnn_complex.py).
    heturun -w 2 python test_model_cnn.py --stages 2 [--pipeline gpipe]"""
import argparse

from cnn_suite import build, ht, run_and_save

ap = argparse.ArgumentParser()
ap.add_argument('--stages', type=int, default=2, choices=[2, 3, 4, 5])
ap.add_argument('--steps', type=int, default=12)
ap.add_argument('--pipeline', default=None, choices=[None, 'gpipe', 'pipedream'])
args = ap.parse_args()
stage_of = [min(i * args.stages // 5, args.stages - 1) for i in range(5)]
x, y_, loss, train = build([ht.gpu(s) for s in stage_of])
run_and_save('pp%d%s' % (args.stages, '_' + args.pipeline if args.pipeline else ''), x, y_, loss, train, args.steps, args.pipeline)
