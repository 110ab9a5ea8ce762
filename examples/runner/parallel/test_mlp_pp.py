#!/usr/bin/env python
"""Layers placed on consecutive devices = pipeline stages linked by send / receive
(reference test_mlp_pp.py).  Without ``--pipeline`` the stages run one batch at a time and the
losses must equal the baseline exactly; with gpipe the 4 micro-batch gradients are summed.
    heturun -w 4 python examples/runner/parallel/test_mlp_pp.py --stages 4"""
import argparse

from mlp_suite import build, ht, run_and_save

ap = argparse.ArgumentParser()
ap.add_argument('--stages', type=int, default=2, choices=[2, 3, 4, 5])
ap.add_argument('--steps', type=int, default=20)
ap.add_argument('--pipeline', default=None, choices=[None, 'gpipe', 'pipedream', 'hetpipe'])
ap.add_argument('--save', default=None)
args = ap.parse_args()
stage_of = [min(i * args.stages // 5, args.stages - 1) for i in range(5)]
x, y_, loss, train = build([ht.gpu(s) for s in stage_of])
run_and_save(args.save or 'pp%d%s' % (args.stages, '_' + args.pipeline if args.pipeline else ''), x, y_, loss, train,
             args.steps, args.pipeline)
