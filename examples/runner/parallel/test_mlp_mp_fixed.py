#!/usr/bin/env python
"""Tensor parallel middle layers on two devices with the suite's fixed weights / feeds
(reference test_mlp_mp.py --split {left,right,middle} as driven by all_mlp_tests.sh).
    heturun -w 2 python examples/runner/parallel/test_mlp_mp_fixed.py --split right"""
import argparse

from mlp_suite import build, ht, run_and_save

ap = argparse.ArgumentParser()
ap.add_argument('--split', default='middle', choices=['left', 'right', 'middle'])
ap.add_argument('--steps', type=int, default=20)
ap.add_argument('--save', default=None)
args = ap.parse_args()
pair = (ht.gpu(0), ht.gpu(1))
x, y_, loss, train = build([ht.gpu(0), (pair, args.split), (pair, args.split), (pair, args.split), ht.gpu(1)])
run_and_save(args.save or 'mp_' + args.split, x, y_, loss, train, args.steps)
