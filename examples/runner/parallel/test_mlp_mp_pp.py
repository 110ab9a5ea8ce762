#!/usr/bin/env python
"""Pipeline stages whose middle stage is tensor parallel over two devices (reference
test_mlp_mp_pp.py): layer 0 on gpu0, layers 1-3 on (gpu1, gpu2) with the given split, layer 4 on gpu3.
    heturun -w 4 python examples/runner/parallel/test_mlp_mp_pp.py --split middle"""
import argparse

from mlp_suite import build, ht, run_and_save

ap = argparse.ArgumentParser()
ap.add_argument('--split', default='middle', choices=['left', 'right', 'middle'])
ap.add_argument('--steps', type=int, default=20)
ap.add_argument('--save', default=None)
args = ap.parse_args()
pair = (ht.gpu(1), ht.gpu(2))
x, y_, loss, train = build([ht.gpu(0), (pair, args.split), (pair, args.split), (pair, args.split), ht.gpu(3)])
run_and_save(args.save or 'mp_pp_' + args.split, x, y_, loss, train, args.steps)
