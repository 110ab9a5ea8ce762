#!/usr/bin/env python
"""Single-device baseline of the equivalence suite (reference test_mlp_base.py).
    python examples/runner/parallel/test_mlp_base.py [--steps 20]"""
import argparse

from mlp_suite import build, ht, run_and_save

ap = argparse.ArgumentParser()
ap.add_argument('--steps', type=int, default=20)
ap.add_argument('--save', default='base')
args = ap.parse_args()
x, y_, loss, train = build([ht.gpu(0)] * 5)
run_and_save(args.save, x, y_, loss, train, args.steps)
