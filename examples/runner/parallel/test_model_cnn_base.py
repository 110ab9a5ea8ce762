#!/usr/bin/env python
"""Single-device baseline of the CNN equivalence suite (reference test_model_cnn_base.py)."""
import argparse

from cnn_suite import build, ht, run_and_save

ap = argparse.ArgumentParser()
ap.add_argument('--steps', type=int, default=12)
args = ap.parse_args()
x, y_, loss, train = build([ht.gpu(0)] * 5)
run_and_save('base', x, y_, loss, train, args.steps)
