#!/usr/bin/env python
"""The conv net under a mixed layout: convolution blocks pipelined over single GPUs, the first dense layer on ONE
tensor-parallel worker of two GPUs with a chosen ``ht.dispatch`` split (left: rows of the activation, right: columns of
the weight, middle: the contraction dimension), the classifier on a last GPU.  Loss curve goes to
``results/cnn_complex_<split>.npy`` for validate_results.py (reference
examples/runner/parallel/test_model_cnn_complex.py: conv layers under ``--split 0..4`` / ``--split2`` layouts on 8 GPUs;
the tensor-parallel lowering of this framework covers matrix products, so the split sits on the dense layer).

    heturun -w 5 python test_model_cnn_complex.py --split middle
"""
import argparse

from cnn_suite import fixed_weights, ht, run_and_save

ap = argparse.ArgumentParser()
ap.add_argument('--split', default='middle', choices=['left', 'right', 'middle'])
ap.add_argument('--steps', type=int, default=12)
ap.add_argument('--learning-rate', type=float, default=0.05)
args = ap.parse_args()

W = fixed_weights()
with ht.context(ht.gpu(0)):
    x = ht.Variable('x')
    a = ht.relu_op(ht.conv2d_op(x, ht.Variable('cnn_c1', value=W['c1'].copy()), padding=1, stride=1))
    a = ht.max_pool2d_op(a, 2, 2, 0, 2)
with ht.context(ht.gpu(1)):
    a = ht.relu_op(ht.conv2d_op(a, ht.Variable('cnn_c2', value=W['c2'].copy()), padding=1, stride=1))
    a = ht.max_pool2d_op(a, 2, 2, 0, 2)
    a = ht.relu_op(ht.conv2d_op(a, ht.Variable('cnn_c3', value=W['c3'].copy()), padding=1, stride=1))
    a = ht.array_reshape_op(a, (-1, 16 * 4 * 4))
with ht.context((ht.gpu(2), ht.gpu(3))):                      # one worker, two GPUs
    w = ht.Variable('cnn_fc1', value=W['fc1'].copy())
    act_parts, w_parts = {'left': ((2, 1), (1, 1)), 'right': ((1, 1), (1, 2)), 'middle': ((1, 2), (2, 1))}[args.split]
    a = ht.relu_op(ht.matmul_op(ht.dispatch(a, act_parts), ht.dispatch(w, w_parts)))
with ht.context(ht.gpu(4)):
    a = ht.dispatch(a, (1, 1))
    y_ = ht.Variable('y_')
    logits = ht.matmul_op(a, ht.Variable('cnn_fc2', value=W['fc2'].copy()))
    loss = ht.reduce_mean_op(ht.softmaxcrossentropy_op(logits, y_), [0])
    train = ht.optim.SGDOptimizer(args.learning_rate).minimize(loss)
run_and_save('complex_' + args.split, x, y_, loss, train, args.steps)
