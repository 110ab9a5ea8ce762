"""Vision / sequence model zoo: LogReg, MLP, 3-layer CNN, LeNet, AlexNet, VGG-16/19, ResNet-18/34,
RNN and LSTM classifiers.

reference: examples/cnn/models/{LogReg,MLP,CNN,LeNet,AlexNet,VGG,ResNet,RNN,LSTM}.py - same call
contract ``model(x, y_) -> (loss, y)`` on MNIST (784) / CIFAR (3x32x32) shaped inputs, so
``examples/cnn/main.py --model resnet18 ...`` works as upstream.  Written against the layer
helpers below; convolutions run as im2col + tcgen05 GEMM, batch-norm / pooling / softmax-CE on the
sm_100a kernels.
"""
from __future__ import annotations

import numpy as np

import hetu as ht
from hetu import init


# ------------------------------------------------------------------------------- helpers
def fc(x, shape, name, with_relu=True):
    w = init.random_normal(shape=shape, stddev=0.1, name=name + '_weight')
    b = init.random_normal(shape=shape[-1:], stddev=0.1, name=name + '_bias')
    x = ht.matmul_op(x, w)
    x = x + ht.broadcastto_op(b, x)
    return ht.relu_op(x) if with_relu else x


def conv_relu(x, in_ch, out_ch, name, kernel=3, padding=1, stride=1, relu=True, bias=True):
    w = init.random_normal(shape=(out_ch, in_ch, kernel, kernel), stddev=0.1, name=name + '_weight')
    x = ht.conv2d_op(x, w, padding=padding, stride=stride)
    if bias:
        b = init.random_normal(shape=(out_ch,), stddev=0.1, name=name + '_bias')
        x = x + ht.conv2d_broadcastto_op(b, x)
    return ht.relu_op(x) if relu else x


def conv_bn_relu(x, in_ch, out_ch, name, kernel=3, padding=1, stride=1, relu=True):
    w = init.he_normal(shape=(out_ch, in_ch, kernel, kernel), name=name + '_weight')
    x = ht.conv2d_op(x, w, padding=padding, stride=stride)
    scale = init.ones(shape=(out_ch,), name=name + '_bn_scale')
    bias = init.zeros(shape=(out_ch,), name=name + '_bn_bias')
    x = ht.batch_normalization_op(x, scale, bias, momentum=0.9, eps=1e-5)
    return ht.relu_op(x) if relu else x


def softmax_loss(logits, y_):
    loss = ht.softmaxcrossentropy_op(logits, y_)
    return ht.reduce_mean_op(loss, [0])


# --------------------------------------------------------------------------------- models
def logreg(x, y_, in_dim=784, num_class=10):
    """Multinomial logistic regression (reference LogReg.py)."""
    w = init.zeros((in_dim, num_class), name='logreg_weight')
    b = init.zeros((num_class,), name='logreg_bias')
    y = ht.matmul_op(x, w)
    y = y + ht.broadcastto_op(b, y)
    return softmax_loss(y, y_), y


def mlp(x, y_, in_dim=3072, hidden=256, num_class=10):
    """3-layer perceptron (reference MLP.py; BASELINE config 1 uses the CIFAR10 shape)."""
    x = fc(x, (in_dim, hidden), 'mlp_fc1')
    x = fc(x, (hidden, hidden), 'mlp_fc2')
    y = fc(x, (hidden, num_class), 'mlp_fc3', with_relu=False)
    return softmax_loss(y, y_), y


def cnn_3_layers(x, y_, num_class=10):
    """conv5-pool, conv5-pool, fc on MNIST (reference CNN.py)."""
    x = ht.array_reshape_op(x, [-1, 1, 28, 28])
    x = conv_relu(x, 1, 32, 'cnn3_conv1', kernel=5, padding=2)
    x = ht.max_pool2d_op(x, 2, 2, padding=0, stride=2)
    x = conv_relu(x, 32, 64, 'cnn3_conv2', kernel=5, padding=2)
    x = ht.max_pool2d_op(x, 2, 2, padding=0, stride=2)
    x = ht.array_reshape_op(x, [-1, 7 * 7 * 64])
    y = fc(x, (7 * 7 * 64, num_class), 'cnn3_fc', with_relu=False)
    return softmax_loss(y, y_), y


def lenet(x, y_, num_class=10):
    """LeNet-5 on MNIST (reference LeNet.py)."""
    x = ht.array_reshape_op(x, [-1, 1, 28, 28])
    x = conv_relu(x, 1, 6, 'lenet_conv1', kernel=5, padding=2)
    x = ht.max_pool2d_op(x, 2, 2, padding=0, stride=2)
    x = conv_relu(x, 6, 16, 'lenet_conv2', kernel=5, padding=2)
    x = ht.max_pool2d_op(x, 2, 2, padding=0, stride=2)
    x = ht.array_reshape_op(x, [-1, 7 * 7 * 16])
    x = fc(x, (7 * 7 * 16, 120), 'lenet_fc1')
    x = fc(x, (120, 84), 'lenet_fc2')
    y = fc(x, (84, num_class), 'lenet_fc3', with_relu=False)
    return softmax_loss(y, y_), y


def alexnet(x, y_, num_class=10):
    """AlexNet scaled to 28x28 / 32x32 inputs (reference AlexNet.py)."""
    def block(x, cin, cout, name, kernel, padding, pool):
        x = conv_bn_relu(x, cin, cout, name, kernel=kernel, padding=padding)
        return ht.max_pool2d_op(x, 2, 2, padding=0, stride=2) if pool else x
    x = ht.array_reshape_op(x, [-1, 1, 28, 28])
    x = block(x, 1, 32, 'alexnet_conv1', 3, 1, True)          # 14
    x = block(x, 32, 64, 'alexnet_conv2', 3, 1, True)         # 7
    x = block(x, 64, 128, 'alexnet_conv3', 3, 1, False)
    x = block(x, 128, 256, 'alexnet_conv4', 3, 1, False)
    x = block(x, 256, 256, 'alexnet_conv5', 3, 1, False)
    x = ht.max_pool2d_op(x, 3, 3, padding=1, stride=2)        # 4
    x = ht.array_reshape_op(x, [-1, 256 * 4 * 4])
    x = ht.dropout_op(fc(x, (256 * 4 * 4, 1024), 'alexnet_fc1'), 0.5)
    x = ht.dropout_op(fc(x, (1024, 512), 'alexnet_fc2'), 0.5)
    y = fc(x, (512, num_class), 'alexnet_fc3', with_relu=False)
    return softmax_loss(y, y_), y


_VGG = {16: (2, 2, 3, 3, 3), 19: (2, 2, 4, 4, 4)}


def vgg(x, y_, num_layers=16, num_class=10):
    """VGG-16 / VGG-19 with batch norm on CIFAR (reference VGG.py)."""
    x = ht.array_reshape_op(x, [-1, 3, 32, 32])
    cin = 3
    for bi, (reps, cout) in enumerate(zip(_VGG[num_layers], (64, 128, 256, 512, 512))):
        for r in range(reps):
            x = conv_bn_relu(x, cin, cout, 'vgg_b%d_conv%d' % (bi + 1, r + 1))
            cin = cout
        x = ht.max_pool2d_op(x, 2, 2, padding=0, stride=2)
    x = ht.array_reshape_op(x, [-1, 512])
    x = fc(x, (512, 4096), 'vgg_fc1')
    x = fc(x, (4096, 4096), 'vgg_fc2')
    y = fc(x, (4096, num_class), 'vgg_fc3', with_relu=False)
    return softmax_loss(y, y_), y


def vgg16(x, y_, num_class=10):
    return vgg(x, y_, 16, num_class)


def vgg19(x, y_, num_class=10):
    return vgg(x, y_, 19, num_class)


def _basic_block(x, cin, cout, stride, name):
    """Two 3x3 conv-bn, identity or 1x1 projection shortcut."""
    out = conv_bn_relu(x, cin, cout, name + '_conv1', stride=stride)
    out = conv_bn_relu(out, cout, cout, name + '_conv2', relu=False)
    if stride != 1 or cin != cout:
        x = conv_bn_relu(x, cin, cout, name + '_short', kernel=1, padding=0, stride=stride, relu=False)
    return ht.relu_op(out + x)


def resnet(x, y_, num_layers=18, num_class=10):
    """ResNet-18 / ResNet-34 for CIFAR (3x3 stem, 4 stages of basic blocks; reference ResNet.py)."""
    reps = {18: (2, 2, 2, 2), 34: (3, 4, 6, 3)}[num_layers]
    x = ht.array_reshape_op(x, [-1, 3, 32, 32])
    x = conv_bn_relu(x, 3, 64, 'resnet_stem')
    cin = 64
    for si, (n, cout) in enumerate(zip(reps, (64, 128, 256, 512))):
        for bi in range(n):
            stride = 2 if (bi == 0 and si > 0) else 1
            x = _basic_block(x, cin, cout, stride, 'resnet_s%d_b%d' % (si + 1, bi + 1))
            cin = cout
    x = ht.avg_pool2d_op(x, 4, 4, padding=0, stride=4)
    x = ht.array_reshape_op(x, [-1, 512])
    y = fc(x, (512, num_class), 'resnet_fc', with_relu=False)
    return softmax_loss(y, y_), y


def resnet18(x, y_, num_class=10):
    return resnet(x, y_, 18, num_class)


def resnet34(x, y_, num_class=10):
    return resnet(x, y_, 34, num_class)


def rnn(x, y_, dimhidden=128, diminput=28, nsteps=28, num_class=10):
    """Elman RNN over the 28 rows of an MNIST digit (reference RNN.py): the recurrence is unrolled
    into the static graph, one matmul pair per step."""
    w_in = init.random_normal((diminput, dimhidden), stddev=0.1, name='rnn_w_in')
    b_in = init.random_normal((dimhidden,), stddev=0.1, name='rnn_b_in')
    w_h = init.random_normal((dimhidden + dimhidden, dimhidden), stddev=0.1, name='rnn_w_h')
    b_h = init.random_normal((dimhidden,), stddev=0.1, name='rnn_b_h')
    w_out = init.random_normal((dimhidden, num_class), stddev=0.1, name='rnn_w_out')
    b_out = init.random_normal((num_class,), stddev=0.1, name='rnn_b_out')
    h = None
    for t in range(nsteps):
        xt = ht.slice_op(x, (0, t * diminput), (-1, diminput))
        a = ht.matmul_op(xt, w_in)
        a = a + ht.broadcastto_op(b_in, a)
        if h is None:
            h = ht.zeroslike_op(a)
        cat = ht.concat_op(a, h, axis=1)
        z = ht.matmul_op(cat, w_h)
        h = ht.relu_op(z + ht.broadcastto_op(b_h, z))
    y = ht.matmul_op(h, w_out)
    y = y + ht.broadcastto_op(b_out, y)
    return softmax_loss(y, y_), y


def lstm(x, y_, dimhidden=128, diminput=28, nsteps=28, num_class=10):
    """LSTM classifier over the rows of an MNIST digit (reference LSTM.py)."""
    def gate(name):
        return (init.random_normal((diminput, dimhidden), stddev=0.1, name='lstm_%s_w' % name),
                init.random_normal((dimhidden, dimhidden), stddev=0.1, name='lstm_%s_u' % name),
                init.random_normal((dimhidden,), stddev=0.1, name='lstm_%s_b' % name))
    params = {g: gate(g) for g in ('forget', 'input', 'output', 'cell')}
    w_out = init.random_normal((dimhidden, num_class), stddev=0.1, name='lstm_out_w')
    b_out = init.random_normal((num_class,), stddev=0.1, name='lstm_out_b')
    h = c = None

    def pre(xt, h, g):
        w, u, b = params[g]
        z = ht.matmul_op(xt, w)
        if h is not None:
            z = z + ht.matmul_op(h, u)
        return z + ht.broadcastto_op(b, z)

    for t in range(nsteps):
        xt = ht.slice_op(x, (0, t * diminput), (-1, diminput))
        f = ht.sigmoid_op(pre(xt, h, 'forget'))
        i = ht.sigmoid_op(pre(xt, h, 'input'))
        o = ht.sigmoid_op(pre(xt, h, 'output'))
        g = ht.tanh_op(pre(xt, h, 'cell'))
        c = ht.mul_op(i, g) if c is None else ht.mul_op(f, c) + ht.mul_op(i, g)
        h = ht.mul_op(o, ht.tanh_op(c))
    y = ht.matmul_op(h, w_out)
    y = y + ht.broadcastto_op(b_out, y)
    return softmax_loss(y, y_), y


MODELS = {'logreg': logreg, 'mlp': mlp, 'cnn_3_layers': cnn_3_layers, 'lenet': lenet, 'alexnet': alexnet,
          'vgg16': vgg16, 'vgg19': vgg19, 'resnet18': resnet18, 'resnet34': resnet34, 'rnn': rnn, 'lstm': lstm}
