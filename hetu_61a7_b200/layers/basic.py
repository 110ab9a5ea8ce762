"""Basic layers.  reference: python/hetu/layers/{linear,conv,normalization,
dropout,embedding,pooling,relu,gelu,reshape,slice,sequence,sum,concatenate,
identity}.py.  Layers only build graph nodes; all compute is in ``ops``."""
from __future__ import annotations

from .. import initializers as init
from ..graph.node import Op
from .. import ops as _ops


class BaseLayer(object):
    def __init__(self):
        pass

    def __call__(self, *args, **kwargs):
        raise NotImplementedError


def _act(activation):
    if isinstance(activation, str):
        table = {'relu': _ops.relu_op, 'gelu': _ops.gelu_op, 'tanh': _ops.tanh_op,
                 'sigmoid': _ops.sigmoid_op}
        if activation not in table:
            raise NotImplementedError(activation)
        return table[activation]
    return activation


class Linear(BaseLayer):
    def __init__(self, in_features, out_features, initializer=init.GenXavierUniform(), bias=True,
                 activation=None, weight_transpose=False, name='linear'):
        self.in_features, self.out_features = in_features, out_features
        self.bias = bias
        self.activation = _act(activation)
        self.weight_transpose = weight_transpose
        self.name = name
        if isinstance(initializer, Op):
            self.weight_var = initializer
        else:
            shape = (out_features, in_features) if weight_transpose else (in_features, out_features)
            self.weight_var = initializer(shape=shape, name=name + '_weight')
        if bias:
            self.bias_var = init.zeros(shape=(out_features,), name=name + '_bias')

    def __call__(self, x):
        if self.bias:
            x = _ops.linear_op(x, self.weight_var, self.bias_var, trans_B=self.weight_transpose)
        else:
            x = _ops.matmul_op(x, self.weight_var, trans_B=self.weight_transpose)
        if self.activation is not None:
            x = self.activation(x)
        return x


class Conv2d(BaseLayer):
    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0,
                 initializer=init.GenXavierUniform(), bias=True, activation=None, name='conv2d'):
        self.in_channels, self.out_channels = in_channels, out_channels
        if isinstance(kernel_size, tuple):
            assert len(kernel_size) == 2 and kernel_size[0] == kernel_size[1]
            kernel_size = kernel_size[0]
        self.kernel_size, self.stride, self.padding = kernel_size, stride, padding
        self.bias = bias
        self.activation = _act(activation)
        self.name = name
        self.weight_var = initializer(shape=(out_channels, in_channels, kernel_size, kernel_size),
                                      name=name + '_weight')
        if bias:
            self.bias_var = init.zeros(shape=(out_channels,), name=name + '_bias')

    def __call__(self, x):
        if self.bias:
            x = _ops.conv2d_add_bias_op(x, self.weight_var, self.bias_var, stride=self.stride,
                                        padding=self.padding)
        else:
            x = _ops.conv2d_op(x, self.weight_var, stride=self.stride, padding=self.padding)
        if self.activation is not None:
            x = self.activation(x)
        return x


class BatchNorm(BaseLayer):
    def __init__(self, num_channels, name='batchnorm'):
        self.num_channels, self.name = num_channels, name
        self.scale_var = init.ones(shape=(num_channels,), name=name + '_weight')
        self.bias_var = init.zeros(shape=(num_channels,), name=name + '_bias')

    def __call__(self, x):
        return _ops.batch_normalization_op(x, self.scale_var, self.bias_var)


class LayerNorm(BaseLayer):
    def __init__(self, num_channels, name='layernorm', eps=1e-05):
        self.num_channels, self.name, self.eps = num_channels, name, eps
        self.scale_var = init.ones(shape=(num_channels,), name=name + '_weight')
        self.bias_var = init.zeros(shape=(num_channels,), name=name + '_bias')

    def __call__(self, x):
        return _ops.layer_normalization_op(x, self.scale_var, self.bias_var, eps=self.eps)


class DropOut(BaseLayer):
    def __init__(self, p=0.5):
        self.p = p
        self.keep_prob = 1 - p

    def __call__(self, x):
        if self.p == 0:
            return x
        return _ops.dropout_op(x, self.keep_prob)


class Embedding(BaseLayer):
    def __init__(self, num_embeddings, embedding_dim, initializer=init.GenXavierNormal(),
                 name='embedding', ctx=None):
        self.num_embeddings, self.embedding_dim = num_embeddings, embedding_dim
        self.name, self.ctx = name, ctx
        self.embedding_table = initializer(shape=(num_embeddings, embedding_dim),
                                           name=name + '.weight', ctx=ctx)

    def __call__(self, x):
        return _ops.embedding_lookup_op(self.embedding_table, x, ctx=self.ctx)


class MaxPool2d(BaseLayer):
    def __init__(self, kernel_size, stride, padding=0):
        self.kernel_size, self.stride, self.padding = kernel_size, stride, padding

    def __call__(self, x):
        return _ops.max_pool2d_op(x, self.kernel_size, self.kernel_size, self.padding, self.stride)


class AvgPool2d(BaseLayer):
    def __init__(self, kernel_size, stride, padding=0):
        self.kernel_size, self.stride, self.padding = kernel_size, stride, padding

    def __call__(self, x):
        return _ops.avg_pool2d_op(x, self.kernel_size, self.kernel_size, self.padding, self.stride)


class Relu(BaseLayer):
    def __call__(self, x):
        return _ops.relu_op(x)


class Gelu(BaseLayer):
    def __call__(self, x):
        return _ops.gelu_op(x)


class Reshape(BaseLayer):
    def __init__(self, shape):
        self.shape = shape

    def __call__(self, x):
        return _ops.array_reshape_op(x, self.shape)


class Slice(BaseLayer):
    def __init__(self, begin, size):
        self.begin, self.size = begin, size

    def __call__(self, x):
        return _ops.slice_op(x, self.begin, self.size)


class Identity(BaseLayer):
    def __call__(self, x):
        return x


class Sequence(BaseLayer):
    def __init__(self, *args):
        self.layers = args

    def __call__(self, x):
        for layer in self.layers:
            x = layer(x)
        return x


class SumLayers(BaseLayer):
    def __init__(self, layers):
        self.layers = layers

    def __call__(self, x):
        return _ops.sum_op([layer(x) for layer in self.layers])


class Concatenate(BaseLayer):
    def __init__(self, axis):
        self.axis = axis

    def __call__(self, xs):
        return _ops.concatenate_op(xs, axis=self.axis)


class ConcatenateLayers(BaseLayer):
    def __init__(self, layers, axis=0):
        self.layers, self.axis = layers, axis

    def __call__(self, x):
        return _ops.concatenate_op([layer(x) for layer in self.layers], axis=self.axis)
