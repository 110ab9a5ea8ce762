"""Graph neural networks over a normalised sparse adjacency (CSR, fed per step).

reference: examples/gnn/gnn_model/layer.py:5-68 (GCN, SageConv) and model.py:6-62 (dense / sparse
front-ends).  ``norm_adj`` is a placeholder fed with ``ht.sparse_array`` (ND_Sparse_Array); the
aggregation runs on ``csrmm_op`` (warp-per-row CSR SpMM, csrc/gemm_sm100.cu) and the feature
transform on the tcgen05 GEMM.
"""
from __future__ import annotations

import hetu as ht
from hetu import init


class GCN(object):
    """H' = act(A_hat (H W + b))."""

    def __init__(self, in_features, out_features, norm_adj, activation=None, dropout=0, name='GCN', custom_init=None):
        if custom_init is not None:
            self.weight = ht.Variable(value=custom_init[0], name=name + '_Weight')
            self.bias = ht.Variable(value=custom_init[1], name=name + '_Bias')
        else:
            self.weight = init.xavier_uniform(shape=(in_features, out_features), name=name + '_Weight')
            self.bias = init.zeros(shape=(out_features,), name=name + '_Bias')
        self.mp, self.activation, self.dropout = norm_adj, activation, dropout
        self.output_width = out_features

    def __call__(self, x):
        if self.dropout > 0:
            x = ht.dropout_op(x, 1 - self.dropout)
        x = ht.matmul_op(x, self.weight)
        x = ht.csrmm_op(self.mp, x + ht.broadcastto_op(self.bias, x))
        if self.activation == 'relu':
            x = ht.relu_op(x)
        elif self.activation is not None:
            raise NotImplementedError(self.activation)
        return x


class SageConv(object):
    """GraphSAGE mean aggregator: concat(act(A_hat H W1 + b), H W2)."""

    def __init__(self, in_features, out_features, norm_adj, activation=None, dropout=0, name='Sage'):
        self.weight = init.xavier_uniform(shape=(in_features, out_features), name=name + '_Weight')
        self.bias = init.zeros(shape=(out_features,), name=name + '_Bias')
        self.weight2 = init.xavier_uniform(shape=(in_features, out_features), name=name + '_SelfWeight')
        self.mp, self.activation, self.dropout = norm_adj, activation, dropout
        self.output_width = 2 * out_features

    def __call__(self, x):
        feat = x
        if self.dropout > 0:
            x = ht.dropout_op(x, 1 - self.dropout)
        x = ht.matmul_op(ht.csrmm_op(self.mp, x), self.weight)
        x = x + ht.broadcastto_op(self.bias, x)
        if self.activation == 'relu':
            x = ht.relu_op(x)
        elif self.activation is not None:
            raise NotImplementedError(self.activation)
        return ht.concat_op(x, ht.matmul_op(feat, self.weight2), axis=1)


def dense_model(feature_dim, hidden_dim, num_classes, lr=0.1, layer=GCN, arch=None):
    """Two-layer node classifier on dense features.  Returns
    [feat, y_, norm_adj, mask], [loss, y, train_op]  (reference model.py:6-30)."""
    feat, y_, mask = ht.Variable('feat'), ht.Variable('y_'), ht.Variable('mask')
    norm_adj = ht.Variable('message_passing', trainable=False)
    l1 = layer(feature_dim, hidden_dim, norm_adj, activation='relu', name='gnn_l1')
    l2 = layer(l1.output_width, num_classes if layer is GCN else hidden_dim, norm_adj, name='gnn_l2')
    y = l2(l1(feat))
    if layer is not GCN:
        w = init.xavier_uniform((l2.output_width, num_classes), name='gnn_classifier')
        y = ht.matmul_op(y, w)
    loss = ht.softmaxcrossentropy_op(y, y_)
    loss = ht.reduce_mean_op(ht.mul_op(loss, mask), [0])
    train_op = ht.optim.SGDOptimizer(lr).minimize(loss)
    return [feat, y_, norm_adj, mask], [loss, y, train_op]


def sparse_model(num_nodes, embed_dim, hidden_dim, num_classes, lr=0.1, layer=GCN):
    """Featureless graphs: a trainable embedding per node replaces the input features
    (reference model.py:33-62)."""
    node_ids, y_, mask = ht.Variable('node_ids'), ht.Variable('y_'), ht.Variable('mask')
    norm_adj = ht.Variable('message_passing', trainable=False)
    table = init.random_normal((num_nodes, embed_dim), stddev=0.1, name='node_embedding')
    feat = ht.embedding_lookup_op(table, node_ids)
    l1 = layer(embed_dim, hidden_dim, norm_adj, activation='relu', name='gnn_l1')
    l2 = layer(l1.output_width, num_classes, norm_adj, name='gnn_l2')
    y = l2(l1(feat))
    if layer is not GCN:
        y = ht.matmul_op(y, init.xavier_uniform((l2.output_width, num_classes), name='gnn_classifier'))
    loss = ht.reduce_mean_op(ht.mul_op(ht.softmaxcrossentropy_op(y, y_), mask), [0])
    train_op = ht.optim.SGDOptimizer(lr).minimize(loss)
    return [node_ids, y_, norm_adj, mask], [loss, y, train_op]
