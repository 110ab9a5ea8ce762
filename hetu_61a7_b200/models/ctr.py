"""CTR models: Wide&Deep, DeepFM, Deep&Cross on Criteo-shaped data, Wide&Deep on Adult.

reference: examples/ctr/models/{wdl_criteo,deepfm_criteo,dcn_criteo,dc_criteo,wdl_adult}.py.
Same graphs, same shapes (13 dense + 26 sparse fields, 33 762 577 x 128 table, MLP
13-256-256-256) and hyper-parameters; the embedding table lives wherever ``ctx``
says (the reference pins it to ``ht.cpu(0)`` and serves it through the parameter
server; on B200 the default is HBM, optionally sharded over NVLink -
``parallel/sharded_embedding.py``).
"""
from __future__ import annotations

from .. import initializers as init
from .. import ops as ht
from ..optimizer import SGDOptimizer

CRITEO_VOCAB = 33762577


def wdl_criteo(dense_input, sparse_input, y_, feature_dimension=CRITEO_VOCAB, embedding_size=128,
               learning_rate=0.01, ctx=None, embed_ctx=None):
    Embedding = init.random_normal([feature_dimension, embedding_size], stddev=0.01,
                                   name="snd_order_embedding", ctx=embed_ctx or ctx)
    sparse = ht.embedding_lookup_op(Embedding, sparse_input, ctx=embed_ctx or ctx)
    sparse = ht.array_reshape_op(sparse, (-1, 26 * embedding_size))
    W1 = init.random_normal([13, 256], stddev=0.01, name="W1")
    W2 = init.random_normal([256, 256], stddev=0.01, name="W2")
    W3 = init.random_normal([256, 256], stddev=0.01, name="W3")
    W4 = init.random_normal([256 + 26 * embedding_size, 1], stddev=0.01, name="W4")
    relu1 = ht.relu_op(ht.matmul_op(dense_input, W1))
    relu2 = ht.relu_op(ht.matmul_op(relu1, W2))
    y3 = ht.matmul_op(relu2, W3)
    y4 = ht.concat_op(sparse, y3, axis=1)
    y = ht.sigmoid_op(ht.matmul_op(y4, W4))
    loss = ht.reduce_mean_op(ht.binarycrossentropy_op(y, y_), [0])
    train_op = SGDOptimizer(learning_rate=learning_rate).minimize(loss)
    return loss, y, y_, train_op


def dfm_criteo(dense_input, sparse_input, y_, feature_dimension=CRITEO_VOCAB, embedding_size=128,
               learning_rate=0.01, ctx=None, embed_ctx=None):
    ectx = embed_ctx or ctx
    Embedding1 = init.random_normal([feature_dimension, 1], stddev=0.01, name="fst_order_embedding", ctx=ectx)
    FM_W = init.random_normal([13, 1], stddev=0.01, name="dense_parameter")
    sparse_1dim = ht.embedding_lookup_op(Embedding1, sparse_input, ctx=ectx)
    y1 = ht.matmul_op(dense_input, FM_W) + ht.reduce_sum_op(sparse_1dim, axes=1)
    Embedding2 = init.random_normal([feature_dimension, embedding_size], stddev=0.01,
                                    name="snd_order_embedding", ctx=ectx)
    sparse_2dim = ht.embedding_lookup_op(Embedding2, sparse_input, ctx=ectx)
    s_sum = ht.reduce_sum_op(sparse_2dim, axes=1)
    s_sum_sq = ht.mul_op(s_sum, s_sum)
    s_sq_sum = ht.reduce_sum_op(ht.mul_op(sparse_2dim, sparse_2dim), axes=1)
    second = (s_sum_sq + -1 * s_sq_sum) * 0.5
    y2 = ht.reduce_sum_op(second, axes=1, keepdims=True)
    flatten = ht.array_reshape_op(sparse_2dim, (-1, 26 * embedding_size))
    W1 = init.random_normal([26 * embedding_size, 256], stddev=0.01, name="W1")
    W2 = init.random_normal([256, 256], stddev=0.01, name="W2")
    W3 = init.random_normal([256, 1], stddev=0.01, name="W3")
    y3 = ht.matmul_op(ht.relu_op(ht.matmul_op(ht.relu_op(ht.matmul_op(flatten, W1)), W2)), W3)
    y = ht.sigmoid_op((y1 + y2) + y3)
    loss = ht.reduce_mean_op(ht.binarycrossentropy_op(y, y_), [0])
    train_op = SGDOptimizer(learning_rate=learning_rate).minimize(loss)
    return loss, y, y_, train_op


def cross_layer(x0, x1, embedding_len):
    weight = init.random_normal(shape=(embedding_len, 1), stddev=0.01, name='weight')
    bias = init.random_normal(shape=(embedding_len,), stddev=0.01, name='bias')
    x1w = ht.matmul_op(x1, weight)
    y = ht.mul_op(x0, ht.broadcastto_op(x1w, x0))
    return y + x1 + ht.broadcastto_op(bias, y)


def build_cross_layer(x0, embedding_len, num_layers=3):
    x1 = x0
    for _ in range(num_layers):
        x1 = cross_layer(x0, x1, embedding_len)
    return x1


def dcn_criteo(dense_input, sparse_input, y_, feature_dimension=CRITEO_VOCAB, embedding_size=128,
               learning_rate=0.003, ctx=None, embed_ctx=None):
    ectx = embed_ctx or ctx
    Embedding = init.random_normal([feature_dimension, embedding_size], stddev=0.01,
                                   name="snd_order_embedding", ctx=ectx)
    sparse = ht.array_reshape_op(ht.embedding_lookup_op(Embedding, sparse_input, ctx=ectx),
                                 (-1, 26 * embedding_size))
    x = ht.concat_op(sparse, dense_input, axis=1)
    emb_len = 26 * embedding_size + 13
    cross_output = build_cross_layer(x, emb_len, num_layers=3)
    W1 = init.random_normal([emb_len, 256], stddev=0.01, name="W1")
    W2 = init.random_normal([256, 256], stddev=0.01, name="W2")
    W3 = init.random_normal([256, 256], stddev=0.01, name="W3")
    W4 = init.random_normal([256 + emb_len, 1], stddev=0.01, name="W4")
    y3 = ht.matmul_op(ht.relu_op(ht.matmul_op(ht.relu_op(ht.matmul_op(x, W1)), W2)), W3)
    y = ht.sigmoid_op(ht.matmul_op(ht.concat_op(cross_output, y3, axis=1), W4))
    loss = ht.reduce_mean_op(ht.binarycrossentropy_op(y, y_), [0])
    train_op = SGDOptimizer(learning_rate=learning_rate).minimize(loss)
    return loss, y, y_, train_op


def residual_layer(x0, input_dim, hidden_dim):
    w1 = init.random_normal(shape=(input_dim, hidden_dim), stddev=0.1, name='dc_w1')
    b1 = init.random_normal(shape=(hidden_dim,), stddev=0.1, name='dc_b1')
    w2 = init.random_normal(shape=(hidden_dim, input_dim), stddev=0.1, name='dc_w2')
    b2 = init.random_normal(shape=(input_dim,), stddev=0.1, name='dc_b2')
    h = ht.relu_op(ht.linear_op(x0, w1, b1))
    return ht.relu_op(ht.linear_op(h, w2, b2) + x0)


def dc_criteo(dense_input, sparse_input, y_, feature_dimension=CRITEO_VOCAB, embedding_size=8,
              learning_rate=0.001, ctx=None, embed_ctx=None, num_layers=5):
    """Deep Crossing: residual MLP over [embeddings, dense]."""
    ectx = embed_ctx or ctx
    Embedding = init.random_normal([feature_dimension, embedding_size], stddev=0.01,
                                   name="snd_order_embedding", ctx=ectx)
    sparse = ht.array_reshape_op(ht.embedding_lookup_op(Embedding, sparse_input, ctx=ectx),
                                 (-1, 26 * embedding_size))
    x = ht.concat_op(sparse, dense_input, axis=1)
    dim = 26 * embedding_size + 13
    for _ in range(num_layers):
        x = residual_layer(x, dim, dim)
    W = init.random_normal([dim, 1], stddev=0.1, name="dc_out")
    y = ht.sigmoid_op(ht.matmul_op(x, W))
    loss = ht.reduce_mean_op(ht.binarycrossentropy_op(y, y_), [0])
    train_op = SGDOptimizer(learning_rate=learning_rate).minimize(loss)
    return loss, y, y_, train_op


def wdl_adult(X_deep, X_wide, y_, dim_wide=809, learning_rate=5.0 / 128):
    """Wide&Deep on Adult: 8 categorical embeddings (dim 8) + 4 continuous -> 50-20-2 deep
    part, linear wide part; softmax cross entropy over 2 classes.
    ``X_deep``: list of 12 input nodes (8 id columns, 4 continuous columns)."""
    vocab = [50, 50, 50, 50, 50, 50, 50, 50]
    parts = []
    for i in range(8):
        table = init.random_normal([vocab[i], 8], stddev=0.1, name="Embedding_deep_%d" % i)
        parts.append(ht.array_reshape_op(ht.embedding_lookup_op(table, X_deep[i]), (-1, 8)))
    for i in range(4):
        parts.append(ht.array_reshape_op(X_deep[8 + i], (-1, 1)))
    deep_in = ht.concatenate_op(parts, axis=1)
    W1 = init.random_normal([68, 50], stddev=0.1, name="W1")
    b1 = init.random_normal([50], stddev=0.1, name="b1")
    W2 = init.random_normal([50, 20], stddev=0.1, name="W2")
    b2 = init.random_normal([20], stddev=0.1, name="b2")
    h = ht.relu_op(ht.linear_op(ht.relu_op(ht.linear_op(deep_in, W1, b1)), W2, b2))
    W_out = init.random_normal([dim_wide + 20, 2], stddev=0.1, name="W")
    y = ht.matmul_op(ht.concat_op(X_wide, h, axis=1), W_out)
    loss = ht.reduce_mean_op(ht.softmaxcrossentropy_op(y, y_), [0])
    train_op = SGDOptimizer(learning_rate=learning_rate).minimize(loss)
    return loss, ht.softmax_op(y), y_, train_op
