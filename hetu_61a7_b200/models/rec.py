"""Recommendation: Neural Collaborative Filtering (NeuMF = GMF branch * MLP branch).

reference: examples/rec/hetu_ncf.py:7-47 (MovieLens; one embedding row per user / item holds both
the GMF and the MLP factors).  ``embedding_ctx=None`` keeps the tables in HBM (they are tiny next
to 180 GB); pass ``ht.cpu(0)`` to reproduce the reference's host placement for PS / Hybrid runs.
"""
from __future__ import annotations

import hetu as ht
from hetu import init


def neural_mf(user_input, item_input, y_, num_users, num_items, embed_dim=8, layers=(64, 32, 16, 8),
              learning_rate=0.01, embedding_ctx=None):
    width = embed_dim + layers[0] // 2
    users = init.random_normal((num_users, width), stddev=0.01, name='user_embed', ctx=embedding_ctx)
    items = init.random_normal((num_items, width), stddev=0.01, name='item_embed', ctx=embedding_ctx)
    u = ht.embedding_lookup_op(users, user_input, ctx=embedding_ctx)
    i = ht.embedding_lookup_op(items, item_input, ctx=embedding_ctx)
    gmf = ht.mul_op(ht.slice_op(u, (0, 0), (-1, embed_dim)), ht.slice_op(i, (0, 0), (-1, embed_dim)))
    h = ht.concat_op(ht.slice_op(u, (0, embed_dim), (-1, -1)), ht.slice_op(i, (0, embed_dim), (-1, -1)), axis=1)
    for k in range(len(layers) - 1):
        w = init.random_normal((layers[k], layers[k + 1]), stddev=0.1, name='ncf_W%d' % (k + 1))
        h = ht.relu_op(ht.matmul_op(h, w))
    w_out = init.random_normal((embed_dim + layers[-1], 1), stddev=0.1, name='ncf_W_out')
    y = ht.sigmoid_op(ht.matmul_op(ht.concat_op(gmf, h, axis=1), w_out))
    loss = ht.reduce_mean_op(ht.binarycrossentropy_op(y, y_), [0])
    train_op = ht.optim.SGDOptimizer(learning_rate=learning_rate).minimize(loss)
    return loss, y, train_op
