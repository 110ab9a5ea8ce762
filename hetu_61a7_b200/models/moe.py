"""MoE benchmark model.  reference: examples/moe/test_moe_{top,hash,ktop1,base,sam}.py
(one MoE layer, batch 16 x 1024 tokens x model_dim 2048, hidden 2048, 2 local experts per
GPU, top-k gate, nll loss over the token axis + auxiliary balance loss)."""
from __future__ import annotations

import numpy as np

from .. import layers as htl
from .. import ops as ht


def moe_model(x, y_, batch_size=16, num_tokens=1024, model_dim=2048, hidden_size=2048,
              num_local_experts=2, world=1, rank=0, top=2, gate='top', num_local_gpus=8,
              hierarchical=None, hash_indices=None):
    total = num_local_experts * world
    experts = [htl.Expert(embed_dim=model_dim, ffn_dim=hidden_size, dropout_rate=0.1,
                          activation='relu', name="expert_%d" % (rank * num_local_experts + i))
               for i in range(num_local_experts)]
    T = batch_size * num_tokens
    extra = None
    if gate == 'top':
        g = htl.TopKGate(embed_dim=model_dim, num_tokens=T, num_experts=total, k=top)
        layer = htl.MoELayer(gate=g, experts=experts, num_tokens=num_tokens, embed_dim=model_dim,
                             all2all_size=world, top=top, hierarchical=hierarchical)
        y, l_aux = layer(x)
    elif gate == 'ktop1':
        g = htl.KTop1Gate(embed_dim=model_dim, num_tokens=T, num_experts=total, k=top)
        layer = htl.KTop1Layer(gate=g, experts=experts, num_tokens=num_tokens, embed_dim=model_dim,
                               all2all_size=world, k=top)
        y, l_aux = layer(x)
    elif gate == 'sam':
        g = htl.SAMGate(embed_dim=model_dim, num_tokens=T, num_experts=total, k=top,
                        num_local_gpus=num_local_gpus)
        layer = htl.SAMLayer(gate=g, experts=experts, num_tokens=num_tokens, embed_dim=model_dim,
                             all2all_size=world, k=top, num_local_gpus=num_local_gpus)
        y, l_aux, extra = layer(x)
    elif gate == 'hash':
        g = htl.HashGate(embed_dim=model_dim, num_tokens=T, num_experts=total)
        layer = htl.HashLayer(gate=g, experts=experts, num_tokens=num_tokens, embed_dim=model_dim,
                              all2all_size=world)
        y, l_aux = layer(x, hash_indices), None
    elif gate == 'base':
        g = htl.BalanceAssignmentGate(embed_dim=model_dim, num_tokens=T, num_experts=total,
                                      device_id=rank)
        layer = htl.MoELayer(gate=g, experts=experts, num_tokens=T, embed_dim=model_dim,
                             all2all_size=world, name='BalanceAssignmentLayer')
        y, l_aux = layer(x), None
    else:
        raise ValueError(gate)
    y = ht.array_reshape_op(y, [-1, num_tokens, model_dim])
    y = ht.reduce_sum_op(y, axes=2)
    y = ht.softmax_op(y)
    loss = ht.nll_loss_op(y, y_, num_tokens)
    if l_aux is not None:
        loss = loss + l_aux
    if extra is not None:
        loss = loss + ht.array_reshape_op(extra, [1])
    return loss, y
