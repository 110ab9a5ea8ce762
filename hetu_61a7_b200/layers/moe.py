"""Mixture-of-experts layers and gates.

reference: python/hetu/layers/{moe_layer,TopGate,KTop1Gate,ktop1_layer,SAMGate,
sam_layer,HashGate,hash_layer,BalanceGate}.py.  The four near-identical
``Expert`` / ``*Layer`` classes of the reference are one implementation here,
parameterised by the gate kind.  Experts are named ``expert_*`` so the optimizer
leaves their gradients out of the data-parallel all-reduce (optimizer.py:150-152).

On B200 the dispatch -> all-to-all -> expert FFN -> all-to-all -> combine chain
can be replaced by the fused peer-store path (``fused=True``; parallel/moe_fused.py):
tokens are written straight into the owner GPU's expert buffer over NVLink and
the expert GEMMs run as one grouped tcgen05 launch.
"""
from __future__ import annotations

import math
import os

import numpy as np

from .. import initializers as init
from .. import ops as ht
from ..ops.variable import Variable
from .basic import BaseLayer, _act


class Expert(BaseLayer):
    """Two-layer FFN: (d -> ffn) -> activation -> (ffn -> d)."""

    def __init__(self, embed_dim, ffn_dim, dropout_rate, initializer=init.GenXavierUniform(),
                 bias=False, activation=None, name="expert"):
        self.embed_dim, self.ffn_dim = embed_dim, ffn_dim
        self.keep_prob = 1 - dropout_rate
        self.bias = bias
        self.activation = _act(activation)
        self.initializer = initializer
        self.name = name
        self.w1 = self.w2 = None

    def build(self):
        if self.w1 is None:
            self.w1 = self.initializer(shape=(self.embed_dim, self.ffn_dim), name=self.name + '_weight_1')
            self.w2 = self.initializer(shape=(self.ffn_dim, self.embed_dim), name=self.name + '_weight_2')
            if self.bias:
                self.b1 = init.zeros(shape=(1, self.ffn_dim), name=self.name + '_bias_1')
                self.b2 = init.zeros(shape=(1, self.embed_dim), name=self.name + '_bias_2')

    def __call__(self, x):
        self.build()
        x = ht.array_reshape_op(x, [-1, self.embed_dim])
        x = ht.matmul_op(x, self.w1)
        if self.bias:
            x = x + ht.broadcastto_op(self.b1, x)
        if self.activation is not None:
            x = self.activation(x)
        x = ht.matmul_op(x, self.w2)
        if self.bias:
            x = x + ht.broadcastto_op(self.b2, x)
        return x


def balance_loss(gates, mask, num_experts):
    """GShard auxiliary loss: E * sum_e mean_t(gates)[e] * mean_t(mask)[e]."""
    me = ht.reduce_mean_op(gates, axes=0)
    ce = ht.reduce_mean_op(mask, axes=0)
    return ht.mul_byconst_op(ht.reducesumaxiszero_op(me * ce), float(num_experts))


def alignment_loss(gates, top1_group, topk_indice, num_local_gpus):
    tmp = ht.sam_max_op(gates, top1_group, topk_indice, num_local_gpus)
    return ht.reduce_sum_op(ht.reduce_sum_op(tmp, axes=0), axes=0)


def _slots(mask_topk, k):
    """Per-route slot index inside the chosen expert's capacity buffer."""
    locations1 = ht.cumsum_with_bias_op(mask_topk[0], bias=-1, dim=0)
    location_s = [ht.reduce_sum_op(locations1 * mask_topk[0], axes=1)]
    acc_base = None
    for i in range(1, k):
        prev = ht.reduce_sum_op(mask_topk[i - 1], axes=0, keepdims=True)
        acc_base = prev if acc_base is None else acc_base + prev
        loc = ht.cumsum_with_bias_op(mask_topk[i], bias=-1, dim=0)
        loc = loc + ht.broadcastto_op(acc_base, loc)
        location_s.append(ht.reduce_sum_op(loc * mask_topk[i], axes=1))
    return location_s


def topkgating(logits, k, capacity_factor, num_tokens, num_experts, embed_dim):
    capacity = k * math.ceil((num_tokens / num_experts) * capacity_factor)
    if os.environ.get('HETU_MOE_FUSED_GATE', '1') == '1' and k <= 4:
        # one op for softmax + top-k + slots + balance loss (ops/moe_gate.py); the pieces are views of its output
        packed = ht.fused_topk_gate_op(logits, k)
        packed.expected_tokens = num_tokens
        indices_s = [ht.gate_piece_op(packed, ('idx', i), num_tokens) for i in range(k)]
        location_s = [ht.gate_piece_op(packed, ('loc', i), num_tokens) for i in range(k)]
        gates_s = [ht.gate_piece_op(packed, ('gate', i), num_tokens) for i in range(k)]
        l_aux = ht.gate_piece_op(packed, 'aux', num_tokens)
        return l_aux, indices_s, location_s, gates_s, capacity
    gates = ht.softmax_op(logits)
    topk_indices = ht.topk_idx_op(gates, topk=k)
    indices_s = [ht.array_reshape_op(ht.split_op(topk_indices, axes=[1], indices=[i], splits=[k]), [-1])
                 for i in range(k)]
    mask_topk = [ht.array_reshape_op(ht.one_hot_op(indices_s[i], num_classes=num_experts),
                                     [-1, num_experts]) for i in range(k)]
    l_aux = balance_loss(gates, mask_topk[0], num_experts)
    for i in range(1, k):
        l_aux = l_aux + balance_loss(gates, mask_topk[i], num_experts)
    location_s = _slots(mask_topk, k)
    gates_s = [ht.reduce_sum_op(ht.mul_op(gates, mask_topk[i]), axes=1) for i in range(k)]
    return l_aux, indices_s, location_s, gates_s, capacity


def ktop1gating(logits, k, capacity_factor, num_tokens, num_experts, embed_dim):
    """k prototypes, top-1 inside each group of E/k experts."""
    per = num_experts // k
    gates = [ht.softmax_op(ht.split_op(logits, axes=[1], indices=[i], splits=[k])) for i in range(k)]
    capacity = k * math.ceil((num_tokens / num_experts) * capacity_factor)
    indices_s = [ht.array_reshape_op(ht.topk_idx_op(gates[i], topk=1), [-1]) for i in range(k)]
    mask_topk = [ht.array_reshape_op(ht.one_hot_op(indices_s[i], num_classes=per), [-1, per])
                 for i in range(k)]
    l_aux = balance_loss(gates[0], mask_topk[0], per)
    for i in range(1, k):
        l_aux = l_aux + balance_loss(gates[i], mask_topk[i], per)
    gates_s = [ht.reduce_sum_op(ht.mul_op(gates[i], mask_topk[i]), axes=1) for i in range(k)]
    location_s = []
    for i in range(k):
        loc = ht.cumsum_with_bias_op(mask_topk[i], bias=-1, dim=0)
        location_s.append(ht.reduce_sum_op(loc * mask_topk[i], axes=1))
    indices_s = [indices_s[i] + (i * per) if i > 0 else indices_s[i] for i in range(k)]
    return l_aux, indices_s, location_s, gates_s, capacity


def samgating(logits, k, capacity_factor, num_tokens, num_experts, embed_dim, num_local_gpus):
    """Switch-and-mix: pick the best GPU group first, then top-k inside it."""
    gates = ht.softmax_op(logits)
    capacity = k * math.ceil((num_tokens / num_experts) * capacity_factor)
    group_sum = ht.sam_group_sum_op(gates, num_local_gpus)
    top1_group = ht.topk_idx_op(group_sum, topk=1)
    topk_indices = ht.group_topk_idx_op(gates, top1_group, topk=k,
                                        num_local_gpus=num_experts // num_local_gpus)
    indices_s = [ht.array_reshape_op(ht.split_op(topk_indices, axes=[1], indices=[i], splits=[k]), [-1])
                 for i in range(k)]
    mask_topk = [ht.array_reshape_op(ht.one_hot_op(indices_s[i], num_classes=num_experts),
                                     [-1, num_experts]) for i in range(k)]
    l_aux = balance_loss(gates, mask_topk[0], num_experts)
    for i in range(1, k):
        l_aux = l_aux + balance_loss(gates, mask_topk[i], num_experts)
    l_alignment = alignment_loss(gates, top1_group, topk_indices, num_experts // num_local_gpus)
    location_s = _slots(mask_topk, k)
    gates_s = [ht.reduce_sum_op(ht.mul_op(gates, mask_topk[i]), axes=1) for i in range(k)]
    return l_aux, l_alignment, indices_s, location_s, gates_s, capacity


def hashgating(logits, capacity_factor, num_tokens, num_experts, embed_dim, indice):
    capacity = math.ceil((num_tokens / num_experts) * capacity_factor)
    indices_s = [indice]
    mask = ht.array_reshape_op(ht.one_hot_op(indice, num_classes=num_experts), [-1, num_experts])
    loc = ht.cumsum_with_bias_op(mask, bias=-1, dim=0)
    return indices_s, [ht.reduce_sum_op(loc * mask, axes=1)], capacity


class _LinearGate(BaseLayer):
    def __init__(self, embed_dim, num_tokens, num_experts, k=1, capacity_factor=1.0,
                 eval_capacity_factor=1.0, initializer=init.GenXavierUniform(), name="Gate"):
        self.embed_dim, self.num_experts, self.top_k = embed_dim, num_experts, k
        self.k = k
        self.num_tokens = num_tokens
        self.capacity_factor, self.eval_capacity_factor = capacity_factor, eval_capacity_factor
        self.initializer, self.name = initializer, name
        self.weight_var = self.bias_var = None

    def logits(self, x):
        if self.weight_var is None:
            self.weight_var = self.initializer(shape=(self.embed_dim, self.num_experts),
                                               name=self.name + '_linear_weight')
            self.bias_var = init.zeros(shape=(self.num_experts,), name=self.name + '_linear_bias')
        return ht.linear_op(x, self.weight_var, self.bias_var)


class TopKGate(_LinearGate):
    def __init__(self, embed_dim, num_tokens, num_experts, k=1, capacity_factor=1.0,
                 eval_capacity_factor=1.0, initializer=init.GenXavierUniform(), name="TopK_Gate"):
        super().__init__(embed_dim, num_tokens, num_experts, k, capacity_factor,
                         eval_capacity_factor, initializer, name)

    def __call__(self, x):
        return topkgating(self.logits(x), self.top_k, self.capacity_factor, self.num_tokens,
                          self.num_experts, self.embed_dim)


class KTop1Gate(_LinearGate):
    def __init__(self, embed_dim, num_tokens, num_experts, k=1, capacity_factor=1.0,
                 eval_capacity_factor=1.0, initializer=init.GenXavierUniform(), name="KTop1_Gate"):
        super().__init__(embed_dim, num_tokens, num_experts, k, capacity_factor,
                         eval_capacity_factor, initializer, name)

    def __call__(self, x):
        return ktop1gating(self.logits(x), self.k, self.capacity_factor, self.num_tokens,
                           self.num_experts, self.embed_dim)


class SAMGate(_LinearGate):
    def __init__(self, embed_dim, num_tokens, num_experts, k=1, capacity_factor=1.0,
                 eval_capacity_factor=1.0, initializer=init.GenXavierUniform(), name="TopK_Gate",
                 num_local_gpus=8):
        super().__init__(embed_dim, num_tokens, num_experts, k, capacity_factor,
                         eval_capacity_factor, initializer, name)
        self.num_local_gpus = num_local_gpus

    def __call__(self, x):
        return samgating(self.logits(x), self.top_k, self.capacity_factor, self.num_tokens,
                         self.num_experts, self.embed_dim, self.num_local_gpus)


class HashGate(BaseLayer):
    def __init__(self, embed_dim, num_tokens, num_experts, capacity_factor=1.0,
                 eval_capacity_factor=1.0, name="Hash_Gate"):
        self.embed_dim, self.num_experts, self.num_tokens = embed_dim, num_experts, num_tokens
        self.capacity_factor, self.eval_capacity_factor = capacity_factor, eval_capacity_factor
        self.name = name

    def __call__(self, x, indice):
        return hashgating(x, self.capacity_factor, self.num_tokens, self.num_experts,
                          self.embed_dim, indice)


def generate_orthogonal(input_shape, gain=0.1):
    rows = input_shape[0]
    cols = int(np.prod(input_shape[1:]))
    flat = np.random.normal(0, 1, (rows, cols))
    if rows < cols:
        flat = flat.T
    q, r = np.linalg.qr(flat)
    q *= np.sign(np.diag(r))
    if rows < cols:
        q = q.T
    return (q * gain).astype(np.float32)


class BalanceAssignmentGate(BaseLayer):
    """BASE-layer gate: balanced (auction) assignment to fixed orthogonal centroids."""

    def __init__(self, embed_dim, num_tokens, num_experts, capacity_factor=1.0,
                 eval_capacity_factor=1.0, initializer=init.GenXavierUniform(),
                 name="BalanceAssignment_Gate", device_id=None):
        self.embed_dim, self.num_experts, self.num_tokens = embed_dim, num_experts, num_tokens
        self.name, self.device_id = name, device_id or 0
        self.expert_centroids = Variable(value=generate_orthogonal((num_experts, embed_dim)),
                                         name='centroids', trainable=False)

    def __call__(self, x):
        scores = ht.matmul_op(x, self.expert_centroids, trans_B=True)
        indice = ht.balance_assignment_op(scores)
        centroid = ht.split_op(self.expert_centroids, axes=[0], indices=[self.device_id],
                               splits=[self.num_experts])
        return indice, ht.array_reshape_op(centroid, [-1, 1])


class MoELayer(BaseLayer):
    """gate -> dispatch -> all-to-all -> local experts -> all-to-all -> combine.

    ``name`` selects the routing family like the reference: 'MoELayer' (top-k with
    gate weights), 'KTop1Layer', 'SAMLayer', 'HashLayer' (no gate weights) and
    'BalanceAssignmentLayer' (BASE layer).  ``hierarchical=(num_nodes,
    num_local_gpus)`` uses ``halltoall_op`` for the two exchanges."""

    def __init__(self, gate=None, experts=None, num_tokens=None, embed_dim=None, all2all_size=None,
                 name='MoELayer', device_id=None, top=None, k=None, num_local_gpus=8,
                 hierarchical=None, fused=None):
        self.name, self.gate, self.experts = name, gate, experts
        self.num_local_experts = len(experts)
        self.num_tokens, self.embed_dim = num_tokens, embed_dim
        self.all2all_size = all2all_size
        self.device_id = device_id
        self.top = top if top is not None else k
        self.num_local_gpus = num_local_gpus
        self.hierarchical = hierarchical
        # fused peer-store dispatch/combine (ops/moe_fused.py) unless the NCCL baseline is requested
        if fused is None:
            import os
            fused = os.environ.get('HETU_MOE_FUSED', '1') == '1' and hierarchical is None
        self.fused = fused and name != 'BalanceAssignmentLayer'
        if name == 'BalanceAssignmentLayer':
            self.arange_array = Variable(value=np.arange(num_tokens).astype(np.float32),
                                         name='arange_array', trainable=False)

    def _plain_ffn(self):
        """Every local expert is the stock two-GEMM relu FFN without bias (what ops/moe_ffn.py computes itself)."""
        return (os.environ.get('HETU_MOE_OVERLAP', '1') == '1'
                and all(type(ex) is Expert and not ex.bias and ex.activation is ht.relu_op for ex in self.experts)
                and len({(ex.embed_dim, ex.ffn_dim) for ex in self.experts}) == 1
                and self.experts[0].embed_dim == self.embed_dim)

    def _a2a(self, x):
        if self.hierarchical is not None:
            return ht.halltoall_op(x, self.hierarchical[0], self.hierarchical[1])
        return ht.alltoall_op(x)

    def _experts(self, dispatched):
        x = ht.array_reshape_op(dispatched, [self.all2all_size, self.num_local_experts, -1,
                                             self.embed_dim])
        outs = []
        for i in range(self.num_local_experts):
            tok = ht.split_op(x, axes=[1], indices=[i], splits=[self.num_local_experts])
            outs.append(ht.array_reshape_op(self.experts[i](tok),
                                            [self.all2all_size, 1, -1, self.embed_dim]))
        y = ht.concatenate_op(outs, axis=1) if len(outs) > 1 else outs[0]
        return ht.array_reshape_op(y, [-1, self.embed_dim])

    def __call__(self, x, indice=None):
        total = self.num_local_experts * self.all2all_size
        reshaped = ht.array_reshape_op(x, [-1, self.embed_dim])
        if self.name == 'BalanceAssignmentLayer':
            return self._base_layer(reshaped)
        extra = None
        if self.name == 'HashLayer':
            indices_s, location_s, capacity = self.gate(reshaped, indice)
            gates_s, l_aux = None, None
        elif self.name == 'SAMLayer':
            l_aux, extra, indices_s, location_s, gates_s, capacity = self.gate(reshaped)
        else:
            l_aux, indices_s, location_s, gates_s, capacity = self.gate(reshaped)
        if self.fused and self._plain_ffn():
            # exchange chunks overlapped with the expert GEMMs, one op (ops/moe_ffn.py)
            from ..ops.moe_ffn import fused_moe_ffn_op
            for ex in self.experts:
                ex.build()
            out = fused_moe_ffn_op(reshaped, indices_s, location_s, gates_s, [ex.w1 for ex in self.experts],
                                   [ex.w2 for ex in self.experts], capacity, self.all2all_size)
            if gates_s is None:
                return out
            if extra is not None:
                return out, l_aux, extra
            return out, l_aux
        if self.fused:
            from ..ops.moe_fused import fused_moe_dispatch_op, fused_moe_combine_op
            E, W = self.num_local_experts, self.all2all_size
            recv = fused_moe_dispatch_op(reshaped, indices_s, location_s, capacity, E, W)
            outs = []
            for i in range(E):
                tok = ht.split_op(recv, axes=[0], indices=[i], splits=[E])          # contiguous slab: a view
                outs.append(ht.array_reshape_op(self.experts[i](tok), [1, W, -1, self.embed_dim]))
            y = ht.concatenate_op(outs, axis=0) if E > 1 else outs[0]
            out = fused_moe_combine_op(y, indices_s, location_s, capacity, E, W, gates=gates_s)
            if gates_s is None:
                return out
            if extra is not None:
                return out, l_aux, extra
            return out, l_aux
        dispatched = ht.layout_transform_op(reshaped, indices_s, location_s, capacity, total)
        dispatched = self._a2a(dispatched)
        y = self._experts(dispatched)
        y = self._a2a(y)
        if gates_s is None:
            out = ht.reverse_layout_transform_no_gate_op(y, indices_s, location_s, capacity, total)
            return out
        out = ht.reverse_layout_transform_op(y, indices_s, location_s, gates_s, capacity, total)
        if extra is not None:
            return out, l_aux, extra
        return out, l_aux

    def _base_layer(self, reshaped):
        indice, centroid = self.gate(reshaped)
        reverse_indice = ht.scatter1d_op(self.arange_array, indice)
        routed = ht.indexing_op(reshaped, indice)
        routed = self._a2a(routed)
        y = self._experts(routed)
        alpha = ht.sigmoid_op(ht.matmul_op(routed, centroid))
        alpha_b = ht.broadcastto_op(alpha, y)
        out = ht.mul_op(alpha_b, y) + ht.mul_op(ht.minus_byconst_op(alpha_b, 1.0), routed)
        out = self._a2a(out)
        return ht.indexing_op(out, reverse_indice)


def KTop1Layer(gate=None, experts=None, num_tokens=None, embed_dim=None, all2all_size=None,
               name='KTop1Layer', k=None):
    return MoELayer(gate, experts, num_tokens, embed_dim, all2all_size, name='KTop1Layer', top=k)


def SAMLayer(gate=None, experts=None, num_tokens=None, embed_dim=None, all2all_size=None,
             name='SAMLayer', k=None, num_local_gpus=8):
    return MoELayer(gate, experts, num_tokens, embed_dim, all2all_size, name='SAMLayer', top=k,
                    num_local_gpus=num_local_gpus)


def HashLayer(gate=None, experts=None, num_tokens=None, embed_dim=None, all2all_size=None,
              name='HashLayer'):
    return MoELayer(gate, experts, num_tokens, embed_dim, all2all_size, name='HashLayer', top=1)
