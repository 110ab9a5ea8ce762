"""Mixture-of-experts ops: dispatch / combine layout transforms and gate helpers.

reference op files: LayoutTransform ReverseLayoutTransform
ReverseLayoutTransformNoGate BalanceAssignment SamGroupSum GroupTopKIdx SamMax
(python/hetu/gpu_ops/*.py; kernels src/ops/LayoutTransform.cu:3-116 - 128 CTAs x
1024 threads, top-2 as two launches, combine via float atomics).

B200 path (csrc/moe.cu): one warp per token row, 128-bit copies, all k routes in
one launch.  Combine is a gather (each output row reads its k expert rows and
sums them in a fixed order), so it is deterministic and atomic-free.  Note: the
reference's top-2 combine launches its overwrite kernel twice, so only the
second expert's contribution survives (LayoutTransform.cu:246-247); here the
result is the documented sum over the k routes.

Tokens whose slot >= capacity or whose expert index < 0 are dropped exactly like
the reference (``location < capacity`` test, LayoutTransform.cu:5).
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from .. import _lib as L
from ..graph.node import Op, as_index

_PTR4 = ctypes.c_void_p * 4


def _k(*ts):
    return L.kernels_on(*ts) and all(t.dtype == torch.float32 for t in ts)


def _ptrs(ts):
    vals = [t.data_ptr() for t in ts] + [0] * (4 - len(ts))
    return _PTR4(*vals)


def moe_dispatch(x, idxs, locs, capacity, total_experts, gates=None):
    """out[idx*C + loc, :] = (gate *) x[t, :]   -> [E*C, d]"""
    T, d = x.shape
    out = torch.zeros((int(capacity * total_experts), d), dtype=x.dtype, device=x.device)
    if _k(x, *idxs, *locs) and len(idxs) <= 4:
        x = x.contiguous()
        idxs = [i.contiguous().reshape(-1) for i in idxs]
        locs = [l.contiguous().reshape(-1) for l in locs]
        g = [t.contiguous().reshape(-1) for t in gates] if gates is not None else []
        L.call('moe_dispatch', L.p(x), _ptrs(idxs), _ptrs(locs), _ptrs(g), L.c_int(len(idxs)),
               L.c_int(1 if gates is not None else 0), L.p(out), L.c_int(T), L.c_int(d),
               L.c_int(int(capacity)), L.c_int(int(total_experts)), L.stream_ptr(x.device))
        return out
    for j, (idx, loc) in enumerate(zip(idxs, locs)):
        idx, loc = as_index(idx).reshape(-1), as_index(loc).reshape(-1)
        ok = (loc < capacity) & (idx >= 0) & (loc >= 0)
        rows = (idx * capacity + loc)[ok]
        src = x[ok] if gates is None else x[ok] * gates[j].reshape(-1)[ok].unsqueeze(1)
        out.index_add_(0, rows, src)
    return out


def moe_combine(y, idxs, locs, capacity, gates=None):
    """out[t, :] = sum_k (gate_k[t] *) y[idx_k[t]*C + loc_k[t], :]   -> [T, d]"""
    T = idxs[0].numel()
    d = y.shape[-1]
    if _k(y, *idxs, *locs) and len(idxs) <= 4:
        y = y.contiguous()
        idxs = [i.contiguous().reshape(-1) for i in idxs]
        locs = [l.contiguous().reshape(-1) for l in locs]
        g = [t.contiguous().reshape(-1) for t in gates] if gates is not None else []
        out = torch.empty((T, d), dtype=y.dtype, device=y.device)
        L.call('moe_combine', L.p(y), _ptrs(idxs), _ptrs(locs), _ptrs(g), L.c_int(len(idxs)),
               L.c_int(1 if gates is not None else 0), L.p(out), L.c_int(T), L.c_int(d),
               L.c_int(int(capacity)), L.c_int(y.shape[0]), L.stream_ptr(y.device))
        return out
    out = torch.zeros((T, d), dtype=y.dtype, device=y.device)
    for j, (idx, loc) in enumerate(zip(idxs, locs)):
        idx, loc = as_index(idx).reshape(-1), as_index(loc).reshape(-1)
        ok = (loc < capacity) & (idx >= 0) & (loc >= 0)
        rows = (idx * capacity + loc).clamp(0, y.shape[0] - 1)
        contrib = y[rows]
        if gates is not None:
            contrib = contrib * gates[j].reshape(-1, 1)
        out = out + contrib * ok.unsqueeze(1).to(y.dtype)
    return out


def moe_gate_grad(dout, y, idx, loc, capacity):
    """dgate[t] = <dout[t,:], y[idx[t]*C + loc[t], :]>"""
    T, d = dout.shape
    if _k(dout, y, idx, loc):
        out = torch.empty((T,), dtype=dout.dtype, device=dout.device)
        L.call('moe_gate_grad', L.p(dout.contiguous()), L.p(y.contiguous()), L.p(idx.contiguous()),
               L.p(loc.contiguous()), L.p(out), L.c_int(T), L.c_int(d), L.c_int(int(capacity)),
               L.c_int(y.shape[0]), L.stream_ptr(dout.device))
        return out
    idx, loc = as_index(idx).reshape(-1), as_index(loc).reshape(-1)
    ok = (loc < capacity) & (idx >= 0) & (loc >= 0)
    rows = (idx * capacity + loc).clamp(0, y.shape[0] - 1)
    return (dout * y[rows]).sum(dim=1) * ok.to(dout.dtype)


class LayoutTransformOp(Op):
    def __init__(self, input, indices_s, location_s, capacity, total_experts, ctx=None):
        super().__init__(LayoutTransformOp, [input] + list(indices_s) + list(location_s), ctx)
        self.capacity, self.topK, self.total_experts = capacity, len(indices_s), total_experts

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        k = self.topK
        if x.device.type == 'meta':
            return torch.empty((int(self.capacity * self.total_experts), x.shape[-1]), device='meta')
        return moe_dispatch(x, input_vals[1:1 + k], input_vals[1 + k:1 + 2 * k], self.capacity,
                            self.total_experts)

    def gradient(self, g):
        k = self.topK
        grad = LayoutTransformGradientOp(g, self.inputs[1:1 + k], self.inputs[1 + k:1 + 2 * k],
                                         self.capacity, ctx=self.raw_ctx)
        return [grad] + [None] * (2 * k)

    def infer_shape(self, input_shapes):
        return (int(self.capacity * self.total_experts), input_shapes[0][-1])


class LayoutTransformGradientOp(Op):
    def __init__(self, input, indice, location, capacity, ctx=None):
        indice = list(indice) if isinstance(indice, (list, tuple)) else [indice]
        location = list(location) if isinstance(location, (list, tuple)) else [location]
        super().__init__(LayoutTransformGradientOp, [input] + indice + location, ctx)
        self.capacity, self.topK = capacity, len(indice)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        g = input_vals[0]
        k = self.topK
        if g.device.type == 'meta':
            return torch.empty((input_vals[1].shape[0], g.shape[-1]), device='meta')
        return moe_combine(g, input_vals[1:1 + k], input_vals[1 + k:1 + 2 * k], self.capacity)

    def infer_shape(self, input_shapes):
        return (input_shapes[1][0], input_shapes[0][-1])


class ReverseLayoutTransformOp(Op):
    def __init__(self, input, indices_s, location_s, gates, capacity, num_experts, ctx=None):
        super().__init__(ReverseLayoutTransformOp,
                         [input] + list(indices_s) + list(location_s) + list(gates), ctx)
        self.capacity, self.topK, self.num_experts = capacity, len(indices_s), num_experts

    def compute(self, input_vals, output_val=None, stream_handle=None):
        y = input_vals[0]
        k = self.topK
        if y.device.type == 'meta':
            return torch.empty((input_vals[1].shape[0], y.shape[-1]), device='meta')
        return moe_combine(y, input_vals[1:1 + k], input_vals[1 + k:1 + 2 * k], self.capacity,
                           gates=input_vals[1 + 2 * k:1 + 3 * k])

    def gradient(self, g):
        k = self.topK
        ins = self.inputs
        idxs, locs, gates = ins[1:1 + k], ins[1 + k:1 + 2 * k], ins[1 + 2 * k:1 + 3 * k]
        gd = reverse_layout_transform_gradient_data_op(g, idxs, locs, gates, self.capacity,
                                                       self.num_experts, ctx=self.raw_ctx)
        gg = [reverse_layout_transform_gradient_gate_op(g, ins[0], idxs[j], locs[j], self.capacity,
                                                        ctx=self.raw_ctx) for j in range(k)]
        return [gd] + [None] * (2 * k) + gg

    def infer_shape(self, input_shapes):
        return (input_shapes[1][0], input_shapes[0][-1])


class ReverseLayoutTransformGradientDataOp(Op):
    def __init__(self, input, indices, locations, gates, capacity, num_experts, ctx=None):
        super().__init__(ReverseLayoutTransformGradientDataOp,
                         [input] + list(indices) + list(locations) + list(gates), ctx)
        self.capacity, self.topK, self.num_experts = capacity, len(indices), num_experts

    def compute(self, input_vals, output_val=None, stream_handle=None):
        g = input_vals[0]
        k = self.topK
        if g.device.type == 'meta':
            return torch.empty((int(self.capacity * self.num_experts), g.shape[-1]), device='meta')
        return moe_dispatch(g, input_vals[1:1 + k], input_vals[1 + k:1 + 2 * k], self.capacity,
                            self.num_experts, gates=input_vals[1 + 2 * k:1 + 3 * k])

    def infer_shape(self, input_shapes):
        return (int(self.capacity * self.num_experts), input_shapes[0][-1])


class ReverseLayoutTransformGradientGateOp(Op):
    def __init__(self, combined_output, expert_output, indices, locations, capacity, ctx=None):
        super().__init__(ReverseLayoutTransformGradientGateOp,
                         [combined_output, expert_output, indices, locations], ctx)
        self.capacity = capacity

    def compute(self, input_vals, output_val=None, stream_handle=None):
        dout, y, idx, loc = input_vals
        if dout.device.type == 'meta':
            return torch.empty((dout.shape[0],), device='meta')
        return moe_gate_grad(dout, y, idx.reshape(-1), loc.reshape(-1), self.capacity)

    def infer_shape(self, input_shapes):
        return (input_shapes[0][0],)


class ReverseLayoutTransformNoGateOp(Op):
    def __init__(self, input, indices_s, location_s, capacity, num_experts, ctx=None):
        super().__init__(ReverseLayoutTransformNoGateOp,
                         [input] + list(indices_s) + list(location_s), ctx)
        self.capacity, self.topK, self.num_experts = capacity, len(indices_s), num_experts

    def compute(self, input_vals, output_val=None, stream_handle=None):
        y = input_vals[0]
        k = self.topK
        if y.device.type == 'meta':
            return torch.empty((input_vals[1].shape[0], y.shape[-1]), device='meta')
        return moe_combine(y, input_vals[1:1 + k], input_vals[1 + k:1 + 2 * k], self.capacity)

    def gradient(self, g):
        k = self.topK
        gd = reverse_layout_transform_no_gate_gradient_op(
            g, self.inputs[1:1 + k], self.inputs[1 + k:1 + 2 * k], self.capacity, self.num_experts,
            ctx=self.raw_ctx)
        return [gd] + [None] * (2 * k)

    def infer_shape(self, input_shapes):
        return (input_shapes[1][0], input_shapes[0][-1])


class ReverseLayoutTransformNoGateGradientOp(Op):
    def __init__(self, input, indices, locations, capacity, num_experts, ctx=None):
        indices = list(indices) if isinstance(indices, (list, tuple)) else [indices]
        locations = list(locations) if isinstance(locations, (list, tuple)) else [locations]
        super().__init__(ReverseLayoutTransformNoGateGradientOp, [input] + indices + locations, ctx)
        self.capacity, self.topK, self.num_experts = capacity, len(indices), num_experts

    def compute(self, input_vals, output_val=None, stream_handle=None):
        g = input_vals[0]
        k = self.topK
        if g.device.type == 'meta':
            return torch.empty((int(self.capacity * self.num_experts), g.shape[-1]), device='meta')
        return moe_dispatch(g, input_vals[1:1 + k], input_vals[1 + k:1 + 2 * k], self.capacity,
                            self.num_experts)

    def infer_shape(self, input_shapes):
        return (int(self.capacity * self.num_experts), input_shapes[0][-1])


# ------------------------------------------------------------------ gate helpers
class BalanceAssignmentOp(Op):
    """BASE-layer balanced assignment (auction algorithm): scores [T, E] ->
    a permutation of the T tokens such that consecutive blocks of T/E tokens go
    to experts 0..E-1 with maximal total score.  reference
    BalanceAssignment.py:11-84 drives the bidding loop from the host; here the
    loop runs on device tensors with a fixed iteration cap."""
    graph_safe = False

    def __init__(self, node, ctx=None):
        super().__init__(BalanceAssignmentOp, [node], ctx)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        scores = input_vals[0]
        if scores.device.type == 'meta':
            return torch.empty((scores.shape[0],), device='meta')
        return balanced_assignment(scores).to(torch.float32)

    def gradient(self, g):
        return [None]

    def infer_shape(self, input_shapes):
        return (input_shapes[0][0],)


def balanced_assignment(scores: torch.Tensor, max_iterations: int = 100) -> torch.Tensor:
    """Auction algorithm (Bertsekas) for the balanced token->expert assignment."""
    T, E = scores.shape
    jobs_per_worker = T // E
    value = scores.clone()
    eps = (scores.max() - scores.min()) / 50
    eps = torch.clamp(eps, min=1e-4)
    cost = scores.new_zeros(1, E)
    iterations = 0
    while True:
        top_values, top_index = torch.topk(value, jobs_per_worker + 1 if jobs_per_worker < T else T, dim=0)
        bid_increments = top_values[:jobs_per_worker] - top_values[-1:] + eps
        bids = torch.scatter(torch.zeros(T, E, device=scores.device, dtype=scores.dtype), 0,
                             top_index[:jobs_per_worker], bid_increments)
        if 0 < iterations < max_iterations:
            bids[top_bidders, jobs_with_bids] = eps
        top_bids, top_bidders = bids.max(dim=1)
        jobs_with_bids = top_bids > 0
        top_bidders = top_bidders[jobs_with_bids]
        cost = cost + torch.scatter_add(torch.zeros_like(cost), 1, top_bidders.unsqueeze(0),
                                        top_bids[jobs_with_bids].unsqueeze(0))
        value = scores - cost
        if bool(jobs_with_bids.all()) or iterations >= max_iterations:
            break
        value[jobs_with_bids, top_bidders] = float('inf')
        iterations += 1
    if not bool(jobs_with_bids.all()):
        # fallback: greedy fill keeps the result a valid permutation
        order = torch.argsort(scores.max(dim=1).values, descending=True)
        return order
    return top_index[:jobs_per_worker].t().reshape(-1)


class SamGroupSumOp(Op):
    """[T, E] -> [T, G]: sum of the gate values of the experts of each GPU group."""

    def __init__(self, node, num_local_gpus, ctx=None):
        super().__init__(SamGroupSumOp, [node], ctx)
        self.num_local_gpus = num_local_gpus

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x = input_vals[0]
        g = self.num_local_gpus
        return x.reshape(x.shape[0], g, x.shape[1] // g).sum(dim=2)

    def gradient(self, g):
        return [None]


class GroupTopKIdxOp(Op):
    """top-k expert indices restricted to the token's chosen group
    (columns [group*G, (group+1)*G) like src/ops/GroupTopKIdx.cu)."""

    def __init__(self, node_A, node_B, topk, num_local_gpus, ctx=None):
        super().__init__(GroupTopKIdxOp, [node_A, node_B], ctx)
        self.k, self.num_local_gpus = topk, num_local_gpus

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x, grp = input_vals
        if x.device.type == 'meta':
            return torch.empty((x.shape[0], self.k), device='meta')
        G = self.num_local_gpus
        group = as_index(grp).reshape(-1)
        cols = torch.arange(x.shape[1], device=x.device).unsqueeze(0)
        lo = (group * G).unsqueeze(1)
        masked = torch.where((cols >= lo) & (cols < lo + G), x, torch.full_like(x, -1e30))
        return torch.topk(masked, self.k, dim=1).indices.to(torch.float32)

    def gradient(self, g):
        return [None, None]


class SamMaxOp(Op):
    """SAM alignment helper: for every token, hinge of out-of-group gate values
    above the k-th in-group gate value: out[t, e] = max(0, g[t,e] - g[t, topk_last])
    for experts e outside the token's group, 0 inside (src/ops/SamMax.cu)."""

    def __init__(self, node_A, node_B, node_C, num_local_gpus, ctx=None):
        super().__init__(SamMaxOp, [node_A, node_B, node_C], ctx)
        self.num_local_gpus = num_local_gpus

    def compute(self, input_vals, output_val=None, stream_handle=None):
        gates, grp, topk_idx = input_vals
        if gates.device.type == 'meta':
            return torch.empty_like(gates)
        G = self.num_local_gpus
        group = as_index(grp).reshape(-1)
        cols = torch.arange(gates.shape[1], device=gates.device).unsqueeze(0)
        lo = (group * G).unsqueeze(1)
        inside = (cols >= lo) & (cols < lo + G)
        last = as_index(topk_idx)[:, -1:]
        thr = torch.gather(gates, 1, last)
        return torch.where(inside, torch.zeros_like(gates), (gates - thr).clamp_min(0))

    def gradient(self, g):
        return [SamMaxGradOp(g, self, ctx=self.raw_ctx), None, None]


class SamMaxGradOp(Op):
    def __init__(self, grad, forward_node, ctx=None):
        super().__init__(SamMaxGradOp, [grad, forward_node] + forward_node.inputs, ctx)
        self.num_local_gpus = forward_node.num_local_gpus

    def compute(self, input_vals, output_val=None, stream_handle=None):
        g, out, gates, grp, topk_idx = input_vals
        if g.device.type == 'meta':
            return torch.empty_like(gates)
        act = (out > 0).to(g.dtype)
        dg = g * act
        last = as_index(topk_idx)[:, -1:]
        return dg.scatter_add(1, last, -(dg.sum(dim=1, keepdim=True)))

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[2])


# ---------------------------------------------------------------- public API
def layout_transform_op(input, indices_s, location_s, capacity, total_experts, ctx=None):
    return LayoutTransformOp(input, indices_s, location_s, capacity, total_experts, ctx=ctx)


def layout_transform_gradient_op(input, indice, location, capacity, ctx=None):
    return LayoutTransformGradientOp(input, indice, location, capacity, ctx=ctx)


def reverse_layout_transform_op(input, indices_s, location_s, gates, capacity, num_experts, ctx=None):
    return ReverseLayoutTransformOp(input, indices_s, location_s, gates, capacity, num_experts, ctx=ctx)


def reverse_layout_transform_gradient_data_op(input, indices, locations, gates, capacity,
                                              num_experts, ctx=None):
    return ReverseLayoutTransformGradientDataOp(input, indices, locations, gates, capacity,
                                                num_experts, ctx=ctx)


def reverse_layout_transform_gradient_gate_op(combined_output, expert_output, indices, locations,
                                              capacity, ctx=None):
    return ReverseLayoutTransformGradientGateOp(combined_output, expert_output, indices, locations,
                                                capacity, ctx=ctx)


def reverse_layout_transform_no_gate_op(input, indices_s, location_s, capacity, num_experts, ctx=None):
    return ReverseLayoutTransformNoGateOp(input, indices_s, location_s, capacity, num_experts, ctx=ctx)


def reverse_layout_transform_no_gate_gradient_op(input, indices, locations, capacity, num_experts,
                                                 ctx=None):
    return ReverseLayoutTransformNoGateGradientOp(input, indices, locations, capacity, num_experts,
                                                  ctx=ctx)


def balance_assignment_op(node, ctx=None):
    return BalanceAssignmentOp(node, ctx=ctx)


def sam_group_sum_op(node, num_local_gpus, ctx=None):
    return SamGroupSumOp(node, num_local_gpus, ctx=ctx)


def group_topk_idx_op(node_A, node_B, topk, num_local_gpus, ctx=None):
    return GroupTopKIdxOp(node_A, node_B, topk, num_local_gpus, ctx=ctx)


def sam_max_op(node_A, node_B, node_C, num_local_gpus, ctx=None):
    return SamMaxOp(node_A, node_B, node_C, num_local_gpus, ctx=ctx)
