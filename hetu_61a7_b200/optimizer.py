"""Optimizers and the OptimizerOp graph node.

reference: python/hetu/optimizer.py:13-555 (one kernel launch per parameter
tensor: src/ops/Optimizers.cu:39-70; sparse rows: src/ops/OptimizersSparse.cu
with the unique done on host in numpy, ndarray.py:535-552).

B200 design:
 * dense step = ONE multi-tensor launch per optimizer (``hetu_multi_tensor_opt``:
   grid.y indexes the tensor, pointers travel by value in the launch params),
   hyper-parameters that change per step (lr, beta^t) are read from device
   memory so the step can live inside a replayed CUDA graph;
 * data-parallel: grads are packed into the symmetric heap and reduced +
   applied by one fused kernel (parallel/symm.py ``fused_allreduce_update``)
   instead of one ncclAllReduce + one update per tensor;
 * sparse (IndexedSlices) step = device sort + segmented reduce (no host
   unique) + fused row update; SGD needs no dedup at all (atomic row adds).
"""
from __future__ import annotations

import ctypes
from copy import copy, deepcopy
from typing import List, Optional

import numpy as np
import torch

from . import _lib as L
from .graph.node import Op
from .lr_scheduler import FixedScheduler
from .ndarray import IndexedSlices, NDArray
from .ops.variable import PlaceholderOp

OPT_SGD, OPT_MOMENTUM, OPT_NESTEROV, OPT_ADAGRAD, OPT_ADAM, OPT_ADAMW, OPT_LAMB = range(7)
_MAX_PER_LAUNCH = 48
_PTRS = ctypes.c_void_p * _MAX_PER_LAUNCH
_SIZES = ctypes.c_int64 * _MAX_PER_LAUNCH


class Optimizer(object):
    """Base class.  ``minimize`` differentiates the loss and returns an OptimizerOp."""
    name = 'Optimizer'
    kind = OPT_SGD
    n_states = 0

    def __init__(self, learning_rate, l2reg=0):
        if isinstance(learning_rate, FixedScheduler):
            self.lr_sched = learning_rate
        else:
            assert learning_rate >= 0, "learning rate must be non-negative"
            self.lr_sched = FixedScheduler(learning_rate)
        assert l2reg >= 0, 'L2 regularizer should be positive or 0.'
        self.l2reg = l2reg
        self.params: Optional[List[PlaceholderOp]] = None
        self.tensors: Optional[List[torch.Tensor]] = None
        self.states: List[List[torch.Tensor]] = []
        self.initiated = False
        self.step_t = None      # device int64 step counter (1-based after first update)
        self.lr_t = None        # device fp32 learning rate
        self._lr_host = None

    @property
    def learning_rate(self):
        return self.lr_sched.get()

    @staticmethod
    def get_var_list(loss):
        visited, out = set(), []
        stack = list(loss) if isinstance(loss, list) else [loss]
        order = []
        # iterative DFS, pre-order parity with the reference is irrelevant: keep
        # discovery order stable so ranks agree on the parameter order.
        def visit(node):
            if node in visited:
                return
            visited.add(node)
            if isinstance(node, PlaceholderOp) and node.trainable:
                out.append(node)
            for n in node.inputs:
                visit(n)
        import sys
        old = sys.getrecursionlimit()
        sys.setrecursionlimit(max(old, 100000))
        try:
            for l in stack:
                visit(l)
        finally:
            sys.setrecursionlimit(old)
        return out

    # -- state ------------------------------------------------------------------
    def initiate_states(self, config):
        assert not self.initiated, "Optimizer already initiated."
        self.tensors = [config.placeholder_to_arr_map.get(node) for node in self.params]
        self.states = []
        for t in self.tensors:
            if isinstance(t, torch.Tensor):
                self.states.append([self._new_state(t, i) for i in range(self.n_states)])
            else:
                self.states.append([])
        dev = next((t.device for t in self.tensors if isinstance(t, torch.Tensor)),
                   config.context.torch_device)
        self.step_t = torch.zeros(1, dtype=torch.int64, device=dev)
        self.lr_t = torch.full((1,), float(self.learning_rate), dtype=torch.float32, device=dev)
        self._lr_host = float(self.learning_rate)
        self.initiated = True

    def _new_state(self, t, i):
        return torch.zeros_like(t)

    def uninitiate_states(self):
        self.tensors, self.states = [], []
        self.initiated = False

    def update_tensors_version(self, tensor_map):
        self.tensors = [tensor_map[node] for node in self.params]

    def sync_lr(self):
        """Host-side, outside any captured graph: push a changed lr to the device."""
        lr = float(self.learning_rate)
        if self.lr_t is not None and lr != self._lr_host:
            self.lr_t.fill_(lr)
            self._lr_host = lr

    def state_dict(self):
        """Optimizer slots + step (the reference does not checkpoint these: SURVEY 5.4)."""
        out = {'step': int(self.step_t.item()) if self.step_t is not None else 0}
        for p, sts in zip(self.params or [], self.states):
            for i, s in enumerate(sts):
                out['%s/slot%d' % (p.name, i)] = s.detach().cpu().numpy()
        return out

    def load_state_dict(self, sd):
        if self.step_t is not None and 'step' in sd:
            self.step_t.fill_(int(sd['step']))
        for p, sts in zip(self.params or [], self.states):
            for i, s in enumerate(sts):
                key = '%s/slot%d' % (p.name, i)
                if key in sd:
                    s.copy_(torch.as_tensor(sd[key]).to(s.device))

    # -- graph construction -----------------------------------------------------------
    def minimize(self, loss, var_list=None):
        from .graph.autodiff import gradients
        self.loss = loss
        if not var_list:
            var_list = self.get_var_list(loss)
        self.params = list(var_list)
        grads, self.backward2forward, self.forward2backward = gradients(
            loss, self.params, return_all=True)
        return OptimizerOp(grads, self)

    def __deepcopy__(self, memo):
        assert not self.initiated, 'Should not deep copy optimizer if already initiated!'
        new_opt = copy(self)
        new_opt.loss = deepcopy(self.loss, memo)
        new_opt.params = [deepcopy(node, memo) for node in self.params]
        return new_opt

    def get_config(self):
        """(type, lr, hyper...) tuple for a server-side optimizer (PS function ParamInit)."""
        return (self.kind, [float(self.learning_rate)] + [float(h) for h in self.hyper()])

    def hyper(self):
        """[h0, h1, h2, h3] passed to the kernels (meaning depends on ``kind``)."""
        return [0.0, 0.0, 0.0, 0.0]

    # -- the step -------------------------------------------------------------------------
    def update(self, grads, stream_handle=None):
        """Apply one step given the gradient of every parameter (None = skip)."""
        dense_idx, sparse_idx = [], []
        for i, g in enumerate(grads):
            if g is None or self.tensors[i] is None:
                continue
            (sparse_idx if isinstance(g, IndexedSlices) else dense_idx).append(i)
        if dense_idx:
            dev = self.tensors[dense_idx[0]].device
            if self.step_t.device != dev:
                self.step_t, self.lr_t = self.step_t.to(dev), self.lr_t.to(dev)
        self._bump_step()
        if dense_idx:
            ps = [self.tensors[i] for i in dense_idx]
            gs = [grads[i] for i in dense_idx]
            sts = [self.states[i] for i in dense_idx]
            if L.kernels_on(*ps, *gs) and all(p.dtype == torch.float32 for p in ps):
                self._dense_kernel(ps, gs, sts)
            else:
                for p, g, st in zip(ps, gs, sts):
                    self._dense_torch(p, g.to(p.device), st)
        for i in sparse_idx:
            self._sparse_update(self.tensors[i], grads[i], self.states[i])

    def _bump_step(self):
        if self.step_t.is_cuda and L.cuda_lib() is not None:
            L.call('increment_i64', L.p(self.step_t), L.stream_ptr(self.step_t.device))
        else:
            self.step_t += 1

    def _dense_kernel(self, ps, gs, sts):
        h = self.hyper()
        dev = ps[0].device
        scratch = None
        if self.kind == OPT_LAMB:
            scratch = torch.zeros(2 * len(ps), dtype=torch.float32, device=dev)
        for s in range(0, len(ps), _MAX_PER_LAUNCH):
            chunk = slice(s, s + _MAX_PER_LAUNCH)
            cp, cg, cs = ps[chunk], [g.contiguous() for g in gs[chunk]], sts[chunk]
            n = len(cp)
            pad = [0] * (_MAX_PER_LAUNCH - n)
            a_p = _PTRS(*([t.data_ptr() for t in cp] + pad))
            a_g = _PTRS(*([t.data_ptr() for t in cg] + pad))
            a_s1 = _PTRS(*([st[0].data_ptr() if len(st) > 0 else 0 for st in cs] + pad))
            a_s2 = _PTRS(*([st[1].data_ptr() if len(st) > 1 else 0 for st in cs] + pad))
            a_n = _SIZES(*([t.numel() for t in cp] + pad))
            L.call('multi_tensor_opt', L.c_int(self.kind), L.c_int(n), a_p, a_g, a_s1, a_s2, a_n,
                   L.p(self.lr_t), L.p(self.step_t), L.c_f32(h[0]), L.c_f32(h[1]), L.c_f32(h[2]),
                   L.c_f32(h[3]), L.c_f32(float(self.l2reg)),
                   L.p(scratch[2 * s:] if scratch is not None else None), L.stream_ptr(dev))

    def _dense_torch(self, p, g, st):
        raise NotImplementedError

    # sparse -------------------------------------------------------------------------------
    def _sparse_update(self, p, slices: IndexedSlices, st):
        ind, val = slices._iv()
        if hasattr(p, 'apply_sparse_grad'):            # sharded / cached table handle
            p.apply_sparse_grad(ind, val, self)
            return
        ind, val = ind.to(p.device), val.to(p.device)
        if L.kernels_on(p, val) and p.dtype == torch.float32:
            h = self.hyper()
            width = p.shape[-1]
            if self.kind == OPT_SGD:
                # linear update: duplicates may be applied independently (atomic row add)
                L.call('sparse_opt', L.c_int(self.kind), L.p(p), L.p(ind), L.p(val.contiguous()),
                       L.p(None), L.p(None), L.c_sz(ind.numel()), L.c_int(width), L.p(self.lr_t),
                       L.p(self.step_t), L.c_f32(0), L.c_f32(0), L.c_f32(0), L.c_f32(0),
                       L.c_f32(float(self.l2reg)), L.c_int(1), L.stream_ptr(p.device))
                return
            sidx, perm = torch.sort(ind)
            dval = torch.empty_like(val)
            didx = torch.empty_like(sidx)
            L.call('dedup_sorted_rows', L.p(sidx), L.p(perm), L.p(val.contiguous()), L.p(didx),
                   L.p(dval), L.c_sz(sidx.numel()), L.c_int(width), L.stream_ptr(p.device))
            L.call('sparse_opt', L.c_int(self.kind), L.p(p), L.p(didx), L.p(dval),
                   L.p(st[0] if len(st) > 0 else None), L.p(st[1] if len(st) > 1 else None),
                   L.c_sz(didx.numel()), L.c_int(width), L.p(self.lr_t), L.p(self.step_t),
                   L.c_f32(h[0]), L.c_f32(h[1]), L.c_f32(h[2]), L.c_f32(h[3]),
                   L.c_f32(float(self.l2reg)), L.c_int(0), L.stream_ptr(p.device))
            return
        uniq, inv = torch.unique(ind, return_inverse=True)
        acc = torch.zeros((uniq.numel(), val.shape[1]), dtype=val.dtype, device=val.device)
        acc.index_add_(0, inv, val)
        rows = p[uniq]
        row_states = [s[uniq] for s in st]
        self._dense_torch(rows, acc, row_states)
        p[uniq] = rows
        for s, rs in zip(st, row_states):
            s[uniq] = rs


class SGDOptimizer(Optimizer):
    name, kind, n_states = 'SGD', OPT_SGD, 0

    def __init__(self, learning_rate=0.01, l2reg=0):
        super().__init__(learning_rate, l2reg)

    def _dense_torch(self, p, g, st):
        if p.device.type == 'cpu' and p.dtype == torch.float32 and p.is_contiguous() and g.shape == p.shape \
                and g.dtype == torch.float32:
            from .ops.elementwise import _cpu_native
            if _cpu_native():                          # CPU executor: OpenMP library (csrc_cpu)
                from . import cpu_links
                cpu_links.sgd_update(p, g.contiguous(), float(self.lr_t), float(self.l2reg))
                return
        if self.l2reg > 0:
            g = g + self.l2reg * p
        p.sub_(self.lr_t.to(p.device) * g)


class MomentumOptimizer(Optimizer):
    name, n_states = 'Momentum', 1

    def __init__(self, learning_rate=0.01, momentum=0.9, nesterov=False, l2reg=0):
        super().__init__(learning_rate, l2reg)
        self.momentum, self.nesterov = momentum, nesterov
        self.kind = OPT_NESTEROV if nesterov else OPT_MOMENTUM

    def hyper(self):
        return [self.momentum, 0.0, 0.0, 0.0]

    def _dense_torch(self, p, g, st):
        v = st[0]
        lr = self.lr_t.to(p.device)
        if self.l2reg > 0:
            g = g + self.l2reg * p
        if self.nesterov:
            temp = lr * g
            v.copy_(self.momentum * (v - temp))
            p.add_(v - temp)
        else:
            v.copy_(self.momentum * v - lr * g)
            p.add_(v)


class AdaGradOptimizer(Optimizer):
    name, kind, n_states = 'AdaGrad', OPT_ADAGRAD, 1

    def __init__(self, learning_rate=0.01, initial_accumulator_value=0.0, eps=1e-7, l2reg=0):
        assert initial_accumulator_value >= 0.0, "initial accumulator value must be non-negative"
        assert eps > 0.0, "epsilon must be positive"
        super().__init__(learning_rate, l2reg)
        self.initial_accumulator_value, self.eps = initial_accumulator_value, eps

    def _new_state(self, t, i):
        return torch.full_like(t, float(self.initial_accumulator_value))

    def hyper(self):
        return [self.eps, 0.0, 0.0, 0.0]

    def _dense_torch(self, p, g, st):
        acc = st[0]
        if self.l2reg > 0:
            g = g + self.l2reg * p
        acc.add_(g * g)
        p.sub_(self.lr_t.to(p.device) * g / (acc.sqrt() + self.eps))


class AdamOptimizer(Optimizer):
    name, kind, n_states = 'Adam', OPT_ADAM, 2

    def __init__(self, learning_rate=0.01, beta1=0.9, beta2=0.999, epsilon=1e-7, l2reg=0):
        super().__init__(learning_rate, l2reg)
        self.beta1, self.beta2, self.epsilon = beta1, beta2, epsilon
        self.weight_decay = 0.0

    def hyper(self):
        return [self.beta1, self.beta2, self.epsilon, self.weight_decay]

    def _moments(self, p, g, st):
        m, v = st
        if self.l2reg > 0:
            g = g + self.l2reg * p
        t = self.step_t.to(p.device).to(torch.float32)
        m.copy_(self.beta1 * m + (1 - self.beta1) * g)
        v.copy_(self.beta2 * v + (1 - self.beta2) * g * g)
        mhat = m / (1 - self.beta1 ** t)
        vhat = v / (1 - self.beta2 ** t)
        return mhat / (vhat.sqrt() + self.epsilon)

    def _dense_torch(self, p, g, st):
        p.sub_(self.lr_t.to(p.device) * self._moments(p, g, st))


class AdamWOptimizer(AdamOptimizer):
    name, kind = 'AdamW', OPT_ADAMW

    def __init__(self, learning_rate=0.01, beta1=0.9, beta2=0.999, epsilon=1e-7, weight_decay=0):
        super().__init__(learning_rate, beta1, beta2, epsilon, 0)
        self.weight_decay = weight_decay

    def _dense_torch(self, p, g, st):
        upd = self._moments(p, g, st)
        p.sub_(self.lr_t.to(p.device) * (upd + self.weight_decay * p))


class LambOptimizer(AdamOptimizer):
    name, kind = 'Lamb', OPT_LAMB

    def __init__(self, learning_rate=0.01, beta1=0.9, beta2=0.999, epsilon=1e-7, weight_decay=0):
        super().__init__(learning_rate, beta1, beta2, epsilon, 0)
        self.weight_decay = weight_decay

    def _dense_torch(self, p, g, st):
        upd = self._moments(p, g, st)
        n_p = torch.linalg.vector_norm(p)
        n_u = torch.linalg.vector_norm(upd)
        p.sub_(self.lr_t.to(p.device) * (n_p / n_u) * (upd + self.weight_decay * p))


class OptimizerOp(Op):
    """Graph node whose inputs are the gradients; computing it applies the step."""

    def __init__(self, grads, optimizer):
        super().__init__(OptimizerOp, grads, None)
        self.name = "Optimizer_%s" % (optimizer.name)
        self.optimizer = optimizer
        self.comm_mode = None
        self.fused_dp = None      # set by backward_hook when the symm engine takes the all-reduce
        self.skip_mask = None

    def compute(self, input_vals, output_val=None, stream_handle=None, new_tensors_map=None):
        if new_tensors_map is not None:
            self.optimizer.update_tensors_version(new_tensors_map)
        if self.comm_mode == 'PS':
            return None
        grads = list(input_vals)
        if self.fused_dp is not None:
            self.fused_dp.step(self.optimizer, grads)
        else:
            self.optimizer.update(grads, stream_handle)
        return None

    def gradient(self, output_grad):
        raise NotImplementedError

    def infer_shape(self, input_shapes):
        return None

    def forward_hook(self, config):
        self.optimizer.initiate_states(config)
        self.comm_mode = config.comm_mode
        self.ctx = config.context

    def backward_hook(self, config):
        """Wrap gradients with their synchronisation: all-reduce for dense DP,
        parameter-server push/pull for PS-resident (embedding) parameters.
        Expert parameters (name contains 'expert') stay local (optimizer.py:150-152)."""
        from .ops.comm import allreduceCommunicate_op
        self.comm_mode = config.comm_mode
        new_inputs = []
        fused_ids = []
        for i, node in enumerate(self.inputs):
            cur_param = self.optimizer.params[i]
            strategy = config.node_strategy.get(cur_param, self.comm_mode)
            if strategy is None:
                strategy = self.comm_mode
            cur_node = node
            if node is None or 'expert' in cur_param.name:
                pass
            elif getattr(cur_param, 'dp_group_ranks', None) is not None:
                # placement pass decided the replica group of this parameter (DP x TP x PP)
                ranks = cur_param.dp_group_ranks
                if len(ranks) > 1:
                    from .parallel.comm import new_group_comm
                    cur_node = allreduceCommunicate_op(node, new_group_comm(list(ranks)))
            elif strategy == 'AllReduce' or (strategy == 'Hybrid' and not node.use_indexed_slices):
                if config.symm is not None and not node.use_indexed_slices and config.fuse_dp_update:
                    fused_ids.append(i)
                else:
                    cur_node = allreduceCommunicate_op(node, config.nccl_comm)
            elif strategy == 'PS' or (strategy == 'Hybrid' and node.use_indexed_slices):
                if config.ps_comm is not None:
                    from .ops.ps_ops import parameterServerCommunicate_op
                    cur_node = parameterServerCommunicate_op(node, cur_param, self.optimizer.get_config())
                else:
                    from .parallel.sharded_embedding import should_shard
                    if config.nccl_comm is not None and config.nccl_comm.multi and not (
                            getattr(cur_param, 'sharded', False) or should_shard(cur_param, config)):
                        cur_node = allreduceCommunicate_op(node, config.nccl_comm)
            new_inputs.append(cur_node)
        self.inputs = new_inputs
        if fused_ids:
            from .parallel.symm import FusedDPUpdate
            self.fused_dp = FusedDPUpdate(config.symm, fused_ids)

    def re_minimize(self):
        from .graph.autodiff import gradients
        self.inputs = gradients(self.optimizer.loss, self.optimizer.params)
