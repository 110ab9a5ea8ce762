"""Pipeline-parallel sub-executors: GPipe, PipeDream (1F1B with weight stashing), HetPipe.

reference: python/hetu/gpu_ops/gpipe_subexecutor.py:33-111 (all forwards, then all backwards,
optimizer applied once), pipedream_subexecutor.py:25-48 (1F1B schedule generator), :130-147
(copy_latest_weight / weight stashing), :149-169,317-328 (HetPipe: local accumulation, PS sync
every ``pp_nrank`` micro-batches).

The per-rank graph comes from ``parallel/placement.py`` (stage cut + send/recv insertion).
Activations and gradients travel over NCCL p2p (BASELINE.json keeps the pipeline on NCCL) as
non-blocking isend/irecv on two separate communicators - one per direction - so a stage that
is sending an activation forward and receiving a gradient backward cannot dead-lock (the
reference hand-places ncclGroupStart/End for the same purpose).
"""
from __future__ import annotations

from typing import Dict, List

import torch

from ..dataloader import DataloaderOp
from ..graph.executor import SubExecutor
from ..ndarray import IndexedSlices, NDArray
from ..ops.comm import PipelineReceiveOp, PipelineSendOp
from ..ops.variable import PlaceholderOp
from ..optimizer import OptimizerOp


def pipedream_scheduler(rank: int, nrank: int, num_batches: int):
    """Yields (micro_batch, phase) with phase 0 = forward, 1 = backward: ``nrank - rank - 1``
    warm-up forwards, then one-forward-one-backward, then the remaining backwards."""
    warmup = min(nrank - rank - 1, num_batches)
    for i in range(warmup):
        yield i, 0
    for i in range(num_batches - warmup):
        yield i + warmup, 0
        yield i, 1
    for i in range(num_batches - warmup, num_batches):
        yield i, 1


class _PipeBase(SubExecutor):
    def __init__(self, name, eval_node_list, config):
        super().__init__(name, eval_node_list, config)
        self.graph_ok = False
        self.opt_nodes = [n for n in self.topo_order if isinstance(n, OptimizerOp)]
        assert self.opt_nodes, 'pipeline executors run the training sub-executor'
        self.opt = self.opt_nodes[0]
        self._split()

    def _split(self):
        """Forward nodes vs gradient-side nodes (autodiff products and what hangs off them)."""
        b2f = getattr(self.opt.optimizer, 'backward2forward', {})
        consumers: Dict = {}
        for n in self.topo_order:
            for i in n.inputs:
                if i is not None:
                    consumers.setdefault(i, []).append(n)
        grad_side = set()
        for n in self.topo_order:
            if isinstance(n, OptimizerOp) or n in b2f or any(i in grad_side for i in n.inputs if i is not None):
                grad_side.add(n)
        changed = True
        while changed:                                     # receives that only feed gradient nodes
            changed = False
            for n in self.topo_order:
                if n not in grad_side and isinstance(n, PipelineReceiveOp):
                    cs = consumers.get(n, [])
                    if cs and all(c in grad_side for c in cs):
                        grad_side.add(n)
                        changed = True
                if n not in grad_side and any(i in grad_side for i in n.inputs if i is not None):
                    grad_side.add(n)
                    changed = True
        if any(hasattr(n, 'pp_forward') for n in self.topo_order):
            # the placement pass classified every node on the global graph: use that (ranks must agree on
            # which pass a collective belongs to); nodes created later follow their inputs
            grad_side = set()
            for n in self.topo_order:
                flag = getattr(n, 'pp_forward', None)
                if flag is None:
                    flag = not (isinstance(n, OptimizerOp) or any(i in grad_side for i in n.inputs if i is not None))
                if not flag:
                    grad_side.add(n)
        self.forward_nodes = [n for n in self.topo_order if n not in grad_side]
        self.backward_nodes = [n for n in self.topo_order if n in grad_side and not isinstance(n, OptimizerOp)]
        for n in self.topo_order:
            if isinstance(n, (PipelineSendOp, PipelineReceiveOp)):
                n.direction = 1 if n in grad_side else 0
        self.param_nodes_mine = [p for p in self.opt.optimizer.params
                                 if self.config.placeholder_to_arr_map.get(p) is not None]

    def _run_nodes(self, nodes, vals):
        cfg = self.config
        for node in nodes:
            if node in vals:
                continue
            if isinstance(node, PlaceholderOp):
                v = cfg.placeholder_to_arr_map.get(node)
                if v is None and node.tensor_value is not None:
                    tv = node.tensor_value
                    v = tv.tensor if isinstance(tv, NDArray) else tv
                vals[node] = v
                continue
            ins = [vals.get(i) if i is not None else None for i in node.inputs]
            if getattr(node, 'takes_inference', False):
                vals[node] = node.compute(ins, None, cfg.comp_stream, inference=False)
            else:
                vals[node] = node.compute(ins, None, cfg.comp_stream)

    def _feeds(self, feed_dict, m):
        vals = {}
        for node, value in feed_dict.items():
            v = value[m] if isinstance(value, (list, tuple)) else value
            vals[node] = self._to_device(self._feed_tensor(node, v))
        for node in self.dataloader_nodes:
            if node not in vals:
                vals[node] = self._to_device(node.get_arr(self.name))
        return vals

    def _grads_of(self, vals):
        return [vals.get(g) if g is not None else None for g in self.opt.inputs]

    @staticmethod
    def _accumulate(acc, grads):
        if acc is None:
            return [g.clone() if isinstance(g, torch.Tensor) else g for g in grads]
        out = []
        for a, g in zip(acc, grads):
            if g is None:
                out.append(a)
            elif a is None:
                out.append(g.clone() if isinstance(g, torch.Tensor) else g)
            elif isinstance(g, IndexedSlices):
                out.append(a.merge(g))
            else:
                out.append(a + g)
        return out

    def _results(self, eval_nodes, per_batch_vals, convert):
        out = []
        for n in eval_nodes:
            if isinstance(n, OptimizerOp) or n in getattr(self.config, 'dropped_eval_nodes', []):
                out.append(None)
                continue
            vs = [v.get(n) for v in per_batch_vals]
            if any(v is None for v in vs):
                out.append(None)
            elif convert:
                out.append([v.detach().cpu().numpy() for v in vs])
            else:
                out.append([NDArray(v) for v in vs])
        return out


def schedule_for_different_dp(dp_rank: int, dp_nrank: int, min_dp_nrank: int):
    """Which micro-batches a replica takes when its stage has MORE data-parallel replicas than the narrowest stage of
    the pipeline (reference pipeline_subexecutor.py:83-106).  The narrow stages emit ``min_dp_nrank`` micro-batches
    per tick, handed round-robin to the ``dp_nrank`` replicas of the wide stage; the pattern repeats after
    ``dp_nrank / gcd(dp_nrank, min_dp_nrank)`` ticks.  [True] when the degrees are equal."""
    if dp_nrank <= min_dp_nrank:
        return [True]
    a, b = dp_nrank, min_dp_nrank
    while b:
        a, b = b, a % b
    cycle = dp_nrank // a
    schedule, index = [], dp_rank
    while len(schedule) < cycle:
        if index < min_dp_nrank:
            schedule.append(True)
            index = index + dp_nrank - min_dp_nrank
        else:
            schedule.append(False)
            index -= min_dp_nrank
    return schedule


class SubExecutor4Gpipe(_PipeBase):
    """All micro-batch forwards, then all backwards, gradients accumulated, one optimizer step.  A replica of a stage
    that is wider (more data-parallel replicas) than the narrowest stage runs only the micro-batches of its turn
    (``schedule_for_different_dp``); the gradient all-reduce at the end of the batch puts the replicas' shares together."""

    def _turns(self, M):
        cfg = self.config
        sched = schedule_for_different_dp(getattr(cfg, 'dp_rank', 0) or 0, getattr(cfg, 'dp_nrank', 1) or 1,
                                          getattr(cfg, 'min_dp_nrank', None) or (getattr(cfg, 'dp_nrank', 1) or 1))
        assert M % len(sched) == 0, ('the number of micro-batches (%d) must be a multiple of the schedule length (%d), '
                                     'so that every replica reaches the gradient all-reduce' % (M, len(sched)))
        start = getattr(self, '_turn_index', 0)
        self._turn_index = (start + M) % len(sched)
        return [sched[(start + m) % len(sched)] for m in range(M)]

    def run(self, eval_node_list=[], feed_dict={}, convert_to_numpy_ret_vals=False, batch_num=None, **kwargs):
        M = batch_num or 1
        eval_nodes = list(eval_node_list) if eval_node_list else getattr(self, 'result_nodes', self.eval_node_list)
        self._pre_step()
        turns = self._turns(M)
        all_vals = []
        for m in range(M):
            if not turns[m]:
                all_vals.append(None)
                continue
            vals = self._feeds(feed_dict, m)
            self._run_nodes(self.forward_nodes, vals)
            all_vals.append(vals)
        acc = None
        for m in range(M):
            if all_vals[m] is None:
                continue
            self._run_nodes(self.backward_nodes, all_vals[m])
            acc = self._accumulate(acc, self._grads_of(all_vals[m]))
        self.opt.compute(acc)
        self.step_count += 1
        return self._results(eval_nodes, [v for v in all_vals if v is not None], convert_to_numpy_ret_vals)


class SubExecutor4Pipedream(_PipeBase):
    """1F1B with weight stashing; ``pipeline='hetpipe'`` additionally accumulates the local
    updates and synchronises replicas every ``pp_nrank`` micro-batches (PS push/pull of the
    weight delta when a parameter server is attached, all-reduce average otherwise)."""

    def __init__(self, name, eval_node_list, config):
        super().__init__(name, eval_node_list, config)
        self.hetpipe = config.pipeline == 'hetpipe'
        self._since_sync = 0
        self._anchor = None

    def _params(self):
        return [self.config.placeholder_to_arr_map[p] for p in self.param_nodes_mine]

    def run(self, eval_node_list=[], feed_dict={}, convert_to_numpy_ret_vals=False, batch_num=None, **kwargs):
        M = batch_num or 1
        eval_nodes = list(eval_node_list) if eval_node_list else getattr(self, 'result_nodes', self.eval_node_list)
        cfg = self.config
        self._pre_step()
        vals_of: Dict[int, dict] = {}
        stash: Dict[int, List[torch.Tensor]] = {}
        order = []
        for m, phase in pipedream_scheduler(cfg.pp_rank, cfg.pp_nrank, M):
            if phase == 0:
                vals = self._feeds(feed_dict, m)
                stash[m] = [p.clone() for p in self._params()]          # weights this micro-batch saw
                self._run_nodes(self.forward_nodes, vals)
                vals_of[m] = vals
                order.append(m)
            else:
                vals = vals_of[m]
                params = self._params()
                latest = [p.clone() for p in params]
                for p, s in zip(params, stash.pop(m)):                   # backward with the stashed version
                    p.copy_(s)
                self._run_nodes(self.backward_nodes, vals)
                for p, l in zip(params, latest):                         # copy_latest_weight
                    p.copy_(l)
                self.opt.compute(self._grads_of(vals))
                if self.hetpipe:
                    self._hetpipe_sync()
        self.step_count += 1
        return self._results(eval_nodes, [vals_of[m] for m in sorted(vals_of)], convert_to_numpy_ret_vals)

    def _hetpipe_sync(self):
        cfg = self.config
        self._since_sync += 1
        if self._since_sync < max(cfg.pp_nrank, 1):
            return
        self._since_sync = 0
        params = self._params()
        if self._anchor is None:
            self._anchor = [p.clone() for p in params]
            return
        for node, p, a in zip(self.param_nodes_mine, params, self._anchor):
            delta = p - a
            if cfg.ps_comm is not None and getattr(node, 'ps_resident', False):
                fresh = cfg.ps_comm.DDPushPull(node.id, delta.detach().cpu().numpy().reshape(-1))
                p.copy_(torch.from_numpy(fresh.reshape(tuple(p.shape))).to(p.device))
            else:
                ranks = getattr(node, 'dp_group_ranks', None)
                if ranks and len(ranks) > 1:
                    from .comm import new_group_comm
                    comm = new_group_comm(list(ranks))
                    comm.allreduce(delta)
                    p.copy_(a + delta / len(ranks))
            a.copy_(p)


# reference spelling
SubExecutor4Pipe = SubExecutor4Gpipe
