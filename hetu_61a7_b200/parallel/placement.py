"""Device placement pass: pipeline stages, model-parallel workers and data-parallel replicas
described with context lists -> per-rank sub-graphs joined by send/recv.

reference contract (the lowering code itself is missing upstream):
``with ht.context([...])`` - GPU entries of a group are DP workers, a *tuple* of GPUs is one
model-parallel worker, different groups on consecutive layers are pipeline stages
(examples/runner/parallel/{gpipe,test_mlp_mp,complex_pipeline_mlp}.py, context.py:27-57).
One process drives one GPU; rank == device id on a single node.

For every graph edge whose producer and consumer live on different ranks the pass inserts a
``PipelineSendOp`` on the producer and a ``PipelineReceiveOp`` on the consumer (NCCL p2p, as
BASELINE.json prescribes); a consumer at position (worker j, member k) of its stage receives
from position (j mod W_p, k mod T_p) of the producer stage.  Nodes that do not belong to this
rank are dropped from its graph; parameters get their data-parallel replica group.
"""
from __future__ import annotations

from typing import Dict, List, Optional

from ..context import DeviceGroup
from ..ndarray import is_gpu_ctx


def _positions(group: DeviceGroup):
    """[(worker index, member index, rank)] for every GPU of the group."""
    out = []
    for j, w in enumerate(group.workers):
        members = w if isinstance(w, tuple) else (w,)
        for k, d in enumerate(members):
            out.append((j, k, d.device_id))
    return out


def _ranks(group: Optional[DeviceGroup]) -> List[int]:
    return [r for _, _, r in _positions(group)] if group is not None else []


def _fix_missing_ctx(topo):
    for node in topo:
        if node.raw_ctx is None or node.raw_ctx.worker_num == 0:
            for i in node.inputs:
                if i is not None and i.raw_ctx is not None and i.raw_ctx.worker_num > 0:
                    node.raw_ctx = i.raw_ctx
                    break
    for node in reversed(topo):                      # sources without inputs: take a consumer's
        for i in node.inputs:
            if i is not None and (i.raw_ctx is None or i.raw_ctx.worker_num == 0) and node.raw_ctx is not None:
                i.raw_ctx = node.raw_ctx


def assign_placement(config, eval_nodes) -> None:
    from ..graph.autodiff import find_topo_sort
    from ..ops.comm import pipeline_receive_op, pipeline_send_op
    from ..ops.variable import PlaceholderOp
    from ..optimizer import OptimizerOp
    from ..dataloader import DataloaderOp
    topo = find_topo_sort(eval_nodes)
    _fix_missing_ctx(topo)
    stages: List[DeviceGroup] = []
    for node in topo:
        rc = node.raw_ctx
        if rc is None or rc.worker_num == 0 or isinstance(node, OptimizerOp):
            continue
        if rc not in stages:
            stages.append(rc)
    pure_dp = len(stages) <= 1 and not any(s.is_mp for s in stages)
    if pure_dp:
        return
    from .comm import make_p2p_comms, new_group_comm
    make_p2p_comms()
    # process groups are created collectively: EVERY rank walks all stages in the same order and creates
    # each tensor-parallel group (one per tuple worker) and each data-parallel replica group (same member
    # index across the workers of a stage), also the ones it is not a member of
    for s in stages:
        pos = _positions(s)
        for w in s.workers:
            if isinstance(w, tuple) and len(w) > 1:
                new_group_comm(sorted(d.device_id for d in w))
        width = max(len(w) if isinstance(w, tuple) else 1 for w in s.workers)
        if s.worker_num > 1:
            for k in range(width):
                ranks = sorted(r for _, kk, r in pos if kk == k)
                if len(ranks) > 1:
                    new_group_comm(ranks)
    me = config.rank
    mine = lambda n: isinstance(n, OptimizerOp) or n.raw_ctx is None or me in _ranks(n.raw_ctx)   # noqa: E731
    my_stage_ids = [i for i, s in enumerate(stages) if me in _ranks(s)]
    config.pp_nrank = len(stages)
    config.min_dp_nrank = min(s.worker_num for s in stages)      # narrowest stage: sets the micro-batch turns of the wider ones
    config.pp_rank = my_stage_ids[0] if my_stage_ids else 0
    for s in stages:
        for j, k, r in _positions(s):
            if r == me:
                config.dp_rank, config.dp_nrank, config.mp_index = j, s.worker_num, k

    # parameters: replica group = same member index across the workers of the stage
    for node in topo:
        if isinstance(node, PlaceholderOp) and node.trainable and node.raw_ctx is not None:
            pos = _positions(node.raw_ctx)
            myk = next((k for _, k, r in pos if r == me), None)
            node.dp_group_ranks = sorted(r for _, k, r in pos if k == myk) if myk is not None else [me]
        if isinstance(node, DataloaderOp) and config.shard_dataloaders and node.raw_ctx is not None:
            pos = _positions(node.raw_ctx)
            for j, k, r in pos:
                if r == me and node.raw_ctx.worker_num > 1:
                    node.set_dp_rank(j, node.raw_ctx.worker_num)

    def source_rank(u, recv_rank) -> int:
        """Which rank of u's stage feeds ``recv_rank`` (position-wise modulo mapping)."""
        prod = _positions(u.raw_ctx)
        cons_stage = next(s for s in stages if recv_rank in _ranks(s) and s != u.raw_ctx) \
            if any(recv_rank in _ranks(s) and s != u.raw_ctx for s in stages) else u.raw_ctx
        j, k = next(((j, k) for j, k, r in _positions(cons_stage) if r == recv_rank), (0, 0))
        nw = u.raw_ctx.worker_num
        w = u.raw_ctx.workers[j % nw]
        members = w if isinstance(w, tuple) else (w,)
        return members[k % len(members)].device_id

    def source_ranks(u, recv_rank):
        """All ranks of u's stage that feed ``recv_rank``.  Stages of equal width pair up position-wise; a WIDER
        producer stage fans in (the replicas j with j % C == i take turns, micro-batch by micro-batch, in replica
        order - parallel/pipeline.py schedule_for_different_dp); a narrower one is covered by source_rank."""
        cons = [s for s in stages if recv_rank in _ranks(s) and s != u.raw_ctx]
        if not cons:
            return [source_rank(u, recv_rank)]
        cons_stage = cons[0]
        P, C = u.raw_ctx.worker_num, cons_stage.worker_num
        if P <= C:
            return [source_rank(u, recv_rank)]
        j, k = next(((j, k) for j, k, r in _positions(cons_stage) if r == recv_rank), (0, 0))
        out = []
        for jj in range(P):
            if jj % C == j:
                w = u.raw_ctx.workers[jj]
                members = w if isinstance(w, tuple) else (w,)
                out.append(members[k % len(members)].device_id)
        return out

    for i, n in enumerate(topo):
        n.global_order = float(i)
    # forward / backward side of every node, decided ONCE on the global graph so that all ranks agree
    # (the pipeline executors run the two sides in separate passes)
    fwd_set = set(find_topo_sort([n for n in eval_nodes if not isinstance(n, OptimizerOp)]))
    for n in topo:
        n.pp_forward = n in fwd_set
    recv_cache: Dict = {}
    sends: List = []
    sent = set()
    for v in topo:
        if isinstance(v, OptimizerOp):
            for idx, g in enumerate(v.inputs):
                p = v.optimizer.params[idx]
                if g is not None and not (p.raw_ctx is None or me in _ranks(p.raw_ctx)):
                    v.inputs[idx] = None
            continue
        v_ranks = _ranks(v.raw_ctx)
        for idx, u in enumerate(v.inputs):
            if u is None or u.raw_ctx is None or v.raw_ctx is None:
                continue
            u_ranks = _ranks(u.raw_ctx)
            # consumer side
            if me in v_ranks and me not in u_ranks:
                if u not in recv_cache:
                    srcs = source_ranks(u, me)
                    src = srcs[0]
                    rv = pipeline_receive_op(src if len(srcs) == 1 else tuple(srcs), config.nccl_comm,
                                             use_indexed_slices=u.use_indexed_slices, ctx=v.raw_ctx)
                    rv.ctx = config.context
                    rv.name = 'recv_%s_from%s' % (u.name, '_'.join(str(x) for x in srcs))
                    rv.global_order = v.global_order - 0.5
                    rv.pp_forward = getattr(v, 'pp_forward', True)
                    recv_cache[u] = rv
                v.inputs[idx] = recv_cache[u]
            # producer side
            if me in u_ranks:
                targets = [r for r in v_ranks if r not in u_ranks and me in source_ranks(u, r)]
                if len(targets) > 1 and v.raw_ctx.worker_num > u.raw_ctx.worker_num:
                    # a wider consumer stage: ONE send node that hands the micro-batches out round-robin, in replica
                    # order (the replicas take them in turns)
                    groups = [tuple(targets)]
                else:
                    groups = [(r,) for r in targets]
                for grp in groups:
                    if (u, grp) in sent:
                        continue
                    sent.add((u, grp))
                    sd = pipeline_send_op(u, grp[0] if len(grp) == 1 else grp, config.nccl_comm, ctx=u.raw_ctx)
                    sd.name = 'send_%s_to%s' % (u.name, '_'.join(str(x) for x in grp))
                    sd.global_order = v.global_order - 0.25
                    sd.pp_forward = getattr(v, 'pp_forward', True)
                    sends.append((topo.index(v), sd))
    # this rank evaluates: its own eval nodes + its sends, interleaved in topological position so
    # that both sides of every p2p pair issue them in the same relative order
    my_eval = [n for n in eval_nodes if mine(n)]
    order = {n: i for i, n in enumerate(topo)}
    # a collective of a model-parallel worker needs every member, also the ones whose own
    # results do not depend on it (e.g. the all-gather that only feeds the next stage)
    from .mp_lowering import MPAllGatherOp, MPAllReduceOp, MPReduceScatterOp
    coll = [n for n in topo if isinstance(n, (MPAllGatherOp, MPAllReduceOp, MPReduceScatterOp))
            and me in _ranks(n.raw_ctx) and n not in my_eval]
    merged = [(order.get(n, len(topo)), 1, n) for n in my_eval] + [(pos, 0, sd) for pos, sd in sends] \
        + [(order[n], 1, n) for n in coll]
    merged.sort(key=lambda t: (t[0], t[1]))
    config.my_eval_nodes = [n for _, _, n in merged]
    config.dropped_eval_nodes = [n for n in eval_nodes if not mine(n)]
    if config.p2p_stream is None:
        from ..stream import Stream
        import torch
        if is_gpu_ctx(config.context) and torch.cuda.is_available():
            config.p2p_stream = Stream(config.context)


def split_stages(node_list, stage_groups, boundaries=None) -> None:
    """Assign ``raw_ctx`` of every node to one of ``stage_groups``: contiguous cuts of the forward
    topological order, balanced by parameter count unless ``boundaries`` (nodes that START a new
    stage) are given.  Gradient nodes follow their forward node through ``gradient(ctx=raw_ctx)``
    when the strategy is applied before ``minimize``; nodes created afterwards inherit from
    their inputs (placement pass)."""
    from ..graph.autodiff import find_topo_sort
    from ..ops.variable import PlaceholderOp
    from ..optimizer import OptimizerOp
    import numpy as np
    topo = [n for n in find_topo_sort(node_list) if not isinstance(n, OptimizerOp)]
    nstage = len(stage_groups)
    if boundaries:
        cut_at = {id(b) for b in boundaries}
        stage = 0
        for n in topo:
            if id(n) in cut_at and stage < nstage - 1:
                stage += 1
            n.raw_ctx = stage_groups[stage]
        return
    weights = [int(np.prod(n.shape)) if isinstance(n, PlaceholderOp) and n.trainable and n.shape else 0 for n in topo]
    total = max(sum(weights), 1)
    acc, stage = 0, 0
    for n, w in zip(topo, weights):
        n.raw_ctx = stage_groups[stage]
        acc += w
        if acc >= total * (stage + 1) / nstage and stage < nstage - 1:
            stage += 1
