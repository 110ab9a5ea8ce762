"""Model / pipeline / hybrid strategies.

The reference snapshot exports only ``DataParallel``; its examples assume
model-parallel (``ht.dispatch`` + tuple contexts), pipeline (different contexts
on consecutive layers) and Galvatron-style hybrid layouts whose lowering code
is absent.  These strategies assign ``raw_ctx`` per node;
the actual graph rewriting lives in ``parallel/placement.py`` and
``parallel/mp_lowering.py``.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

from ..context import DeviceGroup
from ..graph.autodiff import find_topo_sort
from ..ops.variable import PlaceholderOp
from .base import Strategy


class LayerPlan(object):
    """Parallel layout of a span of layers: ``dp`` replicas x ``tp`` shards on
    pipeline stage ``pp_stage`` (Galvatron's per-layer (dp, tp, pp) triple)."""

    def __init__(self, dp: int = 1, tp: int = 1, pp_stage: int = 0):
        self.dp, self.tp, self.pp_stage = dp, tp, pp_stage

    def __repr__(self):
        return 'LayerPlan(dp=%d, tp=%d, pp_stage=%d)' % (self.dp, self.tp, self.pp_stage)


def _stage_group(devices: List[str], dp: int, tp: int) -> DeviceGroup:
    assert len(devices) == dp * tp, (devices, dp, tp)
    units = []
    for r in range(dp):
        shard = devices[r * tp:(r + 1) * tp]
        units.append(tuple(shard) if tp > 1 else shard[0])
    return DeviceGroup(units)


class ModelParallel(Strategy):
    """All workers form one tensor-parallel group (Megatron-style); nodes the
    user marked with ``ht.dispatch`` are sharded, everything else replicated."""

    def __init__(self, tp: Optional[int] = None):
        super().__init__()
        devs = self.worker_devices()
        self.tp = tp or len(devs)
        self.dp = len(devs) // self.tp
        self.raw_ctx = _stage_group(devs, self.dp, self.tp)

    def set_raw_ctxs_n_states(self, node_list, memory_pool=None):
        for node in find_topo_sort(node_list):
            node.raw_ctx = self.raw_ctx
        return self.raw_ctx


class PipelineParallel(Strategy):
    """Cut the forward graph into ``num_stages`` contiguous stages balanced by
    parameter count (or at user-given boundary nodes)."""

    def __init__(self, num_stages: Optional[int] = None, boundaries: Optional[Sequence] = None,
                 dp: int = 1):
        super().__init__()
        devs = self.worker_devices()
        self.dp = dp
        self.num_stages = num_stages or max(len(devs) // dp, 1)
        self.boundaries = list(boundaries) if boundaries else None
        per = len(devs) // self.num_stages
        self.stage_groups = [DeviceGroup(devs[i * per:(i + 1) * per]) for i in range(self.num_stages)]
        self.raw_ctx = self.stage_groups[0]

    def set_raw_ctxs_n_states(self, node_list, memory_pool=None):
        from ..parallel.placement import split_stages
        split_stages(node_list, self.stage_groups, self.boundaries)
        return DeviceGroup([d for g in self.stage_groups for d in g.flat_devices()])


class HybridParallel(Strategy):
    """Per-stage (dp, tp) with ``len(plans)`` pipeline stages, e.g. BERT-large
    on 8 GPUs as 2 stages x (dp=2, tp=2): ``HybridParallel([LayerPlan(2,2,0),
    LayerPlan(2,2,1)])``."""

    def __init__(self, plans: Sequence[LayerPlan], boundaries: Optional[Sequence] = None):
        super().__init__()
        devs = self.worker_devices()
        self.plans = list(plans)
        self.boundaries = list(boundaries) if boundaries else None
        self.stage_groups = []
        off = 0
        for plan in self.plans:
            n = plan.dp * plan.tp
            self.stage_groups.append(_stage_group(devs[off:off + n], plan.dp, plan.tp))
            off += n
        assert off == len(devs), 'plans use %d devices, cluster has %d' % (off, len(devs))
        self.raw_ctx = self.stage_groups[0]

    def set_raw_ctxs_n_states(self, node_list, memory_pool=None):
        from ..parallel.placement import split_stages
        split_stages(node_list, self.stage_groups, self.boundaries)
        return DeviceGroup([u for g in self.stage_groups for u in g])
