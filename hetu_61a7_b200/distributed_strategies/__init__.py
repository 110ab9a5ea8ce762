"""Distribution strategies (``ht.dist``).  reference:
python/hetu/distributed_strategies/{base,simple}.py - only ``DataParallel``
exists upstream; ``ModelParallel4LM``, ``PipelineParallel`` and ``Hybrid``
(Galvatron-style per-layer (dp, tp, pp) description) are provided here because
the examples' API expects them."""
from .base import Strategy
from .simple import DataParallel
from .hybrid import ModelParallel, PipelineParallel, HybridParallel, LayerPlan

__all__ = ['Strategy', 'DataParallel', 'ModelParallel', 'PipelineParallel', 'HybridParallel',
           'LayerPlan']
