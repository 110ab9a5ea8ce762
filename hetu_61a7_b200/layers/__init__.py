"""Layer library (``ht.layers``).  reference: python/hetu/layers/*.py."""
from .basic import (BaseLayer, Linear, Conv2d, BatchNorm, LayerNorm, DropOut, Embedding, MaxPool2d,
                    AvgPool2d, Relu, Gelu, Reshape, Slice, Sequence, SumLayers, Concatenate,
                    ConcatenateLayers, Identity)
from .moe import (Expert, MoELayer, TopKGate, KTop1Gate, KTop1Layer, SAMGate, SAMLayer, HashGate,
                  HashLayer, BalanceAssignmentGate, topkgating, ktop1gating, samgating, hashgating,
                  balance_loss, alignment_loss, generate_orthogonal)
