"""PyTorch helpers for sampled sub-graphs (reference third_party/GraphMix/python/graphmix/torch/)."""
from .gcn import GCN, PytorchGCN
from .utils import mp_matrix

__all__ = ['mp_matrix', 'GCN', 'PytorchGCN']
