"""Helpers of the reference's examples/gnn/gnn_tools live in the package:
part_graph -> ``python -m hetu_61a7_b200.graphmix.partition``; launcher -> ``graphmix.launcher``;
sparse datasets -> ``graphmix.dataset.load_dataset``."""
from hetu_61a7_b200.graphmix import launcher, part_graph          # noqa: F401
from hetu_61a7_b200.graphmix.dataset import load_dataset as load_sparse_dataset   # noqa: F401
