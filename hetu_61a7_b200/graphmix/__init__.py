"""Distributed graph sampling for GNN training (the reference's optional ``graphmix`` package,
third_party/GraphMix): partition a graph, run graph servers with samplers, pull mini-batch
sub-graphs from workers.  Also importable as ``graphmix``."""
from . import dataset
from ._core import cache, sampler
from .client import Client, Graph, NodePack, default_server_port
from .launcher import launcher, start_standalone_server
from .partition import part_graph, partition_nodes
from .server import Server
from .shard import Shard

__version__ = '0.1.0'
