"""``import graphmix`` - alias of :mod:`hetu_61a7_b200.graphmix` so the reference's GraphMix
scripts (third_party/GraphMix/tests, examples/gnn) run unchanged."""
import sys as _sys

import hetu_61a7_b200.graphmix as _impl
from hetu_61a7_b200.graphmix import *            # noqa: F401,F403
from hetu_61a7_b200.graphmix import (Client, Graph, NodePack, Server, Shard, cache, dataset, default_server_port, launcher,
                                     part_graph, sampler, start_standalone_server)

for _name in ('dataset', 'partition', 'shard', 'client', 'server', 'launcher', 'torch', 'tensorflow'):
    _sys.modules['graphmix.' + _name] = getattr(_impl, _name) if hasattr(_impl, _name) else __import__(
        'hetu_61a7_b200.graphmix.' + _name, fromlist=['_'])
__version__ = _impl.__version__
