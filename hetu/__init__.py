"""``import hetu as ht`` - alias of :mod:`hetu_61a7_b200` so reference scripts run unchanged."""
import sys as _sys
import hetu_61a7_b200 as _impl
from hetu_61a7_b200 import *            # noqa: F401,F403
from hetu_61a7_b200 import (optim, lr, init, layers, dist, ndarray, stream, data, metrics, Op,
                            Executor, HetuConfig, gradients)

for _name in ('layers', 'optimizer', 'lr_scheduler', 'initializers', 'ndarray', 'stream', 'context',
              'dataloader', 'distributed_strategies', 'ops', 'graph', 'parallel', 'utils', 'models',
              'launcher', 'runner', 'cpu_links', 'gpu_links', 'onnx', 'graphboard', 'memory_pool'):
    try:
        _mod = __import__('hetu_61a7_b200.' + _name, fromlist=['_'])
        _sys.modules['hetu.' + _name] = _mod
    except ImportError:
        pass
_sys.modules['hetu.optim'] = optim
_sys.modules['hetu.init'] = init
_sys.modules['hetu.lr'] = lr
_sys.modules['hetu.dist'] = dist
_sys.modules['hetu.gpu_ops'] = _sys.modules.get('hetu.ops', _impl)
__version__ = _impl.__version__
