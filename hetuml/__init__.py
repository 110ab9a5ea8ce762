"""``import hetuml`` - alias of :mod:`hetu_61a7_b200.ml` so the reference's HetuML scripts
(third_party/HetuML/examples) run unchanged: ``from hetuml.ensemble import GBDT`` etc."""
import sys as _sys

import hetu_61a7_b200.ml as _impl
from hetu_61a7_b200.ml import *            # noqa: F401,F403

for _name in ('cluster', 'data', 'decomposition', 'ensemble', 'factorization_machines', 'linear', 'model_base',
              'naive_bayes', 'neighbors'):
    _sys.modules['hetuml.' + _name] = getattr(_impl, _name)
_C = _impl._core
