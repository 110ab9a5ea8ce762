"""Per-node timing executor (``Executor(..., timing='gpu'|'cpu')``).
reference: python/hetu/gpu_ops/timer_subexecutor.py:21-178."""
from __future__ import annotations

import time
from collections import defaultdict

import torch

from .executor import SubExecutor


class TimerSubExecutor(SubExecutor):
    def __init__(self, name, eval_node_list, config, timing='gpu'):
        super().__init__(name, eval_node_list, config)
        self.graph_ok = False            # every node is bracketed by events: eager only
        self.timing = timing
        self.use_gpu = timing == 'gpu' and self.device.type == 'cuda'
        self.timer = defaultdict(float)
        self.counts = defaultdict(int)
        self._wrap()

    def _wrap(self):
        for node in self.computing_nodes:
            if getattr(node, '_timed', False):
                continue
            orig = node.compute

            def timed(*a, _orig=orig, _node=node, **kw):
                if self.use_gpu:
                    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    s.record()
                    out = _orig(*a, **kw)
                    e.record()
                    e.synchronize()
                    ms = s.elapsed_time(e)
                else:
                    t0 = time.perf_counter()
                    out = _orig(*a, **kw)
                    ms = (time.perf_counter() - t0) * 1000.0
                self.timer[_node] += ms
                self.counts[_node] += 1
                return out

            node.compute = timed
            node._timed = True

    def logOut(self, path, log_level='node', clear=True):
        assert log_level in ('node', 'type')
        with open(path, 'w') as fw:
            if log_level == 'node':
                for node in self.computing_nodes:
                    n = max(self.counts[node], 1)
                    print(node.name, node.op_type, '%.6f ms' % (self.timer[node] / n), file=fw)
            else:
                agg, cnt = defaultdict(float), defaultdict(int)
                for node in self.computing_nodes:
                    n = max(self.counts[node], 1)
                    agg[node.op_type] += self.timer[node] / n
                    cnt[node.op_type] += 1
                for k in sorted(agg, key=lambda x: -agg[x]):
                    print(k, cnt[k], '%.6f ms' % agg[k], file=fw)
        if clear:
            self.clearTimer()

    def clearTimer(self):
        self.timer.clear()
        self.counts.clear()
