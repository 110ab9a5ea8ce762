"""Training-run metrics: step time, throughput and loss as a JSON-lines log and, optionally, as Prometheus gauges.

The reference prints losses from its example scripts and has no metrics export; this is the small
piece production runs ask for first.  Nothing here touches the device: pass in numbers you already have.

    mon = StepMonitor(samples_per_step=128, log_path='train_metrics.jsonl', prometheus_port=9100)
    for step in range(n):
        t0 = time.time(); loss = float(ex.run(...)[0].asnumpy()); mon.observe(time.time() - t0, loss=loss)
    mon.close()
"""
from __future__ import annotations

import json
import os
import time
from typing import Optional


class StepMonitor(object):
    def __init__(self, samples_per_step: int = 0, log_path: Optional[str] = None, prometheus_port: Optional[int] = None,
                 window: int = 50, rank: Optional[int] = None):
        self.samples_per_step = int(samples_per_step)
        self.rank = int(os.environ.get('RANK', '0')) if rank is None else int(rank)
        self.window = max(1, int(window))
        self.step = 0
        self._recent = []
        self._fw = open(log_path, 'a') if log_path else None
        self._gauges = None
        if prometheus_port is not None:
            from prometheus_client import Gauge, start_http_server
            start_http_server(int(prometheus_port) + self.rank)
            labels = {'rank': str(self.rank)}
            mk = lambda n, d: Gauge(n, d, list(labels)).labels(**labels)      # noqa: E731
            self._gauges = {'step_seconds': mk('hetu_step_seconds', 'duration of the last training step'),
                            'samples_per_second': mk('hetu_samples_per_second', 'throughput over the recent window'),
                            'loss': mk('hetu_loss', 'last reported loss'), 'step': mk('hetu_step', 'steps completed')}

    def observe(self, step_seconds: float, loss: Optional[float] = None, **extra) -> dict:
        self.step += 1
        self._recent.append(float(step_seconds))
        if len(self._recent) > self.window:
            self._recent.pop(0)
        mean = sum(self._recent) / len(self._recent)
        rec = {'time': time.time(), 'rank': self.rank, 'step': self.step, 'step_seconds': float(step_seconds),
               'samples_per_second': self.samples_per_step / mean if (self.samples_per_step and mean > 0) else None}
        if loss is not None:
            rec['loss'] = float(loss)
        rec.update(extra)
        if self._fw is not None:
            self._fw.write(json.dumps(rec) + '\n')
            self._fw.flush()
        if self._gauges is not None:
            self._gauges['step_seconds'].set(rec['step_seconds'])
            self._gauges['step'].set(self.step)
            if rec['samples_per_second'] is not None:
                self._gauges['samples_per_second'].set(rec['samples_per_second'])
            if loss is not None:
                self._gauges['loss'].set(float(loss))
        return rec

    def close(self):
        if self._fw is not None:
            self._fw.close()
            self._fw = None
