"""Learning-rate schedulers (``get()`` / ``step()``).
reference: python/hetu/lr_scheduler.py:2-142."""
from __future__ import annotations


class FixedScheduler(object):
    def __init__(self, learning_rate):
        assert learning_rate >= 0, "learning rate must be non-negative"
        self.learning_rate = learning_rate

    def step(self):
        return self.learning_rate

    def get(self):
        return self.learning_rate


class StepScheduler(FixedScheduler):
    """lr *= gamma every ``step_size`` steps, floored at ``ending``."""

    def __init__(self, learning_rate, step_size, gamma=0.1, ending=1e-8):
        super().__init__(learning_rate)
        assert step_size > 0, "step size must be positive"
        assert 0 < gamma < 1, "gamma must be in (0, 1)"
        self.step_size, self.gamma, self.ending = step_size, gamma, ending
        self.cnt = 0

    def step(self):
        self.cnt += 1
        if self.learning_rate > self.ending and self.cnt >= self.step_size:
            self.learning_rate = max(self.learning_rate * self.gamma, self.ending)
            self.cnt = 0
        return self.learning_rate


class MultiStepScheduler(FixedScheduler):
    """lr *= gamma at each milestone."""

    def __init__(self, learning_rate, milestones, gamma=0.1):
        super().__init__(learning_rate)
        assert len(milestones) > 0, "milestones must be non-empty"
        assert 0 < gamma < 1, "gamma must be in (0, 1)"
        self.milestones = sorted(milestones)
        self.gamma = gamma
        self.cnt = 0
        self.ind = 0

    def step(self):
        self.cnt += 1
        if self.ind < len(self.milestones) and self.cnt >= self.milestones[self.ind]:
            self.learning_rate *= self.gamma
            self.ind += 1
        return self.learning_rate


class ExponentialScheduler(FixedScheduler):
    """lr *= gamma every step, floored at ``ending``."""

    def __init__(self, learning_rate, gamma=0.9, ending=1e-8):
        super().__init__(learning_rate)
        assert 0 < gamma < 1, "gamma must be in (0, 1)"
        self.gamma, self.ending = gamma, ending

    def step(self):
        if self.learning_rate > self.ending:
            self.learning_rate = max(self.learning_rate * self.gamma, self.ending)
        return self.learning_rate


class ReduceOnPlateauScheduler(FixedScheduler):
    """Reduce lr by ``factor`` when the monitored value stops improving for
    ``patience`` steps (then wait ``cooldown`` steps)."""

    def __init__(self, learning_rate, mode='min', factor=0.1, patience=10, threshold=0.0001,
                 threshold_mode='rel', cooldown=0, ending=1e-8):
        super().__init__(learning_rate)
        assert mode in ('min', 'max')
        assert 0 < factor < 1
        assert threshold_mode in ('rel', 'abs')
        self.mode, self.factor, self.patience = mode, factor, patience
        self.threshold, self.threshold_mode = threshold, threshold_mode
        self.cooldown, self.ending = cooldown, ending
        self.best = float('inf') if mode == 'min' else -float('inf')
        self.num_bad = 0
        self.cooldown_cnt = 0

    def _better(self, value):
        if self.mode == 'min':
            ref = self.best * (1 - self.threshold) if self.threshold_mode == 'rel' \
                else self.best - self.threshold
            return value < ref
        ref = self.best * (1 + self.threshold) if self.threshold_mode == 'rel' \
            else self.best + self.threshold
        return value > ref

    def step(self, value):
        if self._better(value):
            self.best = value
            self.num_bad = 0
        else:
            self.num_bad += 1
        if self.cooldown_cnt > 0:
            self.cooldown_cnt -= 1
            self.num_bad = 0
        if self.num_bad > self.patience:
            if self.learning_rate > self.ending:
                self.learning_rate = max(self.learning_rate * self.factor, self.ending)
            self.cooldown_cnt = self.cooldown
            self.num_bad = 0
        return self.learning_rate
