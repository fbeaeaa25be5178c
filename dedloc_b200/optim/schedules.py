"""Learning-rate schedules of the reference recipes. All of them are stepped once per COLLABORATIVE step by
CollaborativeOptimizer (the LR is a function of the global step).

* `get_linear_schedule_with_warmup` -- transformers' linear warm-up / linear decay used at
  /root/reference/albert/run_trainer.py:96-98.
* `LinearWarmupCosineAnnealingLR` -- the schedule defined next to SGDCollaborative
  (/root/reference/swav/ClassyVision/classy_vision/optim/sgd_collaborative.py:25-84): linear warm-up from
  `warmup_start_lr` to the base lr over `warmup_epochs` steps, then cosine annealing to `eta_min` at `max_epochs`.
* `NoOpScheduler` -- the dummy handed to HF Trainer (run_trainer.py:189-207); the real schedule is owned by the
  collaborative optimizer.
"""
from __future__ import annotations

import math
from typing import List

from torch.optim.lr_scheduler import LambdaLR, LRScheduler


def get_linear_schedule_with_warmup(optimizer, num_warmup_steps: int, num_training_steps: int, last_epoch: int = -1):
    def lr_lambda(step: int) -> float:
        if step < num_warmup_steps:
            return float(step) / float(max(1, num_warmup_steps))
        return max(0.0, float(num_training_steps - step) / float(max(1, num_training_steps - num_warmup_steps)))

    return LambdaLR(optimizer, lr_lambda, last_epoch)


class LinearWarmupCosineAnnealingLR(LRScheduler):
    def __init__(self, optimizer, warmup_epochs: int, max_epochs: int, warmup_start_lr: float = 0.0,
                 eta_min: float = 0.0, last_epoch: int = -1):
        self.warmup_epochs, self.max_epochs = warmup_epochs, max_epochs
        self.warmup_start_lr, self.eta_min = warmup_start_lr, eta_min
        super().__init__(optimizer, last_epoch)

    def get_lr(self) -> List[float]:
        return self._closed_form(self.last_epoch)

    def _closed_form(self, t: int) -> List[float]:
        if t < self.warmup_epochs:
            return [self.warmup_start_lr + t * (base - self.warmup_start_lr) / max(1, self.warmup_epochs)
                    for base in self.base_lrs]
        t = min(t, self.max_epochs)
        span = max(1, self.max_epochs - self.warmup_epochs)
        return [self.eta_min + 0.5 * (base - self.eta_min) * (1 + math.cos(math.pi * (t - self.warmup_epochs) / span))
                for base in self.base_lrs]


class NoOpScheduler(LRScheduler):
    """Does nothing on step(); reports the LR that the collaborative optimizer's own scheduler set."""

    def __init__(self, optimizer):
        self.optimizer = optimizer
        self._last_lr = [g["lr"] for g in optimizer.param_groups]
        self._step_count = 0
        self.last_epoch = 0

    def get_lr(self):
        return [g["lr"] for g in self.optimizer.param_groups]

    def get_last_lr(self):
        return self.get_lr()

    def print_lr(self, *args, **kwargs):
        sched = getattr(self.optimizer, "scheduler", None)
        if sched is not None and hasattr(sched, "print_lr"):
            return sched.print_lr(*args, **kwargs)

    def step(self, *args, **kwargs):
        self._last_lr = self.get_lr()

    def state_dict(self):
        return {}

    def load_state_dict(self, *args, **kwargs):
        pass
