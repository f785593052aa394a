"""StepLR with a state_dict interchangeable with ``torch.optim.lr_scheduler.StepLR``.

Spec: torch:optim/lr_scheduler.py:594-660 (closed form ``lr = base * gamma**(epoch//step)``;
the scheduler's state_dict is its ``__dict__`` minus the optimizer, ``:182-199``) as used by
ref:restnet_ddp.py:123,140 — stepped once per *epoch* (quirk Q16).
"""
from __future__ import annotations

from typing import List


class StepLR:
    def __init__(self, optimizer, step_size: int, gamma: float = 0.1, last_epoch: int = -1):
        self.optimizer = optimizer
        self.step_size = int(step_size)
        self.gamma = gamma
        if last_epoch == -1:
            for g in optimizer.param_groups:
                g.setdefault("initial_lr", g["lr"])
        self.base_lrs: List[float] = [g["initial_lr"] for g in optimizer.param_groups]
        self.last_epoch = last_epoch
        self._step_count = 0
        self._is_initial = False
        self._get_lr_called_within_step = False
        self._last_lr: List[float] = [g["lr"] for g in optimizer.param_groups]
        self.step()                      # torch performs an initial step: last_epoch -> 0

    def _closed_form(self) -> List[float]:
        return [b * self.gamma ** (self.last_epoch // self.step_size) for b in self.base_lrs]

    def step(self) -> None:
        self._step_count += 1
        self.last_epoch += 1
        lrs = self._closed_form()
        for g, lr in zip(self.optimizer.param_groups, lrs):
            g["lr"] = lr
        self._last_lr = lrs

    def get_last_lr(self) -> List[float]:
        return self._last_lr

    def state_dict(self):
        return {k: v for k, v in self.__dict__.items() if k != "optimizer"}

    def load_state_dict(self, sd) -> None:
        self.__dict__.update(sd)
        for g, lr in zip(self.optimizer.param_groups, self._last_lr):
            g["lr"] = lr
