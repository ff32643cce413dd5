"""Scheduler capsule (API mirror of the reference's rocket/core/scheduler.py:21-143).

The prepared scheduler steps only on gradient-sync steps and, like accelerate's wrapper,
``world_size`` times per optimizer step.  It writes the host-side
``param_groups[i]["lr"]``; the fused step kernel receives that scalar as a launch argument.
"""
from __future__ import annotations

import torch

from rocket_b200.core.capsule import Attributes, Capsule


class Scheduler(Capsule):
    def __init__(self, scheduler: torch.optim.lr_scheduler.LRScheduler,
                 priority: int = 1000) -> None:
        super().__init__(statefull=False, priority=priority)
        self._scheduler = scheduler

    def setup(self, attrs: Attributes | None = None) -> None:
        Capsule.setup(self, attrs=attrs)
        found = False
        for prepared in self._accelerator._schedulers:
            if prepared.scheduler is not self._scheduler:
                continue
            if found:
                raise RuntimeError(
                    f"{type(self).__name__}: same scheduler has been registered twice. ")
            found = True
            self._scheduler = prepared
        if not found:
            self._scheduler = self._accelerator.prepare(self._scheduler)

    def launch(self, attrs: Attributes | None = None) -> None:
        if torch.is_grad_enabled():
            self._scheduler.step()

    def destroy(self, attrs: Attributes | None = None) -> None:
        prepared = self._accelerator._schedulers
        for i, candidate in enumerate(prepared):
            if candidate is self._scheduler:
                prepared.pop(i)
                break
        Capsule.destroy(self, attrs=attrs)
