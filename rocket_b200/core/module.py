"""Module capsule: forward of the user's network, then its children (Loss, Optimizer, ...).

API mirror of the reference's rocket/core/module.py:24-219.  The forward/backward of the
user's ``nn.Module`` stays in PyTorch; the engine owns only the two contexts entered around
it -- ``autocast`` and ``accumulate`` (gradient-accumulation phase; on multi-GPU it also
arms or disarms the bucketed gradient all-reduce the way DDP's ``no_sync`` would).
"""
from __future__ import annotations

import sys
from contextlib import contextmanager

import torch

from rocket_b200.core.capsule import Attributes, Capsule
from rocket_b200.core.dispatcher import Dispatcher
from rocket_b200.utils.torch import torch_move


class Module(Dispatcher):
    def __init__(self, module: torch.nn.Module, capsules: list[Capsule] = [],
                 priority: int = 1000) -> None:
        super().__init__(capsules=capsules, priority=priority)
        self._module = module

    def setup(self, attrs: Attributes | None = None) -> None:
        self.check_accelerator()
        found = False
        for model in self._accelerator._models:
            if model is not self._module:
                continue
            if found:
                raise RuntimeError(
                    f"{type(self).__name__}: same module has been registered twice.")
            found = True
            self._module = model
        if not found:
            self._module = torch_move(self._module, self._accelerator.device)
            self._module = self._accelerator.prepare(self._module)
        Dispatcher.setup(self, attrs)

    def launch(self, attrs: Attributes | None = None) -> None:
        if attrs is None or attrs.batch is None:
            return
        # grad mode was set by the Looper; it decides train vs eval (reference :130-135)
        if torch.is_grad_enabled():
            self._module.train()
        else:
            self._module.eval()
        with self.runner():
            attrs.batch = self._module.forward(attrs.batch)
            Dispatcher.launch(self, attrs=attrs)

    def destroy(self, attrs: Attributes | None = None) -> None:
        models = self._accelerator._models
        for i, model in enumerate(models):
            if model is self._module:
                models.pop(i)
                break
        Dispatcher.destroy(self, attrs=attrs)

    @contextmanager
    def runner(self):
        """autocast + accumulate, entered in that order, both always exited
        (reference module.py:175-219)."""
        autocast_ctx = self._accelerator.autocast()
        accumulate_ctx = self._accelerator.accumulate(self._module)
        try:
            autocast_ctx.__enter__()
            accumulate_ctx.__enter__()
            yield
        finally:
            autocast_ctx.__exit__(*sys.exc_info())
            accumulate_ctx.__exit__(*sys.exc_info())
