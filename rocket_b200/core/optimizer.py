"""Optimizer capsule: the fused step.

API mirror of the reference's rocket/core/optimizer.py:21-204, plus two additive keywords:
``max_grad_norm`` (default ``None`` = the reference's behaviour: it has no clipping at all,
SURVEY.md section 0.5) and ``sharded`` (default ``None`` = replicated step, as under DDP;
``True`` = each rank updates 1/W of the parameters, Engine.prepare_optimizer).  The reference's ``step(); zero_grad()`` (optimizer.py:129-130)
reaches torch.optim's per-tensor loops; here the prepared optimizer runs, on sync steps,
one deterministic norm kernel (if clipping) and one fused multi-tensor kernel that applies
1/world_size, the clip coefficient, the AdamW/SGD update and the gradient zero-fill in a
single pass over HBM.  The learning rate is read from ``param_groups`` at every step, so
schedulers keep working unchanged.
"""
from __future__ import annotations

import torch

from rocket_b200.core.capsule import Attributes, Capsule


class Optimizer(Capsule):
    def __init__(self, optimizer: torch.optim.Optimizer, tag: str = "opt",
                 priority: int = 1000, max_grad_norm: float | None = None,
                 sharded: bool | None = None) -> None:
        super().__init__(statefull=False, priority=priority)
        self._optimizer = optimizer
        self._tag = tag
        self._iter_idx = 0
        self._max_grad_norm = max_grad_norm
        self._sharded = sharded

    def setup(self, attrs: Attributes | None = None) -> None:
        Capsule.setup(self, attrs=attrs)
        found = False
        for prepared in self._accelerator._optimizers:
            if prepared.optimizer is not self._optimizer:
                continue
            if found:
                raise RuntimeError(
                    f"{type(self).__name__}: same optimizer has been registered twice.")
            found = True
            self._optimizer = prepared
        if not found:
            if self._sharded is None:
                self._optimizer = self._accelerator.prepare(self._optimizer)
            else:
                self._optimizer = self._accelerator.prepare_optimizer(self._optimizer,
                                                                      sharded=self._sharded)
        if self._max_grad_norm is not None:
            self._optimizer.max_grad_norm = float(self._max_grad_norm)

    def launch(self, attrs: Attributes | None = None) -> None:
        if torch.is_grad_enabled():
            # both are no-ops on accumulation steps; on sync steps zero_grad is folded
            # into the fused step kernel
            self._optimizer.step()
            self._optimizer.zero_grad()
        if self._accelerator.sync_gradients:
            state = Attributes(
                step=self._iter_idx,
                data={f"{self._tag}.lr.{i}": group.get("lr")
                      for i, group in enumerate(self._optimizer.param_groups)})
            if attrs.tracker is not None:
                attrs.tracker.scalars.append(state)
            if attrs.looper is not None:
                attrs.looper.state.lr = list(state.data.values())
            self._iter_idx += 1

    def destroy(self, attrs: Attributes | None = None) -> None:
        # leave the user's torch optimizer with ordinary per-parameter state
        export = getattr(self._optimizer, "_export_state", None)
        if export is not None:
            export()
        prepared = self._accelerator._optimizers
        for i, candidate in enumerate(prepared):
            if candidate is self._optimizer:
                prepared.pop(i)
                break
        Capsule.destroy(self, attrs=attrs)

    def state_dict(self) -> dict:
        return dict(iter_idx=self._iter_idx)

    def load_state_dict(self, state: dict) -> None:
        self._iter_idx = state["iter_idx"]
