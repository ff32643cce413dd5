"""Loss capsule: objective, loss bookkeeping, backward.

API mirror of the reference's rocket/core/loss.py:21-150.  The reference does
``gather(loss).mean().item()`` on every micro-step (loss.py:95-98): an all-gather, a D2H
copy and a full stream synchronisation.  Here the running value lives in a device-side
double accumulator (kernel ``rk_scalar_accumulate``), is averaged across ranks and copied to
a pinned ring only on sync steps, and reaches ``attrs.looper.state.loss`` / the tracker as a
:class:`~rocket_b200.deferred.DeferredScalar` that materialises without blocking the step.
Same numbers, same ``step`` indices, no per-step host sync.
"""
from __future__ import annotations

import torch

from rocket_b200.core.capsule import Attributes, Capsule


class Loss(Capsule):
    def __init__(self, objective: torch.nn.Module, tag: str = "train_loss",
                 priority: int = 1100) -> None:
        super().__init__(statefull=True, priority=priority)
        self._objective = objective
        self._value = 0.0          # host mirror; authoritative only between flushes
        self._tag = tag
        self._step = 0
        self._acc = None           # engine-side accumulator handle

    def launch(self, attrs: Attributes | None = None) -> None:
        if attrs is None or attrs.batch is None:
            return
        if not torch.is_grad_enabled():
            return
        engine = self._accelerator
        gp = engine.graph_pass
        # replaying a captured step (Module(graph=True)): objective, accumulation and backward
        # are nodes of the graph; only the host-side bookkeeping below runs again
        loss = None if gp is not None and gp.mode == "replay" else self._objective(attrs.batch)
        if self._acc is None:
            self._acc = engine.scalar_accumulator(initial=self._value)
        # value += mean_over_ranks(loss) / accumulation_steps      (reference loss.py:95-98)
        self._acc.add(loss, divisor=engine.gradient_accumulation_steps)
        if engine.sync_gradients:
            value = self._acc.publish_and_reset()
            if attrs.tracker is not None:
                attrs.tracker.scalars.append(
                    Attributes(step=self._step, data={self._tag: value}))
            if attrs.looper is not None:
                attrs.looper.state.loss = value
            self._value = 0.0
            self._step += 1
        engine.backward(loss)

    def destroy(self, attrs: Attributes | None = None) -> None:
        self._acc = None
        Capsule.destroy(self, attrs=attrs)

    def state_dict(self) -> dict:
        value = float(self._acc.peek()) if self._acc is not None else self._value
        return dict(value=value, step=self._step)

    def load_state_dict(self, state: dict):
        self._value = state["value"]
        self._step = state["step"]
        self._acc = None
