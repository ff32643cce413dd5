"""Meter / Metric capsules (API mirror of the reference's rocket/core/meter.py:23-206).

``Meter`` runs only in eval loops.  The reference all-gathers every keyed tensor on every
eval step (``gather_for_metrics``, meter.py:93 -- for ImageNet logits that is 1 MB per rank
per step) and the example metric then syncs the host with ``.item()``
(examples/mnist.py:34).  Child metrics that declare ``device_reduced = True`` (see
rocket_b200/metrics.py) count on their own rank with integer atomics on the device and
combine the counters once, when they are read; for those the gather is skipped.  Any other
Metric gets the reference behaviour unchanged.
"""
from __future__ import annotations

from typing import List

import torch

from rocket_b200.core.capsule import Attributes, Capsule
from rocket_b200.core.dispatcher import Dispatcher
from rocket_b200.utils.collections import apply_to_collection


def rebuild_batch(lookup_table):
    """Put gathered values back at their keys (reference meter.py:23-27)."""
    def fn(value, key, **kwargs):
        return lookup_table.get(key, value)
    return fn


class Meter(Dispatcher):
    def __init__(self, capsules: List[Capsule], keys: List, priority=1000) -> None:
        super().__init__(capsules=capsules, priority=priority)
        self._keys = sorted(keys)

    def _needs_gather(self) -> bool:
        return not all(getattr(c, "device_reduced", False) for c in self._capsules)

    def launch(self, attrs: Attributes | None = None) -> None:
        if attrs is None or attrs.batch is None:
            return
        if torch.is_grad_enabled():
            return
        if self._needs_gather():
            values = [attrs.batch[key] for key in self._keys]
            gathered = self._accelerator.gather_for_metrics(values)
            table = dict(zip(self._keys, gathered))
            attrs.batch = apply_to_collection(attrs.batch, rebuild_batch(table))
        Dispatcher.launch(self, attrs=attrs)


class Metric(Capsule):
    def __init__(self, priority: int = 1000) -> None:
        super().__init__(priority=priority)

    def set(self, attrs: Attributes | None = None) -> None:
        Capsule.set(self, attrs)
        self._step = attrs.launcher.epoch_idx

    def launch(self, attrs: Attributes | None = None) -> None:
        raise NotImplementedError(f"{type(self).__name__}: metric should implement launch()")

    def reset(self, attrs: Attributes | None = None) -> None:
        raise NotImplementedError(f"{type(self).__name__}: metric should implement reset()")
