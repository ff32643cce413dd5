"""Looper: the per-batch loop that drives the hot path.

API mirror of the reference's rocket/core/loop.py:24-323.  One trip of the loop is one
micro-batch: clear ``attrs.batch``, set the grad mode, LAUNCH the children in priority
order, stop when a child voted ``terminate``.  Nothing in the loop body touches the device
or waits on it: the progress bar prints whatever deferred scalars have already landed.
"""
from __future__ import annotations

from typing import Callable

import torch
from tqdm import tqdm

from rocket_b200.core.capsule import Attributes, Capsule
from rocket_b200.core.dataset import Dataset
from rocket_b200.core.dispatcher import Dispatcher


def _green(text: str) -> str:
    # the reference colours the tag with termcolor (loop.py:203); plain ANSI does the same
    return f"\033[32m{text}\033[0m"


class Looper(Dispatcher):
    def __init__(self, capsules: list[Capsule], tag: str = "Looper", grad_enabled: bool = True,
                 repeats: int | None = None, run_every: int = 1, statefull: bool = True,
                 priority: int = 1000) -> None:
        super().__init__(capsules=capsules, priority=priority)
        self._statefull = statefull
        self._repeats = None
        self._user_defined_repeats = repeats or None
        self._grad_enabled = grad_enabled
        self._run_every = run_every
        self._iter_idx = 0
        self._tag = tag

    def run_if_needed(method: Callable) -> Callable:
        """Skip the handler on epochs that are not a multiple of ``run_every``
        (reference loop.py:91-113)."""
        def wrapper(self, attrs: Attributes | None = None):
            if attrs.launcher.epoch_idx % self._run_every == 0:
                method(self, attrs=attrs)
        return wrapper

    @run_if_needed
    def set(self, attrs: Attributes | None = None) -> None:
        Dispatcher.set(self, attrs=attrs)
        self._repeats = self._user_defined_repeats
        if self._repeats is None:
            self.infer_repeats()
        if self._repeats is None:
            raise RuntimeError(
                f"{type(self).__name__}: infinite loops are not allowed. "
                "Please, specify number of repeats.")
        if attrs.looper is None:
            attrs.looper = Attributes(repeats=self._repeats, state=Attributes(),
                                      terminate=False, tag=self._tag)

    @run_if_needed
    def reset(self, attrs: Attributes | None = None) -> None:
        Dispatcher.reset(self, attrs=attrs)
        self._repeats = None
        del attrs.looper

    @run_if_needed
    def launch(self, attrs: Attributes | None) -> None:
        desc = f"{_green(self._tag)} epoch={attrs.launcher.epoch_idx}, grad={self._grad_enabled}"
        bar = tqdm(range(self._repeats), initial=0, desc=desc,
                   disable=not self._accelerator.is_local_main_process)
        for _ in range(self._repeats):
            attrs.batch = None
            with torch.set_grad_enabled(self._grad_enabled):
                Dispatcher.launch(self, attrs)
            if attrs.looper.terminate:
                break
            bar.set_postfix(attrs.looper.state, refresh=False)
            bar.update(1)
        bar.close()
        self._iter_idx = 0
        self._repeats = -1

    def state_dict(self) -> dict:
        return dict(iter_idx=self._iter_idx)

    def load_state_dict(self, state: dict) -> None:
        self._iter_idx = state.get("iter_idx")

    def guard(self, capsules: list[Capsule]):
        super().guard(capsules)
        for capsule in capsules:
            if isinstance(capsule, Looper):
                raise RuntimeError(f"{type(self).__name__}: internal loopers are not allowed.")

    def infer_repeats(self):
        """Sum the batch counts of the Dataset children (reference loop.py:312-319)."""
        total = sum(c._total for c in self._capsules if isinstance(c, Dataset))
        if total:
            self._repeats = total
        self._logger.info(f"{type(self).__name__} infered {self._repeats} repeats.")
