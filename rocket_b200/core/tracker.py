"""Tracker capsule (API mirror of the reference's rocket/core/tracker.py:21-254).

A logging sink: capsules append ``Attributes(step=..., data={tag: value})`` records to
``attrs.tracker.scalars`` / ``.images`` and this capsule forwards them to a backend on the
main process.  Values may be :class:`~rocket_b200.deferred.DeferredScalar`; records whose
values have not landed on the host yet are kept for a later launch instead of blocking the
step, and everything left is flushed (blocking) at RESET / DESTROY.  Steps and tags are
exactly the reference's.
"""
from __future__ import annotations

from typing import List

from rocket_b200.core.capsule import Attributes, Capsule
from rocket_b200.deferred import DeferredScalar, is_ready, materialize
from rocket_b200.tracking import GeneralTracker


class Tracker(Capsule):
    def __init__(self, backend: str = "tensorboard", config: dict = None,
                 priority: int = 200) -> None:
        super().__init__(priority=priority)
        self._backend = backend
        self._tracker = None
        self._config = config or None
        self._pending: list = []

    def setup(self, attrs: Attributes | None = None) -> None:
        Capsule.setup(self, attrs=attrs)
        if isinstance(self._backend, GeneralTracker):
            self._tracker = self._backend
            return
        self._tracker = self._accelerator.get_tracker(self._backend)
        if type(self._tracker) == GeneralTracker:
            self._logger.warn(
                f"Accelerator has not initialized {self._backend}. Trying to create it...")
            try:
                self._accelerator.log_with.append(self._backend)
                self._accelerator.init_trackers("", self._config)
            except Exception as e:
                raise RuntimeError(f"{type(self).__name__} can't create tracker: {e}")
        self._tracker = self._accelerator.get_tracker(self._backend)

    def set(self, attrs: Attributes | None = None) -> None:
        Capsule.set(self, attrs=attrs)
        attrs.tracker = Attributes(scalars=[], images=[])

    def launch(self, attrs: Attributes | None = None) -> None:
        Capsule.launch(self, attrs=attrs)
        if attrs is None or attrs.tracker is None:
            return
        if not attrs.tracker.images and not attrs.tracker.scalars and not self._pending:
            return
        self.log(attrs.tracker.images, attrs.tracker.scalars)
        attrs.tracker = Attributes(scalars=[], images=[])

    def reset(self, attrs: Attributes = None) -> None:
        Capsule.reset(self, attrs=attrs)
        if attrs is None or attrs.tracker is None:
            return
        self._log(attrs.tracker.images, attrs.tracker.scalars, block=True)
        del attrs.tracker

    def destroy(self, attrs: Attributes = None) -> None:
        self._drain(block=True)
        self._tracker = None
        Capsule.destroy(self, attrs=attrs)

    def _drain(self, block: bool) -> None:
        """Forward queued scalar records in order; stop at the first one still in flight."""
        while self._pending:
            record = self._pending[0]
            if not block and not all(is_ready(v) for v in record.data.values()):
                return
            data = {k: materialize(v) for k, v in record.data.items()}
            self._tracker.log(data, step=record.step)
            self._logger.debug(f"Successfully logged scalars to {self._backend}")
            self._pending.pop(0)

    def log(self, images: List[Attributes] | None, scalars: List[Attributes] | None) -> None:
        """Forward records to the backend (reference tracker.py:201-254); scalar records whose
        values are still in flight are kept for the next call."""
        self._log(images, scalars, block=False)

    def _log(self, images, scalars, block: bool) -> None:
        main = self._accelerator.is_main_process
        if images and main:
            try:
                for image in images:
                    self._tracker.log_images(image.data, step=image.step)
                    self._logger.debug(f"Successfully logged images to {self._backend}")
            except Exception as e:
                raise RuntimeError(f"Can't log images: {e}")
        if not main:
            return
        if scalars:
            self._pending.extend(scalars)
        try:
            self._drain(block=block)
        except Exception as e:
            raise RuntimeError(f"Can't log scalars: {e}")


__all__ = ["Tracker", "DeferredScalar"]
