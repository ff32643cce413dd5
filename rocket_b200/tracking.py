"""Tracker backends: the small part of `accelerate.tracking` the Tracker capsule needs
(reference rocket/core/tracker.py:17,86-105,246-254).  Logging sinks are off the hot path."""
from __future__ import annotations

import os


class GeneralTracker:
    """Base class; a bare instance means "no such backend configured"."""

    name = None
    main_process_only = True

    def log(self, values: dict, step: int | None = None, **kwargs):
        pass

    def log_images(self, values: dict, step: int | None = None, **kwargs):
        pass

    def finish(self):
        pass


class MemoryTracker(GeneralTracker):
    """Keeps records in lists; used by tests and as the sink when tensorboard is absent."""

    def __init__(self, name: str = "memory"):
        self.name = name
        self.scalars: list = []
        self.images: list = []

    def log(self, values, step=None, **kwargs):
        self.scalars.append((step, dict(values)))

    def log_images(self, values, step=None, **kwargs):
        self.images.append((step, values))


class TensorBoardTracker(GeneralTracker):
    name = "tensorboard"

    def __init__(self, run_name: str, logging_dir: str):
        from torch.utils import tensorboard  # raises ImportError when tensorboard is absent
        self.logging_dir = os.path.join(logging_dir or ".", run_name)
        self.writer = tensorboard.SummaryWriter(self.logging_dir)

    def log(self, values, step=None, **kwargs):
        for key, value in values.items():
            if isinstance(value, (int, float)):
                self.writer.add_scalar(key, value, global_step=step)
            elif isinstance(value, str):
                self.writer.add_text(key, value, global_step=step)
            elif isinstance(value, dict):
                self.writer.add_scalars(key, value, global_step=step)
            elif isinstance(value, (list, tuple)):
                for i, v in enumerate(value):
                    self.writer.add_scalar(f"{key}.{i}", v, global_step=step)
        self.writer.flush()

    def log_images(self, values, step=None, **kwargs):
        for key, value in values.items():
            self.writer.add_images(key, value, global_step=step, **kwargs)

    def finish(self):
        self.writer.close()


def make_tracker(name: str, run_name: str, logging_dir: str | None) -> GeneralTracker:
    if name == "tensorboard":
        return TensorBoardTracker(run_name, logging_dir)
    if name == "memory":
        return MemoryTracker(name)
    raise ValueError(f"unknown tracker backend {name!r} (known: tensorboard, memory)")
