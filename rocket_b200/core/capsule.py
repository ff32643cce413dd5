"""Capsule: the unit of the pipeline, and the five lifecycle events it answers.

API mirror of the reference's rocket/core/capsule.py (Events :38-68, Capsule :71-440):
same constructor, same handler names and signatures, same error behaviour.  The handle
stored by ``accelerate()`` is a :class:`rocket_b200.engine.Engine`, which exposes the member
set the reference uses on ``accelerate.Accelerator``.
"""
from __future__ import annotations

import logging
from enum import Enum

from rocket_b200.attributes import Attributes

__all__ = ["Attributes", "Events", "Capsule"]


class Events(Enum):
    """Lifecycle events; the value is the name of the handler method."""

    SETUP = "setup"
    DESTROY = "destroy"
    SET = "set"
    RESET = "reset"
    LAUNCH = "launch"


class _RankLogger(logging.LoggerAdapter):
    """What `accelerate.logging.get_logger` gives the reference (capsule.py:114): a logger
    adapter that also answers the deprecated ``warn`` the Tracker calls (tracker.py:94)."""

    def warn(self, msg, *args, **kwargs):
        self.warning(msg, *args, **kwargs)

    def process(self, msg, kwargs):
        kwargs.pop("main_process_only", None)
        kwargs.pop("in_order", None)
        return msg, kwargs


def get_logger(name: str) -> logging.LoggerAdapter:
    return _RankLogger(logging.getLogger(name), {})


class Capsule:
    """Base of every pipeline component.

    Parameters mirror the reference (capsule.py:104-114): ``statefull`` capsules register
    themselves with the engine for checkpointing; ``priority`` orders siblings inside a
    Dispatcher (higher runs first).
    """

    def __init__(self, statefull: bool = False, logger: logging.Logger | None = None,
                 priority: int = 1000) -> None:
        super().__init__()
        self._priority = priority
        self._statefull = statefull
        self._accelerator = None
        self._logger = logger or get_logger(self.__module__)

    # ---- event handlers -------------------------------------------------------------
    def setup(self, attrs: Attributes | None = None) -> None:
        """SETUP: requires an engine; stateful capsules enrol for checkpointing
        (reference capsule.py:133-141)."""
        self.check_accelerator()
        if self._statefull:
            self._accelerator.register_for_checkpointing(self)
        self._logger.debug(f"{type(self).__name__} initialized.")

    def destroy(self, attrs: Attributes | None = None) -> None:
        """DESTROY: stateful capsules must come off the engine's checkpoint list in LIFO
        order; anything else is an illegal teardown (reference capsule.py:165-176)."""
        if self._statefull:
            last = self._accelerator._custom_objects.pop()
            if last is not self:
                raise RuntimeError(
                    f"{type(self).__name__}: Illegal destroy request. "
                    f"Attempted to remove {type(last).__name__}, "
                    f"but expected {type(self).__name__}."
                )
        self._logger.debug(f"{type(self).__name__} destroyed.")

    def launch(self, attrs: Attributes | None = None) -> None:
        return None

    def set(self, attrs: Attributes | None = None) -> None:
        return None

    def reset(self, attrs: Attributes | None = None) -> None:
        return None

    def dispatch(self, event: Events, attrs: Attributes | None = None) -> None:
        """Route ``event`` to the handler of the same name (reference capsule.py:254)."""
        return getattr(self, event.value)(attrs)

    # ---- engine handle ---------------------------------------------------------------
    def accelerate(self, accelerator) -> None:
        self._accelerator = accelerator

    def clear(self) -> None:
        self._accelerator = None

    def set_logger(self, logger: logging.Logger) -> None:
        self._logger = logger

    def check_accelerator(self) -> None:
        if self._accelerator is None:
            raise RuntimeError(
                f"{type(self).__name__}: accelerator is not defined. "
                "Please, specify it in __init__ function "
                "or set it via .accelerate(accelerator) method."
            )

    # ---- checkpoint contract ------------------------------------------------------------
    def state_dict(self) -> dict:
        if self._statefull:
            raise NotImplementedError("state_dict() must be implemented by stateful subclasses")
        return {}

    def load_state_dict(self, state_dict: dict) -> None:
        if self._statefull:
            raise NotImplementedError("load_state_dict() must be implemented by subclasses")

    def __repr__(self) -> str:
        pad = " " * 4
        body = f"\n{pad}".join(
            f"{k}={str(v).replace(chr(10), chr(10) + pad * 2)}" for k, v in self.__dict__.items()
        )
        return f"{type(self).__name__}(\n{pad}{body}\n)"
