"""Dispatcher: a capsule that owns child capsules and fans every event out to them.

API mirror of the reference's rocket/core/dispatcher.py:22-255.  Children run in
descending ``_priority``; equal priorities keep construction order (the sort is stable,
reference dispatcher.py:54-56).  DESTROY walks the children in reverse so stateful capsules
leave the engine's checkpoint list in LIFO order.
"""
from __future__ import annotations

from rocket_b200.core.capsule import Attributes, Capsule, Events


class Dispatcher(Capsule):
    def __init__(self, capsules: list[Capsule], priority: int = 1000) -> None:
        super().__init__(statefull=False, priority=priority)
        self.guard(capsules)
        self._capsules = sorted(capsules, key=lambda c: c._priority, reverse=True)

    def _fan_out(self, event: Events, attrs, order=None) -> None:
        for capsule in (self._capsules if order is None else order):
            capsule.dispatch(event, attrs)

    def setup(self, attrs: Attributes | None = None) -> None:
        Capsule.setup(self, attrs=attrs)
        self._fan_out(Events.SETUP, attrs)

    def destroy(self, attrs: Attributes | None = None) -> None:
        self._fan_out(Events.DESTROY, attrs, order=reversed(self._capsules))
        Capsule.destroy(self, attrs=attrs)

    def set(self, attrs: Attributes | None = None) -> None:
        Capsule.set(self, attrs=attrs)
        self._fan_out(Events.SET, attrs)

    def reset(self, attrs: Attributes | None = None) -> None:
        Capsule.reset(self, attrs=attrs)
        self._fan_out(Events.RESET, attrs)

    def launch(self, attrs: Attributes | None = None) -> None:
        Capsule.launch(self, attrs=attrs)
        self._fan_out(Events.LAUNCH, attrs)

    def accelerate(self, accelerator) -> None:
        Capsule.accelerate(self, accelerator)
        for capsule in self._capsules:
            capsule.accelerate(accelerator)

    def clear(self) -> None:
        Capsule.clear(self)
        for capsule in self._capsules:
            capsule.clear()

    def guard(self, capsules: list[Capsule]) -> None:
        """Reject anything that is not a Capsule (reference dispatcher.py:219-223)."""
        for capsule in capsules:
            if not isinstance(capsule, Capsule):
                raise ValueError(f"{type(self).__name__} got invalid capsule.")

    def __repr__(self) -> str:
        pad = " " * 4
        own = f"\n{pad}".join(
            f"{k}={str(v).replace(chr(10), chr(10) + pad * 2)}"
            for k, v in self.__dict__.items() if k != "_capsules"
        )
        kids = "\n".join(str(c) for c in self._capsules).replace("\n", f"\n{pad}")
        kids = f"\n_capsules=[\n{pad}{kids}\n]".replace("\n", f"\n{pad}")
        return f"{type(self).__name__}(\n{pad}{own}{kids}\n)"
