"""Deferred scalars: values that are computed on the device and land on the host later.

The reference hands plain Python floats to ``attrs.looper.state`` and to the tracker, which
costs a stream synchronisation per step (reference rocket/core/loss.py:97).  A
``DeferredScalar`` is produced by an async device->pinned-host copy followed by a CUDA
event.  Reading it as a number (``float(x)``, arithmetic, comparisons, pickling) waits for
that event; printing it (``str`` / ``format``, which is what tqdm's postfix does every
iteration) never waits -- it shows the value once it has landed and a placeholder before.
"""
from __future__ import annotations

from typing import Callable


class DeferredScalar:
    __slots__ = ("_slot", "_event", "_finalize", "_value", "_armed")

    def __init__(self, slot, event=None, finalize: Callable | None = None, armed: bool = True):
        self._slot = slot            # 1-element pinned host tensor (or any 0/1-elem tensor)
        self._event = event          # torch.cuda.Event or None (value already valid)
        self._finalize = finalize    # host-side post-processing of the raw number
        self._value = None
        self._armed = armed          # False: the copy has not even been enqueued yet

    def arm(self) -> None:
        """The copy into the slot has been enqueued: record the event that guards it."""
        self._event.record()
        self._armed = True

    # ---- readiness ------------------------------------------------------------------
    def ready(self) -> bool:
        if self._value is not None:
            return True
        return self._armed and (self._event is None or self._event.query())

    def _resolve(self):
        if self._value is None:
            if not self._armed:
                raise RuntimeError("DeferredScalar read before its producer was enqueued")
            if self._event is not None and not self._event.query():
                self._event.synchronize()
            raw = self._slot.item() if self._slot.numel() == 1 else self._slot.tolist()
            self._value = self._finalize(raw) if self._finalize is not None else raw
            self._slot = self._event = self._finalize = None
        return self._value

    # ---- number protocol (blocking) -----------------------------------------------------
    def __float__(self):
        return float(self._resolve())

    def __int__(self):
        return int(self._resolve())

    def item(self):
        return self._resolve()

    def __eq__(self, other):
        return self._resolve() == (other._resolve() if isinstance(other, DeferredScalar) else other)

    def __lt__(self, other):
        return float(self) < float(other)

    def __le__(self, other):
        return float(self) <= float(other)

    def __gt__(self, other):
        return float(self) > float(other)

    def __ge__(self, other):
        return float(self) >= float(other)

    def __hash__(self):
        return id(self)

    def __add__(self, other):
        return float(self) + float(other)

    __radd__ = __add__

    def __sub__(self, other):
        return float(self) - float(other)

    def __rsub__(self, other):
        return float(other) - float(self)

    def __mul__(self, other):
        return float(self) * float(other)

    __rmul__ = __mul__

    def __truediv__(self, other):
        return float(self) / float(other)

    def __rtruediv__(self, other):
        return float(other) / float(self)

    def __reduce__(self):
        return (float, (float(self),))

    # ---- text protocol (never blocks) ---------------------------------------------------
    def __format__(self, spec):
        if not self.ready():
            return "..."
        return format(self._resolve(), spec or ".4g")

    def __str__(self):
        return self.__format__("")

    def __repr__(self):
        return f"DeferredScalar({self.__format__('')})"


def is_ready(value) -> bool:
    return value.ready() if isinstance(value, DeferredScalar) else True


def materialize(value):
    """Plain Python number for a tracker backend (blocks if the copy is still in flight)."""
    if isinstance(value, DeferredScalar):
        return value.item()
    return value
