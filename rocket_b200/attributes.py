"""`Attributes`: the mutable exchange buffer every capsule handler receives.

The reference aliases the third-party `adict` for this (reference rocket/core/capsule.py:23).
The contract the capsules rely on, and that this class provides without the dependency:

* it is a ``dict`` (tqdm's ``set_postfix`` and ``setdefault`` are called on it:
  loop.py:225, launcher.py:230-231);
* fields are read and written as attributes;
* reading a field that does not exist yields ``None`` instead of raising -- the capsules
  test ``attrs.looper is None``, ``attrs.tracker is None``, ``attrs.batch is not None``
  (loop.py:152, loss.py:103, dataset.py:264);
* ``del attrs.field`` removes it (loop.py:180, launcher.py:312, tracker.py:180).
"""
from __future__ import annotations


class Attributes(dict):
    __slots__ = ()

    def __getattr__(self, name):
        # dunder lookups (copy, pickle, ...) must fail normally
        if name[:2] == "__" and name[-2:] == "__":
            raise AttributeError(name)
        return self.get(name)

    def __setattr__(self, name, value):
        self[name] = value

    def __delattr__(self, name):
        self.pop(name, None)

    def __repr__(self):
        inner = ", ".join(f"{k}={v!r}" for k, v in self.items())
        return f"Attributes({inner})"
