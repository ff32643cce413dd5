"""Container mapping helper (mirror of the reference's rocket/utils/collections.py:17-71).

``apply_to_collection(container, fn, **kw)`` calls ``fn(item, key=<key or index>, **kw)``
on every direct child of a Mapping or Sequence and rebuilds a container of the same type
where that is possible: mutable containers are shallow-copied and refilled, immutable ones
are rebuilt through their constructor, and if the constructor refuses (e.g. namedtuple
subclasses with other signatures) a plain dict / list is returned.
"""
from __future__ import annotations

import collections.abc as abc
import copy
from typing import Callable


def is_collection(x) -> bool:
    return isinstance(x, (abc.Mapping, abc.Sequence))


def apply_to_mapping(container: abc.Mapping, fn: Callable, **kwargs):
    mapped = {k: fn(container[k], key=k, **kwargs) for k in container}
    try:
        if isinstance(container, abc.MutableMapping):
            out = copy.copy(container)
            out.update(mapped)
            return out
        return type(container)(mapped)
    except TypeError:
        return mapped


def apply_to_sequence(container: abc.Sequence, fn: Callable, **kwargs):
    try:
        if isinstance(container, abc.MutableSequence):
            out = copy.copy(container)
            for i, item in enumerate(container):
                out[i] = fn(item, key=i, **kwargs)
            return out
        return type(container)([fn(item, key=i, **kwargs) for i, item in enumerate(container)])
    except TypeError:
        return [fn(item, key=i, **kwargs) for i, item in enumerate(container)]


def apply_to_collection(container, fn: Callable, **kwargs):
    if isinstance(container, abc.Mapping):
        return apply_to_mapping(container, fn, **kwargs)
    if isinstance(container, abc.Sequence):
        return apply_to_sequence(container, fn, **kwargs)
    raise TypeError("{} is not a collection.".format(type(container)))
