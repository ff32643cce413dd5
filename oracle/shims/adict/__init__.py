"""TEST-ONLY stand-in for `adict` (github.com/dsenushkin/adict, unpinned, not installed).

The reference aliases it as `Attributes` (rocket/core/capsule.py:23) and relies on:
dict subclass, attribute access, missing key -> None (capsule.py:31-34), `del a.x`
(loop.py:180), `setdefault` (launcher.py:230-231), construction from kwargs
(loop.py:153-158).  That contract is all this restates.
"""


class adict(dict):
    def __getattr__(self, key):
        if key.startswith("__") and key.endswith("__"):
            raise AttributeError(key)
        return self.get(key)

    def __setattr__(self, key, value):
        self[key] = value

    def __delattr__(self, key):
        self.pop(key, None)
