"""TEST-ONLY stand-in for `termcolor` (not installed on this image).

Lets the unmodified reference (`/root/reference/rocket/core/loop.py:18`) import.
Only the one function the reference calls is provided; colours are dropped.
"""


def colored(text, *args, **kwargs):
    return str(text)
