"""The capsule API is unchanged: every public method of every reference class exists here
with the same parameters, kinds and defaults (boundary b1).

Compared against tests/golden/api_signatures.json, frozen from the reference by
tests/golden/make_api_signatures.py; when /root/reference is present (build container) the
frozen file is also re-derived and must still match the reference.
"""
from __future__ import annotations

import inspect
import json
from pathlib import Path

import pytest

import oracle

FROZEN = json.loads((Path(__file__).parent / "golden" / "api_signatures.json").read_text())
CLASSES = [c for c in FROZEN if c not in ("Events", "utils.torch", "utils.collections")]
# the single additive keyword: default None = reference behaviour
ADDITIVE = {("Optimizer", "__init__"): ["max_grad_norm", "sharded"],
            ("Module", "__init__"): ["graph"]}


def _sig(fn):
    return [[p.name, str(p.kind), None if p.default is p.empty else repr(p.default)]
            for p in inspect.signature(fn).parameters.values()]


@pytest.mark.parametrize("cls", CLASSES)
def test_class_surface_matches_reference(cls):
    import rocket_b200 as rk
    mine = getattr(rk, cls)
    for name, want in FROZEN[cls].items():
        assert hasattr(mine, name), f"{cls}.{name} is missing"
        got = _sig(getattr(mine, name))
        extra = ADDITIVE.get((cls, name), [])
        got = [p for p in got if p[0] not in extra]
        assert got == want, f"{cls}.{name}: {got} != {want}"
    for (c, m), names in ADDITIVE.items():
        if c == cls:
            params = inspect.signature(getattr(mine, m)).parameters
            for n in names:
                assert params[n].default in (None, False)   # off = the reference's behaviour


def test_events_utils_and_exports():
    import rocket_b200 as rk
    from rocket_b200.core.capsule import Events
    assert {e.name: e.value for e in Events} == FROZEN["Events"]
    import rocket_b200.utils.collections as rc
    import rocket_b200.utils.torch as rt
    for n, want in FROZEN["utils.torch"].items():
        assert [p for p in inspect.signature(getattr(rt, n)).parameters] == want, n
    for n, want in FROZEN["utils.collections"].items():
        assert [p for p in inspect.signature(getattr(rc, n)).parameters] == want, n
    for name in CLASSES:
        assert name in rk.__all__


def test_install_as_rocket_aliases_submodules():
    import subprocess
    import sys
    code = ("import rocket_b200; rocket_b200.install_as_rocket(); import rocket;"
            "from rocket.core.capsule import Attributes, Capsule;"
            "from rocket.utils.torch import torch_move; import rocket.core.loop;"
            "assert rocket.Looper is rocket_b200.Looper; print('ok')")
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True,
                         cwd=str(Path(__file__).resolve().parent.parent))
    assert out.stdout.strip() == "ok", out.stderr[-500:]


@pytest.mark.skipif(not oracle.reference_available(), reason="/root/reference not present")
def test_frozen_signatures_still_match_the_reference():
    import subprocess
    import sys
    root = Path(__file__).resolve().parent.parent
    out = subprocess.run(
        [sys.executable, "-c",
         "import sys, json; sys.path.insert(0, 'tests/golden'); import make_api_signatures as m;"
         "rk = m.oracle.import_reference();"
         "print(json.dumps({c: m.describe(getattr(rk, c)) for c in m.CLASSES}, sort_keys=True))"],
        capture_output=True, text=True, cwd=str(root))
    assert out.returncode == 0, out.stderr[-800:]
    live = json.loads(out.stdout)
    assert live == {c: FROZEN[c] for c in live}
