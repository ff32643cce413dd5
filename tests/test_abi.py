"""The C-ABI library loads and exports every symbol include/rocket_b200.h declares.
No compute calls: this runs without a GPU."""
from __future__ import annotations

import ctypes
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
HEADER = ROOT / "include" / "rocket_b200.h"


def declared_symbols() -> list[str]:
    text = re.sub(r"/\*.*?\*/", "", HEADER.read_text(), flags=re.S)
    return sorted(set(re.findall(r"\b(rk_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_expected_surface():
    names = declared_symbols()
    for must in ("rk_l2norm_multi", "rk_adamw_step_multi", "rk_sgd_step_multi",
                 "rk_meter_argmax_count", "rk_scalar_accumulate", "rk_stage_create",
                 "rk_plan_build", "rk_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol():
    from rocket_b200._build import build_native
    lib = ctypes.CDLL(str(build_native()))
    missing = [n for n in declared_symbols() if not hasattr(lib, n)]
    assert not missing, missing
    lib.rk_version.restype = ctypes.c_int
    assert lib.rk_version() == 100


def test_python_binding_lists_the_same_symbols():
    from rocket_b200 import _native
    assert sorted(_native.EXPORTS) == declared_symbols()


def test_plan_host_side_is_checked_without_a_gpu():
    """rk_plan_* are host functions: exercise their validation here."""
    from rocket_b200 import _native as N
    lib = N.load()
    numel = (ctypes.c_int64 * 3)(10, 5000, 0)
    pb, sb = ctypes.c_size_t(), ctypes.c_size_t()
    assert lib.rk_plan_query(numel, 3, ctypes.byref(pb), ctypes.byref(sb)) == 0
    assert pb.value > 0 and sb.value > 0
    assert lib.rk_plan_query(numel, 0, ctypes.byref(pb), ctypes.byref(sb)) == -1
    assert b"n_tensors" in lib.rk_last_error()
    bad = (ctypes.c_int64 * 1)(-4)
    assert lib.rk_plan_query(bad, 1, ctypes.byref(pb), ctypes.byref(sb)) == -1
    d = (N.TensorDesc * 1)()
    d[0].numel, d[0].group = 16, 99
    buf = ctypes.create_string_buffer(4096)
    assert lib.rk_plan_build(d, 1, buf, 4096) == -1 and b"group" in lib.rk_last_error()
    g = N.AdamGroup(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=1e-2, step=3,
                    decoupled=1, skip=0)
    out, flags = (ctypes.c_float * 9)(), ctypes.c_int()
    assert lib.rk_adam_derive(ctypes.byref(g), out, ctypes.byref(flags)) == 0
    import numpy as np
    want = [1 - 1e-3 * 1e-2, 1e-2, 1 - 0.9, np.float32(1 - 0.9) - np.float32(1), 0.999,
            1 - 0.999, (1 - 0.999 ** 3.0) ** 0.5, 1e-8, -(1e-3 / (1 - 0.9 ** 3.0))]
    assert [np.float32(x) for x in out] == [np.float32(x) for x in want]


def test_product_package_never_imports_the_oracle_or_the_test_double():
    """oracle/ and tests/host_backend.py are checkers: nothing under rocket_b200/ may import,
    load or execute them (a product path through the CPU restatement would void parity)."""
    import re
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent / "rocket_b200"
    pat = re.compile(r"^\s*(from|import)\s+(oracle|host_backend)\b"
                     r"|#\s*include\s*[<\"][^>\"]*oracle|(CDLL|dlopen)\([^)]*oracle", re.M)
    offenders = [str(p.relative_to(root)) for p in root.rglob("*")
                 if p.suffix in (".py", ".cu", ".cuh", ".h") and pat.search(p.read_text())]
    assert offenders == []
