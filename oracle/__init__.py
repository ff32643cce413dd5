"""oracle/ -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU restatement of the arithmetic of rocket's per-iteration training step, used only as
the checker for the CUDA path: by tests/, by __graft_entry__.smoke() and by bench.py's
cpu_baseline / --impl reference legs.  Nothing under rocket_b200/ imports this package.

  step_oracle.c   plain C restatement of torch's clip / AdamW / Adam / SGD / argmax-count
                  arithmetic (the reference delegates all arithmetic to torch:
                  rocket/core/optimizer.py:129-130), pinned bit-exactly against torch's CPU
                  kernels and against golden vectors produced by the unmodified reference.
  looper_port.py  pure-torch restatement of one reference Looper iteration
                  (rocket/core/loop.py:213-226 and the capsules it drives).
  shims/          test-only stand-ins for accelerate / adict / termcolor so the unmodified
                  reference can be imported in the build container (it cannot travel to
                  the GPU box; golden vectors are committed under tests/golden/).
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
SRC = HERE / "step_oracle.c"
BUILD_DIR = HERE / "_build"
LIB = BUILD_DIR / "liboracle.so"
SHIMS = HERE / "shims"
REFERENCE_ROOT = Path("/root/reference")

_lib = None


def build(force: bool = False) -> Path:
    """gcc the C restatement (a few hundred ms). Idempotent."""
    if not force and LIB.exists() and LIB.stat().st_mtime >= SRC.stat().st_mtime:
        return LIB
    BUILD_DIR.mkdir(exist_ok=True)
    cmd = ["gcc", "-O2", "-ffp-contract=off", "-shared", "-fPIC", "-o", str(LIB), str(SRC), "-lm"]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError(f"gcc failed building the oracle:\n{proc.stderr}")
    return LIB


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        L = ctypes.CDLL(str(build()))
        f32p = ctypes.POINTER(ctypes.c_float)
        i64p = ctypes.POINTER(ctypes.c_int64)
        L.oracle_clip_coef.restype = ctypes.c_float
        L.oracle_clip_coef.argtypes = [ctypes.c_float, ctypes.c_float]
        L.oracle_l2norm_multi.restype = ctypes.c_float
        L.oracle_l2norm_multi.argtypes = [ctypes.POINTER(f32p), i64p, ctypes.c_int,
                                          ctypes.c_float, f32p]
        L.oracle_adam_step.restype = None
        L.oracle_adam_step.argtypes = [f32p, f32p, f32p, f32p, ctypes.c_int64,
                                       ctypes.c_double, ctypes.c_double, ctypes.c_double,
                                       ctypes.c_double, ctypes.c_double, ctypes.c_int64,
                                       ctypes.c_int, ctypes.c_float, ctypes.c_int,
                                       ctypes.c_float, ctypes.c_float, ctypes.c_int]
        L.oracle_sgd_step.restype = None
        L.oracle_sgd_step.argtypes = [f32p, f32p, f32p, ctypes.c_int64, ctypes.c_double,
                                      ctypes.c_double, ctypes.c_double, ctypes.c_double,
                                      ctypes.c_int, ctypes.c_int, ctypes.c_float,
                                      ctypes.c_int, ctypes.c_float, ctypes.c_float,
                                      ctypes.c_int]
        L.oracle_argmax_count.restype = None
        L.oracle_argmax_count.argtypes = [f32p, i64p, ctypes.c_int64, ctypes.c_int64,
                                          ctypes.c_int64, i64p]
        L.oracle_scalar_accumulate.restype = ctypes.c_double
        L.oracle_scalar_accumulate.argtypes = [ctypes.c_double, ctypes.c_float,
                                               ctypes.c_double]
        _lib = L
    return _lib


def _f32(a: np.ndarray):
    assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


# ---- numpy-facing wrappers (arrays are updated in place) ---------------------------------
def clip_coef(total_norm: float, max_norm: float) -> float:
    return float(lib().oracle_clip_coef(total_norm, max_norm))


def l2norm_multi(grads: list[np.ndarray], grad_scale: float = 1.0):
    """-> (total_norm fp32, per-tensor norms fp32[n])."""
    n = len(grads)
    ptrs = (ctypes.POINTER(ctypes.c_float) * n)(*[_f32(g.reshape(-1)) for g in grads])
    numel = (ctypes.c_int64 * n)(*[g.size for g in grads])
    tn = np.zeros(n, dtype=np.float32)
    tot = lib().oracle_l2norm_multi(ptrs, numel, n, grad_scale, _f32(tn))
    return np.float32(tot), tn


def adam_step(p, g, m, v, *, lr, beta1, beta2, eps, weight_decay, step, decoupled=True,
              grad_scale=1.0, total_norm=None, max_norm=0.0, zero_grads=False):
    lib().oracle_adam_step(_f32(p), _f32(g), _f32(m), _f32(v), p.size, lr, beta1, beta2, eps,
                           weight_decay, step, int(decoupled), grad_scale,
                           int(total_norm is not None),
                           0.0 if total_norm is None else float(total_norm), max_norm,
                           int(zero_grads))


def sgd_step(p, g, buf, *, lr, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False,
             first_step=False, grad_scale=1.0, total_norm=None, max_norm=0.0,
             zero_grads=False):
    bufp = _f32(buf) if buf is not None else None
    lib().oracle_sgd_step(_f32(p), _f32(g), bufp, p.size, lr, momentum, dampening,
                          weight_decay, int(nesterov), int(first_step), grad_scale,
                          int(total_norm is not None),
                          0.0 if total_norm is None else float(total_norm), max_norm,
                          int(zero_grads))


def argmax_count(logits: np.ndarray, labels: np.ndarray, counters: np.ndarray) -> None:
    assert logits.dtype == np.float32 and logits.ndim == 2 and labels.dtype == np.int64
    assert counters.dtype == np.int64 and counters.size == 2
    rows, cols = logits.shape
    logits = np.ascontiguousarray(logits)
    labels = np.ascontiguousarray(labels)
    lib().oracle_argmax_count(_f32(logits),
                              labels.ctypes.data_as(ctypes.POINTER(ctypes.c_int64)),
                              rows, cols, cols,
                              counters.ctypes.data_as(ctypes.POINTER(ctypes.c_int64)))


def scalar_accumulate(acc: float, value: float, divisor: float) -> float:
    return float(lib().oracle_scalar_accumulate(acc, value, divisor))


def reference_available() -> bool:
    return (REFERENCE_ROOT / "rocket" / "core" / "loop.py").exists()


def import_reference():
    """Import the UNMODIFIED reference package (build container only) with the shims."""
    import sys
    if not reference_available():
        raise RuntimeError("/root/reference is not present on this machine")
    for p in (str(SHIMS), str(REFERENCE_ROOT)):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.setdefault("ACCELERATE_USE_CPU", "1")
    import rocket  # noqa: F401  (the reference)
    return rocket
