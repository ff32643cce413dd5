"""ctypes binding of librocket_b200.so (declared in include/rocket_b200.h).

This is the only place the package touches the C-ABI.  There is no fallback: if the
shared library is missing or a call fails, a RuntimeError with the library's own message is
raised.  Tensors cross the boundary as raw device pointers (``Tensor.data_ptr()``) and
streams as ``cudaStream_t`` handles (``torch.cuda.current_stream().cuda_stream``).
"""
from __future__ import annotations

import ctypes
from ctypes import (POINTER, Structure, c_char_p, c_double, c_float, c_int, c_int32, c_int64,
                    c_size_t, c_uint64, c_void_p)
from pathlib import Path

LIB_PATH = Path(__file__).resolve().parent / "csrc" / "librocket_b200.so"

RK_MAX_GROUPS = 64
RK_TENSOR_SKIP = 1
RK_COMM_SIGNAL_BYTES = 36864
RK_STEP_TABLE_BYTES = 4096
RK_COMM_MAX_BUCKETS = 256
EXPORTS = [
    "rk_version", "rk_last_error", "rk_launch_count", "rk_plan_query", "rk_plan_build",
    "rk_l2norm_multi", "rk_l2norm_multi_static", "rk_adamw_step_multi", "rk_adam_derive",
    "rk_sgd_step_multi",
    "rk_meter_argmax_count", "rk_scalar_accumulate", "rk_stage_create", "rk_stage_submit",
    "rk_stage_wait", "rk_stage_release", "rk_stage_destroy", "rk_l2norm_finalize",
    "rk_comm_create", "rk_comm_allreduce_norm", "rk_comm_destroy", "rk_adam_table_write",
    "rk_sgd_table_write", "rk_adamw_step_multi_devtable", "rk_sgd_step_multi_devtable",
    "rk_comm_reduce_scatter_norm", "rk_comm_sharded_step", "rk_step_variant_override",
]


class TensorDesc(Structure):
    _fields_ = [("p", c_void_p), ("g", c_void_p), ("m", c_void_p), ("v", c_void_p),
                ("numel", c_int64), ("group", c_int32), ("flags", c_int32)]


class AdamGroup(Structure):
    _fields_ = [("lr", c_double), ("beta1", c_double), ("beta2", c_double), ("eps", c_double),
                ("weight_decay", c_double), ("step", c_int64), ("decoupled", c_int32),
                ("skip", c_int32)]


class SgdGroup(Structure):
    _fields_ = [("lr", c_double), ("momentum", c_double), ("dampening", c_double),
                ("weight_decay", c_double), ("nesterov", c_int32), ("first_step", c_int32),
                ("skip", c_int32), ("_pad", c_int32)]


class NativeError(RuntimeError):
    pass


_lib = None


def load() -> ctypes.CDLL:
    """Load the library (once). Raises if it has not been built: no silent fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise NativeError(
            f"{LIB_PATH} is missing. Build it with `python -c 'import __graft_entry__ as g; "
            "g.build()'` (nvcc, sm_100a). rocket_b200 has no CPU or eager fallback.")
    lib = ctypes.CDLL(str(LIB_PATH))
    lib.rk_version.restype = c_int
    lib.rk_last_error.restype = c_char_p
    lib.rk_launch_count.restype = c_uint64
    lib.rk_plan_query.argtypes = [POINTER(c_int64), c_int, POINTER(c_size_t), POINTER(c_size_t)]
    lib.rk_plan_build.argtypes = [POINTER(TensorDesc), c_int, c_void_p, c_size_t]
    lib.rk_l2norm_multi.argtypes = [c_void_p, c_void_p, c_float, c_void_p, c_void_p, c_void_p,
                                    c_void_p]
    lib.rk_l2norm_multi_static.argtypes = [c_void_p, c_void_p, c_float, c_void_p, c_void_p,
                                           c_void_p]
    lib.rk_adamw_step_multi.argtypes = [c_void_p, c_void_p, POINTER(AdamGroup), c_int, c_void_p,
                                        c_float, c_float, c_int, c_void_p]
    lib.rk_adam_derive.argtypes = [POINTER(AdamGroup), POINTER(c_float), POINTER(c_int)]
    lib.rk_sgd_step_multi.argtypes = [c_void_p, c_void_p, POINTER(SgdGroup), c_int, c_void_p,
                                      c_float, c_float, c_int, c_void_p]
    lib.rk_meter_argmax_count.argtypes = [c_void_p, c_int, c_void_p, c_int64, c_int64, c_int64,
                                          c_void_p, c_void_p]
    lib.rk_scalar_accumulate.argtypes = [c_void_p, c_double, c_void_p, c_void_p]
    lib.rk_stage_create.argtypes = [POINTER(c_void_p), c_void_p, c_void_p, c_size_t, c_int, c_int]
    lib.rk_stage_submit.argtypes = [c_void_p, c_int, POINTER(c_void_p), POINTER(c_size_t),
                                    POINTER(c_size_t), c_int, POINTER(c_uint64)]
    lib.rk_stage_wait.argtypes = [c_void_p, c_uint64, c_void_p]
    lib.rk_stage_release.argtypes = [c_void_p, c_int, c_void_p]
    lib.rk_stage_destroy.argtypes = [c_void_p]
    lib.rk_l2norm_finalize.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]
    lib.rk_comm_create.argtypes = [POINTER(c_void_p), c_int, c_int, POINTER(c_void_p), c_void_p,
                                   c_uint64]
    lib.rk_comm_allreduce_norm.argtypes = [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                           c_int, c_int, c_void_p]
    lib.rk_comm_destroy.argtypes = [c_void_p]
    lib.rk_comm_reduce_scatter_norm.argtypes = lib.rk_comm_allreduce_norm.argtypes
    lib.rk_comm_sharded_step.argtypes = [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p,
                                         c_void_p, c_void_p, c_float, c_float, c_int, c_int,
                                         c_void_p]
    lib.rk_adam_table_write.argtypes = [POINTER(AdamGroup), c_int, c_void_p]
    lib.rk_sgd_table_write.argtypes = [POINTER(SgdGroup), c_int, c_void_p]
    lib.rk_adamw_step_multi_devtable.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_float,
                                                 c_float, c_int, c_void_p]
    lib.rk_sgd_step_multi_devtable.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_float,
                                               c_float, c_int, c_void_p]
    lib.rk_step_variant_override.argtypes = [c_int]
    for name in EXPORTS:
        fn = getattr(lib, name)          # AttributeError here = header/library mismatch
        if name not in ("rk_last_error", "rk_launch_count"):
            fn.restype = c_int
    _lib = lib
    return lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().rk_last_error().decode("utf-8", "replace")
        raise NativeError(f"{what} failed (status {rc}): {msg}")


def launch_count() -> int:
    return int(load().rk_launch_count())
