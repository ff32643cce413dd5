"""Build librocket_b200.so in-tree with nvcc for sm_100a (B200).

The shared object is a plain C-ABI library (see include/rocket_b200.h); it does not link
against torch, so there is no ABI coupling and the same file is what a cgo / JNI /
ctypes host would load.  nvcc cross-compiles without a GPU, so this runs in the build
container; the resulting .so travels to the GPU box with the repository snapshot.
"""
from __future__ import annotations

import os
import shutil
import subprocess
from pathlib import Path

CSRC = Path(__file__).resolve().parent / "csrc"
LIB_PATH = CSRC / "librocket_b200.so"
SOURCES = ["rk_api.cu", "rk_norm.cu", "rk_step.cu", "rk_meter.cu", "rk_stage.cu", "rk_comm.cu"]
HEADERS = [CSRC / "rk_common.cuh", CSRC / "rk_step_ops.cuh", CSRC / "rk_step_bulk.cuh", CSRC.parent.parent / "include" / "rocket_b200.h"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-O3,-Wall",
    "-shared",
]


def find_nvcc() -> str:
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found: cannot build librocket_b200.so")
    return nvcc


def is_stale() -> bool:
    if not LIB_PATH.exists():
        return True
    built = LIB_PATH.stat().st_mtime
    deps = [CSRC / s for s in SOURCES] + HEADERS
    return any(d.stat().st_mtime > built for d in deps)


def build_native(force: bool = False, verbose: bool = False) -> Path:
    """Compile every .cu under csrc/ into one shared library. Returns its path."""
    if not force and not is_stale():
        return LIB_PATH
    cmd = [find_nvcc(), *NVCC_FLAGS]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    cmd += ["-o", str(LIB_PATH), *[str(CSRC / s) for s in SOURCES], "-lpthread"]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError(f"nvcc failed ({proc.returncode}):\n{proc.stdout}\n{proc.stderr}")
    if verbose:
        print(proc.stderr)
    return LIB_PATH


if __name__ == "__main__":
    import sys
    print(build_native(force=True, verbose="-v" in sys.argv))
