"""Small driver for compute-sanitizer: every single-GPU kernel of the library (K1 norm, K2
AdamW + clip + zero, K3 SGD nesterov, device-table variants, K4 argmax-count and accumulate) on
awkward tensor sizes, through the C-ABI.  Results are checked elsewhere (tests/); this only
has to touch every code path cheaply enough to run under memcheck.

    compute-sanitizer --tool memcheck --error-exitcode 1 python tools/sanitize_kernels.py
"""
from __future__ import annotations

import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import torch  # noqa: E402

from rocket_b200 import _native as N  # noqa: E402
from rocket_b200.backend import NativeBackend, TensorSpec  # noqa: E402


def main() -> None:
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    be = NativeBackend(dev)
    sizes = [6, 10, 150, 4097, 1, 70001, 3]
    torch.manual_seed(0)

    def specs(with_v: bool, misalign: bool = False):
        out = []
        for i, n in enumerate(sizes):
            # odd element offsets exercise the scalar (non-vectorised) paths
            def buf():
                t = torch.randn(n + 1, device=dev) * 0.01
                return t[1:] if misalign and i % 2 else t[:n]
            out.append(TensorSpec(p=buf(), g=buf(), m=torch.zeros(n, device=dev),
                                  v=torch.zeros(n, device=dev) if with_v else None,
                                  group=i % 2))
        return out

    for misalign in (False, True):
        plan = be.make_plan(specs(True, misalign))
        adam = [dict(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01 * gi, step=1,
                     decoupled=True, skip=False) for gi in range(2)]
        norm = be.l2norm(plan, 1.0, want_tensor_norms=True)
        be.adam_step(plan, adam, norm, 1.0, 1.0, True)
        table = torch.zeros(N.RK_STEP_TABLE_BYTES, dtype=torch.uint8, device=dev)
        host = torch.zeros(N.RK_STEP_TABLE_BYTES, dtype=torch.uint8).pin_memory()
        be.step_table_write("adamw", adam, host)
        table.copy_(host)
        be.step_devtable("adamw", plan, table, norm, 1.0, 1.0, True)
        plan_s = be.make_plan(specs(False, misalign))
        sgd = [dict(lr=0.05, momentum=0.9, dampening=0.0, weight_decay=1e-4, nesterov=gi == 1,
                    first_step=True, skip=False) for gi in range(2)]
        be.sgd_step(plan_s, sgd, None, 0.0, 0.5, True)
        be.step_table_write("sgd", sgd, host)
        table.copy_(host)
        be.step_devtable("sgd", plan_s, table, None, 0.0, 1.0, True)
    logits = torch.randn(37, 10, device=dev)
    labels = torch.randint(0, 10, (37,), device=dev)
    counters = torch.zeros(2, dtype=torch.int64, device=dev)
    be.argmax_count(logits, labels, counters)
    acc = torch.zeros(1, dtype=torch.float64, device=dev)
    be.scalar_accumulate(torch.tensor(1.5, device=dev), 2.0, acc)
    torch.cuda.synchronize()
    assert int(counters[1]) == 37 and float(acc) == 0.75
    print("kernels ran:", be.launch_count(), "launches")


if __name__ == "__main__":
    main()
