"""Generate golden vectors by running the UNMODIFIED reference.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

Imports dsenushkin/rocket from /root/reference (read-only) with the test-only shims for its
three uninstallable dependencies (oracle/shims: accelerate, adict, termcolor), runs the
corrected mnist-shaped capsule graph (tests/workloads.py) on CPU with
``torch.optim.*(foreach=False)``, and freezes: parameters after training, the loss series
and lr series the Loss / Optimizer capsules posted to the tracker, and the eval counts.
The .npz files next to this script are committed; tests compare both the oracle and the
CUDA engine against them.  torch version and CPU flags are recorded in each file because
torch's CPU kernels (MKL sqrt in particular) define the last bit.
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

import oracle  # noqa: E402

CASES = {
    # name: kwargs of workloads.build_mnist_graph
    "lenet_adamw_accum2": dict(optimizer="adamw", max_norm=None, accum=2, epochs=2,
                               n_train=64 * 25),
    "lenet_adamw_clip": dict(optimizer="adamw", max_norm=0.5, accum=2, epochs=2,
                             n_train=64 * 25),
    "lenet_sgd_clip": dict(optimizer="sgd", max_norm=1.0, accum=1, epochs=1, n_train=64 * 20),
    "lenet_adam_l2": dict(optimizer="adam", max_norm=None, accum=3, epochs=1, n_train=64 * 20),
    # mixed_precision="fp16": torch's GradScaler through the accelerate shim (on the CPU -- real
    # accelerate wants a GPU for fp16; the scaler arithmetic is the same).  Inputs x128 make the
    # fp16 weight gradients overflow until the loss scale has backed off: 4-5 skipped steps.
    "lenet_fp16_adamw_clip": dict(optimizer="adamw", max_norm=0.5, accum=2, epochs=1,
                                  n_train=64 * 20, sched_step=3, input_scale=128.0,
                                  mixed_precision="fp16"),
    "lenet_fp16_sgd": dict(optimizer="sgd", max_norm=None, accum=1, epochs=1, n_train=64 * 20,
                           sched_step=3, lr=2e-5, input_scale=128.0, mixed_precision="fp16"),
}


def run_case(rk, name: str, kwargs: dict) -> dict:
    import workloads
    from accelerate.tracking import MemoryTracker
    tracker = MemoryTracker("memory")
    launcher, h = workloads.build_mnist_graph(rk, tracker_backend=tracker, **kwargs)
    launcher.launch()
    out = {}
    for k, p in h["net"].named_parameters():
        out[f"param/{k}"] = p.detach().cpu().numpy().copy()
    losses = [(s, d["train_loss"]) for s, d in tracker.scalars if "train_loss" in d]
    lrs = [(s, d["opt.lr.0"]) for s, d in tracker.scalars if "opt.lr.0" in d]
    out["loss_steps"] = np.array([s for s, _ in losses], dtype=np.int64)
    out["loss_values"] = np.array([v for _, v in losses], dtype=np.float64)
    out["lr_steps"] = np.array([s for s, _ in lrs], dtype=np.int64)
    out["lr_values"] = np.array([v for _, v in lrs], dtype=np.float64)
    out["accuracy_history"] = np.array(h["accuracy"].history, dtype=np.int64)
    out["meta"] = np.array([f"torch={torch.__version__}", f"case={name}",
                            f"kwargs={sorted(kwargs.items())}",
                            f"cpu_capability={torch.backends.cpu.get_cpu_capability()}"])
    return out


def main() -> None:
    rk = oracle.import_reference()
    assert "/root/reference" in rk.__file__, rk.__file__
    torch.set_num_threads(1)
    torch.use_deterministic_algorithms(True)
    here = Path(__file__).resolve().parent
    only = set(sys.argv[1:])              # optional: regenerate just the named cases
    for name, kwargs in CASES.items():
        if only and name not in only:
            continue
        out = run_case(rk, name, kwargs)
        np.savez_compressed(here / f"{name}.npz", **out)
        print(name, "steps", len(out["loss_values"]), "final loss", out["loss_values"][-1],
              "acc", out["accuracy_history"].tolist())


if __name__ == "__main__":
    main()
