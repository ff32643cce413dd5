"""Two-process (gloo, CPU) workers for the multi-rank tests and golden generation."""
from __future__ import annotations

import os
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
for p in (str(ROOT), str(ROOT / "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

WS2_CASES = {
    "ws2_adamw_clip": dict(optimizer="adamw", max_norm=0.5, accum=2, epochs=2, n_train=64 * 11,
                           n_val=64 * 3 + 13),
    "ws2_sgd": dict(optimizer="sgd", max_norm=None, accum=1, epochs=1, n_train=64 * 9,
                    n_val=64 * 2 + 5),
}


# fp16 + GradScaler under DDP: every rank runs its own scaler on the averaged (still scaled)
# gradients, so all ranks skip the same steps.  Kept apart from WS2_CASES, which the GPU tests
# parametrize over.
WS2_FP16_CASES = {
    "ws2_fp16_sgd": dict(optimizer="sgd", max_norm=None, accum=1, epochs=1, n_train=64 * 24,
                         n_val=64 * 2 + 5, sched_step=3, lr=2e-5, input_scale=128.0,
                         mixed_precision="fp16"),
}
ALL_CASES = {**WS2_CASES, **WS2_FP16_CASES}


def _env(rank: int, world: int, port: int) -> None:
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), TQDM_DISABLE="1",
                      ACCELERATE_USE_CPU="1")
    torch.set_num_threads(1)


def _collect(h, tracker) -> dict:
    out = {f"param/{k}": p.detach().cpu().numpy().copy() for k, p in h["net"].named_parameters()}
    losses = [(s, float(d["train_loss"])) for s, d in tracker.scalars if "train_loss" in d]
    lrs = [(s, d["opt.lr.0"]) for s, d in tracker.scalars if "opt.lr.0" in d]
    out["loss_steps"] = np.array([s for s, _ in losses], dtype=np.int64)
    out["loss_values"] = np.array([v for _, v in losses], dtype=np.float64)
    out["lr_values"] = np.array([v for _, v in lrs], dtype=np.float64)
    out["accuracy_history"] = np.array(h["accuracy"].history, dtype=np.int64)
    return out


def reference_worker(rank: int, world: int, port: int, case: str, out_dir: str) -> None:
    """The UNMODIFIED reference under torch DDP/gloo (build container only)."""
    _env(rank, world, port)
    import oracle
    rk = oracle.import_reference()
    import workloads
    from accelerate.tracking import MemoryTracker
    tracker = MemoryTracker("memory")
    launcher, h = workloads.build_mnist_graph(rk, tracker_backend=tracker, **ALL_CASES[case])
    launcher.launch()
    if rank == 0:
        np.savez_compressed(Path(out_dir) / f"{case}.npz", **_collect(h, tracker))
    import torch.distributed as dist
    if dist.is_initialized():
        dist.destroy_process_group()


def engine_worker(rank: int, world: int, port: int, case: str, out_dir: str,
                  device_metric: bool) -> None:
    """rocket_b200 with the host test double under gloo."""
    _env(rank, world, port)
    import rocket_b200 as rk
    import workloads
    from host_backend import HostBackend
    from rocket_b200.engine import Engine
    from rocket_b200.metrics import Accuracy
    from rocket_b200.tracking import MemoryTracker
    rk.Launcher.engine_factory = staticmethod(
        lambda **kw: Engine(backend=HostBackend(), device="cpu", **kw))
    tracker = MemoryTracker("memory")
    kw = dict(ALL_CASES[case])
    launcher, h = workloads.build_mnist_graph(
        rk, tracker_backend=tracker, clip_via="kwarg" if kw["max_norm"] else "capsule",
        accuracy_cls=Accuracy if device_metric else None, **kw)
    launcher.launch()
    out = _collect(h, tracker)
    np.savez_compressed(Path(out_dir) / f"{case}_rank{rank}.npz", **out)
    import torch.distributed as dist
    if dist.is_initialized():
        dist.destroy_process_group()
