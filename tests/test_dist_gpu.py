"""The N>1 path on real GPUs (NCCL): needs >= 2 devices, otherwise skipped.

Two ranks run the capsule graph on the CUDA engine.  Replicas must end bit-identical
(identical gradients after the in-place SUM all-reduce + deterministic norm kernel => the
same clip coefficient and the same update everywhere); the run must track the golden
produced by the unmodified reference under DDP on CPU up to the GPU-vs-CPU forward/backward
difference (loose bars: the tight arithmetic bars live in test_kernels_gpu / test_engine_gpu /
test_dist_cpu).
"""
from __future__ import annotations

import os
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

import dist_worker

pytestmark = pytest.mark.gpu
GOLDEN = Path(__file__).parent / "golden"


def _gpu_worker(rank: int, world: int, port: int, case: str, out_dir: str) -> None:
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), TQDM_DISABLE="1")
    root = Path(__file__).resolve().parent.parent
    for p in (str(root), str(root / "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cudnn.deterministic = True
    import rocket_b200 as rk
    import workloads
    from rocket_b200.metrics import Accuracy
    from rocket_b200.tracking import MemoryTracker
    tracker = MemoryTracker("memory")
    kw = dict(dist_worker.WS2_CASES[case])
    launcher, h = workloads.build_mnist_graph(
        rk, tracker_backend=tracker, clip_via="kwarg" if kw["max_norm"] else "capsule",
        accuracy_cls=Accuracy, **kw)
    launcher.launch()
    np.savez_compressed(Path(out_dir) / f"{case}_rank{rank}.npz",
                        **dist_worker._collect(h, tracker))
    import torch.distributed as dist
    if dist.is_initialized():
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("case", list(dist_worker.WS2_CASES))
def test_two_gpus_replicas_identical_and_track_reference(case, tmp_path):
    port = 29800 + list(dist_worker.WS2_CASES).index(case)
    mp.spawn(_gpu_worker, args=(2, port, case, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = (np.load(tmp_path / f"{case}_rank{r}.npz") for r in range(2))
    gold = np.load(GOLDEN / f"{case}.npz")
    for k in gold.files:
        if k.startswith("param/"):
            assert np.array_equal(r0[k].view(np.int32), r1[k].view(np.int32)), k
            assert np.abs(r0[k] - gold[k]).max() <= 1e-3 * np.abs(gold[k]).max(), k
    assert r0["loss_steps"].tolist() == gold["loss_steps"].tolist()
    assert r0["lr_values"].tolist() == gold["lr_values"].tolist()
    np.testing.assert_allclose(r0["loss_values"], gold["loss_values"], rtol=1e-4)
    assert r0["accuracy_history"].tolist() == r1["accuracy_history"].tolist()
    assert np.abs(r0["accuracy_history"] - gold["accuracy_history"]).max() <= 2
