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
# cases compared GPU against GPU only (no reference golden)
GPU_ONLY_CASES = {
    "ws2_fp16_adamw_clip": dict(optimizer="adamw", max_norm=0.5, accum=2, epochs=1,
                                n_train=64 * 24, n_val=64 * 2 + 5, sched_step=3,
                                input_scale=128.0, mixed_precision="fp16"),
}


def _gpu_worker(rank: int, world: int, port: int, case: str, out_dir: str,
                sync: str = "p2p", graph: bool = False, sharded: bool = False) -> None:
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), TQDM_DISABLE="1",
                      ROCKET_B200_GRAD_SYNC=sync, ROCKET_B200_SHARDED_STEP=str(int(sharded)))
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
    from rocket_b200.engine import Engine
    engines = []
    rk.Launcher.engine_factory = staticmethod(lambda **k: engines.append(Engine(**k)) or engines[-1])
    tracker = MemoryTracker("memory")
    kw = dict({**dist_worker.ALL_CASES, **GPU_ONLY_CASES}[case])
    launcher, h = workloads.build_mnist_graph(
        rk, tracker_backend=tracker, clip_via="kwarg" if kw["max_norm"] else "capsule",
        accuracy_cls=Accuracy, graph=graph, **kw)
    launcher.launch()
    if graph:
        assert launcher._capsules[0]._capsules[1].graphs_captured >= 1, "no step was captured"
        sync = "p2p_graph"
    # the path under test really ran (no silent fall back to the other exchange)
    want = "GradSync" if sync == "nccl" else "P2PGradSync" + ("+sharded" if sharded else "")
    assert engines[0].grad_sync_kinds == [want], engines[0].grad_sync_kinds
    out = dist_worker._collect(h, tracker)
    # optimizer state as torch.optim holds it after the run (Optimizer.destroy exported it)
    for i, p in enumerate(h["net"].parameters()):
        for name, t in h["opt"].state.get(p, {}).items():
            if torch.is_tensor(t):
                out[f"state/{i}/{name}"] = t.detach().cpu().numpy().copy()
    if sharded:
        sync = "p2p_sharded"
    np.savez_compressed(Path(out_dir) / f"{case}_{sync}_rank{rank}.npz", **out)
    import torch.distributed as dist
    if dist.is_initialized():
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("case", list(dist_worker.WS2_CASES))
@pytest.mark.parametrize("sync", ["p2p", "nccl"])
def test_two_gpus_replicas_identical_and_track_reference(case, sync, tmp_path):
    port = 29800 + 2 * list(dist_worker.WS2_CASES).index(case) + (sync == "nccl")
    mp.spawn(_gpu_worker, args=(2, port, case, str(tmp_path), sync), nprocs=2, join=True)
    r0, r1 = (np.load(tmp_path / f"{case}_{sync}_rank{r}.npz") for r in range(2))
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


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_fused_nvlink_allreduce_equals_nccl_path_bitwise(tmp_path):
    """With two ranks a sum has one possible order, 1/2 is exact, and the fused kernel's norm
    partials use the same chunking as the stand-alone norm kernel: the hand-written NVLink
    exchange (+ fused norm) and the NCCL exchange (+ separate norm kernel) must therefore
    produce bit-identical parameters, losses and counts."""
    case = "ws2_adamw_clip"
    for i, sync in enumerate(("p2p", "nccl")):
        mp.spawn(_gpu_worker, args=(2, 29850 + i, case, str(tmp_path), sync), nprocs=2, join=True)
    a = np.load(tmp_path / f"{case}_p2p_rank0.npz")
    b = np.load(tmp_path / f"{case}_nccl_rank0.npz")
    for k in a.files:
        if k.startswith("param/"):
            assert np.array_equal(a[k].view(np.int32), b[k].view(np.int32)), k
    assert a["loss_values"].tolist() == b["loss_values"].tolist()
    assert a["accuracy_history"].tolist() == b["accuracy_history"].tolist()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("case", ["ws2_fp16_sgd", "ws2_fp16_adamw_clip"])
def test_fp16_fused_nvlink_norm_equals_nccl_path_bitwise(case, tmp_path):
    """mixed_precision="fp16" on two GPUs.  NVLink path: the exchange kernel's fused partials are
    of the loss-scaled gradients; the norm of the unscaled ones (overflow check + clipping) is
    that norm times 1/scale, a power of two.  NCCL path: a separate norm pass with the unscale
    folded into its gradient read.  Same skipped steps, same scale, same parameters, bitwise."""
    for i, sync in enumerate(("p2p", "nccl")):
        mp.spawn(_gpu_worker, args=(2, 29856 + i, case, str(tmp_path), sync), nprocs=2, join=True)
    a = np.load(tmp_path / f"{case}_p2p_rank0.npz")
    b = np.load(tmp_path / f"{case}_nccl_rank0.npz")
    c = np.load(tmp_path / f"{case}_p2p_rank1.npz")
    assert len(set(a["lr_values"][:5].tolist())) == 1, "the run should start with skipped steps"
    assert a["lr_values"].tolist() == b["lr_values"].tolist()
    for k in a.files:
        if k.startswith("param/"):
            assert np.array_equal(a[k].view(np.int32), b[k].view(np.int32)), k
            assert np.array_equal(a[k].view(np.int32), c[k].view(np.int32)), k
    assert a["loss_values"].tolist() == b["loss_values"].tolist()
    assert a["accuracy_history"].tolist() == b["accuracy_history"].tolist()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("case", ["ws2_fp16_sgd", "ws2_fp16_adamw_clip"])
def test_fp16_sharded_step_equals_replicated_bitwise(case, tmp_path):
    """mixed_precision="fp16" with Optimizer(sharded=True): the sharded kernel unscales the
    gradients it reads (grad_scale = 1/loss_scale), the overflow check and the clip norm come
    from the reduce-scatter's fused partials.  Skipped steps, parameters and the optimizer
    state torch ends up holding must equal the replicated fp16 run bit for bit on both ranks."""
    mp.spawn(_gpu_worker, args=(2, 29862, case, str(tmp_path), "p2p", False, False), nprocs=2,
             join=True)
    mp.spawn(_gpu_worker, args=(2, 29863, case, str(tmp_path), "p2p", False, True), nprocs=2,
             join=True)
    a = np.load(tmp_path / f"{case}_p2p_rank0.npz")
    assert len(set(a["lr_values"][:5].tolist())) == 1, "the run should start with skipped steps"
    for r in range(2):
        b = np.load(tmp_path / f"{case}_p2p_sharded_rank{r}.npz")
        for k in a.files:
            if k.startswith(("param/", "state/")):
                assert np.array_equal(a[k].view(np.int32), b[k].view(np.int32)), (r, k)
        assert a["accuracy_history"].tolist() == b["accuracy_history"].tolist()
    b0 = np.load(tmp_path / f"{case}_p2p_sharded_rank0.npz")
    assert a["lr_values"].tolist() == b0["lr_values"].tolist()
    assert a["loss_values"].tolist() == b0["loss_values"].tolist()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_graph_captured_step_with_nvlink_exchange_equals_eager(tmp_path):
    """Module(graph=True) on two GPUs: the fused NVLink exchange kernels (device-side epochs),
    the NCCL loss reduction and the optimizer step are nodes of the replayed graph; results
    must equal the eager two-GPU run bit for bit."""
    case = "ws2_adamw_clip"
    mp.spawn(_gpu_worker, args=(2, 29860, case, str(tmp_path), "p2p", False), nprocs=2, join=True)
    mp.spawn(_gpu_worker, args=(2, 29861, case, str(tmp_path), "p2p", True), nprocs=2, join=True)
    a = np.load(tmp_path / f"{case}_p2p_rank0.npz")
    b = np.load(tmp_path / f"{case}_p2p_graph_rank0.npz")
    c = np.load(tmp_path / f"{case}_p2p_graph_rank1.npz")
    for k in a.files:
        if k.startswith("param/"):
            assert np.array_equal(a[k].view(np.int32), b[k].view(np.int32)), k
            assert np.array_equal(b[k].view(np.int32), c[k].view(np.int32)), k
    assert a["loss_values"].tolist() == b["loss_values"].tolist()
    assert a["accuracy_history"].tolist() == b["accuracy_history"].tolist()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("case", ["ws2_adamw_clip", "ws2_sgd"])
def test_graph_captured_sharded_step_equals_eager_replicated(case, tmp_path):
    """Module(graph=True) together with Optimizer(sharded=True): reduce-scatter, the sharded
    step kernel (constants from the device table, barrier epoch on the device) and the
    parameter push are nodes of the replayed graph; the run must equal the eager replicated
    two-GPU run bit for bit on both ranks."""
    mp.spawn(_gpu_worker, args=(2, 29866, case, str(tmp_path), "p2p", False, False), nprocs=2,
             join=True)
    mp.spawn(_gpu_worker, args=(2, 29867, case, str(tmp_path), "p2p", True, True), nprocs=2,
             join=True)
    a = np.load(tmp_path / f"{case}_p2p_rank0.npz")
    for r in range(2):
        b = np.load(tmp_path / f"{case}_p2p_sharded_rank{r}.npz")
        for k in a.files:
            if k.startswith(("param/", "state/")):
                assert np.array_equal(a[k].view(np.int32), b[k].view(np.int32)), (r, k)
        assert a["accuracy_history"].tolist() == b["accuracy_history"].tolist()
    b0 = np.load(tmp_path / f"{case}_p2p_sharded_rank0.npz")
    assert a["loss_values"].tolist() == b0["loss_values"].tolist()


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("case", list(dist_worker.WS2_CASES))
def test_sharded_optimizer_step_equals_replicated_bitwise(case, world, tmp_path):
    """Sharded step (reduce-scatter + owner update + parameter write to all ranks over NVLink)
    against the replicated step after the fused all-reduce: per element the arithmetic is the
    same, so parameters, losses, counts AND the optimizer state torch.optim ends up holding
    (moments gathered from their owners) must agree bit for bit, on every rank.  Four ranks
    take the NVSwitch multicast variants of both kernels."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    port = 29870 + 2 * list(dist_worker.WS2_CASES).index(case)
    mp.spawn(_gpu_worker, args=(world, port, case, str(tmp_path), "p2p", False, False),
             nprocs=world, join=True)
    mp.spawn(_gpu_worker, args=(world, port + 1, case, str(tmp_path), "p2p", False, True),
             nprocs=world, join=True)
    a = np.load(tmp_path / f"{case}_p2p_rank0.npz")
    n_state = 0
    for r in range(world):
        b = np.load(tmp_path / f"{case}_p2p_sharded_rank{r}.npz")
        assert sorted(a.files) == sorted(b.files)
        for k in a.files:
            if k.startswith(("param/", "state/")):
                assert np.array_equal(a[k].view(np.int32), b[k].view(np.int32)), (r, k)
                n_state += k.startswith("state/")
        if r == 0:                                    # trackers log on the main process only
            assert len(b["loss_values"]) > 0
            assert a["loss_values"].tolist() == b["loss_values"].tolist()
        assert a["accuracy_history"].tolist() == b["accuracy_history"].tolist()
    if case == "ws2_adamw_clip":
        assert n_state > 0


def _native_worker_setup(rank: int, world: int, port: int, extra_env=None):
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), **(extra_env or {}))
    root = Path(__file__).resolve().parent.parent
    for p in (str(root), str(root / "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch.distributed as dist
    from rocket_b200.backend import NativeBackend
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    return dist, NativeBackend(dev), dev


def _allreduce_worker(rank: int, world: int, port: int, out_dir: str, arena_kind: str,
                      multicast: bool) -> None:
    """The checks themselves live in tests/parity_checks.py (bench.py runs the same ones as its
    `parity_check` at N > 1, where the driver's 8-GPU box can see them)."""
    dist, be, dev = _native_worker_setup(rank, world, port, {"ROCKET_B200_ARENA": arena_kind})
    import parity_checks
    out = parity_checks.exchange_check(be, rank, world, dev, multicast)
    if out is None:
        Path(out_dir, f"skipped_{rank}").write_text("no multicast support")
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("path", ["ipc-p2p", "symm-p2p", "symm-multicast"])
def test_fused_allreduce_norm_kernel_against_nccl(world, path, tmp_path):
    """The exchange kernel on its own, with the arena shared through CUDA IPC or torch's
    symmetric memory, using P2P loads/stores or the NVSwitch multicast (multimem) path."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    kind, mode = path.split("-")
    port = 29870 + world * 3 + ["ipc-p2p", "symm-p2p", "symm-multicast"].index(path)
    mp.spawn(_allreduce_worker, args=(world, port, str(tmp_path), kind, mode == "multicast"),
             nprocs=world, join=True)
    if (tmp_path / "skipped_0").exists():
        pytest.skip("this box has no NVSwitch multicast support")


def _sharded_step_worker(rank: int, world: int, port: int, out_dir: str, multicast: bool,
                         kind: str, compact: bool) -> None:
    dist, be, dev = _native_worker_setup(rank, world, port)
    import parity_checks
    out = parity_checks.sharded_step_check(be, rank, world, dev, multicast, kind, compact)
    if out is None:
        Path(out_dir, f"skipped_{rank}").write_text("no multicast support")
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("kind", ["adamw", "sgd"])
@pytest.mark.parametrize("path", ["p2p", "multicast", "p2p_fullsize_moments"])
def test_sharded_step_kernel_against_replicated_step(world, kind, path, tmp_path):
    """rk_comm_reduce_scatter_norm + rk_comm_sharded_step on awkward sizes against
    rk_comm_allreduce_norm + the ordinary step kernel: same norm, same parameters on every
    rank, gradients zero-filled, moments equal to the replicated step's on their owners --
    in the rank's compact shard (1/W of the memory; what the engine uses) or in full-size
    per-tensor buffers -- and nothing else written."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    if path == "p2p_fullsize_moments" and kind == "sgd":
        pytest.skip("one optimizer is enough for the alternative moment layout")
    paths = ["p2p", "multicast", "p2p_fullsize_moments"]
    port = 29930 + world * 6 + 3 * (kind == "sgd") + paths.index(path)
    mp.spawn(_sharded_step_worker,
             args=(world, port, str(tmp_path), path == "multicast", kind,
                   path != "p2p_fullsize_moments"),
             nprocs=world, join=True)
    if (tmp_path / "skipped_0").exists():
        pytest.skip("this box has no NVSwitch multicast support")
