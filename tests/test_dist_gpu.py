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
    kw = dict(dist_worker.WS2_CASES[case])
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


def _allreduce_worker(rank: int, world: int, port: int, out_dir: str, arena_kind: str,
                      multicast: bool) -> None:
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
                      ROCKET_B200_ARENA=arena_kind)
    root = Path(__file__).resolve().parent.parent
    sys.path.insert(0, str(root))
    import torch.distributed as dist
    from rocket_b200 import _native as N
    from rocket_b200.backend import NativeBackend, TensorSpec
    from rocket_b200.comm import alloc_arena, map_arena
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    be = NativeBackend(dev)
    sizes = [6, 10, 150, 4097, 48000, 1, 0, 123457, 3_000_001]       # awkward + multi-chunk
    offs, total = [], 0
    for n in sizes:
        offs.append(total)
        total += (n + 63) // 64 * 64
    sig = N.RK_COMM_SIGNAL_BYTES
    scratch = -(-be.scratch_bytes(sizes) // 256) * 256
    arena = alloc_arena(sig + scratch + 4 * total, dev)
    flat = arena[sig + scratch:].view(torch.float32)
    views = [flat[o:o + n] for o, n in zip(offs, sizes)]
    specs = [TensorSpec(p=torch.zeros(n, device=dev), g=v, m=None, v=None, group=0)
             for n, v in zip(sizes, views)]
    plan = be.make_plan(specs, scratch=arena[sig:sig + scratch])

    class Eng:
        device, process_index, num_processes = dev, rank, world
    ptrs, mc_ptr, peers = map_arena(arena, Eng)
    if multicast and not mc_ptr:
        dist.barrier()
        dist.destroy_process_group()
        Path(out_dir, f"skipped_{rank}").write_text("no multicast support")
        return
    comm = be.comm_create(rank, world, ptrs, sig, mc_ptr)
    dist.barrier()
    stream = torch.cuda.current_stream().cuda_stream
    half = plan.n_chunks_norm // 2
    for it in range(3):                                 # epochs advance, nothing is reset
        torch.manual_seed(100 * it + rank)
        flat.zero_()                                    # padding between tensors stays zero
        for v in views:
            v.normal_(0, 1e-2)
        want = flat.clone()
        dist.all_reduce(want)
        want /= world
        torch.cuda.synchronize()
        dist.barrier()
        be.comm_allreduce_norm(comm, plan, half, plan.n_chunks_norm, 0, 0, stream, multicast)
        be.comm_allreduce_norm(comm, plan, 0, half, 1, 0, stream, multicast)   # two buckets
        norm = be.l2norm_finalize(plan, want_tensor_norms=True).clone()
        torch.cuda.synchronize()
        got = flat.clone()
        # replicas bit-identical
        gathered = [torch.empty_like(got) for _ in range(world)]
        dist.all_gather(gathered, got)
        for q in range(world):
            assert torch.equal(gathered[q], gathered[0]), (it, q)
        # value: sum_r(g_r) / W up to fp32 summation order (exact for W == 2)
        if world == 2:
            assert torch.equal(got, want)
        else:
            assert torch.allclose(got, want, rtol=2e-6, atol=1e-9)
        # fused norm partials == the stand-alone norm kernel on the averaged gradients
        ref_plan = be.make_plan(specs)
        ref = be.l2norm(ref_plan, 1.0, want_tensor_norms=True)
        torch.cuda.synchronize()
        assert torch.equal(norm, ref), (float(norm), float(ref))
        assert torch.equal(plan.tensor_norms, ref_plan.tensor_norms)
    dist.barrier()
    be.comm_destroy(comm)
    del peers
    torch.cuda.ipc_collect()
    dist.barrier()
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
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    root = Path(__file__).resolve().parent.parent
    sys.path.insert(0, str(root))
    import torch.distributed as dist
    from rocket_b200 import _native as N
    from rocket_b200.backend import NativeBackend, TensorSpec
    from rocket_b200.comm import alloc_arena, map_arena
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    be = NativeBackend(dev)
    sizes = [6, 10, 150, 4097, 48000, 1, 0, 123457, 3_000_001]
    groups_of = [0, 1, 0, 1, 0, 1, 0, 1, 0]
    offs, total = [], 0
    for n in sizes:
        offs.append(total)
        total += (n + 63) // 64 * 64
    sig = N.RK_COMM_SIGNAL_BYTES
    scratch = -(-be.scratch_bytes(sizes) // 256) * 256
    arena = alloc_arena(sig + scratch + 8 * total, dev)      # [signals|scratch|g|p]
    flat_g = arena[sig + scratch:sig + scratch + 4 * total].view(torch.float32)
    flat_p = arena[sig + scratch + 4 * total:].view(torch.float32)
    sgd = kind == "sgd"
    # moments: full-size per-tensor buffers in the plan, or this rank's compact shards
    flat_m = None if compact else torch.zeros(total, device=dev)
    flat_v = None if compact or sgd else torch.zeros(total, device=dev)
    ref_p, ref_g, ref_m = (torch.zeros(total, device=dev) for _ in range(3))
    ref_v = None if sgd else torch.zeros(total, device=dev)

    def specs_of(p, g, m, v):
        return [TensorSpec(p=p[o:o + n], g=g[o:o + n], m=None if m is None else m[o:o + n],
                           v=None if v is None else v[o:o + n], group=gi)
                for o, n, gi in zip(offs, sizes, groups_of)]
    plan = be.make_plan(specs_of(flat_p, flat_g, flat_m, flat_v), scratch=arena[sig:sig + scratch])
    ref_plan = be.make_plan(specs_of(ref_p, ref_g, ref_m, ref_v))

    class Eng:
        device, process_index, num_processes = dev, rank, world
    ptrs, mc_ptr, peers = map_arena(arena, Eng)
    if multicast and not mc_ptr:
        dist.barrier()
        dist.destroy_process_group()
        Path(out_dir, f"skipped_{rank}").write_text("no multicast support")
        return
    comm = be.comm_create(rank, world, ptrs, sig, mc_ptr)
    torch.manual_seed(7)
    for o, n in zip(offs, sizes):
        flat_p[o:o + n].normal_(0, 0.1)
    ref_p.copy_(flat_p)
    table = torch.zeros(N.RK_STEP_TABLE_BYTES, dtype=torch.uint8, device=dev)
    host_table = torch.zeros(N.RK_STEP_TABLE_BYTES, dtype=torch.uint8).pin_memory()
    dist.barrier()
    stream = torch.cuda.current_stream().cuda_stream
    half = plan.n_chunks_norm // 2
    chunk = plan.chunk_norm
    rows = -(-plan.n_chunks_norm // world)
    shard_m = torch.zeros(rows * chunk, device=dev) if compact else None
    shard_v = torch.zeros(rows * chunk, device=dev) if compact and not sgd else None
    owner = torch.zeros(total, dtype=torch.bool, device=dev)       # elements this rank owns
    slot_of = torch.zeros(total, dtype=torch.int64, device=dev)    # ... and their shard position
    for i, (o, n) in enumerate(zip(offs, sizes)):
        e = torch.arange(n, device=dev)
        c = e // chunk + int(plan.prefix_norm[i])
        owner[o:o + n] = (c % world) == rank
        slot_of[o:o + n] = (c // world) * chunk + e % chunk
    for it in range(3):
        torch.manual_seed(100 * it + rank)
        flat_g.zero_()
        for o, n in zip(offs, sizes):
            flat_g[o:o + n].normal_(0, 1e-2)
        if it == 1:
            flat_g[offs[4]:offs[4] + 4000].zero_()        # zero gradients: the special values
        mine = flat_g.clone()
        if sgd:
            groups = [dict(lr=0.05, momentum=0.9, dampening=0.0, weight_decay=1e-4,
                           nesterov=gi == 1, first_step=it == 0, skip=False) for gi in range(2)]
        else:
            groups = [dict(lr=1e-3 * (gi + 1), beta1=0.9, beta2=0.999, eps=1e-8,
                           weight_decay=0.01 * gi, step=it + 1, decoupled=True, skip=False)
                      for gi in range(2)]
        max_norm = 0.0 if it == 2 else 0.05
        # replicated arm: fused all-reduce, then the ordinary step kernel on private copies
        torch.cuda.synchronize(); dist.barrier()
        be.comm_allreduce_norm(comm, plan, half, plan.n_chunks_norm, 0, 0, stream, multicast)
        be.comm_allreduce_norm(comm, plan, 0, half, 1, 0, stream, multicast)
        norm_a = be.l2norm_finalize(plan).clone()
        ref_g.copy_(flat_g)
        (be.sgd_step if sgd else be.adam_step)(ref_plan, groups, norm_a if max_norm else None,
                                               max_norm, 1.0, True)
        torch.cuda.synchronize(); dist.barrier()
        # sharded arm on the same inputs
        flat_g.copy_(mine)
        torch.cuda.synchronize(); dist.barrier()
        be.comm_reduce_scatter_norm(comm, plan, half, plan.n_chunks_norm, 0, 0, stream, multicast)
        be.comm_reduce_scatter_norm(comm, plan, 0, half, 1, 0, stream, multicast)
        norm_b = be.l2norm_finalize(plan)
        be.step_table_write(kind, groups, host_table)
        table.copy_(host_table, non_blocking=True)
        be.comm_sharded_step(comm, plan, kind, table.data_ptr(), norm_b if max_norm else None,
                             max_norm, 0, multicast, shard_m, shard_v)
        torch.cuda.synchronize()
        assert torch.equal(norm_a, norm_b), (it, float(norm_a), float(norm_b))
        assert torch.equal(flat_p.view(torch.int32), ref_p.view(torch.int32)), \
            (it, int((flat_p != ref_p).sum()))
        assert not flat_g.any(), it
        if compact:
            pairs = [(shard_m, ref_m)] + ([] if sgd else [(shard_v, ref_v)])
            for shard, ref_s in pairs:
                # owned elements sit in slot c // W of the shard; every other slot stays zero
                assert torch.equal(shard[slot_of[owner]].view(torch.int32),
                                   ref_s[owner].view(torch.int32)), it
                untouched = torch.ones_like(shard, dtype=torch.bool)
                untouched[slot_of[owner]] = False
                assert not shard[untouched].any(), it
        else:
            pairs = [(flat_m, ref_m)] + ([] if sgd else [(flat_v, ref_v)])
            for mine_s, ref_s in pairs:
                assert torch.equal(mine_s[owner].view(torch.int32),
                                   ref_s[owner].view(torch.int32)), it
                assert not mine_s[~owner].any(), it       # other ranks' chunks are never touched
    dist.barrier()
    be.comm_destroy(comm)
    del peers
    torch.cuda.ipc_collect()
    dist.barrier()
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
