"""Sweep of the fused NVLink exchange (K6: all-reduce + 1/W + norm partials) over its store path
(P2P loads/stores vs NVSwitch multimem), grid size and multimem loads in flight, next to
NCCL all_reduce + the norm kernel.  Device-timed, max over ranks, JSON lines.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node W --master-addr 127.0.0.1 \
        --master-port 29730 tools/bench_allreduce.py [n_elements] [--out file.jsonl]

bus GB/s = 2(W-1)/W x bytes / time: what a ring would put on each GPU's links per direction;
with in-switch reduction a GPU really sends and receives about `bytes` (wire GB/s = bytes/time).
"""
import json
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from rocket_b200 import _native as N  # noqa: E402
from rocket_b200.backend import NativeBackend, TensorSpec  # noqa: E402
from rocket_b200.comm import alloc_arena, map_arena  # noqa: E402

args = [a for a in sys.argv[1:] if not a.startswith("--")]
out_path = sys.argv[sys.argv.index("--out") + 1] if "--out" in sys.argv else ""
n = int(args[0]) if args else 1 << 28
iters = 10
rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
be = NativeBackend(dev)
sig = N.RK_COMM_SIGNAL_BYTES
scratch = -(-be.scratch_bytes([n]) // 256) * 256
arena = alloc_arena(sig + scratch + 4 * n, dev)
g = arena[sig + scratch:].view(torch.float32)
p = torch.zeros(n, device=dev)
plan = be.make_plan([TensorSpec(p=p, g=g, m=None, v=None, group=0)],
                    scratch=arena[sig:sig + scratch])


class Eng:
    device, process_index, num_processes = dev, rank, world


ptrs, mc_ptr, peers = map_arena(arena, Eng)
comm = be.comm_create(rank, world, ptrs, sig, mc_ptr)
dist.barrier()
rows = []


def fill():
    torch.manual_seed(1234 + rank)
    g.normal_(0, 1e-3)


def timed(fn):
    ts = []
    for i in range(iters + 2):
        fill()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        if i >= 2:
            ts.append(e0.elapsed_time(e1))
    t = torch.tensor([sum(ts) / len(ts)], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def row(kind, ctas, unroll, t):
    r = {"world": world, "elements": n, "path": kind, "ctas": ctas, "mc_unroll": unroll,
         "ms": round(t, 4), "bus_gbs": round(2 * (world - 1) / world * 4 * n / t / 1e6, 1),
         "wire_gbs_in_switch": round(4 * n / t / 1e6, 1)}
    rows.append(r)
    if rank == 0:
        print(json.dumps(r), flush=True)


stream = torch.cuda.current_stream().cuda_stream
for mc in ([False, True] if mc_ptr else [False]):
    for unroll in ((2, 4, 8) if mc else (0,)):
        if mc:
            os.environ["RK_COMM_MC_UNROLL"] = str(unroll)
        for ctas in (74, 148, 296):
            def fused():
                be.comm_allreduce_norm(comm, plan, 0, plan.n_chunks_norm, 0, ctas, stream, mc)
                be.l2norm_finalize(plan)
            row("multimem" if mc else "p2p", ctas, unroll, timed(fused))
os.environ.pop("RK_COMM_MC_UNROLL", None)


def nccl():
    dist.all_reduce(g)
    be.l2norm(plan, 1.0 / world)


row("nccl+norm_kernel", 0, 0, timed(nccl))
# correctness of the last configuration: result == sum / W
fill()
mine = g.clone()
dist.all_reduce(mine)
mine /= world
fill()
torch.cuda.synchronize()
dist.barrier()
be.comm_allreduce_norm(comm, plan, 0, plan.n_chunks_norm, 0, 0, stream, bool(mc_ptr))
torch.cuda.synchronize()
assert torch.equal(g, mine) if world == 2 else torch.allclose(g, mine, rtol=1e-6, atol=1e-9)
dist.barrier()
be.comm_destroy(comm)
del peers
torch.cuda.ipc_collect()
dist.barrier()
dist.destroy_process_group()
if rank == 0 and out_path:
    Path(out_path).write_text("".join(json.dumps(r) + "\n" for r in rows))
