"""Sweep the sharded optimizer step (rk_comm_reduce_scatter_norm + rk_comm_sharded_step) over
grid sizes and the P2P / NVSwitch-multicast store paths.  Device-timed, max over ranks.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node W --master-addr 127.0.0.1 \
        --master-port 29720 tools/bench_sharded_step.py [--params 500000000]
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--params", type=int, default=500_000_000)
    ap.add_argument("--iters", type=int, default=5)
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    from rocket_b200 import _native as N
    from rocket_b200.backend import NativeBackend, TensorSpec
    from rocket_b200.comm import alloc_arena, map_arena
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    be = NativeBackend(dev)
    n = args.params
    sig = N.RK_COMM_SIGNAL_BYTES
    scratch = -(-be.scratch_bytes([n]) // 256) * 256
    arena = alloc_arena(sig + scratch + 8 * n, dev)
    g = arena[sig + scratch:sig + scratch + 4 * n].view(torch.float32)
    p = arena[sig + scratch + 4 * n:].view(torch.float32)
    p.normal_(0, 0.02)
    plan = be.make_plan([TensorSpec(p=p, g=g, m=None, v=None, group=0)],
                        scratch=arena[sig:sig + scratch])
    shard = -(-plan.n_chunks_norm // world) * plan.chunk_norm      # compact moment shards
    m, v = torch.zeros(shard, device=dev), torch.zeros(shard, device=dev)

    class Eng:
        device, process_index, num_processes = dev, rank, world
    ptrs, mc_ptr, keep = map_arena(arena, Eng)
    comm = be.comm_create(rank, world, ptrs, sig, mc_ptr)
    dist.barrier()
    hp = dict(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=1e-2, decoupled=True)
    table = torch.zeros(N.RK_STEP_TABLE_BYTES, dtype=torch.uint8, device=dev)
    host_table = torch.zeros(N.RK_STEP_TABLE_BYTES, dtype=torch.uint8).pin_memory()
    stream = torch.cuda.current_stream().cuda_stream
    rows = []
    step_no = 0
    for use_mc in ([False, True] if mc_ptr else [False]):
        for ctas in (74, 148, 296, 444):
            tx, ts, ta = [], [], []
            for i in range(args.iters + 1):
                step_no += 1
                g.normal_(0, 1e-3)
                be.step_table_write("adamw", [dict(hp, step=step_no)], host_table)
                table.copy_(host_table, non_blocking=True)
                torch.cuda.synchronize()
                dist.barrier()
                e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
                e[0].record()
                be.comm_reduce_scatter_norm(comm, plan, 0, plan.n_chunks_norm, 0, ctas, stream,
                                            use_mc)
                norm = be.l2norm_finalize(plan)
                e[1].record()
                be.comm_sharded_step(comm, plan, "adamw", table.data_ptr(), norm, 1.0, ctas, use_mc,
                                     m, v)
                e[2].record()
                torch.cuda.synchronize()
                # the all-reduce form of the exchange at the same grid, for comparison
                g.normal_(0, 1e-3)
                torch.cuda.synchronize()
                dist.barrier()
                a = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
                a[0].record()
                be.comm_allreduce_norm(comm, plan, 0, plan.n_chunks_norm, 1, ctas, stream, use_mc)
                be.l2norm_finalize(plan)
                a[1].record()
                torch.cuda.synchronize()
                if i:
                    tx.append(e[0].elapsed_time(e[1]))
                    ts.append(e[1].elapsed_time(e[2]))
                    ta.append(a[0].elapsed_time(a[1]))
            t = torch.tensor([statistics.mean(tx), statistics.mean(ts), statistics.mean(ta)],
                             device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            rows.append({"multicast": use_mc, "ctas": ctas,
                         "reduce_scatter_norm_ms": round(float(t[0]), 4),
                         "sharded_step_ms": round(float(t[1]), 4),
                         "allreduce_norm_ms": round(float(t[2]), 4)})
    dist.barrier()
    be.comm_destroy(comm)
    del keep
    if rank == 0:
        print(json.dumps({"world": world, "params": n, "rows": rows}, indent=1))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
