"""Microbench (N GPUs, torchrun): fused NVLink all-reduce+norm kernel vs NCCL all_reduce + norm kernel
over a flat gradient buffer of n fp32 elements.  Prints per-rank-max times and bus bandwidth."""
import os, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch, torch.distributed as dist
from rocket_b200.backend import NativeBackend, TensorSpec
from rocket_b200 import _native as N
from rocket_b200.comm import alloc_arena, map_arena

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 27
iters = 20
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
dev = torch.device("cuda", local); torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
be = NativeBackend(dev)
sig = N.RK_COMM_SIGNAL_BYTES
scratch = -(-be.scratch_bytes([n]) // 256) * 256
arena = alloc_arena(sig + scratch + 4 * n, dev)
g = arena[sig + scratch:].view(torch.float32)
p = torch.zeros(n, device=dev)
plan = be.make_plan([TensorSpec(p=p, g=g, m=None, v=None, group=0)], scratch=arena[sig:sig + scratch])
class E: pass
eng = E(); eng.device = dev; eng.process_index = rank; eng.num_processes = world
ptrs, mc_ptr, peers = map_arena(arena, eng)
comm = be.comm_create(rank, world, ptrs, sig, mc_ptr)
if rank == 0: print('arena', 'symm' if getattr(arena, '_rk_symm', False) else 'ipc', 'multicast_ptr', hex(mc_ptr))
dist.barrier()
ref = torch.empty(n, device=dev)
def fill():
    torch.manual_seed(1234 + rank); g.normal_(0, 1e-3)
def timed(fn):
    ts = []
    for i in range(iters + 3):
        fill(); torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        if i >= 3: ts.append(e0.elapsed_time(e1))
    t = torch.tensor([sum(ts) / len(ts)], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)
for mc in ([False, True] if mc_ptr else [False]):
  for ctas in (0, 74, 296):
    def fused():
        be.comm_allreduce_norm(comm, plan, 0, plan.n_chunks_norm, 0, ctas, torch.cuda.current_stream().cuda_stream, mc)
        be.l2norm_finalize(plan)
    t_f = timed(fused); nf = float(plan.total_norm)
    if rank == 0: print(f"fused {'multimem' if mc else 'p2p'} ctas={ctas or 148}: {t_f:.3f} ms  busbw {2*(world-1)/world*4*n/t_f/1e6:.1f} GB/s  norm {nf:.6g}")
def nccl():
    dist.all_reduce(g); be.l2norm(plan, 1.0 / world)
t_n = timed(nccl); nn = float(plan.total_norm)
if rank == 0: print(f"nccl+norm: {t_n:.3f} ms  busbw {2*(world-1)/world*4*n/t_n/1e6:.1f} GB/s  norm {nn:.6g}  n={n} W={world}")
# correctness: fused result == sum/W
fill(); mine = g.clone(); dist.all_reduce(mine); mine /= world
fill(); torch.cuda.synchronize(); dist.barrier(); fused(); torch.cuda.synchronize()
if world == 2:
    assert torch.equal(g, mine), float((g - mine).abs().max())
else:
    assert torch.allclose(g, mine, rtol=1e-6, atol=1e-9)
dist.barrier(); be.comm_destroy(comm); dist.destroy_process_group()
if rank == 0: print("ok")
