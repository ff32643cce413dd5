"""Diagnostic (torchrun, N GPUs): where does the fused all-reduce differ from NCCL's?"""
import os, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch, torch.distributed as dist
from rocket_b200 import _native as N
from rocket_b200.backend import NativeBackend, TensorSpec
from rocket_b200.comm import alloc_arena, map_arena
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dev = torch.device("cuda", rank); torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
be = NativeBackend(dev)
sizes = [6, 10, 150, 4097, 48000, 1, 0, 123457, 3_000_001]
offs, total = [], 0
for n in sizes:
    offs.append(total); total += (n + 63) // 64 * 64
sig = N.RK_COMM_SIGNAL_BYTES
scratch = -(-be.scratch_bytes(sizes) // 256) * 256
arena = alloc_arena(sig + scratch + 4 * total, dev)
flat = arena[sig + scratch:].view(torch.float32)
views = [flat[o:o + n] for o, n in zip(offs, sizes)]
specs = [TensorSpec(p=torch.zeros(n, device=dev), g=v, m=None, v=None, group=0) for n, v in zip(sizes, views)]
plan = be.make_plan(specs, scratch=arena[sig:sig + scratch])
class Eng: device, process_index, num_processes = dev, rank, world
ptrs, mc_ptr, peers = map_arena(arena, Eng)
comm = be.comm_create(rank, world, ptrs, sig, mc_ptr)
dist.barrier()
stream = torch.cuda.current_stream().cuda_stream
mode = sys.argv[1] if len(sys.argv) > 1 else "two"
for it in range(2):
    torch.manual_seed(100 * it + rank); flat.normal_(0, 1e-2)
    want = flat.clone(); dist.all_reduce(want); want /= world
    torch.cuda.synchronize(); dist.barrier()
    n = plan.n_chunks_norm; half = n // 2
    if mode == "two":
        be.comm_allreduce_norm(comm, plan, half, n, 0, 0, stream)
        be.comm_allreduce_norm(comm, plan, 0, half, 1, 0, stream)
    else:
        be.comm_allreduce_norm(comm, plan, 0, n, 0, 0, stream)
    torch.cuda.synchronize(); dist.barrier()
    bad = (flat != want).nonzero().flatten()
    msg = f"rank {rank} it {it} mode {mode}: {bad.numel()} mismatches of {total}"
    if bad.numel():
        b = bad.cpu().numpy()
        import numpy as np
        tens = np.searchsorted(np.array(offs), b, side="right") - 1
        msg += f" first {b[:6].tolist()} last {b[-3:].tolist()} tensors {sorted(set(tens.tolist()))} in-tensor-first {(b[:6]-np.array(offs)[tens[:6]]).tolist()}"
        msg += f" vals got {flat[bad[:3]].tolist()} want {want[bad[:3]].tolist()} mine {(want[bad[:3]]*world).tolist()}"
    print(msg, flush=True)
dist.barrier(); be.comm_destroy(comm); del peers; torch.cuda.ipc_collect(); dist.barrier(); dist.destroy_process_group()
