"""Host-side bookkeeping of the sharded optimizer step (flat.py): which rank owns which
elements of a moment buffer.  Runs without a GPU: the plan (chunk tables) is host data built by
the C-ABI library, the ownership rule is the kernels' (norm chunk c belongs to rank c % W,
csrc/rk_comm.cu)."""
from __future__ import annotations

from types import SimpleNamespace

import pytest
import torch

from rocket_b200.backend import Plan, TensorSpec
from rocket_b200.flat import EngineOptimizer, _align

SHAPES = [(5000,), (3, 7), (1,), (4096,), (12289,), (64, 129), (40961,)]


def _layout():
    params = [torch.nn.Parameter(torch.zeros(s)) for s in SHAPES]
    offsets, off = [], 0
    for p in params:
        offsets.append(off)
        off += _align(p.numel())
    flat = torch.zeros(off)
    specs = [TensorSpec(p=flat[o:o + p.numel()], g=flat[o:o + p.numel()], m=None, v=None, group=0)
             for p, o in zip(params, offsets)]
    return params, offsets, off, Plan(specs, torch.device("cpu"), None)


def _opt_for(rank, world, params, offsets, plan):
    opt = EngineOptimizer.__new__(EngineOptimizer)      # only the ownership helpers are used
    opt.engine = SimpleNamespace(num_processes=world, process_index=rank)
    opt._params, opt._offsets = params, offsets
    opt._get_plan = lambda: plan
    return opt


@pytest.mark.parametrize("world", [2, 4, 8])
def test_unowned_elements_are_zeroed_and_every_element_has_one_owner(world):
    params, offsets, total, plan = _layout()
    assert plan.n_chunks_norm > len(SHAPES), "the shapes must span several chunks"
    src = torch.arange(1, total + 1, dtype=torch.float32)
    kept = []
    for rank in range(world):
        flat = src.clone()
        _opt_for(rank, world, params, offsets, plan)._zero_unowned(flat)
        kept.append(flat)
    for i, (p, o) in enumerate(zip(params, offsets)):
        n = p.numel()
        chunk_of = torch.arange(n) // plan.chunk_norm + int(plan.prefix_norm[i])
        for rank in range(world):
            mine = (chunk_of % world) == rank
            seg = kept[rank][o:o + n]
            assert torch.equal(seg[mine], src[o:o + n][mine]), (i, rank)
            assert not seg[~mine].any(), (i, rank)
    # what _gathered does with NCCL: integer SUM of the bit patterns rebuilds the buffer
    bits = sum(k.view(torch.int32) for k in kept)
    for p, o in zip(params, offsets):
        n = p.numel()
        assert torch.equal(bits[o:o + n], src.view(torch.int32)[o:o + n])


def test_sharded_request_without_nvlink_exchange_stays_replicated():
    from host_backend import HostBackend
    from rocket_b200.engine import Engine
    eng = Engine(backend=HostBackend(), device="cpu")
    net = torch.nn.Linear(4, 3)
    opt = eng.prepare_optimizer(torch.optim.AdamW(net.parameters()), sharded=True)
    assert opt.sharded is False and opt.arena is None
