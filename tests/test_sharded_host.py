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
    opt = EngineOptimizer.__new__(EngineOptimizer)      # only the shard helpers are used
    opt.engine = SimpleNamespace(num_processes=world, process_index=rank)
    opt._params, opt._offsets = params, offsets
    opt._get_plan = lambda: plan
    return opt


@pytest.mark.parametrize("world", [2, 4, 8])
def test_moment_shards_scatter_and_reassemble(world):
    """Full per-parameter moments -> each rank's compact shard (1/W of the chunks) -> back.
    Slot of chunk c on rank c % W is c // W: checked element by element against that rule."""
    params, offsets, total, plan = _layout()
    chunk = plan.chunk_norm
    assert plan.n_chunks_norm > len(SHAPES), "the shapes must span several chunks"
    torch.manual_seed(world)
    full = [torch.randn(p.shape) for p in params]
    rows = -(-plan.n_chunks_norm // world)
    shards = torch.zeros(world, rows * chunk)
    opts = [_opt_for(r, world, params, offsets, plan) for r in range(world)]
    for r, opt in enumerate(opts):
        for i in range(len(params)):
            opt._scatter(i, full[i], shards[r])
    # the ownership rule, spelled out independently
    used = torch.zeros(world, rows * chunk, dtype=torch.bool)
    for i, t in enumerate(full):
        flat = t.reshape(-1)
        e = torch.arange(flat.numel())
        c = e // chunk + int(plan.prefix_norm[i])
        owner, pos = c % world, (c // world) * chunk + e % chunk
        assert torch.equal(shards[owner, pos], flat), i
        used[owner, pos] = True
    assert not shards[~used].any()                        # nothing else was written
    # and back (what _gather_chunks + _assemble do after the all-gather)
    chunks = EngineOptimizer._interleave(shards, chunk)
    for i, t in enumerate(full):
        assert torch.equal(opts[0]._assemble(i, chunks), t), i


def test_sharded_request_without_nvlink_exchange_stays_replicated():
    from host_backend import HostBackend
    from rocket_b200.engine import Engine
    eng = Engine(backend=HostBackend(), device="cpu")
    net = torch.nn.Linear(4, 3)
    opt = eng.prepare_optimizer(torch.optim.AdamW(net.parameters()), sharded=True)
    assert opt.sharded is False and opt.arena is None
