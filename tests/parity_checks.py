"""Multi-GPU parity checks shared by tests/test_dist_gpu.py and bench.py's `parity_check`.

Each function runs on every rank of an initialised NCCL process group (one GPU per rank),
raises AssertionError on the first mismatch and returns a dict of what it measured.  They are
checkers: they use the oracle and host arithmetic, and are never part of a timed region.

  exchange_check      K6 (rk_comm_allreduce_norm, P2P or multimem) on an awkward multi-tensor
                      list: replicas bit-identical; every element against the fp64 sum of the W
                      inputs / W computed on the host; fused norm partials against the
                      stand-alone norm kernel (bit-equal) and against oracle.l2norm_multi.
  sharded_step_check  K7 (rk_comm_reduce_scatter_norm + rk_comm_sharded_step) against K6 + the
                      replicated step kernel: same norm, same parameters on every rank, bit for
                      bit; gradients zero-filled; moments equal on their owners.
  replicas_identical  an exact checksum of a list of tensors agrees on all ranks.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist

SIZES = [6, 10, 150, 4097, 48000, 1, 0, 123457, 3_000_001]       # awkward + multi-chunk


def _layout(sizes):
    offs, total = [], 0
    for n in sizes:
        offs.append(total)
        total += (n + 63) // 64 * 64
    return offs, total


class _Eng:
    def __init__(self, dev, rank, world):
        self.device, self.process_index, self.num_processes = dev, rank, world


def replicas_identical(tensors, dev) -> bool:
    """Exact: the int64 sum of every tensor's bit pattern, position-weighted, must agree."""
    acc = torch.zeros(2, dtype=torch.int64, device=dev)
    for i, t in enumerate(tensors):
        bits = t.detach().reshape(-1).view(torch.int32).to(torch.int64)
        acc[0] += bits.sum()
        acc[1] += (bits * ((torch.arange(bits.numel(), device=dev) + i) % 8191 + 1)).sum()
    lo, hi = acc.clone(), acc.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    return bool(torch.equal(lo, hi))


def exchange_check(be, rank: int, world: int, dev, multicast: bool, iters: int = 3,
                   host_sum: bool = True):
    """Returns None if multicast was requested and this box has no multicast mapping."""
    import oracle
    from rocket_b200 import _native as N
    from rocket_b200.backend import TensorSpec
    from rocket_b200.comm import alloc_arena, map_arena
    sizes = SIZES
    offs, total = _layout(sizes)
    sig = N.RK_COMM_SIGNAL_BYTES
    scratch = -(-be.scratch_bytes(sizes) // 256) * 256
    arena = alloc_arena(sig + scratch + 4 * total, dev)
    flat = arena[sig + scratch:].view(torch.float32)
    views = [flat[o:o + n] for o, n in zip(offs, sizes)]
    specs = [TensorSpec(p=torch.zeros(n, device=dev), g=v, m=None, v=None, group=0)
             for n, v in zip(sizes, views)]
    plan = be.make_plan(specs, scratch=arena[sig:sig + scratch])
    ptrs, mc_ptr, peers = map_arena(arena, _Eng(dev, rank, world))
    if multicast and not mc_ptr:
        dist.barrier()
        return None
    comm = be.comm_create(rank, world, ptrs, sig, mc_ptr)
    dist.barrier()
    stream = torch.cuda.current_stream().cuda_stream
    half = plan.n_chunks_norm // 2
    out = {"tensors": len(sizes), "elements": sum(sizes), "iterations": iters,
           "replicas_bit_identical": True, "norm_equals_standalone_kernel": True,
           "max_ulp_vs_fp64_host_sum": 0, "max_err_over_bound": 0.0, "norm_ulp_vs_oracle": 0}
    for it in range(iters):                             # epochs advance, nothing is reset
        torch.manual_seed(100 * it + rank)
        flat.zero_()                                    # padding between tensors stays zero
        for v in views:
            v.normal_(0, 1e-2)
        mine = flat.clone()
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
        if host_sum:
            # ... and against the W inputs summed in fp64 on the host, divided by W.  fp32
            # addition of W numbers in any order is within (W-1) half-ulps of sum|g_r| of the
            # exact sum; in result ulps it can be more where the inputs cancel, so both are
            # reported and the order-independent bound is the assertion.
            inputs = [torch.empty_like(mine) for _ in range(world)]
            dist.all_gather(inputs, mine)
            host = np.stack([t.cpu().numpy().astype(np.float64) for t in inputs])
            exact = host.sum(0) / world
            got_h = got.cpu().numpy()
            bound = (world - 1) * 2.0 ** -24 * np.abs(host).sum(0) / world + 1e-45
            err = np.abs(got_h.astype(np.float64) - exact)
            assert (err <= bound).all(), (it, float((err / bound).max()))
            out["max_err_over_bound"] = max(out["max_err_over_bound"],
                                            float((err / bound).max()))
            exact32 = exact.astype(np.float32)
            ulp = np.abs(got_h.view(np.int32).astype(np.int64)
                         - exact32.view(np.int32).astype(np.int64))
            if world == 2:
                assert int(ulp.max()) == 0, (it, int(ulp.max()))
            out["max_ulp_vs_fp64_host_sum"] = max(out["max_ulp_vs_fp64_host_sum"],
                                                  int(ulp.max()))
            # the norm of the averaged gradients against the CPU oracle (double accumulation)
            want_norm, want_tn = oracle.l2norm_multi(
                [got_h[o:o + n] for o, n in zip(offs, sizes)], 1.0)
            nu = abs(int(norm.cpu().numpy().view(np.int32)[0])
                     - int(np.array([want_norm], np.float32).view(np.int32)[0]))
            assert nu <= 2, (it, float(norm), float(want_norm))
            tn = plan.tensor_norms.cpu().numpy()
            assert np.abs(tn.view(np.int32).astype(np.int64)
                          - want_tn.view(np.int32).astype(np.int64)).max() <= 2
            out["norm_ulp_vs_oracle"] = max(out["norm_ulp_vs_oracle"], nu)
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
    return out


def sharded_step_check(be, rank: int, world: int, dev, multicast: bool, kind: str,
                       compact: bool, iters: int = 3):
    """Returns None if multicast was requested and this box has no multicast mapping."""
    from rocket_b200 import _native as N
    from rocket_b200.backend import TensorSpec
    from rocket_b200.comm import alloc_arena, map_arena
    sizes = SIZES
    groups_of = [0, 1, 0, 1, 0, 1, 0, 1, 0]
    offs, total = _layout(sizes)
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
    ptrs, mc_ptr, peers = map_arena(arena, _Eng(dev, rank, world))
    if multicast and not mc_ptr:
        dist.barrier()
        return None
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
    for it in range(iters):
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
    # every rank ends with the same parameters
    same = replicas_identical([flat_p], dev)
    assert same
    dist.barrier()
    be.comm_destroy(comm)
    del peers
    torch.cuda.ipc_collect()
    dist.barrier()
    return {"tensors": len(sizes), "elements": sum(sizes), "iterations": iters,
            "optimizer": kind, "params_bitwise_equal_replicated_step": True,
            "norm_bitwise_equal": True, "moments_equal_on_owners": True,
            "gradients_zero_filled": True, "replicas_bit_identical": same}
