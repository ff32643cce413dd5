"""rk_l2norm_multi_static: the gradient norm for callers that cannot give it scratch memory.

Who needs it: a host that may not allocate on the device when the norm is due -- a runtime
that gave the whole GPU to its own memory pool at start-up, or a thin binding (cgo / JNI /
plugin ABI) that only holds the pointers its framework passes in, gradients and a result
slot, and cannot create a buffer of its own.  The entry point keeps its scratch in a fixed
array inside the library (RK_STATIC_SCRATCH_BYTES per GPU) and allocates nothing.

Checked here, through the C-ABI (ctypes), against the existing rk_l2norm_multi on the same
plans and against the CPU oracle (oracle.l2norm_multi, double accumulation) on the same
gradients: same total norm and per-tensor norms (the kernel is the same and deterministic, so the
bar is the 2 ulp the other norm tests use and the observed difference is zero), on calls made
back to back -- including a small plan right after a large one, whose completion ticket then
sits where the large plan left partial sums, so the entry point's own clearing is what makes
it work -- and that an over-sized tensor list is refused with a clear message.

This file is named so that pytest collects it last: a run that stops at the first failure
has then already run every other test.  It passes on its own in a fresh process as well.
"""
from __future__ import annotations

import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ulp(a: np.ndarray, b: np.ndarray) -> int:
    ia, ib = a.view(np.int32).astype(np.int64), b.view(np.int32).astype(np.int64)
    return int(np.abs(ia - ib).max()) if ia.size else 0


def _plan(sizes, seed):
    from rocket_b200.backend import NativeBackend, TensorSpec
    dev = torch.device("cuda", torch.cuda.current_device())
    be = NativeBackend(dev)
    g = torch.Generator().manual_seed(seed)
    specs, host = [], []
    for n in sizes:
        grad = torch.randn(n, generator=g) * 1e-2
        host.append(grad.numpy())
        grad = grad.to(dev)
        specs.append(TensorSpec(p=grad.clone(), g=grad, m=None, v=None, group=0))
    plan = be.make_plan(specs)
    plan.host_grads = host            # for the oracle
    return be, plan


def _static_norm(plan, scale):
    """Call the new entry point through the C-ABI; the caller passes only inputs and outputs."""
    from rocket_b200 import _native as N
    lib = N.load()
    dev = plan.dev.device
    total = torch.full((1,), -1.0, dtype=torch.float32, device=dev)
    per_tensor = torch.full((plan.n,), -1.0, dtype=torch.float32, device=dev)
    # count torch's device allocations (not bytes in use: garbage left by earlier tests may be
    # collected at any moment and would make a byte count move on its own)
    before = torch.cuda.memory_stats(dev)["allocation.all.allocated"]
    rc = lib.rk_l2norm_multi_static(plan.host.data_ptr(), plan.dev.data_ptr(),
                                    ctypes.c_float(scale), total.data_ptr(),
                                    per_tensor.data_ptr(),
                                    torch.cuda.current_stream().cuda_stream)
    N.check(rc, "rk_l2norm_multi_static")
    assert torch.cuda.memory_stats(dev)["allocation.all.allocated"] == before   # none made
    torch.cuda.synchronize()
    return total.cpu().numpy(), per_tensor.cpu().numpy()


def _reference_norm(be, plan, scale):
    tot = be.l2norm(plan, scale, want_tensor_norms=True)
    torch.cuda.synchronize()
    return tot.cpu().numpy().copy(), plan.tensor_norms.cpu().numpy().copy()


def _oracle_norm(plan, scale):
    import oracle
    tot, per_tensor = oracle.l2norm_multi(plan.host_grads, scale)
    return np.array([tot], np.float32), np.asarray(per_tensor, np.float32)


def test_static_scratch_norm_matches_existing_entry_point_on_consecutive_calls():
    # large list first: ~25K norm chunks, its partial sums cover the low scratch offsets
    be_big, big = _plan([4096 * 6000 + 11, 4096 * 19000 + 3, 77], seed=1)
    # then small lists: their ticket lives at a low offset, inside what `big` wrote
    be_small, small = _plan([6, 10, 150, 4097, 48000, 1, 0, 123457], seed=2)
    be_one, one = _plan([5], seed=3)
    for be, plan, scale in ((be_big, big, 1.0), (be_small, small, 0.5), (be_small, small, 0.5),
                            (be_one, one, 1.0), (be_big, big, 0.25), (be_small, small, 1.0)):
        want_tot, want_tn = _reference_norm(be, plan, scale)
        got_tot, got_tn = _static_norm(plan, scale)
        assert np.isfinite(got_tot).all() and (got_tot > 0).all()
        assert _ulp(got_tot, want_tot) <= 2, (got_tot, want_tot)
        assert _ulp(got_tn, want_tn) <= 2
        cpu_tot, cpu_tn = _oracle_norm(plan, scale)      # CUDA against the CPU restatement
        assert _ulp(got_tot, cpu_tot) <= 2, (got_tot, cpu_tot)
        assert _ulp(got_tn, cpu_tn) <= 2


def test_static_scratch_norm_two_calls_in_a_row_same_plan():
    be, plan = _plan([1000, 70000, 33], seed=9)
    want_tot, want_tn = _reference_norm(be, plan, 1.0)
    first = _static_norm(plan, 1.0)
    second = _static_norm(plan, 1.0)
    cpu_tot, cpu_tn = _oracle_norm(plan, 1.0)
    for got_tot, got_tn in (first, second):
        assert _ulp(got_tot, want_tot) <= 2 and _ulp(got_tn, want_tn) <= 2
        assert _ulp(got_tot, cpu_tot) <= 2 and _ulp(got_tn, cpu_tn) <= 2


def test_static_scratch_norm_refuses_lists_beyond_its_limit():
    from rocket_b200 import _native as N
    from rocket_b200.backend import NativeBackend, TensorSpec
    lib = N.load()
    dev = torch.device("cuda", torch.cuda.current_device())
    be = NativeBackend(dev)
    n = 140_000                                   # > 1 MiB / 8 B norm chunks
    buf = torch.zeros(n, dtype=torch.float32, device=dev)
    specs = [TensorSpec(p=buf[i:i + 1], g=buf[i:i + 1], m=None, v=None, group=0)
             for i in range(n)]
    plan = be.make_plan(specs)
    out = torch.zeros(1, dtype=torch.float32, device=dev)
    rc = lib.rk_l2norm_multi_static(plan.host.data_ptr(), plan.dev.data_ptr(),
                                    ctypes.c_float(1.0), out.data_ptr(), None,
                                    torch.cuda.current_stream().cuda_stream)
    assert rc == -2                               # RK_ERR_TOO_LARGE
    msg = lib.rk_last_error().decode()
    assert "built-in scratch holds 1048576" in msg and "rk_l2norm_multi" in msg
    # the ordinary entry point, with caller-provided scratch, still handles the same list
    assert float(be.l2norm(plan, 1.0)) == 0.0
    rc = lib.rk_l2norm_multi_static(plan.host.data_ptr(), plan.dev.data_ptr(),
                                    ctypes.c_float(1.0), None, None, None)
    assert rc == -1 and "total_norm_dev is null" in lib.rk_last_error().decode()
