"""Every code shape of the fused step (K2/K3) gives the oracle's bits.

The library ships several kernels for the same update -- per-thread 128-bit and 256-bit
ld/st.global walkers and the bulk-copy (cp.async.bulk + mbarrier ring) kernel -- and picks one
by list size.  All of them run the element code of csrc/rk_step_ops.cuh, so each must match
oracle/step_oracle.c bit for bit on the same inputs, including awkward sizes, tensors off the
16-byte grid (scalar / direct path), lists longer than the shared-memory prefix table, skipped
tensors and guard words around every tensor.
"""
from __future__ import annotations

import contextlib

import numpy as np
import pytest
import torch

import oracle
from test_kernels_gpu import ADAM_CASES, SGD_CASES, _backend, _bits, _make_list

pytestmark = pytest.mark.gpu

# (variant number in csrc/rk_step.cu, what it is)
VARIANTS = [(4, "v4 U2 3cta"), (9, "v4 U1 5cta"), (30, "v8 U1 3cta"), (31, "v8 U1 4cta"),
            (33, "v8 U1 5cta"), (21, "bulk 16w t4096 pdl"), (20, "bulk 16w t2048"),
            (22, "bulk 16w t1024"), (23, "bulk 12w t4096"), (24, "bulk 8w t4096"),
            (25, "bulk 16w t4096 no-pdl"), (26, "bulk 8w t2048 2cta")]
IDS = [f"v{v}-{name.replace(' ', '_')}" for v, name in VARIANTS]


@contextlib.contextmanager
def forced_variant(v: int):
    from rocket_b200 import _native as N
    lib = N.load()
    lib.rk_step_variant_override(v)
    try:
        yield
    finally:
        lib.rk_step_variant_override(-1)


# small list (1024-element step chunks) and one above 2^26 elements (4096-element chunks)
LISTS = {
    "awkward": dict(sizes=[6, 10, 150, 1, 4095, 4096, 4097, 48000, 3, 0, 8191, 123457, 1024,
                           2048, 2052], misalign=(2, 5, 11)),
    "big": dict(sizes=[(1 << 26) + 4099, 70001, 5], misalign=(1,)),
}


# the 2^26-element list is what selects 4096-element step chunks; it costs ~1 GB of host copies
# and a CPU oracle pass per step, so only the shapes that differ there run it
CASES = [(v, "awkward") for v, _ in VARIANTS] + [(v, "big") for v in (4, 30, 20, 21)]


@pytest.mark.parametrize("variant,name", CASES, ids=[f"v{v}-{n}" for v, n in CASES])
def test_adam_variants_bit_exact(variant, name):
    be = _backend()
    spec = LISTS[name]
    hp = ADAM_CASES[1]
    specs, host = _make_list(spec["sizes"], seed=variant + 3, groups=2, misalign=spec["misalign"])
    plan = be.make_plan(specs)
    steps = (1, 2) if name == "big" else (1, 2, 7)
    with forced_variant(variant):
        for step in steps:
            g = torch.Generator().manual_seed(100 + step)
            for s, h in zip(specs, host):
                new = torch.randn(s.g.numel(), generator=g) * 1e-2
                new[::9] = 0.0
                s.g.copy_(new)
                h[1][:] = new.numpy()
            zero = step != 2                      # one step that leaves the gradients alone
            norm = be.l2norm(plan, 0.5)
            groups = [dict(hp, step=step), dict(hp, lr=hp["lr"] * 0.5, step=step)]
            be.adam_step(plan, groups, norm, 0.05, 0.5, zero)
            torch.cuda.synchronize()
            tn = float(norm.item())
            for i, (s, h) in enumerate(zip(specs, host)):
                gh = groups[i % 2]
                oracle.adam_step(h[0], h[1], h[2], h[3], lr=gh["lr"], beta1=gh["beta1"],
                                 beta2=gh["beta2"], eps=gh["eps"],
                                 weight_decay=gh["weight_decay"], step=step,
                                 decoupled=gh["decoupled"], grad_scale=0.5, total_norm=tn,
                                 max_norm=0.05, zero_grads=zero)
                for k, t in enumerate((s.p, s.g, s.m, s.v)):
                    assert np.array_equal(_bits(t.cpu().numpy()), _bits(h[k])), \
                        f"variant {variant}: tensor {i} ({spec['sizes'][i]}) stream {'pgmv'[k]} " \
                        f"step {step}"


@pytest.mark.parametrize("variant", [v for v, _ in VARIANTS], ids=IDS)
@pytest.mark.parametrize("case", [0, 2])          # plain SGD (no state), nesterov momentum
def test_sgd_variants_bit_exact(variant, case):
    be = _backend()
    hp = SGD_CASES[case]
    with_m = hp["momentum"] != 0.0
    sizes = LISTS["awkward"]["sizes"]
    specs, host = _make_list(sizes, seed=case + 41, misalign=(0, 7), with_v=False, with_m=with_m)
    plan = be.make_plan(specs)
    with forced_variant(variant):
        for step in (1, 2):
            g = torch.Generator().manual_seed(300 + step)
            for s, h in zip(specs, host):
                new = torch.randn(s.g.numel(), generator=g) * 1e-2
                s.g.copy_(new)
                h[1][:] = new.numpy()
            be.sgd_step(plan, [dict(hp, first_step=step == 1)], None, 0.0, 1.0, True)
            torch.cuda.synchronize()
            for i, (s, h) in enumerate(zip(specs, host)):
                oracle.sgd_step(h[0], h[1], h[2] if with_m else None, first_step=step == 1,
                                total_norm=None, max_norm=0.0, zero_grads=True, **hp)
                assert np.array_equal(_bits(s.p.cpu().numpy()), _bits(h[0])), (variant, i, step)
                assert not s.g.any()
                if with_m:
                    assert np.array_equal(_bits(s.m.cpu().numpy()), _bits(h[2])), (variant, i)


@pytest.mark.parametrize("variant", [21, 22, 30], ids=["bulk4096", "bulk1024", "v8"])
def test_variants_many_tensors_skips_and_guard_words(variant):
    """5000 seven-element tensors (prefix table searched in global memory), every third one
    skipped, one group skipped, all carved out of one allocation per stream with guard words
    between them: updated tensors match the oracle, everything else keeps its bits."""
    from rocket_b200.backend import TensorSpec
    be = _backend()
    n_t, n, gap = 5000, 7, 3
    total = n_t * (n + gap) + gap
    sentinel = -4321.5
    torch.manual_seed(variant)
    streams = {k: torch.full((total,), sentinel, device="cuda") for k in "pgmv"}
    init = {k: (torch.randn(n_t, n) * (1e-2 if k != "v" else 1e-3)).abs_() if k == "v"
            else torch.randn(n_t, n) * 1e-2 for k in "pgmv"}
    specs = []
    for i in range(n_t):
        off = gap + i * (n + gap)
        views = {k: streams[k][off:off + n] for k in "pgmv"}
        for k in "pgmv":
            views[k].copy_(init[k][i])
        specs.append(TensorSpec(p=views["p"], g=views["g"], m=views["m"], v=views["v"],
                                group=i % 3, skip=i % 3 == 0 and i % 2 == 0))
    plan = be.make_plan(specs)
    hp = dict(ADAM_CASES[0], step=4)
    groups = [hp, dict(hp, lr=2e-3), dict(hp, skip=True)]
    with forced_variant(variant):
        be.adam_step(plan, groups, None, 0.0, 1.0, True)
    torch.cuda.synchronize()
    got = {k: streams[k].cpu().numpy() for k in "pgmv"}
    mask = np.ones(total, dtype=bool)
    for i in range(n_t):
        off = gap + i * (n + gap)
        mask[off:off + n] = False
        h = [init[k][i].numpy().copy() for k in "pgmv"]
        untouched = specs[i].skip or i % 3 == 2
        if not untouched:
            gh = groups[i % 3]
            oracle.adam_step(h[0], h[1], h[2], h[3], lr=gh["lr"], beta1=gh["beta1"],
                             beta2=gh["beta2"], eps=gh["eps"], weight_decay=gh["weight_decay"],
                             step=4, decoupled=gh["decoupled"], zero_grads=True)
        for k, a in zip("pgmv", h):
            assert np.array_equal(_bits(got[k][off:off + n]), _bits(a)), (variant, i, k, untouched)
    for k in "pgmv":
        assert (got[k][mask] == sentinel).all(), f"guard words of stream {k} were written"
