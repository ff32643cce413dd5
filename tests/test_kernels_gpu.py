"""GPU parity tests for the C-ABI kernels (K1-K4) against the CPU oracle.

Every check calls librocket_b200.so through its C-ABI (rocket_b200.backend is a ctypes
shim, no torch extension) and compares with oracle/step_oracle.c on the same seeded
inputs.  Bars: integer results bit-exact; optimizer updates bit-exact given the same clip
norm (the kernels spell torch's CPU rounding sequence); norms within 2 ulp of the
double-accumulated oracle (fp32 summation order is the only freedom).
"""
from __future__ import annotations

import numpy as np
import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu

DEV = "cuda"


def _backend():
    from rocket_b200.backend import NativeBackend
    return NativeBackend(torch.device("cuda", torch.cuda.current_device()))


def _bits(a: np.ndarray) -> np.ndarray:
    return a.view(np.int32)


def _ulp_diff(a: np.ndarray, b: np.ndarray) -> int:
    ia, ib = _bits(a).astype(np.int64), _bits(b).astype(np.int64)
    return int(np.abs(ia - ib).max()) if ia.size else 0


def _make_list(sizes, seed, *, groups=1, misalign=(), gscale=1e-2, with_v=True, with_m=True):
    """Tensor list on the GPU + host copies. ``misalign``: indices whose views start 4 bytes
    off a 16-byte boundary (forces the scalar path)."""
    from rocket_b200.backend import TensorSpec
    g = torch.Generator().manual_seed(seed)
    specs, host = [], []
    for i, n in enumerate(sizes):
        off = 1 if i in misalign else 0
        def mk(scale, positive=False):
            t = torch.randn(n + off, generator=g) * scale
            if positive:
                t = t * t
            return t
        hp, hg, hm, hv = mk(0.02), mk(gscale), mk(1e-3), mk(1e-3, True)
        dev = [t.to(DEV)[off:] for t in (hp, hg, hm, hv)]
        host.append([t[off:].numpy().copy() for t in (hp, hg, hm, hv)])
        specs.append(TensorSpec(p=dev[0], g=dev[1], m=dev[2] if with_m else None,
                                v=dev[3] if with_v else None, group=i % groups))
    return specs, host


SIZE_SETS = {
    "awkward": [6, 10, 150, 1, 4095, 4096, 4097, 48000, 3, 0, 8191, 123457],
    "lenet": [150, 6, 2400, 16, 48000, 120, 10080, 84, 840, 10],
    "one_big": [3_000_000],
    "many_small": [7] * 5000,            # > RK_SMEM_PREFIX_CAP tensors: global prefix search
}


@pytest.mark.parametrize("name", list(SIZE_SETS))
@pytest.mark.parametrize("scale", [1.0, 0.125])
def test_l2norm_matches_oracle(name, scale):
    be = _backend()
    sizes = SIZE_SETS[name]
    specs, host = _make_list(sizes, seed=1, misalign=(1, 4) if name == "awkward" else ())
    plan = be.make_plan(specs)
    for _ in range(2):                                   # second launch: ticket was reset
        tot = be.l2norm(plan, scale, want_tensor_norms=True)
    torch.cuda.synchronize()
    want_tot, want_tn = oracle.l2norm_multi([h[1] for h in host], scale)
    got_tn = plan.tensor_norms.cpu().numpy()
    assert _ulp_diff(got_tn, want_tn) <= 2
    assert _ulp_diff(tot.cpu().numpy(), np.array([want_tot], np.float32)) <= 2


def test_l2norm_huge_single_tensor_block_wide_reduce():
    """> 1024 norm chunks in one tensor takes the block-wide level-2 path."""
    be = _backend()
    n = 4096 * 1500 + 17
    g = torch.Generator().manual_seed(5)
    hg = torch.randn(n, generator=g) * 1e-3
    from rocket_b200.backend import TensorSpec
    dg = hg.to(DEV)
    plan = be.make_plan([TensorSpec(p=dg.clone(), g=dg, m=None, v=None, group=0)])
    tot = be.l2norm(plan, 1.0)
    want, _ = oracle.l2norm_multi([hg.numpy()], 1.0)
    assert _ulp_diff(tot.cpu().numpy(), np.array([want], np.float32)) <= 2


def test_l2norm_is_deterministic():
    be = _backend()
    specs, _ = _make_list([100003, 77, 5_000_001], seed=9)
    plan = be.make_plan(specs)
    ref = be.l2norm(plan, 1.0).clone()
    for _ in range(20):
        assert torch.equal(be.l2norm(plan, 1.0), ref)


ADAM_CASES = [
    dict(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=1e-2, decoupled=True),
    dict(lr=6e-4, beta1=0.9, beta2=0.95, eps=1e-8, weight_decay=0.1, decoupled=True),
    dict(lr=3e-4, beta1=0.3, beta2=0.95, eps=1e-6, weight_decay=1e-3, decoupled=False),
    dict(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, decoupled=True),
]


@pytest.mark.parametrize("case", range(len(ADAM_CASES)))
@pytest.mark.parametrize("clip", [None, 0.05])
@pytest.mark.parametrize("name", ["awkward", "lenet"])
def test_adam_step_bit_exact(case, clip, name):
    be = _backend()
    hp = ADAM_CASES[case]
    sizes = SIZE_SETS[name]
    specs, host = _make_list(sizes, seed=case + 11, groups=2,
                             misalign=(2, 5) if name == "awkward" else ())
    plan = be.make_plan(specs)
    scale = 0.5
    for step in (1, 2, 7):
        # fresh gradients every step (the kernel zero-fills them)
        g = torch.Generator().manual_seed(100 + step)
        for s, h in zip(specs, host):
            new = torch.randn(s.g.numel(), generator=g) * 1e-2
            s.g.copy_(new)
            h[1][:] = new.numpy()
        norm = be.l2norm(plan, scale) if clip is not None else None
        groups = [dict(hp, step=step), dict(hp, lr=hp["lr"] * 0.5, step=step)]
        be.adam_step(plan, groups, norm, clip or 0.0, scale, True)
        torch.cuda.synchronize()
        tn = float(norm.item()) if norm is not None else None
        for i, (s, h) in enumerate(zip(specs, host)):
            gh = groups[i % 2]
            oracle.adam_step(h[0], h[1], h[2], h[3], lr=gh["lr"], beta1=gh["beta1"],
                             beta2=gh["beta2"], eps=gh["eps"], weight_decay=gh["weight_decay"],
                             step=step, decoupled=gh["decoupled"], grad_scale=scale,
                             total_norm=tn, max_norm=clip or 0.0, zero_grads=True)
            for k, t in enumerate((s.p, s.g, s.m, s.v)):
                got = t.cpu().numpy()
                assert np.array_equal(_bits(got), _bits(h[k])), \
                    f"tensor {i} ({sizes[i]} elems) stream {'pgmv'[k]} step {step}"


def test_adam_step_zeros_denormals_and_specials_bit_exact():
    """Exact zeros (dead units), denormals, huge values, infinities and NaNs in every stream:
    the division / sqrt wrappers must give the same bits as the IEEE oracle."""
    be = _backend()
    from rocket_b200.backend import TensorSpec
    n = 8192 + 5
    g = torch.Generator().manual_seed(77)
    hp_, hg, hm, hv = (torch.randn(n, generator=g) * s for s in (0.02, 1e-2, 1e-3, 1e-3))
    hv = hv * hv
    specials = torch.tensor([0.0, -0.0, 1e-45, -1e-45, 1e-39, 3e-38, 1e38, -1e38, 1e-30,
                             float("inf"), float("-inf"), float("nan"), 1.0, -1.0])
    k = specials.numel()
    for j, t in enumerate((hp_, hg, hm, hv)):
        for r in range(4):                       # every special against several partners
            idx = torch.arange(k) * 7 + 100 * (r + 1) + 13 * j
            t[idx] = specials.roll(r * (j + 1)) if t is not hv else specials.roll(r).abs()
    hg[::5] = 0.0                                 # 20 % dead gradients
    hm[::5] = 0.0
    hv[::5] = 0.0
    for hp in (ADAM_CASES[0], dict(ADAM_CASES[2], eps=0.0)):
        host = [t.numpy().copy() for t in (hp_, hg, hm, hv)]
        spec = TensorSpec(*[torch.from_numpy(a.copy()).to(DEV) for a in host], group=0)
        plan = be.make_plan([spec])
        be.adam_step(plan, [dict(hp, step=3)], None, 0.0, 1.0, False)
        torch.cuda.synchronize()
        oracle.adam_step(*host, lr=hp["lr"], beta1=hp["beta1"], beta2=hp["beta2"],
                         eps=hp["eps"], weight_decay=hp["weight_decay"], step=3,
                         decoupled=hp["decoupled"])
        for name, t, a in zip("pgmv", (spec.p, spec.g, spec.m, spec.v), host):
            got = t.cpu().numpy()
            same = (_bits(got) == _bits(a)) | (np.isnan(got) & np.isnan(a))
            assert same.all(), (name, np.nonzero(~same)[0][:5], got[~same][:5], a[~same][:5])


SGD_CASES = [
    dict(lr=0.1, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False),
    dict(lr=0.1, momentum=0.9, dampening=0.0, weight_decay=1e-4, nesterov=False),
    dict(lr=0.05, momentum=0.9, dampening=0.0, weight_decay=5e-4, nesterov=True),
    dict(lr=0.05, momentum=0.8, dampening=0.1, weight_decay=0.0, nesterov=False),
]


@pytest.mark.parametrize("case", range(len(SGD_CASES)))
@pytest.mark.parametrize("clip", [None, 0.05])
def test_sgd_step_bit_exact(case, clip):
    be = _backend()
    hp = SGD_CASES[case]
    with_m = hp["momentum"] != 0.0
    specs, host = _make_list(SIZE_SETS["awkward"], seed=case + 31, misalign=(0, 7),
                             with_v=False, with_m=with_m)
    plan = be.make_plan(specs)
    for step in (1, 2, 3):
        g = torch.Generator().manual_seed(200 + step)
        for s, h in zip(specs, host):
            new = torch.randn(s.g.numel(), generator=g) * 1e-2
            s.g.copy_(new)
            h[1][:] = new.numpy()
        norm = be.l2norm(plan, 1.0) if clip is not None else None
        be.sgd_step(plan, [dict(hp, first_step=step == 1)], norm, clip or 0.0, 1.0, False)
        torch.cuda.synchronize()
        tn = float(norm.item()) if norm is not None else None
        for i, (s, h) in enumerate(zip(specs, host)):
            oracle.sgd_step(h[0], h[1], h[2] if with_m else None, first_step=step == 1,
                            total_norm=tn, max_norm=clip or 0.0, zero_grads=False, **hp)
            assert np.array_equal(_bits(s.p.cpu().numpy()), _bits(h[0])), (i, step)
            if with_m:
                assert np.array_equal(_bits(s.m.cpu().numpy()), _bits(h[2])), (i, step)
            assert np.array_equal(_bits(s.g.cpu().numpy()), _bits(h[1]))  # untouched


def test_skip_flags_leave_tensors_untouched():
    be = _backend()
    specs, host = _make_list([1000, 2000, 3000], seed=3, groups=3)
    specs[1].skip = True
    plan = be.make_plan(specs)
    hp = dict(ADAM_CASES[0], step=1)
    be.adam_step(plan, [hp, hp, dict(hp, skip=True)], None, 0.0, 1.0, True)
    torch.cuda.synchronize()
    for i in (1, 2):
        for k, t in enumerate((specs[i].p, specs[i].g, specs[i].m, specs[i].v)):
            assert np.array_equal(_bits(t.cpu().numpy()), _bits(host[i][k]))
    assert not np.array_equal(specs[0].p.cpu().numpy(), host[0][0])
    tot = be.l2norm(plan, 1.0)
    want, _ = oracle.l2norm_multi([host[0][1] * 0, host[2][1]], 1.0)  # g0 zeroed, g1 skipped
    assert _ulp_diff(tot.cpu().numpy(), np.array([want], np.float32)) <= 2


def test_errors_are_loud():
    from rocket_b200 import _native as N
    be = _backend()
    specs, _ = _make_list([128], seed=0)
    plan = be.make_plan(specs)
    with pytest.raises(N.NativeError, match="step must be >= 1"):
        be.adam_step(plan, [dict(ADAM_CASES[0], step=0)], None, 0.0, 1.0, True)
    specs2, _ = _make_list([128], seed=0, with_v=False)
    plan2 = be.make_plan(specs2)
    with pytest.raises(N.NativeError, match="exp_avg_sq"):
        be.adam_step(plan2, [dict(ADAM_CASES[0], step=1)], None, 0.0, 1.0, True)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("shape", [(1, 1), (7, 10), (256, 10), (333, 1000), (64, 50257)])
def test_argmax_count_bit_exact(dtype, shape):
    be = _backend()
    g = torch.Generator().manual_seed(shape[0] * 7 + shape[1])
    logits = torch.randn(shape, generator=g).to(dtype)
    rows, cols = shape
    if rows > 5 and cols > 4:
        logits[5, 3] = float("nan")
        logits[2, 1] = float("nan"); logits[2, 4] = float("nan")
        logits[4, 2] = 100.0; logits[4, 4] = 100.0                      # tie: first wins
        logits[3, :] = 0.0                                              # all equal
    labels = torch.randint(0, cols, (rows,), generator=g)
    want_idx = logits.float().argmax(1)
    labels[::3] = want_idx[::3]                                         # plant some hits
    counters = torch.tensor([5, 11], dtype=torch.int64, device=DEV)
    be.argmax_count(logits.to(DEV), labels.to(DEV), counters)
    want = np.array([5, 11], dtype=np.int64)
    oracle.argmax_count(logits.float().numpy(), labels.numpy(), want)
    assert counters.cpu().numpy().tolist() == want.tolist()
    assert want[0] - 5 == int((want_idx == labels).sum())               # oracle == torch.argmax


def test_argmax_count_strided_rows_and_empty():
    be = _backend()
    big = torch.randn(50, 64, device=DEV)
    view = big[:, :10]                                                 # row stride 64
    labels = view.argmax(1)
    counters = torch.zeros(2, dtype=torch.int64, device=DEV)
    be.argmax_count(view, labels, counters)
    be.argmax_count(view[:0], labels[:0], counters)                    # empty batch: no-op
    assert counters.tolist() == [50, 50]


def test_scalar_accumulate_matches_python_double():
    be = _backend()
    acc = torch.zeros(1, dtype=torch.float64, device=DEV)
    want = 0.0
    g = torch.Generator().manual_seed(0)
    for i in range(50):
        v = torch.rand((), generator=g) * 3
        be.scalar_accumulate(v.to(DEV), 3.0, acc)
        want = oracle.scalar_accumulate(want, float(v), 3.0)
    assert acc.item() == want


@pytest.mark.parametrize("kind", ["adamw", "sgd"])
def test_kernels_write_nothing_outside_their_tensors(kind):
    """Bounds check with canaries (compute-sanitizer is closed on this GPU pool): every tensor
    of an awkward-sized list sits between guard words inside one allocation per stream; after
    the norm, the fused step (launch-argument and device-table variants) and the zero-fill,
    every guard word must be untouched.  Odd gaps also put most tensors off the 16-byte grid,
    so both the vectorised and the scalar paths run."""
    from rocket_b200 import _native as N
    from rocket_b200.backend import TensorSpec
    be = _backend()
    sizes = [6, 10, 150, 1, 4095, 4096, 4097, 48000, 3, 8191, 123457]
    gaps = [5, 8, 3, 4, 7, 12, 1, 16, 9, 2, 6, 11]            # guard words before / between / after
    total = sum(sizes) + sum(gaps)
    sentinel = -123456.75
    torch.manual_seed(3)
    streams = {k: torch.full((total,), sentinel, device=DEV) for k in "pgmv"}
    guard = torch.ones(total, dtype=torch.bool, device=DEV)
    specs, off = [], 0
    for i, n in enumerate(sizes):
        off += gaps[i]
        views = {}
        for k, t in streams.items():
            views[k] = t[off:off + n]
            views[k].copy_(torch.randn(n, device=DEV) * (1e-2 if k != "v" else 1e-3))
        views["v"].abs_()
        guard[off:off + n] = False
        specs.append(TensorSpec(p=views["p"], g=views["g"], m=views["m"],
                                v=views["v"] if kind != "sgd" else None, group=i % 2))
        off += n
    plan = be.make_plan(specs)
    if kind == "sgd":
        groups = [dict(lr=0.05, momentum=0.9, dampening=0.0, weight_decay=1e-4,
                       nesterov=gi == 1, first_step=False, skip=False) for gi in range(2)]
        step = be.sgd_step
    else:
        groups = [dict(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01, step=3,
                       decoupled=True, skip=False) for _ in range(2)]
        step = be.adam_step
    table = torch.zeros(N.RK_STEP_TABLE_BYTES, dtype=torch.uint8, device=DEV)
    host = torch.zeros(N.RK_STEP_TABLE_BYTES, dtype=torch.uint8).pin_memory()
    norm = be.l2norm(plan, 1.0, want_tensor_norms=True)
    step(plan, groups, norm, 1.0, 1.0, False)                  # keep g for the second variant
    be.step_table_write(kind, groups, host)
    table.copy_(host)
    be.step_devtable(kind, plan, table, norm, 1.0, 1.0, True)  # ... which also zero-fills g
    torch.cuda.synchronize()
    touched = "pgm" if kind == "sgd" else "pgmv"
    for k in touched:
        assert bool((streams[k][guard] == sentinel).all()), f"guard words of stream {k} were written"
    assert not streams["g"][~guard].any()                      # zero-fill reached every element
    assert bool(torch.isfinite(streams["p"][~guard]).all())
    if kind == "sgd":
        assert bool((streams["v"][guard] == sentinel).all())


def _draws(n, seed):
    rng = np.random.default_rng(seed)
    pick = lambda xs: xs[int(rng.integers(len(xs)))]        # noqa: E731
    for _ in range(n):
        yield dict(lr=float(10 ** rng.uniform(-6, 0)), wd=pick([0.0, 1e-4, 1e-2, 0.3]),
                   gscale=pick([0.0, 1e-30, 1e-8, 1e-2, 1.0, 1e4]),
                   clip=pick([None, None, 0.05, 1e3]), grad_scale=pick([1.0, 0.5, 0.125]),
                   beta1=pick([0.0, 0.3, 0.5, 0.9, 0.99]), beta2=pick([0.9, 0.95, 0.999]),
                   eps=pick([1e-8, 1e-6, 1e-3]), decoupled=bool(rng.integers(2)),
                   step=int(rng.integers(1, 2000)),
                   momentum=pick([0.0, 0.5, 0.9, 0.99]), dampening=pick([0.0, 0.1]),
                   nesterov=bool(rng.integers(2)), first=bool(rng.integers(2)),
                   seed=int(rng.integers(1 << 30)))


def _refill_grads(specs, host, gscale, seed):
    g = torch.Generator().manual_seed(seed)
    for s, h in zip(specs, host):
        new = torch.randn(s.g.numel(), generator=g) * gscale
        new[::7] = 0.0                                    # dead units
        if gscale:
            new[3::11] = 1e-42                            # denormals
        s.g.copy_(new)
        h[1][:] = new.numpy()


def test_adam_and_sgd_random_hyperparameters_bit_exact():
    """24 seeded random draws of hyper-parameters, gradient magnitudes (all-zero to 1e4), clip
    thresholds and 1/W factors: the fused kernels against the C oracle, every stream, bit for
    bit (the oracle itself is pinned against torch's CPU kernels with the same kind of draws,
    tests/test_oracle.py)."""
    be = _backend()
    sizes = [6, 150, 4097, 70001, 1]
    for k, d in enumerate(_draws(24, seed=2024)):
        # ---- Adam / AdamW
        specs, host = _make_list(sizes, seed=d["seed"] % 1000, misalign=(1,))
        plan = be.make_plan(specs)
        _refill_grads(specs, host, d["gscale"], d["seed"])
        norm = be.l2norm(plan, d["grad_scale"]) if d["clip"] is not None else None
        hp = dict(lr=d["lr"], beta1=d["beta1"], beta2=d["beta2"], eps=d["eps"],
                  weight_decay=d["wd"], decoupled=d["decoupled"], step=d["step"])
        be.adam_step(plan, [hp], norm, d["clip"] or 0.0, d["grad_scale"], True)
        torch.cuda.synchronize()
        tn = float(norm.item()) if norm is not None else None
        for i, (s, h) in enumerate(zip(specs, host)):
            oracle.adam_step(h[0], h[1], h[2], h[3], lr=d["lr"], beta1=d["beta1"],
                             beta2=d["beta2"], eps=d["eps"], weight_decay=d["wd"], step=d["step"],
                             decoupled=d["decoupled"], grad_scale=d["grad_scale"], total_norm=tn,
                             max_norm=d["clip"] or 0.0, zero_grads=True)
            for name, t, a in zip("pgmv", (s.p, s.g, s.m, s.v), h):
                got = t.cpu().numpy()
                same = (_bits(got) == _bits(a)) | (np.isnan(got) & np.isnan(a))
                assert same.all(), ("adam", k, d, i, name, got[~same][:4], a[~same][:4])
        # ---- SGD
        nesterov = d["nesterov"] and d["momentum"] != 0.0 and d["dampening"] == 0.0
        with_m = d["momentum"] != 0.0
        specs, host = _make_list(sizes, seed=d["seed"] % 997, misalign=(2,), with_v=False,
                                 with_m=with_m)
        plan = be.make_plan(specs)
        _refill_grads(specs, host, d["gscale"], d["seed"] + 1)
        norm = be.l2norm(plan, d["grad_scale"]) if d["clip"] is not None else None
        hp = dict(lr=d["lr"], momentum=d["momentum"], dampening=d["dampening"],
                  weight_decay=d["wd"], nesterov=nesterov)
        be.sgd_step(plan, [dict(hp, first_step=d["first"])], norm, d["clip"] or 0.0,
                    d["grad_scale"], True)
        torch.cuda.synchronize()
        tn = float(norm.item()) if norm is not None else None
        for i, (s, h) in enumerate(zip(specs, host)):
            oracle.sgd_step(h[0], h[1], h[2] if with_m else None, first_step=d["first"],
                            grad_scale=d["grad_scale"], total_norm=tn, max_norm=d["clip"] or 0.0,
                            zero_grads=True, **hp)
            streams = [("p", s.p, h[0]), ("g", s.g, h[1])] + ([("m", s.m, h[2])] if with_m else [])
            for name, t, a in streams:
                got = t.cpu().numpy()
                same = (_bits(got) == _bits(a)) | (np.isnan(got) & np.isnan(a))
                assert same.all(), ("sgd", k, d, i, name, got[~same][:4], a[~same][:4])


def test_fifty_steps_back_to_back_without_host_syncs_bit_exact():
    """The norm's two kernels and the step kernel are chained by programmatic dependent launch
    (the next kernel's set-up and first loads overlap the previous one's tail).  Fifty
    norm + step rounds queued with no host synchronisation in between -- fresh gradients
    written on the device before each round -- must end on the oracle's bits: an ordering
    mistake in that chain (a load of p, m or v issued too early, a norm read too early) shows
    up as a mismatch here."""
    be = _backend()
    sizes = [150, 6, 2400, 16, 48000, 120, 10080, 84, 840, 10, 70001]
    specs, host = _make_list(sizes, seed=5)
    plan = be.make_plan(specs)
    base = [s.g.clone() for s in specs]
    base_h = [h[1].copy() for h in host]
    hp = dict(ADAM_CASES[0])
    steps = 50
    norms = []
    for step in range(1, steps + 1):
        f = 1.0 + 0.03125 * (step % 7)                 # exactly representable factors
        for s, b in zip(specs, base):
            torch.mul(b, f, out=s.g)                   # device-side refill, stream ordered
        norm = be.l2norm(plan, 0.5)
        norms.append(norm.clone())                     # device copy, no sync
        be.adam_step(plan, [dict(hp, step=step)], norm, 0.05, 0.5, True)
    torch.cuda.synchronize()
    for step in range(1, steps + 1):
        f = np.float32(1.0 + 0.03125 * (step % 7))
        tn = float(norms[step - 1].item())
        for h, b in zip(host, base_h):
            h[1][:] = b * f
            oracle.adam_step(h[0], h[1], h[2], h[3], lr=hp["lr"], beta1=hp["beta1"],
                             beta2=hp["beta2"], eps=hp["eps"], weight_decay=hp["weight_decay"],
                             step=step, decoupled=hp["decoupled"], grad_scale=0.5,
                             total_norm=tn, max_norm=0.05, zero_grads=True)
        if step in (1, steps):                         # the norm itself against the oracle
            want, _ = oracle.l2norm_multi([b * f for b in base_h], 0.5)
            assert _ulp_diff(np.array([tn], np.float32), np.array([want], np.float32)) <= 2
    for i, (s, h) in enumerate(zip(specs, host)):
        for k, t in enumerate((s.p, s.g, s.m, s.v)):
            assert np.array_equal(_bits(t.cpu().numpy()), _bits(h[k])), (i, "pgmv"[k])
