"""Pin the oracle (CPU only).

1. oracle/step_oracle.c vs torch's own CPU kernels, i.e. the arithmetic the reference's
   Optimizer capsule reaches (reference rocket/core/optimizer.py:129-130 ->
   torch.optim.*(foreach=False)): SGD bit-exact; Adam moments bit-exact and parameters
   within 1 ulp (torch's CPU sqrt goes through MKL VML, which is not correctly rounded: ~0.7 %
   of elements differ by one ulp from IEEE sqrtf; the oracle and the CUDA kernels use IEEE).
2. oracle/looper_port.py vs golden vectors produced by the UNMODIFIED reference
   (tests/golden/make_golden.py): loss series, lr series, eval counts, final parameters.
"""
from __future__ import annotations

import numpy as np
import pytest
import torch

import oracle
import workloads
from oracle import looper_port

GOLDEN = __import__("pathlib").Path(__file__).parent / "golden"


def _bits(a):
    return a.view(np.int32)


def _ulp(a, b):
    return int(np.abs(_bits(a).astype(np.int64) - _bits(b).astype(np.int64)).max())


@pytest.mark.parametrize("kind,kw", [
    ("adamw", dict(lr=1e-3, weight_decay=1e-2)),
    ("adamw", dict(lr=6e-4, weight_decay=0.1, betas=(0.9, 0.95))),
    ("adamw", dict(lr=1e-3, weight_decay=0.0)),
    ("adam", dict(lr=3e-4, weight_decay=0.1, betas=(0.3, 0.95), eps=1e-6)),
])
@pytest.mark.parametrize("clip", [None, 0.5])
def test_c_oracle_adam_vs_torch_cpu(kind, kw, clip):
    torch.manual_seed(0)
    n = 10007
    p = torch.nn.Parameter(torch.randn(n) * 0.02)
    cls = torch.optim.AdamW if kind == "adamw" else torch.optim.Adam
    opt = cls([p], foreach=False, **kw)
    m, v = np.zeros(n, np.float32), np.zeros(n, np.float32)
    for step in range(1, 9):
        g = torch.randn(n) * 1e-2
        p.grad = g.clone()
        pn = p.detach().numpy().copy()           # resync: compare one step at a time
        p_before = pn.copy()
        tn = None
        if clip is not None:
            tn = torch.nn.utils.clip_grad_norm_([p], clip, foreach=False)
            want_tn, _ = oracle.l2norm_multi([g.numpy()])
            assert abs(float(want_tn) - float(tn)) <= 1e-6 * float(tn)
        opt.step()
        grp = opt.param_groups[0]
        gn = g.numpy().copy()
        oracle.adam_step(pn, gn, m, v, lr=grp["lr"], beta1=grp["betas"][0],
                         beta2=grp["betas"][1], eps=grp["eps"],
                         weight_decay=grp["weight_decay"], step=step,
                         decoupled=kind == "adamw",
                         total_norm=None if tn is None else float(tn), max_norm=clip or 0.0)
        st = opt.state[p]
        assert np.array_equal(_bits(m), _bits(st["exp_avg"].numpy()))
        assert np.array_equal(_bits(v), _bits(st["exp_avg_sq"].numpy()))
        # p: one ulp of sqrt(v) moves the update by 1.2e-7 relative
        got, want = pn, p.detach().numpy()
        assert np.all(np.abs(got - want) <= 4.8e-7 * np.abs(want - p_before)
                      + 2 * np.spacing(np.abs(want)))


@pytest.mark.parametrize("kw", [
    dict(lr=0.1), dict(lr=0.1, momentum=0.9, weight_decay=1e-4),
    dict(lr=0.05, momentum=0.9, nesterov=True, weight_decay=5e-4),
    dict(lr=0.05, momentum=0.8, dampening=0.1),
])
def test_c_oracle_sgd_vs_torch_cpu_bit_exact(kw):
    torch.manual_seed(1)
    n = 10007
    p = torch.nn.Parameter(torch.randn(n) * 0.02)
    opt = torch.optim.SGD([p], foreach=False, **kw)
    pn = p.detach().numpy().copy()
    buf = np.zeros(n, np.float32) if kw.get("momentum") else None
    for step in range(1, 7):
        g = torch.randn(n) * 1e-2
        p.grad = g.clone()
        tn = torch.nn.utils.clip_grad_norm_([p], 0.7, foreach=False)
        opt.step()
        oracle.sgd_step(pn, g.numpy().copy(), buf, lr=kw["lr"], momentum=kw.get("momentum", 0.0),
                        dampening=kw.get("dampening", 0.0),
                        weight_decay=kw.get("weight_decay", 0.0),
                        nesterov=kw.get("nesterov", False), first_step=step == 1,
                        total_norm=float(tn), max_norm=0.7)
        assert np.array_equal(_bits(pn), _bits(p.detach().numpy())), step


def test_c_oracle_argmax_count_vs_torch():
    g = torch.Generator().manual_seed(0)
    lg = torch.randn(300, 10, generator=g)
    lg[5, 3] = float("nan")
    lg[6, 1] = lg[6, 7] = float("nan")
    lg[7, 2] = lg[7, 4] = 100.0
    lg[8, :] = 1.0
    lb = torch.randint(0, 10, (300,), generator=g)
    lb[5], lb[6], lb[7], lb[8] = 3, 1, 2, 0
    c = np.array([2, 3], np.int64)
    oracle.argmax_count(lg.numpy(), lb.numpy(), c)
    assert c.tolist() == [2 + int((lg.argmax(1) == lb).sum()), 303]


def test_c_oracle_two_level_norm_vs_torch():
    g = torch.Generator().manual_seed(3)
    grads = [torch.randn(n, generator=g) * 1e-3 for n in (6, 10, 150, 48000, 1, 123457)]
    params = [torch.nn.Parameter(torch.zeros_like(x)) for x in grads]
    for p, x in zip(params, grads):
        p.grad = x.clone()
    want = torch.nn.utils.clip_grad_norm_(params, 1e9, foreach=False)
    got, tn = oracle.l2norm_multi([x.numpy() for x in grads], 1.0)
    assert abs(float(got) - float(want)) <= 1e-6 * float(want)   # torch sums in fp32
    for t, x in zip(tn, grads):
        assert abs(float(t) - float(torch.linalg.vector_norm(x))) <= 1e-6 * float(t)


CASES = {
    "lenet_adamw_accum2": dict(optimizer="adamw", max_norm=None, accum=2, epochs=2,
                               n_train=64 * 25),
    "lenet_adamw_clip": dict(optimizer="adamw", max_norm=0.5, accum=2, epochs=2,
                             n_train=64 * 25),
    "lenet_sgd_clip": dict(optimizer="sgd", max_norm=1.0, accum=1, epochs=1, n_train=64 * 20),
    "lenet_adam_l2": dict(optimizer="adam", max_norm=None, accum=3, epochs=1, n_train=64 * 20),
    # fp16 + GradScaler, 4-5 overflowing (skipped) steps at the start
    "lenet_fp16_adamw_clip": dict(optimizer="adamw", max_norm=0.5, accum=2, epochs=1,
                                  n_train=64 * 20, sched_step=3, input_scale=128.0,
                                  mixed_precision="fp16"),
    "lenet_fp16_sgd": dict(optimizer="sgd", max_norm=None, accum=1, epochs=1, n_train=64 * 20,
                           sched_step=3, lr=2e-5, input_scale=128.0, mixed_precision="fp16"),
}


@pytest.mark.parametrize("name", list(CASES))
def test_looper_port_matches_reference_golden(name):
    """The travelling oracle reproduces what the unmodified reference produced."""
    torch.set_num_threads(1)
    gold = np.load(GOLDEN / f"{name}.npz")
    got = looper_port.run_mnist_port(device="cpu", workloads=workloads, **CASES[name])
    assert got["loss_steps"] == gold["loss_steps"].tolist()
    assert got["lr_steps"] == gold["lr_steps"].tolist()
    assert got["lr_values"] == gold["lr_values"].tolist()
    np.testing.assert_allclose(got["loss_values"], gold["loss_values"], rtol=2e-6)
    assert [list(x) for x in got["accuracy_history"]] == gold["accuracy_history"].tolist()
    for k, p in got["params"].items():
        want = gold[f"param/{k}"]
        assert np.abs(p.numpy() - want).max() <= 2e-5 * np.abs(want).max(), k


# ---- randomised pinning of the C oracle against torch's CPU kernels (hypothesis) ---------------
from hypothesis import given, settings, strategies as st  # noqa: E402

_lr = st.floats(1e-6, 1.0)
_wd = st.sampled_from([0.0, 1e-4, 1e-2, 0.3])
_gscale = st.sampled_from([0.0, 1e-30, 1e-8, 1e-2, 1.0, 1e4])     # incl. all-zero and huge grads


def _grad(n, scale, seed):
    g = torch.Generator().manual_seed(seed)
    t = torch.randn(n, generator=g) * scale
    t[::7] = 0.0                                  # exact zeros, as real ReLU networks produce
    t[3::11] = 1e-42 * (1 if scale else 0)        # denormals
    return t


@settings(max_examples=60, deadline=None, derandomize=True)
@given(n=st.integers(1, 3000), lr=_lr, wd=_wd, gscale=_gscale, seed=st.integers(0, 2 ** 16),
       momentum=st.sampled_from([0.0, 0.5, 0.9, 0.99]), dampening=st.sampled_from([0.0, 0.1]),
       nesterov=st.booleans(), steps=st.integers(1, 4))
def test_c_oracle_sgd_random_hyperparameters_bit_exact(n, lr, wd, gscale, seed, momentum,
                                                       dampening, nesterov, steps):
    if nesterov and (momentum == 0.0 or dampening != 0.0):
        nesterov = False                           # torch rejects the combination
    torch.manual_seed(seed)
    p = torch.nn.Parameter(torch.randn(n) * 0.05)
    opt = torch.optim.SGD([p], lr=lr, momentum=momentum, dampening=dampening, weight_decay=wd,
                          nesterov=nesterov, foreach=False)
    pn = p.detach().numpy().copy()
    buf = np.zeros(n, np.float32) if momentum else None
    for step in range(1, steps + 1):
        g = _grad(n, gscale, seed + step)
        p.grad = g.clone()
        opt.step()
        oracle.sgd_step(pn, g.numpy().copy(), buf, lr=lr, momentum=momentum, dampening=dampening,
                        weight_decay=wd, nesterov=nesterov, first_step=step == 1)
        assert np.array_equal(_bits(pn), _bits(p.detach().numpy())), step
        if momentum:
            assert np.array_equal(_bits(buf), _bits(opt.state[p]["momentum_buffer"].numpy()))


def _same_bits_up_to_zero_sign(a, b) -> bool:
    """Bit equality, except that +0.0 and -0.0 count as equal.  torch's own lerp is not
    consistent there: its vectorised body computes fmadd(w - 1, end - start, end) and its
    scalar tail end - (end - start) * (1 - w); for w == 1 (beta1 == 0) and a gradient of -0.0
    the first gives -0.0 and the second +0.0, so the sign of a zero moment depends on the
    element's position in the tensor.  The oracle (and the kernels) follow the vector body."""
    nz = (a != 0) | (b != 0)
    return bool(np.array_equal(a, b) and np.array_equal(_bits(a)[nz], _bits(b)[nz]))


@settings(max_examples=60, deadline=None, derandomize=True)
@given(n=st.integers(1, 3000), lr=_lr, wd=_wd, gscale=_gscale, seed=st.integers(0, 2 ** 16),
       beta1=st.sampled_from([0.0, 0.3, 0.5, 0.9]), beta2=st.sampled_from([0.9, 0.95, 0.999]),
       eps=st.sampled_from([1e-8, 1e-6, 1e-3]), decoupled=st.booleans(),
       steps=st.integers(1, 4))
def test_c_oracle_adam_random_hyperparameters(n, lr, wd, gscale, seed, beta1, beta2, eps,
                                              decoupled, steps):
    """Moments bit-exact; parameters within the one sqrt ulp that separates torch's MKL sqrt
    from the correctly rounded one (beta1 < 0.5 takes torch's other lerp branch)."""
    torch.manual_seed(seed)
    p = torch.nn.Parameter(torch.randn(n) * 0.05)
    cls = torch.optim.AdamW if decoupled else torch.optim.Adam
    opt = cls([p], lr=lr, betas=(beta1, beta2), eps=eps, weight_decay=wd, foreach=False)
    m, v = np.zeros(n, np.float32), np.zeros(n, np.float32)
    for step in range(1, steps + 1):
        g = _grad(n, gscale, seed + step)
        p.grad = g.clone()
        pn = p.detach().numpy().copy()            # resync every step: compare single updates
        before = pn.copy()
        opt.step()
        oracle.adam_step(pn, g.numpy().copy(), m, v, lr=lr, beta1=beta1, beta2=beta2, eps=eps,
                         weight_decay=wd, step=step, decoupled=decoupled)
        state = opt.state[p]
        assert _same_bits_up_to_zero_sign(m, state["exp_avg"].numpy()), step
        assert np.array_equal(_bits(v), _bits(state["exp_avg_sq"].numpy())), step
        want = p.detach().numpy()
        assert np.all(np.abs(pn - want) <= 4.8e-7 * np.abs(want - before)
                      + 2 * np.spacing(np.abs(want))), step
        m[:] = state["exp_avg"].numpy()           # carry torch's zero signs into the next step
