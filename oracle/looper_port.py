"""looper_port.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Pure-torch, single-process restatement ("port") of what one epoch of the reference's
Looper does with the example graph, written as straight-line code with no capsules, no
accelerate and no engine.  It is the oracle that can TRAVEL: /root/reference is not present
on the GPU box, so GPU parity tests and bench.py's CPU baseline use this, after
tests/test_oracle.py has pinned it (on CPU, in the build container) against golden vectors
produced by the unmodified reference (tests/golden/make_golden.py).

Order of operations per micro-batch, with the reference line each step restates:

    batch = next(loader); move to device                  rocket/core/dataset.py:267-282
    model.train(); forward                                 rocket/core/module.py:130-139
    accumulate(): step counter / forced sync on last batch  module.py:211 (accelerate _do_sync)
    loss = objective(batch)                                rocket/core/loss.py:92
    value += loss.item() / accum                           loss.py:95-98
    on sync: post (step, value); value = 0; step += 1      loss.py:101-116
    (loss / accum).backward()                              loss.py:119 (accelerate.backward)
    on sync: [clip_grad_norm_]; opt.step(); opt.zero_grad()  rocket/core/optimizer.py:129-130
    on sync: post lr                                       optimizer.py:133-147
    on sync: scheduler.step() x world_size                 rocket/core/scheduler.py:113
    eval loop: forward in eval mode, argmax counts         rocket/core/meter.py:63-105,
                                                           examples/mnist.py:26-34

All arithmetic is torch's own (``torch.optim.*(foreach=False)``,
``clip_grad_norm_(foreach=False)``), on whatever device is passed.
"""
from __future__ import annotations

import contextlib
import time

import torch


class OracleOptimizer:
    """The C oracle (oracle/step_oracle.c) wearing a torch.optim interface.

    Parameters may live on any device; each step copies gradients to the host, runs the
    plain-C restatement of torch's CPU arithmetic there (clip coefficient, AdamW / Adam / SGD
    update) and writes the parameters back.  Slow by construction -- it is the checker for
    the CUDA engine in GPU parity tests, where both arms share the same GPU forward/backward
    and must then agree to the last bit (no clipping) or to the clip-norm's last ulp.
    """

    def __init__(self, inner: torch.optim.Optimizer, max_norm=None):
        import numpy as np
        import oracle
        self._np, self._oracle = np, oracle
        self.inner, self.max_norm = inner, max_norm
        self.param_groups = inner.param_groups
        self.kind = {torch.optim.AdamW: "adamw", torch.optim.Adam: "adam",
                     torch.optim.SGD: "sgd"}[type(inner)]
        self.state = {}
        self.steps = 0

    def step(self):
        np, oracle = self._np, self._oracle
        params = [p for g in self.param_groups for p in g["params"] if p.grad is not None]
        grads = [p.grad.detach().cpu().numpy().reshape(-1).copy() for p in params]
        total = None
        if self.max_norm is not None:
            total, _ = oracle.l2norm_multi(grads, 1.0)
        self.steps += 1
        i = 0
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is None:
                    continue
                g = grads[i]
                i += 1
                st = self.state.setdefault(p, {})
                host = p.detach().cpu().numpy().reshape(-1).copy()
                if self.kind == "sgd":
                    first = "buf" not in st
                    if group["momentum"] != 0 and first:
                        st["buf"] = np.zeros_like(host)
                    oracle.sgd_step(host, g, st.get("buf"), lr=group["lr"],
                                    momentum=group["momentum"], dampening=group["dampening"],
                                    weight_decay=group["weight_decay"],
                                    nesterov=group["nesterov"], first_step=first,
                                    total_norm=total, max_norm=self.max_norm or 0.0)
                else:
                    if "m" not in st:
                        st["m"], st["v"], st["step"] = np.zeros_like(host), np.zeros_like(host), 0
                    st["step"] += 1
                    oracle.adam_step(host, g, st["m"], st["v"], lr=group["lr"],
                                     beta1=group["betas"][0], beta2=group["betas"][1],
                                     eps=group["eps"], weight_decay=group["weight_decay"],
                                     step=st["step"], decoupled=self.kind == "adamw",
                                     total_norm=total, max_norm=self.max_norm or 0.0)
                with torch.no_grad():
                    p.copy_(torch.from_numpy(host).view(p.shape))

    def zero_grad(self, set_to_none=True):
        self.inner.zero_grad(set_to_none=set_to_none)


def train_epoch(net, objective, opt, sched, batches, *, accum: int, max_norm, device,
                state: dict, world: int = 1, sync_loss: bool = True,
                autocast_dtype=None, scaler=None) -> None:
    """One pass over ``batches`` (a sized iterable of CPU batches). Appends to ``state``:
    loss_steps/loss_values, lr_steps/lr_values; keeps the cross-epoch counters there."""
    net.train()
    n = len(batches)
    phase = state.setdefault("phase", 0)
    value = state.setdefault("value", 0.0)
    for i, batch in enumerate(batches):
        batch = [t.to(device) if isinstance(t, torch.Tensor) else t for t in batch]
        last = (i == n - 1)
        if last:
            phase, sync = 0, True
        else:
            phase += 1
            sync = (phase % accum) == 0
        # autocast is entered and left around every forward + objective, as Module.runner does
        # (module.py:175-219): leaving the outermost context drops the cached low-precision
        # weight copies, which would otherwise be stale after the optimizer step
        amp = torch.autocast(device_type=torch.device(device).type, dtype=autocast_dtype) \
            if autocast_dtype is not None else contextlib.nullcontext()
        with torch.enable_grad():
            with amp:
                out = net(batch)
                loss = objective(out)
            if sync_loss:
                value += loss.item() / accum
            if sync:
                state["loss_steps"].append(state["loss_step"])
                state["loss_values"].append(value)
                state["loss_step"] += 1
                value = 0.0
            # accelerate.backward: scaler.scale(loss / accum).backward() under fp16
            scaled = loss / accum
            (scaler.scale(scaled) if scaler is not None else scaled).backward()
        if sync:
            skipped = False
            if scaler is not None:
                # accelerate: clip_grad_norm_ unscales first; optimizer.step goes through
                # scaler.step (skipped on overflow) + scaler.update; an overflowing step also
                # skips the scheduler (AcceleratedScheduler checks step_was_skipped)
                before = scaler.get_scale()
                scaler.unscale_(opt)
            if max_norm is not None and not isinstance(opt, OracleOptimizer):
                torch.nn.utils.clip_grad_norm_(net.parameters(), max_norm, foreach=False)
            if scaler is not None:
                scaler.step(opt)
                scaler.update()
                skipped = scaler.get_scale() < before
            else:
                opt.step()
            opt.zero_grad()
            state["lr_steps"].append(state["opt_step"])
            state["lr_values"].append(opt.param_groups[0]["lr"])
            state["opt_step"] += 1
            if sched is not None and not skipped:
                for _ in range(world):
                    sched.step()
        elif sched is not None:
            sched._step_count += 1
    state["phase"], state["value"] = phase, value


@torch.no_grad()
def eval_epoch(net, batches, *, device, autocast_dtype=None) -> tuple[int, int]:
    net.eval()
    positive = total = 0
    for batch in batches:
        batch = [t.to(device) if isinstance(t, torch.Tensor) else t for t in batch]
        amp = torch.autocast(device_type=torch.device(device).type, dtype=autocast_dtype) \
            if autocast_dtype is not None else contextlib.nullcontext()
        with amp:
            _, gt, logits = net(batch)
        pr = logits.argmax(1)
        total += pr.numel()
        positive += int(torch.sum(gt == pr))
    return positive, total


def new_state() -> dict:
    return dict(loss_steps=[], loss_values=[], lr_steps=[], lr_values=[], loss_step=0,
                opt_step=0, accuracy_history=[])


def wrap_autocast_forward(net, device, autocast_dtype) -> None:
    """What accelerate's prepare_model does to a module under mixed precision (reached from
    rocket/core/module.py:106): forward runs under autocast and its low-precision outputs are
    converted to fp32 -- so the objective sees fp32 logits.  (On CUDA, cross_entropy of fp16
    logits and of the same logits converted to fp32 differ in the last digits.)"""
    if autocast_dtype is None:
        return
    inner = net.forward

    def forward(*a, **k):
        with torch.autocast(device_type=torch.device(device).type, dtype=autocast_dtype):
            out = inner(*a, **k)
        return [t.float() if isinstance(t, torch.Tensor) and t.dtype == autocast_dtype else t
                for t in out]
    net.forward = forward


def run_mnist_port(*, device="cpu", seed=0, n_train=64 * 24, n_val=64 * 5 + 13, batch_size=64,
                   optimizer="adamw", max_norm=None, accum=2, epochs=2, sched_step=5, lr=1e-3,
                   net=None, workloads=None, optimizer_impl="torch", mixed_precision=None,
                   input_scale=1.0) -> dict:
    """The graph of tests/workloads.build_mnist_graph, as straight-line code."""
    if workloads is None:
        import workloads  # tests/workloads.py (on sys.path in tests and bench)
    torch.manual_seed(seed)
    net = net if net is not None else workloads.LeNet()
    if optimizer == "adamw":
        opt = torch.optim.AdamW(net.parameters(), lr=lr, foreach=False)
    elif optimizer == "adam":
        opt = torch.optim.Adam(net.parameters(), lr=lr, weight_decay=1e-3, foreach=False)
    elif optimizer == "sgd":
        opt = torch.optim.SGD(net.parameters(), lr=lr * 50, momentum=0.9, weight_decay=1e-4,
                              nesterov=True, foreach=False)
    else:
        raise ValueError(optimizer)
    sched = torch.optim.lr_scheduler.StepLR(opt, sched_step, gamma=0.5)
    if optimizer_impl == "oracle":
        opt._opt_called = True                      # keep LRScheduler's order check quiet
        opt = OracleOptimizer(opt, max_norm)
    train = workloads.synthetic_classification(n_train, (1, 28, 28), 10, seed + 1, input_scale)
    val = workloads.synthetic_classification(n_val, (1, 28, 28), 10, seed + 2, input_scale)
    amp_dtype = {None: None, "bf16": torch.bfloat16, "fp16": torch.float16}[mixed_precision]
    scaler = torch.amp.GradScaler(torch.device(device).type) if mixed_precision == "fp16" \
        else None
    net.to(device)
    wrap_autocast_forward(net, device, amp_dtype)
    objective = workloads.CrossEntropy()
    tl = torch.utils.data.DataLoader(train, batch_size=batch_size, shuffle=False)
    vl = torch.utils.data.DataLoader(val, batch_size=batch_size, shuffle=False)
    state = new_state()
    for epoch in range(epochs):
        train_epoch(net, objective, opt, sched, tl, accum=accum, max_norm=max_norm,
                    device=device, state=state, autocast_dtype=amp_dtype, scaler=scaler)
        pos, tot = eval_epoch(net, vl, device=device, autocast_dtype=amp_dtype)
        state["accuracy_history"].append((epoch, pos, tot))
        # eval iterations advance accelerate's accumulation phase too (module.py:138,211);
        # the last eval batch resets it
        state["phase"] = 0
    state["params"] = {k: p.detach().cpu().clone() for k, p in net.named_parameters()}
    state["net"], state["opt"], state["scaler"] = net, opt, scaler
    return state


def time_steps(net, objective, opt, batch, *, steps: int, warmup: int, max_norm, device):
    """Seconds per step of the reference arithmetic on ``device`` for one fixed batch
    (bench.py's cpu_baseline / --impl reference)."""
    net.train()

    def one():
        out = net([t.to(device) for t in batch])
        loss = objective(out)
        loss.item()
        loss.backward()
        if max_norm is not None:
            torch.nn.utils.clip_grad_norm_(net.parameters(), max_norm, foreach=False)
        opt.step()
        opt.zero_grad()

    for _ in range(warmup):
        one()
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    return (time.perf_counter() - t0) / steps


def run_generic_port(net, objective, opt, batches, *, accum: int, max_norm, device,
                     optimizer_impl: str = "torch", autocast_dtype=None) -> dict:
    """One train epoch of an arbitrary (net, objective, optimizer) over ``batches`` with the
    reference's step order; optimizer arithmetic = torch's own or the C oracle."""
    if optimizer_impl == "oracle":
        opt._opt_called = True
        opt = OracleOptimizer(opt, max_norm)
    net.to(device)
    wrap_autocast_forward(net, device, autocast_dtype)
    state = new_state()
    train_epoch(net, objective, opt, None, batches, accum=accum, max_norm=max_norm,
                device=device, state=state, autocast_dtype=autocast_dtype)
    state["params"] = {k: p.detach().cpu().clone() for k, p in net.named_parameters()}
    return state
