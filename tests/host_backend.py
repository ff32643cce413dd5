"""TEST DOUBLE, not product code: a host-only stand-in for rocket_b200.backend.NativeBackend.

It implements the backend's methods with the CPU oracle (oracle/step_oracle.c) on CPU
tensors so that the engine's HOST logic -- capsule ordering, accumulation phase, flat
layout, gradient adoption, bucketing over gloo, deferred scalars, checkpoint format -- can
be exercised on machines without a GPU.  It lives under tests/ because only tests may route
through the oracle; the package itself refuses to run without CUDA.
"""
from __future__ import annotations

import numpy as np
import torch

import oracle


class HostPlan:
    def __init__(self, specs):
        self.specs = specs
        self.n = len(specs)
        self.total_norm = torch.zeros(1, dtype=torch.float32)
        self.tensor_norms = torch.zeros(self.n, dtype=torch.float32)
        self.numel = sum(s.p.numel() for s in specs)


class HostBackend:
    name = "host-oracle (tests only)"

    def __init__(self):
        self.launches = 0

    def make_plan(self, specs):
        return HostPlan(specs)

    def l2norm(self, plan, grad_scale=1.0, want_tensor_norms=False):
        grads = [s.g.numpy() if not s.skip else np.zeros(0, np.float32) for s in plan.specs]
        tot, tn = oracle.l2norm_multi(grads, grad_scale)
        plan.total_norm[0] = float(tot)
        plan.tensor_norms.copy_(torch.from_numpy(tn))
        self.launches += 1
        return plan.total_norm

    def adam_step(self, plan, groups, total_norm, max_norm, grad_scale, zero_grads):
        tn = float(total_norm[0]) if total_norm is not None else None
        for s in plan.specs:
            g = groups[s.group]
            if s.skip or g.get("skip"):
                continue
            oracle.adam_step(s.p.numpy(), s.g.numpy(), s.m.numpy(), s.v.numpy(), lr=g["lr"],
                             beta1=g["beta1"], beta2=g["beta2"], eps=g["eps"],
                             weight_decay=g["weight_decay"], step=g["step"],
                             decoupled=g["decoupled"], grad_scale=grad_scale, total_norm=tn,
                             max_norm=max_norm, zero_grads=zero_grads)
        self.launches += 1

    def sgd_step(self, plan, groups, total_norm, max_norm, grad_scale, zero_grads):
        tn = float(total_norm[0]) if total_norm is not None else None
        for s in plan.specs:
            g = groups[s.group]
            if s.skip or g.get("skip"):
                continue
            oracle.sgd_step(s.p.numpy(), s.g.numpy(), s.m.numpy() if s.m is not None else None,
                            lr=g["lr"], momentum=g["momentum"], dampening=g["dampening"],
                            weight_decay=g["weight_decay"], nesterov=g["nesterov"],
                            first_step=g["first_step"], grad_scale=grad_scale, total_norm=tn,
                            max_norm=max_norm, zero_grads=zero_grads)
        self.launches += 1

    def argmax_count(self, logits, labels, counters):
        c = counters.numpy()
        oracle.argmax_count(logits.detach().float().reshape(-1, logits.shape[-1]).numpy(),
                            labels.reshape(-1).to(torch.int64).numpy(), c)
        self.launches += 1

    def scalar_accumulate(self, value, divisor, acc):
        acc[0] = oracle.scalar_accumulate(float(acc[0]), float(value.detach().float()), divisor)
        self.launches += 1

    def launch_count(self):
        return self.launches
