"""Device backend of the engine: torch tensors in, C-ABI kernel launches out.

``NativeBackend`` is the product path and the only backend this package ships: every method
is one call into librocket_b200.so on the current CUDA stream.  It raises at construction if
CUDA or the library is unavailable -- there is no CPU or eager-PyTorch fallback.  (The test
suite substitutes a host-only double that implements the same five methods with the
oracle, to exercise the engine's host logic on machines without a GPU; see
tests/host_backend.py.)
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass

import torch

from rocket_b200 import _native as N


@dataclass
class TensorSpec:
    """One parameter tensor and its state, as torch tensors (flat fp32 views)."""
    p: torch.Tensor
    g: torch.Tensor
    m: torch.Tensor | None
    v: torch.Tensor | None
    group: int
    skip: bool = False


class Plan:
    """Tensor-list metadata for the multi-tensor kernels: host blob, device copy, scratch."""

    def __init__(self, specs: list[TensorSpec], device: torch.device,
                 scratch: torch.Tensor | None = None):
        lib = N.load()
        self.specs = specs
        self.n = len(specs)
        descs = (N.TensorDesc * self.n)()
        for i, s in enumerate(specs):
            for t in (s.p, s.g, s.m, s.v):
                if t is not None and (t.dtype != torch.float32 or not t.is_contiguous()):
                    raise ValueError("plan tensors must be contiguous fp32")
            d = descs[i]
            d.p, d.g = s.p.data_ptr(), s.g.data_ptr()
            d.m = s.m.data_ptr() if s.m is not None else None
            d.v = s.v.data_ptr() if s.v is not None else None
            d.numel, d.group = s.p.numel(), s.group
            d.flags = N.RK_TENSOR_SKIP if s.skip else 0
        numel = (ctypes.c_int64 * self.n)(*[s.p.numel() for s in specs])
        plan_bytes, scratch_bytes = ctypes.c_size_t(), ctypes.c_size_t()
        N.check(lib.rk_plan_query(numel, self.n, ctypes.byref(plan_bytes),
                                  ctypes.byref(scratch_bytes)), "rk_plan_query")
        self.host = torch.empty(plan_bytes.value, dtype=torch.uint8).pin_memory() \
            if device.type == "cuda" else torch.empty(plan_bytes.value, dtype=torch.uint8)
        N.check(lib.rk_plan_build(descs, self.n, self.host.data_ptr(), plan_bytes.value),
                "rk_plan_build")
        self.dev = self.host.to(device, non_blocking=False)
        if scratch is None:
            scratch = torch.zeros(scratch_bytes.value, dtype=torch.uint8, device=device)
        elif scratch.numel() < scratch_bytes.value:
            raise ValueError(f"scratch of {scratch.numel()} bytes, plan needs {scratch_bytes.value}")
        self.scratch = scratch
        # header fields the host side needs (struct rk_plan_header, csrc/rk_common.cuh):
        # int32[3] = chunk_norm, int32[5] = n_chunks_norm, uint32[8] = off_prefix_norm
        words = self.host.numpy().view("<i4")
        self.chunk_norm, self.n_chunks_norm = int(words[3]), int(words[5])
        off = int(words[8]) // 4
        self.prefix_norm = words[off:off + self.n + 1].copy()
        self.total_norm = torch.zeros(1, dtype=torch.float32, device=device)
        self.tensor_norms = torch.zeros(self.n, dtype=torch.float32, device=device)
        self.numel = sum(s.p.numel() for s in specs)


class NativeBackend:
    name = "sm_100a"

    def __init__(self, device: torch.device):
        if device.type != "cuda" or not torch.cuda.is_available():
            raise RuntimeError(
                "rocket_b200 runs on CUDA (B200, sm_100a) only: no CUDA device is available "
                "and there is no CPU fallback.")
        self.lib = N.load()
        self.device = device

    @staticmethod
    def _stream() -> int:
        return torch.cuda.current_stream().cuda_stream

    def make_plan(self, specs: list[TensorSpec], scratch: torch.Tensor | None = None) -> Plan:
        return Plan(specs, self.device, scratch)

    @staticmethod
    def scratch_bytes(numels: list[int]) -> int:
        lib = N.load()
        arr = (ctypes.c_int64 * len(numels))(*numels)
        pb, sb = ctypes.c_size_t(), ctypes.c_size_t()
        N.check(lib.rk_plan_query(arr, len(numels), ctypes.byref(pb), ctypes.byref(sb)),
                "rk_plan_query")
        return sb.value

    # K1, level 2 only: partials are already in plan.scratch (written by the fused all-reduce)
    def l2norm_finalize(self, plan: Plan, want_tensor_norms: bool = False):
        N.check(self.lib.rk_l2norm_finalize(
            plan.host.data_ptr(), plan.dev.data_ptr(), plan.scratch.data_ptr(),
            plan.total_norm.data_ptr(),
            plan.tensor_norms.data_ptr() if want_tensor_norms else None, self._stream()),
            "rk_l2norm_finalize")
        return plan.total_norm

    # K6
    def comm_create(self, rank: int, world: int, arena_ptrs: list[int], scratch_offset: int,
                    multicast_ptr: int = 0):
        handle = ctypes.c_void_p()
        arr = (ctypes.c_void_p * world)(*arena_ptrs)
        N.check(self.lib.rk_comm_create(ctypes.byref(handle), rank, world, arr,
                                        multicast_ptr or None, scratch_offset), "rk_comm_create")
        return handle

    def comm_allreduce_norm(self, comm, plan: Plan, chunk_lo: int, chunk_hi: int, bucket: int,
                            max_ctas: int, stream: int, use_multicast: bool = True) -> None:
        N.check(self.lib.rk_comm_allreduce_norm(comm, plan.host.data_ptr(), plan.dev.data_ptr(),
                                                chunk_lo, chunk_hi, bucket, max_ctas,
                                                int(use_multicast), stream),
                "rk_comm_allreduce_norm")

    def comm_reduce_scatter_norm(self, comm, plan: Plan, chunk_lo: int, chunk_hi: int,
                                 bucket: int, max_ctas: int, stream: int,
                                 use_multicast: bool = True) -> None:
        N.check(self.lib.rk_comm_reduce_scatter_norm(
            comm, plan.host.data_ptr(), plan.dev.data_ptr(), chunk_lo, chunk_hi, bucket,
            max_ctas, int(use_multicast), stream), "rk_comm_reduce_scatter_norm")

    # K7
    def comm_sharded_step(self, comm, plan: Plan, kind: str, table_dev_ptr: int,
                          total_norm=None, max_norm: float = 0.0, max_ctas: int = 0,
                          use_multicast: bool = True, shard_m=None, shard_v=None, *,
                          grad_scale: float = 1.0) -> None:
        N.check(self.lib.rk_comm_sharded_step(
            comm, plan.host.data_ptr(), plan.dev.data_ptr(), 1 if kind == "sgd" else 0,
            table_dev_ptr, shard_m.data_ptr() if shard_m is not None else None,
            shard_v.data_ptr() if shard_v is not None else None,
            total_norm.data_ptr() if total_norm is not None else None,
            float(max_norm), float(grad_scale), max_ctas, int(use_multicast), self._stream()),
            "rk_comm_sharded_step")

    def comm_destroy(self, comm) -> None:
        self.lib.rk_comm_destroy(comm)

    # K1
    def l2norm(self, plan: Plan, grad_scale: float = 1.0, want_tensor_norms: bool = False):
        N.check(self.lib.rk_l2norm_multi(
            plan.host.data_ptr(), plan.dev.data_ptr(), grad_scale, plan.scratch.data_ptr(),
            plan.total_norm.data_ptr(),
            plan.tensor_norms.data_ptr() if want_tensor_norms else None, self._stream()),
            "rk_l2norm_multi")
        return plan.total_norm

    # K2
    @staticmethod
    def _adam_groups(groups: list[dict]):
        arr = (N.AdamGroup * len(groups))()
        for a, g in zip(arr, groups):
            a.lr, a.beta1, a.beta2 = g["lr"], g["beta1"], g["beta2"]
            a.eps, a.weight_decay, a.step = g["eps"], g["weight_decay"], g["step"]
            a.decoupled, a.skip = int(g["decoupled"]), int(g.get("skip", False))
        return arr

    @staticmethod
    def _sgd_groups(groups: list[dict]):
        arr = (N.SgdGroup * len(groups))()
        for a, g in zip(arr, groups):
            a.lr, a.momentum, a.dampening = g["lr"], g["momentum"], g["dampening"]
            a.weight_decay, a.nesterov = g["weight_decay"], int(g["nesterov"])
            a.first_step, a.skip = int(g["first_step"]), int(g.get("skip", False))
        return arr

    def step_table_write(self, kind: str, groups: list[dict], out: torch.Tensor) -> None:
        """Per-group constants of one step into RK_STEP_TABLE_BYTES of host memory (graph
        replays: the table is then copied to the device in stream order)."""
        if kind == "sgd":
            N.check(self.lib.rk_sgd_table_write(self._sgd_groups(groups), len(groups),
                                                out.data_ptr()), "rk_sgd_table_write")
        else:
            N.check(self.lib.rk_adam_table_write(self._adam_groups(groups), len(groups),
                                                 out.data_ptr()), "rk_adam_table_write")

    def step_devtable(self, kind: str, plan: Plan, table_dev: torch.Tensor, total_norm,
                      max_norm: float, grad_scale: float, zero_grads: bool) -> None:
        fn = self.lib.rk_sgd_step_multi_devtable if kind == "sgd" \
            else self.lib.rk_adamw_step_multi_devtable
        N.check(fn(plan.host.data_ptr(), plan.dev.data_ptr(), table_dev.data_ptr(),
                   total_norm.data_ptr() if total_norm is not None else None, max_norm,
                   grad_scale, int(zero_grads), self._stream()), "rk_*_step_multi_devtable")

    def adam_step(self, plan: Plan, groups: list[dict], total_norm, max_norm: float,
                  grad_scale: float, zero_grads: bool) -> None:
        arr = self._adam_groups(groups)
        N.check(self.lib.rk_adamw_step_multi(
            plan.host.data_ptr(), plan.dev.data_ptr(), arr, len(groups),
            total_norm.data_ptr() if total_norm is not None else None, max_norm, grad_scale,
            int(zero_grads), self._stream()), "rk_adamw_step_multi")

    # K3
    def sgd_step(self, plan: Plan, groups: list[dict], total_norm, max_norm: float,
                 grad_scale: float, zero_grads: bool) -> None:
        arr = self._sgd_groups(groups)
        N.check(self.lib.rk_sgd_step_multi(
            plan.host.data_ptr(), plan.dev.data_ptr(), arr, len(groups),
            total_norm.data_ptr() if total_norm is not None else None, max_norm, grad_scale,
            int(zero_grads), self._stream()), "rk_sgd_step_multi")

    # K4
    _DTYPES = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2}

    def argmax_count(self, logits: torch.Tensor, labels: torch.Tensor,
                     counters: torch.Tensor) -> None:
        if logits.dim() != 2:
            logits = logits.reshape(-1, logits.shape[-1])
        if logits.stride(1) != 1:
            logits = logits.contiguous()
        labels = labels.reshape(-1)
        if labels.dtype != torch.int64 or not labels.is_contiguous():
            labels = labels.to(torch.int64).contiguous()
        if labels.numel() != logits.shape[0]:
            raise ValueError("argmax_count: one label per row expected")
        code = self._DTYPES.get(logits.dtype)
        if code is None:
            raise TypeError(f"argmax_count: unsupported logits dtype {logits.dtype}")
        N.check(self.lib.rk_meter_argmax_count(
            logits.data_ptr(), code, labels.data_ptr(), logits.shape[0], logits.shape[1],
            logits.stride(0), counters.data_ptr(), self._stream()), "rk_meter_argmax_count")

    def scalar_accumulate(self, value: torch.Tensor, divisor: float, acc: torch.Tensor) -> None:
        if value.dtype != torch.float32:
            value = value.float()
        N.check(self.lib.rk_scalar_accumulate(value.data_ptr(), float(divisor), acc.data_ptr(),
                                              self._stream()), "rk_scalar_accumulate")

    def launch_count(self) -> int:
        return N.launch_count()
