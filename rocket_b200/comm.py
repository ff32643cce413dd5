"""Gradient exchange for data parallelism (kernel row K6).

The reference gets this implicitly: ``accelerator.prepare(module)`` wraps the model in torch
DDP (reference rocket/core/module.py:106), whose C++ reducer copies every gradient into
25 MiB buckets (dividing by the world size), all-reduces each bucket as backward fills it,
and copies the result back.  Here gradients already live in the optimizer's flat buffer
(rocket_b200/flat.py), so a bucket is just a contiguous slice of it:

* no copy-in, no copy-out -- NCCL reduces the gradient memory in place;
* plain SUM on the wire (lets NVSwitch reduce in-network); the 1/world_size factor is
  folded into the norm and step kernels' gradient read (``grad_scale``), which is exact for
  power-of-two world sizes and therefore identical to DDP's divide-then-sum;
* buckets are launched from ``post_accumulate_grad`` hooks in reverse parameter order, on
  NCCL's own stream, so the exchange overlaps the rest of backward; on accumulation
  micro-steps the hooks do nothing (DDP ``no_sync`` semantics);
* after the exchange every rank holds bit-identical gradients, so the deterministic norm
  kernel yields bit-identical clip coefficients with no further collective.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


class GradSync:
    def __init__(self, opt, engine, bucket_cap_mb: float = 32.0, first_bucket_mb: float = 1.0):
        self.opt = opt
        self.engine = engine
        self.buckets: list[tuple[int, int]] = []       # element ranges of flat_g
        self.bucket_of: list[int] = [0] * len(opt._params)
        self._members: list[list[int]] = []
        cap = int(first_bucket_mb * (1 << 20)) // 4
        big = int(bucket_cap_mb * (1 << 20)) // 4
        hi = opt._total
        members: list[int] = []
        size = 0
        # walk parameters backwards: gradients become ready roughly in reverse order
        for i in range(len(opt._params) - 1, -1, -1):
            members.append(i)
            size += opt._params[i].numel()
            if size >= cap or i == 0:
                lo = opt._offsets[i]
                b = len(self.buckets)
                self.buckets.append((lo, hi))
                self._members.append(members)
                for m in members:
                    self.bucket_of[m] = b
                hi, members, size, cap = lo, [], 0, big
        self._pending = [len(m) for m in self._members]
        self._launched = [False] * len(self.buckets)
        self._works: list = []
        self._handles = [p.register_post_accumulate_grad_hook(self._make_hook(i))
                         for i, p in enumerate(opt._params)]
        self.bytes_reduced = 0

    def _make_hook(self, i: int):
        def hook(param):
            if not self.engine.sync_gradients or not self.opt._adopted[i]:
                return
            if param.grad.data_ptr() != self.opt._grad_ptr[i]:
                return                      # adopted later, reduced in finish()
            b = self.bucket_of[i]
            self._pending[b] -= 1
            if self._pending[b] == 0:
                self._launch(b)
        return hook

    def _launch(self, b: int) -> None:
        lo, hi = self.buckets[b]
        view = self.opt.flat_g[lo:hi]
        self._works.append(dist.all_reduce(view, op=dist.ReduceOp.SUM, async_op=True))
        self._launched[b] = True
        self.bytes_reduced += (hi - lo) * 4

    def finish(self) -> None:
        """Launch whatever backward did not, then order the current stream after all of it."""
        # buckets must be issued in the same order on every rank: launch leftovers by index
        for b in range(len(self.buckets)):
            if not self._launched[b]:
                self._launch(b)
        for w in self._works:
            w.wait()
        self._works.clear()
        self._pending = [len(m) for m in self._members]
        self._launched = [False] * len(self.buckets)

    def remove(self) -> None:
        for h in self._handles:
            h.remove()
        self._handles = []


def broadcast_module_state(module: torch.nn.Module, src: int = 0) -> None:
    """Start every rank from rank 0's parameters and buffers (what DDP does at wrap time)."""
    with torch.no_grad():
        for t in list(module.parameters()) + list(module.buffers()):
            dist.broadcast(t.data, src=src)


class BufferSync:
    """DDP's ``broadcast_buffers=True``: rank 0's module buffers (BatchNorm statistics)
    overwrite everyone's before a training forward.  Buffers are re-homed into one flat
    tensor per dtype so this is one collective per dtype, without packing copies."""

    def __init__(self, module: torch.nn.Module):
        self.flats: list[torch.Tensor] = []
        by_dtype: dict = {}
        for b in module.buffers():
            by_dtype.setdefault(b.dtype, []).append(b)
        with torch.no_grad():
            for dtype, bufs in by_dtype.items():
                total = sum(b.numel() for b in bufs)
                if total == 0:
                    continue
                flat = torch.empty(total, dtype=dtype, device=bufs[0].device)
                off = 0
                for b in bufs:
                    n = b.numel()
                    view = flat[off:off + n].view(b.shape)
                    view.copy_(b)
                    b.data = view
                    off += n
                self.flats.append(flat)
        self._handle = module.register_forward_pre_hook(self._hook)

    def _hook(self, module, args):
        if torch.is_grad_enabled() and module.training:
            for flat in self.flats:
                dist.broadcast(flat, src=0)
        return None

    def remove(self) -> None:
        self._handle.remove()
