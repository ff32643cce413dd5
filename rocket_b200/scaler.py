"""Dynamic loss scaling for ``mixed_precision="fp16"``.

The reference accepts ``fp16`` in ``Launcher`` (rocket/core/launcher.py:100,187) and leaves the
work to accelerate, which drives ``torch.amp.GradScaler``: ``scaler.scale(loss).backward()``,
then per optimizer ``scaler.step(opt); scaler.update()`` [accelerate optimizer.py, from memory].
GradScaler makes two passes over the gradients per step (``_amp_foreach_non_finite_check_and_
unscale_`` and the optimizer) and reads ``found_inf`` back to the host.

Here the unscale is the ``grad_scale`` factor the norm (K1) and step (K2/K3) kernels already
apply when they read a gradient, and the overflow check is the norm itself: a non-finite
element makes the sum of squares non-finite.  What is left on the host is GradScaler's
bookkeeping, restated from ``$TORCH/amp/grad_scaler.py`` and the ``_amp_update_scale_`` kernel
(``$TORCH/../aten/src/ATen/native/cuda/AmpKernels.cu``): fp32 scale, int32 growth tracker.

One host read of the 4-byte norm per sync step decides whether the step is skipped, as
``GradScaler.step`` does with ``found_inf.item()``.
"""
from __future__ import annotations

import numpy as np
import torch


class LossScaler:
    def __init__(self, device: torch.device, init_scale: float = 2.0 ** 16,
                 growth_factor: float = 2.0, backoff_factor: float = 0.5,
                 growth_interval: int = 2000) -> None:
        self.growth_factor = float(growth_factor)
        self.backoff_factor = float(backoff_factor)
        self.growth_interval = int(growth_interval)
        self._scale = np.float32(init_scale)
        self._growth_tracker = 0
        # what backward multiplies the loss with; refreshed by fill_ (a launch, no sync)
        self._dev_scale = torch.full((), float(self._scale), dtype=torch.float32, device=device)

    # ---- GradScaler surface -------------------------------------------------------------------
    def get_scale(self) -> float:
        return float(self._scale)

    def scale(self, loss: torch.Tensor) -> torch.Tensor:
        return loss * self._dev_scale          # grad_scaler.py scale(): outputs * self._scale

    @property
    def inv_scale(self) -> float:
        """``scale.double().reciprocal().float()`` (grad_scaler.py ``_unscale_grads_`` caller)."""
        return float(np.float32(1.0 / np.float64(self._scale)))

    def update(self, found_inf: bool) -> None:
        """``torch._amp_update_scale_``: back off on overflow, grow after ``growth_interval``
        clean steps unless that would overflow the scale itself."""
        if found_inf:
            self._scale = np.float32(np.float64(self._scale) * self.backoff_factor)
            self._growth_tracker = 0
        else:
            successful = self._growth_tracker + 1
            if successful == self.growth_interval:
                grown = np.float32(np.float64(self._scale) * self.growth_factor)
                if np.isfinite(grown):
                    self._scale = grown
                self._growth_tracker = 0
            else:
                self._growth_tracker = successful
        self._dev_scale.fill_(float(self._scale))

    # ---- torch.amp.GradScaler's checkpoint format ---------------------------------------------
    def state_dict(self) -> dict:
        return {"scale": float(self._scale), "growth_factor": self.growth_factor,
                "backoff_factor": self.backoff_factor, "growth_interval": self.growth_interval,
                "_growth_tracker": int(self._growth_tracker)}

    def load_state_dict(self, state: dict) -> None:
        self._scale = np.float32(state["scale"])
        self.growth_factor = float(state["growth_factor"])
        self.backoff_factor = float(state["backoff_factor"])
        self.growth_interval = int(state["growth_interval"])
        self._growth_tracker = int(state["_growth_tracker"])
        self._dev_scale.fill_(float(self._scale))
