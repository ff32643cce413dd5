"""Dataset capsule: one batch per LAUNCH, delivered on the device.

API mirror of the reference's rocket/core/dataset.py:22-361.  What differs is below the
API: the loader the engine hands back from ``prepare`` (rocket_b200/engine.py,
EngineDataLoader) shards batches across ranks and, on a CUDA engine, runs each batch through
the pinned-ring stager (async H2D on a side stream, one batch ahead) instead of the blocking
pageable ``Tensor.to(device)`` the reference issues here (reference dataset.py:279-282 ->
utils/torch.py:50-51).  ``torch_move`` is still called: it moves whatever the stager did not
(modules, registered custom types) and is a no-op for tensors already on the device.
"""
from __future__ import annotations

from typing import Iterable

import torch
import torch.utils.data

from rocket_b200.core.capsule import Attributes, Capsule
from rocket_b200.utils.torch import torch_collate, torch_move


class Dataset(Capsule):
    def __init__(self, dataset: Iterable, statefull: bool = True, priority: int = 1000,
                 **kwargs):
        super().__init__(statefull=statefull, priority=priority)
        self._dataset = dataset
        self._dataloader = None           # the prepared loader
        self._active_dataloader = None    # differs from it only when resuming mid-epoch
        self._iterator = None
        self._kwargs = kwargs
        self._kwargs.setdefault("collate_fn", torch_collate)
        self._batch_idx = 0
        self._total = 0

    def setup(self, attrs: Attributes | None = None) -> None:
        Capsule.setup(self, attrs=attrs)
        # Two capsules over the same dataset object share one prepared loader
        # (reference dataset.py:158-171); a third registration is an error.
        found = False
        for loader in self._accelerator._dataloaders:
            if loader.dataset is not self._dataset:
                continue
            if found:
                raise RuntimeError(
                    f"{type(self).__name__}: same dataset has been registered twice.")
            found = True
            self._dataloader = loader
        if not found:
            loader = torch.utils.data.DataLoader(self._dataset, **self._kwargs)
            self._dataloader = self._accelerator.prepare(loader, device_placement=[False])

    def set(self, attrs: Attributes | None = None) -> None:
        Capsule.set(self, attrs=attrs)
        if torch.is_grad_enabled() and self._batch_idx > 0:
            # resumed from a checkpoint taken mid-epoch (reference dataset.py:205-208)
            self._active_dataloader = self._accelerator.skip_first_batches(
                self._dataloader, self._batch_idx)
        else:
            self._active_dataloader = self._dataloader
        self._total = len(self._active_dataloader)
        self._iterator = iter(self._active_dataloader)

    def reset(self, attrs: Attributes | None = None) -> None:
        Capsule.reset(self, attrs=attrs)
        self._batch_idx = 0
        self._total = 0
        self._iterator = None

    def launch(self, attrs: Attributes | None = None) -> None:
        Capsule.launch(self, attrs=attrs)
        if attrs is None or attrs.batch is not None:
            return
        data = next(self._iterator, None)
        if data is None:
            attrs.batch = None
            if attrs.looper is not None:       # exhausted: vote to leave the loop
                attrs.looper.terminate = True
            return
        attrs.batch = torch_move(data, self._accelerator.device)
        if attrs.looper is not None:
            attrs.looper.terminate = False
        self._batch_idx += 1

    def destroy(self, attrs: Attributes | None = None) -> None:
        Capsule.destroy(self, attrs=attrs)
        loader, self._dataloader, self._active_dataloader = self._dataloader, None, None
        # NOTE: the reference clears its handle before searching and therefore never
        # removes the loader; here the loader is really released.
        loaders = self._accelerator._dataloaders
        for i, candidate in enumerate(loaders):
            if candidate is loader:
                loaders.pop(i)
                break

    def state_dict(self) -> dict:
        return dict(batch_idx=self._batch_idx)

    def load_state_dict(self, state: dict) -> None:
        self._batch_idx = state["batch_idx"]
