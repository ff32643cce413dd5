"""Batch collation and device placement helpers (mirror of the reference's
rocket/utils/torch.py:17-95).

``torch_collate`` stacks tensors and leaves every other leaf type as a plain list.
``torch_move`` walks a nested batch and moves tensors / modules; other leaves pass through.
Types can be added with ``register_move_hook``.  In the engine the hot-path H2D copy does
not go through here: batches arrive on the device from the pinned-ring stager
(rocket_b200/staging.py) and ``torch_move`` then finds nothing left to move.
"""
from __future__ import annotations

import collections
from typing import Callable, Dict, Type

import torch
from torch.utils.data._utils.collate import collate, collate_tensor_fn

from rocket_b200.utils.collections import apply_to_collection, is_collection

MapType = Dict[Type, Callable]
BUILTIN_TYPES = [int, float, str, bool, complex, bytes]


def _keep(batch, *, collate_fn_map=None):
    return batch


# every type that has no entry collates to "leave the list of samples alone"
COLLATE_MAPPINGS = collections.defaultdict(lambda: _keep)
COLLATE_MAPPINGS[torch.Tensor] = collate_tensor_fn


def torch_collate(batch):
    return collate(batch, collate_fn_map=COLLATE_MAPPINGS)


def _stay(batch, device, *, move_fn_map=None):
    return batch


def _to_device(batch, device, *, move_fn_map=None):
    return batch.to(device)


MOVE_MAPPINGS = collections.defaultdict(lambda: _stay)
MOVE_MAPPINGS[torch.Tensor] = _to_device
MOVE_MAPPINGS[torch.nn.Module] = _to_device


def move(batch, device, *, move_fn_map: MapType | None = None, **kwargs):
    kind = type(batch)
    if move_fn_map is not None:
        if kind in move_fn_map or kind in BUILTIN_TYPES:
            return move_fn_map[kind](batch, device, move_fn_map=move_fn_map)
        for registered in move_fn_map:
            if isinstance(batch, registered):
                return move_fn_map[registered](batch, device, move_fn_map=move_fn_map)
    if is_collection(batch):
        return apply_to_collection(batch, move, device=device, move_fn_map=move_fn_map)
    return batch


def torch_move(batch, device):
    return move(batch, device, move_fn_map=MOVE_MAPPINGS)


def register_move_hook(dtype: type, hook: Callable) -> None:
    if not isinstance(type(dtype), type):
        raise RuntimeError("The provided dtype is not a type.")
    MOVE_MAPPINGS[dtype] = hook


def register_default_move_hook(dtype: type) -> None:
    register_move_hook(dtype=dtype, hook=_to_device)
