"""TEST-ONLY: accelerate.utils pieces used at rocket/core/launcher.py:20,150,161,189."""
import torch
import torch.distributed as dist


class ProjectConfiguration:
    def __init__(self, project_dir=None, logging_dir=None, **_):
        self.project_dir = project_dir
        self.logging_dir = logging_dir or project_dir


def _initialized():
    return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1


def broadcast_object_list(object_list, from_process=0):
    if _initialized():
        dist.broadcast_object_list(object_list, src=from_process)
    return object_list


def recursively_apply(fn, data):
    if isinstance(data, (list, tuple)):
        return type(data)(recursively_apply(fn, d) for d in data)
    if isinstance(data, dict):
        return type(data)({k: recursively_apply(fn, v) for k, v in data.items()})
    if isinstance(data, torch.Tensor):
        return fn(data)
    return data


def _gather_one(t):
    if t.ndim == 0:
        t = t.clone()[None]
    t = t.contiguous()
    out = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return torch.cat(out, dim=0)


def gather(data):
    if not _initialized():
        return data
    return recursively_apply(_gather_one, data)


def convert_to_fp32(data):
    return recursively_apply(
        lambda t: t.float() if t.is_floating_point() and t.dtype in (torch.float16, torch.bfloat16) else t,
        data)
