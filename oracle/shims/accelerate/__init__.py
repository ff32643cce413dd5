"""TEST-ONLY stand-in for HuggingFace `accelerate` (>=0.34.2 per the reference's
requirements.txt:2; not installed on this image and not installable offline).

PURPOSE: let the UNMODIFIED reference under /root/reference import and run in this
container so that golden vectors can be generated from the reference's own Looper step
(tests/golden/make_golden.py).  It is NOT part of the product and nothing in
rocket_b200/ imports it.

It restates, from the accelerate 0.34 sources as remembered (they are not available here:
"[accelerate, from memory]" in SURVEY.md Appendix A), exactly the surface the reference
touches:

  Accelerator(device_placement, mixed_precision, gradient_accumulation_steps,
              project_config, project_dir)                     rocket/core/launcher.py:185-193
  .prepare(obj, device_placement=[...])   dataset.py:178, module.py:106, optimizer.py:109,
                                          scheduler.py:92
  .accumulate(model) / .autocast() / .backward(loss)           module.py:210-211, loss.py:119
  .gather / .gather_for_metrics                                loss.py:95, meter.py:93
  .sync_gradients .gradient_accumulation_steps .device .is_main_process
  .is_local_main_process .project_dir ._models ._optimizers ._schedulers ._dataloaders
  ._custom_objects .register_for_checkpointing .skip_first_batches .save_state
  .load_state .get_tracker .log_with .init_trackers .end_training

All arithmetic is torch's own (torch.optim with foreach=False is chosen by the caller).
"""
from __future__ import annotations

import contextlib
import os
import pickle

import torch
import torch.distributed as dist

from . import logging, tracking, utils  # noqa: F401
from .utils import ProjectConfiguration  # noqa: F401


def _world():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


class GradientState:
    """accelerate.state.GradientState: shared accumulation bookkeeping."""

    def __init__(self, num_steps: int):
        self.num_steps = num_steps
        self.sync_gradients = True
        self.adjust_scheduler = True
        self.sync_with_dataloader = True
        self.active_dataloader = None
        self.dataloader_references = [None]

    @property
    def in_dataloader(self):
        return self.active_dataloader is not None

    @property
    def end_of_dataloader(self):
        if not self.in_dataloader:
            return False
        return self.active_dataloader.end_of_dataloader

    @property
    def remainder(self):
        if not self.in_dataloader:
            return -1
        return self.active_dataloader.remainder

    def _add_dataloader(self, dl):
        self.active_dataloader = dl
        self.dataloader_references.append(dl)

    def _remove_dataloader(self, dl):
        self.dataloader_references.remove(dl)
        self.active_dataloader = self.dataloader_references[-1]


class BatchSamplerShard(torch.utils.data.Sampler):
    """accelerate.data_loader.BatchSamplerShard (split_batches=False, even_batches=True),
    restated: rank r takes batches r, r+W, ...; an incomplete last round is completed, at the
    index level, by cycling through the samples of the first round."""

    def __init__(self, batch_sampler, world, rank):
        self.batch_sampler, self.world, self.rank = batch_sampler, world, rank
        self.batch_size = batch_sampler.batch_size
        self.drop_last = batch_sampler.drop_last

    def __len__(self):
        n = len(self.batch_sampler)
        return n // self.world if self.drop_last else -(-n // self.world)

    def __iter__(self):
        W, rank, bs = self.world, self.rank, self.batch_size
        initial, mine, idx, batch = [], [], -1, []
        for idx, batch in enumerate(self.batch_sampler):
            batch = list(batch)
            if not self.drop_last and idx < W:
                initial += batch
            if idx % W == rank:
                mine = batch
            if idx % W == W - 1 and len(batch) == bs:
                yield mine
                mine = []
        if self.drop_last or not initial:
            return
        if len(mine) == bs:
            yield mine
        while len(initial) < W * bs:
            initial += initial
        if len(batch) == bs:
            batch = []
            idx += 1
        cycle = 0
        while idx % W != 0 or len(batch) > 0:
            end = cycle + bs - len(batch)
            batch = batch + initial[cycle:end]
            if idx % W == rank:
                yield batch
            cycle, batch = end, []
            idx += 1


class DataLoaderShard:
    """accelerate.data_loader.DataLoaderShard, restated: iterates the (rank-sharded) loader
    one batch ahead so `end_of_dataloader` is True while the last batch is out.  No device
    placement (the reference passes device_placement=[False])."""

    def __init__(self, base, gradient_state, skip_batches: int = 0):
        self.base = base
        self.gradient_state = gradient_state
        self.skip_batches = skip_batches
        self.end_of_dataloader = False
        self.remainder = -1
        self.dataset = base.dataset
        self.batch_size = base.batch_size
        rank, world = _world()
        if world > 1:
            self.loader = torch.utils.data.DataLoader(
                base.dataset, batch_sampler=BatchSamplerShard(base.batch_sampler, world, rank),
                collate_fn=base.collate_fn, num_workers=base.num_workers)
        else:
            self.loader = base

    def __len__(self):
        return len(self.loader) - self.skip_batches

    def __iter__(self):
        self.end_of_dataloader = False
        self.gradient_state._add_dataloader(self)
        rank, world = _world()
        total = len(self.base.dataset) if hasattr(self.base.dataset, "__len__") else None
        if total is not None and self.base.batch_size:
            self.remainder = total % (self.base.batch_size * world)
        it = iter(self.loader)
        try:
            current = next(it)
        except StopIteration:
            self.gradient_state._remove_dataloader(self)
            return
        index = 0
        while True:
            try:
                nxt = next(it)
                if index >= self.skip_batches:
                    yield current
                index += 1
                current = nxt
            except StopIteration:
                self.end_of_dataloader = True
                if index >= self.skip_batches:
                    yield current
                break
        self.gradient_state._remove_dataloader(self)


class AcceleratedOptimizer:
    """accelerate.optimizer.AcceleratedOptimizer without the XLA paths."""

    def __init__(self, optimizer, gradient_state, scaler=None):
        self.optimizer = optimizer
        self.gradient_state = gradient_state
        self.scaler = scaler
        self._is_overflow = False

    @property
    def param_groups(self):
        return self.optimizer.param_groups

    @property
    def state(self):
        return self.optimizer.state

    @property
    def step_was_skipped(self):
        return self._is_overflow

    def zero_grad(self, set_to_none=None):
        if self.gradient_state.sync_gradients:
            if set_to_none is None:
                self.optimizer.zero_grad()
            else:
                self.optimizer.zero_grad(set_to_none=set_to_none)

    def step(self, closure=None):
        if not self.gradient_state.sync_gradients:
            return
        if self.scaler is None:
            self.optimizer.step(closure)
            return
        # fp16: scaler.step skips optimizer.step on overflow; accelerate learns that by
        # patching the step method and looking whether it was called
        called = []
        original = self.optimizer.step

        def patched(*a, **k):
            called.append(True)
            return original(*a, **k)
        self.optimizer.step = patched
        try:
            self.scaler.step(self.optimizer, closure)
            self.scaler.update()
        finally:
            del self.optimizer.step
        self._is_overflow = not called

    def state_dict(self):
        return self.optimizer.state_dict()

    def load_state_dict(self, sd):
        self.optimizer.load_state_dict(sd)


class AcceleratedScheduler:
    """accelerate.scheduler.AcceleratedScheduler (step_with_optimizer=True,
    split_batches=False)."""

    def __init__(self, scheduler, optimizers, gradient_state):
        self.scheduler = scheduler
        self.optimizers = optimizers
        self.gradient_state = gradient_state

    def step(self, *args, **kwargs):
        if not self.gradient_state.sync_gradients:
            if self.gradient_state.adjust_scheduler:
                self.scheduler._step_count += 1
            return
        for opt in self.optimizers:
            if opt.step_was_skipped:
                return
        _, world = _world()
        for _ in range(world):
            if hasattr(self.scheduler, "total_steps"):
                if self.scheduler._step_count <= self.scheduler.total_steps:
                    self.scheduler.step(*args, **kwargs)
            else:
                self.scheduler.step(*args, **kwargs)

    def get_last_lr(self):
        return self.scheduler.get_last_lr()

    def state_dict(self):
        return self.scheduler.state_dict()

    def load_state_dict(self, sd):
        self.scheduler.load_state_dict(sd)


class PartialState:
    def destroy_process_group(self):
        if dist.is_available() and dist.is_initialized():
            dist.destroy_process_group()


class Accelerator:
    def __init__(self, device_placement=True, mixed_precision=None,
                 gradient_accumulation_steps=1, project_config=None, project_dir=None,
                 cpu=None, **_):
        if "RANK" in os.environ and "WORLD_SIZE" in os.environ and not dist.is_initialized() \
                and int(os.environ["WORLD_SIZE"]) > 1:
            backend = "nccl" if torch.cuda.is_available() and not cpu else "gloo"
            dist.init_process_group(backend)
        rank, world = _world()
        self.process_index = rank
        self.num_processes = world
        self.local_process_index = int(os.environ.get("LOCAL_RANK", rank))
        force_cpu = cpu if cpu is not None else os.environ.get("ACCELERATE_USE_CPU") == "1"
        if torch.cuda.is_available() and not force_cpu:
            self.device = torch.device("cuda", self.local_process_index)
            torch.cuda.set_device(self.device)
        else:
            self.device = torch.device("cpu")
        self.device_placement = device_placement
        self.mixed_precision = mixed_precision if mixed_precision not in (None, "no") else "no"
        self.gradient_state = GradientState(gradient_accumulation_steps)
        # real accelerate refuses fp16 without an accelerator device; the shim lets torch's own
        # GradScaler run on the CPU so that the unmodified reference can produce fp16 goldens
        self.scaler = torch.amp.GradScaler(self.device.type) \
            if self.mixed_precision == "fp16" else None
        self.project_dir = project_dir if project_dir is not None else \
            getattr(project_config, "project_dir", None)
        self.step = 0
        self._models, self._optimizers, self._schedulers = [], [], []
        self._dataloaders, self._custom_objects = [], []
        self.log_with = []
        self.trackers = []

    # -- plain attributes -------------------------------------------------------------
    @property
    def sync_gradients(self):
        return self.gradient_state.sync_gradients

    @property
    def gradient_accumulation_steps(self):
        return self.gradient_state.num_steps

    @property
    def is_main_process(self):
        return self.process_index == 0

    @property
    def is_local_main_process(self):
        return self.local_process_index == 0

    # -- prepare ------------------------------------------------------------------------
    def prepare(self, *objs, device_placement=None):
        out = tuple(self._prepare_one(o) for o in objs)
        return out[0] if len(out) == 1 else out

    def _prepare_one(self, obj):
        if isinstance(obj, torch.utils.data.DataLoader):
            dl = DataLoaderShard(obj, self.gradient_state)
            self._dataloaders.append(dl)
            return dl
        if isinstance(obj, torch.nn.Module):
            model = obj
            if self.num_processes > 1:
                kw = {}
                if self.device.type == "cuda":
                    kw = dict(device_ids=[self.device.index], output_device=self.device.index)
                model = torch.nn.parallel.DistributedDataParallel(model, **kw)
            if self.mixed_precision in ("bf16", "fp16"):
                dtype = torch.bfloat16 if self.mixed_precision == "bf16" else torch.float16
                inner = model.forward
                dev_type = self.device.type

                def forward(*a, **k):
                    with torch.autocast(device_type=dev_type, dtype=dtype):
                        out = inner(*a, **k)
                    return utils.convert_to_fp32(out)
                model.forward = forward
            self._models.append(model)
            return model
        if isinstance(obj, torch.optim.Optimizer):
            opt = AcceleratedOptimizer(obj, self.gradient_state, self.scaler)
            self._optimizers.append(opt)
            return opt
        if isinstance(obj, torch.optim.lr_scheduler.LRScheduler):
            sch = AcceleratedScheduler(obj, self._optimizers, self.gradient_state)
            self._schedulers.append(sch)
            return sch
        return obj

    def skip_first_batches(self, dataloader, num_batches=0):
        return DataLoaderShard(dataloader.base, self.gradient_state,
                               skip_batches=dataloader.skip_batches + num_batches)

    # -- step contexts ----------------------------------------------------------------------
    def _do_sync(self):
        gs = self.gradient_state
        if gs.sync_with_dataloader and gs.end_of_dataloader:
            self.step = 0
            gs.sync_gradients = True
        else:
            self.step += 1
            gs.sync_gradients = (self.step % gs.num_steps) == 0

    @contextlib.contextmanager
    def accumulate(self, *models):
        self._do_sync()
        with contextlib.ExitStack() as stack:
            for m in models:
                if not self.sync_gradients and hasattr(m, "no_sync") and self.num_processes > 1:
                    stack.enter_context(m.no_sync())
            yield

    @contextlib.contextmanager
    def autocast(self, autocast_handler=None):
        if self.mixed_precision in ("bf16", "fp16"):
            dtype = torch.bfloat16 if self.mixed_precision == "bf16" else torch.float16
            with torch.autocast(device_type=self.device.type, dtype=dtype):
                yield
        else:
            yield

    def backward(self, loss, **kwargs):
        loss = loss / self.gradient_accumulation_steps
        if self.scaler is not None:
            loss = self.scaler.scale(loss)
        loss.backward(**kwargs)

    def clip_grad_norm_(self, parameters, max_norm, norm_type=2):
        if self.scaler is not None:          # accelerate: unscale_gradients() first
            for opt in self._optimizers:
                self.scaler.unscale_(opt.optimizer)
        return torch.nn.utils.clip_grad_norm_(parameters, max_norm, norm_type=norm_type)

    # -- collectives ----------------------------------------------------------------------------
    def gather(self, tensor):
        return utils.gather(tensor)

    def gather_for_metrics(self, input_data):
        data = utils.gather(input_data)
        gs = self.gradient_state
        if gs.end_of_dataloader and gs.remainder > 0 and self.num_processes > 1:
            return utils.recursively_apply(lambda t: t[: gs.remainder], data)
        return data

    # -- checkpointing ---------------------------------------------------------------------------
    def register_for_checkpointing(self, *objects):
        for o in objects:
            if not (hasattr(o, "state_dict") and hasattr(o, "load_state_dict")):
                raise ValueError("object must have state_dict and load_state_dict")
        self._custom_objects.extend(objects)

    def save_state(self, output_dir=None, **_):
        os.makedirs(output_dir, exist_ok=True)
        for i, m in enumerate(self._models):
            torch.save(m.state_dict(), os.path.join(output_dir, f"model_{i}.bin"))
        for i, o in enumerate(self._optimizers):
            torch.save(o.state_dict(), os.path.join(output_dir, f"optimizer_{i}.bin"))
        for i, s in enumerate(self._schedulers):
            torch.save(s.state_dict(), os.path.join(output_dir, f"scheduler_{i}.bin"))
        for i, c in enumerate(self._custom_objects):
            with open(os.path.join(output_dir, f"custom_checkpoint_{i}.pkl"), "wb") as f:
                pickle.dump(c.state_dict(), f)
        if self.scaler is not None:
            torch.save(self.scaler.state_dict(), os.path.join(output_dir, "scaler.pt"))
        torch.save({"torch": torch.get_rng_state()},
                   os.path.join(output_dir, f"random_states_{self.process_index}.pkl"))
        return output_dir

    def load_state(self, input_dir=None, **_):
        for i, m in enumerate(self._models):
            m.load_state_dict(torch.load(os.path.join(input_dir, f"model_{i}.bin")))
        for i, o in enumerate(self._optimizers):
            o.load_state_dict(torch.load(os.path.join(input_dir, f"optimizer_{i}.bin")))
        for i, s in enumerate(self._schedulers):
            s.load_state_dict(torch.load(os.path.join(input_dir, f"scheduler_{i}.bin")))
        for i, c in enumerate(self._custom_objects):
            with open(os.path.join(input_dir, f"custom_checkpoint_{i}.pkl"), "rb") as f:
                c.load_state_dict(pickle.load(f))
        if self.scaler is not None:
            self.scaler.load_state_dict(torch.load(os.path.join(input_dir, "scaler.pt")))

    # -- trackers --------------------------------------------------------------------------------
    def get_tracker(self, name, unwrap=False):
        for t in self.trackers:
            if getattr(t, "name", None) == name:
                return t
        return tracking.GeneralTracker(_blank=True)

    def init_trackers(self, project_name, config=None, init_kwargs=None):
        for name in self.log_with:
            self.trackers.append(tracking.MemoryTracker(name))

    def end_training(self):
        for t in self.trackers:
            t.finish()

    def wait_for_everyone(self):
        if self.num_processes > 1:
            dist.barrier()


def notebook_launcher(function, args=(), num_processes=None, num_nodes=1, **_):
    function(*args)
