"""Engine: the handle every capsule holds in ``self._accelerator``.

It provides the member set the reference uses on ``accelerate.Accelerator`` (SURVEY.md
Appendix A; reference call sites are cited per method) and implements the engine-owned work
of the per-iteration step with the sm_100a kernels behind the C-ABI
(include/rocket_b200.h) plus NCCL for the gradient exchange:

    prepare(DataLoader)  -> EngineDataLoader : rank sharding + pinned-ring H2D staging (K5)
    prepare(Module)      -> the module itself : param/buffer broadcast, autocast forward
    prepare(Optimizer)   -> EngineOptimizer   : flat HBM layout, norm (K1) + fused step (K2/K3)
    prepare(LRScheduler) -> EngineScheduler
    accumulate / autocast / backward           : accumulation phase, bucketed all-reduce (K6)
    scalar_accumulator / argmax_count          : device-side Meter and loss bookkeeping (K4)

One process per GPU; ``torch.distributed`` (NCCL over NVLink) is initialised from the usual
RANK / WORLD_SIZE / LOCAL_RANK / MASTER_* environment when WORLD_SIZE > 1.
"""
from __future__ import annotations

import contextlib
import os

import torch
import torch.distributed as dist

from rocket_b200 import tracking
from rocket_b200.comm import (BufferSync, GradSync, P2PGradSync, broadcast_module_state,
                              p2p_possible)
from rocket_b200.deferred import DeferredScalar
from rocket_b200.flat import EngineOptimizer
from rocket_b200.scaler import LossScaler
from rocket_b200.staging import EngineDataLoader, PinnedRingStager


def _env_int(name: str, default: int) -> int:
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


class EngineScheduler:
    """Role of accelerate's AcceleratedScheduler (reference rocket/core/scheduler.py:92,113):
    steps only on sync steps, ``world_size`` times, never past a OneCycle ``total_steps``."""

    def __init__(self, scheduler, engine) -> None:
        self.scheduler = scheduler
        self.engine = engine

    def step(self, *args, **kwargs) -> None:
        eng = self.engine
        if not eng.sync_gradients:
            self.scheduler._step_count += 1     # accelerate's adjust_scheduler bookkeeping
            return
        if any(o.step_was_skipped for o in eng._optimizers):
            return
        for _ in range(eng.num_processes):
            if hasattr(self.scheduler, "total_steps"):
                if self.scheduler._step_count <= self.scheduler.total_steps:
                    self.scheduler.step(*args, **kwargs)
            else:
                self.scheduler.step(*args, **kwargs)

    def get_last_lr(self):
        return self.scheduler.get_last_lr()

    def state_dict(self):
        return self.scheduler.state_dict()

    def load_state_dict(self, state):
        self.scheduler.load_state_dict(state)


class GraphPass:
    """One pass of a ``Module(graph=True)`` capsule over its children while a CUDA graph of
    the step is being captured (``mode == "capture"``: device work is recorded, host logic
    runs) or is about to be replayed (``mode == "replay"``: host logic only, device work is
    already in the graph).  ``pre`` actions run on the stream right before the replay (uploads
    of this step's scalars), ``post`` actions right after it (read-backs).  ``slots`` keeps the
    graph's static tensors that participants need again on later replays."""

    def __init__(self, mode: str, slots: dict):
        self.mode = mode
        self.slots = slots
        self.pre: list = []
        self.post: list = []


class ScalarAccumulator:
    """Device-side running sum for the Loss capsule (reference rocket/core/loss.py:95-116)."""

    _CHUNK = 1024

    def __init__(self, engine, initial: float = 0.0) -> None:
        self.engine = engine
        self.acc = torch.full((1,), float(initial), dtype=torch.float64, device=engine.device)
        self._ring = None
        self._ring_pos = 0

    def add(self, value, divisor: float) -> None:
        gp = self.engine.graph_pass
        if gp is not None and gp.mode == "replay":
            return                                   # the accumulate kernel is in the graph
        self.engine.backend.scalar_accumulate(value.detach(), divisor, self.acc)

    def _slot(self) -> torch.Tensor:
        if self._ring is None or self._ring_pos == self._CHUNK:
            ring = torch.empty(self._CHUNK, dtype=torch.float64)
            self._ring = ring.pin_memory() if self.engine.device.type == "cuda" else ring
            self._ring_pos = 0
        slot = self._ring[self._ring_pos:self._ring_pos + 1]
        self._ring_pos += 1
        return slot

    def publish_and_reset(self) -> DeferredScalar:
        eng = self.engine
        gp = eng.graph_pass
        if gp is not None:
            return self._publish_in_graph(gp)
        out = self.acc.clone()
        self.acc.zero_()
        if eng.num_processes > 1:                 # mean over ranks, as gather(loss).mean()
            dist.all_reduce(out, op=dist.ReduceOp.SUM)
            out /= eng.num_processes
        slot = self._slot()
        slot.copy_(out, non_blocking=True)
        event = None
        if eng.device.type == "cuda":
            event = torch.cuda.Event()
            event.record()
        return DeferredScalar(slot, event)

    def _publish_in_graph(self, gp: GraphPass) -> DeferredScalar:
        """Graph passes: snapshot + reset are nodes of the graph writing a static tensor; the
        copy to the pinned ring happens after each replay."""
        eng = self.engine
        if gp.mode == "capture":
            out = self.acc.clone()
            self.acc.zero_()
            if eng.num_processes > 1:
                dist.all_reduce(out, op=dist.ReduceOp.SUM)
                out /= eng.num_processes
            gp.slots[("loss", id(self))] = out
        out = gp.slots[("loss", id(self))]
        slot = self._slot()
        value = DeferredScalar(slot, torch.cuda.Event(), armed=False)

        def read_back():
            slot.copy_(out, non_blocking=True)
            value.arm()
        gp.post.append(read_back)
        return value

    def peek(self) -> float:
        return float(self.acc.item())


class Engine:
    def __init__(self, device_placement: bool = True, mixed_precision: str | None = None,
                 gradient_accumulation_steps: int = 1, project_config=None,
                 project_dir: str | None = None, *, backend=None, device=None,
                 stage_slots: int = 3, stage_threads: int = 4, bucket_cap_mb: float = 16.0,
                 stage_batches: bool = True) -> None:
        world = _env_int("WORLD_SIZE", 1)
        self.local_process_index = _env_int("LOCAL_RANK", 0)
        if device is None:
            if torch.cuda.is_available():
                device = torch.device("cuda", self.local_process_index)
            elif backend is not None:
                device = torch.device("cpu")      # host-only test double
            else:
                raise RuntimeError(
                    "rocket_b200 runs on CUDA (B200, sm_100a) only: no CUDA device is "
                    "available and there is no CPU fallback.")
        self.device = torch.device(device)
        if self.device.type == "cuda":
            torch.cuda.set_device(self.device)
        if world > 1 and not dist.is_initialized():
            if self.device.type == "cuda":
                dist.init_process_group("nccl", device_id=self.device)
            else:
                dist.init_process_group("gloo")
        self._owns_pg = world > 1
        if dist.is_available() and dist.is_initialized():
            self.process_index, self.num_processes = dist.get_rank(), dist.get_world_size()
        else:
            self.process_index, self.num_processes = 0, 1
        if backend is None:
            from rocket_b200.backend import NativeBackend
            backend = NativeBackend(self.device)
        self.backend = backend
        self.device_placement = device_placement
        self.mixed_precision = mixed_precision if mixed_precision not in (None, "no") else "no"
        if self.mixed_precision not in ("no", "bf16", "fp16"):
            raise ValueError(f"unknown mixed_precision {mixed_precision!r}")
        # fp16: dynamic loss scaling, GradScaler's defaults as accelerate constructs it
        self.scaler = LossScaler(self.device) if self.mixed_precision == "fp16" else None
        self._accum_steps = int(gradient_accumulation_steps)
        self.sync_gradients = True
        self.step = 0
        self.project_dir = project_dir if project_dir is not None else \
            getattr(project_config, "project_dir", None)
        self._models, self._optimizers, self._schedulers = [], [], []
        self._dataloaders, self._custom_objects = [], []
        self.log_with: list = []
        self.trackers: list = []
        self._active_dataloaders: list = []
        self._grad_sync: dict = {}
        self._buffer_sync: list = []
        self._bucket_cap_mb = bucket_cap_mb
        self._p2p = None
        self.graph_pass: GraphPass | None = None    # set by Module(graph=True) around a pass
        self.replayed_kernel_launches = 0           # engine kernels run by graph replays
        self.grad_sync_kinds: list[str] = []      # which exchange each optimizer got
        self.stager = None
        if self.device.type == "cuda" and stage_batches:
            self.stager = PinnedRingStager(self.device, slots=stage_slots,
                                           copy_threads=stage_threads)

    # ---- plain attributes the capsules read -----------------------------------------------------
    @property
    def gradient_accumulation_steps(self) -> int:
        return self._accum_steps

    @property
    def is_main_process(self) -> bool:
        return self.process_index == 0

    @property
    def is_local_main_process(self) -> bool:
        return self.local_process_index == 0

    # ---- prepare (reference dataset.py:178, module.py:106, optimizer.py:109, scheduler.py:92) ----
    def prepare(self, *objects, device_placement=None):
        out = tuple(self._prepare_one(o) for o in objects)
        return out[0] if len(out) == 1 else out

    def _prepare_one(self, obj):
        if isinstance(obj, torch.utils.data.DataLoader):
            loader = EngineDataLoader(obj, self)
            self._dataloaders.append(loader)
            return loader
        if isinstance(obj, torch.nn.Module):
            return self._prepare_model(obj)
        if isinstance(obj, torch.optim.Optimizer):
            return self.prepare_optimizer(obj)
        if isinstance(obj, torch.optim.lr_scheduler.LRScheduler):
            sched = EngineScheduler(obj, self)
            self._schedulers.append(sched)
            return sched
        return obj

    def prepare_optimizer(self, obj: torch.optim.Optimizer, sharded: bool | None = None):
        """``prepare`` for an optimizer, with the one choice ``prepare`` has no argument for:
        ``sharded=True`` makes each rank update 1/W of the parameters and write them to all
        ranks over NVLink (rk_comm_sharded_step) instead of repeating the whole step on every
        replica.  Same results bit for bit; default from ROCKET_B200_SHARDED_STEP, else off."""
        if sharded is None:
            sharded = os.environ.get("ROCKET_B200_SHARDED_STEP", "0") == "1"
        opt = EngineOptimizer(obj, self, sharded=sharded)
        if self.num_processes > 1:
            cls = P2PGradSync if opt.arena is not None else GradSync
            self._grad_sync[id(opt)] = cls(opt, self, self._bucket_cap_mb)
            self.grad_sync_kinds.append(cls.__name__ + ("+sharded" if opt.sharded else ""))
        self._optimizers.append(opt)
        return opt

    def _prepare_model(self, model: torch.nn.Module) -> torch.nn.Module:
        if self.num_processes > 1:
            broadcast_module_state(model, src=0)
            if any(True for _ in model.buffers()):
                self._buffer_sync.append(BufferSync(model))
        if self.mixed_precision in ("bf16", "fp16") and not getattr(model, "_rk_autocast", False):
            inner = model.forward
            dev_type = self.device.type
            dtype = self._autocast_dtype

            def forward(*args, **kwargs):
                with torch.autocast(device_type=dev_type, dtype=dtype):
                    out = inner(*args, **kwargs)
                return _to_fp32(out)

            model.forward = forward
            model._rk_autocast = True
        self._models.append(model)
        return model

    def skip_first_batches(self, dataloader, num_batches: int = 0):
        """reference dataset.py:206-208: loader that skips the first batches (resume)."""
        return EngineDataLoader(dataloader.base, self,
                                skip_batches=dataloader.skip_batches + num_batches)

    # ---- step contexts (reference module.py:210-211) ------------------------------------------------
    def _begin_dataloader(self, loader) -> None:
        self._active_dataloaders.append(loader)

    def _end_dataloader(self, loader) -> None:
        if loader in self._active_dataloaders:
            self._active_dataloaders.remove(loader)

    @property
    def end_of_dataloader(self) -> bool:
        return bool(self._active_dataloaders) and self._active_dataloaders[-1].end_of_dataloader

    @property
    def remainder(self) -> int:
        return self._active_dataloaders[-1].remainder if self._active_dataloaders else -1

    @contextlib.contextmanager
    def accumulate(self, *models):
        """Advance the accumulation phase: every N-th call, and always on the last batch of
        the active loader, is a gradient-sync step [accelerate ``_do_sync``, from memory]."""
        if self.end_of_dataloader:
            self.step = 0
            self.sync_gradients = True
        else:
            self.step += 1
            self.sync_gradients = (self.step % self._accum_steps) == 0
        yield

    @contextlib.contextmanager
    def autocast(self, autocast_handler=None):
        if self.mixed_precision in ("bf16", "fp16"):
            with torch.autocast(device_type=self.device.type, dtype=self._autocast_dtype):
                yield
        else:
            yield

    @property
    def _autocast_dtype(self):
        return torch.bfloat16 if self.mixed_precision == "bf16" else torch.float16

    def backward(self, loss: torch.Tensor, **kwargs) -> None:
        """reference loss.py:119: scale by 1/accumulation_steps, then autograd.  Bucketed
        all-reduces start from gradient hooks while this runs (sync steps only)."""
        if self.graph_pass is not None and self.graph_pass.mode == "replay":
            return                                   # backward is part of the graph
        loss = loss / self._accum_steps
        if self.scaler is not None:                  # accelerate: scaler.scale(loss).backward()
            loss = self.scaler.scale(loss)
        loss.backward(**kwargs)

    # ---- gradient exchange -----------------------------------------------------------------------------
    def wants_p2p(self) -> bool:
        """Use the fused NVLink all-reduce kernel (comm.P2PGradSync) for gradient exchange?"""
        if self._p2p is None:
            self._p2p = hasattr(self.backend, "comm_create") and p2p_possible(self)
            if self.num_processes > 1:           # all ranks must take the same path
                flag = torch.tensor([int(self._p2p)], device=self.device)
                dist.all_reduce(flag, op=dist.ReduceOp.MIN)
                self._p2p = bool(flag.item())
        return self._p2p

    def grad_scale(self, opt=None) -> float:
        """Factor the kernels still apply when they read a gradient: 1/world_size after a
        plain SUM on the wire (NCCL path), 1 when the exchange kernel already averaged."""
        sync = self._grad_sync.get(id(opt)) if opt is not None else None
        if sync is not None:
            return sync.grad_scale
        return 1.0 / self.num_processes if opt is None else 1.0

    def grad_sync_of(self, opt):
        return self._grad_sync.get(id(opt))

    def finish_grad_sync(self, opt) -> bool:
        """Complete the gradient exchange of this step.  True if it also left the norm's
        per-chunk partials in the plan's scratch (fused path)."""
        sync = self._grad_sync.get(id(opt))
        return sync.finish() if sync is not None else False

    def grad_sync_step_done(self, opt) -> None:
        sync = self._grad_sync.get(id(opt))
        if sync is not None:
            sync.step_done()

    def clip_grad_norm_(self, parameters=None, max_norm: float = 1.0, norm_type: float = 2.0):
        """Global-norm clipping for user capsules (accelerate: ``clip_grad_norm_``).  The
        coefficient is applied inside the next fused step; returns the device norm."""
        if float(norm_type) != 2.0:
            raise NotImplementedError("only the L2 norm is implemented")
        if not self.sync_gradients:
            return None
        norms = []
        for opt in self._optimizers:
            self.finish_grad_sync(opt)
            norms.append(opt.clip_grad_norm_(max_norm))
        if len(norms) != 1:
            raise NotImplementedError("clip_grad_norm_ with several optimizers: set "
                                      "max_grad_norm on each Optimizer capsule instead")
        return norms[0]

    # ---- collectives the capsules call (reference loss.py:95, meter.py:93) ----------------------------------
    def gather(self, tensor):
        if self.num_processes == 1:
            return tensor
        return _recursively(self._gather_one, tensor)

    def _gather_one(self, t: torch.Tensor) -> torch.Tensor:
        t = t.detach()
        if t.ndim == 0:
            t = t[None]
        t = t.contiguous()
        out = torch.empty((self.num_processes * t.shape[0],) + tuple(t.shape[1:]),
                          dtype=t.dtype, device=t.device)
        dist.all_gather_into_tensor(out, t) if t.device.type == "cuda" else \
            dist.all_gather(list(out.chunk(self.num_processes)), t)
        return out

    def gather_for_metrics(self, input_data):
        data = self.gather(input_data)
        if self.num_processes > 1 and self.end_of_dataloader and self.remainder > 0:
            rem = self.remainder
            return _recursively(lambda t: t[:rem], data)
        return data

    def valid_rows(self, n_rows: int) -> int:
        """How many of this rank's ``n_rows`` samples in the current batch are real (not
        ``even_batches`` padding): the local view of gather_for_metrics' truncation."""
        if self.num_processes == 1 or not self.end_of_dataloader or self.remainder <= 0:
            return n_rows
        lo = self.process_index * n_rows
        return max(0, min(n_rows, self.remainder - lo))

    def broadcast_object_list(self, object_list, from_process: int = 0):
        if self.num_processes > 1:
            dist.broadcast_object_list(object_list, src=from_process)
        return object_list

    def wait_for_everyone(self) -> None:
        if self.num_processes > 1:
            dist.barrier()

    # ---- device-side Meter / loss bookkeeping ---------------------------------------------------------------
    def scalar_accumulator(self, initial: float = 0.0) -> ScalarAccumulator:
        return ScalarAccumulator(self, initial)

    def new_counters(self, n: int = 2) -> torch.Tensor:
        return torch.zeros(n, dtype=torch.int64, device=self.device)

    def argmax_count(self, logits, labels, counters) -> None:
        self.backend.argmax_count(logits, labels, counters)

    def read_counters(self, counters: torch.Tensor, reduce: bool = True) -> DeferredScalar:
        """Async copy of integer counters to the host (summed over ranks when ``reduce``)."""
        out = counters.clone()
        if reduce and self.num_processes > 1:
            dist.all_reduce(out, op=dist.ReduceOp.SUM)
        host = torch.empty(out.shape, dtype=out.dtype)
        event = None
        if self.device.type == "cuda":
            host = host.pin_memory()
            host.copy_(out, non_blocking=True)
            event = torch.cuda.Event()
            event.record()
        else:
            host.copy_(out)
        return DeferredScalar(host, event)

    # ---- checkpointing (reference capsule.py:139, checkpoint.py:129, launcher.py:356-363) ----------------------
    def register_for_checkpointing(self, *objects) -> None:
        bad = [o for o in objects
               if not (hasattr(o, "state_dict") and hasattr(o, "load_state_dict"))]
        if bad:
            raise ValueError("objects registered for checkpointing need state_dict() and "
                             f"load_state_dict(): {[type(o).__name__ for o in bad]}")
        self._custom_objects.extend(objects)

    # File names and serializers follow accelerate's `save_state` / `load_state` layout, so that
    # a checkpoint directory means the same to either library [accelerate >= 0.34, from memory:
    # checkpointing.py save_accelerator_state / load_accelerator_state; the package is not on
    # this image, so the layout is restated, not verified against an install]:
    #   model.safetensors, model_1.safetensors ...   (read side also accepts pytorch_model*.bin)
    #   optimizer.bin, optimizer_1.bin ...           torch.save of the torch-format state_dict
    #   scheduler.bin ...                            torch.save
    #   scaler.pt                                    GradScaler's state_dict (fp16 only)
    #   random_states_<rank>.pkl                     torch.save of python / numpy / torch / cuda RNG
    #   custom_checkpoint_<i>.pkl                    torch.save of each registered object's state
    @staticmethod
    def _numbered(stem: str, i: int, ext: str) -> str:
        return f"{stem}{'' if i == 0 else f'_{i}'}.{ext}"

    @staticmethod
    def _model_tensors(model: torch.nn.Module) -> dict:
        """state_dict as safetensors wants it: host copies, one name per aliased tensor (tied
        weights are written once, under their first name, like safetensors' save_model)."""
        out, seen = {}, set()
        for name, t in model.state_dict().items():
            key = (t.data_ptr(), tuple(t.shape), t.dtype) if t.numel() else (name,)
            if key in seen:
                continue
            seen.add(key)
            out[name] = t.detach().cpu().contiguous()
        return out

    def save_state(self, output_dir: str | None = None, **_) -> str:
        if output_dir is None:
            if self.project_dir is None:
                raise ValueError("save_state needs an output_dir or a project_dir")
            output_dir = os.path.join(self.project_dir, "checkpoints")
        from safetensors.torch import save_file
        # collective parts first (a sharded optimizer gathers its moments): every rank gets here
        states = {
            "optimizers": [o.state_dict() for o in self._optimizers],
            "schedulers": [s.state_dict() for s in self._schedulers],
            "custom": [c.state_dict() for c in self._custom_objects],
        }
        os.makedirs(output_dir, exist_ok=True)
        if self.is_main_process:
            for i, m in enumerate(self._models):
                save_file(self._model_tensors(m),
                          os.path.join(output_dir, self._numbered("model", i, "safetensors")))
            for i, sd in enumerate(states["optimizers"]):
                torch.save(sd, os.path.join(output_dir, self._numbered("optimizer", i, "bin")))
            for i, sd in enumerate(states["schedulers"]):
                torch.save(sd, os.path.join(output_dir, self._numbered("scheduler", i, "bin")))
            for i, sd in enumerate(states["custom"]):
                torch.save(sd, os.path.join(output_dir, f"custom_checkpoint_{i}.pkl"))
            if self.scaler is not None:          # accelerate: scaler.pt, GradScaler's format
                torch.save(self.scaler.state_dict(), os.path.join(output_dir, "scaler.pt"))
        import random

        import numpy as np
        rng = {"step": self.step, "random_state": random.getstate(),
               "numpy_random_seed": np.random.get_state(),
               "torch_manual_seed": torch.get_rng_state()}
        if self.device.type == "cuda":
            rng["torch_cuda_manual_seed"] = torch.cuda.get_rng_state_all()
        torch.save(rng, os.path.join(output_dir, f"random_states_{self.process_index}.pkl"))
        self.wait_for_everyone()
        return output_dir

    def load_state(self, input_dir: str | None = None, **_) -> None:
        if input_dir is None or not os.path.isdir(input_dir):
            raise RuntimeError(f"checkpoint directory {input_dir!r} does not exist")
        dev = self.device
        for i, m in enumerate(self._models):
            safe = os.path.join(input_dir, self._numbered("model", i, "safetensors"))
            if os.path.exists(safe):
                from safetensors.torch import load_file
                sd = load_file(safe, device=str(dev))
                missing, unexpected = m.load_state_dict(sd, strict=False)
                # names left out on purpose: aliases of a tensor that was written
                written = {t.data_ptr() for k, t in m.state_dict().items() if k in sd}
                lost = [k for k in missing if m.state_dict()[k].data_ptr() not in written]
                if lost or unexpected:
                    raise RuntimeError(f"{safe}: missing {lost}, unexpected {list(unexpected)}")
            else:                                 # accelerate with safe_serialization=False
                m.load_state_dict(torch.load(
                    os.path.join(input_dir, self._numbered("pytorch_model", i, "bin")),
                    map_location=dev, weights_only=True))
        for i, o in enumerate(self._optimizers):
            o.load_state_dict(torch.load(
                os.path.join(input_dir, self._numbered("optimizer", i, "bin")),
                map_location=dev, weights_only=True))
        for i, s in enumerate(self._schedulers):
            s.load_state_dict(torch.load(
                os.path.join(input_dir, self._numbered("scheduler", i, "bin")),
                weights_only=False))              # scheduler state may hold plain objects
        for i, c in enumerate(self._custom_objects):
            c.load_state_dict(torch.load(os.path.join(input_dir, f"custom_checkpoint_{i}.pkl"),
                                         map_location="cpu", weights_only=False))
        if self.scaler is not None:
            self.scaler.load_state_dict(torch.load(os.path.join(input_dir, "scaler.pt")))
        rng_path = os.path.join(input_dir, f"random_states_{self.process_index}.pkl")
        if os.path.exists(rng_path):
            import random

            import numpy as np
            rng = torch.load(rng_path, weights_only=False)
            self.step = rng.get("step", self.step)
            random.setstate(rng["random_state"])
            np.random.set_state(rng["numpy_random_seed"])
            torch.set_rng_state(rng["torch_manual_seed"])
            if "torch_cuda_manual_seed" in rng and dev.type == "cuda":
                torch.cuda.set_rng_state_all(rng["torch_cuda_manual_seed"])

    # ---- trackers (reference tracker.py:89-105) -----------------------------------------------------------------
    def get_tracker(self, name: str, unwrap: bool = False):
        for t in self.trackers:
            if t.name == name:
                return t
        return tracking.GeneralTracker()

    def init_trackers(self, project_name: str, config: dict | None = None, init_kwargs=None):
        for name in self.log_with:
            if any(t.name == name for t in self.trackers):
                continue
            if isinstance(name, tracking.GeneralTracker):
                self.trackers.append(name)
            else:
                self.trackers.append(
                    tracking.make_tracker(name, project_name, self.project_dir))

    def end_training(self) -> None:
        for t in self.trackers:
            t.finish()
        for sync in self._grad_sync.values():
            sync.remove()
        self._grad_sync.clear()
        for b in self._buffer_sync:
            b.remove()
        self._buffer_sync.clear()
        if self.stager is not None:
            self.stager.close()

    def destroy_process_group(self) -> None:
        if dist.is_available() and dist.is_initialized():
            dist.destroy_process_group()


def _recursively(fn, data):
    if isinstance(data, torch.Tensor):
        return fn(data)
    if isinstance(data, dict):
        return type(data)({k: _recursively(fn, v) for k, v in data.items()})
    if isinstance(data, tuple) and hasattr(data, "_fields"):
        return type(data)(*[_recursively(fn, v) for v in data])
    if isinstance(data, (list, tuple)):
        return type(data)(_recursively(fn, v) for v in data)
    return data


def _to_fp32(data):
    return _recursively(
        lambda t: t.float() if t.dtype in (torch.bfloat16, torch.float16) else t, data)


def notebook_launcher(function, args=(), num_processes=None, num_nodes=1, **_):
    """Jupyter entry (reference launcher.py:238-244).  Single process: call through;
    multi-process spawning from a notebook is not part of the hot path."""
    if num_processes not in (None, 1):
        raise NotImplementedError("use torchrun for multi-process runs")
    function(*args)
