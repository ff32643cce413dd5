"""Prepared optimizer: flat HBM layout + the fused step.

What ``Engine.prepare(optimizer)`` returns.  It plays the role of accelerate's
``AcceleratedOptimizer`` for the reference's Optimizer capsule (reference
rocket/core/optimizer.py:92-109,129-130,138): ``.optimizer``, ``.param_groups``, ``.step()``
and ``.zero_grad()``, the last two being no-ops on gradient-accumulation steps.

HBM layout (built once): every trainable parameter of every param group is re-homed into
one flat fp32 buffer ``flat_p`` at a 256-byte aligned offset, and ``flat_g`` / ``flat_m`` /
``flat_v`` mirror it (sharded step: ``flat_p`` joins ``flat_g`` in the peer-mapped arena and
the moments shrink to this rank's compact shards ``shard_m`` / ``shard_v``).  ``param.data``, ``param.grad`` and the optimizer's
``state[p]["exp_avg"]`` ... are views into those buffers, so user code, ``state_dict()`` and
checkpoints see ordinary per-parameter tensors while the kernels see four aligned streams
and the gradient all-reduce sees contiguous buckets (no copy-in / copy-out).

Gradients stay allocated: the fused kernel zero-fills them, which is the in-place form of
``zero_grad()``.  (The reference's ``zero_grad`` frees them, ``set_to_none=True``; code
that tests ``p.grad is None`` after a step sees the difference.)  Parameters that have never
received a gradient are skipped, as torch.optim skips ``grad is None``.
"""
from __future__ import annotations

import math

import torch

from rocket_b200.backend import TensorSpec

_ALIGN = 64  # elements: 256-byte aligned segment starts


def _align(n: int) -> int:
    return (n + _ALIGN - 1) // _ALIGN * _ALIGN


class EngineOptimizer:
    def __init__(self, optimizer: torch.optim.Optimizer, engine, sharded: bool = False) -> None:
        self.optimizer = optimizer
        self.engine = engine
        # sharded step (csrc/rk_comm.cu, K7): every rank updates 1/W of the chunks and writes
        # the new parameters to all ranks.  Needs the NVLink exchange path; otherwise the step
        # stays replicated (which is what one GPU is anyway).
        self.sharded = bool(sharded) and engine.num_processes > 1 and engine.wants_p2p()
        self._skipped = False            # fp16: the last step overflowed and was skipped
        self.max_grad_norm: float | None = None
        self.last_grad_norm: torch.Tensor | None = None   # device scalar, for logging
        self._zeroed_by_step = False
        self._pending_clip: float | None = None
        self._kind = self._classify(optimizer)
        self._build()

    # ---- construction --------------------------------------------------------------------
    @staticmethod
    def _classify(opt) -> str:
        kind = {torch.optim.AdamW: "adamw", torch.optim.Adam: "adam",
                torch.optim.SGD: "sgd"}.get(type(opt))
        if kind is None:
            raise NotImplementedError(
                f"rocket_b200 fuses torch.optim.AdamW, Adam and SGD; got {type(opt).__name__}. "
                "There is no unfused fallback path.")
        for group in opt.param_groups:
            for flag in ("amsgrad", "maximize", "capturable", "differentiable"):
                if group.get(flag):
                    raise NotImplementedError(f"{type(opt).__name__}({flag}=True) is not supported")
            if isinstance(group["lr"], torch.Tensor):
                raise NotImplementedError("tensor learning rates are not supported")
        if len(opt.param_groups) > 64:
            raise NotImplementedError("at most 64 param groups (RK_MAX_GROUPS)")
        return kind

    def _build(self) -> None:
        dev = self.engine.device
        self._params: list[torch.nn.Parameter] = []
        self._group_of: list[int] = []
        self._offsets: list[int] = []
        off = 0
        self._frozen: list[torch.nn.Parameter] = []     # requires_grad=False at prepare time
        self._n_group_params = sum(len(g["params"]) for g in self.optimizer.param_groups)
        for gi, group in enumerate(self.optimizer.param_groups):
            for p in group["params"]:
                if not p.requires_grad:
                    self._frozen.append(p)
                    continue
                if p.dtype != torch.float32:
                    raise NotImplementedError(
                        f"parameters must be fp32 (master weights); got {p.dtype}")
                if p.device != dev:
                    raise RuntimeError(
                        f"parameter on {p.device}, engine on {dev}: prepare the module first")
                self._params.append(p)
                self._group_of.append(gi)
                self._offsets.append(off)
                off += _align(p.numel())
        if not self._params:
            raise ValueError("optimizer has no trainable parameters")
        self._total = off
        # Gradients live in an "arena" when the fused NVLink all-reduce will be used: one
        # allocation [signal pad | norm scratch | gradients] that the peer ranks map (comm.py).
        self.arena = self._scratch_view = None
        if self.engine.wants_p2p():
            from rocket_b200 import _native as N
            sig = N.RK_COMM_SIGNAL_BYTES
            scratch = -(-self.engine.backend.scratch_bytes([p.numel() for p in self._params])
                        // 256) * 256
            from rocket_b200.comm import alloc_arena
            # sharded step: the parameters follow the gradients, so that owners can write
            # them into every rank's copy      [signals | scratch | gradients | parameters]
            self.arena = alloc_arena(sig + scratch + 4 * off * (2 if self.sharded else 1), dev)
            self._scratch_offset = sig
            self._scratch_view = self.arena[sig:sig + scratch]
            g0 = sig + scratch
            self.flat_g = self.arena[g0:g0 + 4 * off].view(torch.float32)
            if self.sharded:
                self.flat_p = self.arena[g0 + 4 * off:].view(torch.float32)
        else:
            self.flat_g = torch.zeros(off, dtype=torch.float32, device=dev)
        if not self.sharded:
            self.flat_p = torch.zeros(off, dtype=torch.float32, device=dev)
        self._needs_m = self._kind != "sgd" or any(
            g.get("momentum", 0) != 0 for g in self.optimizer.param_groups)
        # Moments: full-size flat buffers, or -- sharded step -- this rank's compact shards
        # (1/W of that), allocated below once the chunk layout is known.
        self.flat_m = self.flat_v = self.shard_m = self.shard_v = None
        if not self.sharded:
            if self._needs_m:
                self.flat_m = torch.zeros(off, dtype=torch.float32, device=dev)
            if self._kind != "sgd":
                self.flat_v = torch.zeros(off, dtype=torch.float32, device=dev)
        self._views_p, self._views_g, self._views_m, self._views_v = [], [], [], []
        with torch.no_grad():
            for p, o in zip(self._params, self._offsets):
                n = p.numel()
                vp = self.flat_p[o:o + n].view(p.shape)
                vp.copy_(p.data)
                p.data = vp
                self._views_p.append(vp)
                self._views_g.append(self.flat_g[o:o + n].view(p.shape))
                self._views_m.append(self.flat_m[o:o + n].view(p.shape)
                                     if self.flat_m is not None else None)
                self._views_v.append(self.flat_v[o:o + n].view(p.shape)
                                     if self.flat_v is not None else None)
        self._grad_ptr = [v.data_ptr() for v in self._views_g]
        self._adopted = [False] * len(self._params)      # grad lives in flat_g
        # torch keeps Adam's `step` (and SGD's "no momentum buffer yet") per PARAMETER
        # ($TORCH/optim/adam.py:168-175,414; sgd.py:359-365): a parameter whose first gradient
        # arrives after its group-mates' gets its own bias correction.  Here: steps taken per
        # group, and per parameter how many of them it missed before its gradient first showed
        # up; parameters of one group with the same lag share one row of the kernels'
        # constant table (normally: one row per group).
        self._group_step = [0] * len(self.optimizer.param_groups)
        self._lag = [0] * len(self._params)
        self._lag_known = [False] * len(self._params)    # taken from imported optimizer state
        self._rows: list[tuple[int, int]] = []           # (param group, lag) per table row
        self._row_of = [0] * len(self._params)
        self._plan = None
        self._table_ring = None
        self._dev_table = torch.zeros(4096, dtype=torch.uint8, device=dev)
        if self.sharded:
            plan = self._get_plan()        # the chunk layout depends on the sizes only
            self._shard_rows = -(-plan.n_chunks_norm // self.engine.num_processes)
            n = self._shard_rows * plan.chunk_norm
            if self._needs_m:
                self.shard_m = torch.zeros(n, dtype=torch.float32, device=dev)
            if self._kind != "sgd":
                self.shard_v = torch.zeros(n, dtype=torch.float32, device=dev)
        self._import_state()

    def _import_state(self) -> None:
        """Adopt state the wrapped optimizer already holds (resume / prior training)."""
        for i, p in enumerate(self._params):
            st = self.optimizer.state.get(p)
            if not st:
                continue
            gi = self._group_of[i]
            if "step" in st:
                self._group_step[gi] = max(self._group_step[gi], int(float(st["step"])))
                self._lag[i] = -int(float(st["step"]))    # fixed up below, once the max is known
                self._lag_known[i] = True
            if self._kind == "sgd":
                buf = st.get("momentum_buffer")
                if buf is not None and self._needs_m:
                    if self.sharded:
                        self._scatter(i, buf, self.shard_m)
                    else:
                        self._views_m[i].copy_(buf)
                        st["momentum_buffer"] = self._views_m[i]
                    self._group_step[gi] = max(self._group_step[gi], 1)
                    self._lag_known[i] = True
            elif "exp_avg" in st:
                if self.sharded:
                    self._scatter(i, st["exp_avg"], self.shard_m)
                    self._scatter(i, st["exp_avg_sq"], self.shard_v)
                else:
                    self._views_m[i].copy_(st["exp_avg"])
                    self._views_v[i].copy_(st["exp_avg_sq"])
                    st["exp_avg"], st["exp_avg_sq"] = self._views_m[i], self._views_v[i]
        for i in range(len(self._params)):
            if self._lag[i] < 0:                 # had a `step`: lag = group's steps - its own
                self._lag[i] += self._group_step[self._group_of[i]]
            elif self._lag_known[i] and self._kind == "sgd":
                self._lag[i] = 0                 # has a momentum buffer: not its first step

    # ---- sharded step: compact moment shards ---------------------------------------------------
    # Norm chunk c (global index over all tensors, plan.prefix_norm) belongs to rank c % W, as
    # in the exchange kernel, and its moments occupy slot c // W of that rank's shard; a slot
    # is chunk_norm floats (a tensor's short last chunk leaves the rest of its slot unused).
    def _scatter(self, i: int, full: torch.Tensor, shard: torch.Tensor) -> None:
        """This rank's chunks of parameter i's full moment tensor -> its shard slots."""
        plan = self._get_plan()
        world, rank = self.engine.num_processes, self.engine.process_index
        chunk, c0, n = plan.chunk_norm, int(plan.prefix_norm[i]), self._params[i].numel()
        n_chunks = -(-n // chunk)
        if n_chunks == 0:
            return
        padded = torch.zeros(n_chunks * chunk, dtype=torch.float32, device=shard.device)
        padded[:n] = full.detach().reshape(-1).to(shard.device)
        j0 = (rank - c0) % world
        own = padded.view(n_chunks, chunk)[j0::world]
        if own.shape[0]:
            k0 = (c0 + j0) // world
            shard.view(-1, chunk)[k0:k0 + own.shape[0]] = own

    @staticmethod
    def _interleave(shards: torch.Tensor, chunk: int) -> torch.Tensor:
        """[W, rows * chunk] (one shard per rank) -> [rows * W, chunk] in global chunk order."""
        world = shards.shape[0]
        return shards.view(world, -1, chunk).permute(1, 0, 2).reshape(-1, chunk)

    def _gather_chunks(self, shard: torch.Tensor) -> torch.Tensor:
        """Collective: every rank's shard, as one table of chunks in global order."""
        import torch.distributed as dist
        world = self.engine.num_processes
        out = torch.empty(world, shard.numel(), dtype=shard.dtype, device=shard.device)
        dist.all_gather_into_tensor(out, shard)
        return self._interleave(out, self._get_plan().chunk_norm)

    def _assemble(self, i: int, chunks: torch.Tensor) -> torch.Tensor:
        """Parameter i's full moment tensor out of the table ``_gather_chunks`` returns."""
        plan = self._get_plan()
        p = self._params[i]
        c0, c1 = int(plan.prefix_norm[i]), int(plan.prefix_norm[i + 1])
        return chunks[c0:c1].reshape(-1)[:p.numel()].view(p.shape).clone()

    # ---- the AcceleratedOptimizer surface ---------------------------------------------------
    @property
    def param_groups(self):
        return self.optimizer.param_groups

    @property
    def state(self):
        """torch's per-parameter state.  With a sharded step this gathers the moment shards
        (a collective: every rank has to read it, like ``state_dict()``)."""
        self._export_state()
        return self.optimizer.state

    @property
    def step_was_skipped(self) -> bool:
        return self._skipped

    def zero_grad(self, set_to_none: bool | None = None) -> None:
        if not self.engine.sync_gradients:
            return
        if self._zeroed_by_step:       # the fused step already cleared flat_g
            self._zeroed_by_step = False
            return
        self.flat_g.zero_()
        if set_to_none:
            for i, p in enumerate(self._params):
                if not self._adopted[i]:
                    p.grad = None

    def clip_grad_norm_(self, max_norm: float) -> torch.Tensor:
        """Engine-level ``clip_grad_norm_``: computes the global norm now (one kernel) and
        applies the coefficient inside the next fused step.  Returns the device norm."""
        gp = self.engine.graph_pass
        if gp is not None and gp.mode == "replay":     # the norm kernel is part of the graph
            self._pending_clip = float(max_norm)
            return self.last_grad_norm
        self._adopt_grads()
        plan = self._get_plan()
        if self.engine.finish_grad_sync(self):
            norm = self.engine.backend.l2norm_finalize(plan)
            if self.engine.scaler is not None:       # partials of loss-scaled gradients (step())
                norm.mul_(self.engine.scaler.inv_scale)
        else:
            norm = self.engine.backend.l2norm(plan, self._read_scale())
        self._pending_clip = float(max_norm)
        self.last_grad_norm = norm
        return norm

    def step(self, closure=None) -> None:
        if closure is not None:
            raise NotImplementedError("closures are not supported by the fused step")
        eng = self.engine
        if not eng.sync_gradients:
            return
        gp = eng.graph_pass                      # None, or a Module's CUDA-graph pass
        replay = gp is not None and gp.mode == "replay"
        scaler = eng.scaler                      # fp16 loss scaler, or None
        if scaler is not None and gp is not None:
            raise NotImplementedError("Module(graph=True) with mixed_precision='fp16'")
        max_norm = self._pending_clip if self._pending_clip is not None else self.max_grad_norm
        if not replay:
            self._adopt_grads()
            if gp is not None and self._plan is None:
                raise RuntimeError("CUDA-graph capture before the optimizer's first eager step")
            plan = self._get_plan()
            fused_norm = eng.finish_grad_sync(self)  # True: norm partials are in plan.scratch
            scale = self._read_scale()
            # fp16 on the NVLink path: the exchange's fused partials are of the still loss-scaled
            # (averaged) gradients.  1/scale is a power of two, so the norm of the unscaled
            # gradients is that norm times 1/scale, exactly -- no second pass over the gradients.
            rescale = scaler.inv_scale if (scaler is not None and fused_norm) else 1.0
            norm = None
            if max_norm is not None or scaler is not None:
                if self._pending_clip is not None:
                    norm = self.last_grad_norm
                elif fused_norm:
                    norm = eng.backend.l2norm_finalize(plan)
                    if rescale != 1.0:
                        norm.mul_(rescale)
                else:
                    norm = eng.backend.l2norm(plan, scale)
                self.last_grad_norm = norm
        self._pending_clip = None
        if scaler is not None:
            # fp16: the norm of the unscaled gradients is also GradScaler's overflow check (any
            # non-finite element makes it non-finite); reading it is the one host sync
            # GradScaler.step makes too (found_inf.item()).  Overflow: skip the step, keep the
            # step counters, clear the gradients, back the scale off.
            self._skipped = not math.isfinite(float(norm))
            scaler.update(self._skipped)
            if self._skipped:
                self.flat_g.zero_()
                eng.grad_sync_step_done(self)
                self._zeroed_by_step = True
                return
            if max_norm is None:
                norm = None
        groups = self._step_groups()
        if not groups:
            # no parameter has received a gradient yet: torch.optim skips every `grad is None`
            # parameter, i.e. the step is a no-op
            eng.grad_sync_step_done(self)
            self._zeroed_by_step = True
            return
        if self.sharded:
            # the sharded kernel always reads its constants from the device table; inside a
            # graph pass the upload moves in front of the replay (gp.pre), as for the replicated
            # step, and the launch itself is captured once (its barrier epoch lives on the device)
            slot, uploaded = self._table_slot()
            eng.backend.step_table_write(self._kind, groups, slot)
            table = self._dev_table
            if gp is None:
                table.copy_(slot, non_blocking=True)
                uploaded.record()
            else:
                gp.post.append(uploaded.record)
                gp.pre.append(lambda: table.copy_(slot, non_blocking=True))
            if not replay:
                sync = eng.grad_sync_of(self)
                eng.backend.comm_sharded_step(
                    sync._comm, plan, self._kind, table.data_ptr(), norm,
                    float(max_norm) if max_norm is not None else 0.0, sync.step_ctas,
                    sync.use_multicast, self.shard_m, self.shard_v, grad_scale=scale)
        elif gp is None:
            fn = eng.backend.sgd_step if self._kind == "sgd" else eng.backend.adam_step
            fn(plan, groups, norm, float(max_norm) if max_norm is not None else 0.0, scale, True)
        else:
            # Kernel arguments are frozen inside a graph: this step's constants go through a
            # pinned ring slot into a device table, uploaded in stream order before the replay.
            slot, uploaded = self._table_slot()
            gp.post.append(uploaded.record)
            eng.backend.step_table_write(self._kind, groups, slot)
            table = self._dev_table
            gp.pre.append(lambda: table.copy_(slot, non_blocking=True))
            if not replay:
                eng.backend.step_devtable(self._kind, plan, table, norm,
                                          float(max_norm) if max_norm is not None else 0.0,
                                          scale, True)
        eng.grad_sync_step_done(self)
        self._zeroed_by_step = True
        self.optimizer._opt_called = True   # what torch's LRScheduler checks before warning

    def _read_scale(self) -> float:
        """Factor applied to every gradient as the kernels read it: 1/world_size where the
        exchange left a plain sum, times 1/loss_scale under fp16 (GradScaler's unscale pass,
        folded in; both are powers of two, so the product is exact)."""
        scale = self.engine.grad_scale(self)
        scaler = getattr(self.engine, "scaler", None)
        return scale * scaler.inv_scale if scaler is not None else scale

    def _check_param_set(self) -> None:
        """The flat layout is fixed at prepare time; changes torch.optim would pick up silently
        are refused loudly instead of being ignored."""
        if sum(len(g["params"]) for g in self.optimizer.param_groups) != self._n_group_params:
            raise NotImplementedError(
                "param groups changed after prepare (add_param_group): prepare a new optimizer")
        if self._frozen and any(p.requires_grad for p in self._frozen):
            raise NotImplementedError(
                "a parameter that had requires_grad=False when the optimizer was prepared now "
                "requires grad: it has no slot in the flat layout; prepare a new optimizer")

    def _assign_rows(self) -> None:
        """One table row per (param group, lag) among the adopted parameters."""
        keys = sorted({(self._group_of[i], self._lag[i])
                       for i in range(len(self._params)) if self._adopted[i]})
        if len(keys) > 64:
            raise NotImplementedError(
                f"{len(keys)} distinct (param group, first-gradient step) combinations; the "
                "kernels' constant table holds 64 (RK_MAX_GROUPS)")
        index = {k: r for r, k in enumerate(keys)}
        self._rows = keys
        for i in range(len(self._params)):
            self._row_of[i] = index.get((self._group_of[i], self._lag[i]), 0)

    def _step_groups(self) -> list[dict]:
        """This step's hyper-parameters per table row (advances the step counters)."""
        active = set(gi for gi, _ in self._rows)
        for gi in active:
            self._group_step[gi] += 1
        rows = []
        for gi, lag in self._rows:
            g = self.optimizer.param_groups[gi]
            step = self._group_step[gi] - lag            # this row's parameters' own count
            if self._kind == "sgd":
                rows.append(dict(lr=float(g["lr"]), momentum=float(g["momentum"]),
                                 dampening=float(g["dampening"]),
                                 weight_decay=float(g["weight_decay"]),
                                 nesterov=bool(g["nesterov"]),
                                 first_step=step <= 1, skip=False))
            else:
                rows.append(dict(lr=float(g["lr"]), beta1=float(g["betas"][0]),
                                 beta2=float(g["betas"][1]), eps=float(g["eps"]),
                                 weight_decay=float(g["weight_decay"]),
                                 step=max(step, 1),
                                 decoupled=self._kind == "adamw"
                                 or bool(g.get("decoupled_weight_decay", False)),
                                 skip=False))
        return rows

    _TABLE_SLOTS = 32

    def _table_slot(self):
        """Next slot of the pinned ring the step tables are staged in, and the event the
        caller records once the slot's upload has been issued.  A slot is reused only after
        that upload has executed, which also bounds how far the host can run ahead of the
        device."""
        from rocket_b200 import _native as N
        if self._table_ring is None:
            self._table_ring = torch.zeros(self._TABLE_SLOTS, N.RK_STEP_TABLE_BYTES,
                                           dtype=torch.uint8).pin_memory()
            self._table_events = [None] * self._TABLE_SLOTS
            self._table_pos = 0
        i = self._table_pos
        self._table_pos = (i + 1) % self._TABLE_SLOTS
        if self._table_events[i] is not None:
            self._table_events[i].synchronize()
        ev = torch.cuda.Event()
        self._table_events[i] = ev
        return self._table_ring[i], ev

    # ---- gradient adoption -----------------------------------------------------------------------
    def _adopt_grads(self) -> None:
        """Make sure every gradient that exists lives in ``flat_g``.

        Steady state: one pointer compare per parameter.  First sync step (or whenever user
        code replaced ``p.grad``): copy the gradient into its flat view and re-point
        ``p.grad`` there, so later backward passes accumulate in place.
        """
        self._check_param_set()
        changed = False
        for i, p in enumerate(self._params):
            g = p.grad
            if g is None:
                if self._adopted[i]:
                    # user set it to None: treat as zero gradient, keep the slot
                    self._views_g[i].zero_()
                    p.grad = self._views_g[i]
                continue
            if g.data_ptr() != self._grad_ptr[i]:
                if g.is_sparse:
                    raise NotImplementedError("sparse gradients are not supported")
                self._views_g[i].copy_(g)
                p.grad = self._views_g[i]
            if not self._adopted[i]:
                self._adopted[i] = True
                if not self._lag_known[i]:
                    self._lag[i] = self._group_step[self._group_of[i]]   # steps it did not take
                    self._lag_known[i] = True
                changed = True
        if changed:
            self._assign_rows()
            self._plan = None

    def _get_plan(self):
        if self._plan is None:
            specs = [TensorSpec(p=self._views_p[i].reshape(-1), g=self._views_g[i].reshape(-1),
                                m=self._views_m[i].reshape(-1)
                                if self._views_m[i] is not None else None,
                                v=self._views_v[i].reshape(-1)
                                if self._views_v[i] is not None else None,
                                group=self._row_of[i], skip=not self._adopted[i])
                     for i in range(len(self._params))]
            if self._scratch_view is not None:
                self._plan = self.engine.backend.make_plan(specs, scratch=self._scratch_view)
            else:
                self._plan = self.engine.backend.make_plan(specs)
        return self._plan

    # ---- state in torch's own format (checkpoints) ----------------------------------------------------
    def _export_state(self) -> None:
        views_m, views_v = self._views_m, self._views_v
        if self.sharded:     # collective: every rank has to get here
            def views(shard):
                if shard is None:
                    return [None] * len(self._params)
                chunks = self._gather_chunks(shard)
                return [self._assemble(i, chunks) for i in range(len(self._params))]
            views_m, views_v = views(self.shard_m), views(self.shard_v)
        for i, p in enumerate(self._params):
            if not self._adopted[i]:
                continue
            step = self._group_step[self._group_of[i]] - self._lag[i]
            if step <= 0:
                continue
            st = self.optimizer.state[p]
            if self._kind == "sgd":
                st["momentum_buffer"] = views_m[i] if self._needs_m else None
            else:
                st["step"] = torch.tensor(float(step), dtype=torch.float32)
                st["exp_avg"], st["exp_avg_sq"] = views_m[i], views_v[i]

    def state_dict(self) -> dict:
        self._export_state()
        return self.optimizer.state_dict()

    def load_state_dict(self, state_dict: dict) -> None:
        self.optimizer.load_state_dict(state_dict)
        self._group_step = [0] * len(self.optimizer.param_groups)
        self._lag = [0] * len(self._params)
        self._lag_known = [False] * len(self._params)
        with torch.no_grad():
            self._import_state()
            for i, p in enumerate(self._params):
                if self.optimizer.state.get(p):
                    if not self._adopted[i]:
                        self._adopted[i] = True
                        if p.grad is None or p.grad.data_ptr() != self._grad_ptr[i]:
                            self._views_g[i].zero_()
                            p.grad = self._views_g[i]
                elif self._adopted[i]:
                    # adopted here but without state in the checkpoint: it starts counting now
                    self._lag[i] = self._group_step[self._group_of[i]]
                    self._lag_known[i] = True
            self._assign_rows()
            self._plan = None

    # ---- introspection used by bench / tests ---------------------------------------------------------------
    @property
    def numel(self) -> int:
        return sum(p.numel() for p in self._params)
