"""Gradient exchange for data parallelism (kernel row K6).

The reference gets this implicitly: ``accelerator.prepare(module)`` wraps the model in torch
DDP (reference rocket/core/module.py:106), whose C++ reducer copies every gradient into
25 MiB buckets (dividing by the world size), all-reduces each bucket as backward fills it,
and copies the result back.  Here gradients already live in the optimizer's flat buffer
(rocket_b200/flat.py), so a bucket is a contiguous slice of it and nothing is copied in or out.
Two implementations share the bucketing and the hook logic:

``P2PGradSync`` (default on one node of NVLink-connected GPUs)
    One hand-written kernel per bucket and rank (csrc/rk_comm.cu) that does the whole
    exchange over NVLink peer memory: every rank reduces the chunks it owns by loading them
    from all peers, scales by 1/world_size, writes the result back into every peer's buffer
    and, in the same pass, produces the sum-of-squares partials of the global gradient norm
    (so clipping needs no separate pass over the gradients).  Launched on a side stream from
    ``post_accumulate_grad`` hooks, it overlaps the rest of backward.  With a sharded
    optimizer step (``Optimizer(sharded=True)``, flat.py) the same kernel runs without its
    write-back half -- a reduce-scatter: only the owner keeps the averaged chunk, updates it
    (csrc/rk_comm.cu, rk_sharded_step_kernel) and writes the new parameters to all ranks.

``GradSync`` (NCCL; used across nodes, on CPU/gloo in tests, or with ROCKET_B200_GRAD_SYNC=nccl)
    In-place ``all_reduce(SUM)`` of each bucket on NCCL's stream; the 1/world_size factor is
    folded into the norm and step kernels' gradient read (``grad_scale``), which is exact for
    power-of-two world sizes and therefore identical to DDP's divide-then-sum.

Both: buckets are issued in reverse parameter order as their last gradient arrives; on
accumulation micro-steps the hooks do nothing (DDP ``no_sync`` semantics); after the exchange
every rank holds bit-identical gradients, so the deterministic norm yields bit-identical
clip coefficients with no further collective.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


class _BucketedSync:
    """Bucketing over the flat gradient buffer + readiness hooks (shared by both backends)."""

    def __init__(self, opt, engine, bucket_cap_mb: float, first_bucket_mb: float = 1.0):
        self.opt = opt
        self.engine = engine
        self.param_ranges: list[tuple[int, int]] = []   # [first_param, last_param] inclusive
        self.buckets: list[tuple[int, int]] = []        # element ranges of flat_g
        self.bucket_of: list[int] = [0] * len(opt._params)
        self._members: list[list[int]] = []
        cap = int(first_bucket_mb * (1 << 20)) // 4
        big = int(float(os.environ.get("ROCKET_B200_BUCKET_MB", bucket_cap_mb)) * (1 << 20)) // 4
        # the bucket that holds parameter 0 completes last, when backward is over: whatever it
        # carries is exchanged with nothing left to hide behind, so it is kept small
        tail = int(float(os.environ.get("ROCKET_B200_TAIL_BUCKET_MB", 1.0)) * (1 << 20)) // 4
        sizes = [p.numel() for p in opt._params]
        tail_first = 0                      # parameters [0, tail_first] form the tail bucket
        acc = 0
        for i, n in enumerate(sizes):
            acc += n
            tail_first = i
            if acc >= tail:
                break
        hi = opt._total
        members: list[int] = []
        size = 0
        # walk parameters backwards: gradients become ready roughly in reverse order
        for i in range(len(opt._params) - 1, -1, -1):
            members.append(i)
            size += sizes[i]
            at_tail = i == tail_first + 1 and tail > 0
            # elements still to come before the tail bucket: do not leave a sliver behind
            left = sum(sizes[tail_first + 1:i]) if i > tail_first + 1 else 0
            close = i == 0 or at_tail or (size >= cap and (left >= big // 8 or i <= tail_first))
            if close:
                lo = opt._offsets[i]
                b = len(self.buckets)
                self.buckets.append((lo, hi))
                self.param_ranges.append((i, members[0]))
                self._members.append(members)
                for m in members:
                    self.bucket_of[m] = b
                hi, members, size, cap = lo, [], 0, big
        self._pending = [len(m) for m in self._members]
        self._launched = [False] * len(self.buckets)
        self._finished = False
        self._handles = [p.register_post_accumulate_grad_hook(self._make_hook(i))
                         for i, p in enumerate(opt._params)]
        self.bytes_reduced = 0

    def _make_hook(self, i: int):
        def hook(param):
            if not self.engine.sync_gradients or not self.opt._adopted[i]:
                return
            if param.grad.data_ptr() != self.opt._grad_ptr[i] or not self._can_launch():
                return                      # reduced later, from finish()
            b = self.bucket_of[i]
            self._pending[b] -= 1
            if self._pending[b] == 0:
                self._launch(b)
        return hook

    def _can_launch(self) -> bool:
        return True

    def _launch(self, b: int) -> None:
        raise NotImplementedError

    def _wait(self) -> None:
        raise NotImplementedError

    #: factor the norm / step kernels still have to apply when they read a gradient
    grad_scale = 1.0
    #: True when finish() leaves the norm's per-chunk partials in the plan's scratch
    fused_norm = False

    def finish(self) -> bool:
        """Launch whatever backward did not, then order the current stream after all of it.
        Idempotent within a step; ``step_done`` re-arms it."""
        if not self._finished:
            # every rank must issue the buckets in the same order: leftovers go by index
            for b in range(len(self.buckets)):
                if not self._launched[b]:
                    self._launch(b)
            self._wait()
            self._finished = True
        return self.fused_norm

    def step_done(self) -> None:
        self._pending = [len(m) for m in self._members]
        self._launched = [False] * len(self.buckets)
        self._finished = False

    def remove(self) -> None:
        for h in self._handles:
            h.remove()
        self._handles = []


class GradSync(_BucketedSync):
    """NCCL (or gloo) in-place SUM all-reduce per bucket."""

    def __init__(self, opt, engine, bucket_cap_mb: float = 16.0):
        super().__init__(opt, engine, bucket_cap_mb)
        self._works: list = []
        self.grad_scale = 1.0 / engine.num_processes

    def _launch(self, b: int) -> None:
        lo, hi = self.buckets[b]
        view = self.opt.flat_g[lo:hi]
        self._works.append(dist.all_reduce(view, op=dist.ReduceOp.SUM, async_op=True))
        self._launched[b] = True
        self.bytes_reduced += (hi - lo) * 4

    def _wait(self) -> None:
        for w in self._works:
            w.wait()
        self._works.clear()


class P2PGradSync(_BucketedSync):
    """Fused all-reduce + 1/W + norm partials over NVLink peer memory (csrc/rk_comm.cu)."""

    fused_norm = True
    grad_scale = 1.0           # the kernel already wrote averaged gradients

    def __init__(self, opt, engine, bucket_cap_mb: float = 16.0):
        super().__init__(opt, engine, bucket_cap_mb)
        # reduce-scatter instead of all-reduce when the optimizer step is sharded: the owner
        # alone needs the averaged chunk, everybody gets the new parameters instead
        self.broadcast = not opt.sharded
        if len(self.buckets) > 256 - int(opt.sharded):     # the sharded step's barrier slot
            raise NotImplementedError("more than 255 gradient buckets")
        self.backend = engine.backend
        self.max_ctas = int(os.environ.get("ROCKET_B200_COMM_CTAS", "0"))
        self.stream = torch.cuda.Stream(device=engine.device)
        self._event = torch.cuda.Event()
        self.profile = False        # bench.py: bracket every exchange kernel with events
        self.timings: list = []     # (bucket bytes, start event, end event) on self.stream
        # graph-captured steps: one external event pair around all buckets of the step, as
        # event-record nodes of the graph (re-recorded by every replay); bench.py samples it
        self.graph_ev: list = []    # (bucket bytes, start, end) per bucket
        self.graph_ms: list = []
        self.graph_bytes = 0
        ptrs, self.multicast_ptr, self._peers = map_arena(opt.arena, engine)
        # in-switch reduction moves ~G bytes per GPU instead of 2(W-1)/W * G: a win from W = 4
        want_mc = os.environ.get("ROCKET_B200_MULTICAST", "auto")
        # (the sharded step's two kernels each move (W-1)/W of the data in ONE direction, which
        # the switch cannot shrink: P2P measured faster there, 4 GPUs: 4.74 vs 5.73 ms per 5e8)
        self.use_multicast = bool(self.multicast_ptr) and (
            want_mc == "1" or (want_mc == "auto" and engine.num_processes >= 4
                               and not opt.sharded))
        if self.use_multicast and self.max_ctas == 0:
            self.max_ctas = 74      # the switch does the work: half the SMs issue enough
        self.step_ctas = int(os.environ.get("ROCKET_B200_SHARDED_STEP_CTAS", "0"))
        self._comm = self.backend.comm_create(
            engine.process_index, engine.num_processes, ptrs, opt._scratch_offset,
            self.multicast_ptr)
        dist.barrier()         # every rank has mapped every arena (and they are zero-filled)

    def _can_launch(self) -> bool:
        # Also while a CUDA graph of the step is being captured: stream capture follows streams,
        # not threads, so a bucket launched from autograd's hook thread onto the side stream --
        # forked from the capturing stream by the event in _launch, joined back in finish() --
        # becomes a node that runs beside the rest of backward in every replay.  (On replays no
        # hook fires: backward is inside the graph.)  ROCKET_B200_GRAPH_OVERLAP=0 issues the
        # buckets from finish() instead, after backward.
        gp = self.engine.graph_pass
        if gp is not None and (gp.mode != "capture" or not self.graph_overlap):
            return False
        return self.opt._plan is not None

    def _chunk_range(self, b: int) -> tuple[int, int]:
        first, last = self.param_ranges[b]
        prefix = self.opt._plan.prefix_norm
        return int(prefix[first]), int(prefix[last + 1])

    def _launch(self, b: int) -> None:
        plan = self.opt._get_plan()
        lo, hi = self._chunk_range(b)
        # the comm stream starts this bucket once the gradients produced so far are final
        self._event.record(torch.cuda.current_stream(self.engine.device))
        self.stream.wait_event(self._event)
        e0, e1 = self.buckets[b]
        gp = self.engine.graph_pass
        capturing = gp is not None and gp.mode == "capture"
        timed = self.graph_timing if capturing else self.profile
        if capturing and not any(self._launched):
            self.graph_ev = []           # first bucket of the step being captured
        if timed:
            # inside a capture: external events = event-record nodes, re-recorded by every replay
            kw = dict(enable_timing=True, external=True) if capturing else dict(enable_timing=True)
            t0, t1 = torch.cuda.Event(**kw), torch.cuda.Event(**kw)
            t0.record(self.stream)
        exchange = self.backend.comm_allreduce_norm if self.broadcast \
            else self.backend.comm_reduce_scatter_norm
        exchange(self._comm, plan, lo, hi, b, self.max_ctas, self.stream.cuda_stream,
                 self.use_multicast)
        if timed:
            t1.record(self.stream)
            if capturing:
                self.graph_ev.append(((e1 - e0) * 4, t0, t1))
            else:
                self.timings.append(((e1 - e0) * 4, t0, t1))
        self._launched[b] = True
        self.bytes_reduced += (e1 - e0) * 4

    def _wait(self) -> None:
        torch.cuda.current_stream(self.engine.device).wait_stream(self.stream)

    #: bench.py sets this to bracket every bucket's kernel inside captured steps with events
    graph_timing = False
    #: launch buckets from the gradient hooks during capture (overlap with backward in replays)
    graph_overlap = os.environ.get("ROCKET_B200_GRAPH_OVERLAP", "1") == "1"

    def sample_graph_events(self) -> None:
        """One sample of the exchange's kernel time per step from the in-graph event pairs (sum
        over the buckets), if the last replay has completed and the pairs are consistent."""
        if not self.graph_ev:
            return
        try:
            total = 0.0
            for _, s, e in self.graph_ev:
                if not e.query():
                    return
                dt = s.elapsed_time(e)
                if not 0.0 < dt < 50.0:          # a torn pair (replay in flight) is negative
                    return
                total += dt
            self.graph_ms.append(total)
            self.graph_bytes = sum(nb for nb, _, _ in self.graph_ev)
        except Exception:
            pass

    def remove(self) -> None:
        super().remove()
        if self._comm is not None:
            torch.cuda.synchronize(self.engine.device)
            dist.barrier()     # nobody unmaps while a peer may still be writing
            self.backend.comm_destroy(self._comm)
            self._comm = None
            self._peers = []
            torch.cuda.ipc_collect()   # drop our mappings of the peers' arenas ...
            dist.barrier()             # ... before any producer may free its arena


def alloc_arena(nbytes: int, device: torch.device) -> torch.Tensor:
    """Zero-filled device memory that peers can map.  Preferred: torch's symmetric-memory
    allocator (CUDA VMM allocations that also get an NVSwitch multicast mapping); otherwise a
    plain allocation shared through CUDA IPC."""
    if os.environ.get("ROCKET_B200_ARENA", "symm") == "symm":
        try:
            import torch.distributed._symmetric_memory as symm_mem
            arena = symm_mem.empty(nbytes, dtype=torch.uint8, device=device)
            arena.zero_()
            arena._rk_symm = True
            return arena
        except Exception:
            pass
    return torch.zeros(nbytes, dtype=torch.uint8, device=device)


def map_arena(arena: torch.Tensor, engine):
    """Collective.  Returns (pointers to every rank's arena as mapped here, multicast pointer
    or 0, objects that keep the mappings alive)."""
    world, rank = engine.num_processes, engine.process_index
    symm = torch.tensor([int(getattr(arena, "_rk_symm", False))], device=engine.device)
    dist.all_reduce(symm, op=dist.ReduceOp.MIN)
    torch.cuda.synchronize(engine.device)
    if symm.item():
        import torch.distributed._symmetric_memory as symm_mem
        hdl = symm_mem.rendezvous(arena, dist.group.WORLD)
        ptrs = [int(p) for p in hdl.buffer_ptrs]
        assert ptrs[rank] == arena.data_ptr(), "symmetric memory handle does not match the arena"
        mc = int(hdl.multicast_ptr) if hdl.has_multicast_support else 0
        return ptrs, mc, [hdl]
    # CUDA IPC through torch's tensor-sharing plumbing; the handle is opened on THIS rank's
    # device so that peer access towards the owner is enabled
    from torch.multiprocessing import reductions
    _, meta = reductions.reduce_tensor(arena)
    metas = [None] * world
    dist.all_gather_object(metas, meta)
    peers = []
    for q, m in enumerate(metas):
        if q == rank:
            peers.append(arena)
            continue
        m = list(m)
        m[6] = engine.device.index
        peers.append(reductions.rebuild_cuda_tensor(*m))
    return [t.data_ptr() for t in peers], 0, peers


def p2p_possible(engine) -> bool:
    """Fused NVLink path: CUDA, one node, 2/4/8 ranks, every pair of devices peer-accessible."""
    if engine.device.type != "cuda" or engine.num_processes not in (2, 4, 8):
        return False
    if os.environ.get("ROCKET_B200_GRAD_SYNC", "p2p") != "p2p":
        return False
    if int(os.environ.get("LOCAL_WORLD_SIZE", engine.num_processes)) != engine.num_processes:
        return False
    me = engine.device.index
    return all(q == me or torch.cuda.can_device_access_peer(me, q)
               for q in range(engine.num_processes))


def broadcast_module_state(module: torch.nn.Module, src: int = 0) -> None:
    """Start every rank from rank 0's parameters and buffers (what DDP does at wrap time)."""
    with torch.no_grad():
        for t in list(module.parameters()) + list(module.buffers()):
            dist.broadcast(t.data, src=src)


class BufferSync:
    """DDP's ``broadcast_buffers=True``: rank 0's module buffers (BatchNorm statistics)
    overwrite everyone's before a training forward.  All buffers, whatever their dtype, are
    re-homed into ONE flat byte buffer (16-byte aligned segments), so this is a single small
    collective per forward without packing copies (it was one per dtype: float statistics and
    the int64 batch counters)."""

    def __init__(self, module: torch.nn.Module):
        bufs = [b for b in module.buffers() if b.numel() > 0]
        self.flat = None
        if bufs:
            offs, total = [], 0
            for b in bufs:
                offs.append(total)
                total += -(-(b.numel() * b.element_size()) // 16) * 16
            self.flat = torch.zeros(total, dtype=torch.uint8, device=bufs[0].device)
            with torch.no_grad():
                for b, off in zip(bufs, offs):
                    nbytes = b.numel() * b.element_size()
                    view = self.flat[off:off + nbytes].view(b.dtype).view(b.shape)
                    view.copy_(b)
                    b.data = view
        self.flats = [self.flat] if self.flat is not None else []
        self._handle = module.register_forward_pre_hook(self._hook)

    def _hook(self, module, args):
        if self.flat is not None and torch.is_grad_enabled() and module.training:
            dist.broadcast(self.flat, src=0)
        return None

    def remove(self) -> None:
        self._handle.remove()
