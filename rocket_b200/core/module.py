"""Module capsule: forward of the user's network, then its children (Loss, Optimizer, ...).

API mirror of the reference's rocket/core/module.py:24-219.  The forward/backward of the
user's ``nn.Module`` stays in PyTorch; the engine owns only the two contexts entered around
it -- ``autocast`` and ``accumulate`` (gradient-accumulation phase; on multi-GPU it also
arms or disarms the bucketed gradient all-reduce the way DDP's ``no_sync`` would).

One additive keyword, ``graph`` (default ``False`` = the reference's behaviour).  With
``graph=True`` the training step of this capsule -- forward, objective, backward, gradient
norm, fused optimizer step -- is captured once per input signature into a CUDA graph after
a few eager iterations and replayed afterwards with one launch, which removes the eager
dispatch cost that bounds small models (a ResNet-18 step is 2.3 ms of GPU work behind
4.3 ms of Python).  Requirements: static shapes and control flow in the network and the
objective, and children that are engine capsules (Loss / Optimizer / Scheduler) or do host
work only.  Per-step scalars (learning rate, Adam bias corrections) reach the graph through a
device table refreshed before every replay; loss values are read back after it.
"""
from __future__ import annotations

import sys
from contextlib import contextmanager

import torch

from rocket_b200.core.capsule import Attributes, Capsule
from rocket_b200.core.dispatcher import Dispatcher
from rocket_b200.utils.torch import torch_move


class Module(Dispatcher):
    GRAPH_WARMUP = 3     # eager iterations per input signature before capture

    def __init__(self, module: torch.nn.Module, capsules: list[Capsule] = [],
                 priority: int = 1000, graph: bool = False) -> None:
        super().__init__(capsules=capsules, priority=priority)
        self._module = module
        self._graph = graph
        self._graphs: dict = {}      # signature -> _CapturedStep
        self._warm: dict = {}        # signature -> eager iterations seen
        self.graphs_captured = 0     # how many step graphs this capsule has captured

    def setup(self, attrs: Attributes | None = None) -> None:
        self.check_accelerator()
        found = False
        for model in self._accelerator._models:
            if model is not self._module:
                continue
            if found:
                raise RuntimeError(
                    f"{type(self).__name__}: same module has been registered twice.")
            found = True
            self._module = model
        if not found:
            self._module = torch_move(self._module, self._accelerator.device)
            self._module = self._accelerator.prepare(self._module)
        Dispatcher.setup(self, attrs)

    def launch(self, attrs: Attributes | None = None) -> None:
        if attrs is None or attrs.batch is None:
            return
        # grad mode was set by the Looper; it decides train vs eval (reference :130-135)
        if torch.is_grad_enabled():
            self._module.train()
        else:
            self._module.eval()
        with self.runner():
            if self._graph and torch.is_grad_enabled() and self._graphable():
                self._launch_graphed(attrs)
            else:
                attrs.batch = self._module.forward(attrs.batch)
                Dispatcher.launch(self, attrs=attrs)

    # ---- CUDA-graph path (graph=True) ---------------------------------------------------
    def _graphable(self) -> bool:
        engine = self._accelerator
        if engine.device.type != "cuda" or not hasattr(engine, "graph_pass"):
            return False
        if getattr(engine, "scaler", None) is not None:
            return False     # fp16: the overflow decision is a host read between norm and step
        # multi-GPU: only the hand-written NVLink exchange is replayable (device-side epochs),
        # with the replicated or the sharded optimizer step
        return engine.num_processes == 1 or (
            bool(engine.grad_sync_kinds)
            and all(k.startswith("P2PGradSync") for k in engine.grad_sync_kinds))

    def _launch_graphed(self, attrs: Attributes) -> None:
        from rocket_b200.engine import GraphPass
        engine = self._accelerator
        leaves: list[torch.Tensor] = []
        template = _split(attrs.batch, leaves)
        key = (engine.sync_gradients,
               tuple((tuple(t.shape), t.dtype) for t in leaves))
        step = self._graphs.get(key)
        if step is None and self._warm.get(key, 0) < self.GRAPH_WARMUP:
            self._warm[key] = self._warm.get(key, 0) + 1
            attrs.batch = self._module.forward(attrs.batch)
            Dispatcher.launch(self, attrs=attrs)
            return
        if step is None:
            step = _CapturedStep([torch.empty_like(t) for t in leaves])
            for dst, src in zip(step.inputs, leaves):
                dst.copy_(src)
            engine.graph_pass = gp = GraphPass("capture", step.slots)
            launches_before = engine.backend.launch_count()
            try:
                torch.cuda.synchronize(engine.device)
                with torch.cuda.graph(step.graph, capture_error_mode="thread_local"):
                    attrs.batch = self._module.forward(_join(template, step.inputs))
                    step.outputs = attrs.batch
                    Dispatcher.launch(self, attrs=attrs)
            finally:
                engine.graph_pass = None
            # how many of the engine's kernels one replay of this graph runs
            step.kernel_launches = engine.backend.launch_count() - launches_before
            self._graphs[key] = step
            self.graphs_captured += 1
        else:
            for dst, src in zip(step.inputs, leaves):
                dst.copy_(src)
            attrs.batch = step.outputs
            engine.graph_pass = gp = GraphPass("replay", step.slots)
            try:
                Dispatcher.launch(self, attrs=attrs)      # host-side bookkeeping only
            finally:
                engine.graph_pass = None
        for action in gp.pre:
            action()
        step.graph.replay()
        engine.replayed_kernel_launches += step.kernel_launches
        for action in gp.post:
            action()

    def destroy(self, attrs: Attributes | None = None) -> None:
        if self._graphs:
            # Captured steps hold NCCL / NVLink-exchange kernels and pool memory: release them
            # while the process group and the peers' mappings are still alive (a graph that
            # outlives them hangs the interpreter at exit).
            if self._accelerator.device.type == "cuda":
                torch.cuda.synchronize(self._accelerator.device)
            self._graphs.clear()
            self._warm.clear()
        models = self._accelerator._models
        for i, model in enumerate(models):
            if model is self._module:
                models.pop(i)
                break
        Dispatcher.destroy(self, attrs=attrs)

    @contextmanager
    def runner(self):
        """autocast + accumulate, entered in that order, both always exited
        (reference module.py:175-219)."""
        autocast_ctx = self._accelerator.autocast()
        accumulate_ctx = self._accelerator.accumulate(self._module)
        try:
            autocast_ctx.__enter__()
            accumulate_ctx.__enter__()
            yield
        finally:
            autocast_ctx.__exit__(*sys.exc_info())
            accumulate_ctx.__exit__(*sys.exc_info())


class _CapturedStep:
    """A captured training step: the graph, its static inputs / outputs, and the static
    tensors its participants re-read after every replay."""

    def __init__(self, inputs: list[torch.Tensor]):
        self.graph = torch.cuda.CUDAGraph()
        self.inputs = inputs
        self.outputs = None
        self.slots: dict = {}
        self.kernel_launches = 0


class _Leaf:
    __slots__ = ("i",)

    def __init__(self, i: int):
        self.i = i


def _split(batch, leaves: list):
    """Replace device tensors in a nested batch by placeholders, collecting the tensors."""
    if isinstance(batch, torch.Tensor):
        if batch.device.type != "cuda":
            return batch
        leaves.append(batch)
        return _Leaf(len(leaves) - 1)
    if isinstance(batch, dict):
        return {k: _split(v, leaves) for k, v in batch.items()}
    if isinstance(batch, (list, tuple)):
        return type(batch)(_split(v, leaves) for v in batch)
    return batch


def _join(template, leaves: list):
    if isinstance(template, _Leaf):
        return leaves[template.i]
    if isinstance(template, dict):
        return {k: _join(v, leaves) for k, v in template.items()}
    if isinstance(template, (list, tuple)):
        return type(template)(_join(v, leaves) for v in template)
    return template
