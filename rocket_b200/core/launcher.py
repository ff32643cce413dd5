"""Launcher: the single entry point; owns the engine and the epoch loop.

API mirror of the reference's rocket/core/launcher.py:41-448.  Where the reference
constructs an ``accelerate.Accelerator`` (launcher.py:185-193) this constructs a
:class:`rocket_b200.engine.Engine` with the same arguments; everything else -- project
directory versioning ``logs/<tag>/vN``, the epoch loop calling ``set / launch / reset`` on
each child, resume, teardown -- follows the reference's behaviour.
"""
from __future__ import annotations

import os
from typing import Callable

from typing_extensions import Self

from rocket_b200.core.capsule import Attributes, Capsule
from rocket_b200.core.dispatcher import Dispatcher


def in_notebook() -> bool:
    try:
        from IPython import get_ipython
        return "IPKernelApp" in get_ipython().config
    except (ImportError, AttributeError):
        return False


def _default_engine_factory(**kwargs):
    from rocket_b200.engine import Engine
    return Engine(**kwargs)


class Launcher(Dispatcher):
    #: how the engine is built; tests substitute a host-only test double here
    engine_factory = staticmethod(_default_engine_factory)

    def __init__(self, capsules: list[Capsule], tag: str = None, logging_dir: str = "./logs",
                 experiment_versioning: bool = True, mixed_precision: str | None = None,
                 gradient_accumulation_steps: int = 1, num_procs: int | None = 1,
                 num_nodes: int | None = 1, num_epochs: int = 1,
                 destroy_process_group_after_launch: bool = True,
                 statefull: bool = False) -> None:
        super().__init__(capsules=capsules)
        self._num_epochs = num_epochs
        self._epoch_idx = 0
        self._statefull = statefull
        self._num_procs = num_procs
        self._num_nodes = num_nodes
        self._mixed_precision = mixed_precision
        self._gradient_accumulation_steps = gradient_accumulation_steps
        self._tag = tag
        self._logging_dir = logging_dir
        self._experiment_versioning = experiment_versioning
        self._project_dir = None
        self._destroy_process_group_after_launch = destroy_process_group_after_launch
        self._resume_from = None
        self._load_capsules = True

    # ---- project directory --------------------------------------------------------------
    def _resolve_project_dir(self) -> None:
        """``logs/<tag>`` or, with versioning, ``logs/<tag>/v<last+1>``; rank 0 decides and
        broadcasts (reference launcher.py:125-150)."""
        if self._tag is None:
            return
        root = os.path.join(self._logging_dir, self._tag)
        if not self._experiment_versioning:
            if os.path.isdir(root):
                raise ValueError(
                    "Project directory already exists and versioning is switched off. "
                    "Change experiment name or enable experiment versioning")
            self._project_dir = root
        else:
            last = -1
            if os.path.isdir(root):
                versions = sorted(int(n[1:]) for n in os.listdir(root)
                                  if n.startswith("v") and n[1:].isdigit())
                if versions:
                    last = versions[-1]
            self._project_dir = os.path.join(root, f"v{last + 1}")
        self._project_dir = self._accelerator.broadcast_object_list(
            [self._project_dir], from_process=0)[0]

    def _create_project_dir(self) -> None:
        if self._tag is None:
            return
        if self._accelerator.is_main_process:
            os.makedirs(self._project_dir, exist_ok=True)
        self._accelerator.broadcast_object_list([None])  # barrier, as in the reference

    # ---- events ----------------------------------------------------------------------------
    def setup(self, attrs: Attributes | None = None) -> None:
        engine = type(self).engine_factory(
            device_placement=True,
            mixed_precision=self._mixed_precision,
            gradient_accumulation_steps=self._gradient_accumulation_steps,
            project_dir=None)
        self.accelerate(engine)
        self._resolve_project_dir()
        engine.project_dir = self._project_dir
        self._create_project_dir()
        self._num_procs = attrs.launcher.num_procs
        self._num_nodes = attrs.launcher.num_nodes
        Dispatcher.setup(self, attrs)

    def notebook(method: Callable) -> Callable:
        """Fill ``attrs.launcher`` defaults; inside Jupyter, spawn the processes
        (reference launcher.py:202-247)."""
        def wrapper(self, attrs: Attributes | None = None) -> None:
            attrs = attrs or Attributes()
            if attrs.launcher is None:
                attrs.launcher = Attributes()
            attrs.launcher.setdefault("num_procs", self._num_procs)
            attrs.launcher.setdefault("num_nodes", self._num_nodes)
            if in_notebook():
                from rocket_b200.engine import notebook_launcher
                notebook_launcher(method, args=(self, attrs),
                                  num_processes=attrs.launcher.num_procs,
                                  num_nodes=attrs.launcher.num_nodes)
            else:
                method(self, attrs=attrs)
        return wrapper

    def set(self, attrs: Attributes | None = None) -> None:
        pass

    def reset(self, attrs: Attributes | None = None) -> None:
        pass

    @notebook
    def launch(self, attrs: Attributes | None = None) -> None:
        Capsule.launch(self, attrs)
        self.setup(attrs)
        self._resume(attrs)
        for epoch in range(self._epoch_idx, self._num_epochs):
            attrs.launcher.epoch_idx = epoch
            self._epoch_idx = epoch
            for capsule in self._capsules:
                capsule.set(attrs)
                capsule.launch(attrs)
                capsule.reset(attrs)
        self.destroy(attrs)

    def destroy(self, attrs: Attributes | None = None) -> None:
        Dispatcher.destroy(self, attrs=attrs)
        del attrs.launcher
        engine = self._accelerator
        engine.end_training()
        self.clear()
        if self._destroy_process_group_after_launch:
            engine.destroy_process_group()

    @staticmethod
    def destroy_process_group():
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            dist.destroy_process_group()

    # ---- resume --------------------------------------------------------------------------------
    def _resume(self, attrs: Attributes) -> None:
        if self._resume_from is None:
            return
        engine = self._accelerator
        if not self._load_capsules:
            saved, engine._custom_objects = engine._custom_objects, []
            try:
                engine.load_state(self._resume_from)
            finally:
                engine._custom_objects = saved
        else:
            try:
                engine.load_state(self._resume_from)
            except Exception:
                raise RuntimeError(
                    "Failed to load state from resume checkpoint. Please check if the "
                    "checkpoint is valid and compatible with the current launcher "
                    "configuration.")
        if (self._num_procs != attrs.launcher.num_procs
                or self._num_nodes != attrs.launcher.num_nodes):
            raise RuntimeError(
                "You need to resume your training in the exact same distributed setup.")

    def resume(self, path: str, load_capsules: bool = True) -> Self:
        self._resume_from = path
        self._load_capsules = load_capsules
        return self

    def state_dict(self) -> dict:
        return dict(epoch_idx=self._epoch_idx, num_procs=self._num_procs,
                    num_nodes=self._num_nodes)

    def load_state_dict(self, state: Attributes) -> None:
        self._epoch_idx = state["epoch_idx"]
        self._num_procs = state["num_procs"]
        self._num_nodes = state["num_nodes"]
