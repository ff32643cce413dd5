"""Checkpointer capsule (API mirror of the reference's rocket/core/checkpoint.py:20-169).

Cold path: every ``save_every`` launches the main process asks the engine to write its
state (models, optimizers -- with the flat moment buffers exported per parameter in torch's
own state_dict format --, schedulers, RNG and the registered capsules).
"""
from __future__ import annotations

import os

from rocket_b200.core.capsule import Attributes, Capsule


class Checkpointer(Capsule):
    def __init__(self, output_dir_format: str = "weights/{:03d}", save_every: int | None = None,
                 overwrite: bool = True, statefull: bool = True, priority: int = 100) -> None:
        super().__init__(statefull=statefull, priority=priority)
        self._save_every = save_every or -1
        self._output_dir_format = output_dir_format
        self._overwrite = overwrite
        self._iter_idx = 0

    def setup(self, attrs: Attributes | None = None) -> None:
        # Unlike the reference (checkpoint.py:74-81, which skips Capsule.setup and then trips
        # the LIFO check in destroy -- SURVEY.md Appendix C) the capsule registers itself.
        Capsule.setup(self, attrs=attrs)
        if self._accelerator.project_dir is None:
            raise ValueError(
                "Checkpointer can be used only when project directory is configured. "
                "Current project directory is None. This might be due to the `tag=None` "
                "set when creating `rocket.Launcher`. Set `tag` parameter of "
                "`rocket.Launcher` to a specific experiment name")

    def launch(self, attrs: Attributes | None = None) -> None:
        Capsule.launch(self, attrs=attrs)
        if self._save_every < 0:
            return
        if (self._iter_idx + 1) % self._save_every == 0:
            output_dir = os.path.join(self._accelerator.project_dir,
                                      self._output_dir_format.format(self._iter_idx))
            if self._accelerator.is_main_process:
                if not self._overwrite and os.path.exists(output_dir):
                    raise RuntimeError(
                        f"{type(self).__name__}: Cannot overwrite existing directory. "
                        f"'overwrite' is set to False and '{output_dir}' already exists.")
            # every rank calls save_state (it may synchronise); only rank 0 writes
            self._accelerator.save_state(output_dir=output_dir)
            if self._accelerator.is_main_process:
                self._logger.info(f"{type(self).__name__}: saved {output_dir}")
        self._iter_idx += 1

    def state_dict(self) -> dict:
        return dict(iter_idx=self._iter_idx + 1)

    def load_state_dict(self, state: dict) -> None:
        self._iter_idx = state["iter_idx"]
