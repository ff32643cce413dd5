"""rocket_b200 -- a B200-native implementation of rocket's per-iteration training step behind
rocket's unchanged capsule API.

    import rocket_b200 as rocket          # or rocket_b200.install_as_rocket()
    rocket.Launcher([rocket.Looper([rocket.Dataset(...), rocket.Module(net, capsules=[
        rocket.Loss(...), rocket.Optimizer(opt), rocket.Scheduler(sched)])])]).launch()

The capsule classes mirror the reference (dsenushkin/rocket, rocket/core/*.py) signature for
signature; the engine they talk to replaces HuggingFace accelerate and torch.optim's step
with hand-written sm_100a kernels behind a C-ABI (include/rocket_b200.h).
"""
from rocket_b200.attributes import Attributes  # noqa: F401
from rocket_b200.core import (  # noqa: F401
    Capsule, Checkpointer, Dataset, Dispatcher, Events, Launcher, Looper, Loss, Meter, Metric,
    Module, Optimizer, Scheduler, Tracker)

__version__ = "0.1.0"

__all__ = ["Attributes", "Capsule", "Checkpointer", "Dataset", "Dispatcher", "Events",
           "Launcher", "Looper", "Loss", "Meter", "Metric", "Module", "Optimizer",
           "Scheduler", "Tracker", "install_as_rocket"]


def install_as_rocket() -> None:
    """Register this package under the name ``rocket`` so that unmodified user scripts
    (``import rocket``, ``from rocket.core.capsule import Attributes`` ...) run on it."""
    import importlib
    import sys
    if "rocket" in sys.modules and sys.modules["rocket"] is not sys.modules[__name__]:
        raise RuntimeError("another package named 'rocket' is already imported")
    sys.modules["rocket"] = sys.modules[__name__]
    for sub in ("core", "core.capsule", "core.checkpoint", "core.dataset", "core.dispatcher",
                "core.launcher", "core.loop", "core.loss", "core.meter", "core.module",
                "core.optimizer", "core.scheduler", "core.tracker", "utils",
                "utils.collections", "utils.torch"):
        sys.modules[f"rocket.{sub}"] = importlib.import_module(f"rocket_b200.{sub}")
