from rocket_b200.core.capsule import Attributes, Capsule, Events  # noqa: F401
from rocket_b200.core.checkpoint import Checkpointer  # noqa: F401
from rocket_b200.core.dataset import Dataset  # noqa: F401
from rocket_b200.core.dispatcher import Dispatcher  # noqa: F401
from rocket_b200.core.launcher import Launcher  # noqa: F401
from rocket_b200.core.loop import Looper  # noqa: F401
from rocket_b200.core.loss import Loss  # noqa: F401
from rocket_b200.core.meter import Meter, Metric  # noqa: F401
from rocket_b200.core.module import Module  # noqa: F401
from rocket_b200.core.optimizer import Optimizer  # noqa: F401
from rocket_b200.core.scheduler import Scheduler  # noqa: F401
from rocket_b200.core.tracker import Tracker  # noqa: F401
