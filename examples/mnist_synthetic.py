"""The reference's examples/mnist.py, runnable -- written against `import rocket` only.

Same capsule graph as the reference example (LeNet, AdamW, StepLR(100), gradient accumulation
2, batch 1024, 4 epochs, a Checkpointer), plus the eval Looper the example's own `Accuracy`
metric was written for.  Everything below the imports is user code in the reference's style:
the network, the objective and the user-written `Accuracy(rocket.Metric)` are defined here,
not taken from this repository.  Differences from the reference file, all forced:

* its four call sites that no longer match the reference's own API are corrected (SURVEY.md
  section 0.4: `Checkpointer(output_dir_format=...)`, no `accelerator=` argument to
  `Launcher`, `Metric.__init__(priority)`, `Attributes` imported);
* MNIST is replaced by synthetic 1x28x28 data (there is no network here);
* it calls `launcher.launch()` (the reference file only prints the launcher).

Which package answers `import rocket` is decided outside the script: the reference if it is
installed, otherwise this repository registered under that name (`install_as_rocket()`).

    python examples/mnist_synthetic.py                       # one B200
    torchrun --nproc-per-node 8 examples/mnist_synthetic.py
    python examples/mnist_synthetic.py --epochs 1 --samples 4096 --json out.json
"""
import argparse
import json
import sys
from pathlib import Path

import numpy as np
import torch
import torch.nn.functional as F
from torch import nn

try:
    import rocket                                   # the reference package, when installed
except ImportError:
    sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
    import rocket_b200
    rocket_b200.install_as_rocket()                 # this repository, under the same name
    import rocket
from rocket.core.capsule import Attributes          # noqa: E402


class Accuracy(rocket.Metric):
    """The example's metric (reference examples/mnist.py:20-39): user code on the public
    Metric API.  It goes through Meter's gather and reads a ratio back every eval step, exactly
    as written there; `history` is added so that a run can report its counts."""

    def __init__(self, priority: int = 1000) -> None:
        super().__init__(priority)
        self.positive = 0.0
        self.total = 0.0
        self.history = []

    def launch(self, attrs: Attributes = None):
        batch = attrs.batch
        gt, pr = batch[1], batch[2]
        pr = pr.argmax(1)

        self.total += pr.numel()
        self.positive += torch.sum(gt == pr)

        attrs.looper.state.accuracy = (self.positive / self.total).item()

    def reset(self, attrs: Attributes = None):
        self.history.append((int(self.positive), int(self.total)))
        self.total = 0.0
        self.positive = 0.0


class LeNet(nn.Module):
    def __init__(self):
        super(LeNet, self).__init__()
        self.conv1 = nn.Conv2d(1, 6, 5, padding=2)
        self.conv2 = nn.Conv2d(6, 16, 5)
        self.fc1 = nn.Linear(16 * 5 * 5, 120)
        self.fc2 = nn.Linear(120, 84)
        self.fc3 = nn.Linear(84, 10)

    def forward(self, x):
        inp = x
        x = F.max_pool2d(F.relu(self.conv1(x[0])), (2, 2))
        x = F.max_pool2d(F.relu(self.conv2(x)), (2, 2))
        x = x.view(-1, self.num_flat_features(x))
        x = F.relu(self.fc1(x))
        x = F.relu(self.fc2(x))
        x = self.fc3(x)
        return inp[0], inp[1], x

    def num_flat_features(self, x):
        size = x.size()[1:]
        return int(np.prod(size))


class CrossEntropy(torch.nn.Module):
    def forward(self, batch):
        return torch.nn.CrossEntropyLoss()(batch[2], batch[1])


def synthetic_mnist(n: int, seed: int):
    g = torch.Generator().manual_seed(seed)
    return torch.utils.data.TensorDataset(torch.randn((n, 1, 28, 28), generator=g),
                                          torch.randint(0, 10, (n,), generator=g))


def build(epochs: int = 4, samples: int = 1024 * 8, batch_size: int = 1024, tag=None,
          accuracy=None):
    torch.manual_seed(0)
    net = LeNet()
    opt = torch.optim.AdamW(net.parameters())
    sched = torch.optim.lr_scheduler.StepLR(opt, 100)
    loss = CrossEntropy()
    accuracy = accuracy if accuracy is not None else Accuracy()
    train_capsules = [
        rocket.Dataset(synthetic_mnist(samples, 1), batch_size=batch_size),
        rocket.Module(net, capsules=[
            rocket.Loss(objective=loss),
            rocket.Optimizer(opt),
            rocket.Scheduler(sched)
        ]),
    ]
    if tag is not None:                               # a project directory exists only with a tag
        train_capsules.append(rocket.Checkpointer(output_dir_format="chkpt/{:03d}",
                                                  overwrite=True, save_every=50))
    launcher = rocket.Launcher([
            rocket.Looper(train_capsules, tag="train"),
            rocket.Looper([
                rocket.Dataset(synthetic_mnist(samples // 4, 2), batch_size=batch_size),
                rocket.Module(net),
                rocket.Meter([accuracy], keys=[1, 2]),
            ], tag="val", grad_enabled=False),
        ],
        tag=tag,
        statefull=False,
        num_epochs=epochs,
        gradient_accumulation_steps=2,
    )
    return launcher, net, accuracy


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--epochs", type=int, default=4)
    ap.add_argument("--samples", type=int, default=1024 * 8)
    ap.add_argument("--batch-size", type=int, default=1024)
    ap.add_argument("--tag", default=None, help="experiment name: enables the Checkpointer")
    ap.add_argument("--deterministic", action="store_true",
                    help="cuDNN deterministic algorithms (for run-to-run comparisons)")
    ap.add_argument("--json", default=None, help="write the eval counts and a parameter checksum here")
    args = ap.parse_args()
    if args.deterministic:
        torch.backends.cudnn.deterministic = True
        torch.backends.cudnn.benchmark = False
        torch.backends.cudnn.allow_tf32 = False
    launcher, net, accuracy = build(args.epochs, args.samples, args.batch_size, args.tag)
    print(launcher)
    launcher.launch()
    print("package:", rocket.__name__, "| eval (correct, total) per epoch:", accuracy.history)
    if args.json:
        checksum = float(sum(p.detach().double().sum() for p in net.parameters()))
        Path(args.json).write_text(json.dumps(
            {"package": rocket.__name__, "history": accuracy.history, "checksum": checksum}))
