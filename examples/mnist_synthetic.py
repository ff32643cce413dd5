"""The reference's examples/mnist.py, runnable.

Same capsule graph (LeNet, AdamW, StepLR(100), gradient accumulation 2, 4 epochs, train +
eval Looper with an accuracy Meter); the four call sites of the reference example that no
longer match its own API are corrected and MNIST is replaced by
synthetic 1x28x28 data because there is no network.  Runs unchanged on the reference
(`import rocket`) and on this package:

    python examples/mnist_synthetic.py                 # one B200
    torchrun --nproc-per-node 8 examples/mnist_synthetic.py
"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "tests"))

import rocket_b200 as rocket  # noqa: E402   (or: import rocket)
from rocket_b200.metrics import Accuracy  # noqa: E402
from workloads import CrossEntropy, LeNet, synthetic_classification  # noqa: E402

net = LeNet()
opt = torch.optim.AdamW(net.parameters())
sched = torch.optim.lr_scheduler.StepLR(opt, 100)
train = synthetic_classification(1024 * 8, (1, 28, 28), 10, seed=1)
val = synthetic_classification(1024 * 2, (1, 28, 28), 10, seed=2)
accuracy = Accuracy()

launcher = rocket.Launcher([
    rocket.Looper([
        rocket.Dataset(train, batch_size=1024, shuffle=True),
        rocket.Module(net, capsules=[
            rocket.Loss(objective=CrossEntropy()),
            rocket.Optimizer(opt),
            rocket.Scheduler(sched),
        ]),
    ], tag="train"),
    rocket.Looper([
        rocket.Dataset(val, batch_size=1024),
        rocket.Module(net),
        rocket.Meter([accuracy], keys=[1, 2]),
    ], tag="val", grad_enabled=False),
], statefull=False, num_epochs=4, gradient_accumulation_steps=2)

if __name__ == "__main__":
    launcher.launch()
    print("eval (epoch, correct, total):", accuracy.history)
