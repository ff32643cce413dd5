"""Shared workload definitions for parity tests, golden generation, smoke and bench.

``build_mnist_graph(rk, ...)`` builds the capsule graph of the reference's example
(reference examples/mnist.py:81-105) with its four stale call sites corrected (SURVEY.md
section 0.4) and MNIST replaced by synthetic 1x28x28 data.  ``rk`` is either the reference
package (imported from /root/reference through the test shims) or ``rocket_b200``: the
script is the same for both, which is what "drop-in" means here.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F
from torch import nn


class LeNet(nn.Module):
    """The network of reference examples/mnist.py:42-74 (takes the batch, returns
    ``(images, labels, logits)``)."""

    def __init__(self):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 6, 5, padding=2)
        self.conv2 = nn.Conv2d(6, 16, 5)
        self.fc1 = nn.Linear(16 * 5 * 5, 120)
        self.fc2 = nn.Linear(120, 84)
        self.fc3 = nn.Linear(84, 10)

    def forward(self, batch):
        x = F.max_pool2d(F.relu(self.conv1(batch[0])), (2, 2))
        x = F.max_pool2d(F.relu(self.conv2(x)), (2, 2))
        x = x.flatten(1)
        x = F.relu(self.fc1(x))
        x = F.relu(self.fc2(x))
        return batch[0], batch[1], self.fc3(x)


class BatchModel(nn.Module):
    """Adapts a plain ``images -> logits`` network to rocket's batch-in / batch-out style."""

    def __init__(self, net: nn.Module):
        super().__init__()
        self.net = net

    def forward(self, batch):
        return batch[0], batch[1], self.net(batch[0])


class CrossEntropy(nn.Module):
    def forward(self, batch):
        return F.cross_entropy(batch[2], batch[1])


def synthetic_classification(n: int, shape, classes: int, seed: int, input_scale: float = 1.0):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn((n, *shape), generator=g)
    if input_scale != 1.0:     # large inputs: fp16 weight gradients overflow until the loss
        x = x * input_scale    # scale has backed off (exercises GradScaler's skipped steps)
    y = torch.randint(0, classes, (n,), generator=g)
    return torch.utils.data.TensorDataset(x, y)


def make_user_accuracy(rk):
    """The example's Accuracy metric (reference examples/mnist.py:20-39), corrected: it is
    user code written against the public Metric API, so it runs on either package."""

    class Accuracy(rk.Metric):
        def __init__(self, priority: int = 1000):
            super().__init__(priority=priority)
            self.positive = 0
            self.total = 0
            self.history = []

        def launch(self, attrs=None):
            gt, pr = attrs.batch[1], attrs.batch[2].argmax(1)
            self.total += pr.numel()
            self.positive += int(torch.sum(gt == pr))

        def reset(self, attrs=None):
            self.history.append((self._step, self.positive, self.total))
            self.positive = 0
            self.total = 0

    return Accuracy


def make_user_clip(rk):
    """Gradient clipping the way a user of the reference would add it: a capsule between
    Loss (1100) and Optimizer (1000) that calls the accelerator on sync steps."""

    class Clip(rk.Capsule):
        def __init__(self, module, max_norm, priority=1050):
            super().__init__(priority=priority)
            self._module, self._max_norm = module, max_norm

        def launch(self, attrs=None):
            if torch.is_grad_enabled() and self._accelerator.sync_gradients:
                self._accelerator.clip_grad_norm_(self._module.parameters(), self._max_norm)

    return Clip


def build_mnist_graph(rk, *, seed=0, n_train=64 * 24, n_val=64 * 5 + 13, batch_size=64,
                      optimizer="adamw", max_norm=None, accum=2, epochs=2, sched_step=5,
                      tracker_backend=None, accuracy_cls=None, clip_via="capsule",
                      lr=1e-3, net=None, opt_kwargs=None, graph=False, mixed_precision=None,
                      input_scale=1.0):
    """Returns (launcher, handles) -- handles: net, opt, sched, accuracy, tracker_backend."""
    torch.manual_seed(seed)
    net = net if net is not None else LeNet()
    opt_kwargs = dict(opt_kwargs or {})
    if optimizer == "adamw":
        opt = torch.optim.AdamW(net.parameters(), lr=lr, foreach=False, **opt_kwargs)
    elif optimizer == "adam":
        opt = torch.optim.Adam(net.parameters(), lr=lr, weight_decay=1e-3, foreach=False,
                               **opt_kwargs)
    elif optimizer == "sgd":
        opt = torch.optim.SGD(net.parameters(), lr=lr * 50, momentum=0.9, weight_decay=1e-4,
                              nesterov=True, foreach=False, **opt_kwargs)
    else:
        raise ValueError(optimizer)
    sched = torch.optim.lr_scheduler.StepLR(opt, sched_step, gamma=0.5)
    train = synthetic_classification(n_train, (1, 28, 28), 10, seed + 1, input_scale)
    val = synthetic_classification(n_val, (1, 28, 28), 10, seed + 2, input_scale)
    accuracy = (accuracy_cls or make_user_accuracy(rk))()
    inner = [rk.Loss(objective=CrossEntropy())]
    if max_norm is not None and clip_via == "capsule":
        inner.append(make_user_clip(rk)(net, max_norm))
    if max_norm is not None and clip_via == "kwarg":
        inner.append(rk.Optimizer(opt, max_grad_norm=max_norm))
    else:
        inner.append(rk.Optimizer(opt))
    inner.append(rk.Scheduler(sched))
    module_kw = dict(graph=True) if graph else {}          # additive keyword of rocket_b200
    train_caps = [rk.Dataset(train, batch_size=batch_size, shuffle=False),
                  rk.Module(net, capsules=inner, **module_kw)]
    val_caps = [rk.Dataset(val, batch_size=batch_size, shuffle=False),
                rk.Module(net, capsules=[]),
                rk.Meter([accuracy], keys=[1, 2])]
    if tracker_backend is not None:
        train_caps.append(rk.Tracker(backend=tracker_backend))
    launcher = rk.Launcher(
        [rk.Looper(train_caps, tag="train"),
         rk.Looper(val_caps, tag="val", grad_enabled=False)],
        num_epochs=epochs, gradient_accumulation_steps=accum,
        destroy_process_group_after_launch=False,
        **(dict(mixed_precision=mixed_precision) if mixed_precision else {}))
    return launcher, dict(net=net, opt=opt, sched=sched, accuracy=accuracy,
                          tracker=tracker_backend)
