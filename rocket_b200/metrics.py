"""Device-reduced metrics.

``Accuracy`` is the reference example's metric (reference examples/mnist.py:20-39:
``pr = logits.argmax(1); total += pr.numel(); positive += sum(gt == pr)``) with the
reduction moved onto the device: one ``rk_meter_argmax_count`` launch per eval step adds to
two int64 counters with integer atomics (bit-exact, order independent), nothing is gathered
across ranks per step and nothing synchronises the host.  The counters are summed over
ranks and copied to the host when the value is read -- at RESET, i.e. once per epoch -- and
``attrs.looper.state.accuracy`` carries a non-blocking view for the progress bar.
"""
from __future__ import annotations

from rocket_b200.core.capsule import Attributes
from rocket_b200.core.meter import Metric
from rocket_b200.deferred import DeferredScalar


class Accuracy(Metric):
    """Top-1 accuracy over an eval loop.

    ``target`` / ``prediction`` index the batch (the reference example uses ``batch[1]`` and
    ``batch[2]``).  After each epoch the exact integer counts are available as
    ``.positive`` / ``.total`` and ``history`` holds ``(epoch, positive, total)``.
    """

    device_reduced = True   # tells Meter it need not all-gather for us

    def __init__(self, target=1, prediction=2, tag: str = "accuracy", priority: int = 1000,
                 show_every: int = 16) -> None:
        super().__init__(priority=priority)
        self._target, self._prediction, self._tag = target, prediction, tag
        self._counters = None
        self._show_every = show_every
        self._seen = 0
        self.positive = 0
        self.total = 0
        self.history: list[tuple[int, int, int]] = []

    def launch(self, attrs: Attributes = None):
        if attrs is None or attrs.batch is None:
            return
        engine = self._accelerator
        if self._counters is None:
            self._counters = engine.new_counters(2)
        batch = attrs.batch
        labels, logits = batch[self._target], batch[self._prediction]
        valid = engine.valid_rows(logits.shape[0])     # drop even_batches padding
        if valid < logits.shape[0]:
            labels, logits = labels[:valid], logits[:valid]
        if valid > 0:
            engine.argmax_count(logits, labels, self._counters)
        self._seen += 1
        if attrs.looper is not None and self._seen % self._show_every == 0:
            # progress-bar value: rank-local running ratio, copied without blocking
            snap = engine.read_counters(self._counters, reduce=False)
            attrs.looper.state.accuracy = DeferredScalar(
                snap._slot, snap._event, lambda c: c[0] / max(c[1], 1))

    def reset(self, attrs: Attributes = None):
        if self._counters is None:
            return
        engine = self._accelerator
        pos, tot = engine.read_counters(self._counters, reduce=True).item()
        self.positive, self.total = int(pos), int(tot)
        self.history.append((getattr(self, "_step", 0), self.positive, self.total))
        if attrs is not None and attrs.tracker is not None and self.total:
            attrs.tracker.scalars.append(Attributes(
                step=getattr(self, "_step", 0),
                data={self._tag: self.positive / self.total}))
        self._counters.zero_()
        self._seen = 0

    @property
    def value(self) -> float:
        return self.positive / self.total if self.total else float("nan")
