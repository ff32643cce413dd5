"""TEST-ONLY: accelerate.tracking.GeneralTracker (rocket/core/tracker.py:17,86-105)."""


class GeneralTracker:
    main_process_only = True

    def __init__(self, _blank=False):
        self._blank = _blank

    def log(self, values, step=None, **kwargs):
        pass

    def log_images(self, values, step=None, **kwargs):
        pass

    def finish(self):
        pass


class MemoryTracker(GeneralTracker):
    """Keeps what was logged, so tests can read it back."""

    def __init__(self, name):
        super().__init__()
        self.name = name
        self.scalars = []
        self.images = []

    def log(self, values, step=None, **kwargs):
        self.scalars.append((step, dict(values)))

    def log_images(self, values, step=None, **kwargs):
        self.images.append((step, values))
