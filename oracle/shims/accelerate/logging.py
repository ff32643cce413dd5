"""TEST-ONLY: accelerate.logging.get_logger (rocket/core/capsule.py:20,114)."""
import logging as _logging


class _Adapter(_logging.LoggerAdapter):
    def warn(self, msg, *a, **k):
        self.warning(msg, *a, **k)

    def process(self, msg, kwargs):
        kwargs.pop("main_process_only", None)
        kwargs.pop("in_order", None)
        return msg, kwargs


def get_logger(name, log_level=None):
    logger = _logging.getLogger(name)
    if log_level is not None:
        logger.setLevel(log_level.upper())
    return _Adapter(logger, {})
