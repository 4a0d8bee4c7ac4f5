"""Key-value event logging.

The reference logs through ``structlog.get_logger`` (reference src/trainer.py:6,19,25)
and treats structlog as a hard dependency.  Here structlog is optional: when it is
missing the same ``logger.info("event", key=value)`` calls are rendered by a small
stdlib adapter, so importing the trainer never fails on a bare box.
"""
from __future__ import annotations

import logging
import os
import sys

__all__ = ["get_logger", "set_verbosity"]

_LEVEL = os.environ.get("MLT_LOG_LEVEL", "INFO").upper()


class _KVLogger:
    """Minimal structlog look-alike: positional event + keyword context."""

    def __init__(self, name: str):
        self._log = logging.getLogger(name)
        if not self._log.handlers:
            handler = logging.StreamHandler(sys.stderr)
            handler.setFormatter(logging.Formatter("%(asctime)s [%(levelname)-7s] %(message)s"))
            self._log.addHandler(handler)
            self._log.propagate = False
        self._log.setLevel(getattr(logging, _LEVEL, logging.INFO))

    @staticmethod
    def _render(event, kv):
        if not kv:
            return str(event)
        tail = " ".join(f"{k}={v!r}" for k, v in kv.items())
        return f"{event} {tail}"

    def debug(self, event, **kv):
        self._log.debug(self._render(event, kv))

    def info(self, event, **kv):
        self._log.info(self._render(event, kv))

    def warning(self, event, **kv):
        self._log.warning(self._render(event, kv))

    warn = warning

    def error(self, event, **kv):
        self._log.error(self._render(event, kv))

    def setLevel(self, level):  # noqa: N802 (stdlib spelling)
        self._log.setLevel(level)


_loggers: dict[str, object] = {}


def get_logger(name: str = "ml_trainer_b200"):
    if name in _loggers:
        return _loggers[name]
    try:
        import structlog  # type: ignore

        lg = structlog.get_logger(name)
    except Exception:
        lg = _KVLogger(name)
    _loggers[name] = lg
    return lg


def set_verbosity(level: str | int) -> None:
    """Set the level of every stdlib-backed logger handed out so far."""
    if isinstance(level, str):
        level = getattr(logging, level.upper())
    for lg in _loggers.values():
        if isinstance(lg, _KVLogger):
            lg.setLevel(level)
