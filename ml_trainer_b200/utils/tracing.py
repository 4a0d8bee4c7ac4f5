"""Tracing hooks (the reference has none beyond tqdm's batch/s).

* NVTX ranges around the phases of a step (``h2d``, ``forward+loss``, ``backward``, ``reduce+update``, ``eval``), visible in
  Nsight Systems / ``ncu --nvtx``.  Off by default; ``MLT_NVTX=1`` or ``Trainer(trace=True)`` turns them on.  Under CUDA
  graph replay the host only sees one range per step (the phases were recorded at capture time).
* ``PhaseTimer``: CUDA-event timing of the same phases for eager steps, summarised per epoch - device time, no profiler.
"""
from __future__ import annotations

import os
from collections import defaultdict
from contextlib import contextmanager
from typing import Dict

import torch

_state = {"nvtx": os.environ.get("MLT_NVTX", "0") == "1", "timer": None}


def enable(on: bool = True, timer: "PhaseTimer | None" = None) -> None:
    _state["nvtx"] = bool(on)
    _state["timer"] = timer


def enabled() -> bool:
    return _state["nvtx"]


@contextmanager
def phase(name: str):
    """``with phase("backward"): ...`` - an NVTX range and, when a PhaseTimer is installed, a CUDA-event pair."""
    nvtx = _state["nvtx"] and torch.cuda.is_available()
    timer = _state["timer"]
    if nvtx:
        torch.cuda.nvtx.range_push(name)
    tok = timer.start(name) if timer is not None else None
    try:
        yield
    finally:
        if tok is not None:
            timer.stop(tok)
        if nvtx:
            torch.cuda.nvtx.range_pop()


class PhaseTimer:
    """Accumulates device time per phase from CUDA events (a no-op on CPU and while a CUDA graph is being captured)."""

    def __init__(self):
        self._pending = []
        self.total_ms: Dict[str, float] = defaultdict(float)
        self.count: Dict[str, int] = defaultdict(int)

    def _live(self) -> bool:
        return torch.cuda.is_available() and not torch.cuda.is_current_stream_capturing()

    def start(self, name: str):
        if not self._live():
            return None
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        return (name, a, b)

    def stop(self, tok) -> None:
        name, a, b = tok
        b.record()
        self._pending.append((name, a, b))

    def summary(self, reset: bool = True) -> Dict[str, float]:
        """Per-phase mean milliseconds since the last call (synchronises once)."""
        if self._pending:
            torch.cuda.synchronize()
            for name, a, b in self._pending:
                self.total_ms[name] += a.elapsed_time(b)
                self.count[name] += 1
            self._pending.clear()
        out = {k: self.total_ms[k] / max(self.count[k], 1) for k in self.total_ms}
        if reset:
            self.total_ms.clear()
            self.count.clear()
        return out
