"""Tracing helpers (the reference has none beyond a wall-clock delta, distribute_training.py:218,:230).

* :func:`nvtx_range` — NVTX ranges around the host-side phases of a step (feed, launch, fetch, checkpoint), visible in
  Nsight Systems / ncu when ``--nvtx`` is on; a no-op otherwise and on machines without CUDA.
* :class:`DeviceStepTimer` — CUDA-event device time over a window of steps (what ``bench.py`` reports, for use
  inside a training run): no host synchronisation until the window is read.
"""
from __future__ import annotations

import contextlib
from typing import Optional

_ENABLED = False


def enable_nvtx(on: bool = True) -> None:
    global _ENABLED
    _ENABLED = bool(on)


def nvtx_enabled() -> bool:
    return _ENABLED


@contextlib.contextmanager
def nvtx_range(name: str):
    if not _ENABLED:
        yield
        return
    try:
        import torch
        nvtx = torch.cuda.nvtx if torch.cuda.is_available() else None
    except Exception:  # noqa: BLE001
        nvtx = None
    if nvtx is None:
        yield
        return
    nvtx.range_push(name)
    try:
        yield
    finally:
        nvtx.range_pop()


class DeviceStepTimer:
    """``start()`` before a window of steps, ``stop(n)`` after: microseconds per step on the device."""

    def __init__(self, stream=None):
        import torch
        self._torch = torch
        self._stream = stream
        self._e0 = torch.cuda.Event(enable_timing=True)
        self._e1 = torch.cuda.Event(enable_timing=True)
        self.last_us_per_step: Optional[float] = None

    def start(self) -> None:
        self._e0.record(self._stream)

    def stop(self, steps: int) -> float:
        self._e1.record(self._stream)
        self._e1.synchronize()
        self.last_us_per_step = self._e0.elapsed_time(self._e1) * 1e3 / max(steps, 1)
        return self.last_us_per_step
