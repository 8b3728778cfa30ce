"""Meters.

``AverageMeter`` / ``TimeMeter`` / ``StopwatchMeter`` keep the attribute names and semantics of the
reference (meters.py:4-66) because instances are pickled into checkpoints
(``extra_state['train_meters']``, controller.py:131) and formatted by the log line. The reference
only has host wall-clock timing; ``CudaEventTimer`` adds the device-side timing the benchmarks need
(CUDA events on the launching stream, no host sync on the hot path, max over ranks at read-out).
"""
from __future__ import annotations

import time
from typing import List, Optional


class AverageMeter(object):
    """Running weighted mean; ``val`` is the last value (attributes val / avg / sum / count as in the reference)."""

    def __init__(self):
        self.val = self.avg = self.sum = self.count = 0

    reset = __init__

    def update(self, val, n=1):
        self.val, self.sum, self.count = val, self.sum + val * n, self.count + n
        self.avg = self.sum / self.count if self.count else 0


class TimeMeter(object):
    """Events per second of host wall clock since ``reset`` (attributes init / start / n)."""

    def __init__(self, init=0):
        self.init, self.start, self.n = init, time.time(), 0

    reset = __init__

    def update(self, val=1):
        self.n += val

    @property
    def elapsed_time(self):
        return self.init + (time.time() - self.start)

    @property
    def avg(self):
        t = self.elapsed_time
        return self.n / t if t > 0 else 0.0


class StopwatchMeter(object):
    """Accumulated host seconds between ``start()`` and ``stop()`` (attributes sum / n / start_time)."""

    def __init__(self):
        self.sum, self.n, self.start_time = 0, 0, None

    reset = __init__

    def start(self):
        self.start_time = time.time()

    def stop(self, n=1):
        began, self.start_time = self.start_time, None
        if began is not None:
            self.sum += time.time() - began
            self.n += n

    @property
    def avg(self):
        return self.sum / self.n if self.n else 0.0


class CudaEventTimer(object):
    """Device-side interval timer.

    ``start()``/``stop()`` record CUDA events on the current stream (no synchronisation); the
    elapsed milliseconds are resolved lazily in ``flush()``/``elapsed_ms``. On CPU it falls back to
    ``time.perf_counter``.
    """

    def __init__(self, use_cuda: Optional[bool] = None):
        import torch

        self._torch = torch
        self.use_cuda = torch.cuda.is_available() if use_cuda is None else use_cuda
        self._pending: List = []
        self._open = None
        self.total_ms = 0.0
        self.n = 0

    def start(self):
        if self.use_cuda:
            ev = self._torch.cuda.Event(enable_timing=True)
            ev.record()
            self._open = ev
        else:
            self._open = time.perf_counter()

    def stop(self):
        if self._open is None:
            return
        if self.use_cuda:
            ev = self._torch.cuda.Event(enable_timing=True)
            ev.record()
            self._pending.append((self._open, ev))
        else:
            self.total_ms += (time.perf_counter() - self._open) * 1e3
            self.n += 1
        self._open = None

    def flush(self):
        """Resolve recorded intervals (synchronises on the last stop event only)."""
        if self.use_cuda and self._pending:
            self._pending[-1][1].synchronize()
            for a, b in self._pending:
                self.total_ms += a.elapsed_time(b)
                self.n += 1
            self._pending = []

    @property
    def elapsed_ms(self) -> float:
        self.flush()
        return self.total_ms

    @property
    def avg_ms(self) -> float:
        self.flush()
        return self.total_ms / self.n if self.n else 0.0

    def reset(self):
        self._pending, self._open, self.total_ms, self.n = [], None, 0.0, 0
