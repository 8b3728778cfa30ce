"""NVTX ranges for ``--profile-nvtx``.

The reference has exactly one range, around the MLM decoder GEMM (bert_modeling.py:546-548), and formats
tensor sizes into its label on every forward. Here ranges are off unless requested (a disabled range is one
attribute test), labels are constant strings, and the phases of an update are covered: forward+backward per
micro-batch, every encoder layer's forward and backward, gradient reduction, norm/clip, update + all-gather.
They show up in Nsight Systems / Nsight Compute (``--nvtx``) timelines.
"""
from __future__ import annotations

import contextlib

import torch

_enabled = False


def enable(flag: bool = True) -> None:
    global _enabled
    _enabled = bool(flag) and torch.cuda.is_available()


def enabled() -> bool:
    return _enabled


def push(name: str) -> None:
    if _enabled:
        torch.cuda.nvtx.range_push(name)


def pop() -> None:
    if _enabled:
        torch.cuda.nvtx.range_pop()


@contextlib.contextmanager
def range(name: str):   # noqa: A001 - mirrors torch.cuda.nvtx.range
    if _enabled:
        torch.cuda.nvtx.range_push(name)
        try:
            yield
        finally:
            torch.cuda.nvtx.range_pop()
    else:
        yield
