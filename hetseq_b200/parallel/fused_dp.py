"""Fused NVLink data-parallel backend (placeholder until the kernels land in this round)."""
from __future__ import annotations


def is_supported(args) -> bool:
    return False


def symmetric_allocator(args):
    raise RuntimeError('--dp-backend fused is not available in this build')


class FusedBackend(object):
    def __init__(self, *a, **k):
        raise RuntimeError('--dp-backend fused is not available in this build')
