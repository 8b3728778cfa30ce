"""Decides, per call site, whether the native sm_100a kernels or the PyTorch restatement runs.

There is exactly one GPU backend (sm_100a); the PyTorch path exists because the framework must
also run on CPU (MNIST config, unit tests, the numerics oracle). On a CUDA device the native
library is *required* unless the user explicitly asked for ``--kernels torch``: a missing or
unloadable library is an error, never a silent fallback.
"""
from __future__ import annotations

import os

import torch

_FORCE = os.environ.get('HETSEQ_KERNELS', '').lower()  # '', 'native', 'torch'


def mode(args=None) -> str:
    m = getattr(args, 'kernels', None) or 'auto'
    if _FORCE in ('native', 'torch'):
        m = _FORCE
    return m


def use_native(args, device) -> bool:
    device = torch.device(device)
    m = mode(args)
    if m == 'torch' or device.type != 'cuda':
        if m == 'native' and device.type != 'cuda':
            raise RuntimeError("--kernels native needs a CUDA device")
        return False
    # CUDA device and mode in {auto, native}: the library must load.
    from hetseq_b200.ops import lib
    lib.load()  # raises with a clear message if the .so is missing / unloadable
    return True
