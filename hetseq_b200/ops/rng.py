"""Per-step seed shared by every native dropout site.

The reference reseeds torch's global generators with ``seed + num_updates`` before each update
(controller.py:426-432) so that a resumed run replays the same dropout masks. The native kernels
draw from a counter-based Philox stream keyed by (step seed, site id, element index): the mask is
a pure function of those, so backward recomputes it instead of storing it, and resume reproduces it.
The seed lives in a 2-element int64 device tensor ``[seed, rank]`` that kernels read through a
pointer, which keeps captured CUDA graphs valid across steps.
"""
from __future__ import annotations

import torch

_host = [0, 0]
_dev = {}
_site_counter = [0]


def set_step_seed(seed: int, rank: int = 0) -> None:
    _host[0], _host[1] = int(seed), int(rank)
    for t in _dev.values():
        t.copy_(torch.tensor(_host, dtype=torch.int64), non_blocking=True)


def set_micro_batch(index: int) -> None:
    """Fold the micro-batch number of the current update into the seed (``--update-freq`` > 1): the masks are a
    pure function of (seed, site, element), so without this every micro-batch of an update would draw the very
    same masks, whereas the reference draws fresh ones from the torch generator for each (controller.py:235).
    Forward and backward of one micro-batch run between two calls, so recomputation stays consistent."""
    base = _host[0]
    for t in _dev.values():
        t.copy_(torch.tensor([base + 2654435761 * int(index), _host[1]], dtype=torch.int64), non_blocking=True)


def seed_tensor(device) -> torch.Tensor:
    device = torch.device(device)
    key = (device.type, device.index)
    if key not in _dev:
        _dev[key] = torch.tensor(_host, dtype=torch.int64, device=device)
    return _dev[key]


def new_site() -> int:
    """A unique id per dropout site (layer x position in the layer), assigned at model attach time."""
    _site_counter[0] += 1
    return _site_counter[0]
