"""Small helpers used across the framework.

Counterparts of the *live* parts of the reference's utils.py (apply_to_sample :12-29,
move_to_cuda :32-37, item :86-91, get_perplexity :167-171, activation lookup :179-206,
import_user_module :137-150, has_parameters :209-214). The dead fairseq residue that references
undefined names (utils.py:40-52,99-135) is not reproduced.
"""
from __future__ import annotations

import importlib
import math
import os
import sys
import warnings
from typing import Callable, List

import torch
import torch.nn.functional as F


def apply_to_sample(f, sample):
    if sample is None or (hasattr(sample, '__len__') and len(sample) == 0):
        return {}

    def _apply(x):
        if torch.is_tensor(x):
            return f(x)
        if isinstance(x, dict):
            return {k: _apply(v) for k, v in x.items()}
        if isinstance(x, (list, tuple)):
            return type(x)(_apply(v) for v in x) if isinstance(x, tuple) else [_apply(v) for v in x]
        return x

    return _apply(sample)


class PackedSample(list):
    """A micro-batch whose tensors are views of ONE flat buffer (``.flat``), so that it crosses PCIe -- and is
    copied into the static input buffers of a captured CUDA graph -- as a single copy instead of one per tensor
    (a BERT batch is 7 small int32 tensors: the per-copy latency, not the 70 KB, was what the step waited for).
    Behaves like the plain list of tensors everywhere else."""

    ALIGN = 4      # elements; keeps every view 16-byte aligned for 4-byte dtypes

    def __init__(self, flat: torch.Tensor, shapes):
        self.flat = flat
        self.shapes = [tuple(s) for s in shapes]
        views, off = [], 0
        for shp in self.shapes:
            n = 1
            for d in shp:
                n *= d
            views.append(flat[off:off + n].view(shp))
            off += (n + self.ALIGN - 1) // self.ALIGN * self.ALIGN
        super().__init__(views)

    @classmethod
    def pack(cls, tensors, pin: bool = False):
        """Pack CPU tensors of one common dtype; returns None when they cannot be packed."""
        tensors = list(tensors)
        if not tensors or any((not torch.is_tensor(t)) or t.device.type != 'cpu' or t.dtype != tensors[0].dtype
                              for t in tensors):
            return None
        total = sum((t.numel() + cls.ALIGN - 1) // cls.ALIGN * cls.ALIGN for t in tensors)
        flat = torch.zeros(total, dtype=tensors[0].dtype, pin_memory=bool(pin))
        out = cls(flat, [t.shape for t in tensors])
        for dst, src in zip(out, tensors):
            dst.copy_(src)
        return out

    def to(self, device, non_blocking: bool = True) -> 'PackedSample':
        device = torch.device(device)
        if self.flat.device == device:
            return self
        return PackedSample(self.flat.to(device, non_blocking=non_blocking), self.shapes)

    def clone(self) -> 'PackedSample':
        return PackedSample(self.flat.clone(), self.shapes)

    def same_layout(self, other) -> bool:
        return isinstance(other, PackedSample) and other.shapes == self.shapes and other.flat.dtype == self.flat.dtype


def move_to_device(sample, device, non_blocking: bool = True):
    """Copy every tensor of a (nested) sample to ``device``.

    The reference does a blocking ``.cuda()`` from pageable memory (utils.py:32-37); here tensors
    are pinned first when they are not already, so the copy can overlap with compute.
    """
    device = torch.device(device)
    if isinstance(sample, PackedSample):
        return sample.to(device, non_blocking=non_blocking)

    def _move(t):
        if t.device == device:
            return t
        if device.type == 'cuda' and t.device.type == 'cpu' and non_blocking:
            if not t.is_pinned():
                t = t.pin_memory()
            return t.to(device, non_blocking=True)
        return t.to(device)

    return apply_to_sample(_move, sample)


def move_to_cuda(sample):
    return move_to_device(sample, torch.device('cuda', torch.cuda.current_device()))


def item(tensor):
    if hasattr(tensor, 'item'):
        return tensor.item()
    if hasattr(tensor, '__getitem__'):
        return tensor[0]
    return tensor


def get_perplexity(loss):
    try:
        return float('{:.2f}'.format(math.pow(2, loss)))
    except (OverflowError, ValueError, TypeError):
        return float('inf')


def gelu_ref(x: torch.Tensor) -> torch.Tensor:
    """GELU exactly as the reference writes it (bert_modeling.py:104-106): note 1.41421, not sqrt(2)."""
    return x * 0.5 * (1.0 + torch.erf(x / 1.41421))


def gelu_accurate(x: torch.Tensor) -> torch.Tensor:
    return 0.5 * x * (1 + torch.tanh(math.sqrt(2 / math.pi) * (x + 0.044715 * torch.pow(x, 3))))


def swish(x: torch.Tensor) -> torch.Tensor:
    return x * torch.sigmoid(x)


def get_activation_fn(activation: str) -> Callable:
    table = {
        'relu': F.relu,
        'gelu': gelu_ref,
        'gelu_fast': gelu_accurate,
        'gelu_accurate': gelu_accurate,
        'tanh': torch.tanh,
        'swish': swish,
        'linear': lambda x: x,
    }
    if activation not in table:
        raise RuntimeError('--activation-fn {} not supported'.format(activation))
    if activation == 'gelu_fast':
        deprecation_warning('--activation-fn=gelu_fast has been renamed to gelu_accurate')
    return table[activation]


def get_available_activation_fns() -> List[str]:
    return ['relu', 'gelu', 'gelu_fast', 'gelu_accurate', 'tanh', 'swish', 'linear']


def deprecation_warning(message, stacklevel=3):
    warnings.warn(message, stacklevel=stacklevel)


def import_user_module(args):
    """Import ``--user-dir`` so that its ``@register_*`` decorators run."""
    module_path = getattr(args, 'user_dir', None)
    if module_path is None:
        return
    module_path = os.path.abspath(module_path)
    module_parent, module_name = os.path.split(module_path)
    if module_name not in sys.modules:
        sys.path.insert(0, module_parent)
        try:
            importlib.import_module(module_name)
        finally:
            sys.path.pop(0)


def has_parameters(module) -> bool:
    try:
        next(module.parameters())
        return True
    except StopIteration:
        return False


def fill_with_neg_inf(t):
    return t.float().fill_(float('-inf')).type_as(t)
