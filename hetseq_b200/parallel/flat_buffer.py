"""Flat parameter / gradient buffers.

The reference never flattens anything: it relies on DDP's internal 25 MiB buckets for the gradient
all-reduce (controller.py:78-86, options.py:122) and loops over tensors for ``multiply_grads``,
``clip_grad_norm`` and ``Adam.step`` (optim.py:59-70,162-231) -- thousands of tiny launches per
step. Here every trainable parameter is a *view* into one contiguous buffer
and so is its gradient ("main grad"), which makes the gradient path a handful of kernels over
flat memory: one norm, one fused unscale/clip/Adam, and -- in the fused data-parallel backend --
reduce-scatter / all-gather over rank-contiguous shards of the very same buffers.

Layout: parameters in registration order, each start aligned to ``ALIGN`` elements, total length
padded to a multiple of ``world_size * CHUNK``. Ownership for the sharded optimizer is *cyclic by
chunk*: element ``i`` belongs to rank ``(i // CHUNK) % world_size``. Any bucket whose bounds are
multiples of ``world_size * CHUNK`` is therefore spread evenly over all ranks, so a bucket-wise
reduce-scatter loads every GPU's NVLink ingress equally no matter how buckets are chosen. A rank's
optimizer state is compact (``total / world_size`` elements); ``owned(t)`` gives the strided view
of a flat tensor that corresponds to it. Tied parameters (one ``nn.Parameter`` reachable
through several modules, e.g. BERT's decoder/word-embedding weight, bert_modeling.py:541) appear
once.
"""
from __future__ import annotations

from typing import Callable, Iterable, List, Optional, Tuple

import torch

ALIGN = 128    # elements; 256 B for 16-bit params, 512 B for fp32
CHUNK = 2048   # elements; ownership granule of the sharded optimizer (8 KiB of fp32 gradient)
SMALL_PARAM_NUMEL = 16384   # biases / LayerNorm vectors: zeroed eagerly at the start of an update


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class FlatParamBuffer(object):
    """Owns ``flat_param`` (model dtype) and ``flat_grad`` (``grad_dtype``, fp32 by default).

    After construction ``p.data`` of every parameter aliases ``flat_param`` and ``p.main_grad``
    aliases ``flat_grad``. Gradients reach ``main_grad`` in one of three ways:

    * native fused ops write (or accumulate) it directly and return no autograd gradient;
    * parameters whose dtype equals ``grad_dtype`` get ``p.grad = p.main_grad`` so autograd
      accumulates in place;
    * any other parameter gets a post-accumulate hook that folds ``p.grad`` into ``main_grad``.
    """

    def __init__(self, params: Iterable[torch.nn.Parameter], world_size: int = 1, rank: int = 0,
                 grad_dtype: torch.dtype = torch.float32,
                 allocator: Optional[Callable[[int, torch.dtype, torch.device], torch.Tensor]] = None):
        seen = set()
        self.params: List[torch.nn.Parameter] = []
        for p in params:
            if p.requires_grad and id(p) not in seen:
                seen.add(id(p))
                self.params.append(p)
        if not self.params:
            raise ValueError('FlatParamBuffer needs at least one trainable parameter')
        p0 = self.params[0]
        self.device, self.param_dtype = p0.device, p0.dtype
        for p in self.params:
            if p.device != self.device or p.dtype != self.param_dtype:
                raise ValueError('all parameters must share one device and dtype '
                                 f'(got {p.device}/{p.dtype} vs {self.device}/{self.param_dtype})')
        self.grad_dtype = grad_dtype
        self.world_size, self.rank = int(world_size), int(rank)

        self.offsets: List[int] = []
        self.numels: List[int] = []
        off = 0
        for p in self.params:
            self.offsets.append(off)
            self.numels.append(p.numel())
            off = _round_up(off + p.numel(), ALIGN)
        self.used = off
        self.total = _round_up(max(off, 1), self.world_size * CHUNK)
        self.shard_size = self.total // self.world_size
        self.num_blocks = self.total // (self.world_size * CHUNK)

        #: where the two flat buffers came from (a SymmetricPool under the fused NVLink backend, else None)
        self.allocator = allocator
        alloc = allocator or (lambda n, dt, dev: torch.zeros(n, dtype=dt, device=dev))
        self.flat_param = alloc(self.total, self.param_dtype, self.device)
        self.flat_grad = alloc(self.total, self.grad_dtype, self.device)
        self.main_grads: List[torch.Tensor] = []
        self._written: List[bool] = [False] * len(self.params)
        self._hooks = []
        with torch.no_grad():
            for i, p in enumerate(self.params):
                view = self.flat_param[self.offsets[i]: self.offsets[i] + self.numels[i]].view(p.shape)
                view.copy_(p.data)
                p.data = view
                g = self.flat_grad[self.offsets[i]: self.offsets[i] + self.numels[i]].view(p.shape)
                self.main_grads.append(g)
                p.main_grad = g
                p._hs_flat = (self, i)
                self._hooks.append(p.register_post_accumulate_grad_hook(self._make_hook(i)))
        #: called with a parameter index once that parameter's gradient is final for the current
        #: micro-batch; data-parallel backends use it to start reducing a bucket during backward
        self.ready_callback: Optional[Callable[[int], None]] = None
        #: how many gradient contributions a parameter receives from ops that write main_grad themselves
        #: (2 for BERT's tied decoder / word-embedding matrix); the callback fires on the last one
        self._expected_notifications: List[int] = [1] * len(self.params)
        self._notifications: List[int] = [0] * len(self.params)
        self._is_small = [p.numel() <= SMALL_PARAM_NUMEL for p in self.params]
        self._small_grads = [g for g, sm in zip(self.main_grads, self._is_small) if sm]
        #: hooks of a backend whose parameter update overlaps the next forward pass (parallel/fused_dp.py):
        #: ``param_gate(flat_end)`` is called by a gated forward pass in front of the first kernel that reads
        #: parameters below flat offset ``flat_end``; ``before_backward()`` once before gradients are written
        self.param_gate: Optional[Callable[[int], None]] = None
        self.before_backward: Optional[Callable[[], None]] = None
        #: set by models whose forward calls ``gate`` (ops/bert_native.py); others get everything up front
        self.gated_forward = False
        self._small_zero_pending = False
        self.zero_grad()

    # -- gradient plumbing ----------------------------------------------------------------------
    def _make_hook(self, i: int):
        fold = self.param_dtype != self.grad_dtype

        def hook(p):
            if fold:
                # autograd also runs this hook when an op returned no gradient for the parameter (the
                # native ops write main_grad themselves and report through notify_ready): nothing to do
                if p.grad is None:
                    return
                g = self.main_grads[i]
                if self._written[i]:
                    g.add_(p.grad)
                else:
                    g.copy_(p.grad)
                    self._written[i] = True
                p.grad = None
            if self.ready_callback is not None:
                self.ready_callback(i)
        return hook

    def notify_ready(self, p: torch.nn.Parameter) -> None:
        """For ops that wrote ``main_grad`` themselves (no autograd hook fires for them). A parameter
        used by several such ops (tied weights) is reported once, after its last contribution."""
        i = p._hs_flat[1]
        self._notifications[i] += 1
        if self._notifications[i] >= self._expected_notifications[i] and self.ready_callback is not None:
            self.ready_callback(i)

    def set_expected_notifications(self, p: torch.nn.Parameter, n: int) -> None:
        self._expected_notifications[p._hs_flat[1]] = int(n)

    def reset_ready_counts(self) -> None:
        """Called at the start of every micro-batch's backward."""
        for i in range(len(self._notifications)):
            self._notifications[i] = 0

    def gate(self, params) -> None:
        """Forward-side wait for the (asynchronously arriving) parameters ``params`` and everything before them
        in the flat buffer. A no-op unless a backend installed ``param_gate``."""
        if self.param_gate is None:
            return
        end = 0
        for p in params:
            i = p._hs_flat[1]
            end = max(end, self.offsets[i] + self.numels[i])
        self.param_gate(end)

    def begin_backward(self) -> None:
        """Called once per micro-batch right before ``loss.backward()``: lets the backend finish reading the
        previous update's gradients, then clears the small (atomically accumulated) gradients if this is the
        first backward pass of the accumulation window."""
        if self.before_backward is not None:
            self.before_backward()
        if self._small_zero_pending:
            self._small_zero_pending = False
            torch._foreach_zero_(self._small_grads)

    def grad_target(self, p: torch.nn.Parameter) -> Tuple[torch.Tensor, bool]:
        """(main_grad view, accumulate?) for an op that writes the gradient itself."""
        i = p._hs_flat[1]
        acc = self._written[i]
        self._written[i] = True
        return self.main_grads[i], acc

    def ensure_initialized(self, p: torch.nn.Parameter) -> torch.Tensor:
        """main_grad view guaranteed to hold a valid partial sum (for ops that can only accumulate)."""
        i = p._hs_flat[1]
        if not self._written[i]:
            self.main_grads[i].zero_()
            self._written[i] = True
        return self.main_grads[i]

    def zero_grad(self, set_to_zero: Optional[bool] = None) -> None:
        """Start a new accumulation window.

        With ``set_to_zero`` the whole flat gradient is cleared with one memset (1.3 GB for BERT-large:
        0.2 ms of HBM time per update). That is only needed when autograd accumulates straight into the
        views (same parameter and gradient dtype). With separate fp32 main gradients every producer -- the
        native ops through ``grad_target`` and the fold hook -- OVERWRITES on its first write of a window,
        so the default (``None``) skips the memset there; ``finalize_grads`` clears whatever no producer
        touched. Padding between parameters is never written by anyone and stays zero from allocation.
        """
        same = self.param_dtype == self.grad_dtype
        if set_to_zero is None:
            set_to_zero = same
        if set_to_zero:
            self.flat_grad.zero_()
        elif self._small_grads:
            # small parameters (biases, LayerNorm) are produced by kernels that accumulate with atomics; one
            # multi-tensor fill is cheaper than a memset node in front of each of those kernels. With a backend
            # whose update still reads the previous gradients during the forward pass the fill waits until
            # ``begin_backward``.
            if self.before_backward is not None:
                self._small_zero_pending = True
            else:
                torch._foreach_zero_(self._small_grads)
        for i, p in enumerate(self.params):
            # after a memset every view is a valid partial sum, so later writes must accumulate
            self._written[i] = bool(set_to_zero) or self._is_small[i]
            p.grad = self.main_grads[i] if same else None

    def finalize_grads(self) -> None:
        """End of the accumulation window: parameters that received no gradient at all (unused in this
        update) get zeros instead of the previous window's values, and are reported ready."""
        if self._small_zero_pending:       # an update without any backward pass
            self.begin_backward()
        for i, w in enumerate(self._written):
            if not w:
                self.main_grads[i].zero_()
                self._written[i] = True
                if self.ready_callback is not None:
                    self.ready_callback(i)

    def mark_all_written(self) -> None:
        for i in range(len(self._written)):
            self._written[i] = True

    # -- shards ---------------------------------------------------------------------------------
    def owned(self, flat: torch.Tensor, rank: Optional[int] = None) -> torch.Tensor:
        """[num_blocks, CHUNK] strided view of the elements of ``flat`` that ``rank`` owns."""
        r = self.rank if rank is None else rank
        return flat.view(self.num_blocks, self.world_size, CHUNK)[:, r, :]

    def compact_to_flat(self, per_rank: torch.Tensor) -> torch.Tensor:
        """[world, shard_size] compact shards (rank-major) -> flat tensor in natural element order."""
        return per_rank.view(self.world_size, self.num_blocks, CHUNK).permute(1, 0, 2).reshape(-1)

    def bucket_ranges(self, bucket_elems: int) -> List[Tuple[int, int]]:
        """Split [0, total) into contiguous buckets of ~bucket_elems whose bounds are multiples of
        world_size * CHUNK, ordered from the END of the buffer (parameters registered last receive
        their gradients first in backward)."""
        gran = self.world_size * CHUNK
        bucket_elems = max(_round_up(bucket_elems, gran), gran)
        out = []
        end = self.total
        while end > 0:
            start = max(0, end - bucket_elems)
            out.append((start, end))
            end = start
        return out

    def params_in_range(self, start: int, end: int) -> List[int]:
        """Indices of the parameters that overlap [start, end)."""
        return [i for i, (off, n) in enumerate(zip(self.offsets, self.numels)) if off < end and off + n > start]

    def numel(self) -> int:
        return sum(self.numels)
