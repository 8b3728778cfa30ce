"""Data-parallel gradient backends over a FlatParamBuffer.

The reference wraps the model in ``torch.nn.parallel.DistributedDataParallel`` (controller.py:74-89)
and skips the all-reduce on all but the last micro-batch with ``no_sync()`` (controller.py:245-258).
Here the model is never wrapped. A backend owns the question "how do the flat gradients of W ranks
become an update of the flat parameters":

``LocalBackend``       world size 1 -- nothing to communicate.
``LocalNormBackend``   world size 1 on a GPU with the native kernels, opt-in (``HETSEQ_NORM_OVERLAP=1``): the sum of
                       squares for the gradient norm is taken bucket by bucket on a side stream while backward
                       is still running (the same gradient-ready callbacks as below), so the 1.3 GB read of the
                       flat gradient is off the critical path. Measured on BERT-large (same box, same call): 13.28 ms
                       with it, 13.25 ms without -- the 0.2 ms it hides are paid back as slower GEMMs next to the
                       bandwidth-bound kernels, so the default stays the single pass after backward.
``AllReduceBackend``   the in-framework baseline (``--dp-backend nccl``): bucketed asynchronous
                       all-reduce of the flat gradient over torch.distributed (NCCL on GPUs, Gloo in
                       CPU tests), launched bucket by bucket from gradient-ready callbacks so it
                       overlaps with backward, followed by a replicated (unsharded) update. This is
                       what DDP + the reference's optimizer loop amount to, minus the launches.
``FusedBackend``       the product (``--dp-backend fused``, parallel/fused_dp.py): in-kernel
                       reduce-scatter over NVLink peer memory, sharded update, in-kernel all-gather.

Common protocol, driven by the Controller for each optimizer update:
    begin_update(n_micro) -> [micro-batch backward passes; ready callbacks fire on the last one]
    -> finish_gradients() -> optimizer.clip_grad_norm()/step() -> (sharded: all_gather_params)
Gradient semantics match DDP: after ``finish_gradients`` the optimizer sees the MEAN over ranks
(the 1/W is folded into the optimizer's pending multiplier, no extra pass over memory).
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.distributed as dist

from hetseq_b200.parallel.flat_buffer import FlatParamBuffer


class LocalBackend(object):
    sharded = False
    name = 'local'

    def __init__(self, args, buffer: FlatParamBuffer):
        self.args, self.buffer = args, buffer
        self.world_size, self.rank = 1, 0

    def broadcast_params(self):
        pass

    def begin_update(self, num_micro_batches: int):
        pass

    def begin_micro_batch(self, index: int, is_last: bool):
        self.buffer.reset_ready_counts()

    def finish_gradients(self, optimizer):
        pass

    def all_reduce_scalar(self, t: torch.Tensor) -> torch.Tensor:
        return t

    def check_params_consistent(self) -> bool:
        return True

    def flush(self, sync: bool = True):
        """Make every queued parameter update visible in ``flat_param`` (only the fused backend defers any)."""
        pass

    def close(self):
        pass


class _ReadyBuckets(object):
    """Bookkeeping shared by the backends that act on a gradient bucket as soon as backward has finished with it:
    the flat gradient is cut into contiguous buckets (last-registered parameters first), parameters report through
    ``FlatParamBuffer.ready_callback`` during the LAST micro-batch of an update, and ``_launch(b)`` (the backend's)
    runs when every parameter that overlaps bucket b has reported."""

    def _setup_buckets(self, buffer: FlatParamBuffer, bucket_elems: int):
        self.buckets = buffer.bucket_ranges(bucket_elems)
        # bucket -> parameters that must be ready before it can be reduced
        self._bucket_params = [buffer.params_in_range(s, e) for s, e in self.buckets]
        self._param_buckets: List[List[int]] = [[] for _ in buffer.params]
        for b, plist in enumerate(self._bucket_params):
            for i in plist:
                self._param_buckets[i].append(b)
        self._pending = [0] * len(self.buckets)
        self._launched = [False] * len(self.buckets)
        self._armed = False
        # debugging aid: act on everything after backward instead of bucket by bucket during it
        import os
        self._no_overlap = bool(int(os.environ.get('HETSEQ_DP_NO_OVERLAP', '0')))
        buffer.ready_callback = self._on_param_ready

    def _arm(self, is_last: bool):
        self._armed = bool(is_last)
        self._seen = set()
        if is_last:
            self._pending = [len(p) for p in self._bucket_params]
            self._launched = [False] * len(self.buckets)

    def _on_param_ready(self, i: int):
        """Called (from autograd hooks or native ops) when parameter i's gradient is final."""
        if not self._armed or self._no_overlap:
            return
        if i in self._seen:
            raise RuntimeError('parameter {} reported gradient-ready twice in one backward pass'.format(i))
        self._seen.add(i)
        for b in self._param_buckets[i]:
            self._pending[b] -= 1
            if self._pending[b] == 0 and not self._launched[b]:
                self._launch(b)

    def _launch_rest(self):
        """Buckets whose parameters never reported (unused parameters, or a model without hooks)."""
        for b in range(len(self.buckets)):
            if not self._launched[b]:
                self._launch(b)
        self._armed = False


class LocalNormBackend(_ReadyBuckets, LocalBackend):
    """One GPU, native kernels: per-bucket sums of squares of the fp32 flat gradient on a side stream, started from
    the gradient-ready callbacks while backward continues; ``take_local_sumsq`` (asked by the optimizer's
    ``clip_grad_norm``) adds the bucket results in bucket order -- one fixed order of additions, same bits every
    run. What the reference does after backward with one norm kernel per parameter
    (/root/reference/optim.py:65-70, controller.py:338-347). Off by default: see the module docstring."""

    name = 'local'
    NORM_BUCKET_MB = 96

    def __init__(self, args, buffer: FlatParamBuffer):
        LocalBackend.__init__(self, args, buffer)
        from hetseq_b200.ops import lib
        from hetseq_b200.parallel.flat_buffer import CHUNK
        self._lib, self._L, self._chunk = lib, lib.load(), CHUNK
        self._setup_buckets(buffer, self.NORM_BUCKET_MB * (1 << 20) // buffer.flat_grad.element_size())
        dev = buffer.device
        n = len(self.buckets)
        self.side = torch.cuda.Stream(device=dev)
        per = int(self._L.hs_sumsq_scratch_elems())
        self._scratch = torch.empty(n * per, dtype=torch.float32, device=dev)
        self._per = per
        self.partials = torch.zeros(n, dtype=torch.float32, device=dev)
        self._valid = False

    @staticmethod
    def usable(args, buffer: FlatParamBuffer) -> bool:
        import os
        if os.environ.get('HETSEQ_NORM_OVERLAP', '0') != '1' or buffer.device.type != 'cuda':
            return False
        if buffer.flat_grad.dtype != torch.float32 or getattr(args, 'kernels', 'auto') == 'torch':
            return False
        from hetseq_b200.ops import lib
        return lib.available()

    def begin_micro_batch(self, index: int, is_last: bool):
        self.buffer.reset_ready_counts()
        self._valid = False
        self._arm(is_last)

    def _launch(self, b: int):
        s, e = self.buckets[b]
        self._launched[b] = True
        ev = torch.cuda.Event()
        ev.record()
        self.side.wait_event(ev)
        g = self.buffer.flat_grad
        rc = self._L.hs_sumsq(g.data_ptr() + s * g.element_size(), self._lib.dtype_code(g.dtype), (e - s) // self._chunk,
                              self._chunk, self._chunk, self._scratch.data_ptr() + b * self._per * 4,
                              self.partials.data_ptr() + b * 4, 0, self.side.cuda_stream)
        self._lib.check(rc, 'hs_sumsq')

    def finish_gradients(self, optimizer):
        self._launch_rest()
        torch.cuda.current_stream().wait_stream(self.side)
        self._valid = True

    def take_local_sumsq(self) -> Optional[torch.Tensor]:
        """Sum of squares of the whole flat gradient if this update's buckets were all measured, else None (the
        optimizer then measures the buffer itself, e.g. when it is driven without the backend protocol)."""
        if not self._valid:
            return None
        self._valid = False
        return self.partials.sum()


class AllReduceBackend(_ReadyBuckets, LocalBackend):
    """Bucketed all-reduce of the flat gradient + replicated optimizer (the NCCL baseline)."""

    name = 'nccl'

    def __init__(self, args, buffer: FlatParamBuffer, group=None):
        LocalBackend.__init__(self, args, buffer)
        self.group = group
        self.world_size = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        elem = buffer.flat_grad.element_size()
        self._setup_buckets(buffer, int(getattr(args, 'bucket_cap_mb', 25)) * (1 << 20) // elem)
        self._handles = []

    def broadcast_params(self):
        """Rank 0's parameters become everybody's (what the DDP constructor does)."""
        dist.broadcast(self.buffer.flat_param, src=0, group=self.group)

    def begin_update(self, num_micro_batches: int):
        self._handles = []

    def begin_micro_batch(self, index: int, is_last: bool):
        self.buffer.reset_ready_counts()
        self._arm(is_last)

    def _launch(self, b: int):
        s, e = self.buckets[b]
        self._launched[b] = True
        self._handles.append(dist.all_reduce(self.buffer.flat_grad[s:e], group=self.group, async_op=True))

    def finish_gradients(self, optimizer):
        self._launch_rest()
        for h in self._handles:
            h.wait()
        self._handles = []
        optimizer.multiply_grads(1.0 / self.world_size)

    def all_reduce_scalar(self, t: torch.Tensor) -> torch.Tensor:
        t = t.clone()
        dist.all_reduce(t, group=self.group)
        return t

    def check_params_consistent(self) -> bool:
        """Cross-rank invariant replacing the reference's grad-norm equality assert
        (controller.py:325-328): every replica must hold bit-identical parameters."""
        p = self.buffer.flat_param
        chk = torch.stack([p.float().sum(), p.float().abs().sum()]).to(torch.float64)
        lo, hi = chk.clone(), chk.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=self.group)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=self.group)
        return bool(torch.equal(lo, hi))


class ShardedAllReduceBackend(AllReduceBackend):
    """Reference implementation of the *sharded* protocol on top of torch.distributed collectives
    (reduce = all-reduce, gather = all-gather of compact shards). Used by the CPU/Gloo tests to
    exercise the partitioned optimizer and checkpoint gathering without GPUs; the fused NVLink
    backend implements the same interface with in-kernel transfers."""

    sharded = True
    name = 'sharded-allreduce'

    def all_gather_compact(self, compact: torch.Tensor) -> torch.Tensor:
        """[shard_size] per rank -> [world, shard_size]."""
        compact = compact.contiguous()
        out = [torch.empty_like(compact) for _ in range(self.world_size)]
        dist.all_gather(out, compact, group=self.group)
        return torch.stack(out)

    def all_gather_params(self, buffer: FlatParamBuffer):
        mine = buffer.owned(buffer.flat_param).reshape(-1).contiguous()
        full = self.all_gather_compact(mine)
        buffer.flat_param.copy_(buffer.compact_to_flat(full))


def build_backend(args, buffer: FlatParamBuffer):
    """Pick the gradient backend for this run (``--dp-backend``)."""
    world = getattr(args, 'distributed_world_size', 1)
    if world <= 1 or not dist.is_initialized():
        return LocalNormBackend(args, buffer) if LocalNormBackend.usable(args, buffer) else LocalBackend(args, buffer)
    choice = getattr(args, 'dp_backend', 'auto')
    on_cuda = buffer.device.type == 'cuda'
    if choice == 'auto':
        from hetseq_b200.parallel import fused_dp
        choice = 'fused' if (on_cuda and fused_dp.is_supported(args)) else 'nccl'
    if choice == 'fused':
        if not on_cuda:
            raise RuntimeError('--dp-backend fused needs CUDA devices in one NVLink domain')
        from hetseq_b200.parallel import fused_dp
        return fused_dp.FusedBackend(args, buffer)
    if choice == 'sharded':
        return ShardedAllReduceBackend(args, buffer)
    return AllReduceBackend(args, buffer)
