"""Fused NVLink data-parallel backend (``--dp-backend fused``): the gradient path without NCCL.

Counterpart of DDP's bucketed all-reduce + the reference's optimizer loop (controller.py:74-89,
338-347; optim.py:59-70,162-231). The flat gradient and parameter buffers of all ranks of
one node are allocated from CUDA-IPC memory and mapped into every peer, then the kernels of
``csrc/comm_sm100.cu`` do, per update:

  backward ............ as soon as a bucket's gradients are final (gradient-ready callbacks from autograd /
                        the native wgrad ops) its ``hs_dp_reduce_scatter`` is queued on a side stream:
                        announce to peers -> wait for peers -> sum the W copies of the owned chunks
                        straight out of peer memory -> norm partial. Overlaps with the rest of backward.
  clip_grad_norm ...... ``hs_dp_norm_exchange``: W floats through the signal pads; identical norm, clip
                        coefficient and overflow flag on all ranks, no host round trip.
  step ................ ``hs_dp_adam_allgather``: Adam on the 1/W owned chunks, new 16-bit parameters stored
                        into every rank's buffer; ``hs_dp_wait_params`` orders the next forward after it.

torch.distributed is used only to exchange the IPC handles at start-up (and, in the Controller, for the
48-byte statistics all-reduce, which the north star allows to stay on NCCL).
"""
from __future__ import annotations

import ctypes
from typing import List, Optional

import torch
import torch.distributed as dist

from hetseq_b200.ops import lib
from hetseq_b200.parallel.data_parallel import AllReduceBackend
from hetseq_b200.parallel.flat_buffer import CHUNK, FlatParamBuffer

MAX_BUCKETS = 1024
_allocs = []   # (tensor, device pointer, ipc handle bytes) of every symmetric allocation of this process


class _IpcBlock(object):
    """Owner object behind a tensor that aliases an hs_ipc_alloc block (CUDA array interface)."""

    def __init__(self, ptr: int, nbytes: int):
        self.ptr, self.nbytes = ptr, nbytes
        self.__cuda_array_interface__ = {'shape': (nbytes,), 'typestr': '|u1', 'data': (ptr, False), 'version': 3,
                                         'strides': None}


def ipc_tensor(numel: int, dtype: torch.dtype, device) -> torch.Tensor:
    """Zero-initialised tensor in IPC-shareable device memory (its own cudaMalloc block)."""
    L = lib.load()
    nbytes = max(numel * torch.empty(0, dtype=dtype).element_size(), 256)
    ptr = ctypes.c_void_p()
    handle = ctypes.create_string_buffer(64)
    with torch.cuda.device(device):
        lib.check(L.hs_ipc_alloc(nbytes, ctypes.byref(ptr), handle), 'hs_ipc_alloc')
    owner = _IpcBlock(ptr.value, nbytes)
    raw = torch.as_tensor(owner, device=torch.device(device))
    t = raw.view(dtype)[:numel]
    _allocs.append((t, ptr.value, handle.raw))
    return t


def _handle_of(t: torch.Tensor) -> bytes:
    for tt, ptr, handle in _allocs:
        if tt.data_ptr() == t.data_ptr():
            return handle
    raise RuntimeError('tensor was not allocated with the symmetric allocator')


def is_supported(args) -> bool:
    """Single node, CUDA, all ranks' GPUs peer-accessible, kernel library present, Adam, 16-bit params."""
    if not torch.cuda.is_available() or getattr(args, 'cpu', False) or not lib.available():
        return False
    world = getattr(args, 'distributed_world_size', 1)
    if world < 2 or world > min(16, torch.cuda.device_count()):
        return False
    if getattr(args, 'optimizer', 'adam') != 'adam' or not (getattr(args, 'fp16', False) or getattr(args, 'bf16', False)):
        return False
    if getattr(args, 'kernels', 'auto') == 'torch':
        return False
    dev = torch.cuda.current_device()
    return all(d == dev or torch.cuda.can_device_access_peer(dev, d) for d in range(world))


def symmetric_allocator(args):
    return ipc_tensor


class FusedBackend(AllReduceBackend):
    sharded = True
    fused = True
    name = 'fused-nvlink'

    def __init__(self, args, buffer: FlatParamBuffer, group=None):
        super().__init__(args, buffer, group)
        self.L = lib.load()
        if buffer.flat_grad.dtype != torch.float32:
            raise RuntimeError('fused backend expects fp32 main gradients')
        if len(self.buckets) > MAX_BUCKETS:
            raise RuntimeError('too many gradient buckets ({}); raise --bucket-cap-mb'.format(len(self.buckets)))
        dev = buffer.device
        W, r = self.world_size, self.rank
        self.sig = ipc_tensor(int(self.L.hs_dp_signal_words(W, MAX_BUCKETS)), torch.int32, dev)
        # exchange IPC handles; rank-indexed tables of device pointers
        mine = {'grad': _handle_of(buffer.flat_grad), 'param': _handle_of(buffer.flat_param), 'sig': _handle_of(self.sig)}
        everyone: List[Optional[dict]] = [None] * W
        dist.all_gather_object(everyone, mine, group=self.group)
        self._opened = []
        self.grad_ptrs = self._table(everyone, 'grad', buffer.flat_grad)
        self.param_ptrs = self._table(everyone, 'param', buffer.flat_param)
        self.sig_ptrs = self._table(everyone, 'sig', self.sig)
        self.sumsq = torch.zeros(1, dtype=torch.float32, device=dev)
        self.scalars = torch.zeros(4, dtype=torch.float32, device=dev)
        self.skip = torch.zeros(1, dtype=torch.int32, device=dev)
        self.done_counter = torch.zeros(1, dtype=torch.int32, device=dev)
        import os
        # default priority (the lowest): the reduce kernels must never delay the placement of backward's CTAs
        self.comm_stream = torch.cuda.Stream(device=dev, priority=int(os.environ.get('HETSEQ_COMM_PRIORITY', '0')))
        # the update number lives in device memory and is bumped by a tiny device op per update, so the
        # very same kernel arguments stay valid when the step is replayed from a CUDA graph
        self.epoch_t = torch.zeros(1, dtype=torch.int32, device=dev)
        # reduce-scatter grid: 2 CTAs per SM measured best next to backward on 8 GPUs (14.2 ms/step vs 14.7 with
        # 148 and 15.4 with one CTA per chunk); HETSEQ_RS_BLOCKS pins another size, -1 = one CTA per owned chunk
        # (2 GPUs: 148 CTAs 13.85 ms vs 296 CTAs 14.24 ms -- with one peer there are few loads in flight per
        # thread anyway and the second CTA per SM only takes slots from backward)
        rb = int(os.environ.get('HETSEQ_RS_BLOCKS', '0') or getattr(args, 'fused_rs_blocks', 0)
                 or (148 if W <= 2 else 296))
        self.rs_blocks = 0 if rb < 0 else rb
        torch.cuda.synchronize()
        dist.barrier(group=self.group, device_ids=[dev.index])

    def _table(self, everyone, key, local: torch.Tensor):
        W = self.world_size
        arr = (ctypes.c_void_p * W)()
        for p in range(W):
            if p == self.rank:
                arr[p] = local.data_ptr()
            else:
                ptr = ctypes.c_void_p()
                lib.check(self.L.hs_ipc_open(everyone[p][key], ctypes.byref(ptr)), 'hs_ipc_open')
                self._opened.append(ptr.value)
                arr[p] = ptr.value
        return arr

    # -- protocol --------------------------------------------------------------------------------------
    def broadcast_params(self):
        dist.broadcast(self.buffer.flat_param, src=0, group=self.group)
        torch.cuda.synchronize()

    def begin_update(self, num_micro_batches: int):
        self.epoch_t.add_(1)
        self._handles = []

    def _launch(self, b: int):
        """Queue bucket b's in-kernel reduce-scatter behind everything enqueued so far on the main stream."""
        s, e = self.buckets[b]
        self._launched[b] = True
        ev = torch.cuda.Event()
        ev.record()
        self.comm_stream.wait_event(ev)
        lib.check(self.L.hs_dp_reduce_scatter(self.grad_ptrs, self.sig_ptrs, self.rank, self.world_size, s // CHUNK,
                                              e // CHUNK, CHUNK, b, 0, self.epoch_t.data_ptr(), self.sumsq.data_ptr(), self.rs_blocks,
                                              self.comm_stream.cuda_stream), 'hs_dp_reduce_scatter')

    def finish_gradients(self, optimizer):
        for b in range(len(self.buckets)):
            if not self._launched[b]:
                self._launch(b)
        self._armed = False
        torch.cuda.current_stream().wait_stream(self.comm_stream)
        optimizer.multiply_grads(1.0 / self.world_size)

    def norm_exchange(self, mult_host: float, mult_dev: Optional[torch.Tensor], max_norm: float):
        """Global gradient norm + clip coefficient + overflow flag, identical on every rank. Returns the
        0-dim norm tensor; ``self.scalars`` / ``self.skip`` feed the update kernel."""
        md = None if mult_dev is None else mult_dev.to(torch.float32).reshape(1).contiguous()
        self._mult_keepalive = md
        lib.check(self.L.hs_dp_norm_exchange(self.sig_ptrs, self.rank, self.world_size, 0, self.epoch_t.data_ptr(),
                                             self.sumsq.data_ptr(),
                                             float(mult_host), lib.ptr(md), float(max_norm), self.scalars.data_ptr(),
                                             self.skip.data_ptr(), lib.stream_ptr()), 'hs_dp_norm_exchange')
        return self.scalars[0]

    def adam_allgather(self, opt, step_factor: float, hyper_dev: Optional[torch.Tensor] = None):
        h = opt.hyper
        b1, b2 = h['betas']
        buf = self.buffer
        lib.check(self.L.hs_dp_adam_allgather(
            buf.flat_grad.data_ptr(), opt.master.data_ptr(), opt.state['exp_avg'].data_ptr(),
            opt.state['exp_avg_sq'].data_ptr(), self.param_ptrs, self.sig_ptrs, self.rank, self.world_size,
            buf.num_blocks, CHUNK, lib.dtype_code(buf.param_dtype), float(h['lr']), float(b1), float(b2), float(h['eps']),
            float(h['weight_decay']), float(step_factor), self.scalars.data_ptr(), lib.ptr(hyper_dev), 0,
            self.epoch_t.data_ptr(), self.done_counter.data_ptr(), 0, lib.stream_ptr()), 'hs_dp_adam_allgather')
        lib.check(self.L.hs_dp_wait_params(self.sig_ptrs, self.rank, self.world_size, 0, self.epoch_t.data_ptr(),
                                           lib.stream_ptr()), 'hs_dp_wait_params')

    # -- sharded-optimizer interface (checkpoint gathering etc.) ---------------------------------------------
    def all_gather_compact(self, compact: torch.Tensor) -> torch.Tensor:
        compact = compact.contiguous()
        out = torch.empty(self.world_size, compact.numel(), dtype=compact.dtype, device=compact.device)
        dist.all_gather_into_tensor(out.view(-1), compact, group=self.group)
        return out

    def all_gather_params(self, buffer: FlatParamBuffer):
        mine = buffer.owned(buffer.flat_param).reshape(-1).contiguous()
        buffer.flat_param.copy_(buffer.compact_to_flat(self.all_gather_compact(mine)))

    def close(self):
        try:
            torch.cuda.synchronize()
            if dist.is_initialized():
                dist.barrier(group=self.group, device_ids=[self.buffer.device.index])
            for p in self._opened:
                self.L.hs_ipc_close(p)
            self._opened = []
        except Exception:
            pass
