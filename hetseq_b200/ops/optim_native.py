"""Binding of the fused flat-buffer optimizer kernels (csrc/optim_sm100.cu) to ``_Optimizer``."""
from __future__ import annotations

import math

import torch

from hetseq_b200.ops import lib
from hetseq_b200.parallel.flat_buffer import CHUNK


class NativeFlatUpdate(object):
    """Runs ``clip_grad_norm``'s reduction and ``step``'s update for one optimizer as native kernels.

    Layout handed to the kernels: (nchunks, CHUNK, stride) with stride = CHUNK * world_size when the
    optimizer is sharded (base pointers offset by rank * CHUNK) and CHUNK otherwise.
    """

    def __init__(self, opt):
        self.opt = opt
        self.lib = lib.load()
        buf = opt.buffer
        if buf.flat_grad.dtype != torch.float32:
            raise TypeError('native optimizer kernels expect an fp32 flat gradient buffer')
        self.world = buf.world_size if opt.sharded else 1
        self.rank = buf.rank if opt.sharded else 0
        self.nchunks = buf.num_blocks if opt.sharded else buf.total // CHUNK
        self.stride = CHUNK * self.world
        dev = buf.device
        self.scratch = torch.empty(int(self.lib.hs_sumsq_scratch_elems()), dtype=torch.float32, device=dev)
        self.sumsq_out = torch.zeros(1, dtype=torch.float32, device=dev)
        self.kind = 'adam' if 'exp_avg' in opt.state else 'adadelta'

    def _base(self, flat: torch.Tensor) -> int:
        return flat.data_ptr() + self.rank * CHUNK * flat.element_size()

    def sumsq(self, g_view: torch.Tensor) -> torch.Tensor:
        g = self.opt.buffer.flat_grad
        lib.check(self.lib.hs_sumsq(self._base(g), lib.dtype_code(g.dtype), self.nchunks, CHUNK, self.stride,
                                    self.scratch.data_ptr(), self.sumsq_out.data_ptr(), 0, lib.stream_ptr()),
                  'hs_sumsq')
        return self.sumsq_out[0].clone()

    def update(self, mult, skip=None, hyper_dev=None) -> None:
        """``skip``: 0-dim bool tensor; when true on the device the kernel leaves everything untouched.
        ``hyper_dev``: device [lr, step_factor] overriding the host values (CUDA-graph replays)."""
        opt, buf = self.opt, self.opt.buffer
        skip_i = None if skip is None else skip.to(torch.int32).reshape(1)
        h = opt.hyper
        mult_dev = None
        mult_host = 1.0
        if torch.is_tensor(mult):
            mult_dev = mult.to(torch.float32).reshape(1).contiguous()
        else:
            mult_host = float(mult)
        master = opt.master
        out_dtype = lib.dtype_code(buf.param_dtype)
        args_common = (self._base(buf.flat_grad), master.data_ptr())
        if self.kind == 'adam':
            b1, b2 = h['betas']
            t = opt.step_count + 1      # the update being queued (finish_update() commits the count)
            step_factor = math.sqrt(1 - b2 ** t) / (1 - b1 ** t)
            rc = self.lib.hs_adam_flat(
                *args_common, opt.state['exp_avg'].data_ptr(), opt.state['exp_avg_sq'].data_ptr(),
                self._base(buf.flat_param), out_dtype, self.nchunks, CHUNK, self.stride,
                float(h['lr']), float(b1), float(b2), float(h['eps']), float(h['weight_decay']), float(step_factor),
                mult_host, lib.ptr(mult_dev), lib.ptr(hyper_dev), lib.ptr(skip_i), lib.stream_ptr())
            lib.check(rc, 'hs_adam_flat')
        else:
            rc = self.lib.hs_adadelta_flat(
                *args_common, opt.state['square_avg'].data_ptr(), opt.state['acc_delta'].data_ptr(),
                self._base(buf.flat_param), out_dtype, self.nchunks, CHUNK, self.stride,
                float(h['lr']), float(h['rho']), float(h['eps']), float(h['weight_decay']),
                mult_host, lib.ptr(mult_dev), lib.ptr(hyper_dev), lib.ptr(skip_i), lib.stream_ptr())
            lib.check(rc, 'hs_adadelta_flat')
        if opt.sharded:
            opt.dp.all_gather_params(buf)
