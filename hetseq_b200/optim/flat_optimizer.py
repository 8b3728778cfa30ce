"""Optimizer wrappers over flat buffers.

Public API = the reference's ``_Optimizer`` (optim.py:6-80): ``get_lr/set_lr/state_dict/
load_state_dict/backward/multiply_grads/clip_grad_norm/step/zero_grad`` and the concrete
``_Adam`` (optim.py:83-108) / ``_Adadelta`` (optim.py:110-131). The implementation is different:

* gradients live in one flat fp32 buffer (parallel/flat_buffer.py);
* ``multiply_grads`` and ``clip_grad_norm`` do not touch gradient memory -- they fold into one
  scalar that the update applies while it reads the gradient (the reference makes two extra
  passes over every tensor, optim.py:59-70);
* the update is one pass over flat memory -- a native fused kernel on CUDA (csrc/optim_sm100.cu)
  or the same math in flat torch ops elsewhere -- with fp32 master weights and state when the
  model is fp16/bf16, and optional dynamic loss scaling whose overflow check rides on the
  gradient-norm reduction;
* with a sharded data-parallel backend each rank keeps state for, and updates, only its 1/W shard.

``state_dict()`` still produces the reference's per-parameter layout, so
checkpoints are interchangeable with a plain per-tensor ``Adam``.
"""
from __future__ import annotations

import math
from typing import Dict, Iterable, Optional

import torch

from hetseq_b200.options import literal
from hetseq_b200.parallel.flat_buffer import FlatParamBuffer
from hetseq_b200.registry import register_optimizer


class DynamicLossScaler(object):
    """fairseq-style dynamic loss scale. The reference has none (options.py:11 is commented out and
    nothing ever raises the OverflowError that controller.py:368-371 catches); this is the
    north-star ``--fp16`` capability."""

    def __init__(self, init_scale=2. ** 15, scale_factor=2., scale_window=1000, min_loss_scale=1e-4):
        self.loss_scale = float(init_scale)
        self.scale_factor = float(scale_factor)
        self.scale_window = int(scale_window)
        self.min_loss_scale = float(min_loss_scale)
        self._since_overflow = 0

    def update(self, overflow: bool) -> None:
        if overflow:
            self.loss_scale = max(self.loss_scale / self.scale_factor, 0.0)
            self._since_overflow = 0
        else:
            self._since_overflow += 1
            if self._since_overflow >= self.scale_window:
                self.loss_scale *= self.scale_factor
                self._since_overflow = 0

    def state_dict(self):
        return {'loss_scale': self.loss_scale, 'since_overflow': self._since_overflow}

    def load_state_dict(self, sd):
        self.loss_scale = float(sd['loss_scale'])
        self._since_overflow = int(sd.get('since_overflow', 0))


class _Optimizer(object):
    """Flat-buffer optimizer with the reference wrapper API."""

    #: names of the fp32 state buffers, in checkpoint order
    state_names = ()
    #: which update rule the native kernels run ('adam' | 'adadelta')
    kind = ''

    def __init__(self, args, params, dp=None):
        super().__init__()
        self.args = args
        if isinstance(params, FlatParamBuffer):
            self.buffer = params
        else:
            self.buffer = FlatParamBuffer(list(params))
        self.dp = dp
        self.sharded = bool(dp is not None and getattr(dp, 'sharded', False))
        buf = self.buffer
        n = buf.shard_size if self.sharded else buf.total
        dev = buf.device
        self.hyper: Dict = dict(self.optimizer_config)
        # compact fp32 state for the owned elements (all of them when not sharded)
        self.state = {name: self._like_owned(torch.zeros(n, dtype=torch.float32, device=dev))
                      for name in self.state_names}
        # fp32 master copy of the owned elements; aliases the parameters themselves when that is free
        if buf.param_dtype == torch.float32 and not self.sharded:
            self.master = buf.flat_param
        else:
            self.master = self._owned(buf.flat_param).float().contiguous()
        self.step_count = 0
        self._mult = 1.0          # pending host-side gradient multiplier
        self._mult_dev = None     # pending device-side multiplier (0-dim tensor) if any
        self._skip = None         # 0-dim bool tensor: non-finite gradients -> skip this update
        self._scaled = False      # gradients in the buffer carry the loss scale (divide it out in the update)
        self.scaler: Optional[DynamicLossScaler] = None
        if getattr(args, 'fp16', False):
            self.scaler = DynamicLossScaler(args.fp16_init_scale, args.fp16_scale_factor,
                                            args.fp16_scale_window, args.min_loss_scale)
        # Device mirrors of the host scalars that change from update to update. Kernels and tensor ops
        # read these instead of baked-in Python floats, so one captured CUDA graph stays valid:
        #   scale_t  = [loss scale]           hyper_t = [lr, sqrt(1-b2^t)/(1-b1^t)]
        self.scale_t = torch.full((1,), self.loss_scale, dtype=torch.float32, device=dev)
        self.hyper_t = torch.zeros(2, dtype=torch.float32, device=dev)
        #: what hyper_t held during the previous update: the fused backend applies update n while step n+1 runs
        self.hyper_prev_t = torch.zeros(2, dtype=torch.float32, device=dev)
        self._ss_dev = None       # this rank's sample size (device scalar) for the in-kernel gradient denominator
        self._host_scalars = torch.zeros(3, dtype=torch.float32)
        if dev.type == 'cuda':
            self._host_scalars = self._host_scalars.pin_memory()
        self._native = None
        self._init_native()
        if getattr(dp, 'fused', False):
            dp.bind_optimizer(self)

    # -- configuration --------------------------------------------------------------------------
    @property
    def optimizer_config(self) -> Dict:
        raise NotImplementedError

    @staticmethod
    def add_args(parser):
        pass

    def _init_native(self):
        """Bind the fused CUDA kernels when the buffers live on a GPU and the library is usable."""
        from hetseq_b200.ops import dispatch
        if dispatch.use_native(self.args, self.buffer.device):
            from hetseq_b200.ops import optim_native
            self._native = optim_native.NativeFlatUpdate(self)

    @property
    def params(self):
        return iter(self.buffer.params)

    @property
    def optimizer(self):
        """The object that carries ``param_groups``-like hyper-parameters (here: this wrapper)."""
        return self

    @property
    def param_groups(self):
        return [self.hyper]

    def get_lr(self):
        return self.hyper['lr']

    def set_lr(self, lr):
        self.hyper['lr'] = lr

    @property
    def loss_scale(self) -> float:
        return self.scaler.loss_scale if self.scaler is not None else 1.0

    # -- gradient handling ----------------------------------------------------------------------
    def sync_device_scalars(self):
        """Push lr, the bias-correction factor of the NEXT update and the loss scale to the device: the
        host values go into a pinned staging buffer, two tiny async copies move them. Runs at the start
        of every update (zero_grad). When the update is replayed from a CUDA graph the copies are part of
        the graph and only ``write_host_scalars`` runs on the host before each replay."""
        self.write_host_scalars()
        self.hyper_prev_t.copy_(self.hyper_t)
        self.hyper_t.copy_(self._host_scalars[:2], non_blocking=True)
        self.scale_t.copy_(self._host_scalars[2:3], non_blocking=True)

    def write_host_scalars(self):
        t = self.step_count + 1
        step_factor = 1.0
        if 'betas' in self.hyper:
            b1, b2 = self.hyper['betas']
            step_factor = math.sqrt(1 - b2 ** t) / (1 - b1 ** t)
        self._host_scalars[0] = float(self.hyper['lr'])
        self._host_scalars[1] = step_factor
        self._host_scalars[2] = self.loss_scale

    def backward(self, loss):
        """loss.backward() (optim.py:55-57), scaled by the current loss scale under --fp16. The scale is
        read from device memory (``scale_t``)."""
        if self.scaler is not None:
            self._scaled = True
            loss = loss * self.scale_t[0]
        self.buffer.begin_backward()
        loss.backward()

    def set_grad_denominator(self, sample_size):
        """This rank's sample size as a device scalar. The fused backend sums it over ranks inside the norm
        exchange and divides the summed gradients by it (controller.py:338-339 without a collective or a host
        read); the other backends get an explicit ``multiply_grads(world / sample_size)`` instead."""
        self._ss_dev = sample_size

    def multiply_grads(self, c):
        """Multiply gradients by ``c`` (float or 0-dim tensor) -- deferred into the update."""
        if torch.is_tensor(c):
            self._mult_dev = c if self._mult_dev is None else self._mult_dev * c
        else:
            self._mult *= float(c)

    def _pending_mult(self):
        """Host factor x device factor (clip coefficient, world/sample_size, 1/loss_scale)."""
        md = self._mult_dev
        if getattr(self, '_scaled', False):
            inv = 1.0 / self.scale_t[0]
            md = inv if md is None else md.to(torch.float32) * inv
        if md is not None:
            return md.to(torch.float32) * self._mult
        return self._mult

    def _owned(self, flat: torch.Tensor) -> torch.Tensor:
        """Owned elements of a flat buffer: [num_blocks, CHUNK] strided view, or the buffer itself."""
        return self.buffer.owned(flat) if self.sharded else flat

    def _like_owned(self, compact: torch.Tensor) -> torch.Tensor:
        from hetseq_b200.parallel.flat_buffer import CHUNK
        return compact.view(-1, CHUNK) if self.sharded else compact

    def _local_sumsq(self) -> torch.Tensor:
        take = getattr(self.dp, 'take_local_sumsq', None)     # measured bucket by bucket during backward
        if take is not None:
            done = take()
            if done is not None:
                return done
        g = self._owned(self.buffer.flat_grad)
        if self._native is not None:
            return self._native.sumsq(g)
        g = g.float()
        return (g * g).sum()

    def clip_grad_norm(self, max_norm):
        """Global L2 norm of the (multiplied, unscaled) gradients; folds the clip coefficient into
        the pending multiplier. Mirrors optim.py:65-70 / torch.nn.utils.clip_grad_norm_
        (coef = max_norm / (norm + 1e-6), clamped to 1). Returns a 0-dim tensor; no host sync."""
        if getattr(self.dp, 'fused', False):
            # norm partials were accumulated by the reduce-scatter kernels; one exchange kernel gives
            # every rank the same norm / clip coefficient / overflow flag
            md = self._pending_mult()
            norm = self.dp.norm_exchange(1.0 if torch.is_tensor(md) else float(md), md if torch.is_tensor(md) else None,
                                         float(max_norm), sample_size=self._ss_dev)
            self._skip = self.dp.scalars[2] != 0
            self._fused_ready = True
            return norm
        sumsq = self._local_sumsq()
        if self.sharded:
            sumsq = self.dp.all_reduce_scalar(sumsq)
        mult = self._pending_mult()
        norm = torch.sqrt(sumsq) * (mult.abs() if torch.is_tensor(mult) else abs(mult))
        self._skip = ~torch.isfinite(norm)
        if max_norm > 0:
            coef = torch.clamp(max_norm / (norm + 1e-6), max=1.0)
            coef = torch.where(self._skip, torch.ones_like(coef), coef)
            self._mult_dev = coef if self._mult_dev is None else self._mult_dev * coef
        return norm

    def zero_grad(self):
        self.buffer.zero_grad()
        self.sync_device_scalars()
        self._mult, self._mult_dev, self._skip = 1.0, None, None
        self._fused_ready = False
        self._scaled = False
        self._ss_dev = None

    # -- update ---------------------------------------------------------------------------------
    def step(self, closure=None):
        """launch_update() + finish_update(): the reference's ``optimizer.step()``. Under --fp16 a
        non-finite gradient norm leaves parameters and state untouched, halves the loss scale and raises
        OverflowError like fairseq, so that Controller.train_step's handler (reference
        controller.py:368-371) finally has a raiser."""
        self.launch_update()
        overflow = bool(self._skip.item()) if self.scaler is not None else False
        self.finish_update(overflow)

    def launch_update(self):
        """Queue the update on the device; nothing here synchronises or depends on host-side values that
        change between updates (safe to capture in a CUDA graph). The kernels take the overflow flag
        from device memory and skip themselves when it is set."""
        fused = getattr(self.dp, 'fused', False)
        if self._skip is None or (fused and not getattr(self, '_fused_ready', False)):
            self.clip_grad_norm(0.0)
        if fused:
            self.dp.queue_update(self)      # applied next to the next forward pass (or by dp.flush())
        elif self._native is not None:
            self._native.update(self._pending_mult(), self._skip if self.scaler is not None else None,
                                hyper_dev=self.hyper_t)
        else:
            # PyTorch path (CPU / --kernels torch): needs the host value of the overflow flag first
            overflow = bool(self._skip.item()) if self.scaler is not None else False
            if not overflow:
                self.step_count += 1
                mult = self._pending_mult()
                g = self._owned(self.buffer.flat_grad)
                out = self._owned(self.buffer.flat_param)
                g32 = g.float() * mult if (torch.is_tensor(mult) or mult != 1.0) else g.float()
                self._update_torch(g32)
                if self.master is not self.buffer.flat_param:
                    out.copy_(self.master)
                if self.sharded:
                    self.dp.all_gather_params(self.buffer)
                self.step_count -= 1

    def finish_update(self, overflow: bool):
        """Host bookkeeping once the overflow flag of the queued update is known."""
        self._mult, self._mult_dev, self._skip = 1.0, None, None
        self._scaled = False
        if self.scaler is not None:
            self.scaler.update(overflow)
            if overflow:
                if self.scaler.loss_scale <= self.scaler.min_loss_scale:
                    raise FloatingPointError(
                        'Minimum loss scale reached ({}). Your loss is probably exploding.'.format(
                            self.scaler.min_loss_scale))
                raise OverflowError('setting loss scale to: ' + str(self.scaler.loss_scale))
        self.step_count += 1

    def _update_torch(self, g32: torch.Tensor) -> None:
        raise NotImplementedError

    # -- (de)serialisation in the reference's per-parameter schema ---------------------------------
    def _full_state(self, name: str) -> torch.Tensor:
        """State buffer in natural flat order (collective when sharded)."""
        t = self.state[name]
        if self.sharded:
            t = self.buffer.compact_to_flat(self.dp.all_gather_compact(t.reshape(-1)))
        return t

    def state_dict(self):
        """{'state': {idx: {'step', <state tensors shaped like the parameter>}}, 'param_groups': [...]}
        -- what ``torch.optim.Optimizer.state_dict()`` yields for the reference's Adam/Adadelta.
        Collective when sharded (every rank must call it)."""
        buf = self.buffer
        full = {name: self._full_state(name) for name in self.state_names}
        state = {}
        if self.step_count > 0:
            for i, p in enumerate(buf.params):
                off, n = buf.offsets[i], buf.numels[i]
                entry = {'step': self.step_count}
                for name in self.state_names:
                    entry[name] = full[name][off:off + n].view(p.shape).clone()
                state[i] = entry
        group = dict(self.hyper)
        group['params'] = list(range(len(buf.params)))
        out = {'state': state, 'param_groups': [group]}
        if self.scaler is not None:
            out['loss_scaler'] = self.scaler.state_dict()
        return out

    def load_state_dict(self, state_dict, optimizer_overrides=None):
        """Load a per-parameter state dict; the live hyper-parameters win over the stored ones
        (same policy as optim.py:41-53), then ``optimizer_overrides`` are applied."""
        buf = self.buffer
        st = state_dict.get('state', {})
        steps = [int(v['step']) for v in st.values() if 'step' in v]
        self.step_count = max(steps) if steps else 0
        for name in self.state_names:
            full = torch.zeros(buf.total, dtype=torch.float32)
            for key, entry in st.items():
                i = int(key)
                if i >= len(buf.params):
                    raise ValueError('optimizer state has more parameters than the model')
                off, n = buf.offsets[i], buf.numels[i]
                if entry[name].numel() != n:
                    raise ValueError('optimizer state for parameter {} has {} elements, expected {}'.format(
                        i, entry[name].numel(), n))
                full[off:off + n].copy_(entry[name].reshape(-1).to(torch.float32))
            self.state[name].copy_(self._owned(full))
        groups = state_dict.get('param_groups') or [{}]
        stored = {k: v for k, v in groups[0].items() if k != 'params'}
        for k, v in stored.items():
            self.hyper.setdefault(k, v)
        if 'loss_scaler' in state_dict and self.scaler is not None:
            self.scaler.load_state_dict(state_dict['loss_scaler'])
        if optimizer_overrides:
            self.hyper.update(optimizer_overrides)
        self.refresh_master()

    def refresh_master(self):
        """Re-derive the fp32 master copy from the (possibly just loaded) model parameters."""
        if self.master is not self.buffer.flat_param:
            self.master.copy_(self._owned(self.buffer.flat_param))


@register_optimizer('adam')
class _Adam(_Optimizer):
    """fairseq Adam (AdamW-style decoupled decay, bias correction) -- math of optim.py:162-231."""

    state_names = ('exp_avg', 'exp_avg_sq')
    kind = 'adam'

    @staticmethod
    def add_args(parser):
        parser.add_argument('--adam-betas', default='(0.9, 0.999)', metavar='B', help='betas for Adam optimizer')
        parser.add_argument('--adam-eps', type=float, default=1e-8, metavar='D', help='epsilon for Adam optimizer')
        parser.add_argument('--weight-decay', '--wd', default=0.0, type=float, metavar='WD', help='weight decay')

    @property
    def optimizer_config(self):
        betas = self.args.adam_betas
        if isinstance(betas, str):
            betas = literal(betas)
        return {'lr': self.args.lr[0], 'betas': tuple(betas), 'eps': self.args.adam_eps,
                'weight_decay': self.args.weight_decay, 'amsgrad': False}

    def _update_torch(self, g32):
        h = self.hyper
        beta1, beta2 = h['betas']
        m, v, p32 = self.state['exp_avg'], self.state['exp_avg_sq'], self.master
        m.mul_(beta1).add_(g32, alpha=1 - beta1)
        v.mul_(beta2).addcmul_(g32, g32, value=1 - beta2)
        denom = v.sqrt().add_(h['eps'])
        bc1 = 1 - beta1 ** self.step_count
        bc2 = 1 - beta2 ** self.step_count
        step_size = h['lr'] * math.sqrt(bc2) / bc1
        if h['weight_decay'] != 0:
            p32.add_(p32, alpha=-h['weight_decay'] * h['lr'])
        p32.addcdiv_(m, denom, value=-step_size)


@register_optimizer('adadelta')
class _Adadelta(_Optimizer):
    """Adadelta with coupled L2 decay -- math of optim.py:263-304."""

    state_names = ('square_avg', 'acc_delta')
    kind = 'adadelta'

    @staticmethod
    def add_args(parser):
        parser.add_argument('--adadelta_rho', '--adadelta-rho', default=0.9, type=float)
        parser.add_argument('--adadelta_eps', '--adadelta-eps', default=1e-6, type=float)
        # (sic) the reference spells it without the leading 'a' (options.py:181)
        parser.add_argument('--dadelta_weight_decay', '--adadelta-weight-decay', default=0.0, type=float)

    @property
    def optimizer_config(self):
        return {'lr': self.args.lr[0], 'rho': self.args.adadelta_rho, 'eps': self.args.adadelta_eps,
                'weight_decay': self.args.dadelta_weight_decay}

    def _update_torch(self, g32):
        h = self.hyper
        rho, eps = h['rho'], h['eps']
        sq, acc, p32 = self.state['square_avg'], self.state['acc_delta'], self.master
        if h['weight_decay'] != 0:
            g32 = g32.add(p32, alpha=h['weight_decay'])
        sq.mul_(rho).addcmul_(g32, g32, value=1 - rho)
        std = sq.add(eps).sqrt_()
        delta = acc.add(eps).sqrt_().div_(std).mul_(g32)
        p32.add_(delta, alpha=-h['lr'])
        acc.mul_(rho).addcmul_(delta, delta, value=1 - rho)
