"""Controller: the training core (reference: controller.py:21-439).

Same public surface as the reference -- ``train_step``, ``get_train_iterator``,
``save_checkpoint``/``load_checkpoint``, ``lr_step``, ``get_meter``, lazy ``model`` / ``optimizer`` /
``lr_scheduler`` properties -- over a different engine:

* the model is never wrapped in DDP; its parameters are views into a ``FlatParamBuffer`` and a
  data-parallel *backend* (parallel/data_parallel.py, parallel/fused_dp.py) turns the flat gradients
  of all ranks into the update (bucket-wise, overlapped with backward);
* gradient scaling, clipping and the optimizer update are one pass (optim/flat_optimizer.py);
* per-step statistics stay on the device until ONE device->host read at the end of the step; the
  reference synchronises mid-step (``.tolist()`` before ``multiply_grads``, controller.py:304) and the
  slow stats path reads the byte buffer element by element (distributed_utils.py:120-122);
* ``sample_size`` / ``nsentences`` / ``ntokens`` are real numbers;
  ``--sample-size-mode reference`` restores the reference's gradient denominator;
* dropout seeding keeps the resume contract ``seed + num_updates`` (controller.py:426-432) and
  additionally folds in the rank so replicas draw different masks (Q14).
"""
from __future__ import annotations

import contextlib
import math
import os
from collections import OrderedDict
from itertools import chain

import torch

from hetseq_b200 import registry
from hetseq_b200.parallel import data_parallel, distributed_utils
from hetseq_b200.parallel.flat_buffer import FlatParamBuffer
from hetseq_b200.utils import checkpoint_utils, nvtx
from hetseq_b200.utils.meters import AverageMeter, CudaEventTimer, StopwatchMeter, TimeMeter
from hetseq_b200.utils.utils import PackedSample


class Controller(object):
    """Synchronous data-parallel trainer for one model replica per process / GPU."""

    def __init__(self, args, task, model, criterion=None, dummy_batch=None, oom_batch=None):
        self.args = args
        self.task = task

        self.cuda = torch.cuda.is_available() and not args.cpu
        self.device = torch.device('cuda', torch.cuda.current_device()) if self.cuda else torch.device('cpu')
        if getattr(args, 'fp16', False):
            model = model.half()
        elif getattr(args, 'bf16', False):
            model = model.bfloat16()
        self._model = model.to(self.device)

        self._dummy_batch = dummy_batch
        self._oom_batch = oom_batch or dummy_batch
        self._lr_scheduler = None
        self._num_updates = 0
        self._optim_history = None
        self._optimizer = None
        self._buffer = None
        self._dp = None
        self._prev_grad_norm = None
        self._native_attached = False

        self.fast_stat_sync = args.fast_stat_sync
        self.step_timer = CudaEventTimer(use_cuda=self.cuda)
        # HETSEQ_PHASE_TIMING=1: CUDA-event timing of the phases of every update (printed by close())
        self._phase_timing = self.cuda and bool(int(os.environ.get('HETSEQ_PHASE_TIMING', '0')))
        nvtx.enable(getattr(args, 'profile_nvtx', False) and self.cuda)
        self._phase_events = []
        # --cuda-graph: capture the device part of an update once, replay it afterwards
        self._graph_enabled = bool(getattr(args, 'cuda_graph', False)) and self.cuda
        self._graph, self._graph_key, self._graph_warmup, self._capturing = None, None, 0, False
        self._static_sample = None
        self._static_sample, self._static_out = None, None
        self._stats_host = None
        #: updates skipped because of non-finite gradients (fp16 loss scaling); bench.py reports it
        self.overflow_updates = 0
        if self._graph_enabled and getattr(args, 'max_sentences', None):
            # static shapes: short last batches and padding batches are grown to full size on the host
            if hasattr(task, 'static_batch_size'):
                task.static_batch_size = int(args.max_sentences)
        self.init_meters(args)

    # -- meters -----------------------------------------------------------------------------------
    #: name -> (meter class, what it measures); order = order of the columns in the training log line. 'oom' is
    #: looked up by the reference's train loop but never created there; 'loss_scale' only exists under --fp16.
    METER_SCHEMA = (
        ('train_loss', AverageMeter, 'training loss per sample, base 2'),
        ('train_nll_loss', AverageMeter, 'same, negative log-likelihood part'),
        ('valid_loss', AverageMeter, 'validation loss'),
        ('valid_nll_loss', AverageMeter, 'validation nll'),
        ('wps', TimeMeter, 'words per second'),
        ('ups', TimeMeter, 'updates per second'),
        ('wpb', AverageMeter, 'words per batch'),
        ('bsz', AverageMeter, 'sentences per batch'),
        ('gnorm', AverageMeter, 'gradient norm'),
        ('clip', AverageMeter, 'share of updates that were clipped'),
        ('oom', AverageMeter, 'out-of-memory batches'),
        ('loss_scale', AverageMeter, 'dynamic loss scale'),
        ('wall', TimeMeter, 'wall time in seconds'),
        ('train_wall', StopwatchMeter, 'time spent inside train_step'),
    )

    def init_meters(self, args):
        fp16 = bool(getattr(args, 'fp16', False))
        self.meters = OrderedDict((name, cls()) for name, cls, _ in self.METER_SCHEMA
                                  if name != 'loss_scale' or fp16)

    # -- lazily built engine ------------------------------------------------------------------------
    @property
    def model(self):
        """The replica itself: there is no wrapper module (the reference returns a DDP wrapper)."""
        if self._optimizer is None:
            self._build_optimizer()
        return self._model

    @property
    def optimizer(self):
        if self._optimizer is None:
            self._build_optimizer()
        return self._optimizer

    @property
    def lr_scheduler(self):
        if self._lr_scheduler is None:
            self._build_optimizer()
        return self._lr_scheduler

    @property
    def buffer(self) -> FlatParamBuffer:
        if self._buffer is None:
            self._build_optimizer()
        return self._buffer

    @property
    def dp(self):
        if self._dp is None:
            self._build_optimizer()
        return self._dp

    def _trainable_params(self):
        model = self._model
        if hasattr(model, 'param_order'):
            params = model.param_order()
        else:
            params = list(model.parameters())
        return [p for p in params if p.requires_grad]

    def _build_optimizer(self):
        args = self.args
        world = args.distributed_world_size if torch.distributed.is_initialized() else 1
        rank = args.distributed_rank if world > 1 else 0

        if self._buffer is None:
            allocator = None
            kind = getattr(args, 'dp_backend', 'auto')
            if world > 1 and self.cuda and kind in ('auto', 'fused'):
                from hetseq_b200.parallel import fused_dp
                if kind == 'fused' or fused_dp.is_supported(args):
                    allocator = fused_dp.symmetric_allocator(args)
            canonical = [p for p in self._model.parameters() if p.requires_grad]
            self._buffer = FlatParamBuffer(self._trainable_params(), world_size=world, rank=rank,
                                           allocator=allocator)
            pos = {id(p): i for i, p in enumerate(canonical)}
            self._buffer.canonical_index = [pos[id(p)] for p in self._buffer.params]
            self._dp = data_parallel.build_backend(args, self._buffer)
            if self._graph_enabled and self._dp.name in ('nccl', 'sharded-allreduce'):
                print('| --cuda-graph is only used with the local and fused-nvlink gradient paths; running eagerly')
                self._graph_enabled = False
            self._dp.broadcast_params()
            self._attach_native()

        self._optimizer = registry.lookup('optimizer', args.optimizer)(args, self._buffer, dp=self._dp)
        # the LR scheduler is built right after the optimizer so that the initial LR is set
        self._lr_scheduler = registry.lookup('lr_scheduler', args.lr_scheduler)(args, self._optimizer)
        self._lr_scheduler.step_update(0)

    def _attach_native(self):
        """Swap the model's hot ops for the sm_100a kernels when running on a GPU."""
        if self._native_attached:
            return
        self._native_attached = True
        from hetseq_b200.ops import dispatch
        if dispatch.use_native(self.args, self.device):
            from hetseq_b200.ops import bert_native
            bert_native.attach(self._model, self._buffer, self.args)

    # -- checkpoints ------------------------------------------------------------------------------
    def save_checkpoint(self, filename, extra_state):
        """Write all training state to ``filename``. Every rank must call this when the optimizer
        is sharded (state gathering is collective); only the master writes."""
        extra_state['train_meters'] = self.meters
        self.dp.flush()      # an update that is still riding next to the forward pass must land first
        opt_state = None
        if not self.args.no_save_optimizer_state:
            if getattr(self.optimizer, 'sharded', False) or distributed_utils.is_master(self.args):
                opt_state = self._buffer_to_canonical_order(self.optimizer.state_dict())
        if distributed_utils.is_master(self.args):
            state = checkpoint_utils.build_state(
                self.args, self.get_model().state_dict(), self.optimizer.__class__.__name__, opt_state,
                self.lr_scheduler.state_dict(), self.get_num_updates(), self._optim_history, extra_state)
            checkpoint_utils.torch_persistent_save(state, filename)

    def load_checkpoint(self, filename, reset_optimizer=False, reset_lr_scheduler=False,
                        optimizer_overrides=None, reset_meters=False):
        """Restore training state from ``filename`` (every rank reads the file itself) in three steps -- weights,
        optimizer + schedule, progress -- and return the saved ``extra_state`` (None when there is nothing to
        resume from). Same messages and semantics as the reference (controller.py:138-200)."""
        if self._dp is not None:
            self._dp.flush()
        self._optim_history = []
        if not os.path.exists(filename):
            print('| no existing checkpoint found {}'.format(filename))
            return None
        state = checkpoint_utils.load_checkpoint_to_cpu(filename)
        self._restore_weights(state, filename)
        self._optim_history = state['optimizer_history']
        if state.get('last_optimizer_state') is not None and not reset_optimizer:
            self._restore_optimizer(state['last_optimizer_state'], reset_lr_scheduler, optimizer_overrides)
        return self._restore_progress(state['extra_state'], filename, reset_meters)

    def _restore_weights(self, state, filename):
        try:
            self.get_model().load_state_dict(state['model'], strict=True)
        except Exception:
            raise Exception('Cannot load model parameters from checkpoint {}; please ensure that the '
                            'architectures match.'.format(filename))
        if self._optimizer is not None:
            self._optimizer.refresh_master()

    def _restore_optimizer(self, optim_state, reset_lr_scheduler, overrides):
        self._build_optimizer()   # after the weights: the fp32 master copy is taken from them
        last = self._optim_history[-1]
        assert last['optimizer_name'] == self.optimizer.__class__.__name__, \
            'Optimizer does not match; please reset the optimizer (--reset-optimizer).'
        if not reset_lr_scheduler:
            self.lr_scheduler.load_state_dict(last['lr_scheduler_state'])
        self.optimizer.load_state_dict(self._canonical_to_buffer_order(optim_state), overrides)
        self.set_num_updates(last['num_updates'])

    def _restore_progress(self, extra_state, filename, reset_meters):
        if not extra_state or 'train_iterator' not in extra_state:
            # a file written by the reference itself: extra_state == {} (checkpoint_utils.py:204)
            print('| loaded checkpoint {} without iterator state (@ {} updates)'.format(
                filename, self.get_num_updates()))
            return None
        epoch = extra_state['train_iterator']['epoch']
        print('| loaded checkpoint {} (epoch {} @ {} updates)'.format(filename, epoch, self.get_num_updates()))
        self.lr_step(epoch)
        saved = extra_state.pop('train_meters', None)
        if saved is not None and not reset_meters:
            self.meters.update(saved)
            for meter in self.meters.values():
                if isinstance(meter, TimeMeter):
                    meter.reset()
        return extra_state

    def _buffer_to_canonical_order(self, opt_state):
        canon = getattr(self.buffer, 'canonical_index', None)
        if canon is None or canon == list(range(len(canon))):
            return opt_state
        out = dict(opt_state)
        out['state'] = {canon[i]: v for i, v in opt_state.get('state', {}).items()}
        out['state'] = dict(sorted(out['state'].items()))
        return out

    def _canonical_to_buffer_order(self, opt_state):
        """Checkpoints index optimizer state by ``model.parameters()`` order (what a plain torch
        optimizer would write); the flat buffer may order parameters differently."""
        canon = getattr(self.buffer, 'canonical_index', None)
        if canon is None or canon == list(range(len(canon))):
            return opt_state
        st = opt_state.get('state', {})
        remapped = {}
        for buf_i, can_i in enumerate(canon):
            if can_i in st:
                remapped[buf_i] = st[can_i]
            elif str(can_i) in st:
                remapped[buf_i] = st[str(can_i)]
        out = dict(opt_state)
        out['state'] = remapped
        return out

    # -- data -------------------------------------------------------------------------------------
    def get_train_iterator(self, epoch, combine=True, load_dataset=True):
        if load_dataset:
            print('| loading train data for epoch {}'.format(epoch))
            self.task.load_dataset(self.args.train_subset)
        return self.task.get_batch_iterator(
            dataset=self.task.dataset(self.args.train_subset),
            max_tokens=self.args.max_tokens, max_sentences=self.args.max_sentences, max_positions=None,
            ignore_invalid_inputs=True,
            required_batch_size_multiple=self.args.required_batch_size_multiple,
            seed=self.args.seed, num_shards=self.args.distributed_world_size,
            shard_id=self.args.distributed_rank, num_workers=self.args.num_workers, epoch=epoch)

    # -- the hot loop -------------------------------------------------------------------------------
    def train_step(self, samples, dummy_batch=False, raise_oom=False):
        """forward / backward over the micro-batches of one update, then reduce, clip and step.

        Device work and host bookkeeping are separated: ``_device_step`` only queues work (no host
        synchronisation, no host values that change between updates baked into kernels), which is what
        allows ``--cuda-graph`` to capture it once and replay it; the host then makes ONE read of a small
        result vector and updates meters, loss scale and the LR schedule.

        Every rank queues the same cross-rank operations in the same order whether it replays the graph or
        runs eagerly (with the fused backend: in-kernel reduce-scatter, norm + sample-size exchange and update;
        the statistics all-reduce comes after the step in both modes), so an odd batch on one rank -- a padding
        batch of an uneven shard, a short last batch -- never needs agreement between ranks. Such batches are
        padded to the static shapes and weighted on the device (weight 0 for a padding batch), the invariant of
        the reference's dummy-batch handling (controller.py:237-241)."""
        if self._dummy_batch is None:
            self._dummy_batch = samples[0]

        self._set_seed()
        model = self.model
        model.train()
        if not dummy_batch:
            self.meters['train_wall'].start()
            self.step_timer.start()

        # ---- host -> device ---------------------------------------------------------------------
        prepared, host_stats = [], [0, 0, 0]          # sample_size, nsentences, ntokens
        allow_host = self._graph_enabled and not dummy_batch and len(samples) == 1
        for sample in samples:
            p = self._prepare_sample(sample, allow_host)
            ignore = p is None
            if ignore:
                # uneven shards: run on the cached dummy batch and ignore its gradients so that every rank
                # still takes part in the reduction (controller.py:237-241)
                p = self._prepare_sample(self._dummy_batch, allow_host)
            sample_dev, stats = p
            prepared.append((sample_dev, stats, ignore))
            if not ignore:
                host_stats[0] += self.task.sample_size(sample_dev, stats[0])
                host_stats[1] += stats[0]
                host_stats[2] += stats[1]

        # ---- device ---------------------------------------------------------------------------------
        graphable = self._graph_enabled and not dummy_batch and len(prepared) == 1
        try:
            if dummy_batch:
                # forward + backward only, like the reference, which returns before stats / clip / step
                # (controller.py:288-289); nothing is reduced and nothing is applied
                self._device_step(prepared, host_stats, update=False)
                return None
            if graphable:
                readback, reduced = self._graphed_device_step(prepared[0], host_stats)
            else:
                readback, reduced = self._device_step(prepared, host_stats)
        except RuntimeError as e:
            if 'out of memory' in str(e):
                raise RuntimeError('ran out of memory with exception') from e
            raise

        # ---- the single device->host read of the update --------------------------------------------------
        if not reduced and self._sync_stats():
            readback = self._reduce_logging_stats(readback)
        sample_size, nsentences, loss_total, nll_total, ntokens, ooms, gnorm, skip = readback.tolist()
        opt = self.optimizer
        logging_output = None
        try:
            if skip and opt.scaler is not None:
                self.overflow_updates += 1
            opt.finish_update(bool(skip) and opt.scaler is not None)
            self.set_num_updates(self.get_num_updates() + 1)
            every = int(getattr(self.args, 'check_params_interval', 0) or 0)
            if every > 0 and self._num_updates % every == 0 and not self.dp.check_params_consistent():
                raise RuntimeError('Fatal error: parameters differ between workers after update {}'.format(
                    self._num_updates))
            self.task.update_step(self._num_updates)
            self._prev_grad_norm = gnorm
            # like the fast path of the reference (controller.py:299-304) losses are logged as
            # sum / (sample_size * ln 2); in 'batch' mode the sum is weighted by sentences so this is
            # the true per-sentence mean in base 2
            denom = max(sample_size, 1e-12) * math.log(2)
            loss_avg = loss_total / denom
            logging_output = {'loss': loss_avg, 'nll_loss': nll_total / denom, 'ntokens': ntokens,
                              'nsentences': nsentences, 'sample_size': sample_size}
            self.meters['wps'].update(ntokens)
            self.meters['ups'].update(1.)
            self.meters['wpb'].update(ntokens)
            self.meters['bsz'].update(nsentences)
            self.meters['gnorm'].update(gnorm)
            self.meters['clip'].update(1. if gnorm > self.args.clip_norm and self.args.clip_norm > 0 else 0.)
            self.meters['oom'].update(ooms)
            self.meters['train_loss'].update(loss_avg, sample_size)
            if 'loss_scale' in self.meters:
                self.meters['loss_scale'].reset()
                self.meters['loss_scale'].update(opt.loss_scale)
        except OverflowError as e:
            print('| WARNING: overflow detected, ' + str(e))
            logging_output = None

        self.step_timer.stop()
        self.meters['train_wall'].stop()
        return logging_output

    def _reduce_logging_stats(self, readback):
        """Sum [sample_size, nsentences, loss, nll_loss, ntokens, ooms] over ranks (controller.py:292-334).
        With ``--fast-stat-sync`` one 48-byte all-reduce; otherwise the general path: gather every rank's
        numbers, aggregate them (the reference gathers and then forgets to aggregate) and run the cross-rank
        gradient-norm consistency check on the previous step's norm (controller.py:325-328)."""
        stats = readback[:6]
        if self.fast_stat_sync:
            comm = stats.clone() if stats.device == distributed_utils.comm_device() else stats.cpu()
            torch.distributed.all_reduce(comm)
            stats = comm.to(self.device)
        else:
            gathered = distributed_utils.all_gather_list([stats.tolist(), self._prev_grad_norm])
            prev_norms = [g[1] for g in gathered]
            assert (all(n == prev_norms[0] for n in prev_norms)
                    or all(n is None or math.isnan(n) or math.isinf(n) for n in prev_norms)), \
                'Fatal error: gradients are inconsistent between workers'
            stats = torch.tensor([sum(g[0][k] for g in gathered) for k in range(6)],
                                 dtype=torch.float64, device=self.device)
        return torch.cat([stats, readback[6:]])

    def _device_step(self, prepared, host_stats, update=True, static_weight=False):
        """Queue one whole update on the device. Returns (the 8-element result vector [sample_size, nsentences,
        loss_sum, nll_sum, ntokens, ooms, grad_norm, overflow] (fp64, device), whether its first six entries
        are already summed over ranks)."""
        model, dp, opt = self._model, self.dp, self.optimizer
        fused = bool(getattr(dp, 'fused', False))
        multi = self._sync_stats()
        self.zero_grad()
        marks = [self._mark()] if (self._phase_timing and not self._capturing) else None
        # host integers travel through a pinned staging vector (re-read by a captured graph on each replay)
        if self._stats_host is None:
            self._stats_host = torch.zeros(6, dtype=torch.float64)
            if self.cuda:
                self._stats_host = self._stats_host.pin_memory()
        self._stats_host[0], self._stats_host[1], self._stats_host[4] = host_stats[0], host_stats[1], host_stats[2]
        stats_in = self._stats_host.to(self.device, non_blocking=True)
        batch_mode = self.args.sample_size_mode == 'batch'
        loss_sum = torch.zeros((), dtype=torch.float32, device=self.device)
        if update:
            dp.begin_update(len(prepared))
        for i, (sample_dev, stats, ignore_grad) in enumerate(prepared):
            if len(prepared) > 1:
                from hetseq_b200.ops import rng
                rng.set_micro_batch(i)        # fresh dropout masks per micro-batch, like the reference
            if update:
                dp.begin_micro_batch(i, is_last=(i == len(prepared) - 1))
            else:
                opt.buffer.reset_ready_counts()
            weight = None
            if static_weight:
                # captured graph: the weight of the (single) micro-batch is read from device memory -- the number
                # of real sentences (0 for a padding batch), or 1/0 in 'reference' mode
                weight = (stats_in[1] if batch_mode else (stats_in[1] > 0).to(torch.float64)).to(torch.float32)
                ignore_grad = False
            with nvtx.range('fwd+bwd micro-batch'):
                loss, sample_size, logging_output = self.task.train_step(sample_dev, model, opt, ignore_grad,
                                                                         stats=stats, weight=weight)
            if weight is not None:
                loss_sum = loss_sum + logging_output['loss'].float() * (weight if batch_mode else (weight > 0).float())
            elif not ignore_grad:
                w = logging_output.get('nsentences', 1) if batch_mode else 1.0
                loss_sum = loss_sum + logging_output['loss'].float() * w
        if not update:
            return None, False
        if marks is not None:
            marks.append(self._mark())      # forward + backward enqueued
        opt.buffer.finalize_grads()
        with nvtx.range('gradient reduction (exposed part)'):
            dp.finish_gradients(opt)
        if marks is not None:
            marks.append(self._mark())      # gradient reduction complete

        # ---- stats: [sample_size, nsentences, loss, nll_loss, ntokens, ooms] like controller.py:292-314
        loss64 = loss_sum.double()
        stats_vec = torch.cat([stats_in[:2], loss64.reshape(1), loss64.reshape(1), stats_in[4:]])
        reduced = False
        if fused:
            # the ranks' sample sizes are summed inside the norm-exchange kernel: no collective, no host read
            opt.set_grad_denominator(stats_vec[0])
        else:
            world = 1.0
            if multi:
                stats_vec = self._reduce_logging_stats(torch.cat([stats_vec, stats_vec[:2]]))[:6]
                reduced = True
                world = float(self.args.distributed_world_size)
            sample_size_dev = stats_vec[0]
            # normalise grads by sample size (controller.py:338-339); the backend already made them the mean
            # over ranks, so the factor is world / sample_size like the reference's
            opt.multiply_grads(torch.where(sample_size_dev > 0, world / sample_size_dev.clamp(min=1e-12),
                                           torch.ones_like(sample_size_dev)).float())
        with nvtx.range('grad norm + clip'):
            grad_norm = opt.clip_grad_norm(self.args.clip_norm)
        if marks is not None:
            marks.append(self._mark())      # stats sync + norm
        with nvtx.range('optimizer update (+ all-gather)'):
            opt.launch_update()
        if marks is not None:
            marks.append(self._mark())      # update (+ all-gather)
            self._phase_events.append(marks)
        skip = opt._skip.double().reshape(1) if opt._skip is not None else torch.zeros(1, dtype=torch.float64, device=self.device)
        return torch.cat([stats_vec, grad_norm.detach().double().reshape(1), skip]), reduced

    # -- CUDA graph capture of the device step -----------------------------------------------------------
    def _graphed_device_step(self, prepared_one, host_stats):
        """Replay (after a few eager warm-up updates: capture) the whole device step as one CUDA graph.
        Inputs are copied into static buffers; per-update scalars (lr, bias correction, loss scale, dropout
        seed, NVLink epoch, batch weight) live in device memory that the graph re-reads."""
        sample_dev, stats, ignore = prepared_one
        key = tuple((tuple(t.shape), t.dtype) for t in sample_dev)
        if self._graph is not None and key != self._graph_key:
            # a batch that does not fit the captured shapes: eager (same cross-rank operations, so peers that
            # replay their graph are not disturbed)
            prepared_one = (self.task.prepare_sample(sample_dev, self.device), stats, ignore)
            return self._device_step([prepared_one], host_stats)
        if self._graph is None:
            self._graph_warmup += 1
            if self._graph_warmup <= 3 or ignore:
                return self._device_step([prepared_one], host_stats)
            # capture
            self._graph_key = key
            if isinstance(sample_dev, PackedSample):
                sample_dev = sample_dev.to(self.device)
                self._static_sample = sample_dev.clone()
            else:
                self._static_sample = [t.clone() for t in sample_dev]
            torch.cuda.synchronize()
            self._graph = torch.cuda.CUDAGraph()
            self._capturing = True
            try:
                # capture on a HIGH-priority stream: kernel nodes inherit it, so at every kernel boundary the
                # compute kernels' CTAs are placed before pending CTAs of the (default-priority) gradient
                # reduce-scatter / update kernels, which then only use what forward and backward leave free
                hp = torch.cuda.Stream(device=self.device, priority=-1)
                with torch.cuda.graph(self._graph, stream=hp):
                    self._static_out, _ = self._device_step([(self._static_sample, stats, False)], host_stats,
                                                            static_weight=True)
            finally:
                self._capturing = False
            # the captured update was only recorded, not executed: fall through to a replay
        if isinstance(self._static_sample, PackedSample) and self._static_sample.same_layout(sample_dev):
            # ONE copy (host-pinned -> static buffer directly when the batch was left on the host, else D2D)
            self._static_sample.flat.copy_(sample_dev.flat, non_blocking=True)
        else:
            for dst, src in zip(self._static_sample, sample_dev):
                dst.copy_(src, non_blocking=True)
        self.optimizer.write_host_scalars()
        self._stats_host[0], self._stats_host[1], self._stats_host[4] = host_stats[0], host_stats[1], host_stats[2]
        self._graph.replay()
        # host-side Python state that _device_step normally leaves behind
        self.optimizer._skip = self._static_out[7] != 0
        return self._static_out, False

    def valid_step(self, sample):
        """Forward only (the reference never validates: train.py:93; available via --enable-validation)."""
        prepared = self._prepare_sample(sample)
        if prepared is None:
            return None
        sample_dev, _ = prepared
        self.dp.flush()
        loss, sample_size, logging_output = self.task.valid_step(sample_dev, self._model)
        return float(loss), sample_size, logging_output

    def zero_grad(self):
        self.optimizer.zero_grad()

    def lr_step(self, epoch, val_loss=None):
        self.lr_scheduler.step(epoch, val_loss)
        return self.lr_step_update()

    def lr_step_update(self):
        return self.lr_scheduler.step_update(self.get_num_updates())

    def get_lr(self):
        return self.optimizer.get_lr()

    def get_model(self):
        return self._model

    def get_meter(self, name):
        if name not in self.meters:
            return None
        return self.meters[name]

    def get_num_updates(self):
        return self._num_updates

    def set_num_updates(self, num_updates):
        self._num_updates = num_updates
        self.lr_step_update()

    def _prepare_sample(self, sample, allow_host: bool = False):
        """Returns (device sample, host stats) or None for an empty (padding) batch. With ``allow_host`` a pinned
        packed batch that matches the captured graph's inputs is left on the host: the replay path copies it
        straight into the graph's static buffers."""
        if sample is None or len(sample) == 0:
            return None
        stats = self.task.sample_stats(sample)
        if (allow_host and self._graph is not None and isinstance(sample, PackedSample) and sample.flat.device.type == 'cpu'
                and sample.flat.is_pinned() and self._static_sample is not None
                and isinstance(self._static_sample, PackedSample) and self._static_sample.same_layout(sample)):
            return sample, stats        # goes host -> static graph inputs in one copy (_graphed_device_step)
        return self.task.prepare_sample(sample, self.device), stats

    def _set_seed(self):
        """seed + num_updates keeps results reproducible across resume (controller.py:426-432);
        the rank offset gives every replica its own dropout stream."""
        seed = self.args.seed + self.get_num_updates()
        rank = self.args.distributed_rank if self._sync_stats() else 0
        torch.manual_seed(seed + 1000003 * rank)
        if self.cuda:
            torch.cuda.manual_seed(seed + 1000003 * rank)
        from hetseq_b200.ops import rng
        rng.set_step_seed(seed, rank)

    def _sync_stats(self):
        return self.args.distributed_world_size > 1 and torch.distributed.is_initialized()

    def _mark(self):
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        return ev

    def phase_report(self, skip=3):
        """Mean milliseconds per phase over the recorded updates (first ``skip`` dropped)."""
        if not self._phase_events:
            return None
        torch.cuda.synchronize()
        names = ['fwd+bwd', 'reduce(exposed)', 'stats+norm', 'update+gather']
        rows = [[m[k].elapsed_time(m[k + 1]) for k in range(len(m) - 1)] for m in self._phase_events[skip:] if len(m) == 5]
        if not rows:
            return None
        return {n: sum(r[k] for r in rows) / len(rows) for k, n in enumerate(names)}

    def close(self):
        if self._phase_timing:
            rep = self.phase_report()
            if rep is not None:
                print('| phase timing (ms, rank {}): {}'.format(self.args.distributed_rank, ', '.join(
                    '{}={:.3f}'.format(k, v) for k, v in rep.items())), force=True, flush=True) \
                    if torch.distributed.is_initialized() else print('| phase timing (ms): {}'.format(rep), flush=True)
        if self._dp is not None:
            self._dp.close()
