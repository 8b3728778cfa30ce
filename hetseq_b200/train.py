"""Command-line trainer: ``python -m hetseq_b200.train --task {bert,mnist} ...`` (reference: train.py).

Launch modes (train.py:213-239):
  A) ``--distributed-init-method tcp://host:port | file:///path`` given: this node runs
     ``--distributed-gpus`` workers with ranks ``--distributed-rank .. +gpus-1`` (spawned unless
     ``--distributed-no-spawn``); start the same command on every node by hand -- that is HetSeq's
     "heterogeneous infrastructure" story (README.md:50-56);
  B) no init method, ``--distributed-world-size`` > 1: single node, ``tcp://127.0.0.1:<random port>``;
  C) world size 1: plain single process, no process group.
Additionally, when started under ``torchrun`` the rendezvous is taken from the environment.

Epoch loop (train.py:74-107): until lr <= --min-lr, --max-epoch or --max-update is reached: train
one epoch, step the LR schedule, save a checkpoint every --save-interval epochs. On top of the
reference: mid-epoch checkpoints (--save-interval-updates; the reference parses the flag but never
saves mid-epoch) and optional validation (--enable-validation; the reference never validates).
"""
from __future__ import annotations

import collections
import dataclasses
import math
import os
import random
from typing import Callable, List, Optional, Tuple

import numpy as np
import torch

from hetseq_b200 import options, registry
from hetseq_b200.controller import Controller
from hetseq_b200.data import iterators
from hetseq_b200.parallel import distributed_utils
from hetseq_b200.utils import checkpoint_utils, progress_bar, utils
from hetseq_b200.utils.meters import AverageMeter, StopwatchMeter


def _setup(args, init_distributed: bool):
    """Process-level set-up in the order the reference uses (train.py:22-72): device, seeds, process group,
    save-dir probe, task, validation splits, model, controller, checkpoint."""
    assert args.max_tokens is not None or args.max_sentences is not None, \
        'Must specify batch size either with --max-tokens or --max-sentences'
    registry.ensure_builtin_components_loaded()
    if torch.cuda.is_available() and not args.cpu:
        torch.cuda.set_device(args.device_id)
    np.random.seed(args.seed)
    torch.manual_seed(args.seed)
    if init_distributed:
        args.distributed_rank = distributed_utils.distributed_init(args)
    if distributed_utils.is_master(args):
        checkpoint_utils.verify_checkpoint_directory(args.save_dir)
    print(args, flush=True)

    task = registry.lookup('task', args.task).setup_task(args)
    # validation splits are loaded up front like the reference (train.py:50-51), but a missing split
    # is only fatal when validation is actually enabled
    for split in args.valid_subset.split(','):
        try:
            task.load_dataset(split, combine=False, epoch=0)
        except (AssertionError, FileNotFoundError) as e:
            if getattr(args, 'enable_validation', False):
                raise
            print('| validation split {} not loaded ({})'.format(split, e))

    model = task.build_model(args)
    n_all = sum(p.numel() for p in model.parameters())
    n_trained = sum(p.numel() for p in model.parameters() if p.requires_grad)
    print('| num. model params: {} (num. trained: {})'.format(n_all, n_trained))
    controller = Controller(args, task, model)
    print('| training on {} GPUs'.format(args.distributed_world_size))
    print('| max tokens per GPU = {} and max sentences per GPU = {}'.format(args.max_tokens, args.max_sentences))
    _, epoch_itr = checkpoint_utils.load_checkpoint(args, controller)
    return task, controller, epoch_itr


class _Budget(object):
    """When training stops (train.py:74-88): the learning rate fell to --min-lr, --max-epoch epochs are done (a
    resumed, half-finished last epoch still counts as pending), or --max-update updates were made; 0 = no limit."""

    def __init__(self, args):
        self.min_lr = args.min_lr
        self.max_epoch = args.max_epoch or math.inf
        self.max_update = args.max_update or math.inf

    def updates_left(self, controller) -> bool:
        return controller.get_num_updates() < self.max_update

    def epochs_left(self, epoch_itr) -> bool:
        resumed_mid_epoch = epoch_itr.epoch == self.max_epoch and epoch_itr._next_epoch_itr is not None
        return epoch_itr.epoch < self.max_epoch or resumed_mid_epoch

    def keep_going(self, controller, epoch_itr, lr) -> bool:
        return lr > self.min_lr and self.epochs_left(epoch_itr) and self.updates_left(controller)


def main(args, init_distributed=False):
    task, controller, epoch_itr = _setup(args, init_distributed)
    budget = _Budget(args)
    validating = getattr(args, 'enable_validation', False) and not args.disable_validation
    wall = StopwatchMeter()
    wall.start()
    lr = controller.get_lr()
    while budget.keep_going(controller, epoch_itr, lr):
        train(args, controller, task, epoch_itr)
        epoch = epoch_itr.epoch
        val_loss = None
        if validating and epoch % args.validate_interval == 0:
            val_loss = validate(args, controller, task, epoch_itr, args.valid_subset.split(','))[0]
        lr = controller.lr_step(epoch, val_loss)
        if epoch % args.save_interval == 0:
            checkpoint_utils.save_checkpoint(args, controller, epoch_itr, val_loss)
        # sharded corpora ("dir1:dir2:...") are re-listed between epochs, like train.py:102-104
        epoch_itr = controller.get_train_iterator(epoch, load_dataset=':' in (getattr(args, 'data', '') or ''))
    wall.stop()
    print('| done training in {:.1f} seconds'.format(wall.sum))
    controller.close()
    return controller


# -- one epoch ---------------------------------------------------------------------------------------------
def _per_epoch(values: List, epoch: int):
    """--update-freq / --lr style lists: entry ``epoch`` (1-based), the last entry for later epochs."""
    return values[min(epoch, len(values)) - 1]


def _update_stream(args, controller, task, epoch_itr):
    """Iterator over the updates of this epoch: shuffled, sharded micro-batches in chunks of update-freq; on a
    GPU the host side of batch preparation (MLM row index, int32 narrowing, pinning) runs one update ahead in a
    background thread, leaving the training thread the asynchronous host->device copies."""
    itr = epoch_itr.next_epoch_itr(fix_batches_to_gpus=args.fix_batches_to_gpus,
                                   shuffle=epoch_itr.epoch >= args.curriculum)
    itr = iterators.GroupedIterator(itr, _per_epoch(args.update_freq, epoch_itr.epoch))
    if controller.cuda or os.environ.get('HETSEQ_PREFETCH', '') == '1':
        itr = iterators.PrefetchIterator(itr, depth=2, transform=lambda chunk: [task.host_prepare(s) for s in chunk])
        itr.attach_to(epoch_itr)
    else:
        epoch_itr.position_hook = None
    return itr


_CORE_KEYS = frozenset(('loss', 'nll_loss', 'ntokens', 'nsentences', 'sample_size'))


def _fold_extra(extra_meters, log_output, stats):
    """Task-specific entries of a logging output (anything beyond the core keys) get running averages and are
    appended to the line; '*loss*' and 'accuracy' entries are weighted by the sample size."""
    for key, value in log_output.items():
        if key in _CORE_KEYS:
            continue
        weighted = 'loss' in key or key == 'accuracy'
        extra_meters[key].update(value, log_output['sample_size']) if weighted else extra_meters[key].update(value)
        stats[key] = extra_meters[key].avg


def _mid_epoch_save_due(args, controller, epoch_itr) -> bool:
    n = controller.get_num_updates()
    return (not args.no_save and args.save_interval_updates > 0 and n > 0
            and n % args.save_interval_updates == 0 and not epoch_itr.end_of_epoch())


def train(args, controller, task, epoch_itr):
    """One epoch of updates (train.py:110-161) + mid-epoch checkpoints (--save-interval-updates)."""
    budget = _Budget(args)
    progress = progress_bar.build_progress_bar(args, _update_stream(args, controller, task, epoch_itr),
                                               epoch_itr.epoch, no_progress_bar='simple')
    extra_meters = collections.defaultdict(AverageMeter)
    first = True
    for samples in progress:
        log_output = controller.train_step(samples)
        if log_output is not None:
            stats = get_training_stats(controller)
            _fold_extra(extra_meters, log_output, stats)
            progress.log(stats, tag='train', step=stats['num_updates'])
            if first and epoch_itr.iterations_in_epoch <= 1:
                # the first update of an epoch carries start-up cost: keep it out of wps / ups
                controller.get_meter('wps').reset()
                controller.get_meter('ups').reset()
            first = False
            if _mid_epoch_save_due(args, controller, epoch_itr):
                checkpoint_utils.save_checkpoint(args, controller, epoch_itr, None)
            if not budget.updates_left(controller):
                break


def validate(args, controller, task, epoch_itr, subsets):
    """Average loss per validation subset (base 2, like the training log)."""
    valid_losses = []
    for subset in subsets:
        if not task.has_dataset(subset):
            valid_losses.append(None)
            continue
        itr = task.get_batch_iterator(
            dataset=task.dataset(subset), max_tokens=args.max_tokens_valid,
            max_sentences=args.max_sentences_valid, required_batch_size_multiple=args.required_batch_size_multiple,
            seed=args.seed, num_shards=args.distributed_world_size, shard_id=args.distributed_rank,
            num_workers=args.num_workers).next_epoch_itr(shuffle=False)
        total, count = 0.0, 0
        for sample in itr:
            out = controller.valid_step(sample)
            if out is None:
                continue
            loss, sample_size, log = out
            n = log.get('nsentences', 1)
            total += loss * n
            count += n
        if args.distributed_world_size > 1 and torch.distributed.is_initialized():
            t = torch.tensor([total, count], dtype=torch.float64, device=distributed_utils.comm_device())
            torch.distributed.all_reduce(t)
            total, count = t.tolist()
        loss = total / max(count, 1) / math.log(2)
        controller.get_meter('valid_loss').update(loss, count)
        print('| epoch {:03d} | valid on \'{}\' subset | loss {:.3f} | ppl {:.2f} | num_updates {}'.format(
            epoch_itr.epoch, subset, loss, utils.get_perplexity(loss), controller.get_num_updates()))
        valid_losses.append(loss)
    return valid_losses


# -- the log line ------------------------------------------------------------------------------------------
def _ppl(controller):
    nll = controller.get_meter('train_nll_loss')
    src = nll if nll.count > 0 else controller.get_meter('train_loss')
    return utils.get_perplexity(src.avg)


#: the training log line, in print order: (key, how to get it, only-if). Keys and order are what log scrapers of
#: the reference see (train.py:164-186); 'oom' and 'loss_scale' are real meters here (the reference prints
#: oom=None and never creates loss_scale).
STAT_SCHEMA: Tuple[Tuple[str, Callable, Optional[Callable]], ...] = (
    ('loss', lambda c: c.get_meter('train_loss'), None),
    ('nll_loss', lambda c: c.get_meter('train_nll_loss'), lambda c: c.get_meter('train_nll_loss').count > 0),
    ('ppl', _ppl, None),
    ('wps', lambda c: c.get_meter('wps'), None),
    ('ups', lambda c: c.get_meter('ups'), None),
    ('wpb', lambda c: c.get_meter('wpb'), None),
    ('bsz', lambda c: c.get_meter('bsz'), None),
    ('num_updates', lambda c: c.get_num_updates(), None),
    ('lr', lambda c: c.get_lr(), None),
    ('gnorm', lambda c: c.get_meter('gnorm'), None),
    ('clip', lambda c: c.get_meter('clip'), None),
    ('oom', lambda c: c.get_meter('oom'), None),
    ('loss_scale', lambda c: c.get_meter('loss_scale'), lambda c: c.get_meter('loss_scale') is not None),
    ('wall', lambda c: round(c.get_meter('wall').elapsed_time), None),
    ('train_wall', lambda c: c.get_meter('train_wall'), None),
)


def get_training_stats(controller):
    return collections.OrderedDict((key, get(controller)) for key, get, cond in STAT_SCHEMA
                                   if cond is None or cond(controller))


# -- launching ---------------------------------------------------------------------------------------------
def distributed_main(i, args, start_rank=0):
    """Entry point of one worker process: local device ``i``, global rank ``start_rank + i``."""
    args.device_id = i
    if args.distributed_rank is None:  # torch.multiprocessing.spawn
        args.distributed_rank = start_rank + i
    main(args, init_distributed=True)


@dataclasses.dataclass
class LaunchPlan(object):
    """How this invocation turns into worker processes (the three modes of train.py:213-239).

    ``mode``: 'single' (no process group), 'inline' (this process is one worker of a group set up elsewhere:
    --distributed-no-spawn, one GPU per node, torchrun), 'spawn' (``nprocs`` workers forked here).
    ``first_rank`` is the global rank of local worker 0."""
    mode: str
    nprocs: int = 1
    init_method: Optional[str] = None
    first_rank: int = 0


def plan_launch(args, n_devices: int, port: Optional[int] = None) -> LaunchPlan:
    on_gpu = not args.cpu
    if args.distributed_init_method is not None:
        # multi-node (or explicit rendezvous): this node contributes --distributed-gpus workers
        want = args.distributed_gpus
        if on_gpu and not args.distributed_no_spawn and want > max(n_devices, 1):
            raise AssertionError('--distributed-gpus {} exceeds the {} visible GPUs'.format(want, n_devices))
        if want > 1 and not args.distributed_no_spawn:
            return LaunchPlan('spawn', want, args.distributed_init_method, args.distributed_rank)
        return LaunchPlan('inline', 1, args.distributed_init_method, args.distributed_rank)
    if args.distributed_world_size > 1:
        # single node, no rendezvous given: make one up on the loopback interface
        world = args.distributed_world_size
        if on_gpu and world > n_devices:
            raise AssertionError('--distributed-world-size {} exceeds the {} visible GPUs'.format(world, n_devices))
        port = port if port is not None else random.randint(10000, 20000)
        return LaunchPlan('spawn', world, 'tcp://127.0.0.1:{}'.format(port), 0)
    return LaunchPlan('single')


def cli_main(argv=None):
    args = options.parse_cli(argv)
    distributed_utils.infer_init_method(args)      # torchrun: rendezvous from the environment
    plan = plan_launch(args, torch.cuda.device_count())
    if plan.mode == 'single':
        return main(args)
    args.distributed_init_method = plan.init_method
    if plan.mode == 'inline':
        return distributed_main(args.device_id, args)
    args.distributed_rank = None                   # every spawned worker derives its own: first_rank + index
    torch.multiprocessing.spawn(fn=distributed_main, args=(args, plan.first_rank), nprocs=plan.nprocs)


if __name__ == '__main__':
    cli_main()
