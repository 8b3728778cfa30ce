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
import math
import os
import random
import sys

import numpy as np
import torch

from hetseq_b200 import options, registry
from hetseq_b200.controller import Controller
from hetseq_b200.data import iterators
from hetseq_b200.parallel import distributed_utils
from hetseq_b200.utils import checkpoint_utils, progress_bar, utils
from hetseq_b200.utils.meters import AverageMeter, StopwatchMeter


def main(args, init_distributed=False):
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
    for valid_sub_split in args.valid_subset.split(','):
        try:
            task.load_dataset(valid_sub_split, combine=False, epoch=0)
        except (AssertionError, FileNotFoundError) as e:
            if getattr(args, 'enable_validation', False):
                raise
            print('| validation split {} not loaded ({})'.format(valid_sub_split, e))

    model = task.build_model(args)
    print('| num. model params: {} (num. trained: {})'.format(
        sum(p.numel() for p in model.parameters()),
        sum(p.numel() for p in model.parameters() if p.requires_grad)))

    controller = Controller(args, task, model)
    print('| training on {} GPUs'.format(args.distributed_world_size))
    print('| max tokens per GPU = {} and max sentences per GPU = {}'.format(args.max_tokens, args.max_sentences))

    extra_state, epoch_itr = checkpoint_utils.load_checkpoint(args, controller)

    max_epoch = args.max_epoch or math.inf
    max_update = args.max_update or math.inf
    lr = controller.get_lr()
    train_meter = StopwatchMeter()
    train_meter.start()

    while (lr > args.min_lr
           and (epoch_itr.epoch < max_epoch
                or (epoch_itr.epoch == max_epoch and epoch_itr._next_epoch_itr is not None))
           and controller.get_num_updates() < max_update):
        train(args, controller, task, epoch_itr)

        valid_losses = [None]
        if getattr(args, 'enable_validation', False) and not args.disable_validation \
                and epoch_itr.epoch % args.validate_interval == 0:
            valid_losses = validate(args, controller, task, epoch_itr, args.valid_subset.split(','))

        lr = controller.lr_step(epoch_itr.epoch, valid_losses[0])

        if epoch_itr.epoch % args.save_interval == 0:
            checkpoint_utils.save_checkpoint(args, controller, epoch_itr, valid_losses[0])

        reload_dataset = ':' in (getattr(args, 'data', '') or '')
        epoch_itr = controller.get_train_iterator(epoch_itr.epoch, load_dataset=reload_dataset)

    train_meter.stop()
    print('| done training in {:.1f} seconds'.format(train_meter.sum))
    controller.close()
    return controller


def train(args, controller, task, epoch_itr):
    """Train the model for one epoch (train.py:110-161)."""
    update_freq = args.update_freq[epoch_itr.epoch - 1] \
        if epoch_itr.epoch <= len(args.update_freq) else args.update_freq[-1]

    itr = epoch_itr.next_epoch_itr(fix_batches_to_gpus=args.fix_batches_to_gpus,
                                   shuffle=(epoch_itr.epoch >= args.curriculum))
    itr = iterators.GroupedIterator(itr, update_freq)
    if controller.cuda or os.environ.get('HETSEQ_PREFETCH', '') == '1':
        # batch preparation (MLM row index, int32 narrowing, pinning) runs one update ahead in a background
        # thread; the training thread is left with the asynchronous host->device copies
        itr = iterators.PrefetchIterator(itr, depth=2, transform=lambda chunk: [task.host_prepare(s) for s in chunk])
        itr.attach_to(epoch_itr)
    else:
        epoch_itr.position_hook = None
    progress = progress_bar.build_progress_bar(args, itr, epoch_itr.epoch, no_progress_bar='simple')

    extra_meters = collections.defaultdict(lambda: AverageMeter())
    max_update = args.max_update or math.inf

    for i, samples in enumerate(progress, start=epoch_itr.iterations_in_epoch):
        log_output = controller.train_step(samples)
        if log_output is None:
            continue

        stats = get_training_stats(controller)
        for k, v in log_output.items():
            if k in ['loss', 'nll_loss', 'ntokens', 'nsentences', 'sample_size']:
                continue  # these are already logged above
            if 'loss' in k or k == 'accuracy':
                extra_meters[k].update(v, log_output['sample_size'])
            else:
                extra_meters[k].update(v)
            stats[k] = extra_meters[k].avg
        progress.log(stats, tag='train', step=stats['num_updates'])

        # ignore the first mini-batch in words-per-second and updates-per-second calculation
        if i == 0:
            controller.get_meter('wps').reset()
            controller.get_meter('ups').reset()

        num_updates = controller.get_num_updates()
        if (not args.no_save and args.save_interval_updates > 0
                and num_updates % args.save_interval_updates == 0 and num_updates > 0
                and not epoch_itr.end_of_epoch()):
            checkpoint_utils.save_checkpoint(args, controller, epoch_itr, None)
        if num_updates >= max_update:
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


def get_training_stats(controller):
    stats = collections.OrderedDict()
    stats['loss'] = controller.get_meter('train_loss')
    if controller.get_meter('train_nll_loss').count > 0:
        nll_loss = controller.get_meter('train_nll_loss')
        stats['nll_loss'] = nll_loss
    else:
        nll_loss = controller.get_meter('train_loss')
    stats['ppl'] = utils.get_perplexity(nll_loss.avg)
    stats['wps'] = controller.get_meter('wps')
    stats['ups'] = controller.get_meter('ups')
    stats['wpb'] = controller.get_meter('wpb')
    stats['bsz'] = controller.get_meter('bsz')
    stats['num_updates'] = controller.get_num_updates()
    stats['lr'] = controller.get_lr()
    stats['gnorm'] = controller.get_meter('gnorm')
    stats['clip'] = controller.get_meter('clip')
    stats['oom'] = controller.get_meter('oom')
    if controller.get_meter('loss_scale') is not None:
        stats['loss_scale'] = controller.get_meter('loss_scale')
    stats['wall'] = round(controller.get_meter('wall').elapsed_time)
    stats['train_wall'] = controller.get_meter('train_wall')
    return stats


def distributed_main(i, args, start_rank=0):
    args.device_id = i
    if args.distributed_rank is None:  # torch.multiprocessing.spawn
        args.distributed_rank = start_rank + i
    main(args, init_distributed=True)


def cli_main(argv=None):
    args = options.parse_cli(argv)
    distributed_utils.infer_init_method(args)
    n_dev = torch.cuda.device_count()

    if args.distributed_init_method is not None:
        if not args.cpu:
            assert args.distributed_gpus <= max(n_dev, 1) or args.distributed_no_spawn, \
                '--distributed-gpus {} exceeds the {} visible GPUs'.format(args.distributed_gpus, n_dev)
        if args.distributed_gpus > 1 and not args.distributed_no_spawn:
            start_rank = args.distributed_rank
            args.distributed_rank = None  # assigned per spawned worker
            torch.multiprocessing.spawn(fn=distributed_main, args=(args, start_rank), nprocs=args.distributed_gpus)
        else:
            distributed_main(args.device_id, args)
    elif args.distributed_world_size > 1:
        if not args.cpu:
            assert args.distributed_world_size <= n_dev, \
                '--distributed-world-size {} exceeds the {} visible GPUs'.format(args.distributed_world_size, n_dev)
        port = random.randint(10000, 20000)
        args.distributed_init_method = 'tcp://127.0.0.1:{port}'.format(port=port)
        args.distributed_rank = None  # set based on device id
        torch.multiprocessing.spawn(fn=distributed_main, args=(args,), nprocs=args.distributed_world_size)
    else:
        main(args)


if __name__ == '__main__':
    cli_main()
