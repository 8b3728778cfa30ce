"""Checkpoint save / load in the reference's fairseq-style file layout (checkpoint_utils.py).

One ``torch.save``d dict per file: ``args``, ``model``, ``optimizer_history``, ``extra_state``,
``last_optimizer_state``; names ``checkpoint{epoch}.pt``,
``checkpoint_{epoch}_{updates}.pt``, ``checkpoint_best.pt``, ``checkpoint_last.pt``; retention via
``--keep-interval-updates`` / ``--keep-last-epochs``.

Fixes relative to the reference:
* ``extra_state`` is actually written (the reference stores ``{}``, checkpoint_utils.py:204, so its
  own files cannot be resumed: controller.py:183 raises KeyError) -- resume works here;
* files are loaded with ``weights_only=False`` explicitly (they pickle an ``argparse.Namespace``
  and meter objects, which torch >= 2.6 rejects by default);
* ``--optimizer-overrides`` is parsed with ``ast.literal_eval`` instead of ``eval``;
* writes go to a temporary file first and are renamed, so a crash cannot leave a torn
  ``checkpoint_last.pt``;
* with a sharded optimizer every rank takes part in gathering the state, only rank 0 writes.
"""
from __future__ import annotations

import logging
import os
import re
import shutil
import traceback
from collections import OrderedDict

import torch

from hetseq_b200.options import literal


EPOCH_PATTERN = r'checkpoint(\d+)\.pt'
UPDATES_PATTERN = r'checkpoint_\d+_(\d+)\.pt'


def planned_files(args, epoch: int, updates: int, end_of_epoch: bool, val_loss, best) -> list:
    """File names one call to ``save_checkpoint`` writes, first = the one actually serialised (the others are
    copies of it): end-of-epoch file every ``--save-interval`` epochs, mid-epoch file every
    ``--save-interval-updates`` updates, ``checkpoint_best.pt`` when the validation loss ties or improves on the
    best so far (direction per ``--maximize-best-checkpoint-metric``), ``checkpoint_last.pt`` always -- each unless
    its ``--no-*`` flag says otherwise."""
    names = []
    if end_of_epoch:
        if not args.no_epoch_checkpoints and epoch % args.save_interval == 0:
            names.append('checkpoint{}.pt'.format(epoch))
    elif args.save_interval_updates > 0 and updates % args.save_interval_updates == 0:
        names.append('checkpoint_{}_{}.pt'.format(epoch, updates))
    if val_loss is not None:
        improved = best is None or (val_loss >= best if args.maximize_best_checkpoint_metric else val_loss <= best)
        if improved:
            names.append('checkpoint_best.pt')
    if not args.no_last_checkpoints:
        names.append('checkpoint_last.pt')
    return names


def prune(save_dir: str, pattern: str, keep: int) -> None:
    """Delete all but the ``keep`` newest checkpoints matching ``pattern`` (no-op for keep <= 0)."""
    if keep > 0:
        for stale in checkpoint_paths(save_dir, pattern=pattern)[keep:]:
            if os.path.lexists(stale):
                os.remove(stale)


def save_checkpoint(args, controller, epoch_itr, val_loss):
    """Write whatever is due now (see ``planned_files``) and apply the retention flags. The best validation
    loss lives on as the function attribute ``save_checkpoint.best`` (restored by ``load_checkpoint``)."""
    from hetseq_b200.parallel import distributed_utils
    from hetseq_b200.utils import meters

    had_best = hasattr(save_checkpoint, 'best')
    best_before = save_checkpoint.best if had_best else None
    if val_loss is not None:
        pick = max if args.maximize_best_checkpoint_metric else min
        save_checkpoint.best = pick(val_loss, best_before if had_best else val_loss)
    if args.no_save:
        return

    timer = meters.StopwatchMeter()
    timer.start()
    epoch, updates, at_end = epoch_itr.epoch, controller.get_num_updates(), epoch_itr.end_of_epoch()
    # "best" is judged against the value INCLUDING this validation loss, i.e. ties with the running best count
    names = planned_files(args, epoch, updates, at_end, val_loss, getattr(save_checkpoint, 'best', None))
    extra_state = {'train_iterator': epoch_itr.state_dict(), 'val_loss': val_loss}
    if hasattr(save_checkpoint, 'best'):
        extra_state['best'] = save_checkpoint.best

    master = distributed_utils.is_master(args)
    if names:
        first = os.path.join(args.save_dir, names[0])
        controller.save_checkpoint(first, extra_state)      # collective with a sharded optimizer; rank 0 writes
        if master:
            for other in names[1:]:
                shutil.copyfile(first, os.path.join(args.save_dir, other))
            timer.stop()
            print('| saved checkpoint {} (epoch {} @ {} updates) (writing took {} seconds)'.format(
                first, epoch, updates, timer.sum))
    if master:
        if not at_end:
            prune(args.save_dir, UPDATES_PATTERN, args.keep_interval_updates)
        prune(args.save_dir, EPOCH_PATTERN, args.keep_last_epochs)
    if names and distributed_utils.get_world_size() > 1:
        # Only the master writes and copies (possibly for many seconds, to slow storage). Hold the other ranks
        # here instead of letting them queue kernels that spin on the master's flags: a host-side barrier is
        # bounded by the process-group timeout (--distributed-timeout), like every wait of the reference.
        distributed_utils.barrier()


def load_checkpoint(args, controller):
    """Restore from ``--restore-file`` when it exists and return ``(extra_state, epoch iterator)``; the iterator is
    positioned where the checkpoint left off unless ``--reset-dataloader``."""
    if args.distributed_rank == 0:
        os.makedirs(args.save_dir, exist_ok=True)
    in_save_dir = args.restore_file in ('checkpoint_last.pt', 'checkpoint_best.pt')
    path = os.path.join(args.save_dir, args.restore_file) if in_save_dir else args.restore_file
    overrides = literal(args.optimizer_overrides) if isinstance(args.optimizer_overrides, str) \
        else args.optimizer_overrides
    extra_state = controller.load_checkpoint(path, args.reset_optimizer, args.reset_lr_scheduler, overrides,
                                             reset_meters=args.reset_meters)
    extra = extra_state or {}
    if 'best' in extra and not (args.reset_optimizer or args.reset_meters):
        save_checkpoint.best = extra['best']
    position = None if args.reset_dataloader else extra.get('train_iterator')
    epoch_itr = controller.get_train_iterator(epoch=position['epoch'] if position else 0, load_dataset=True)
    if position:
        epoch_itr.load_state_dict(position)
    controller.lr_step(epoch_itr.epoch)
    return extra_state, epoch_itr


def load_checkpoint_to_cpu(path, arg_overrides=None):
    """torch.load onto the CPU; ``arg_overrides`` patches the stored args namespace."""
    state = torch.load(path, map_location='cpu', weights_only=False)
    args = state.get('args')
    if arg_overrides is not None and args is not None:
        for arg_name, arg_val in arg_overrides.items():
            setattr(args, arg_name, arg_val)
    state.setdefault('extra_state', {})
    state.setdefault('optimizer_history', [])
    return state


def checkpoint_paths(path, pattern=EPOCH_PATTERN):
    """Checkpoints in ``path`` whose name matches ``pattern``, newest first (ordered by the pattern's first group)."""
    rx = re.compile(pattern)
    found = []
    for position, name in enumerate(os.listdir(path)):
        m = rx.fullmatch(name)
        if m:
            found.append((int(m.group(1)) if m.groups() else position, name))
    found.sort(reverse=True)
    return [os.path.join(path, name) for _, name in found]


def torch_persistent_save(obj, filename, attempts: int = 3):
    """``torch.save`` with retries like the reference (checkpoint_utils.py:161-167), but through a temporary
    file that is renamed into place: a crash never leaves a torn ``checkpoint_last.pt``."""
    tmp = '{}.tmp.{}'.format(filename, os.getpid())
    for attempt in range(1, attempts + 1):
        try:
            torch.save(obj, tmp)
            os.replace(tmp, filename)
            return True
        except Exception:
            if attempt == attempts:
                logging.error(traceback.format_exc())
    if os.path.exists(tmp):
        try:
            os.remove(tmp)
        except OSError:
            pass
    return False


def convert_state_dict_type(obj, ttype=torch.FloatTensor):
    """Copy of a nested state (dicts / lists) with every tensor on the CPU as ``ttype`` (fp32 by default, which is
    what the reference stores in ``last_optimizer_state``)."""
    if torch.is_tensor(obj):
        return obj.detach().to('cpu').type(ttype)
    if isinstance(obj, dict):
        return OrderedDict((k, convert_state_dict_type(v, ttype)) for k, v in obj.items())
    if isinstance(obj, list):
        return [convert_state_dict_type(v, ttype) for v in obj]
    return obj


def build_state(args, model_state_dict, optimizer_name, optimizer_state, lr_scheduler_state, num_updates,
                optim_history=None, extra_state=None):
    """Assemble the checkpoint dict (schema of checkpoint_utils.py:193-208, extra_state populated)."""
    state_dict = {
        'args': args,
        'model': model_state_dict if model_state_dict else {},
        'optimizer_history': (optim_history or []) + [{
            'optimizer_name': optimizer_name,
            'lr_scheduler_state': lr_scheduler_state,
            'num_updates': num_updates,
        }],
        'extra_state': extra_state if extra_state is not None else {},
    }
    if optimizer_state is not None and not getattr(args, 'no_save_optimizer_state', False):
        state_dict['last_optimizer_state'] = convert_state_dict_type(optimizer_state)
    return state_dict


def save_state(filename, args, model_state_dict, criterion, optimizer, lr_scheduler, num_updates,
               optim_history=None, extra_state=None):
    """Reference-compatible entry point (checkpoint_utils.py:184-208): builds and writes the file."""
    opt_state = None if getattr(args, 'no_save_optimizer_state', False) else optimizer.state_dict()
    state = build_state(args, model_state_dict, optimizer.__class__.__name__, opt_state,
                        lr_scheduler.state_dict(), num_updates, optim_history, extra_state)
    torch_persistent_save(state, filename)


def verify_checkpoint_directory(save_dir: str) -> None:
    """Fail early (before any training) when ``save_dir`` cannot be created or written."""
    os.makedirs(save_dir, exist_ok=True)
    probe = os.path.join(save_dir, 'dummy')
    try:
        open(probe, 'w').close()
    except OSError:
        print('| Unable to access checkpoint save directory: {}'.format(save_dir))
        raise
    os.remove(probe)
