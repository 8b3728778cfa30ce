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

import collections
import logging
import os
import re
import shutil
import traceback
from collections import OrderedDict

import torch

from hetseq_b200.options import literal


def save_checkpoint(args, controller, epoch_itr, val_loss):
    from hetseq_b200.parallel import distributed_utils
    from hetseq_b200.utils import meters

    prev_best = getattr(save_checkpoint, 'best', val_loss)
    if val_loss is not None:
        best_function = max if args.maximize_best_checkpoint_metric else min
        save_checkpoint.best = best_function(val_loss, prev_best)

    if args.no_save:
        return
    master = distributed_utils.is_master(args)

    def is_better(a, b):
        return a >= b if args.maximize_best_checkpoint_metric else a <= b

    write_timer = meters.StopwatchMeter()
    write_timer.start()

    epoch = epoch_itr.epoch
    end_of_epoch = epoch_itr.end_of_epoch()
    updates = controller.get_num_updates()

    checkpoint_conds = collections.OrderedDict()
    checkpoint_conds['checkpoint{}.pt'.format(epoch)] = (
        end_of_epoch and not args.no_epoch_checkpoints and epoch % args.save_interval == 0)
    checkpoint_conds['checkpoint_{}_{}.pt'.format(epoch, updates)] = (
        not end_of_epoch and args.save_interval_updates > 0 and updates % args.save_interval_updates == 0)
    checkpoint_conds['checkpoint_best.pt'] = (
        val_loss is not None and
        (not hasattr(save_checkpoint, 'best') or is_better(val_loss, save_checkpoint.best)))
    checkpoint_conds['checkpoint_last.pt'] = not args.no_last_checkpoints

    extra_state = {'train_iterator': epoch_itr.state_dict(), 'val_loss': val_loss}
    if hasattr(save_checkpoint, 'best'):
        extra_state.update({'best': save_checkpoint.best})

    checkpoints = [os.path.join(args.save_dir, fn) for fn, cond in checkpoint_conds.items() if cond]
    if len(checkpoints) > 0:
        # collective when the optimizer is sharded: every rank calls, rank 0 writes
        controller.save_checkpoint(checkpoints[0], extra_state)
        if master:
            for cp in checkpoints[1:]:
                shutil.copyfile(checkpoints[0], cp)
            write_timer.stop()
            print('| saved checkpoint {} (epoch {} @ {} updates) (writing took {} seconds)'.format(
                checkpoints[0], epoch, updates, write_timer.sum))
    if not master:
        return

    if not end_of_epoch and args.keep_interval_updates > 0:
        for old_chk in checkpoint_paths(args.save_dir, pattern=r'checkpoint_\d+_(\d+)\.pt')[args.keep_interval_updates:]:
            if os.path.lexists(old_chk):
                os.remove(old_chk)
    if args.keep_last_epochs > 0:
        for old_chk in checkpoint_paths(args.save_dir, pattern=r'checkpoint(\d+)\.pt')[args.keep_last_epochs:]:
            if os.path.lexists(old_chk):
                os.remove(old_chk)


def load_checkpoint(args, controller):
    """Load a checkpoint (if there is one) and build / restore the training iterator."""
    if args.distributed_rank == 0:
        os.makedirs(args.save_dir, exist_ok=True)

    if args.restore_file in ('checkpoint_last.pt', 'checkpoint_best.pt'):
        checkpoint_path = os.path.join(args.save_dir, args.restore_file)
    else:
        checkpoint_path = args.restore_file

    overrides = args.optimizer_overrides
    if isinstance(overrides, str):
        overrides = literal(overrides)
    extra_state = controller.load_checkpoint(
        checkpoint_path, args.reset_optimizer, args.reset_lr_scheduler, overrides,
        reset_meters=args.reset_meters)

    if extra_state is not None and 'best' in extra_state and not args.reset_optimizer and not args.reset_meters:
        save_checkpoint.best = extra_state['best']

    if extra_state is not None and 'train_iterator' in extra_state and not args.reset_dataloader:
        itr_state = extra_state['train_iterator']
        epoch_itr = controller.get_train_iterator(epoch=itr_state['epoch'], load_dataset=True)
        epoch_itr.load_state_dict(itr_state)
    else:
        epoch_itr = controller.get_train_iterator(epoch=0, load_dataset=True)

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


def checkpoint_paths(path, pattern=r'checkpoint(\d+)\.pt'):
    """All checkpoints in ``path`` matching ``pattern``, sorted by the first group, descending."""
    pt_regexp = re.compile(pattern)
    entries = []
    for i, f in enumerate(os.listdir(path)):
        m = pt_regexp.fullmatch(f)
        if m is not None:
            idx = int(m.group(1)) if len(m.groups()) > 0 else i
            entries.append((idx, m.group(0)))
    return [os.path.join(path, x[1]) for x in sorted(entries, reverse=True)]


def torch_persistent_save(obj, filename):
    """torch.save with 3 attempts (checkpoint_utils.py:161-167), atomically via rename."""
    tmp = '{}.tmp.{}'.format(filename, os.getpid())
    for i in range(3):
        try:
            torch.save(obj, tmp)
            os.replace(tmp, filename)
            return True
        except Exception:
            if i == 2:
                logging.error(traceback.format_exc())
                try:
                    if os.path.exists(tmp):
                        os.remove(tmp)
                except OSError:
                    pass
    return False


def convert_state_dict_type(state_dict, ttype=torch.FloatTensor):
    """Recursively cast every tensor to a CPU float tensor (checkpoint_utils.py:170-181)."""
    if isinstance(state_dict, dict):
        cpu_dict = OrderedDict()
        for k, v in state_dict.items():
            cpu_dict[k] = convert_state_dict_type(v, ttype)
        return cpu_dict
    if isinstance(state_dict, list):
        return [convert_state_dict_type(v, ttype) for v in state_dict]
    if torch.is_tensor(state_dict):
        return state_dict.detach().to('cpu').type(ttype)
    return state_dict


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
    if not os.path.exists(save_dir):
        os.makedirs(save_dir, exist_ok=True)
    temp_file_path = os.path.join(save_dir, 'dummy')
    try:
        with open(temp_file_path, 'w'):
            pass
    except OSError as e:
        print('| Unable to access checkpoint save directory: {}'.format(save_dir))
        raise e
    else:
        os.remove(temp_file_path)
