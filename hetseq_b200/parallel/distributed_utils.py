"""Process-group bootstrap and small host-side collectives (reference: distributed_utils.py).

Control plane only: rendezvous (``tcp://`` / ``file://`` / ``env://``), rank bookkeeping, rank-0
printing and the pickled stats exchange. The gradient data plane lives in
``parallel/data_parallel.py`` (NCCL baseline) and ``parallel/fused_dp.py`` (in-kernel NVLink path).

Unlike the reference, nothing here assumes CUDA tensors (it allocates ``torch.cuda.ByteTensor`` /
``DoubleTensor``, distributed_utils.py:95, controller.py:294), so the same code runs under Gloo
on CPU -- which is how the multi-process tests exercise it.
"""
from __future__ import annotations

import datetime
import os
import pickle
import socket
import warnings
from typing import Any, List, Optional

import torch
import torch.distributed as dist

_builtin_print = None


def infer_init_method(args) -> None:
    """Fill ``distributed_*`` from torchrun-style environment variables when present and the user
    gave no explicit ``--distributed-init-method`` (lets ``torchrun ... -m hetseq_b200.train`` work)."""
    if args.distributed_init_method is not None:
        return
    if all(k in os.environ for k in ('MASTER_ADDR', 'MASTER_PORT', 'WORLD_SIZE', 'RANK')):
        args.distributed_init_method = 'env://'
        args.distributed_world_size = int(os.environ['WORLD_SIZE'])
        args.distributed_rank = int(os.environ['RANK'])
        args.device_id = int(os.environ.get('LOCAL_RANK', args.device_id))
        args.distributed_no_spawn = True


def comm_device(args=None) -> torch.device:
    """Device that tensors handed to the process group must live on."""
    if dist.is_initialized() and dist.get_backend() == 'nccl':
        return torch.device('cuda', torch.cuda.current_device())
    return torch.device('cpu')


def distributed_init(args):
    """init_process_group + communicator warm-up + rank-0-only printing (distributed_utils.py:11-41)."""
    if args.distributed_world_size == 1:
        raise ValueError('Cannot initialize distributed with distributed_world_size=1')
    if dist.is_initialized():
        warnings.warn('Distributed is already initialized, cannot initialize twice!')
    else:
        backend = args.distributed_backend
        if backend == 'nccl' and (getattr(args, 'cpu', False) or not torch.cuda.is_available()):
            backend = 'gloo'   # CPU runs: the reference would crash here
            args.distributed_backend = backend
        print('| distributed init (rank {}): {}'.format(args.distributed_rank, args.distributed_init_method),
              flush=True)
        kwargs = {}
        if backend == 'nccl':
            kwargs['device_id'] = torch.device('cuda', torch.cuda.current_device())
        dist.init_process_group(
            backend=backend, init_method=args.distributed_init_method,
            world_size=args.distributed_world_size, rank=args.distributed_rank,
            timeout=datetime.timedelta(seconds=getattr(args, 'distributed_timeout', 1800)), **kwargs)
        print('| initialized host {} as rank {}'.format(socket.gethostname(), args.distributed_rank), flush=True)
        # dummy all-reduce: creates the NCCL communicator before the first timed step
        dist.all_reduce(torch.zeros(1, device=comm_device(args)))
        suppress_output(is_master(args))
    args.distributed_rank = dist.get_rank()
    return args.distributed_rank


def is_master(args) -> bool:
    return args.distributed_rank == 0


def suppress_output(is_master_: bool) -> None:
    """Replace ``print`` so that only rank 0 prints unless ``force=True`` is passed."""
    import builtins
    global _builtin_print
    if _builtin_print is None:
        _builtin_print = builtins.print
    base = _builtin_print

    def print(*args, **kwargs):
        force = kwargs.pop('force', False)
        if is_master_ or force:
            base(*args, **kwargs)

    builtins.print = print


def restore_output() -> None:
    import builtins
    if _builtin_print is not None:
        builtins.print = _builtin_print


def get_rank() -> int:
    return dist.get_rank() if dist.is_initialized() else 0


def get_world_size() -> int:
    return dist.get_world_size() if dist.is_initialized() else 1


def get_default_group():
    return dist.group.WORLD


def all_reduce(tensor, group=None):
    if group is None:
        group = get_default_group()
    return dist.all_reduce(tensor, group=group)


def barrier():
    if dist.is_initialized():
        if dist.get_backend() == 'nccl':
            dist.barrier(device_ids=[torch.cuda.current_device()])
        else:
            dist.barrier()


def all_gather_list(data: Any, group=None, max_size: int = 16384) -> List[Any]:
    """Gather arbitrary picklable data from every rank into a list ordered by rank.

    Same contract as distributed_utils.py:79-132 -- fixed ``max_size`` bytes per rank (ValueError
    beyond it), every rank must enter together -- but implemented as a real all-gather of
    length-prefixed byte slots instead of an all-reduce(SUM) over a zeroed ``max_size x world``
    buffer with a base-255 header, and without per-byte device->host syncs.
    """
    if not dist.is_initialized():
        return [data]
    if group is None:
        group = get_default_group()
    world_size = dist.get_world_size(group=group)
    enc = pickle.dumps(data)
    header = 4
    if len(enc) + header > max_size:
        raise ValueError('encoded data exceeds max_size: {}'.format(len(enc) + header))
    slot = torch.zeros(max_size, dtype=torch.uint8)
    slot[:header] = torch.tensor(list(len(enc).to_bytes(header, 'little')), dtype=torch.uint8)
    slot[header:header + len(enc)] = torch.frombuffer(bytearray(enc), dtype=torch.uint8)
    dev = comm_device()
    slot = slot.to(dev)
    out = torch.empty(max_size * world_size, dtype=torch.uint8, device=dev)
    dist.all_gather_into_tensor(out, slot, group=group) if dev.type == 'cuda' else \
        dist.all_gather(list(out.view(world_size, max_size).unbind(0)), slot, group=group)
    out = out.cpu().view(world_size, max_size)
    result = []
    try:
        for r in range(world_size):
            size = int.from_bytes(bytes(out[r, :header].tolist()), 'little')
            if size > max_size - header:
                raise pickle.UnpicklingError('corrupt length header')
            if size > 0:
                result.append(pickle.loads(out[r, header:header + size].numpy().tobytes()))
        return result
    except pickle.UnpicklingError:
        raise Exception(
            'Unable to unpickle data from other workers. all_gather_list requires all workers to enter the '
            'function together, so this error usually indicates that the workers have fallen out of sync '
            'somehow. Workers can fall out of sync if one of them runs out of memory, or if there are other '
            'conditions in your training script that can cause one worker to finish an epoch while other '
            'workers are still iterating over their portions of the data.')
