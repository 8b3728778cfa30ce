"""Command-line flags.

Every flag of the reference (options.py:5-290, train.py:197-205; complete list in SURVEY.md
section 5.6) is accepted with the same spelling and default, including the ones the reference
parses but never reads. Differences, all deliberate:

* component-specific flags come from the registered component's ``add_args`` (registry.py)
  instead of if/elif chains (options.py:61-85,166-202);
* ``--lr``, ``--update-freq``, ``--adam-betas`` and ``--optimizer-overrides`` are parsed with
  ``ast.literal_eval`` instead of Python ``eval`` (options.py:263-271) -- same accepted syntax;
* both ``--lr-scheduler`` and ``--lr_scheduler`` select the scheduler (train.py:202 vs options.py:190);
* new: the mixed-precision family (``--fp16``/``--bf16`` + loss-scale flags; the reference has
  ``--fp16`` commented out at options.py:11), ``--dp-backend``, ``--sample-size-mode``,
  ``--kernels``, ``--cuda-graph``, ``--synthetic-*`` and ``--log-format json``.
"""
from __future__ import annotations

import argparse
import ast
from typing import List, Optional, Sequence

import torch

from hetseq_b200 import registry


# ------------------------------------------------------------------------------------------------
# value parsers
# ------------------------------------------------------------------------------------------------
def literal(x: str):
    """Safe replacement for the reference's ``eval`` on CLI strings."""
    try:
        return ast.literal_eval(x)
    except (ValueError, SyntaxError) as exc:
        raise argparse.ArgumentTypeError(f"cannot parse {x!r}: {exc}") from None


def eval_str_list(x, type=float) -> Optional[List]:
    """'0.1' -> [0.1]; '0.1,0.2' / '[0.1, 0.2]' / '(0.1, 0.2)' -> [0.1, 0.2]  (options.py:263-271)."""
    if x is None:
        return None
    if isinstance(x, str):
        x = literal(x)
    try:
        return list(map(type, x))
    except TypeError:
        return [type(x)]


def eval_bool(x, default=False) -> bool:
    if x is None:
        return default
    try:
        return bool(literal(x))
    except (TypeError, argparse.ArgumentTypeError):
        return default


# ------------------------------------------------------------------------------------------------
# flag groups
# ------------------------------------------------------------------------------------------------
def add_general_args(parser):
    parser.add_argument('--no-progress-bar', action='store_true', help='disable progress bar')
    parser.add_argument('--seed', default=19940802, type=int, metavar='N',
                        help='pseudo random number generator seed')
    parser.add_argument('--cpu', action='store_true', help='use CPU instead of CUDA')
    parser.add_argument('--log-interval', type=int, default=1, metavar='N',
                        help='log progress every N batches (when progress bar is disabled)')
    parser.add_argument('--log-format', default='simple', choices=['none', 'simple', 'json', 'tqdm'],
                        help='log format to use')
    parser.add_argument('--log-file', default=None, metavar='PATH',
                        help='also append one JSON object per logged step to this file')

    group = parser.add_argument_group('Precision and kernels')
    group.add_argument('--fp16', action='store_true',
                       help='fp16 parameters/activations/gradients with fp32 master weights and '
                            'optimizer state and dynamic loss scaling')
    group.add_argument('--bf16', action='store_true',
                       help='bf16 parameters/activations with fp32 master weights (no loss scaling)')
    group.add_argument('--fp16-init-scale', default=2 ** 15, type=float,
                       help='initial loss scale for --fp16')
    group.add_argument('--fp16-scale-window', default=1000, type=int,
                       help='double the loss scale after this many overflow-free updates')
    group.add_argument('--fp16-scale-factor', default=2.0, type=float,
                       help='factor by which the loss scale grows / shrinks')
    group.add_argument('--min-loss-scale', default=1e-4, type=float,
                       help='stop training if the loss scale falls below this value')
    group.add_argument('--kernels', default='auto', choices=['auto', 'native', 'torch'],
                       help="'native': hand-written sm_100a kernels (error if unavailable); "
                            "'torch': plain PyTorch ops (the CPU path and the oracle); "
                            "'auto': native on CUDA, torch on CPU")
    group.add_argument('--cuda-graph', action='store_true',
                       help='capture forward+backward+update of a step in a CUDA graph and replay it')
    group.add_argument('--profile-nvtx', action='store_true',
                       help='emit NVTX ranges per phase (forward/backward/reduce/update)')
    group.add_argument('--check-params-interval', type=int, default=0, metavar='N',
                       help='every N updates verify that all ranks hold bit-identical parameters (debug aid for '
                            'the in-kernel gradient path; replaces the grad-norm equality assert of the reference)')
    return parser


def add_dataset_args(parser, train=True):
    group = parser.add_argument_group('Dataset and data loading')
    group.add_argument('--num-workers', default=0, type=int, metavar='N',
                       help='how many subprocesses to use for data loading')
    group.add_argument('--max-tokens', type=int, metavar='N',
                       help='maximum number of tokens in a batch')
    group.add_argument('--max-sentences', '--batch-size', type=int, metavar='N',
                       help='maximum number of sentences in a batch')
    group.add_argument('--required-batch-size-multiple', default=1, type=int, metavar='N',
                       help='batch size will be a multiplier of this value')
    if train:
        group.add_argument('--train-subset', default='train', metavar='SPLIT',
                           choices=['train', 'valid', 'test'],
                           help='data subset to use for training (train, valid, test)')
        group.add_argument('--valid-subset', default='valid', metavar='SPLIT',
                           help='comma separated list of data subsets to use for validation')
        group.add_argument('--validate-interval', type=int, default=1, metavar='N',
                           help='validate every N epochs (only with --enable-validation)')
        group.add_argument('--disable-validation', action='store_true', help='disable validation')
        group.add_argument('--enable-validation', action='store_true',
                           help='run validation at the end of each epoch (the reference never does: '
                                'train.py:93)')
        group.add_argument('--max-tokens-valid', type=int, metavar='N',
                           help='maximum number of tokens in a validation batch (defaults to --max-tokens)')
        group.add_argument('--max-sentences-valid', type=int, metavar='N',
                           help='maximum number of sentences in a validation batch '
                                '(defaults to --max-sentences)')
        group.add_argument('--curriculum', default=0, type=int, metavar='N',
                           help='don\'t shuffle batches for first N epochs')
        group.add_argument('--sample-size-mode', default='batch', choices=['reference', 'batch'],
                           help="gradient denominator. 'reference' reproduces tasks.py:161-166 "
                                "(len of the first example: seq-len for BERT, 1 for MNIST); "
                                "'batch' uses the number of sequences")
    return group


def add_distributed_training_args(parser):
    group = parser.add_argument_group('Distributed training')
    group.add_argument('--distributed-world-size', type=int, metavar='N',
                       default=max(1, torch.cuda.device_count()),
                       help='total number of GPUs across all nodes (default: all visible GPUs)')
    group.add_argument('--distributed-rank', default=0, type=int, help='rank of the current GPU')
    group.add_argument('--distributed-gpus', default=4, type=int,
                       help='number of gpus used in the current worker/node')
    group.add_argument('--distributed-backend', default='nccl', type=str, help='distributed backend')
    group.add_argument('--distributed-init-method', default=None, type=str,
                       help='typically tcp://hostname:port (or file:///shared/path) used to '
                            'establish the initial connection')
    group.add_argument('--device-id', '--local_rank', default=0, type=int,
                       help='which GPU to use (usually configured automatically)')
    group.add_argument('--distributed-no-spawn', action='store_true',
                       help='do not spawn multiple processes even if multiple GPUs are visible')
    group.add_argument('--ddp-backend', default='c10d', type=str, choices=['c10d'],
                       help='accepted for compatibility; see --dp-backend')
    group.add_argument('--dp-backend', default='auto', choices=['auto', 'fused', 'nccl', 'sharded'],
                       help="gradient path. 'fused': in-kernel reduce-scatter over NVLink peer memory + "
                            "clip/unscale + sharded Adam + all-gather (no NCCL call); 'nccl': bucketed "
                            "all-reduce + replicated optimizer (the baseline); 'sharded': the fused protocol "
                            "emulated with torch.distributed collectives (CPU tests); 'auto': fused when "
                            "the ranks share one NVLink domain and CUDA, else nccl")
    group.add_argument('--fused-transport', default='auto', choices=['auto', 'multimem', 'p2p'],
                       help="how the fused gradient path moves data: 'multimem' = NVLS (multimem.ld_reduce / multimem.st "
                            "through the NVSwitch), 'p2p' = peer loads / stores; 'auto' = multimem from 4 GPUs on when "
                            "the symmetric buffers have a multicast mapping")
    group.add_argument('--fused-rs-blocks', default=0, type=int, metavar='N',
                       help='CTAs of a reduce-scatter launch of the fused gradient path (0 = chosen by world size)')
    group.add_argument('--bucket-cap-mb', default=25, type=int, metavar='MB',
                       help='bucket size for reduction')
    group.add_argument('--fix-batches-to-gpus', action='store_true',
                       help='don\'t shuffle batches between GPUs')
    group.add_argument('--find-unused-parameters', default=False, action='store_true',
                       help='tolerate parameters that receive no gradient in a step')
    group.add_argument('--fast-stat-sync', default=False, action='store_true',
                       help='sync the fixed set of stats with one 6-element all-reduce')
    group.add_argument('--distributed-timeout', default=1800, type=int, metavar='SEC',
                       help='process-group timeout in seconds; also bounds every in-kernel wait on a peer of the '
                            'fused gradient path (beyond it the kernel reports the peer and traps)')
    return group


def add_optimization_args(parser):
    group = parser.add_argument_group('Optimization')
    group.add_argument('--max-epoch', '--me', default=0, type=int, metavar='N',
                       help='force stop training at specified epoch')
    group.add_argument('--max-update', '--mu', default=0, type=int, metavar='N',
                       help='force stop training at specified update')
    group.add_argument('--clip-norm', default=25, type=float, metavar='NORM',
                       help='clip threshold of gradients')
    group.add_argument('--update-freq', default='1', metavar='N1,N2,...,N_K',
                       type=lambda uf: eval_str_list(uf, type=int),
                       help='update parameters every N_i batches, when in epoch i')
    group.add_argument('--lr', '--learning-rate', default='0.25', type=eval_str_list,
                       metavar='LR_1,LR_2,...,LR_N',
                       help='learning rate for the first N epochs; all epochs >N using LR_N')
    group.add_argument('--min-lr', default=-1, type=float, metavar='LR',
                       help='stop training when the learning rate reaches this minimum')
    group.add_argument('--use-bmuf', default=False, action='store_true',
                       help='rejected: the reference flag only disables gradient sync (controller.py:77)')
    return group


def add_checkpoint_args(parser):
    group = parser.add_argument_group('Checkpointing')
    group.add_argument('--save-dir', metavar='DIR', default='checkpoints', help='path to save checkpoints')
    group.add_argument('--restore-file', default='checkpoint_last.pt',
                       help='filename from which to load checkpoint (default: <save-dir>/checkpoint_last.pt')
    group.add_argument('--reset-dataloader', action='store_true',
                       help='if set, does not reload dataloader state from the checkpoint')
    group.add_argument('--reset-lr-scheduler', action='store_true',
                       help='if set, does not load lr scheduler state from the checkpoint')
    group.add_argument('--reset-meters', action='store_true',
                       help='if set, does not load meters from the checkpoint')
    group.add_argument('--reset-optimizer', action='store_true',
                       help='if set, does not load optimizer state from the checkpoint')
    group.add_argument('--optimizer-overrides', default="{}", type=str, metavar='DICT',
                       help='a dictionary used to override optimizer args when loading a checkpoint')
    group.add_argument('--save-interval', type=int, default=1, metavar='N',
                       help='save a checkpoint every N epochs')
    group.add_argument('--save-interval-updates', type=int, default=0, metavar='N',
                       help='save a checkpoint every N updates')
    group.add_argument('--keep-interval-updates', type=int, default=-1, metavar='N',
                       help='keep the last N checkpoints saved with --save-interval-updates')
    group.add_argument('--keep-last-epochs', type=int, default=-1, metavar='N',
                       help='keep last N epoch checkpoints')
    group.add_argument('--no-save', action='store_true', help='don\'t save models or checkpoints')
    group.add_argument('--no-epoch-checkpoints', action='store_true',
                       help='only store last and best checkpoints')
    group.add_argument('--no-last-checkpoints', action='store_true', help='don\'t store last checkpoints')
    group.add_argument('--no-save-optimizer-state', action='store_true',
                       help='don\'t save optimizer-state as part of checkpoint')
    group.add_argument('--best-checkpoint-metric', type=str, default='loss',
                       help='metric to use for saving "best" checkpoints')
    group.add_argument('--maximize-best-checkpoint-metric', action='store_true',
                       help='select the largest metric value for saving "best" checkpoints')
    return group


# ------------------------------------------------------------------------------------------------
# parser assembly
# ------------------------------------------------------------------------------------------------
DEFAULT_LR_SCHEDULER = 'PolynomialDecayScheduler'


def get_preparser() -> argparse.ArgumentParser:
    """Stage 1 (train.py:197-205): find out which components contribute flags."""
    registry.ensure_builtin_components_loaded()
    pre = argparse.ArgumentParser(allow_abbrev=False, add_help=False)
    pre.add_argument('--task', type=str, default='bert', choices=sorted(registry.TASKS))
    pre.add_argument('--optimizer', type=str, default='adam', choices=sorted(registry.OPTIMIZERS))
    pre.add_argument('--lr-scheduler', '--lr_scheduler', dest='lr_scheduler', type=str,
                     default=DEFAULT_LR_SCHEDULER)
    return pre


def get_training_parser(task='bert', optimizer='adam', lr_scheduler=DEFAULT_LR_SCHEDULER):
    """Stage 2 (options.py:5-25): the full parser for the chosen components."""
    registry.ensure_builtin_components_loaded()
    parser = argparse.ArgumentParser(allow_abbrev=False, prog='hetseq_b200.train')
    add_general_args(parser)
    add_dataset_args(parser, train=True)
    add_distributed_training_args(parser)
    add_optimization_args(parser)
    add_checkpoint_args(parser)

    parser.add_argument('--task', type=str, default=task)
    parser.add_argument('--optimizer', type=str, default=optimizer)
    parser.add_argument('--lr-scheduler', '--lr_scheduler', dest='lr_scheduler', type=str,
                        default=lr_scheduler)
    registry.lookup('task', task).add_args(parser)
    registry.lookup('optimizer', optimizer).add_args(parser)
    registry.lookup('lr_scheduler', lr_scheduler).add_args(parser)
    return parser


def parse_args_and_arch(parser, input_args: Optional[Sequence[str]] = None):
    """Post-processing of options.py:283-290 plus validation of the new flags."""
    args = parser.parse_args(input_args)
    if getattr(args, 'max_sentences_valid', None) is None:
        args.max_sentences_valid = args.max_sentences
    if getattr(args, 'max_tokens_valid', None) is None:
        args.max_tokens_valid = args.max_tokens
    if args.fp16 and args.bf16:
        parser.error('--fp16 and --bf16 are mutually exclusive')
    if args.use_bmuf:
        parser.error('--use-bmuf is not supported: in the reference it only switches gradient '
                     'synchronisation off (controller.py:77,324), so ranks silently diverge')
    return args


def parse_cli(argv: Optional[Sequence[str]] = None):
    """Two-stage parse used by ``python -m hetseq_b200.train`` (train.py:196-211)."""
    pre_args, rest = get_preparser().parse_known_args(argv)
    parser = get_training_parser(pre_args.task, pre_args.optimizer, pre_args.lr_scheduler)
    args = parse_args_and_arch(parser, rest)
    # values eaten by the pre-parser win over the main parser's defaults
    args.task, args.optimizer, args.lr_scheduler = pre_args.task, pre_args.optimizer, pre_args.lr_scheduler
    return args


def make_args(task='mnist', optimizer='adam', lr_scheduler=DEFAULT_LR_SCHEDULER,
              argv: Optional[Sequence[str]] = None, **overrides):
    """Programmatic construction of a fully-defaulted args namespace (tests, bench.py)."""
    parser = get_training_parser(task, optimizer, lr_scheduler)
    argv = list(argv or [])
    args = parse_args_and_arch(parser, argv)
    args.task, args.optimizer, args.lr_scheduler = task, optimizer, lr_scheduler
    for k, v in overrides.items():
        setattr(args, k, v)
    return args
