"""Learning-rate schedules (behaviour of the reference's lr_scheduler.py:1-105, checked against it update by update in
tests/test_reference_parity.py).

Design: a schedule is a pure function ``lr(base_lr, num_updates) -> float`` (``polynomial_lr`` / ``warmup_constant_lr``)
plus a small amount of epoch state (which entry of the per-epoch ``--lr`` list is the base). The classes below only
hold that state, push the value into the optimizer (from where it reaches the fused update kernels through a
device scalar, so a schedule never forces a CUDA-graph re-capture) and carry ``best`` for checkpoints.

Observable contract kept from the reference:
* the constructor sets ``lr = lr[0] / warmup_updates`` when there is a warm-up (else ``lr[0]``);
* ``step(epoch)`` selects ``lr[min(epoch, len - 1)]`` unless ``--force-anneal`` has been reached, re-applies the LAST
  warm-up factor and returns the learning rate; it also tracks the best validation loss;
* ``step_update(n)``: linear warm-up while ``n <= warmup``, ``--end-learning-rate`` from ``--total-num-update`` on,
  otherwise ``(lr - end) * (1 - (n - warmup) / (total - warmup)) ** power + end``;
* ``state_dict() == {'best': ...}``.
"""
from __future__ import annotations

from typing import Optional, Sequence

from hetseq_b200.registry import register_lr_scheduler


# ---- pure schedule math -------------------------------------------------------------------------------
def warmup_fraction(num_updates: int, warmup: int) -> Optional[float]:
    """Fraction of the base LR during linear warm-up, or None once warm-up is over (or absent)."""
    if warmup > 0 and num_updates <= warmup:
        return num_updates / float(warmup)
    return None


def polynomial_lr(base_lr: float, num_updates: int, warmup: int, total: int, end_lr: float, power: float) -> float:
    frac = warmup_fraction(num_updates, warmup)
    if frac is not None:
        return frac * base_lr
    if num_updates >= total:
        return end_lr
    remaining = 1 - (num_updates - warmup) / (total - warmup)
    return (base_lr - end_lr) * remaining ** power + end_lr


def epoch_base_lr(lrs: Sequence[float], epoch: int, force_anneal: Optional[int], current: float) -> float:
    """Entry of the per-epoch ``--lr`` list for ``epoch``; frozen at ``current`` once force-anneal is reached."""
    if force_anneal is not None and epoch >= force_anneal:
        return current
    return lrs[min(epoch, len(lrs) - 1)]


# ---- stateful wrappers ----------------------------------------------------------------------------------
class _LRScheduler(object):
    """Holds the optimizer, the best validation loss seen, and the checkpointable state."""

    def __init__(self, args, optimizer):
        for needed in ('set_lr', 'get_lr'):
            if not hasattr(optimizer, needed):
                raise ValueError('optimizer must expose get_lr()/set_lr() (an hetseq_b200.optim._Optimizer)')
        self.args, self.optimizer, self.best = args, optimizer, None

    @staticmethod
    def add_args(parser):
        pass

    def _apply(self, lr: float) -> float:
        self.optimizer.set_lr(lr)
        return self.optimizer.get_lr()

    def _track_best(self, val_loss) -> None:
        if val_loss is not None:
            self.best = val_loss if self.best is None else min(self.best, val_loss)

    def state_dict(self):
        return {'best': self.best}

    def load_state_dict(self, state_dict):
        self.best = state_dict['best']

    def step(self, epoch, val_loss=None):
        """End of an epoch."""
        self._track_best(val_loss)

    def step_update(self, num_updates):
        """After every optimizer update."""
        return self.optimizer.get_lr()


def _common_args(parser, polynomial: bool) -> None:
    parser.add_argument('--force-anneal', '--fa', type=int, metavar='N', help='force annealing at specified epoch')
    parser.add_argument('--warmup-updates', default=0, type=int, metavar='N',
                        help='warmup the learning rate linearly for the first N updates')
    if polynomial:
        parser.add_argument('--end-learning-rate', default=0.0, type=float)
        parser.add_argument('--power', default=1.0, type=float)
        parser.add_argument('--total-num-update', default=1000000, type=int)


@register_lr_scheduler('PolynomialDecayScheduler')
class PolynomialDecayScheduler(_LRScheduler):
    def __init__(self, args, optimizer):
        super().__init__(args, optimizer)
        args.warmup_updates = getattr(args, 'warmup_updates', 0) or 0
        self.warmup = args.warmup_updates
        self.lr = args.lr[0]                              # base LR of the current epoch
        self.end_learning_rate, self.total_num_update, self.power = \
            args.end_learning_rate, args.total_num_update, args.power
        # factor applied by step(): 1/warmup before the first update, then the last warm-up fraction
        self.warmup_factor = 1. / self.warmup if self.warmup > 0 else 1
        self._apply(self.warmup_factor * self.lr)

    @staticmethod
    def add_args(parser):
        _common_args(parser, polynomial=True)

    def get_next_lr(self, epoch):
        return epoch_base_lr(self.args.lr, epoch, self.args.force_anneal, self.optimizer.get_lr())

    def step(self, epoch, val_loss=None):
        self._track_best(val_loss)
        self.lr = self.get_next_lr(epoch)
        return self._apply(self.warmup_factor * self.lr)

    def step_update(self, num_updates):
        frac = warmup_fraction(num_updates, self.warmup)
        if frac is not None:
            self.warmup_factor = frac
        return self._apply(polynomial_lr(self.lr, num_updates, self.warmup, self.total_num_update,
                                         self.end_learning_rate, self.power))


@register_lr_scheduler('fixed')
class FixedScheduler(_LRScheduler):
    """Constant epoch-indexed LR with optional linear warm-up (handy for tests and MNIST); not in the reference."""

    def __init__(self, args, optimizer):
        super().__init__(args, optimizer)
        self.lr = args.lr[0]
        self.warmup = getattr(args, 'warmup_updates', 0) or 0
        self._apply(self.lr if self.warmup == 0 else self.lr / self.warmup)

    @staticmethod
    def add_args(parser):
        _common_args(parser, polynomial=False)

    def step(self, epoch, val_loss=None):
        self._track_best(val_loss)
        self.lr = self.args.lr[min(epoch, len(self.args.lr) - 1)]
        return self._apply(self.lr)

    def step_update(self, num_updates):
        frac = warmup_fraction(num_updates, self.warmup)
        if frac is not None:
            self.optimizer.set_lr(self.lr * frac)
        return self.optimizer.get_lr()
