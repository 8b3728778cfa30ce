"""Learning-rate schedules (reference: lr_scheduler.py:1-105).

``PolynomialDecayScheduler``: linear warm-up over ``--warmup-updates`` then polynomial decay to
``--end-learning-rate`` at ``--total-num-update``; ``step(epoch)`` switches to the epoch-indexed
``--lr`` list unless ``--force-anneal`` has been reached.
"""
from __future__ import annotations

from hetseq_b200.registry import register_lr_scheduler


class _LRScheduler(object):
    def __init__(self, args, optimizer):
        super().__init__()
        if not hasattr(optimizer, 'set_lr') or not hasattr(optimizer, 'get_lr'):
            raise ValueError('optimizer must expose get_lr()/set_lr() (an hetseq_b200.optim._Optimizer)')
        self.args = args
        self.optimizer = optimizer
        self.best = None

    @staticmethod
    def add_args(parser):
        pass

    def state_dict(self):
        return {'best': self.best}

    def load_state_dict(self, state_dict):
        self.best = state_dict['best']

    def step(self, epoch, val_loss=None):
        """Update the learning rate at the end of the given epoch."""
        if val_loss is not None:
            if self.best is None:
                self.best = val_loss
            else:
                self.best = min(self.best, val_loss)

    def step_update(self, num_updates):
        """Update the learning rate after each update."""
        return self.optimizer.get_lr()


@register_lr_scheduler('PolynomialDecayScheduler')
class PolynomialDecayScheduler(_LRScheduler):
    def __init__(self, args, optimizer):
        super().__init__(args, optimizer)
        args.warmup_updates = getattr(args, 'warmup_updates', 0) or 0
        self.lr = args.lr[0]
        self.warmup_factor = 1. / args.warmup_updates if args.warmup_updates > 0 else 1
        self.end_learning_rate = args.end_learning_rate
        self.total_num_update = args.total_num_update
        self.power = args.power
        self.optimizer.set_lr(self.warmup_factor * self.lr)

    @staticmethod
    def add_args(parser):
        parser.add_argument('--force-anneal', '--fa', type=int, metavar='N',
                            help='force annealing at specified epoch')
        parser.add_argument('--warmup-updates', default=0, type=int, metavar='N',
                            help='warmup the learning rate linearly for the first N updates')
        parser.add_argument('--end-learning-rate', default=0.0, type=float)
        parser.add_argument('--power', default=1.0, type=float)
        parser.add_argument('--total-num-update', default=1000000, type=int)

    def get_next_lr(self, epoch):
        lrs = self.args.lr
        if self.args.force_anneal is None or epoch < self.args.force_anneal:
            return lrs[min(epoch, len(lrs) - 1)]
        return self.optimizer.get_lr()

    def step(self, epoch, val_loss=None):
        super().step(epoch, val_loss)
        self.lr = self.get_next_lr(epoch)
        self.optimizer.set_lr(self.warmup_factor * self.lr)
        return self.optimizer.get_lr()

    def step_update(self, num_updates):
        warmup = self.args.warmup_updates
        if warmup > 0 and num_updates <= warmup:
            self.warmup_factor = num_updates / float(warmup)
            lr = self.warmup_factor * self.lr
        elif num_updates >= self.total_num_update:
            lr = self.end_learning_rate
        else:
            lr_range = self.lr - self.end_learning_rate
            pct_remaining = 1 - (num_updates - warmup) / (self.total_num_update - warmup)
            lr = lr_range * pct_remaining ** self.power + self.end_learning_rate
        self.optimizer.set_lr(lr)
        return self.optimizer.get_lr()


@register_lr_scheduler('fixed')
class FixedScheduler(_LRScheduler):
    """Constant epoch-indexed LR with optional linear warm-up (handy for tests and MNIST)."""

    def __init__(self, args, optimizer):
        super().__init__(args, optimizer)
        self.lr = args.lr[0]
        self.warmup = getattr(args, 'warmup_updates', 0) or 0
        self.optimizer.set_lr(self.lr if self.warmup == 0 else self.lr / self.warmup)

    @staticmethod
    def add_args(parser):
        parser.add_argument('--force-anneal', '--fa', type=int, metavar='N')
        parser.add_argument('--warmup-updates', default=0, type=int, metavar='N')

    def step(self, epoch, val_loss=None):
        super().step(epoch, val_loss)
        self.lr = self.args.lr[min(epoch, len(self.args.lr) - 1)]
        self.optimizer.set_lr(self.lr)
        return self.optimizer.get_lr()

    def step_update(self, num_updates):
        if self.warmup > 0 and num_updates <= self.warmup:
            self.optimizer.set_lr(self.lr * num_updates / float(self.warmup))
        return self.optimizer.get_lr()
