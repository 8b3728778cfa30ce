"""Per-step logging.

Keeps the reference's line format (progress_bar.py:113-138; sample in SURVEY.md section 5.5) so
existing log scrapers keep working, implements the ``json`` sink whose class the reference forgot
to define (progress_bar.py:20-21), and adds an optional JSON-lines file sink (``--log-file``).
"""
from __future__ import annotations

import json
import sys
from collections import OrderedDict
from numbers import Number

from hetseq_b200.utils.meters import AverageMeter, StopwatchMeter, TimeMeter


# How each kind of stat is rendered in the text line and what number the machine-readable sinks get. The text
# forms are the reference's (progress_bar.py:34-44): plain numbers '{:g}', an AverageMeter its average with three
# decimals, a TimeMeter its rounded rate, a StopwatchMeter its accumulated seconds with four decimals.
_RENDER = (
    (AverageMeter, lambda m: '{:.3f}'.format(m.avg), lambda m: m.avg),
    (TimeMeter, lambda m: '{:g}'.format(round(m.avg)), lambda m: m.avg),
    (StopwatchMeter, lambda m: '{:.4f}'.format(m.sum), lambda m: m.sum),
    (Number, lambda x: '{:g}'.format(x), lambda x: x),
)


def format_stat(stat):
    """Text form of one stat (anything that is neither a number nor a meter is passed through)."""
    for kind, text, _ in _RENDER:
        if isinstance(stat, kind):
            return text(stat)
    return stat


def stat_value(stat):
    """Numeric value of a stat for machine-readable sinks."""
    for kind, _, value in _RENDER:
        if isinstance(stat, kind):
            return value(stat)
    return None if stat is None else str(stat)


def render(stats, pair='{}={}', sep=', '):
    """'k=v, k=v' (step lines) or, with pair='{} {}' and sep=' | ', 'k v | k v' (summary lines)."""
    return sep.join(pair.format(k, str(format_stat(v)).strip()) for k, v in stats.items())


def build_progress_bar(args, iterator, epoch=None, prefix=None, default='simple', no_progress_bar='none'):
    fmt = getattr(args, 'log_format', None) or (no_progress_bar if getattr(args, 'no_progress_bar', False) else default)
    common = dict(epoch=epoch, prefix=prefix, log_file=getattr(args, 'log_file', None))
    if fmt in ('json', 'simple'):
        common['log_interval'] = args.log_interval
    try:
        cls = {'json': json_progress_bar, 'none': noop_progress_bar, 'simple': simple_progress_bar,
               'tqdm': tqdm_progress_bar}[fmt]
    except KeyError:
        raise ValueError('Unknown log format: {}'.format(fmt)) from None
    return cls(iterator, **common)


class progress_bar(object):
    """Base class: iterates the wrapped iterable, subclasses decide what to print. ``offset`` is where a resumed
    epoch starts counting."""

    def __init__(self, iterable, epoch=None, prefix=None, log_file=None):
        self.iterable = iterable
        self.offset = getattr(iterable, 'offset', 0)
        self.epoch = epoch
        parts = []
        if epoch is not None:
            parts.append('| epoch {:03d}'.format(epoch))
        if prefix is not None:
            parts.append(' | {}'.format(prefix))
        self.prefix = ''.join(parts)
        self.log_file = log_file

    def __len__(self):
        return len(self.iterable)

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False

    def __iter__(self):
        raise NotImplementedError

    def log(self, stats, tag='', step=None):
        raise NotImplementedError

    def print(self, stats, tag='', step=None):
        raise NotImplementedError

    def _format_stats(self, stats):
        return OrderedDict((k, str(format_stat(v))) for k, v in stats.items())

    def _json_record(self, stats, i=None, tag=''):
        rec = OrderedDict()
        if self.epoch is not None:
            rec['epoch'] = self.epoch
        if i is not None:
            rec['update'] = i
        if tag:
            rec['tag'] = tag
        for k, v in stats.items():
            rec[k] = stat_value(v)
        return rec

    def _sink(self, rec):
        if self.log_file:
            with open(self.log_file, 'a') as f:
                f.write(json.dumps(rec) + '\n')

    def _steps(self, log_interval):
        """Yield the items; after item number i (counted from the resume offset) call ``self._emit(i, size)`` when a
        line is due: stats have been logged, i > 0 and i is a multiple of the interval."""
        size = len(self.iterable)
        for i, obj in enumerate(self.iterable, start=self.offset):
            yield obj
            if self.stats is not None and i > 0 and log_interval is not None and i % log_interval == 0:
                self._emit(i, size)


class noop_progress_bar(progress_bar):
    def __iter__(self):
        return iter(self.iterable)

    def log(self, stats, tag='', step=None):
        if self.log_file:
            self._sink(self._json_record(stats, step, tag))

    def print(self, stats, tag='', step=None):
        pass


class simple_progress_bar(progress_bar):
    """'| epoch 001:      1 / 4 loss=3.676, ppl=12.78, ...' every ``log_interval`` steps."""

    def __init__(self, iterable, epoch=None, prefix=None, log_interval=1000, log_file=None):
        super().__init__(iterable, epoch, prefix, log_file)
        self.log_interval = log_interval
        self.stats = None

    def __iter__(self):
        return self._steps(self.log_interval)

    def _emit(self, i, size):
        print('{}:  {:5d} / {:d} {}'.format(self.prefix, i, size, render(self.stats)), flush=True)
        if self.log_file:
            self._sink(self._json_record(self.stats, i))

    def log(self, stats, tag='', step=None):
        self.stats = stats

    def print(self, stats, tag='', step=None):
        print('{} | {}'.format(self.prefix, render(stats, '{} {}', ' | ')), flush=True)


class json_progress_bar(progress_bar):
    """One JSON object per logged step on stdout."""

    def __init__(self, iterable, epoch=None, prefix=None, log_interval=1000, log_file=None):
        super().__init__(iterable, epoch, prefix, log_file)
        self.log_interval = log_interval
        self.stats = None

    def __iter__(self):
        return self._steps(self.log_interval)

    def _emit(self, i, size):
        rec = self._json_record(self.stats, i)
        if self.epoch is not None:
            rec['epoch_progress'] = round(self.epoch - 1 + i / float(size), 4)
        print(json.dumps(rec), flush=True)
        self._sink(rec)

    def log(self, stats, tag='', step=None):
        self.stats = stats

    def print(self, stats, tag='', step=None):
        rec = self._json_record(stats, tag=tag)
        print(json.dumps(rec), flush=True)
        self._sink(rec)


class tqdm_progress_bar(progress_bar):
    """Interactive bar on stderr with the current stats as postfix. The reference lists ``tqdm`` among its log
    formats but never defines the class (progress_bar.py:24-27); needs the optional ``tqdm`` package."""

    def __init__(self, iterable, epoch=None, prefix=None, log_file=None):
        super().__init__(iterable, epoch, prefix, log_file)
        try:
            from tqdm import tqdm
        except ImportError as e:
            raise ImportError('--log-format tqdm needs the tqdm package') from e
        self.tqdm = tqdm(iterable, self.prefix, leave=False, file=sys.stderr, initial=self.offset,
                         total=len(iterable) if hasattr(iterable, '__len__') else None)

    def __iter__(self):
        return iter(self.tqdm)

    def log(self, stats, tag='', step=None):
        self.tqdm.set_postfix(self._format_stats(stats), refresh=False)
        if self.log_file:
            self._sink(self._json_record(stats, step, tag))

    def print(self, stats, tag='', step=None):
        self.tqdm.write('{} | {}'.format(self.tqdm.desc, render(stats, '{} {}', ' | ')), file=sys.stdout)
