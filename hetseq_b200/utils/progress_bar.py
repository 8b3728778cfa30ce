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


def build_progress_bar(args, iterator, epoch=None, prefix=None, default='simple', no_progress_bar='none'):
    fmt = getattr(args, 'log_format', None)
    if fmt is None:
        fmt = no_progress_bar if getattr(args, 'no_progress_bar', False) else default
    log_file = getattr(args, 'log_file', None)
    if fmt == 'json':
        return json_progress_bar(iterator, epoch, prefix, args.log_interval, log_file)
    if fmt == 'none':
        return noop_progress_bar(iterator, epoch, prefix, log_file)
    if fmt == 'simple':
        return simple_progress_bar(iterator, epoch, prefix, args.log_interval, log_file)
    if fmt == 'tqdm':
        return tqdm_progress_bar(iterator, epoch, prefix, log_file)
    raise ValueError('Unknown log format: {}'.format(fmt))


def format_stat(stat):
    """Numbers '{:g}', AverageMeter '.3f' of avg, TimeMeter rounded avg, StopwatchMeter '.4f' of sum."""
    if isinstance(stat, Number):
        stat = '{:g}'.format(stat)
    elif isinstance(stat, AverageMeter):
        stat = '{:.3f}'.format(stat.avg)
    elif isinstance(stat, TimeMeter):
        stat = '{:g}'.format(round(stat.avg))
    elif isinstance(stat, StopwatchMeter):
        stat = '{:.4f}'.format(stat.sum)
    return stat


def stat_value(stat):
    """Numeric value of a stat for machine-readable sinks."""
    if isinstance(stat, AverageMeter):
        return stat.avg
    if isinstance(stat, TimeMeter):
        return stat.avg
    if isinstance(stat, StopwatchMeter):
        return stat.sum
    if isinstance(stat, Number):
        return stat
    return None if stat is None else str(stat)


class progress_bar(object):
    """Base class: iterates the wrapped iterable, subclasses decide what to print."""

    def __init__(self, iterable, epoch=None, prefix=None, log_file=None):
        self.iterable = iterable
        self.offset = getattr(iterable, 'offset', 0)
        self.epoch = epoch
        self.prefix = ''
        if epoch is not None:
            self.prefix += '| epoch {:03d}'.format(epoch)
        if prefix is not None:
            self.prefix += ' | {}'.format(prefix)
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

    def _str_commas(self, stats):
        return ', '.join(key + '=' + stats[key].strip() for key in stats.keys())

    def _str_pipes(self, stats):
        return ' | '.join(key + ' ' + stats[key].strip() for key in stats.keys())

    def _format_stats(self, stats):
        postfix = OrderedDict(stats)
        for key in postfix.keys():
            postfix[key] = str(format_stat(postfix[key]))
        return postfix

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


class noop_progress_bar(progress_bar):
    def __iter__(self):
        for obj in self.iterable:
            yield obj

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
        self._raw = None

    def __iter__(self):
        size = len(self.iterable)
        for i, obj in enumerate(self.iterable, start=self.offset):
            yield obj
            if self.stats is not None and i > 0 and self.log_interval is not None \
                    and i % self.log_interval == 0:
                postfix = self._str_commas(self.stats)
                print('{}:  {:5d} / {:d} {}'.format(self.prefix, i, size, postfix), flush=True)
                if self.log_file:
                    self._sink(self._json_record(self._raw, i))

    def log(self, stats, tag='', step=None):
        self._raw = stats
        self.stats = self._format_stats(stats)

    def print(self, stats, tag='', step=None):
        postfix = self._str_pipes(self._format_stats(stats))
        print('{} | {}'.format(self.prefix, postfix), flush=True)


class json_progress_bar(progress_bar):
    """One JSON object per logged step on stdout."""

    def __init__(self, iterable, epoch=None, prefix=None, log_interval=1000, log_file=None):
        super().__init__(iterable, epoch, prefix, log_file)
        self.log_interval = log_interval
        self.stats = None

    def __iter__(self):
        size = float(len(self.iterable))
        for i, obj in enumerate(self.iterable, start=self.offset):
            yield obj
            if self.stats is not None and i > 0 and self.log_interval is not None \
                    and i % self.log_interval == 0:
                rec = self._json_record(self.stats, i)
                if self.epoch is not None:
                    rec['epoch_progress'] = round(self.epoch - 1 + float(i / size), 4)
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
        postfix = self._str_pipes(self._format_stats(stats))
        self.tqdm.write('{} | {}'.format(self.tqdm.desc, postfix), file=sys.stdout)
