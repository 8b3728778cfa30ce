"""Resumable, epoch- and shard-aware iterators (reference: data/iterators.py).

Same public classes and state contract as the reference --
``EpochBatchIterator.state_dict() == {'epoch', 'iterations_in_epoch'}``, per-epoch batch shuffle
seeded with ``seed + epoch`` identically on every rank, round-robin sharding padded with empty
batches so every rank takes the same number of steps -- built on an index plan (numpy arrays)
instead of lists of lists, plus a pinned-memory prefetch thread for the host->device hand-off
(SURVEY.md W13).
"""
from __future__ import annotations

import itertools
import math
import os
import queue
import threading
from typing import Optional

import numpy as np
import torch

from hetseq_b200.data import data_utils


class CountingIterator(object):
    """Iterator wrapper that counts consumed elements (data/iterators.py:10-42)."""

    def __init__(self, iterable, start=0, length=None):
        self.iterable = iterable
        self.count = start
        self.itr = iter(self)
        self.len = start + (len(iterable) if length is None else length)

    def __len__(self):
        return self.len

    def __iter__(self):
        for x in self.iterable:
            self.count += 1
            yield x

    def __next__(self):
        return next(self.itr)

    def has_next(self):
        return self.count < len(self)

    def skip(self, num_to_skip):
        next(itertools.islice(self.itr, num_to_skip, num_to_skip), None)
        return self


class _BatchLoader(object):
    """Turns a list of index batches into collated samples; optional DataLoader workers."""

    def __init__(self, dataset, collate_fn, batches, num_workers=0):
        self.dataset, self.collate_fn, self.batches, self.num_workers = dataset, collate_fn, batches, num_workers

    def __len__(self):
        return len(self.batches)

    def __iter__(self):
        if self.num_workers > 0:
            os.environ['PYTHONWARNINGS'] = 'ignore:semaphore_tracker:UserWarning'
            yield from torch.utils.data.DataLoader(
                self.dataset, collate_fn=self.collate_fn,
                batch_sampler=[list(map(int, b)) for b in self.batches], num_workers=self.num_workers)
            return
        fetch_batch = getattr(self.dataset, 'fetch_batch', None)
        for b in self.batches:
            if fetch_batch is not None:
                yield fetch_batch(b)           # vectorised path: dataset builds the whole batch
            else:
                yield self.collate_fn([self.dataset[int(i)] for i in b])


class EpochBatchIterator(object):
    """Multi-epoch iterator over a dataset with frozen batches (data/iterators.py:67-211)."""

    def __init__(self, dataset, collate_fn, batch_sampler, seed=1, num_shards=1, shard_id=0,
                 num_workers=0, epoch=0):
        self.dataset = dataset
        self.collate_fn = collate_fn
        if isinstance(batch_sampler, data_utils.BatchPlan):
            self.frozen_batches = batch_sampler
        else:
            self.frozen_batches = tuple(batch_sampler)
        self.seed = seed
        self.num_shards = num_shards
        self.shard_id = shard_id
        self.num_workers = num_workers
        self.epoch = epoch
        self._cur_epoch_itr = None
        self._next_epoch_itr = None
        #: set by a PrefetchIterator that reads ahead of the trainer: callable(counting_iterator) -> number of
        #: batches the TRAINER has consumed (the raw count includes batches still waiting in the prefetch queue)
        self.position_hook = None

    def __len__(self):
        return len(self.frozen_batches)

    def next_epoch_itr(self, shuffle=True, fix_batches_to_gpus=False):
        if self._next_epoch_itr is not None:
            self._cur_epoch_itr = self._next_epoch_itr
            self._next_epoch_itr = None
        else:
            self.epoch += 1
            self._cur_epoch_itr = self._get_iterator_for_epoch(
                self.epoch, shuffle, fix_batches_to_gpus=fix_batches_to_gpus)
        if hasattr(self.dataset, 'set_epoch'):
            self.dataset.set_epoch(self.epoch)
        return self._cur_epoch_itr

    def end_of_epoch(self) -> bool:
        if self._cur_epoch_itr.has_next():
            return False
        # exhausted by a prefetcher does not mean trained: batches may still sit in its queue
        return self.position_hook is None or self.position_hook(self._cur_epoch_itr) >= self._cur_epoch_itr.count

    @property
    def iterations_in_epoch(self):
        if self._cur_epoch_itr is not None:
            if self.position_hook is not None:
                return self.position_hook(self._cur_epoch_itr)
            return self._cur_epoch_itr.count
        if self._next_epoch_itr is not None:
            return self._next_epoch_itr.count
        return 0

    def state_dict(self):
        return {'epoch': self.epoch, 'iterations_in_epoch': self.iterations_in_epoch}

    def load_state_dict(self, state_dict):
        self.epoch = state_dict['epoch']
        itr_pos = state_dict.get('iterations_in_epoch', 0)
        if itr_pos > 0:
            # fast-forward inside the epoch; a finished epoch yields None and the next
            # next_epoch_itr() call simply starts epoch+1
            self._next_epoch_itr = self._get_iterator_for_epoch(
                self.epoch, shuffle=state_dict.get('shuffle', True), offset=itr_pos)

    def epoch_batch_ids(self, epoch, shuffle=True, fix_batches_to_gpus=False) -> np.ndarray:
        """Batch ids this shard visits in ``epoch`` (-1 = padding batch)."""
        n = len(self.frozen_batches)
        order = None
        if shuffle and not fix_batches_to_gpus:
            order = np.arange(n, dtype=np.int64)
            with data_utils.numpy_seed(self.seed + epoch):
                np.random.shuffle(order)
        ids = data_utils.shard_plan(order, n, self.num_shards, self.shard_id)
        if shuffle and fix_batches_to_gpus:
            with data_utils.numpy_seed(self.seed + epoch + self.shard_id):
                np.random.shuffle(ids)
        return ids

    def _get_iterator_for_epoch(self, epoch, shuffle, fix_batches_to_gpus=False, offset=0):
        ids = self.epoch_batch_ids(epoch, shuffle, fix_batches_to_gpus)
        if offset > 0 and offset >= len(ids):
            return None
        empty = np.empty(0, dtype=np.int64)
        batches = [self.frozen_batches[int(b)] if b >= 0 else empty for b in ids[offset:]]
        return CountingIterator(_BatchLoader(self.dataset, self.collate_fn, batches, self.num_workers),
                                start=offset)


class GroupedIterator(object):
    """One optimizer update at a time: lists of up to ``chunk_size`` consecutive micro-batches (the last list of
    an epoch may be shorter). ``offset`` is the number of updates already consumed when the wrapped iterator was
    fast-forwarded on resume (data/iterators.py:214-241 of the reference)."""

    def __init__(self, iterable, chunk_size):
        self.chunk_size = max(1, int(chunk_size))
        self.itr = iter(iterable)
        self._len = -(-len(iterable) // self.chunk_size)
        self.offset = -(-int(getattr(iterable, 'count', 0)) // self.chunk_size)

    def __len__(self):
        return self._len

    def __iter__(self):
        return self

    def __next__(self):
        chunk = list(itertools.islice(self.itr, self.chunk_size))
        if not chunk:
            raise StopIteration
        return chunk


class ShardedIterator(object):
    """Shard ``shard_id`` of ``num_shards``: items shard_id, shard_id + num_shards, ... of a sized iterable,
    followed by ``fill_value`` until every shard has ceil(N / num_shards) items, so all ranks take the same
    number of steps (data/iterators.py:244-275 of the reference). Positions come from the same index plan as
    ``EpochBatchIterator`` (``data_utils.shard_plan``): -1 marks a padding slot."""

    def __init__(self, iterable, num_shards, shard_id, fill_value=None):
        if not 0 <= shard_id < num_shards:
            raise ValueError('shard_id must be between 0 and num_shards')
        n = len(iterable)
        self._plan = data_utils.shard_plan(None, n, num_shards, shard_id)
        self._fill = fill_value
        self._src = iter(iterable)
        self._cursor = 0      # items of the source consumed so far
        self._i = 0           # position in the plan

    def __len__(self):
        return len(self._plan)

    def __iter__(self):
        return self

    def __next__(self):
        if self._i >= len(self._plan):
            raise StopIteration
        want = int(self._plan[self._i])
        self._i += 1
        if want < 0:
            return self._fill
        skip = want - self._cursor
        item = next(itertools.islice(self._src, skip, skip + 1))
        self._cursor = want + 1
        return item


class PrefetchIterator(object):
    """Runs the wrapped iterator in a background thread, ``depth`` items ahead, pinning CPU tensors
    so the trainer's host->device copies are asynchronous (the reference copies synchronously from
    pageable memory, utils.py:32-37). ``len`` / ``count`` / ``offset`` are forwarded so it can sit
    between CountingIterator and GroupedIterator or wrap a GroupedIterator."""

    _END = object()

    def __init__(self, iterable, depth=2, pin_memory: Optional[bool] = None, transform=None):
        self.iterable = iterable
        #: optional callable applied to every item in the worker thread (e.g. Task.host_prepare on each
        #: micro-batch of a GroupedIterator chunk): host-side batch preparation then overlaps with the GPU
        self.transform = transform
        self.pin = torch.cuda.is_available() if pin_memory is None else pin_memory
        self.q: queue.Queue = queue.Queue(maxsize=max(1, depth))
        self._err = None
        # batches handed to the consumer so far (a GroupedIterator chunk counts as len(chunk)); only the consumer
        # thread touches these, so no lock is needed
        self._handed = 0
        self._base = 0
        self.offset = getattr(iterable, 'offset', 0)
        self._thread = threading.Thread(target=self._worker, daemon=True)
        self._started = False

    def __len__(self):
        return len(self.iterable)

    @property
    def count(self):
        return getattr(self.iterable, 'count', 0)

    def _pin(self, x):
        from hetseq_b200.utils.utils import PackedSample
        if isinstance(x, PackedSample):      # one flat buffer: pin that, keep the views consistent
            return x if x.flat.is_pinned() else PackedSample(x.flat.pin_memory(), x.shapes)
        if torch.is_tensor(x):
            return x.pin_memory() if (x.device.type == 'cpu' and not x.is_pinned()) else x
        if isinstance(x, dict):
            return {k: self._pin(v) for k, v in x.items()}
        if isinstance(x, (list, tuple)):
            return type(x)(self._pin(v) for v in x)
        return x

    @staticmethod
    def _size(item) -> int:
        return len(item) if isinstance(item, (list, tuple)) else 1

    def consumed_count(self, counter=None) -> int:
        """Position of the CONSUMER in the epoch: where the underlying CountingIterator stood when this prefetcher
        was attached plus the batches handed out since. The raw count of that iterator also includes batches
        still waiting in the queue, which a mid-epoch checkpoint must not skip on resume."""
        return self._base + self._handed

    def attach_to(self, epoch_itr) -> 'PrefetchIterator':
        """Make ``epoch_itr.state_dict()`` / ``end_of_epoch()`` see the trainer's position, not the reader's.
        Call before the first ``next()`` (the worker thread has not pulled anything yet)."""
        assert not self._started, 'attach_to() must be called before iteration starts'
        self._base = getattr(epoch_itr._cur_epoch_itr, 'count', 0)     # the resume offset, or 0
        epoch_itr.position_hook = self.consumed_count
        return self

    def _worker(self):
        try:
            for item in self.iterable:
                if self.transform is not None:
                    item = self.transform(item)
                self.q.put(self._pin(item) if self.pin else item)
        except BaseException as e:  # surfaced in the consumer thread
            self._err = e
        finally:
            self.q.put(self._END)

    def __iter__(self):
        return self

    def __next__(self):
        if not self._started:
            self._started = True
            self._thread.start()
        item = self.q.get()
        if item is self._END:
            if self._err is not None:
                raise self._err
            raise StopIteration
        self._handed += self._size(item)
        return item
