"""Datasets, batch planning and resumable epoch/shard-aware iterators."""
from hetseq_b200.data.data_utils import numpy_seed, batch_by_size, BatchPlan  # noqa: F401
from hetseq_b200.data.iterators import (  # noqa: F401
    CountingIterator, EpochBatchIterator, GroupedIterator, ShardedIterator, PrefetchIterator)
