"""Batch planning helpers (reference: data/data_utils.py:14-61 + data/data_utils_fast.pyx).

``batch_by_size`` keeps the reference's signature (``num_tokens_fn`` callback) but also accepts a
precomputed ``sizes`` array, which is what the built-in datasets pass; the work happens in the
native planner (hetseq_b200/native/batch_planner.cpp) with no per-sample Python call. A pure-numpy
implementation of the same rule is the fallback when the native library has not been built and the
oracle in tests.
"""
from __future__ import annotations

import contextlib
import ctypes
import sys
from typing import Callable, List, Optional, Sequence

import numpy as np


@contextlib.contextmanager
def numpy_seed(seed, *addl_seeds):
    """Seed the NumPy global PRNG inside the block and restore the previous state afterwards."""
    if seed is None:
        yield
        return
    if len(addl_seeds) > 0:
        seed = int(hash((seed, *addl_seeds)) % 1e6)
    state = np.random.get_state()
    np.random.seed(seed)
    try:
        yield
    finally:
        np.random.set_state(state)


class BatchPlan(Sequence):
    """Batches as contiguous slices of an ordered index array: batch b = indices[offsets[b]:offsets[b+1]].

    Behaves like the reference's list of lists (len, indexing, iteration) without materialising
    millions of small Python lists.
    """

    def __init__(self, indices: np.ndarray, offsets: np.ndarray):
        self.indices = np.ascontiguousarray(indices, dtype=np.int64)
        self.offsets = np.ascontiguousarray(offsets, dtype=np.int64)

    def __len__(self):
        return len(self.offsets) - 1

    def __getitem__(self, b):
        if isinstance(b, slice):
            return [self[i] for i in range(*b.indices(len(self)))]
        if b < 0:
            b += len(self)
        if b < 0 or b >= len(self):
            raise IndexError(b)
        return self.indices[self.offsets[b]:self.offsets[b + 1]]

    def tolist(self) -> List[List[int]]:
        return [self[b].tolist() for b in range(len(self))]


_native = None


def _native_lib():
    global _native
    if _native is None:
        from hetseq_b200.ops.build import NATIVE_LIB, build_native
        try:
            path = NATIVE_LIB if NATIVE_LIB.exists() else build_native(verbose=False)
            lib = ctypes.CDLL(str(path))
            i64p = ctypes.POINTER(ctypes.c_int64)
            lib.hs_batch_by_size.restype = ctypes.c_int64
            lib.hs_batch_by_size.argtypes = [i64p, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64,
                                             ctypes.c_int64, i64p, i64p]
            lib.hs_shard_plan.restype = ctypes.c_int64
            lib.hs_shard_plan.argtypes = [i64p, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, i64p]
            _native = lib
        except Exception:  # pragma: no cover - toolchain missing
            _native = False
    return _native or None


def _as_i64p(a: np.ndarray):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_int64))


def batch_offsets_numpy(sizes: np.ndarray, max_tokens: int, max_sentences: int, bsz_mult: int) -> np.ndarray:
    """Pure-Python restatement of the batching rule (data/data_utils_fast.pyx:21-62)."""
    offsets = [0]
    start, sample_len = 0, 0
    n = len(sizes)
    for i in range(n):
        sample_len = max(sample_len, int(sizes[i]))
        assert sample_len <= max_tokens, (
            'sentence at position {} of size {} exceeds max_tokens limit of {}!'.format(i, sample_len, max_tokens))
        length = i - start
        if length > 0 and (length == max_sentences or (length + 1) * sample_len > max_tokens):
            mod_len = max(bsz_mult * (length // bsz_mult), length % bsz_mult)
            start += mod_len
            offsets.append(start)
            sample_len = int(sizes[start:i + 1].max())
    if n - start > 0:
        offsets.append(n)
    return np.asarray(offsets, dtype=np.int64)


def batch_by_size(indices, num_tokens_fn: Optional[Callable[[int], int]] = None, max_tokens=None,
                  max_sentences=None, required_batch_size_multiple=1, sizes: Optional[np.ndarray] = None,
                  use_native: bool = True) -> BatchPlan:
    """Group ``indices`` (ordered) into batches bounded by ``max_tokens`` (= batch size x longest
    sample) and ``max_sentences``, sizes rounded down to ``required_batch_size_multiple``.

    ``sizes[k]`` is the token count of dataset item ``k`` (indexed by dataset index, not position);
    if omitted it is gathered by calling ``num_tokens_fn`` once per index like the reference.
    """
    indices = np.ascontiguousarray(np.asarray(indices, dtype=np.int64))
    max_tokens = max_tokens if max_tokens is not None else sys.maxsize
    max_sentences = max_sentences if max_sentences is not None else sys.maxsize
    bsz_mult = max(1, int(required_batch_size_multiple))
    if sizes is not None:
        pos_sizes = np.ascontiguousarray(np.asarray(sizes, dtype=np.int64)[indices])
    elif num_tokens_fn is not None:
        pos_sizes = np.fromiter((num_tokens_fn(int(i)) for i in indices), dtype=np.int64, count=len(indices))
    else:
        raise ValueError('batch_by_size needs sizes or num_tokens_fn')
    lib = _native_lib() if use_native else None
    if lib is not None and len(indices) > 0:
        offsets = np.empty(len(indices) + 1, dtype=np.int64)
        bad = ctypes.c_int64(-1)
        nb = lib.hs_batch_by_size(_as_i64p(pos_sizes), len(indices), min(max_tokens, 2 ** 62),
                                  min(max_sentences, 2 ** 62), bsz_mult, _as_i64p(offsets), ctypes.byref(bad))
        if nb < 0:
            pos = bad.value
            raise AssertionError('sentence at index {} of size {} exceeds max_tokens limit of {}!'.format(
                int(indices[pos]), int(pos_sizes[pos]), max_tokens))
        offsets = offsets[:nb + 1].copy()
    else:
        offsets = batch_offsets_numpy(pos_sizes, max_tokens, max_sentences, bsz_mult)
    return BatchPlan(indices, offsets)


def shard_plan(order: Optional[np.ndarray], num_batches: int, num_shards: int, shard_id: int) -> np.ndarray:
    """Batch ids for one shard: round-robin over ``order`` padded with -1 (data/iterators.py:244-275)."""
    if shard_id < 0 or shard_id >= num_shards:
        raise ValueError('shard_id must be between 0 and num_shards')
    per = (num_batches + num_shards - 1) // num_shards
    out = np.full(per, -1, dtype=np.int64)
    lib = _native_lib()
    if lib is not None and per > 0:
        o = None if order is None else np.ascontiguousarray(order, dtype=np.int64)
        lib.hs_shard_plan(None if o is None else _as_i64p(o), num_batches, num_shards, shard_id, _as_i64p(out))
    else:
        src = np.arange(num_batches, dtype=np.int64) if order is None else np.asarray(order, dtype=np.int64)
        mine = src[shard_id::num_shards]
        out[:len(mine)] = mine
    return out
