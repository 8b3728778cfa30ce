"""BERT pre-training datasets in the NVIDIA HDF5 shard layout (reference: data/h5pyDataset.py).

Shard keys: ``input_ids, input_mask, segment_ids, masked_lm_positions, masked_lm_ids,
next_sentence_labels``. A sample is the list ``[input_ids, segment_ids, input_mask,
masked_lm_labels, next_sentence_labels]`` -- the positional order of
``BertForPreTraining.forward`` (bert_modeling.py:875) -- where ``masked_lm_labels`` is -1
everywhere except ``labels[positions[:k]] = ids[:k]``, ``k`` = index of the first zero in
``masked_lm_positions`` (data/h5pyDataset.py:42-48).

Differences from the reference: a shard is opened once per process and read in
whole arrays, not re-opened for every sample; ``size()`` reports the true sequence length of the
shard instead of a constant 512; ``--max_pred_length`` is actually forwarded; ``.npz`` shards with
the same keys work without h5py; ``SyntheticBertData`` generates HDF5-shaped samples for
benchmarks (there is no corpus on the benchmark box).
"""
from __future__ import annotations

import bisect
import os
from typing import List, Optional, Sequence

import numpy as np
import torch
import torch.utils.data
from torch.utils.data.dataloader import default_collate

BERT_KEYS = ('input_ids', 'input_mask', 'segment_ids', 'masked_lm_positions', 'masked_lm_ids',
             'next_sentence_labels')


def _labels_from_positions(seq_len: int, positions: np.ndarray, ids: np.ndarray, max_pred_length: int) -> np.ndarray:
    labels = np.full(seq_len, -1, dtype=np.int64)
    k = min(max_pred_length, len(positions))
    zeros = np.nonzero(positions == 0)[0]
    if len(zeros) != 0:
        k = min(k, int(zeros[0]))
    labels[positions[:k]] = ids[:k]
    return labels


def batch_labels(seq_len: int, positions: np.ndarray, ids: np.ndarray, max_pred_length: int) -> np.ndarray:
    """Vectorised ``_labels_from_positions`` for a [B, P] batch."""
    B, P = positions.shape
    P = min(P, max_pred_length)
    positions, ids = positions[:, :P], ids[:, :P]
    is_zero = positions == 0
    first_zero = np.where(is_zero.any(axis=1), is_zero.argmax(axis=1), P)
    valid = np.arange(P)[None, :] < first_zero[:, None]
    labels = np.full((B, seq_len), -1, dtype=np.int64)
    rows = np.broadcast_to(np.arange(B)[:, None], (B, P))
    labels[rows[valid], positions[valid]] = ids[valid]
    return labels


class BertH5pyData(torch.utils.data.Dataset):
    """One shard (``.hdf5``/``.h5`` through h5py, or ``.npz``)."""

    def __init__(self, path, max_pred_length=512):
        super().__init__()
        self.keys = BERT_KEYS
        self.max_pred_length = max_pred_length
        self.path = path
        self._arrays = None          # loaded lazily, once per process (fork-safe for workers)
        self._len, self._seq_len, self._pred_width = self._probe()

    def _open(self):
        if self.path.endswith('.npz'):
            z = np.load(self.path)
            return {k: z[k] for k in self.keys}
        try:
            import h5py
        except ImportError as e:  # pragma: no cover - depends on the environment
            raise ImportError('reading {} needs h5py (not installed); convert the shard to .npz with the same '
                              'keys or use --synthetic-data'.format(self.path)) from e
        with h5py.File(self.path, 'r', libver='latest', swmr=True) as f:
            return {k: np.asarray(f[k]) for k in self.keys}

    def _probe(self):
        if self.path.endswith('.npz'):
            z = np.load(self.path)
            arr = z['input_ids']
            return int(arr.shape[0]), int(arr.shape[1]), int(z['masked_lm_positions'].shape[1])
        import h5py
        with h5py.File(self.path, 'r', libver='latest', swmr=True) as f:
            shp = f[self.keys[0]].shape
            return int(shp[0]), int(shp[1]), int(f['masked_lm_positions'].shape[1])

    @property
    def max_predictions(self) -> int:
        """Most labelled positions a sample of this shard can carry (width of ``masked_lm_positions``, capped by
        ``--max_pred_length``): fixes the capacity of the MLM row index before the first batch is seen."""
        return min(self._pred_width, self.max_pred_length)

    @property
    def arrays(self):
        if self._arrays is None:
            self._arrays = self._open()
        return self._arrays

    def check_index(self, i):
        if i < 0 or i >= self._len:
            raise IndexError('index out of range')

    def __getitem__(self, index):
        self.check_index(index)
        a = self.arrays
        input_ids = a['input_ids'][index].astype(np.int64)
        labels = _labels_from_positions(len(input_ids), a['masked_lm_positions'][index].astype(np.int64),
                                        a['masked_lm_ids'][index].astype(np.int64), self.max_pred_length)
        return [torch.from_numpy(input_ids),
                torch.from_numpy(a['segment_ids'][index].astype(np.int64)),
                torch.from_numpy(a['input_mask'][index].astype(np.int64)),
                torch.from_numpy(labels),
                torch.from_numpy(np.asarray(a['next_sentence_labels'][index]).astype(np.int64))]

    def fetch_rows(self, rows: np.ndarray):
        """Vectorised read of several samples -> the five [B, ...] int64 tensors."""
        a = self.arrays
        rows = np.asarray(rows, dtype=np.int64)
        ids = a['input_ids'][rows].astype(np.int64)
        labels = batch_labels(ids.shape[1], a['masked_lm_positions'][rows].astype(np.int64),
                              a['masked_lm_ids'][rows].astype(np.int64), self.max_pred_length)
        return [torch.from_numpy(ids), torch.from_numpy(a['segment_ids'][rows].astype(np.int64)),
                torch.from_numpy(a['input_mask'][rows].astype(np.int64)), torch.from_numpy(labels),
                torch.from_numpy(np.asarray(a['next_sentence_labels'][rows]).astype(np.int64).reshape(len(rows)))]

    def __len__(self):
        return self._len

    def size(self, idx: int):
        return self._seq_len

    def set_epoch(self, epoch):
        pass


class SyntheticBertData(torch.utils.data.Dataset):
    """HDF5-shaped random samples, deterministic per index (phase-1: seq 128 / 20 predictions,
    phase-2: seq 512 / 80 predictions -- the NVIDIA convention; the reference does not state it)."""

    def __init__(self, num_samples=4096, seq_len=128, max_predictions=20, vocab_size=30522, seed=0,
                 max_pred_length=512):
        super().__init__()
        self.num_samples, self.seq_len, self.max_predictions = num_samples, seq_len, max_predictions
        self.vocab_size, self.seed, self.max_pred_length = vocab_size, seed, max_pred_length

    def __len__(self):
        return self.num_samples

    def _raw(self, rows: np.ndarray):
        B, S, P = len(rows), self.seq_len, self.max_predictions
        ids = np.empty((B, S), dtype=np.int64)
        mask = np.zeros((B, S), dtype=np.int64)
        seg = np.zeros((B, S), dtype=np.int64)
        pos = np.zeros((B, P), dtype=np.int64)
        mids = np.zeros((B, P), dtype=np.int64)
        nsp = np.empty(B, dtype=np.int64)
        for j, r in enumerate(rows):
            g = np.random.default_rng((self.seed, int(r)))
            length = int(g.integers(S // 2, S + 1))
            lo = min(1000, max(1, self.vocab_size // 10))   # skip the special-token range of real vocabularies
            ids[j] = g.integers(lo, self.vocab_size, size=S)
            ids[j, length:] = 0
            mask[j, :length] = 1
            split = int(g.integers(1, max(2, length - 1)))
            seg[j, split:length] = 1
            k = int(min(P, max(1, round(0.15 * length))))
            p = np.sort(g.choice(np.arange(1, length), size=k, replace=False))
            pos[j, :k] = p
            mids[j, :k] = g.integers(lo, self.vocab_size, size=k)
            nsp[j] = int(g.integers(0, 2))
        return ids, seg, mask, pos, mids, nsp

    def fetch_rows(self, rows):
        ids, seg, mask, pos, mids, nsp = self._raw(np.asarray(rows, dtype=np.int64))
        labels = batch_labels(self.seq_len, pos, mids, self.max_pred_length)
        return [torch.from_numpy(x) for x in (ids, seg, mask, labels, nsp)]

    def __getitem__(self, index):
        if index < 0 or index >= self.num_samples:
            raise IndexError('index out of range')
        return [t[0] for t in self.fetch_rows([index])]

    def size(self, idx: int):
        return self.seq_len

    @property
    def max_predictions_per_sample(self) -> int:
        return min(self.max_predictions, self.max_pred_length)

    def set_epoch(self, epoch):
        pass

    def write_npz(self, path, rows: Optional[Sequence[int]] = None):
        """Dump (a slice of) the synthetic corpus as an .npz shard with the HDF5 key layout."""
        rows = np.arange(self.num_samples) if rows is None else np.asarray(rows)
        ids, seg, mask, pos, mids, nsp = self._raw(rows)
        np.savez(path, input_ids=ids.astype(np.int32), input_mask=mask.astype(np.int32),
                 segment_ids=seg.astype(np.int32), masked_lm_positions=pos.astype(np.int32),
                 masked_lm_ids=mids.astype(np.int32), next_sentence_labels=nsp.astype(np.int32))


class ConBertH5pyData(torch.utils.data.Dataset):
    """Concatenation of shards with optional per-shard up/down-sampling (data/h5pyDataset.py:72-136)."""

    @staticmethod
    def cumsum(sequence, sample_ratios):
        r, s = [], 0
        for e, ratio in zip(sequence, sample_ratios):
            curr_len = int(ratio * len(e))
            r.append(curr_len + s)
            s += curr_len
        return r

    def __init__(self, datasets, sample_ratios=1):
        super().__init__()
        assert len(datasets) > 0, 'datasets should not be an empty iterable'
        self.datasets = list(datasets)
        if isinstance(sample_ratios, (int, float)):
            sample_ratios = [sample_ratios] * len(self.datasets)
        self.sample_ratios = sample_ratios
        self.cumulative_sizes = self.cumsum(self.datasets, sample_ratios)
        self.real_sizes = [len(d) for d in self.datasets]

    def __len__(self):
        return self.cumulative_sizes[-1]

    @property
    def max_predictions(self) -> int:
        """Largest number of labelled positions any sample can carry (0 when the shards do not say)."""
        out = 0
        for d in self.datasets:
            out = max(out, int(getattr(d, 'max_predictions_per_sample', getattr(d, 'max_predictions', 0)) or 0))
        return out

    def _get_dataset_and_sample_index(self, idx: int):
        dataset_idx = bisect.bisect_right(self.cumulative_sizes, idx)
        sample_idx = idx if dataset_idx == 0 else idx - self.cumulative_sizes[dataset_idx - 1]
        return dataset_idx, sample_idx % self.real_sizes[dataset_idx]

    def __getitem__(self, idx):
        d, s = self._get_dataset_and_sample_index(int(idx))
        return self.datasets[d][s]

    def fetch_batch(self, indices):
        """Vectorised batch assembly: group the requested rows by shard, read each group at once."""
        if len(indices) == 0:
            return None
        idx = np.asarray(indices, dtype=np.int64)
        cum = np.asarray(self.cumulative_sizes, dtype=np.int64)
        d = np.searchsorted(cum, idx, side='right')
        starts = np.concatenate([[0], cum[:-1]])
        s = (idx - starts[d]) % np.asarray(self.real_sizes, dtype=np.int64)[d]
        if not all(hasattr(ds, 'fetch_rows') for ds in self.datasets):
            return self.collater([self[int(i)] for i in idx])
        parts, order = [], []
        for di in np.unique(d):
            sel = np.nonzero(d == di)[0]
            parts.append(self.datasets[int(di)].fetch_rows(s[sel]))
            order.append(sel)
        if len(parts) == 1 and np.array_equal(order[0], np.arange(len(idx))):
            return parts[0]
        order = np.concatenate(order)
        inv = torch.from_numpy(np.argsort(order))
        return [torch.cat([p[k] for p in parts], dim=0)[inv] for k in range(len(parts[0]))]

    def collater(self, samples):
        if len(samples) == 0:
            return None
        if hasattr(self.datasets[0], 'collater'):
            return self.datasets[0].collater(samples)
        return default_collate(samples)

    def ordered_indices(self):
        return np.arange(len(self))

    def num_tokens(self, index: int):
        return np.max(self.size(index))

    def size(self, idx: int):
        d, s = self._get_dataset_and_sample_index(int(idx))
        return self.datasets[d].size(s)

    def sizes_array(self) -> np.ndarray:
        """Token count of every item, without a Python call per item."""
        out = np.empty(len(self), dtype=np.int64)
        prev = 0
        for ds, end in zip(self.datasets, self.cumulative_sizes):
            out[prev:end] = ds.size(0) if len(ds) else 0
            prev = end
        return out

    def set_epoch(self, epoch):
        for d in self.datasets:
            if hasattr(d, 'set_epoch'):
                d.set_epoch(epoch)
