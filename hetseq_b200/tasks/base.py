"""Task base class (reference: tasks.py:13-183).

A task loads dataset splits, turns a dataset into an ``EpochBatchIterator``, builds the model and
runs one forward/backward on a micro-batch. The model contract is the reference's:
``model(*sample)`` returns a scalar loss (tasks.py:158, docs/source/extending.rst:52).

To add a task: subclass, decorate with ``@register_task('name')``, implement ``add_args``,
``setup_task``, ``load_dataset`` and ``build_model``. The dataset needs ``__getitem__/__len__/
ordered_indices/num_tokens/collater/set_epoch`` (optionally ``sizes_array`` and ``fetch_batch`` for
the vectorised fast paths).
"""
from __future__ import annotations

import collections

import numpy as np
import torch

from hetseq_b200.data import data_utils, iterators


class Task(object):
    def __init__(self, args):
        self.args = args
        self.datasets = {}
        self.dataset_to_epoch_iter = {}

    @staticmethod
    def add_args(parser):
        """Task-specific flags."""
        pass

    @classmethod
    def setup_task(cls, args, **kwargs):
        return cls(args)

    @staticmethod
    def load_dictionary(vocab_file):
        """vocab.txt -> OrderedDict token -> index (tasks.py:23-36)."""
        vocab = collections.OrderedDict()
        with open(vocab_file, 'r', encoding='utf-8') as reader:
            for index, token in enumerate(reader):
                vocab[token.rstrip('\n').strip()] = index
        print('| loaded dictionary with {} subwords  from: {}'.format(len(vocab), vocab_file))
        return vocab

    def load_dataset(self, split, **kwargs):
        raise NotImplementedError

    def has_dataset(self, split) -> bool:
        return split in self.datasets

    def dataset(self, split):
        if split not in self.datasets:
            raise KeyError('Dataset not loaded: ' + split)
        if not isinstance(self.datasets[split], torch.utils.data.Dataset):
            raise TypeError('Datasets are expected to be of type torch.utils.data.Dataset')
        return self.datasets[split]

    def get_batch_iterator(self, dataset, max_tokens=None, max_sentences=None, max_positions=None,
                           ignore_invalid_inputs=False, required_batch_size_multiple=1, seed=1,
                           num_shards=1, shard_id=0, num_workers=0, epoch=0):
        """Batch ``dataset`` by size and wrap it in a (cached) EpochBatchIterator (tasks.py:59-126)."""
        if dataset in self.dataset_to_epoch_iter:
            return self.dataset_to_epoch_iter[dataset]
        with data_utils.numpy_seed(seed):
            indices = dataset.ordered_indices()
        print('| build batch sampler')
        sizes = dataset.sizes_array() if hasattr(dataset, 'sizes_array') else None
        batch_sampler = data_utils.batch_by_size(
            indices, None if sizes is not None else dataset.num_tokens, max_tokens=max_tokens,
            max_sentences=max_sentences, required_batch_size_multiple=required_batch_size_multiple,
            sizes=sizes)
        print('| finish building batch sampler')
        epoch_iter = iterators.EpochBatchIterator(
            dataset=dataset, collate_fn=dataset.collater, batch_sampler=batch_sampler, seed=seed,
            num_shards=num_shards, shard_id=shard_id, num_workers=num_workers, epoch=epoch)
        self.dataset_to_epoch_iter[dataset] = epoch_iter
        return epoch_iter

    def build_model(self, args):
        raise NotImplementedError

    # -- per-batch hooks --------------------------------------------------------------------------
    def sample_stats(self, sample):
        """(nsentences, ntokens) of a collated *host* sample; called before the device copy so it
        costs no synchronisation. The reference hard-codes ntokens = 0 (tasks.py:172)."""
        first = sample[0]
        return int(first.shape[0]), int(first.numel())

    def sample_size(self, sample, nsentences):
        """Gradient denominator. ``reference`` mode reproduces tasks.py:161-166: the length of the
        first example (sequence length for BERT, 1 for MNIST images), not the batch size."""
        if getattr(self.args, 'sample_size_mode', 'batch') == 'reference':
            if sample is None or len(sample) == 0 or len(sample[0][0]) == 0:
                return 0
            return len(sample[0][0])
        return nsentences

    def host_prepare(self, sample):
        """CPU-side part of ``prepare_sample`` (dtype narrowing, index building, pinning). Safe to run in the
        prefetch thread, one update ahead of the GPU (``data/iterators.py::PrefetchIterator(transform=...)``);
        ``prepare_sample`` must accept its output. Default: nothing to do."""
        return sample

    def prepare_sample(self, sample, device):
        """Host -> device hand-off for one collated micro-batch."""
        from hetseq_b200.utils import utils
        return utils.move_to_device(sample, device)

    def train_step(self, sample, model, optimizer, ignore_grad=False, stats=None, weight=None):
        """forward + backward on one micro-batch (tasks.py:139-177). Returns
        (loss, sample_size, logging_output); ``loss`` stays on the device (no sync). ``weight`` (device
        scalar) replaces the host-side weighting below: a captured CUDA graph reads the batch weight -- the
        number of real sentences, 0 for a padding batch -- from device memory on every replay."""
        model.train()
        loss = model(*sample)
        if ignore_grad:
            loss = loss * 0
        nsentences, ntokens = stats if stats is not None else self.sample_stats(sample)
        sample_size = self.sample_size(sample, nsentences)
        logging_output = {
            'nsentences': nsentences,
            'loss': loss.detach(),
            'nll_loss': loss.detach(),
            'ntokens': ntokens,
            'sample_size': sample_size,
        }
        # 'batch' mode: the model returns a batch mean; weight it by the batch size so that dividing
        # the accumulated gradient by the global sample_size yields the true mean over sentences.
        # 'reference' mode: back-propagate the mean itself, as tasks.py:176 does.
        if weight is not None:
            optimizer.backward(loss * weight)
        elif getattr(self.args, 'sample_size_mode', 'batch') == 'batch' and nsentences != 1:
            optimizer.backward(loss * float(nsentences))
        else:
            optimizer.backward(loss)
        return loss, sample_size, logging_output

    def valid_step(self, sample, model):
        model.eval()
        with torch.no_grad():
            loss = model(*sample)
        nsentences, ntokens = self.sample_stats(sample)
        return loss, self.sample_size(sample, nsentences), {
            'nsentences': nsentences, 'loss': loss.detach(), 'nll_loss': loss.detach(), 'ntokens': ntokens}

    def update_step(self, num_updates):
        """Hook called after every optimizer update."""
        pass
