"""BERT tasks (reference: tasks.py:186-285).

``LanguageModelingTask`` (``--task bert``): BertForPreTraining on NVIDIA-format shards. Every file
under ``--data`` whose name contains the split string is one shard (tasks.py:238-246);
``--num_file`` truncates the list; ``--synthetic-data`` swaps in HDF5-shaped random batches
(phase-1: 128 tokens / 20 predictions, phase-2: 512 / 80).

``BertFineTuningTask`` (``--task bert_token_classification``): the reference's stub
(tasks.py:261-285) references undefined names and cannot run; this is a working equivalent that
trains ``BertForTokenClassification`` on ``.npz`` shards with ``input_ids, segment_ids,
input_mask, labels``.
"""
from __future__ import annotations

import os

import numpy as np
import torch

from hetseq_b200.data.bert_dataset import BertH5pyData, ConBertH5pyData, SyntheticBertData
from hetseq_b200.registry import register_task
from hetseq_b200.tasks.base import Task


@register_task('bert')
class LanguageModelingTask(Task):
    def __init__(self, args, dictionary=None):
        super().__init__(args)
        self.dictionary = dictionary
        #: most labelled positions per sequence seen / announced by the shards: capacity of the MLM row index
        self._max_pred_seen = 0
        #: rows every device batch is padded to (set by the Controller under --cuda-graph: static shapes)
        self.static_batch_size = 0

    @staticmethod
    def add_args(parser):
        parser.add_argument('--data', type=str, help='path including data')
        parser.add_argument('--dict', type=str, metavar='PATH of a file', help='PATH to dictionary')
        parser.add_argument('--config_file', '--config-file', type=str, metavar='PATH of a file',
                            help='PATH to bert model configuration', required=True)
        parser.add_argument('--max_pred_length', '--max-pred-length', type=int, default=512,
                            help='max number of tokens in a sentence')
        parser.add_argument('--num_file', '--num-file', type=int, default=0,
                            help='number of file to run, 0 for all')
        parser.add_argument('--synthetic-data', action='store_true',
                            help='train on HDF5-shaped random batches instead of --data')
        parser.add_argument('--synthetic-seq-len', type=int, default=128)
        parser.add_argument('--synthetic-max-predictions', type=int, default=20)
        parser.add_argument('--synthetic-num-samples', type=int, default=8192)
        parser.add_argument('--checkpoint-activations', action='store_true',
                            help='recompute encoder activations in backward in chunks of ceil(sqrt(L)) layers '
                                 '(the reference has the code but never enables it: bert_modeling.py:470-476)')
        parser.add_argument('--init-linear-activation', default='normal', choices=['normal', 'reference'],
                            help="initial weights of the bias+activation Linear layers (FFN1, pooler, MLM transform): "
                                 "'normal' = N(0, initializer_range) like every other Linear; 'reference' = left at "
                                 "the constructor's kaiming-uniform, which is what the reference's init_bert_weights "
                                 "does (bert_modeling.py:599-610) -- use it to compare from-scratch loss curves")
        parser.add_argument('--mlm-all-positions', action='store_true',
                            help='run the MLM decoder + loss over all B*S positions like the reference '
                                 '(bert_modeling.py:878-882) instead of the labelled rows only')

    @classmethod
    def setup_task(cls, args, **kwargs):
        dictionary = cls.load_dictionary(args.dict) if getattr(args, 'dict', None) else None
        return cls(args, dictionary)

    def build_model(self, args):
        from hetseq_b200.models.bert import BertConfig, BertForPreTraining
        config = BertConfig.from_json_file(args.config_file)
        config.linear_activation_init = getattr(args, 'init_linear_activation', 'normal')
        return BertForPreTraining(config, args=args)

    def load_dataset(self, split, **kwargs):
        args = self.args
        if getattr(args, 'synthetic_data', False):
            from hetseq_b200.models.bert import BertConfig
            vocab = BertConfig.from_json_file(args.config_file).vocab_size
            n = args.synthetic_num_samples if split == args.train_subset else max(64, args.synthetic_num_samples // 16)
            ds = SyntheticBertData(n, args.synthetic_seq_len, args.synthetic_max_predictions, vocab,
                                   seed=args.seed + (0 if split == args.train_subset else 1),
                                   max_pred_length=args.max_pred_length)
            self.datasets[split] = ConBertH5pyData([ds])
            self._max_pred_seen = max(self._max_pred_seen, self.datasets[split].max_predictions)
            print('| synthetic {} sentences for split {}'.format(len(ds), split), flush=True)
            return
        path = args.data
        if path is None or not os.path.exists(path):
            raise FileNotFoundError('Dataset not found: ({})'.format(path))
        files = [os.path.join(path, f) for f in os.listdir(path)] if os.path.isdir(path) else [path]
        files = sorted([f for f in files if split in f])
        if args.num_file > 0:
            files = files[0:args.num_file]
        assert len(files) > 0, 'no suitable file in split ***{}***'.format(split)
        datasets = [BertH5pyData(f, max_pred_length=args.max_pred_length) for f in files]
        dataset = ConBertH5pyData(datasets)
        print('| loaded {} sentences from: {}'.format(len(dataset), path), flush=True)
        self.datasets[split] = dataset
        # every rank derives the same MLM capacity from the shards themselves, before the first batch
        self._max_pred_seen = max(self._max_pred_seen, dataset.max_predictions)
        print('| loading finished')

    def sample_stats(self, sample):
        # real tokens = attended positions (input_mask is sample[2]); computed on the host copy. Rows added by
        # pad_batch attend to nothing and are not counted as sentences.
        input_ids, mask = sample[0], sample[2]
        if mask.device.type != 'cpu':
            return int(input_ids.shape[0]), int(input_ids.numel())
        return int((mask.sum(dim=1) > 0).sum()), int(mask.sum())

    def pad_batch(self, sample, rows: int):
        """Grow a collated host batch to ``rows`` sequences with inert ones: no attended token, every MLM label
        and the NSP label at the ignore index (-1), so they add nothing to either loss (both are means over the
        non-ignored rows, bert_modeling.py:881-886) or to any gradient. Keeps the shapes of a short last batch
        equal to those of a full one (static shapes for the captured CUDA graph)."""
        B = int(sample[0].shape[0])
        if rows <= B or len(sample) != 5:
            return sample
        pad = rows - B
        out = []
        for k, t in enumerate(sample):
            fill = -1 if k >= 3 else 0
            out.append(torch.cat([t, torch.full((pad,) + tuple(t.shape[1:]), fill, dtype=t.dtype)], dim=0))
        return out

    def mlm_index(self, masked_lm_labels: torch.Tensor):
        """Host-side (numpy) flat indices + labels of the labelled positions, padded with -1 to a fixed
        capacity of ``batch x max predictions per sequence`` (rounded up to 64). Replaces the on-device
        ``nonzero`` (a sync) and makes the MLM head's shapes static."""
        lab = masked_lm_labels.reshape(-1).numpy()
        rows = np.nonzero(lab != -1)[0]
        B = masked_lm_labels.shape[0]
        per_seq = int((masked_lm_labels != -1).sum(1).max()) if rows.size else 1
        # normally fixed up front by load_dataset (the shards announce their width); only a hand-made batch with
        # more labels than that can still grow it
        self._max_pred_seen = max(self._max_pred_seen, per_seq,
                                  int(getattr(self.args, 'synthetic_max_predictions', 0) or 0)
                                  if getattr(self.args, 'synthetic_data', False) else 0)
        cap = (max(B * self._max_pred_seen, rows.size, 1) + 63) // 64 * 64
        idx = np.full(cap, -1, dtype=np.int32)
        idx[:rows.size] = rows
        row_labels = np.full(cap, -1, dtype=np.int32)
        row_labels[:rows.size] = lab[rows]
        return torch.from_numpy(idx), torch.from_numpy(row_labels)

    @staticmethod
    def embedding_plan(input_ids: torch.Tensor):
        """Host-side plan of the deterministic word-embedding gradient scatter: positions sorted by token id
        (stable) and, for the first slot of each run of equal ids, one past the run's end (0 elsewhere). Fixed
        shape [B*S] each; costs a numpy argsort in the prefetch thread instead of a device sort per step."""
        flat = input_ids.reshape(-1).numpy()
        order = np.argsort(flat, kind='stable')
        sorted_ids = flat[order]
        n = flat.size
        start = np.ones(n, dtype=bool)
        start[1:] = sorted_ids[1:] != sorted_ids[:-1]
        starts = np.nonzero(start)[0]
        run_end = np.zeros(n, dtype=np.int32)
        run_end[starts] = np.append(starts[1:], n)
        return torch.from_numpy(order.astype(np.int32)), torch.from_numpy(run_end)

    def prepare_sample(self, sample, device):
        """int64 on disk, int32 on the wire: token ids and labels fit in 32 bits and the native
        kernels consume int32 directly; PyTorch-path modules widen as needed. Appends
        the host-computed MLM row index (see ``mlm_index``) unless ``--mlm-all-positions``."""
        from hetseq_b200.utils import utils
        cuda = torch.device(device).type == 'cuda'
        sample = self._host_part(sample, narrow=cuda)
        if cuda and not isinstance(sample, utils.PackedSample):
            sample = utils.PackedSample.pack(sample, pin=True) or sample     # one H2D copy instead of seven
        return utils.move_to_device(sample, device)

    def _host_part(self, sample, narrow):
        from hetseq_b200.utils import utils
        if isinstance(sample, utils.PackedSample):
            return sample                    # already prepared by host_prepare()
        sample = list(sample)
        if self.static_batch_size and len(sample) == 5 and sample[0].device.type == 'cpu':
            sample = self.pad_batch(sample, self.static_batch_size)
        if len(sample) == 5 and sample[3].device.type == 'cpu' and not getattr(self.args, 'mlm_all_positions', False):
            sample += list(self.mlm_index(sample[3]))
            sample += list(self.embedding_plan(sample[0]))
        if narrow:
            sample = [t.to(torch.int32) if t.dtype == torch.int64 else t for t in sample]
        return sample

    def host_prepare(self, sample):
        """Everything of ``prepare_sample`` that does not need the device: the MLM row index, int64 -> int32 and
        pinning, so that only the (asynchronous) copies are left for the training thread."""
        if sample is None or len(sample) == 0:
            return sample
        from hetseq_b200.utils import utils
        cuda = torch.cuda.is_available() and not getattr(self.args, 'cpu', False)
        sample = self._host_part(sample, narrow=cuda)
        if cuda and not isinstance(sample, utils.PackedSample):
            packed = utils.PackedSample.pack(sample, pin=True)     # all int32 after narrowing: one pinned buffer
            sample = packed if packed is not None else \
                [t.pin_memory() if (t.device.type == 'cpu' and not t.is_pinned()) else t for t in sample]
        return sample


@register_task('bert_token_classification')
class BertFineTuningTask(Task):
    def __init__(self, args, dictionary=None):
        super().__init__(args)
        self.dictionary = dictionary

    @staticmethod
    def add_args(parser):
        parser.add_argument('--data', type=str, help='directory of .npz shards')
        parser.add_argument('--dict', type=str, help='PATH to dictionary')
        parser.add_argument('--config_file', '--config-file', type=str, required=True)
        parser.add_argument('--num_label', '--num-label', type=int, default=3)

    @classmethod
    def setup_task(cls, args, **kwargs):
        dictionary = cls.load_dictionary(args.dict) if getattr(args, 'dict', None) else None
        return cls(args, dictionary)

    def build_model(self, args):
        from hetseq_b200.models.bert import BertConfig, BertForTokenClassification
        config = BertConfig.from_json_file(args.config_file)
        return _TokenClassificationLoss(BertForTokenClassification(config, args.num_label))

    def load_dataset(self, split, **kwargs):
        path = self.args.data
        files = [os.path.join(path, f) for f in os.listdir(path)] if os.path.isdir(path) else [path]
        files = sorted(f for f in files if split in f and f.endswith('.npz'))
        assert len(files) > 0, 'no suitable file in split ***{}***'.format(split)
        self.datasets[split] = _NpzTokenDataset(files)


class _TokenClassificationLoss(torch.nn.Module):
    """Adapts the head to the framework's ``model(*sample) -> loss`` contract."""

    def __init__(self, head):
        super().__init__()
        self.head = head

    def forward(self, input_ids, segment_ids, input_mask, labels):
        return self.head(input_ids.long(), segment_ids.long(), input_mask.long(), labels.long())


class _NpzTokenDataset(torch.utils.data.Dataset):
    def __init__(self, files):
        arrs = [np.load(f) for f in files]
        self.a = {k: np.concatenate([z[k] for z in arrs]) for k in ('input_ids', 'segment_ids', 'input_mask', 'labels')}

    def __len__(self):
        return len(self.a['input_ids'])

    def __getitem__(self, i):
        return [torch.from_numpy(self.a[k][i].astype(np.int64)) for k in ('input_ids', 'segment_ids', 'input_mask', 'labels')]

    def ordered_indices(self):
        return np.arange(len(self))

    def num_tokens(self, i):
        return self.a['input_ids'].shape[1]

    def sizes_array(self):
        return np.full(len(self), self.a['input_ids'].shape[1], dtype=np.int64)

    def collater(self, samples):
        if len(samples) == 0:
            return None
        return [torch.stack([s[k] for s in samples]) for k in range(4)]

    def set_epoch(self, epoch):
        pass
