"""CPU tests added in round 2: the HDF5 reader (through a stub ``h5py`` backed by an .npz fixture), activation
checkpointing, static-shape batch padding, the launch plan and the declarative log line."""
import argparse
import json
import sys
import types

import numpy as np
import pytest
import torch

from hetseq_b200.data.bert_dataset import BertH5pyData, ConBertH5pyData, SyntheticBertData


# ---------------------------------------------------------------------------------------------------------
class _StubH5File(object):
    """The slice of h5py.File the reader uses: context manager + ``f[key]`` -> array with ``.shape``."""

    opened = []

    def __init__(self, path, mode='r', **kwargs):
        assert mode == 'r' and kwargs.get('swmr') is True and kwargs.get('libver') == 'latest', \
            'shards are opened read-only in SWMR mode like the reference (data/h5pyDataset.py:33)'
        self._z = np.load(path + '.npz')
        _StubH5File.opened.append(path)

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False

    def __getitem__(self, key):
        return self._z[key]


def test_hdf5_reader_through_stub_h5py(tmp_path, monkeypatch):
    """The h5py branch of BertH5pyData (reference: data/h5pyDataset.py:31-51) against the .npz branch on the same
    shard: identical samples and labels, the file is opened once for the probe and once for the data (the
    reference re-opens it for every sample), and the prediction width is read from the shard."""
    stub = types.ModuleType('h5py')
    stub.File = _StubH5File
    monkeypatch.setitem(sys.modules, 'h5py', stub)
    src = SyntheticBertData(24, seq_len=32, max_predictions=6, vocab_size=500, seed=3)
    npz = str(tmp_path / 'shard_train.npz')
    src.write_npz(npz)
    h5 = str(tmp_path / 'shard_train.hdf5')
    import shutil
    shutil.copyfile(npz, h5 + '.npz')            # what the stub reads when asked for shard_train.hdf5
    _StubH5File.opened.clear()
    a, b = BertH5pyData(h5, max_pred_length=512), BertH5pyData(npz, max_pred_length=512)
    assert len(a) == len(b) == 24 and a.size(0) == 32 and a.max_predictions == 6
    for i in (0, 7, 23):
        for x, y in zip(a[i], b[i]):
            assert torch.equal(x, y)
    rows = np.array([3, 1, 20])
    for x, y in zip(a.fetch_rows(rows), b.fetch_rows(rows)):
        assert torch.equal(x, y)
    assert _StubH5File.opened.count(h5) == 2, _StubH5File.opened
    # labels follow data/h5pyDataset.py:42-48: -1 everywhere except the masked positions before the first zero
    ids, seg, mask, labels, nsp = a[5]
    pos = a.arrays['masked_lm_positions'][5]
    k = int(np.nonzero(pos == 0)[0][0]) if (pos == 0).any() else len(pos)
    assert (labels != -1).sum().item() == k
    assert ConBertH5pyData([a]).max_predictions == 6


# ---------------------------------------------------------------------------------------------------------
def _tiny_bert(tmp_path, dropout=0.0, extra=()):
    from hetseq_b200 import options, registry
    cfg = dict(vocab_size=120, hidden_size=32, num_hidden_layers=5, num_attention_heads=2, intermediate_size=64,
               hidden_act='gelu', hidden_dropout_prob=dropout, attention_probs_dropout_prob=dropout,
               max_position_embeddings=64, type_vocab_size=2, initializer_range=0.02)
    path = str(tmp_path / 'cfg.json')
    with open(path, 'w') as f:
        json.dump(cfg, f)
    args = options.make_args('bert', 'adam', 'PolynomialDecayScheduler',
                             ['--config_file', path, '--max-sentences', '4', '--synthetic-data', '--cpu', '--lr', '1e-3',
                              '--no-save', '--log-format', 'none', '--distributed-world-size', '1',
                              '--save-dir', str(tmp_path), *extra])
    task = registry.lookup('task', 'bert').setup_task(args)
    torch.manual_seed(0)
    return args, task, task.build_model(args)


@pytest.mark.parametrize('dropout', [0.0, 0.1])
def test_activation_checkpointing_gives_the_same_gradients(tmp_path, dropout):
    """--checkpoint-activations (reachable here; dead code in the reference, bert_modeling.py:470-476): recomputing
    the encoder in ceil(sqrt(L))-layer chunks must not change the loss or any gradient, dropout included (the
    recomputation replays the RNG state)."""
    _, task, model = _tiny_bert(tmp_path, dropout)
    ds = SyntheticBertData(8, 16, 3, 120, seed=1)
    batch = ds.fetch_rows(list(range(4)))
    out = {}
    for ckpt in (False, True):
        model.zero_grad()
        torch.manual_seed(7)
        loss = model(*batch, checkpoint_activations=ckpt)
        loss.backward()
        out[ckpt] = (loss.detach().clone(), [p.grad.clone() for p in model.parameters() if p.grad is not None])
    assert torch.allclose(out[False][0], out[True][0], rtol=0, atol=1e-6)
    assert len(out[False][1]) == len(out[True][1]) > 50
    for a, b in zip(out[False][1], out[True][1]):
        torch.testing.assert_close(a, b, rtol=1e-5, atol=1e-7)


def test_checkpoint_activations_flag_reaches_the_model(tmp_path):
    _, _, model = _tiny_bert(tmp_path, extra=('--checkpoint-activations',))
    assert model.checkpoint_activations is True
    _, _, model = _tiny_bert(tmp_path)
    assert model.checkpoint_activations is False


def test_init_linear_activation_flag(tmp_path):
    """Default: every Linear-like weight ~ N(0, 0.02). 'reference': the bias+activation layers keep the
    constructor's kaiming-uniform weights and non-zero biases, as /root/reference/bert_modeling.py:599-610 leaves
    them."""
    _, _, normal = _tiny_bert(tmp_path)
    _, _, ref = _tiny_bert(tmp_path, extra=('--init-linear-activation', 'reference'))
    w_n = normal.bert.encoder.layer[0].intermediate.dense_act
    w_r = ref.bert.encoder.layer[0].intermediate.dense_act
    assert float(w_n.weight.detach().std()) < 0.03 and float(w_n.bias.detach().abs().max()) == 0.0
    assert float(w_r.weight.detach().std()) > 0.05 and float(w_r.bias.detach().abs().max()) > 0.0
    # everything else is initialised identically in both modes
    assert float(ref.bert.encoder.layer[0].output.dense.weight.detach().std()) < 0.03


def test_pad_batch_is_inert(tmp_path):
    """A short batch grown to the static batch size: same loss, same gradients, padded rows not counted."""
    _, task, model = _tiny_bert(tmp_path)
    ds = SyntheticBertData(8, 16, 3, 120, seed=2)
    short = ds.fetch_rows([0, 1, 2])
    padded = task.pad_batch(short, 4)
    assert [t.shape[0] for t in padded] == [4] * 5
    assert task.sample_stats(padded) == task.sample_stats(short)
    assert task.sample_stats(short)[0] == 3
    grads = []
    for b in (short, padded):
        model.zero_grad()
        loss = model(*b)
        loss.backward()
        grads.append((loss.detach().clone(), [p.grad.clone() for p in model.parameters() if p.grad is not None]))
    torch.testing.assert_close(grads[0][0], grads[1][0], rtol=1e-6, atol=1e-6)
    for a, b in zip(grads[0][1], grads[1][1]):
        torch.testing.assert_close(a, b, rtol=1e-5, atol=1e-7)
    # with a static batch size the host-side preparation pads on its own and fixes the MLM index capacity
    task.static_batch_size = 4
    task._max_pred_seen = 3
    task.args.synthetic_max_predictions = 3
    prepared = task._host_part(short, narrow=False)
    assert prepared[0].shape[0] == 4 and prepared[5].numel() == 64       # 4 x 3 rounded up to 64


# ---------------------------------------------------------------------------------------------------------
def _launch_args(**kw):
    base = dict(cpu=False, distributed_init_method=None, distributed_world_size=1, distributed_gpus=4,
                distributed_rank=0, distributed_no_spawn=False, device_id=0)
    base.update(kw)
    return argparse.Namespace(**base)


def test_launch_plan_modes():
    """The three launch modes of /root/reference/train.py:213-239 as data."""
    from hetseq_b200.train import plan_launch
    assert plan_launch(_launch_args(), 8).mode == 'single'
    p = plan_launch(_launch_args(distributed_world_size=4), 8, port=12345)
    assert (p.mode, p.nprocs, p.init_method, p.first_rank) == ('spawn', 4, 'tcp://127.0.0.1:12345', 0)
    with pytest.raises(AssertionError):
        plan_launch(_launch_args(distributed_world_size=4), 2)
    p = plan_launch(_launch_args(distributed_init_method='tcp://10.0.0.1:9000', distributed_world_size=16,
                                 distributed_gpus=4, distributed_rank=8), 4)
    assert (p.mode, p.nprocs, p.init_method, p.first_rank) == ('spawn', 4, 'tcp://10.0.0.1:9000', 8)
    p = plan_launch(_launch_args(distributed_init_method='env://', distributed_world_size=8, distributed_gpus=4,
                                 distributed_rank=3, distributed_no_spawn=True), 1)
    assert (p.mode, p.nprocs, p.first_rank) == ('inline', 1, 3)
    with pytest.raises(AssertionError):
        plan_launch(_launch_args(distributed_init_method='tcp://h:1', distributed_gpus=8), 4)
    # CPU runs are not limited by the number of visible GPUs
    assert plan_launch(_launch_args(cpu=True, distributed_world_size=3), 0, port=1).nprocs == 3


def test_stat_schema_order_matches_reference_line():
    from hetseq_b200.train import STAT_SCHEMA
    keys = [k for k, _, _ in STAT_SCHEMA]
    assert keys == ['loss', 'nll_loss', 'ppl', 'wps', 'ups', 'wpb', 'bsz', 'num_updates', 'lr', 'gnorm', 'clip',
                    'oom', 'loss_scale', 'wall', 'train_wall']      # /root/reference/train.py:164-186


def test_sharded_iterator_padding_and_generators():
    from hetseq_b200.data.iterators import GroupedIterator, ShardedIterator

    class Sized(object):          # a sized iterable that can only be consumed once, like a DataLoader iterator
        def __init__(self, n):
            self.n = n

        def __len__(self):
            return self.n

        def __iter__(self):
            return (i * 10 for i in range(self.n))
    got = [list(ShardedIterator(Sized(7), 3, r, fill_value='pad')) for r in range(3)]
    assert got == [[0, 30, 60], [10, 40, 'pad'], [20, 50, 'pad']]
    assert [len(ShardedIterator(Sized(7), 3, r)) for r in range(3)] == [3, 3, 3]
    with pytest.raises(ValueError):
        ShardedIterator(Sized(3), 2, 2)
    g = GroupedIterator(Sized(7), 3)
    assert len(g) == 3 and list(g) == [[0, 10, 20], [30, 40, 50], [60]]


def test_training_budget_and_per_epoch_lists():
    """Stop rules of the epoch driver (reference train.py:74-88) and the per-epoch flag lists (--update-freq 4 2 1
    means 4 in epoch 1, 2 in epoch 2, 1 from epoch 3 on; reference train.py:113-116)."""
    import argparse
    from hetseq_b200.train import _Budget, _per_epoch

    assert [_per_epoch([4, 2, 1], e) for e in (1, 2, 3, 7)] == [4, 2, 1, 1]
    assert _per_epoch([3], 5) == 3

    class Ctl(object):
        def __init__(self, n):
            self.n = n

        def get_num_updates(self):
            return self.n

    class Itr(object):
        def __init__(self, epoch, pending=None):
            self.epoch, self._next_epoch_itr = epoch, pending

    b = _Budget(argparse.Namespace(min_lr=1e-5, max_epoch=3, max_update=100))
    assert b.keep_going(Ctl(10), Itr(2), 1e-3)
    assert not b.keep_going(Ctl(100), Itr(2), 1e-3)            # --max-update reached
    assert not b.keep_going(Ctl(10), Itr(3), 1e-3)             # --max-epoch epochs done
    assert b.keep_going(Ctl(10), Itr(3, pending=object()), 1e-3)   # resumed in the middle of the last epoch
    assert not b.keep_going(Ctl(10), Itr(1), 1e-5)             # learning rate fell to --min-lr
    unlimited = _Budget(argparse.Namespace(min_lr=-1.0, max_epoch=0, max_update=0))
    assert unlimited.keep_going(Ctl(10 ** 9), Itr(10 ** 6), 0.0)
