"""Batch planner, iterators, datasets (reference semantics: data/data_utils_fast.pyx, data/iterators.py,
data/h5pyDataset.py)."""
import numpy as np
import pytest
import torch

from hetseq_b200.data import data_utils, iterators
from hetseq_b200.data.bert_dataset import (BertH5pyData, ConBertH5pyData, SyntheticBertData, batch_labels,
                                           _labels_from_positions)
from hetseq_b200.data.mnist_dataset import MNISTDataset, make_synthetic_mnist


def ref_batch_by_size(indices, sizes, max_tokens, max_sentences, bsz_mult):
    """Literal restatement of data_utils_fast.pyx:21-62 with Python lists."""
    sample_len, sample_lens, batch, batches = 0, [], [], []
    for idx in indices:
        nt = sizes[idx]
        sample_lens.append(nt)
        sample_len = max(sample_len, nt)
        assert sample_len <= max_tokens
        num_tokens = (len(batch) + 1) * sample_len
        full = len(batch) > 0 and (len(batch) == max_sentences or num_tokens > max_tokens)
        if full:
            mod_len = max(bsz_mult * (len(batch) // bsz_mult), len(batch) % bsz_mult)
            batches.append(batch[:mod_len])
            batch = batch[mod_len:]
            sample_lens = sample_lens[mod_len:]
            sample_len = max(sample_lens) if len(sample_lens) > 0 else 0
        batch.append(idx)
    if len(batch) > 0:
        batches.append(batch)
    return batches


@pytest.mark.parametrize('native', [True, False])
@pytest.mark.parametrize('cfg', [(10 ** 9, 4, 1), (64, 10 ** 9, 1), (100, 7, 4), (48, 5, 8), (512, 3, 2)])
def test_batch_by_size_matches_reference_rule(native, cfg):
    max_tokens, max_sentences, mult = cfg
    rng = np.random.RandomState(0)
    sizes = rng.randint(1, 40, size=257).astype(np.int64)
    indices = rng.permutation(257).astype(np.int64)
    plan = data_utils.batch_by_size(indices, None, max_tokens, max_sentences, mult, sizes=sizes, use_native=native)
    assert plan.tolist() == ref_batch_by_size(indices.tolist(), sizes.tolist(), max_tokens, max_sentences, mult)
    # the callback form of the reference API gives the same plan
    plan2 = data_utils.batch_by_size(indices, lambda i: int(sizes[i]), max_tokens, max_sentences, mult, use_native=native)
    assert plan2.tolist() == plan.tolist()


def test_batch_by_size_rejects_oversized_sample():
    with pytest.raises(AssertionError):
        data_utils.batch_by_size(np.arange(3), None, max_tokens=8, sizes=np.array([3, 9, 2]))


def test_numpy_seed_restores_state():
    np.random.seed(5)
    a = np.random.rand()
    np.random.seed(5)
    with data_utils.numpy_seed(123):
        inside = np.random.rand()
    assert np.random.rand() == a
    with data_utils.numpy_seed(123):
        assert np.random.rand() == inside


def test_sharded_grouped_counting_iterators():
    sh = [list(iterators.ShardedIterator(list(range(10)), 4, r, fill_value=[])) for r in range(4)]
    assert sh == [[0, 4, 8], [1, 5, 9], [2, 6, []], [3, 7, []]]
    with pytest.raises(ValueError):
        iterators.ShardedIterator([1, 2], 2, 2)
    c = iterators.CountingIterator(list(range(7)))
    g = iterators.GroupedIterator(c, 3)
    assert len(g) == 3 and list(g) == [[0, 1, 2], [3, 4, 5], [6]]
    c = iterators.CountingIterator(list(range(7)))
    c.skip(3)
    assert next(c) == 3 and c.count == 4 and c.has_next()
    np.testing.assert_array_equal(data_utils.shard_plan(None, 10, 4, 2), [2, 6, -1])


class _ToyDataset(torch.utils.data.Dataset):
    def __init__(self, n):
        self.n = n
        self.epoch = None

    def __len__(self):
        return self.n

    def __getitem__(self, i):
        return torch.tensor(i)

    def collater(self, samples):
        return None if len(samples) == 0 else torch.stack(samples)

    def set_epoch(self, e):
        self.epoch = e


def _epoch_itr(n=23, bsz=4, shards=1, shard=0, seed=7):
    ds = _ToyDataset(n)
    plan = data_utils.batch_by_size(np.arange(n), None, max_sentences=bsz, sizes=np.ones(n, dtype=np.int64))
    return iterators.EpochBatchIterator(ds, ds.collater, plan, seed=seed, num_shards=shards, shard_id=shard)


def test_epoch_iterator_shuffle_is_seeded_and_identical_across_ranks():
    a, b = _epoch_itr(), _epoch_itr()
    e1a = [x.tolist() for x in a.next_epoch_itr(shuffle=True)]
    e1b = [x.tolist() for x in b.next_epoch_itr(shuffle=True)]
    assert e1a == e1b and a.epoch == 1 and a.dataset.epoch == 1
    e2a = [x.tolist() for x in a.next_epoch_itr(shuffle=True)]
    assert sorted(sum(e2a, [])) == list(range(23)) and e2a != e1a
    # shards of one epoch partition the batches; the short shard is padded with None (empty batch)
    parts = []
    for r in range(3):
        it = _epoch_itr(shards=3, shard=r)
        parts.append(list(it.next_epoch_itr(shuffle=True)))
    assert len({len(p) for p in parts}) == 1
    flat = sorted(i for p in parts for x in p if x is not None for i in x.tolist())
    assert flat == list(range(23))
    assert sum(x is None for p in parts for x in p) == 3 * len(parts[0]) - 6


def test_epoch_iterator_resume_fast_forward():
    a = _epoch_itr()
    it = a.next_epoch_itr(shuffle=True)
    consumed = [next(it).tolist() for _ in range(2)]
    state = a.state_dict()
    assert state == {'epoch': 1, 'iterations_in_epoch': 2}
    rest = []
    while it.has_next():
        rest.append(next(it).tolist())
    b = _epoch_itr()
    b.load_state_dict(state)
    resumed = b.next_epoch_itr(shuffle=True)
    got = []
    while resumed.has_next():
        got.append(next(resumed).tolist())
    assert b.epoch == 1 and got == rest and resumed.count == len(consumed) + len(rest)
    # a finished epoch resumes at the next one
    c = _epoch_itr()
    c.load_state_dict({'epoch': 1, 'iterations_in_epoch': 6})
    c.next_epoch_itr(shuffle=True)
    assert c.epoch == 2


def test_prefetch_iterator_preserves_order_and_errors():
    src = iterators.CountingIterator([torch.tensor([i]) for i in range(9)])
    out = [int(x) for x in iterators.PrefetchIterator(src, depth=2, pin_memory=False)]
    assert out == list(range(9))

    def boom():
        yield 1
        raise RuntimeError('x')
    with pytest.raises(RuntimeError):
        list(iterators.PrefetchIterator(boom(), pin_memory=False))


def test_bert_labels_and_shards(tmp_path):
    pos = np.array([3, 7, 9, 0, 0]); ids = np.array([11, 12, 13, 0, 0])
    lab = _labels_from_positions(12, pos, ids, 512)
    assert lab.tolist() == [-1, -1, -1, 11, -1, -1, -1, 12, -1, 13, -1, -1]
    np.testing.assert_array_equal(batch_labels(12, pos[None], ids[None], 512)[0], lab)
    np.testing.assert_array_equal(batch_labels(12, pos[None], ids[None], 2)[0][[3, 7, 9]], [11, 12, -1])
    syn = SyntheticBertData(40, 32, 5, 500, seed=1)
    syn.write_npz(str(tmp_path / 'train_a.npz'), rows=range(0, 25))
    syn.write_npz(str(tmp_path / 'train_b.npz'), rows=range(25, 40))
    shards = [BertH5pyData(str(tmp_path / f)) for f in ('train_a.npz', 'train_b.npz')]
    con = ConBertH5pyData(shards)
    assert len(con) == 40 and con.size(3) == 32 and con.sizes_array().tolist() == [32] * 40
    ref = syn.fetch_rows([2, 30, 24, 25])
    got = con.fetch_batch([2, 30, 24, 25])
    one = con.collater([con[i] for i in [2, 30, 24, 25]])
    for r, g, o in zip(ref, got, one):
        assert torch.equal(r, g) and torch.equal(r, o)
    assert con.collater([]) is None and con.fetch_batch([]) is None
    up = ConBertH5pyData(shards, sample_ratios=[2, 1])      # up-sampling wraps around inside the shard
    assert len(up) == 65 and torch.equal(up[26][0], up[1][0])


def test_mnist_dataset(tmp_path):
    root = make_synthetic_mnist(str(tmp_path), 20, 8)
    ds = MNISTDataset(root + '/training.pt')
    img, target = ds[3]
    assert img.shape == (1, 28, 28) and isinstance(target, int) and ds.num_tokens(0) == 1
    raw = ds.image[3].float() / 255
    torch.testing.assert_close(img[0], (raw - 0.1307) / 0.3081)
    b = ds.fetch_batch([0, 3])
    assert b[0].shape == (2, 1, 28, 28) and torch.equal(b[0][1], img)


def test_prefetch_iterator_keeps_resume_position_exact():
    """A prefetcher reads ahead of the trainer; the checkpointed position must be the trainer's. Also checks the
    worker-thread transform and that resuming from the recorded position yields exactly the untrained batches."""
    import torch
    from hetseq_b200.data import iterators

    class DS(torch.utils.data.Dataset):
        def __len__(self):
            return 40

        def __getitem__(self, i):
            return torch.tensor([i])

        def collater(self, samples):
            return None if len(samples) == 0 else [torch.stack(samples)]

    ds = DS()
    batches = [[2 * i, 2 * i + 1] for i in range(20)]

    def make():
        return iterators.EpochBatchIterator(ds, ds.collater, batches, seed=3, num_shards=1, shard_id=0)

    ref = make()
    expected = [b[0].flatten().tolist() for b in ref.next_epoch_itr(shuffle=True)]

    ebi = make()
    itr = iterators.GroupedIterator(ebi.next_epoch_itr(shuffle=True), 2)
    seen_transform = []
    pre = iterators.PrefetchIterator(itr, depth=3, pin_memory=False,
                                     transform=lambda chunk: (seen_transform.append(len(chunk)) or chunk)).attach_to(ebi)
    got = []
    for k, chunk in enumerate(pre):
        got += [b[0].flatten().tolist() for b in chunk]
        if k == 2:
            break
    import time
    time.sleep(0.2)                                  # let the worker run ahead
    assert ebi._cur_epoch_itr.count > 6              # the reader is ahead of the trainer ...
    state = ebi.state_dict()
    assert state['iterations_in_epoch'] == 6         # ... but the recorded position is the trainer's
    assert not ebi.end_of_epoch()
    assert got == expected[:6] and seen_transform[:3] == [2, 2, 2]

    resumed = make()
    resumed.load_state_dict(state)
    rest = [b[0].flatten().tolist() for b in resumed.next_epoch_itr(shuffle=True)]
    assert rest == expected[6:]

    # consuming everything through a prefetcher ends the epoch
    ebi2 = make()
    pre2 = iterators.PrefetchIterator(iterators.GroupedIterator(ebi2.next_epoch_itr(shuffle=True), 3), depth=2,
                                      pin_memory=False).attach_to(ebi2)
    n = sum(len(c) for c in pre2)
    assert n == 20 and ebi2.end_of_epoch() and ebi2.iterations_in_epoch == 20


def test_packed_sample_views_one_buffer():
    import torch
    from hetseq_b200.utils.utils import PackedSample, move_to_device
    parts = [torch.arange(6, dtype=torch.int32).view(2, 3), torch.tensor([7], dtype=torch.int32),
             torch.arange(5, dtype=torch.int32)]
    p = PackedSample.pack(parts)
    assert isinstance(p, list) and len(p) == 3 and all(torch.equal(a, b) for a, b in zip(p, parts))
    assert all(t.data_ptr() % 16 == p.flat.data_ptr() % 16 for t in p)          # views stay 16-byte aligned
    p.flat += 1                                                               # views alias the flat buffer
    assert p[1].item() == 8
    c = p.clone()
    assert c.same_layout(p) and c.flat.data_ptr() != p.flat.data_ptr() and torch.equal(c[2], p[2])
    assert move_to_device(p, 'cpu') is p
    assert PackedSample.pack([torch.ones(2), torch.ones(2, dtype=torch.int32)]) is None     # mixed dtypes
    assert not p.same_layout([t for t in p])
    # the BERT task packs its 7 int32 tensors when preparing for a CUDA device
    from hetseq_b200.data.iterators import PrefetchIterator
    pre = PrefetchIterator(iter([[p]]), depth=1, pin_memory=False)
    assert next(pre)[0] is p


def test_batch_planner_fuzz_native_numpy_and_reference_rule_agree():
    """Property test: for random sample sizes, orders and limits the C++ planner, the NumPy fallback and the literal
    restatement of the reference's Cython rule produce the same batches; every sample lands in exactly one batch, in
    order, and no batch breaks the sentence limit (nor, without a batch-size multiple, the token limit)."""
    pytest.importorskip('hypothesis')
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=150, deadline=None)
    @given(st.integers(1, 120), st.integers(0, 2 ** 31 - 1), st.integers(1, 9), st.sampled_from([1, 2, 4, 8]),
           st.integers(1, 6))
    def prop(n, seed, max_sentences, mult, tok_factor):
        rng = np.random.RandomState(seed)
        sizes = rng.randint(1, 33, size=n).astype(np.int64)
        indices = rng.permutation(n).astype(np.int64)
        max_tokens = int(sizes.max()) * tok_factor
        want = ref_batch_by_size(indices.tolist(), sizes.tolist(), max_tokens, max_sentences, mult)
        for native in (True, False):
            plan = data_utils.batch_by_size(indices, None, max_tokens, max_sentences, mult, sizes=sizes,
                                            use_native=native).tolist()
            assert plan == want
        flat = [i for b in want for i in b]
        assert flat == indices.tolist()
        for b in want:
            assert 0 < len(b) <= max_sentences
            if mult == 1:
                # with a batch-size multiple the reference rule re-opens a batch with the samples beyond the last
                # multiple and appends to it unchecked (data_utils_fast.pyx:48-56), so the token limit can be
                # exceeded there; the planners reproduce that faithfully (asserted above) rather than "fix" it
                assert len(b) * int(sizes[b].max()) <= max_tokens

    prop()
