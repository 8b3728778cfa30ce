"""Executable specification of the segmented embedding-gradient reduction (csrc/embedding_sm100.cu:
embed_bwd_segment_kernel / embed_bwd_combine_kernel), in NumPy, one Python "CTA" per segment.

The CUDA kernels follow this model decision for decision (same shared arrays, same enters / leaves rules, same
owner rule for runs that cross segments); tests/test_gpu_kernels.py::test_embedding_kernels checks the kernels
themselves on a GPU. Here the boundary logic is exercised on the CPU against ``index_add_``: long [PAD]-like runs
over many segments, runs that end exactly on a segment edge, a ragged last segment, ids outside the table.
Replaces the atomic scatter of the reference's embedding backward (/root/reference/bert_modeling.py:306-320).
"""
import numpy as np
import pytest
import torch

from hetseq_b200.tasks.bert import LanguageModelingTask

ESEG = 16


def _segment_ids(ids, order, s, M, vocab):
    sid = np.full(ESEG + 2, -1, dtype=np.int64)
    pos = np.full(ESEG, -1, dtype=np.int64)
    for k in range(ESEG + 2):
        j = s * ESEG - 1 + k
        if 0 <= j < M:
            p = int(order[j])
            if 0 <= p < M:
                i = int(ids[p])
                sid[k] = i if 0 <= i < vocab else -1
                if 1 <= k <= ESEG:
                    pos[k - 1] = p
    return sid, pos


def model_backward(de, ids, seg, order, run_end, dW, vocab):
    """dW += scatter of de by ids; returns the [2, H] token-type sums. float32 throughout, the kernels' order."""
    M, H = de.shape
    nseg = (M + ESEG - 1) // ESEG
    partW = np.full((nseg, 2, H), np.nan, dtype=np.float32)      # NaN: reading a piece nobody wrote shows up
    partT = np.zeros((nseg, 2, H), dtype=np.float32)
    # ---- phase 1
    for s in range(nseg):
        sid, pos = _segment_ids(ids, order, s, M, vocab)
        acc = np.zeros(H, dtype=np.float32)
        a = 0
        for u in range(ESEG):
            row = de[pos[u]].astype(np.float32) if pos[u] >= 0 else np.zeros(H, dtype=np.float32)
            acc = acc + row
            if pos[u] >= 0 and 0 <= int(seg[pos[u]]) < 2:
                partT[s, int(seg[pos[u]])] += row
            i = sid[u + 1]
            if u == ESEG - 1 or sid[u + 2] != i:
                if i >= 0:
                    enters = a == 0 and sid[0] == i
                    leaves = u == ESEG - 1 and sid[ESEG + 1] == i
                    if not enters and not leaves:
                        dW[i] += acc
                    else:
                        partW[s, 0 if enters else 1] = acc
                acc = np.zeros(H, dtype=np.float32)
                a = u + 1
    # ---- phase 2
    for s in range(nseg):
        if (s + 1) * ESEG >= M:
            continue
        sid, _ = _segment_ids(ids, order, s, M, vocab)
        i = sid[ESEG]
        if i < 0 or sid[ESEG + 1] != i:
            continue
        a = ESEG - 1
        while a > 0 and sid[a] == i:
            a -= 1
        if a == 0 and sid[0] == i:
            continue
        e = min(int(run_end[s * ESEG + a]), M)
        s_last = (e - 1) // ESEG
        if s_last <= s:
            continue
        x = partW[s, 1].copy()
        for s2 in range(s + 1, s_last + 1):
            x = x + partW[s2, 0]
        dW[i] += x
    dT = np.zeros((2, H), dtype=np.float32)
    groups = [sum((partT[s] for s in range(g, nseg, 8)), np.zeros((2, H), dtype=np.float32)) for g in range(8)]
    for g in range(8):
        dT = dT + groups[g]
    return dT


def _case(M, vocab, H, seed, pad_frac, edge_runs=False, bad_ids=False):
    g = np.random.default_rng(seed)
    ids = g.integers(1, vocab, size=M).astype(np.int64)
    if edge_runs:
        # runs engineered to start / end exactly on segment edges in sorted order: ids 0 (x16), 1 (x32), 2 (x15), 3 (x1)
        ids[:64] = np.repeat([0, 1, 2, 3], [16, 32, 15, 1])
        ids[64:] = g.integers(10, vocab, size=M - 64)
        g.shuffle(ids)
    npad = int(pad_frac * M)
    if npad:
        ids[g.choice(M, size=npad, replace=False)] = 0
    if bad_ids:
        ids[g.choice(M, size=9, replace=False)] = vocab + 5
        ids[g.choice(M, size=5, replace=False)] = -3
    seg = g.integers(0, 2, size=M).astype(np.int64)
    de = g.standard_normal((M, H)).astype(np.float16)
    return ids, seg, de


@pytest.mark.parametrize('M,pad,edge,bad', [(4096, 0.25, False, False), (200, 0.0, True, False), (77, 0.5, False, True),
                                            (16, 1.0, False, False), (17, 1.0, False, False), (5, 0.0, False, False),
                                            (1024, 0.9, True, True)])
def test_segment_model_matches_index_add(M, pad, edge, bad):
    vocab, H = 300, 8
    ids, seg, de = _case(M, vocab, H, seed=M, pad_frac=pad, edge_runs=edge and M >= 64, bad_ids=bad)
    order, run_end = (t.numpy() for t in LanguageModelingTask.embedding_plan(torch.from_numpy(ids)))
    dW = np.ones((vocab, H), dtype=np.float32)
    dT = model_backward(de, ids, seg, order, run_end, dW, vocab)
    ok = (ids >= 0) & (ids < vocab)
    refW = torch.ones(vocab, H, dtype=torch.float64).index_add_(0, torch.from_numpy(ids[ok]),
                                                                torch.from_numpy(de[ok].astype(np.float64)))
    refT = torch.zeros(2, H, dtype=torch.float64).index_add_(0, torch.from_numpy(seg), torch.from_numpy(de.astype(np.float64)))
    assert not np.isnan(dW).any(), 'a piece of a crossing run was read but never written'
    np.testing.assert_allclose(dW, refW.numpy(), rtol=1e-4, atol=2e-3)
    np.testing.assert_allclose(dT, refT.numpy(), rtol=1e-4, atol=2e-3)


def test_device_plan_and_host_plan_agree():
    """The static-shape tensor plan (usable inside a captured graph) and the numpy plan of the prefetch thread."""
    from hetseq_b200.ops import native
    ids = torch.from_numpy(_case(333, 50, 8, 1, 0.3)[0])
    a = native.embedding_scatter_plan(ids)
    b = LanguageModelingTask.embedding_plan(ids)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])


def test_segment_model_fuzz():
    """Property test over random batches: any length, tiny vocabularies (so runs of every length and alignment
    relative to the 16-slot segments occur), ids outside the table mixed in."""
    hypothesis = pytest.importorskip('hypothesis')
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=120, deadline=None)
    @given(st.integers(1, 200), st.integers(1, 12), st.integers(0, 2 ** 31 - 1), st.booleans())
    def prop(M, vocab, seed, with_bad):
        g = np.random.default_rng(seed)
        lo, hi = (-2, vocab + 2) if with_bad else (0, vocab)
        ids = g.integers(lo, hi, size=M).astype(np.int64)
        seg = g.integers(0, 2, size=M).astype(np.int64)
        de = g.integers(-8, 9, size=(M, 8)).astype(np.float16)     # small integers: every order of additions is exact
        order, run_end = (t.numpy() for t in LanguageModelingTask.embedding_plan(torch.from_numpy(ids)))
        dW = np.zeros((vocab, 8), dtype=np.float32)
        dT = model_backward(de, ids, seg, order, run_end, dW, vocab)
        ok = (ids >= 0) & (ids < vocab)
        refW = np.zeros((vocab, 8), dtype=np.float32)
        np.add.at(refW, ids[ok], de[ok].astype(np.float32))
        refT = np.zeros((2, 8), dtype=np.float32)
        np.add.at(refT, seg, de.astype(np.float32))
        assert np.array_equal(dW, refW) and np.array_equal(dT, refT)

    prop()
