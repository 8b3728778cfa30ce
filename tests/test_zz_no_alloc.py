"""hs_sumsq_noalloc: the gradient-norm reduction for hosts that may not allocate device memory.

A process that has handed the whole GPU to its own pool, or a C/C++ program that links the C ABI
without any caching allocator, cannot give hs_sumsq the scratch buffer it wants (and a cudaMalloc in
the middle of a step loop would synchronise the device). hs_sumsq_noalloc keeps its block partials in
an array inside the library instead. This test drives it through the C ABI exactly as such a host
would (ctypes, raw pointers) and checks it against hs_sumsq with caller-provided scratch, twice in a
row: the array is shared by all calls and keeps the previous call's contents, so the second call --
on different data and a different size -- must not be affected by what the first one left behind.

Collected last on purpose (file name), so that a run stopping at the first failure has already run
every other test. It must pass both alone in a fresh process and after the rest of the suite.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from hetseq_b200.ops.build import CUDA_LIB
    L = ctypes.CDLL(str(CUDA_LIB))
    vp, ci, cl = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong
    L.hs_sumsq.argtypes = [vp, ci, cl, ci, cl, vp, vp, ci, vp]
    L.hs_sumsq.restype = ci
    L.hs_sumsq_noalloc.argtypes = [vp, ci, cl, ci, cl, vp, ci, vp]
    L.hs_sumsq_noalloc.restype = ci
    L.hs_sumsq_scratch_elems.restype = ci
    L.hs_sumsq_noalloc_max_partials.restype = ci
    L.hs_last_error.restype = ctypes.c_char_p
    return L


def _both(L, g, nchunks, chunk, stride, dtype_code):
    """(reference via hs_sumsq with our own scratch, result of hs_sumsq_noalloc) on the current stream."""
    stream = torch.cuda.current_stream().cuda_stream
    scratch = torch.empty(L.hs_sumsq_scratch_elems(), dtype=torch.float32, device='cuda')
    ref = torch.zeros(1, dtype=torch.float32, device='cuda')
    out = torch.full((1,), -1.0, dtype=torch.float32, device='cuda')
    rc = L.hs_sumsq(g.data_ptr(), dtype_code, nchunks, chunk, stride, scratch.data_ptr(), ref.data_ptr(), 0, stream)
    assert rc == 0, L.hs_last_error().decode()
    rc = L.hs_sumsq_noalloc(g.data_ptr(), dtype_code, nchunks, chunk, stride, out.data_ptr(), 0, stream)
    assert rc == 0, L.hs_last_error().decode()
    torch.cuda.synchronize()
    return ref.item(), out.item()


def test_noalloc_matches_scratch_variant_twice_in_a_row():
    L = _lib()
    torch.manual_seed(0)
    chunk = 2048
    # call 1: fp32, many chunks, large values -> leaves big partials behind in the shared array
    g1 = torch.randn(3000 * chunk, device='cuda') * 50.0
    ref1, got1 = _both(L, g1, 3000, chunk, chunk, 2)
    assert abs(got1 - ref1) <= 1e-6 * abs(ref1), (got1, ref1)
    assert abs(got1 - float((g1.double() ** 2).sum())) <= 1e-4 * abs(ref1)
    # call 2: fp16, few chunks (fewer partials than call 1 wrote), strided like a 4-way sharded buffer
    g2 = (torch.randn(4 * 37 * chunk, device='cuda') * 0.01).half()
    ref2, got2 = _both(L, g2, 37, chunk, 4 * chunk, 0)
    assert abs(got2 - ref2) <= 1e-6 * abs(ref2), (got2, ref2)
    own = g2.view(37, 4, chunk)[:, 0, :].double()
    assert abs(got2 - float((own ** 2).sum())) <= 2e-3 * abs(ref2)
    # accumulate semantics are the same as hs_sumsq's
    stream = torch.cuda.current_stream().cuda_stream
    acc = torch.full((1,), 5.0, dtype=torch.float32, device='cuda')
    assert L.hs_sumsq_noalloc(g2.data_ptr(), 0, 37, chunk, 4 * chunk, acc.data_ptr(), 1, stream) == 0
    torch.cuda.synchronize()
    assert abs(acc.item() - (5.0 + ref2)) <= 1e-5 * (5.0 + abs(ref2))


def test_noalloc_rejects_bad_layout_with_a_message():
    L = _lib()
    assert L.hs_sumsq_noalloc_max_partials() == 4096
    g = torch.zeros(4096, device='cuda')
    out = torch.zeros(1, device='cuda')
    rc = L.hs_sumsq_noalloc(g.data_ptr(), 2, 2, 2047, 2047, out.data_ptr(), 0, None)   # chunk not a multiple of 8
    assert rc != 0 and b'hs_sumsq_noalloc' in L.hs_last_error()
