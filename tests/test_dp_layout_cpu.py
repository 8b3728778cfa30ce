"""Host-side layout rules of the fused data-parallel path, on CPU (the kernel library loads without a GPU; these
entry points launch nothing): which parameter slice ("gather bucket") a block of the flat buffer belongs to, how
many such slices exist, how many signal words a rank needs. The forward-side parameter waits
(parallel/fused_dp.py::FusedBackend.gate) and the gather kernel (csrc/comm_sm100.cu::gather_body) must agree on
these, or a layer would wait for a slice that is never published -- the bug class behind an in-kernel wait that
never ends. Plays the role of DDP's bucket assignment in the reference (/root/reference/controller.py:74-89)."""
import pytest

from hetseq_b200.ops import lib

pytestmark = pytest.mark.skipif(not lib.available(), reason='kernel library not built')


def test_gather_buckets_partition_the_blocks():
    pytest.importorskip('hypothesis')
    from hypothesis import given, settings, strategies as st
    L = lib.load()
    cap = int(L.hs_dp_max_gather_buckets())

    @settings(max_examples=300, deadline=None)
    @given(st.integers(1, 5000), st.integers(1, cap))          # the launch entry points refuse more than `cap`
    def prop(nblocks, req):
        n = int(L.hs_dp_gather_count(nblocks, req))
        assert 1 <= n <= min(req, nblocks)
        owners = [int(L.hs_dp_gather_bucket_of(nblocks, req, b)) for b in (0, nblocks // 3, nblocks // 2, nblocks - 1)]
        assert owners == sorted(owners) and owners[0] == 0 and owners[-1] == n - 1   # monotone, first and last used
        if nblocks <= 400:
            full = [int(L.hs_dp_gather_bucket_of(nblocks, req, b)) for b in range(nblocks)]
            assert full == sorted(full) and set(full) == set(range(n)), 'every counted bucket owns a block'
            sizes = [full.count(j) for j in range(n)]
            assert max(sizes[:-1] or sizes) == sizes[0] and sizes[-1] <= sizes[0]     # equal slices, last may be short

    prop()


def test_signal_words_grow_with_world_and_buckets():
    L = lib.load()
    last = 0
    for W in (1, 2, 4, 8):
        words = int(L.hs_dp_signal_words(W, 64))
        # 3 words per rank for the norm / sample-size exchange, one arrival flag per (gather bucket, source rank),
        # one per gradient bucket
        assert words >= 3 * W + int(L.hs_dp_max_gather_buckets()) * W + 64
        assert words > last
        last = words


def test_scratch_size_helpers():
    L = lib.load()
    assert int(L.hs_embed_backward_scratch_elems(4096, 1024)) == (4096 // 16) * 4 * 1024
    assert int(L.hs_embed_backward_scratch_elems(17, 64)) == 2 * 4 * 64
    assert int(L.hs_version()) >= 100


def test_launch_without_a_gpu_is_a_loud_error():
    """No silent fallback: on a machine without a usable GPU a kernel entry point returns an error that the binding
    turns into an exception naming the kernel and the CUDA reason."""
    import torch
    if torch.cuda.is_available():
        pytest.skip('needs a machine without a GPU')
    L = lib.load()
    x, out, sc = torch.zeros(4096), torch.zeros(1), torch.zeros(int(L.hs_sumsq_scratch_elems()) + 8)
    rc = L.hs_sumsq(x.data_ptr(), lib.dtype_code(torch.float32), 2, 2048, 2048, sc.data_ptr(), out.data_ptr(), 0, None)
    assert rc != 0
    with pytest.raises(RuntimeError, match='hs_sumsq failed'):
        lib.check(rc, 'hs_sumsq')
