"""Bucket-readiness bookkeeping shared by the gradient backends (parallel/data_parallel.py::_ReadyBuckets): a bucket's
action runs exactly once, as soon as every parameter that overlaps it has reported in the LAST micro-batch, in
backward order; buckets whose parameters never report are swept up by ``_launch_rest``. The same rules DDP's reducer
follows for the reference (/root/reference/controller.py:74-89, bucket_cap_mb at options.py:122)."""
import pytest
import torch

from hetseq_b200.parallel.data_parallel import _ReadyBuckets
from hetseq_b200.parallel.flat_buffer import CHUNK, FlatParamBuffer


class _Recorder(_ReadyBuckets):
    def __init__(self, buffer, bucket_elems):
        self.launched_order = []
        self._setup_buckets(buffer, bucket_elems)

    def _launch(self, b):
        self._launched[b] = True
        self.launched_order.append(b)


def _buffer():
    torch.manual_seed(0)
    # sizes chosen so that parameters straddle bucket bounds (a bucket is 2 chunks here)
    params = [torch.nn.Parameter(torch.randn(n)) for n in (CHUNK // 2, 3 * CHUNK, CHUNK + 7, 5, 2 * CHUNK)]
    return FlatParamBuffer(params, grad_dtype=torch.float32), params


def test_buckets_fire_once_when_all_their_parameters_reported():
    buf, params = _buffer()
    r = _Recorder(buf, 2 * CHUNK)
    assert r.buckets[0][1] == buf.total and r.buckets[-1][0] == 0          # from the end of the buffer to its start
    assert all(e - s <= 2 * CHUNK and s % CHUNK == 0 for s, e in r.buckets)
    r._arm(True)
    fired_after = {}
    for i in reversed(range(len(params))):                                  # backward order: last parameter first
        r._on_param_ready(i)
        fired_after[i] = list(r.launched_order)
    # every bucket fired exactly once, each only after all of its parameters had reported
    assert sorted(r.launched_order) == list(range(len(r.buckets)))
    for b, plist in enumerate(r._bucket_params):
        first_seen = max(i for i in fired_after if b in fired_after[i])     # reports went from high i to low
        assert first_seen == min(plist), (b, plist, first_seen)
    # and in backward order: a bucket nearer the end of the buffer never fires after one nearer the start
    assert r.launched_order == sorted(r.launched_order)


def test_not_armed_means_no_launch_and_rest_is_swept():
    buf, params = _buffer()
    r = _Recorder(buf, 2 * CHUNK)
    r._arm(False)                                   # an earlier micro-batch of an accumulation window
    for i in range(len(params)):
        r._on_param_ready(i)
    assert r.launched_order == []
    r._arm(True)
    r._on_param_ready(len(params) - 1)              # only the last parameter reports (others unused in this step)
    done = list(r.launched_order)
    r._launch_rest()
    assert sorted(r.launched_order) == list(range(len(r.buckets))) and r.launched_order[:len(done)] == done
    assert not r._armed


def test_double_report_is_an_error():
    buf, params = _buffer()
    r = _Recorder(buf, 2 * CHUNK)
    r._arm(True)
    r._on_param_ready(0)
    with pytest.raises(RuntimeError):
        r._on_param_ready(0)
