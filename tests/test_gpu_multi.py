"""Multi-GPU (>= 2 B200s): the fused NVLink gradient path against the NCCL all-reduce baseline."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, 'tests', 'dp_worker.py')


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _run(tmp_path, nproc, backend, tag, extra=()):
    out = os.path.join(str(tmp_path), 'out_%s.pt' % tag)
    port = 29600 + (os.getpid() * 7 + abs(hash(tag))) % 2000
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', str(nproc),
           '--master-addr', '127.0.0.1', '--master-port', str(port), WORKER, '--backend', backend, '--out', out,
           '--precision', 'fp16', '--hidden', '256', '--seq', '128', '--steps', '4', *extra]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-6000:]
    return torch.load(out, weights_only=False)


@pytest.mark.skipif(_ngpu() < 2, reason='needs at least 2 GPUs')
def test_fused_nvlink_matches_nccl_baseline(tmp_path):
    n = 2 if _ngpu() < 4 else (4 if _ngpu() < 8 else 8)
    base = _run(tmp_path, n, 'nccl', 'nccl')
    fused = _run(tmp_path, n, 'fused', 'fused')
    assert base['backend'] == 'nccl' and fused['backend'] == 'fused-nvlink'
    assert base['consistent'] and fused['consistent'], 'replicas diverged'
    # same math, different summation order: fp16 parameters after 4 Adam updates
    diff = (fused['params'] - base['params']).abs()
    assert diff.max().item() < 2e-3, diff.max().item()
    assert (diff > 1e-3).float().mean().item() < 1e-3
    for a, b in zip(fused['log'], base['log']):
        assert (a is None) == (b is None)
        if a is not None:
            assert abs(a['loss'] - b['loss']) < 2e-2, (a, b)
            assert abs(a['gnorm'] - b['gnorm']) < 2e-2 * max(1.0, b['gnorm']), (a, b)
    for k in base['opt_state']:
        torch.testing.assert_close(fused['opt_state'][k]['exp_avg'], base['opt_state'][k]['exp_avg'], rtol=5e-2, atol=1e-5)
