"""Multi-process data parallelism on CPU (Gloo): baseline all-reduce backend, sharded-protocol backend,
equivalence with a single process over the concatenated batch, uneven shards, stats exchange."""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, 'tests', 'dp_worker.py')


def _run(tmp_path, nproc, backend, tag, extra=()):
    out = os.path.join(str(tmp_path), 'out_%s.pt' % tag)
    port = 29500 + (os.getpid() * 7 + hash(tag)) % 2000
    cmd = [sys.executable]
    if nproc > 1:
        cmd += ['-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', str(nproc), '--master-addr',
                '127.0.0.1', '--master-port', str(port)]
    cmd += [WORKER, '--backend', backend, '--out', out, '--cpu', *extra]
    env = dict(os.environ, OMP_NUM_THREADS='2')
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    return torch.load(out, weights_only=False)


def test_allreduce_and_sharded_backends_match_single_process(tmp_path):
    """2 ranks x 4 sequences == 1 process x (2 micro-batches of 4): same parameters after 3 updates."""
    single = _run(tmp_path, 1, 'nccl', 'single', ['--update-freq', '2'])
    ar = _run(tmp_path, 2, 'nccl', 'ar')
    sh = _run(tmp_path, 2, 'sharded', 'sh')
    assert ar['backend'] == 'nccl' and sh['backend'] == 'sharded-allreduce'
    assert ar['consistent'] and sh['consistent']
    torch.testing.assert_close(ar['params'], single['params'], rtol=1e-4, atol=2e-5)
    torch.testing.assert_close(sh['params'], ar['params'], rtol=1e-5, atol=1e-6)
    for a, b in zip(ar['log'], single['log']):
        assert abs(a['loss'] - b['loss']) < 1e-4 and abs(a['gnorm'] - b['gnorm']) < 1e-3 * max(1.0, b['gnorm'])
        assert a['nsentences'] == 8
    # the sharded optimizer's gathered state equals the replicated one (checkpoint interchangeability)
    for k in ar['opt_state']:
        for name in ('exp_avg', 'exp_avg_sq'):
            torch.testing.assert_close(sh['opt_state'][k][name], ar['opt_state'][k][name], rtol=1e-5, atol=1e-7)


def test_adadelta_sharded(tmp_path):
    ar = _run(tmp_path, 2, 'nccl', 'ar_ad', ['--optimizer', 'adadelta'])
    sh = _run(tmp_path, 2, 'sharded', 'sh_ad', ['--optimizer', 'adadelta'])
    torch.testing.assert_close(sh['params'], ar['params'], rtol=1e-5, atol=1e-6)
