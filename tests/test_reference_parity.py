"""This framework against the UNMODIFIED reference (baseline/_ref) used as an oracle, component by component, on
CPU: LR schedule, Adam / Adadelta math, iterators, the batch planner vs the reference's Cython extension, the BERT
pre-training loss and every parameter gradient, BertModel and the six task heads (logits, loss, gradients), checkpoint files in both directions, meters / log formatting, flag names and defaults.

The checks run in a separate interpreter (tests/ref_parity_worker.py) because the reference uses flat module names
(``utils``, ``data``, ``options`` ...). Skipped when the reference is not installed (baseline/install_reference.sh)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHECKS = ['check_bert_training_curve', 'check_mnist_training_curve', 'check_checkpoint_interchange', 'check_lr_schedule', 'check_optimizers', 'check_iterators', 'check_batch_by_size',
          'check_bert_loss_and_grads', 'check_all_heads', 'check_meters_and_format', 'check_flag_defaults']
_cache = {}


def _results():
    if 'r' not in _cache:
        if not os.path.exists(os.path.join(ROOT, 'baseline', '_ref', 'controller.py')):
            _cache['r'] = None
        else:
            env = dict(os.environ, CUDA_VISIBLE_DEVICES='')
            p = subprocess.run([sys.executable, os.path.join(ROOT, 'tests', 'ref_parity_worker.py')], capture_output=True,
                               text=True, timeout=900, env=env, cwd=ROOT)
            line = [l for l in p.stdout.splitlines() if l.startswith('RESULT ')]
            assert line, 'worker produced no result:\n' + p.stdout[-2000:] + p.stderr[-2000:]
            _cache['r'] = json.loads(line[-1][len('RESULT '):])
    return _cache['r']


@pytest.mark.parametrize('check', CHECKS)
def test_against_reference(check):
    res = _results()
    if res is None:
        pytest.skip('baseline/_ref not installed')
    assert res.get(check) == 'ok', res.get(check)
