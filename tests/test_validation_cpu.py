"""``--enable-validation`` through the command line on CPU: the validation subset is scored after every epoch, the
score drives ``checkpoint_best.pt`` and the LR scheduler's ``step(epoch, val_loss)``. In the reference the whole
validation loop is commented out (/root/reference/train.py:96-100 passes ``valid_losses = [None]``), so
``checkpoint_best.pt`` is never written there; the flags (--valid-subset, --validate-interval, --disable-validation,
--max-sentences-valid) are its own."""
import json
import os
import re
import subprocess
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CFG = dict(vocab_size=80, hidden_size=32, num_hidden_layers=1, num_attention_heads=4, intermediate_size=64,
           hidden_act='gelu', hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, max_position_embeddings=16,
           type_vocab_size=2, initializer_range=0.02)


def _run(tmp, extra):
    cfg = os.path.join(tmp, 'cfg.json')
    json.dump(CFG, open(cfg, 'w'))
    save = os.path.join(tmp, 'ckpt')
    cmd = [sys.executable, '-m', 'hetseq_b200.train', '--task', 'bert', '--optimizer', 'adam', '--lr-scheduler',
           'PolynomialDecayScheduler', '--synthetic-data', '--synthetic-seq-len', '16', '--synthetic-num-samples', '64',
           '--synthetic-max-predictions', '3', '--config_file', cfg, '--max-sentences', '8', '--lr', '2e-3',
           '--max-epoch', '3', '--cpu', '--save-dir', save, '--log-format', 'simple', '--log-interval', '100',
           '--distributed-world-size', '1', '--valid-subset', 'valid'] + extra
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT, env=dict(os.environ, OMP_NUM_THREADS='2'))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    return r.stdout, save


def test_validation_scores_every_epoch_and_keeps_the_best_checkpoint(tmp_path):
    out, save = _run(str(tmp_path), ['--enable-validation'])
    lines = re.findall(r"\| epoch (\d+) \| valid on 'valid' subset \| loss ([0-9.]+) \| ppl ([0-9.]+) \| num_updates (\d+)", out)
    assert [int(l[0]) for l in lines] == [1, 2, 3]
    losses = [float(l[1]) for l in lines]
    assert all(0 < x < 20 for x in losses) and [int(l[3]) for l in lines] == [8, 16, 24]
    files = sorted(os.listdir(save))
    assert 'checkpoint_best.pt' in files and 'checkpoint_last.pt' in files and 'checkpoint3.pt' in files
    best = torch.load(os.path.join(save, 'checkpoint_best.pt'), map_location='cpu', weights_only=False)
    assert abs(best['extra_state']['best'] - min(losses)) < 5e-4      # the line prints three decimals
    # the best file is the epoch with the lowest validation loss
    assert best['extra_state']['train_iterator']['epoch'] == 1 + losses.index(min(losses))


def test_validation_interval_and_disable(tmp_path):
    out, save = _run(os.path.join(str(tmp_path)), ['--enable-validation', '--validate-interval', '2'])
    assert re.findall(r"\| epoch (\d+) \| valid on", out) == ['002']
    d2 = os.path.join(str(tmp_path), 'off')
    os.makedirs(d2)
    out, save = _run(d2, ['--enable-validation', '--disable-validation'])
    assert 'valid on' not in out and 'checkpoint_best.pt' not in os.listdir(save)
