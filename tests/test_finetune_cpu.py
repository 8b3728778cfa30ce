"""The fine-tuning task (``--task bert_token_classification``) end to end on CPU through the command line: .npz shards
-> batches -> BertForTokenClassification -> Adam updates -> checkpoint -> resume. The reference registers a
``BertForTokenClassificationTask`` (/root/reference/tasks/tasks.py:213-260, bert_modeling.py:1149-1220) whose dataset
loader was left unfinished; here the same flags drive a working one."""
import json
import os
import subprocess
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CFG = dict(vocab_size=60, hidden_size=32, num_hidden_layers=2, num_attention_heads=4, intermediate_size=64,
           hidden_act='gelu', hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, max_position_embeddings=16,
           type_vocab_size=2, initializer_range=0.02)


def _shards(d, n=96, S=12, labels=3, seed=0):
    """A learnable toy problem: the label of a token is its id modulo ``labels``."""
    g = np.random.default_rng(seed)
    for split, m in (('train', n), ('valid', 24)):
        ids = g.integers(1, CFG['vocab_size'], size=(m, S)).astype(np.int64)
        mask = np.ones((m, S), dtype=np.int64)
        mask[:, S - 2:] = 0
        np.savez(os.path.join(d, '%s_0.npz' % split), input_ids=ids, segment_ids=np.zeros_like(ids), input_mask=mask,
                 labels=ids % labels)


def _run(data, cfg, save, max_update, extra=()):
    cmd = [sys.executable, '-m', 'hetseq_b200.train', '--task', 'bert_token_classification', '--optimizer', 'adam',
           '--lr-scheduler', 'PolynomialDecayScheduler', '--data', data, '--config_file', cfg, '--num_label', '3',
           '--max-sentences', '8', '--lr', '5e-3', '--max-update', str(max_update), '--max-epoch', '50', '--cpu',
           '--save-dir', save, '--log-format', 'json', '--log-interval', '1', '--distributed-world-size', '1',
           '--seed', '3'] + list(extra)
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT, env=dict(os.environ, OMP_NUM_THREADS='2'))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    rows = [json.loads(l) for l in r.stdout.splitlines() if l.startswith('{') and '"loss"' in l]
    return r.stdout, rows


def test_token_classification_trains_saves_and_resumes(tmp_path):
    d = str(tmp_path)
    _shards(d)
    cfg = os.path.join(d, 'cfg.json')
    json.dump(CFG, open(cfg, 'w'))
    save = os.path.join(d, 'ckpt')
    out, rows = _run(d, cfg, save, 36)
    assert '| done training' in out and len(rows) >= 20
    first, last = np.mean([r['loss'] for r in rows[:3]]), np.mean([r['loss'] for r in rows[-3:]])
    assert last < 0.7 * first, (first, last)              # the toy labelling is learnable
    state = torch.load(os.path.join(save, 'checkpoint_last.pt'), map_location='cpu', weights_only=False)
    assert state['optimizer_history'][-1]['num_updates'] == 36
    assert any(k.startswith('head.classifier.') for k in state['model'])
    # resume: the update counter and the position in the data continue, nothing starts over
    out2, rows2 = _run(d, cfg, save, 40)
    assert '| loaded checkpoint' in out2 and '@ 36 updates' in out2
    assert rows2 and rows2[0]['num_updates'] > 36 and rows2[-1]['num_updates'] <= 41
    assert np.mean([r['loss'] for r in rows2]) < 0.8 * first
