"""Pretrained-archive plumbing: download cache (file_utils), from_pretrained from a directory / tar.gz / URL,
old gamma/beta key names, TensorFlow checkpoint name mapping (reference: file_utils.py, bert_modeling.py:43-101,
584-733). No network: URLs are file:// and take the same path through the cache after the opener."""
import json
import os
import tarfile

import numpy as np
import pytest
import torch

from hetseq_b200.models.bert import (BertConfig, BertForPreTraining, BertForSequenceClassification, BertModel,
                                      load_tf_weights_in_bert, tf_variable_to_key)
from hetseq_b200.utils import file_utils

CFG = dict(vocab_size=64, hidden_size=32, num_hidden_layers=2, num_attention_heads=2, intermediate_size=64,
           hidden_act='gelu', hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0,
           max_position_embeddings=16, type_vocab_size=2, initializer_range=0.02)


def _save_model_dir(path, legacy_ln_names=False):
    os.makedirs(path, exist_ok=True)
    with open(os.path.join(path, file_utils.CONFIG_NAME), 'w') as f:
        json.dump(CFG, f)
    torch.manual_seed(0)
    model = BertForPreTraining(BertConfig.from_dict(CFG))
    sd = model.state_dict()
    if legacy_ln_names:
        sd = {k.replace('LayerNorm.weight', 'LayerNorm.gamma').replace('LayerNorm.bias', 'LayerNorm.beta'): v
              for k, v in sd.items()}
    torch.save(sd, os.path.join(path, file_utils.WEIGHTS_NAME))
    return model


def _same(a, b):
    sa, sb = a.state_dict(), b.state_dict()
    return all(torch.equal(sa[k], sb[k]) for k in sa if k in sb)


def test_url_to_filename_and_metadata_roundtrip(tmp_path):
    n1 = file_utils.url_to_filename('https://example.org/a.tar.gz')
    n2 = file_utils.url_to_filename('https://example.org/a.tar.gz', etag='"abc"')
    assert len(n1) == 64 and n2.startswith(n1 + '.') and len(n2) == 129
    src = tmp_path / 'payload.bin'
    src.write_bytes(b'x' * 1000)
    url = 'file://' + str(src)
    cache = str(tmp_path / 'cache')
    p = file_utils.cached_path(url, cache_dir=cache)
    assert open(p, 'rb').read() == b'x' * 1000
    assert file_utils.filename_to_url(os.path.basename(p), cache_dir=cache)[0] == url
    assert file_utils.cached_path(url, cache_dir=cache) == p            # second call: cache hit, same entry
    assert not [f for f in os.listdir(cache) if f.startswith('.download-')]
    assert file_utils.cached_path(str(src)) == str(src)                  # plain paths pass through
    with pytest.raises(EnvironmentError):
        file_utils.cached_path(str(tmp_path / 'missing'))
    with pytest.raises(EnvironmentError):
        file_utils.filename_to_url('nope', cache_dir=cache)
    assert file_utils.split_s3_path('s3://bucket/some/key.bin') == ('bucket', 'some/key.bin')
    assert file_utils.get_file_extension('/a/B.TXT') == '.txt' and file_utils.get_file_extension('a.Tar', dot=False) == 'tar'
    (tmp_path / 'set.txt').write_text('a\nb\na\n')
    assert file_utils.read_set_from_file(str(tmp_path / 'set.txt')) == {'a', 'b'}


def test_from_pretrained_directory_archive_and_url(tmp_path):
    d = str(tmp_path / 'model')
    ref = _save_model_dir(d)
    m1 = BertForPreTraining.from_pretrained(d)
    assert _same(m1, ref)
    # the bare trunk loads the 'bert.'-prefixed keys of a pre-training checkpoint
    trunk = BertModel.from_pretrained(d)
    assert torch.equal(trunk.embeddings.word_embeddings.weight, ref.bert.embeddings.word_embeddings.weight)
    # a fine-tuning head takes the trunk and reports its own head as not initialised
    clf = BertForSequenceClassification.from_pretrained(d, num_labels=3)
    assert torch.equal(clf.bert.pooler.dense_act.weight, ref.bert.pooler.dense_act.weight)
    # tar.gz with a top-level directory, through a file:// URL and the cache
    tgz = str(tmp_path / 'model.tar.gz')
    with tarfile.open(tgz, 'w:gz') as t:
        t.add(d, arcname='bert-tiny')
    cache = str(tmp_path / 'cache')
    m2 = BertForPreTraining.from_pretrained('file://' + tgz, cache_dir=cache)
    assert _same(m2, ref)
    m3 = BertForPreTraining.from_pretrained(tgz)            # local archive, unpacked next to it once
    assert _same(m3, ref) and os.path.isdir(tgz + '.extracted')


def test_from_pretrained_renames_gamma_beta(tmp_path):
    d = str(tmp_path / 'legacy')
    ref = _save_model_dir(d, legacy_ln_names=True)
    m = BertForPreTraining.from_pretrained(d)
    assert _same(m, ref)


def test_tf_checkpoint_mapping_and_load():
    assert tf_variable_to_key('bert/encoder/layer_1/attention/self/key/kernel') == \
        ('bert.encoder.layer.1.attention.self.key.weight', True)
    assert tf_variable_to_key('bert/encoder/layer_0/intermediate/dense/bias') == \
        ('bert.encoder.layer.0.intermediate.dense_act.bias', False)
    assert tf_variable_to_key('cls/predictions/output_bias') == ('cls.predictions.bias', False)
    assert tf_variable_to_key('cls/seq_relationship/output_weights') == ('cls.seq_relationship.weight', False)
    assert tf_variable_to_key('bert/pooler/dense/kernel/adam_v') is None
    torch.manual_seed(1)
    src = BertForPreTraining(BertConfig.from_dict(CFG))
    sd = src.state_dict()
    # build the TensorFlow-side view of the same weights: inverse mapping of every key
    inv = []
    for key, t in sd.items():
        if key == 'cls.predictions.decoder.weight':
            continue        # tied to the word embeddings; TF stores it once
        name = key.replace('.dense_act.', '.dense.')
        parts = name.split('.')
        leaf = parts[-1]
        scope = []
        i = 0
        body = parts[:-1]
        while i < len(body):
            if i + 1 < len(body) and body[i + 1].isdigit():
                scope.append(body[i] + '_' + body[i + 1]); i += 2
            else:
                scope.append(body[i]); i += 1
        arr = t.numpy()
        if scope[-1].endswith('_embeddings'):
            tf_name = '/'.join(scope)
        elif key == 'cls.predictions.bias':
            tf_name = 'cls/predictions/output_bias'
        elif key.startswith('cls.seq_relationship'):
            tf_name = 'cls/seq_relationship/' + ('output_weights' if leaf == 'weight' else 'output_bias')
        elif 'LayerNorm' in key:
            tf_name = '/'.join(scope + ['gamma' if leaf == 'weight' else 'beta'])
        elif leaf == 'weight':
            tf_name, arr = '/'.join(scope + ['kernel']), arr.T
        else:
            tf_name = '/'.join(scope + ['bias'])
        inv.append((tf_name, np.ascontiguousarray(arr)))
        inv.append((tf_name + '/adam_m', np.zeros_like(arr)))     # optimizer slots must be skipped
    torch.manual_seed(2)
    dst = BertForPreTraining(BertConfig.from_dict(CFG))
    load_tf_weights_in_bert(dst, 'unused', reader=inv)
    assert _same(dst, src)
    with pytest.raises(KeyError):
        load_tf_weights_in_bert(dst, 'unused', reader=[('bert/encoder/layer_9/output/dense/kernel', np.zeros((2, 2)))])
