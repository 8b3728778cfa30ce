"""BERT (reference: bert_modeling.py, NVIDIA DeepLearningExamples lineage).

The module tree exists to give parameters the reference's state-dict names
(``bert.encoder.layer.N.attention.self.query.weight`` ... ``cls.predictions.decoder.weight``,
SURVEY.md section 2.2; 206 tensors / 110,106,428 params for BERT-base, 398 / 336,226,108 for
BERT-large) -- checkpoints are interchangeable. The computation has two providers:

* the ``forward`` methods of the modules in this file: plain PyTorch, fp32-exact restatement of the
  reference math -- GELU with the constant 1.41421 (bert_modeling.py:104-106), TF-style LayerNorm with
  eps=1e-12 inside the sqrt (:285-289), additive -10000 attention mask (:798-806), scores / sqrt(d)
  (:361-362), mean CE with ignore_index=-1 (:881-886). It is the CPU path, the ``--kernels torch`` path
  and the oracle for the GPU tests.
* ``hetseq_b200/ops/bert_native.py`` (attached to the model by ``attach()`` on CUDA devices with 16-bit
  parameters, ``--kernels auto|native``): the sm_100a kernels -- one autograd node per encoder layer
  (packed QKV tcgen05 GEMM, fused attention, GEMM epilogues with bias/GELU, fused
  dropout+residual+LayerNorm), gathered MLM head + fused cross-entropy. ``BertLayer.forward``,
  ``BertModel.forward`` and ``BertForPreTraining.forward`` hand over to it when it is attached.

Deliberate deviations from the reference, all loss-preserving:
* the MLM decoder and its cross-entropy run on the labelled rows only (<= 15 % of positions) instead
  of all B*S (bert_modeling.py:878-882): ignored rows contribute 0 to a mean over non-ignored rows
  (SURVEY.md Q21); ``--mlm-all-positions`` restores the reference shape of the computation;
* no ``boto3`` / ``file_utils`` import at module import time -- ``from_pretrained``
  imports the download cache lazily;
* ``BertLayerNorm`` is always defined (the reference loses it when apex is importable, Q5);
* activation checkpointing is reachable (``--checkpoint-activations``); the reference's copy is dead code.
"""
from __future__ import annotations

import copy
import json
import math
import os
from typing import List, Optional

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.nn import CrossEntropyLoss
from torch.utils import checkpoint as torch_checkpoint

from hetseq_b200.utils.utils import gelu_ref, swish

ACT2FN = {'gelu': gelu_ref, 'relu': F.relu, 'swish': swish, 'tanh': torch.tanh}


def bias_gelu(bias, y):
    return gelu_ref(y + bias)


def bias_tanh(bias, y):
    return torch.tanh(y + bias)


class BertConfig(object):
    """Model hyper-parameters; JSON-compatible with the reference's (bert_modeling.py:180-266)."""

    def __init__(self, vocab_size_or_config_json_file=30522, hidden_size=768, num_hidden_layers=12,
                 num_attention_heads=12, intermediate_size=3072, hidden_act='gelu', hidden_dropout_prob=0.1,
                 attention_probs_dropout_prob=0.1, max_position_embeddings=512, type_vocab_size=2,
                 initializer_range=0.02):
        if isinstance(vocab_size_or_config_json_file, str):
            with open(vocab_size_or_config_json_file, 'r', encoding='utf-8') as reader:
                for key, value in json.loads(reader.read()).items():
                    self.__dict__[key] = value
        elif isinstance(vocab_size_or_config_json_file, int):
            self.vocab_size = vocab_size_or_config_json_file
            self.hidden_size = hidden_size
            self.num_hidden_layers = num_hidden_layers
            self.num_attention_heads = num_attention_heads
            self.hidden_act = hidden_act
            self.intermediate_size = intermediate_size
            self.hidden_dropout_prob = hidden_dropout_prob
            self.attention_probs_dropout_prob = attention_probs_dropout_prob
            self.max_position_embeddings = max_position_embeddings
            self.type_vocab_size = type_vocab_size
            self.initializer_range = initializer_range
        else:
            raise ValueError('First argument must be either a vocabulary size (int) or the path to a '
                             'pretrained model config file (str)')

    @classmethod
    def from_dict(cls, json_object):
        config = BertConfig(vocab_size_or_config_json_file=-1)
        for key, value in json_object.items():
            config.__dict__[key] = value
        return config

    @classmethod
    def from_json_file(cls, json_file):
        with open(json_file, 'r', encoding='utf-8') as reader:
            return cls.from_dict(json.loads(reader.read()))

    @classmethod
    def bert_base(cls, **kw):
        return cls(30522, 768, 12, 12, 3072, **kw)

    @classmethod
    def bert_large(cls, **kw):
        return cls(30522, 1024, 24, 16, 4096, **kw)

    def __repr__(self):
        return str(self.to_json_string())

    def to_dict(self):
        return copy.deepcopy(self.__dict__)

    def to_json_string(self):
        return json.dumps(self.to_dict(), indent=2, sort_keys=True) + '\n'


# ------------------------------------------------------------------------------------------------
# parameter holders (names = state-dict keys of the reference)
# ------------------------------------------------------------------------------------------------
class LinearActivation(nn.Module):
    """Linear whose bias add is fused with the activation (bert_modeling.py:132-177)."""

    def __init__(self, in_features, out_features, act='gelu', bias=True):
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        self.act_name = act if isinstance(act, str) else getattr(act, '__name__', 'custom')
        self.act_fn = ACT2FN[act] if isinstance(act, str) else act
        self.weight = nn.Parameter(torch.empty(out_features, in_features))
        self.bias = nn.Parameter(torch.empty(out_features)) if bias else None
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        if bias:
            bound = 1 / math.sqrt(in_features)
            nn.init.uniform_(self.bias, -bound, bound)

    def forward(self, x):
        return self.act_fn(F.linear(x, self.weight, self.bias))

    def extra_repr(self):
        return 'in_features={}, out_features={}, act={}'.format(self.in_features, self.out_features, self.act_name)


class BertLayerNorm(nn.Module):
    """TF-style LayerNorm, epsilon inside the square root (bert_modeling.py:277-289)."""

    def __init__(self, hidden_size, eps=1e-12):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(hidden_size))
        self.bias = nn.Parameter(torch.zeros(hidden_size))
        self.variance_epsilon = eps

    def forward(self, x):
        xf = x.float()
        u = xf.mean(-1, keepdim=True)
        s = (xf - u).pow(2).mean(-1, keepdim=True)
        y = (xf - u) / torch.sqrt(s + self.variance_epsilon)
        return (self.weight.float() * y + self.bias.float()).to(x.dtype)


class BertEmbeddings(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.word_embeddings = nn.Embedding(config.vocab_size, config.hidden_size)
        self.position_embeddings = nn.Embedding(config.max_position_embeddings, config.hidden_size)
        self.token_type_embeddings = nn.Embedding(config.type_vocab_size, config.hidden_size)
        self.LayerNorm = BertLayerNorm(config.hidden_size, eps=1e-12)
        self.dropout = nn.Dropout(config.hidden_dropout_prob)

    def forward(self, input_ids, token_type_ids):
        seq_length = input_ids.size(1)
        position_ids = torch.arange(seq_length, dtype=torch.long, device=input_ids.device)
        position_ids = position_ids.unsqueeze(0).expand_as(input_ids)
        e = self.word_embeddings(input_ids.long()) + self.position_embeddings(position_ids) \
            + self.token_type_embeddings(token_type_ids.long())
        return self.dropout(self.LayerNorm(e))


class BertSelfAttention(nn.Module):
    def __init__(self, config):
        super().__init__()
        if config.hidden_size % config.num_attention_heads != 0:
            raise ValueError('The hidden size (%d) is not a multiple of the number of attention heads (%d)'
                             % (config.hidden_size, config.num_attention_heads))
        self.num_attention_heads = config.num_attention_heads
        self.attention_head_size = config.hidden_size // config.num_attention_heads
        self.all_head_size = self.num_attention_heads * self.attention_head_size
        self.query = nn.Linear(config.hidden_size, self.all_head_size)
        self.key = nn.Linear(config.hidden_size, self.all_head_size)
        self.value = nn.Linear(config.hidden_size, self.all_head_size)
        self.dropout = nn.Dropout(config.attention_probs_dropout_prob)

    def forward(self, hidden_states, attention_mask):
        B, S, _ = hidden_states.shape
        h, d = self.num_attention_heads, self.attention_head_size
        q = self.query(hidden_states).view(B, S, h, d).permute(0, 2, 1, 3)
        k = self.key(hidden_states).view(B, S, h, d).permute(0, 2, 3, 1)
        v = self.value(hidden_states).view(B, S, h, d).permute(0, 2, 1, 3)
        scores = torch.matmul(q, k) / math.sqrt(d)
        scores = scores + attention_mask
        probs = self.dropout(F.softmax(scores, dim=-1))
        ctx = torch.matmul(probs, v).permute(0, 2, 1, 3).contiguous()
        return ctx.view(B, S, self.all_head_size)


class BertSelfOutput(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.dense = nn.Linear(config.hidden_size, config.hidden_size)
        self.LayerNorm = BertLayerNorm(config.hidden_size, eps=1e-12)
        self.dropout = nn.Dropout(config.hidden_dropout_prob)

    def forward(self, hidden_states, input_tensor):
        return self.LayerNorm(self.dropout(self.dense(hidden_states)) + input_tensor)


class BertAttention(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.self = BertSelfAttention(config)
        self.output = BertSelfOutput(config)

    def forward(self, input_tensor, attention_mask):
        return self.output(self.self(input_tensor, attention_mask), input_tensor)


class BertIntermediate(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.dense_act = LinearActivation(config.hidden_size, config.intermediate_size, act=config.hidden_act)

    def forward(self, hidden_states):
        return self.dense_act(hidden_states)


class BertOutput(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.dense = nn.Linear(config.intermediate_size, config.hidden_size)
        self.LayerNorm = BertLayerNorm(config.hidden_size, eps=1e-12)
        self.dropout = nn.Dropout(config.hidden_dropout_prob)

    def forward(self, hidden_states, input_tensor):
        return self.LayerNorm(self.dropout(self.dense(hidden_states)) + input_tensor)


class BertLayer(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.attention = BertAttention(config)
        self.intermediate = BertIntermediate(config)
        self.output = BertOutput(config)
        self.native = None  # set by ops.bert_native.attach()

    def forward(self, hidden_states, attention_mask):
        if self.native is not None:
            return self.native.layer_forward(self, hidden_states, attention_mask)
        attention_output = self.attention(hidden_states, attention_mask)
        return self.output(self.intermediate(attention_output), attention_output)


class BertEncoder(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.layer = nn.ModuleList([BertLayer(config) for _ in range(config.num_hidden_layers)])

    def forward(self, hidden_states, attention_mask, output_all_encoded_layers=True, checkpoint_activations=False):
        all_encoder_layers = []
        if checkpoint_activations:
            # recompute in chunks of ceil(sqrt(L)) layers (bert_modeling.py:470-476)
            L = len(self.layer)
            chunk = int(math.ceil(math.sqrt(L)))

            def run(start, end):
                def fn(x, m):
                    for layer in self.layer[start:end]:
                        x = layer(x, m)
                    return x
                return fn

            l = 0
            while l < L:
                hidden_states = torch_checkpoint.checkpoint(run(l, min(L, l + chunk)), hidden_states,
                                                            attention_mask, use_reentrant=False)
                l += chunk
        else:
            for layer_module in self.layer:
                hidden_states = layer_module(hidden_states, attention_mask)
                if output_all_encoded_layers:
                    all_encoder_layers.append(hidden_states)
        if not output_all_encoded_layers or checkpoint_activations:
            all_encoder_layers.append(hidden_states)
        return all_encoder_layers


class BertPooler(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.dense_act = LinearActivation(config.hidden_size, config.hidden_size, act='tanh')

    def forward(self, hidden_states):
        return self.dense_act(hidden_states[:, 0])


class BertPredictionHeadTransform(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.dense_act = LinearActivation(config.hidden_size, config.hidden_size, act=config.hidden_act)
        self.LayerNorm = BertLayerNorm(config.hidden_size, eps=1e-12)

    def forward(self, hidden_states):
        return self.LayerNorm(self.dense_act(hidden_states))


class BertLMPredictionHead(nn.Module):
    """transform -> decoder tied to the word embeddings + separate bias (bert_modeling.py:531-549)."""

    def __init__(self, config, bert_model_embedding_weights):
        super().__init__()
        self.transform = BertPredictionHeadTransform(config)
        self.decoder = nn.Linear(bert_model_embedding_weights.size(1), bert_model_embedding_weights.size(0),
                                 bias=False)
        self.decoder.weight = bert_model_embedding_weights
        self.bias = nn.Parameter(torch.zeros(bert_model_embedding_weights.size(0)))

    def forward(self, hidden_states):
        return self.decoder(self.transform(hidden_states)) + self.bias


class BertOnlyMLMHead(nn.Module):
    def __init__(self, config, bert_model_embedding_weights):
        super().__init__()
        self.predictions = BertLMPredictionHead(config, bert_model_embedding_weights)

    def forward(self, sequence_output):
        return self.predictions(sequence_output)


class BertOnlyNSPHead(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.seq_relationship = nn.Linear(config.hidden_size, 2)

    def forward(self, pooled_output):
        return self.seq_relationship(pooled_output)


class BertPreTrainingHeads(nn.Module):
    def __init__(self, config, bert_model_embedding_weights):
        super().__init__()
        self.predictions = BertLMPredictionHead(config, bert_model_embedding_weights)
        self.seq_relationship = nn.Linear(config.hidden_size, 2)

    def forward(self, sequence_output, pooled_output):
        return self.predictions(sequence_output), self.seq_relationship(pooled_output)


class BertPreTrainedModel(nn.Module):
    """Weight initialisation + pretrained-archive loading (bert_modeling.py:584-733)."""

    def __init__(self, config, *inputs, **kwargs):
        super().__init__()
        if not isinstance(config, BertConfig):
            raise ValueError('Parameter config in `{}(config)` should be an instance of class `BertConfig`.'.format(
                self.__class__.__name__))
        self.config = config

    def init_bert_weights(self, module):
        """Linear / Embedding / LinearActivation weights ~ N(0, initializer_range); LayerNorm (1, 0);
        biases 0 (bert_modeling.py:599-610). In the reference ``LinearActivation`` (FFN1, pooler, MLM
        transform) is not an ``nn.Linear``, so those layers silently keep their constructor's kaiming-uniform
        weights and uniform biases; normal init for them too is the evident intent and the default here.
        ``config.linear_activation_init = 'reference'`` (``--init-linear-activation reference``) reproduces the
        reference's from-scratch initialisation exactly."""
        like_reference = getattr(self.config, 'linear_activation_init', 'normal') == 'reference'
        if isinstance(module, LinearActivation) and like_reference:
            return
        if isinstance(module, (nn.Linear, nn.Embedding, LinearActivation)):
            module.weight.data.normal_(mean=0.0, std=self.config.initializer_range)
        elif isinstance(module, BertLayerNorm):
            module.bias.data.zero_()
            module.weight.data.fill_(1.0)
        if isinstance(module, (nn.Linear, LinearActivation)) and module.bias is not None:
            module.bias.data.zero_()

    @classmethod
    def from_pretrained(cls, pretrained_model_name_or_path, state_dict=None, cache_dir=None, from_tf=False,
                        *inputs, **kwargs):
        """Instantiate from a model name (PRETRAINED_MODEL_ARCHIVE_MAP), a directory or ``.tar.gz`` archive
        holding ``bert_config.json`` + ``pytorch_model.bin``, or a URL resolved through the download cache
        (hetseq_b200/utils/file_utils.py). ``from_tf`` reads a TensorFlow ``model.ckpt`` instead
        (bert_modeling.py:43-101). Old ``gamma``/``beta`` LayerNorm key names are renamed like
        bert_modeling.py:690-701."""
        from hetseq_b200.utils import file_utils
        resolved = file_utils.resolve_archive(pretrained_model_name_or_path, cache_dir=cache_dir,
                                              archive_map=PRETRAINED_MODEL_ARCHIVE_MAP)
        config = BertConfig.from_json_file(os.path.join(resolved, file_utils.CONFIG_NAME))
        model = cls(config, *inputs, **kwargs)
        if from_tf:
            return load_tf_weights_in_bert(model, os.path.join(resolved, file_utils.TF_WEIGHTS_NAME))
        if state_dict is None:
            state_dict = torch.load(os.path.join(resolved, 'pytorch_model.bin'), map_location='cpu',
                                    weights_only=True)
        renamed = {}
        for key, value in state_dict.items():
            new_key = key.replace('gamma', 'weight').replace('beta', 'bias')
            renamed[new_key] = value
        own = model.state_dict()
        # a checkpoint of BertForPreTraining has 'bert.'-prefixed trunk keys, a bare BertModel checkpoint has
        # none; either must load into either kind of model (bert_modeling.py:703-723 does this with a prefix walk)
        loadable = {}
        for key, value in renamed.items():
            for cand in (key, 'bert.' + key, key[5:] if key.startswith('bert.') else None):
                if cand is not None and cand in own and tuple(own[cand].shape) == tuple(value.shape):
                    loadable[cand] = value
                    break
        missing = [k for k in own if k not in loadable]
        model.load_state_dict(loadable, strict=False)
        if missing:
            print('| weights of {} not initialized from pretrained model: {}'.format(cls.__name__, missing))
        return model


# short names accepted by from_pretrained (bert_modeling.py:33-41); resolved through the download cache
_HF_S3 = 'https://s3.amazonaws.com/models.huggingface.co/bert/'
PRETRAINED_MODEL_ARCHIVE_MAP = {name: _HF_S3 + name + '.tar.gz' for name in (
    'bert-base-uncased', 'bert-large-uncased', 'bert-base-cased', 'bert-large-cased',
    'bert-base-multilingual-uncased', 'bert-base-multilingual-cased', 'bert-base-chinese')}

_TF_LEAF = {'kernel': 'weight', 'gamma': 'weight', 'beta': 'bias', 'bias': 'bias', 'output_bias': 'bias',
            'output_weights': 'weight'}
# TF scope names whose module has another name in this model tree (LinearActivation modules)
_TF_RENAMES = (('intermediate/dense/', 'intermediate/dense_act/'), ('pooler/dense/', 'pooler/dense_act/'),
               ('predictions/transform/dense/', 'predictions/transform/dense_act/'))


def tf_variable_to_key(name: str):
    """Google-BERT TensorFlow variable name -> (state-dict key, transpose?) or None for optimizer slots.

    ``bert/encoder/layer_3/attention/self/query/kernel`` -> (``bert.encoder.layer.3.attention.self.query.weight``,
    True): dense kernels are stored [in, out] in TF and [out, in] here (bert_modeling.py:66-101 walks the
    module tree attribute by attribute; this is the same mapping as a pure function, so it can be tested
    without TensorFlow)."""
    parts = name.split('/')
    if any(p in ('adam_v', 'adam_m', 'global_step') or p.startswith('AdamWeightDecayOptimizer') for p in parts):
        return None
    name = name + '/'
    for old, new in _TF_RENAMES:
        name = name.replace(old, new)
    parts = name.rstrip('/').split('/')
    leaf = parts[-1]
    out = []
    for p in parts[:-1]:
        head, sep, num = p.rpartition('_')
        if sep and num.isdigit() and head.isalpha():
            out += [head, num]          # layer_7 -> layer.7
        else:
            out.append(p)
    if leaf.endswith('_embeddings'):
        return '.'.join(out + [leaf, 'weight']), False
    if parts[-2:] == ['predictions', 'output_bias']:
        return '.'.join(out + ['bias']), False
    if leaf not in _TF_LEAF:
        raise KeyError('unrecognised TensorFlow variable {}'.format(name.rstrip('/')))
    return '.'.join(out + [_TF_LEAF[leaf]]), leaf == 'kernel'


def load_tf_weights_in_bert(model, tf_checkpoint_path, reader=None):
    """Copy a TensorFlow BERT checkpoint into ``model`` (bert_modeling.py:43-101). ``reader`` (tests) is an
    iterable of (variable name, numpy array); by default the checkpoint is read with TensorFlow, which is
    imported only here."""
    import numpy as np
    if reader is None:
        try:
            import tensorflow as tf
        except ImportError:
            print('Loading a TensorFlow model in PyTorch requires TensorFlow to be installed.')
            raise
        path = os.path.abspath(tf_checkpoint_path)
        reader = ((n, tf.train.load_variable(path, n)) for n, _ in tf.train.list_variables(path))
    own = model.state_dict()
    loaded = set()
    for name, array in reader:
        mapped = tf_variable_to_key(name)
        if mapped is None:
            continue
        key, transpose = mapped
        if key not in own and hasattr(model, 'bert') and not key.startswith(('bert.', 'cls.')):
            key = 'bert.' + key
        if key not in own:
            raise KeyError('TensorFlow variable {} maps to {}, which this model does not have'.format(name, key))
        array = np.asarray(array)
        if transpose:
            array = array.T
        if tuple(own[key].shape) != tuple(array.shape):
            raise ValueError('shape mismatch for {}: checkpoint {} vs model {}'.format(key, array.shape,
                                                                                      tuple(own[key].shape)))
        with torch.no_grad():
            own[key].copy_(torch.from_numpy(np.ascontiguousarray(array)))
        loaded.add(key)
    return model


def extended_attention_mask(attention_mask, dtype):
    """[B,S] 0/1 -> [B,1,1,S] additive mask, (1-mask)*-10000 (bert_modeling.py:798-806)."""
    m = attention_mask.unsqueeze(1).unsqueeze(2).to(dtype=dtype)
    return (1.0 - m) * -10000.0


class BertModel(BertPreTrainedModel):
    def __init__(self, config):
        super().__init__(config)
        self.embeddings = BertEmbeddings(config)
        self.encoder = BertEncoder(config)
        self.pooler = BertPooler(config)
        self.apply(self.init_bert_weights)
        self.native = None  # set by ops.bert_native.attach()

    def forward(self, input_ids, token_type_ids=None, attention_mask=None, output_all_encoded_layers=True,
                checkpoint_activations=False):
        if attention_mask is None:
            attention_mask = torch.ones_like(input_ids)
        if token_type_ids is None:
            token_type_ids = torch.zeros_like(input_ids)
        if self.native is not None:
            return self.native.model_forward(self, input_ids, token_type_ids, attention_mask,
                                             output_all_encoded_layers, checkpoint_activations)
        dtype = next(self.parameters()).dtype
        ext_mask = extended_attention_mask(attention_mask, dtype)
        embedding_output = self.embeddings(input_ids, token_type_ids)
        encoded_layers = self.encoder(embedding_output, ext_mask,
                                      output_all_encoded_layers=output_all_encoded_layers,
                                      checkpoint_activations=checkpoint_activations)
        sequence_output = encoded_layers[-1]
        pooled_output = self.pooler(sequence_output)
        if not output_all_encoded_layers:
            encoded_layers = encoded_layers[-1]
        return encoded_layers, pooled_output


class BertForPreTraining(BertPreTrainedModel):
    """MLM + NSP pre-training (bert_modeling.py:819-888). ``forward(*sample)`` with the five dataset
    tensors returns the scalar loss = mean CE over labelled MLM positions + mean CE over NSP labels."""

    def __init__(self, config, args=None):
        super().__init__(config)
        self.bert = BertModel(config)
        self.cls = BertPreTrainingHeads(config, self.bert.embeddings.word_embeddings.weight)
        self.apply(self.init_bert_weights)
        self.checkpoint_activations = bool(getattr(args, 'checkpoint_activations', False))
        self.mlm_all_positions = bool(getattr(args, 'mlm_all_positions', False))
        self.native = None  # set by ops.bert_native.attach()

    def forward(self, input_ids, token_type_ids=None, attention_mask=None, masked_lm_labels=None,
                next_sentence_label=None, mlm_rows=None, mlm_row_labels=None, emb_order=None, emb_run_end=None,
                checkpoint_activations=None):
        """``mlm_rows`` / ``mlm_row_labels`` (optional, fixed capacity, padded with -1): flat indices of the
        labelled positions and their labels, precomputed on the host by the task. They remove the
        data-dependent ``nonzero`` (a device->host sync) from the step and give the MLM head a static shape,
        which is what lets the whole step be captured in a CUDA graph. ``emb_order`` / ``emb_run_end`` (optional):
        positions sorted by token id and the run boundaries, for the deterministic scatter of the word-embedding
        gradient (native path; ignored by the PyTorch path)."""
        ckpt = self.checkpoint_activations if checkpoint_activations is None else checkpoint_activations
        if self.native is not None and masked_lm_labels is not None and next_sentence_label is not None:
            return self.native.pretraining_loss(self, input_ids, token_type_ids, attention_mask,
                                                masked_lm_labels, next_sentence_label, ckpt,
                                                mlm_rows=mlm_rows, mlm_row_labels=mlm_row_labels,
                                                emb_plan=None if emb_order is None or emb_run_end is None
                                                else (emb_order, emb_run_end))
        sequence_output, pooled_output = self.bert(input_ids, token_type_ids, attention_mask,
                                                   output_all_encoded_layers=False,
                                                   checkpoint_activations=ckpt)
        if masked_lm_labels is None or next_sentence_label is None:
            return self.cls(sequence_output, pooled_output)
        labels = masked_lm_labels.reshape(-1).long()
        V = self.config.vocab_size
        if self.mlm_all_positions:
            prediction_scores, seq_relationship_score = self.cls(sequence_output, pooled_output)
            masked_lm_loss = F.cross_entropy(prediction_scores.view(-1, V).float(), labels, ignore_index=-1)
        else:
            if mlm_rows is not None:
                rows, row_labels = mlm_rows.long().clamp(min=0), mlm_row_labels.long()
            else:
                rows = torch.nonzero(labels != -1, as_tuple=False).squeeze(1)
                row_labels = labels.index_select(0, rows)
            hidden = sequence_output.reshape(-1, sequence_output.size(-1)).index_select(0, rows)
            prediction_scores = self.cls.predictions(hidden)
            seq_relationship_score = self.cls.seq_relationship(pooled_output)
            masked_lm_loss = F.cross_entropy(prediction_scores.float(), row_labels,
                                             ignore_index=-1) if rows.numel() > 0 else prediction_scores.sum() * 0
        next_sentence_loss = F.cross_entropy(seq_relationship_score.view(-1, 2).float(),
                                             next_sentence_label.view(-1).long(), ignore_index=-1)
        return masked_lm_loss + next_sentence_loss

    def param_order(self) -> List[nn.Parameter]:
        """Flat-buffer order: per encoder layer the q/k/v weights (then biases) are adjacent so the
        packed [3H, H] projection is a *view* of the flat buffer -- no pack/unpack copies."""
        return bert_param_order(self)


def bert_param_order(model: nn.Module) -> List[nn.Parameter]:
    """model.parameters() order, except that each attention block's q/k/v weights and then its q/k/v
    biases are emitted together at the position of ``query.weight``."""
    groups, members = {}, set()
    for m in model.modules():
        if isinstance(m, BertSelfAttention):
            grp = [getattr(lin, name) for name in ('weight', 'bias') for lin in (m.query, m.key, m.value)]
            groups[id(m.query.weight)] = grp
            members.update(id(p) for p in grp)
    out, placed = [], set()
    for p in model.parameters():
        if id(p) in placed:
            continue
        if id(p) in groups:
            for q in groups[id(p)]:
                out.append(q)
                placed.add(id(q))
        elif id(p) not in members:
            out.append(p)
            placed.add(id(p))
    return out


class BertForMaskedLM(BertPreTrainedModel):
    def __init__(self, config):
        super().__init__(config)
        self.bert = BertModel(config)
        self.cls = BertOnlyMLMHead(config, self.bert.embeddings.word_embeddings.weight)
        self.apply(self.init_bert_weights)

    def forward(self, input_ids, token_type_ids=None, attention_mask=None, masked_lm_labels=None,
                checkpoint_activations=False):
        sequence_output, _ = self.bert(input_ids, token_type_ids, attention_mask, output_all_encoded_layers=False,
                                       checkpoint_activations=checkpoint_activations)
        prediction_scores = self.cls(sequence_output)
        if masked_lm_labels is not None:
            return CrossEntropyLoss(ignore_index=-1)(prediction_scores.view(-1, self.config.vocab_size).float(),
                                                     masked_lm_labels.view(-1).long())
        return prediction_scores


class BertForNextSentencePrediction(BertPreTrainedModel):
    def __init__(self, config):
        super().__init__(config)
        self.bert = BertModel(config)
        self.cls = BertOnlyNSPHead(config)
        self.apply(self.init_bert_weights)

    def forward(self, input_ids, token_type_ids=None, attention_mask=None, next_sentence_label=None,
                checkpoint_activations=False):
        _, pooled_output = self.bert(input_ids, token_type_ids, attention_mask, output_all_encoded_layers=False,
                                     checkpoint_activations=checkpoint_activations)
        seq_relationship_score = self.cls(pooled_output)
        if next_sentence_label is not None:
            return CrossEntropyLoss(ignore_index=-1)(seq_relationship_score.view(-1, 2).float(),
                                                     next_sentence_label.view(-1).long())
        return seq_relationship_score


class BertForSequenceClassification(BertPreTrainedModel):
    def __init__(self, config, num_labels):
        super().__init__(config)
        self.num_labels = num_labels
        self.bert = BertModel(config)
        self.dropout = nn.Dropout(config.hidden_dropout_prob)
        self.classifier = nn.Linear(config.hidden_size, num_labels)
        self.apply(self.init_bert_weights)

    def forward(self, input_ids, token_type_ids=None, attention_mask=None, labels=None, checkpoint_activations=False):
        _, pooled_output = self.bert(input_ids, token_type_ids, attention_mask, output_all_encoded_layers=False,
                                     checkpoint_activations=checkpoint_activations)
        logits = self.classifier(self.dropout(pooled_output))
        if labels is not None:
            return CrossEntropyLoss()(logits.view(-1, self.num_labels).float(), labels.view(-1).long())
        return logits


class BertForMultipleChoice(BertPreTrainedModel):
    def __init__(self, config, num_choices):
        super().__init__(config)
        self.num_choices = num_choices
        self.bert = BertModel(config)
        self.dropout = nn.Dropout(config.hidden_dropout_prob)
        self.classifier = nn.Linear(config.hidden_size, 1)
        self.apply(self.init_bert_weights)

    def forward(self, input_ids, token_type_ids=None, attention_mask=None, labels=None, checkpoint_activations=False):
        flat = lambda t: None if t is None else t.view(-1, t.size(-1))
        _, pooled_output = self.bert(flat(input_ids), flat(token_type_ids), flat(attention_mask),
                                     output_all_encoded_layers=False, checkpoint_activations=checkpoint_activations)
        logits = self.classifier(self.dropout(pooled_output)).view(-1, self.num_choices)
        if labels is not None:
            return CrossEntropyLoss()(logits.float(), labels.long())
        return logits


class BertForTokenClassification(BertPreTrainedModel):
    def __init__(self, config, num_labels):
        super().__init__(config)
        self.num_labels = num_labels
        self.bert = BertModel(config)
        self.dropout = nn.Dropout(config.hidden_dropout_prob)
        self.classifier = nn.Linear(config.hidden_size, num_labels)
        self.apply(self.init_bert_weights)

    def forward(self, input_ids, token_type_ids=None, attention_mask=None, labels=None, checkpoint_activations=False):
        sequence_output, _ = self.bert(input_ids, token_type_ids, attention_mask, output_all_encoded_layers=False,
                                       checkpoint_activations=checkpoint_activations)
        logits = self.classifier(self.dropout(sequence_output))
        if labels is not None:
            loss_fct = CrossEntropyLoss()
            if attention_mask is not None:
                active = attention_mask.view(-1) == 1
                return loss_fct(logits.view(-1, self.num_labels)[active].float(), labels.view(-1)[active].long())
            return loss_fct(logits.view(-1, self.num_labels).float(), labels.view(-1).long())
        return logits


class BertForQuestionAnswering(BertPreTrainedModel):
    def __init__(self, config):
        super().__init__(config)
        self.bert = BertModel(config)
        self.qa_outputs = nn.Linear(config.hidden_size, 2)
        self.apply(self.init_bert_weights)

    def forward(self, input_ids, token_type_ids=None, attention_mask=None, start_positions=None,
                end_positions=None, checkpoint_activations=False):
        sequence_output, _ = self.bert(input_ids, token_type_ids, attention_mask, output_all_encoded_layers=False,
                                       checkpoint_activations=checkpoint_activations)
        logits = self.qa_outputs(sequence_output)
        start_logits, end_logits = [t.squeeze(-1) for t in logits.split(1, dim=-1)]
        if start_positions is not None and end_positions is not None:
            if start_positions.dim() > 1:
                start_positions = start_positions.squeeze(-1)
            if end_positions.dim() > 1:
                end_positions = end_positions.squeeze(-1)
            ignored_index = start_logits.size(1)
            start_positions = start_positions.clamp(0, ignored_index)
            end_positions = end_positions.clamp(0, ignored_index)
            loss_fct = CrossEntropyLoss(ignore_index=ignored_index)
            return (loss_fct(start_logits.float(), start_positions.long())
                    + loss_fct(end_logits.float(), end_positions.long())) / 2
        return start_logits, end_logits
