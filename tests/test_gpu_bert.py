"""GPU: the native BERT path (kernels + autograd glue) against the PyTorch restatement, end to end."""
import copy
import json
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


def _cfg(tmp_path, **kw):
    cfg = dict(vocab_size=2000, hidden_size=256, num_hidden_layers=2, num_attention_heads=4, intermediate_size=1024,
               hidden_act='gelu', hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0,
               max_position_embeddings=128, type_vocab_size=2, initializer_range=0.02)
    cfg.update(kw)
    p = os.path.join(tmp_path, 'cfg.json')
    with open(p, 'w') as f:
        json.dump(cfg, f)
    return p


def _build(tmp_path, dtype_flag, kernels, seed=0, **cfgkw):
    from hetseq_b200 import options, registry
    from hetseq_b200.controller import Controller
    argv = ['--config_file', _cfg(tmp_path, **cfgkw), '--max-sentences', '8', '--synthetic-data', '--lr', '1e-3',
            '--no-save', '--log-format', 'none', '--distributed-world-size', '1', '--save-dir', str(tmp_path),
            '--kernels', kernels, '--clip-norm', '0']
    if dtype_flag:
        argv.append(dtype_flag)
    args = options.make_args('bert', 'adam', 'PolynomialDecayScheduler', argv)
    torch.manual_seed(seed)
    task = registry.lookup('task', 'bert').setup_task(args)
    model = task.build_model(args)
    ctl = Controller(args, task, model)
    ctl.optimizer
    return ctl, task


def _batch(vocab=2000, B=8, S=128):
    from hetseq_b200.data.bert_dataset import SyntheticBertData
    return SyntheticBertData(B, S, 10, vocab, seed=3).fetch_rows(list(range(B)))


@pytest.mark.parametrize('flag,S', [('--fp16', 128), ('--bf16', 128), ('--fp16', 256)])
def test_native_loss_and_grads_match_torch(tmp_path, flag, S):
    """Same 16-bit weights, dropout off: native loss/gradients vs an fp32 PyTorch model. S = 256 goes through the
    multi-tile attention kernels inside the fused encoder-layer node."""
    ctl, task = _build(str(tmp_path), flag, 'native', max_position_embeddings=max(128, S))
    assert ctl.get_model().native is not None, 'native kernels were not attached'
    from hetseq_b200.models.bert import BertConfig, BertForPreTraining
    ref = BertForPreTraining(BertConfig.from_json_file(os.path.join(str(tmp_path), 'cfg.json'))).cuda().float()
    ref.load_state_dict({k: v.float() for k, v in ctl.get_model().state_dict().items()})
    sample = _batch(S=S)
    dev = task.prepare_sample(sample, ctl.device)
    ctl.get_model().train()
    ctl.zero_grad()
    loss = ctl.get_model()(*dev)
    # fp16 needs a loss scale like in real training (the attention gradients are ~1e-6 and would sit in
    # fp16's subnormal range otherwise); compare unscaled
    scale = 4096.0 if flag == '--fp16' else 1.0
    (loss * scale).backward()
    ref.train()
    ref_loss = ref(*[t.cuda() for t in sample])
    ref_loss.backward()
    assert abs(loss.item() - ref_loss.item()) < (0.05 if flag == '--bf16' else 0.02), (loss.item(), ref_loss.item())
    named = dict(ref.named_parameters())
    errs = {}
    for name, p in ctl.get_model().named_parameters():
        g, rg = p.main_grad.float() / scale, named[name].grad
        # relative error in the Frobenius norm: 16-bit activations vs the fp32 oracle
        if 'key.bias' in name:
            continue   # mathematically zero (softmax is shift invariant per row): only noise to compare
        errs[name] = ((g - rg).norm() / (rg.norm() + 1e-12)).item()
    top = sorted(errs.items(), key=lambda kv: -kv[1])[:8]
    print('largest relative gradient errors:', top)
    assert top[0][1] < (0.15 if flag == '--bf16' else 0.10), top


def test_fused_layer_node_matches_per_op_nodes(tmp_path):
    """EncoderLayerFn (one autograd node per layer, fused bias gradients / residual adds) against the
    one-node-per-op path, dropout ON: both recompute the same Philox masks, so they must agree closely."""
    from hetseq_b200.ops import rng
    ctl, task = _build(str(tmp_path), '--fp16', 'native', hidden_dropout_prob=0.1, attention_probs_dropout_prob=0.1)
    model = ctl.get_model()
    nb = model.native
    dev = task.prepare_sample(_batch(), ctl.device)
    model.train()
    grads, losses = [], []
    for fused in (True, False):
        nb.fused_layer = fused
        rng.set_step_seed(4321, 0)
        ctl.zero_grad()
        loss = model(*dev)
        (loss * 1024.0).backward()
        ctl.optimizer.buffer.finalize_grads()
        losses.append(loss.item())
        grads.append(ctl.optimizer.buffer.flat_grad.clone())
    nb.fused_layer = True
    assert abs(losses[0] - losses[1]) < 1e-3, losses
    rel = ((grads[0] - grads[1]).norm() / grads[1].norm()).item()
    assert rel < 5e-3, rel
    for name, p in model.named_parameters():
        buf, i = p._hs_flat
        a = grads[0][buf.offsets[i]:buf.offsets[i] + p.numel()]
        b = grads[1][buf.offsets[i]:buf.offsets[i] + p.numel()]
        if 'key.bias' in name:
            continue
        e = ((a - b).norm() / (b.norm() + 1e-6)).item()
        assert e < 2e-2, (name, e)


def test_native_training_reduces_loss(tmp_path):
    ctl, task = _build(str(tmp_path), '--fp16', 'native', hidden_dropout_prob=0.1, attention_probs_dropout_prob=0.1)
    sample = _batch()
    losses = []
    for _ in range(12):
        out = ctl.train_step([sample])
        if out is not None:
            losses.append(out['loss'])
    assert len(losses) >= 8 and losses[-1] < losses[0] - 0.5, losses


def test_smoke_entry():
    import __graft_entry__ as g
    g.smoke()


def test_cuda_graph_step_matches_eager(tmp_path):
    """--cuda-graph replays the captured update; with dropout off it must track the eager run."""
    import copy
    from hetseq_b200.data.bert_dataset import SyntheticBertData
    ds = SyntheticBertData(64, 128, 10, 2000, seed=5)
    batches = [ds.fetch_rows(list(range(i * 8, i * 8 + 8))) for i in range(8)]
    results = {}
    for mode in ('eager', 'graph'):
        d = os.path.join(str(tmp_path), mode)
        os.makedirs(d)
        ctl, task = _build(d, '--fp16', 'native', seed=0)
        ctl.args.cuda_graph = (mode == 'graph')
        ctl._graph_enabled = (mode == 'graph')
        losses = []
        for b in batches:
            out = ctl.train_step([b])
            losses.append(None if out is None else out['loss'])
        if mode == 'graph':
            assert ctl._graph is not None, 'the step was never captured'
        results[mode] = (losses, ctl.buffer.flat_param.detach().float().clone(), ctl.get_num_updates())
    (le, pe, ne), (lg, pg, ng) = results['eager'], results['graph']
    assert ne == ng
    for a, b in zip(le, lg):
        assert (a is None) == (b is None)
        if a is not None:
            assert abs(a - b) < 2e-2, (le, lg)
    assert (pe - pg).abs().max().item() < 3e-3


def test_gradient_norm_taken_during_backward_matches_single_pass(tmp_path, monkeypatch):
    """One GPU: the norm's sum of squares comes from per-bucket kernels on a side stream, started by the
    gradient-ready callbacks while backward runs (parallel/data_parallel.py::LocalNormBackend). Against the single
    pass over the flat gradient after backward (HETSEQ_NORM_OVERLAP=0): same gnorm up to the order of the fp32
    additions, same parameters after a few updates, in eager and in CUDA-graph steps; the graph must contain the
    side stream's work (a fork that is not joined would fail the capture)."""
    from hetseq_b200.data.bert_dataset import SyntheticBertData
    from hetseq_b200.parallel import data_parallel
    ds = SyntheticBertData(64, 128, 10, 2000, seed=7)
    batches = [ds.fetch_rows(list(range(i * 8, i * 8 + 8))) for i in range(6)]
    res = {}
    for mode in ('single', 'overlap', 'overlap-graph'):
        monkeypatch.setenv('HETSEQ_NORM_OVERLAP', '0' if mode == 'single' else '1')
        monkeypatch.setattr(data_parallel.LocalNormBackend, 'NORM_BUCKET_MB', 1)    # several buckets on this small model
        d = os.path.join(str(tmp_path), mode)
        os.makedirs(d)
        ctl, task = _build(d, '--fp16', 'native', seed=0)
        graph = mode.endswith('graph')
        ctl.args.cuda_graph = graph
        ctl._graph_enabled = graph
        assert isinstance(ctl.dp, data_parallel.LocalNormBackend) == (mode != 'single')
        if mode != 'single':
            assert len(ctl.dp.buckets) > 3
        norms = []
        for b in batches:
            out = ctl.train_step([b])
            norms.append(None if out is None else float(ctl._prev_grad_norm))
        if graph:
            assert ctl._graph is not None, 'the step was never captured'
        res[mode] = (norms, ctl.buffer.flat_param.detach().float().clone(), ctl.get_num_updates())
    n0, p0, u0 = res['single']
    for mode in ('overlap', 'overlap-graph'):
        n1, p1, u1 = res[mode]
        assert u0 == u1
        for a, b in zip(n0, n1):
            assert (a is None) == (b is None)
            if a is not None:
                assert abs(a - b) <= 2e-3 * max(1.0, abs(a)), (mode, n0, n1)
        assert (p0 - p1).abs().max().item() < 3e-3


def test_native_activation_checkpointing_matches(tmp_path):
    """--checkpoint-activations on the native path: the encoder is recomputed in chunks of ceil(sqrt(L)) layers during
    backward (reference code path bert_modeling.py:470-476, unreachable there). The recomputation redraws the same
    Philox dropout masks, so loss and gradients must equal the run without checkpointing (up to the order of the fp32
    atomics behind the bias / LayerNorm gradients)."""
    from hetseq_b200.ops import rng
    ctl, task = _build(str(tmp_path), '--fp16', 'native', num_hidden_layers=5, hidden_dropout_prob=0.1,
                       attention_probs_dropout_prob=0.1)
    model = ctl.get_model()
    dev = task.prepare_sample(_batch(), ctl.device)
    model.train()
    out = []
    for ckpt in (False, True):
        rng.set_step_seed(777, 0)
        ctl.zero_grad()
        loss = model(*dev, checkpoint_activations=ckpt)
        (loss * 1024.0).backward()
        ctl.optimizer.buffer.finalize_grads()
        out.append((loss.item(), ctl.optimizer.buffer.flat_grad.clone()))
    assert abs(out[0][0] - out[1][0]) < 1e-4, (out[0][0], out[1][0])
    rel = ((out[0][1] - out[1][1]).norm() / out[0][1].norm()).item()
    assert rel < 1e-4, rel


def test_unsupported_sequence_length_falls_back_with_a_warning(tmp_path, capsys):
    """S = 72 is not a multiple of 16: the layer keeps the native GEMMs / LayerNorm and takes the PyTorch attention
    composite, announcing it once (no silent cliff)."""
    ctl, task = _build(str(tmp_path), '--fp16', 'native')
    dev = task.prepare_sample(_batch(S=72), ctl.device)
    model = ctl.get_model()
    model.train()
    ctl.zero_grad()
    loss = model(*dev)
    (loss * 1024.0).backward()
    assert loss.item() == loss.item() and 5.0 < loss.item() < 12.0
    assert 'no fused attention kernel for sequence length 72' in capsys.readouterr().out
