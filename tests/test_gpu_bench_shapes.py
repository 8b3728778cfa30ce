"""GPU tests at the shapes every quoted number comes from (BERT-large: H = 1024, 16 heads, I = 4096, 32 x 128
tokens, dropout on), plus training-behaviour checks of the native path: a 100-update loss curve against the fp32
PyTorch model and the loss-scale overflow skip."""
import math
import os

import pytest
import torch

from philox_ref import keep_mask
from test_gpu_bert import _batch, _build

pytestmark = pytest.mark.gpu


def _ln(z, gamma, beta):
    mu = z.mean(-1, keepdim=True)
    var = (z - mu).pow(2).mean(-1, keepdim=True)
    return (z - mu) / torch.sqrt(var + 1e-12) * gamma + beta      # TF-style, bert_modeling.py:285-289


def _gelu(x):
    return x * 0.5 * (1.0 + torch.erf(x / 1.41421))               # bert_modeling.py:104-116


def test_encoder_layer_at_benchmark_shape_with_dropout(tmp_path):
    """One BertLayer (/root/reference/bert_modeling.py:351-441) through EncoderLayerFn at H=1024 / 16 heads /
    I=4096 / B=32 / S=128 in fp16 with dropout 0.1 at all three sites, against an fp32 restatement that applies
    the SAME masks (rebuilt on the host from the Philox stream): output and every gradient within 2 %."""
    from hetseq_b200.ops import bert_native as bn, rng
    B, S, H, heads, I = 32, 128, 1024, 16, 4096
    p = 0.1
    ctl, task = _build(str(tmp_path), '--fp16', 'native', hidden_size=H, num_hidden_layers=1, num_attention_heads=heads,
                       intermediate_size=I, vocab_size=2000, hidden_dropout_prob=p, attention_probs_dropout_prob=p)
    model = ctl.get_model()
    nb, layer = model.native, model.bert.encoder.layer[0]
    torch.manual_seed(21)
    x = torch.randn(B * S, H, device='cuda').half().requires_grad_(True)
    add_mask = torch.zeros(B, S, device='cuda')
    for b in range(0, B, 3):
        add_mask[b, S - 5 * (b % 7 + 1):] = -10000.0
    dy = torch.randn(B * S, H, device='cuda').half()
    seed, rank = 987654, 0
    rng.set_step_seed(seed, rank)
    nb._bs = (B, S)
    model.train()
    ctl.zero_grad()
    y = bn.EncoderLayerFn.apply(x, add_mask, layer.output.LayerNorm.weight, nb, layer, True)
    y.backward(dy)
    ctl.optimizer.buffer.finalize_grads()

    # ---- fp32 oracle with the same masks ----------------------------------------------------------------
    s_attn, s_ao, s_ffn = (nb.sites[(id(layer), t)] for t in ('attn', 'ao', 'ffn'))
    m_attn = torch.from_numpy(keep_mask(seed, rank, s_attn, B * heads * S * S, p)).cuda().view(B, heads, S, S)
    m_ao = torch.from_numpy(keep_mask(seed, rank, s_ao, B * S * H, p)).cuda().view(B * S, H)
    m_ffn = torch.from_numpy(keep_mask(seed, rank, s_ffn, B * S * H, p)).cuda().view(B * S, H)
    assert abs(float((m_attn == 0).float().mean()) - p) < 2e-3
    names = ['attention.self.query', 'attention.self.key', 'attention.self.value', 'attention.output.dense',
             'attention.output.LayerNorm', 'intermediate.dense_act', 'output.dense', 'output.LayerNorm']
    mods = dict(layer.named_modules())
    W = {n: (mods[n].weight.detach().float().requires_grad_(True), mods[n].bias.detach().float().requires_grad_(True))
         for n in names}
    x32 = x.detach().float().requires_grad_(True)
    d = H // heads

    def heads_of(t):
        return t.view(B, S, heads, d).permute(0, 2, 1, 3)
    q = heads_of(x32 @ W[names[0]][0].t() + W[names[0]][1])
    k = heads_of(x32 @ W[names[1]][0].t() + W[names[1]][1])
    v = heads_of(x32 @ W[names[2]][0].t() + W[names[2]][1])
    probs = torch.softmax(q @ k.transpose(-1, -2) / math.sqrt(d) + add_mask.view(B, 1, 1, S), dim=-1)
    ctxv = ((probs * m_attn) @ v).permute(0, 2, 1, 3).reshape(B * S, H)
    ao = ctxv @ W[names[3]][0].t() + W[names[3]][1]
    x1 = _ln(x32 + ao * m_ao, *W[names[4]])
    hmid = _gelu(x1 @ W[names[5]][0].t() + W[names[5]][1])
    f = hmid @ W[names[6]][0].t() + W[names[6]][1]
    x2 = _ln(x1 + f * m_ffn, *W[names[7]])
    x2.backward(dy.float())

    def rel(a, r):
        return float((a.float() - r).norm() / (r.norm() + 1e-20))
    errs = {'y': rel(y.detach(), x2.detach()), 'dx': rel(x.grad, x32.grad)}
    for n in names:
        for kind, ref in zip(('weight', 'bias'), W[n]):
            if n.endswith('key') and kind == 'bias':
                continue       # mathematically zero: softmax is invariant to a per-row shift of the scores
            errs[n + '.' + kind] = rel(getattr(mods[n], kind).main_grad, ref.grad)
    worst = sorted(errs.items(), key=lambda kv: -kv[1])[:6]
    print('largest relative errors at the benchmark shape:', worst)
    assert worst[0][1] < 0.02, worst


def test_loss_curve_native_fp16_tracks_torch_fp32(tmp_path):
    """100 updates from the same initial weights on the same batches, dropout off: the native fp16 path (tcgen05
    GEMMs, fused attention / LayerNorm / cross-entropy, flat Adam with dynamic loss scaling) against the PyTorch
    fp32 model with the per-tensor-equivalent flat Adam. Counterpart of the reference's only end-to-end check,
    "does the loss go down the same way" (/root/reference/train.py:135-161)."""
    from hetseq_b200.data.bert_dataset import SyntheticBertData
    ds = SyntheticBertData(8 * 16, 128, 10, 2000, seed=17)
    batches = [ds.fetch_rows(list(range(i * 8, i * 8 + 8))) for i in range(16)]
    curves = {}
    for mode, flag, kernels in (('native', '--fp16', 'native'), ('torch', None, 'torch')):
        d = os.path.join(str(tmp_path), mode)
        os.makedirs(d)
        ctl, _ = _build(d, flag, kernels, seed=0)
        if mode == 'torch':
            ctl.get_model().load_state_dict({k: v.float() for k, v in init.items()})
            ctl.optimizer.refresh_master()
        else:
            init = {k: v.detach().clone() for k, v in ctl.get_model().state_dict().items()}
            ctl.optimizer.scaler.loss_scale = 1024.0     # start below the overflow point: no skipped updates
        losses = []
        for it in range(100):
            out = ctl.train_step([batches[it % len(batches)]])
            if out is not None:
                losses.append(out['loss'])
        curves[mode] = losses
    a, b = curves['native'], curves['torch']
    assert len(b) == 100 and len(a) >= 98, (len(a), len(b))        # hardly any loss-scale skip
    tail_a, tail_b = sum(a[-10:]) / 10, sum(b[-10:]) / 10
    assert tail_b < b[0] - 1.0, 'the fp32 oracle itself must have learned something'
    assert abs(tail_a - tail_b) < 0.05 * tail_b, (tail_a, tail_b)
    k = len(a)
    worst = max(abs(x - y) for x, y in zip(a[-k:], b[-k:]))
    assert worst < 0.35, worst          # log2 units; fp16 activations, and a skipped update shifts the curve


def test_overflow_skips_the_native_update(tmp_path):
    """A loss scale far too large makes the gradients overflow: the fused Adam kernel reads the overflow flag from
    device memory (skip_dev) and must leave parameters, master weights and both moments untouched; the host halves
    the scale and does not count the update (the OverflowError path of /root/reference/controller.py:368-371)."""
    ctl, _ = _build(str(tmp_path), '--fp16', 'native', seed=0)
    ctl.optimizer.scaler.loss_scale = 2.0 ** 40
    sample = _batch()
    before = ctl.buffer.flat_param.clone()
    out = ctl.train_step([sample])
    assert out is None and ctl.overflow_updates == 1 and ctl.get_num_updates() == 0
    assert ctl.optimizer.step_count == 0 and ctl.optimizer.scaler.loss_scale == 2.0 ** 39
    assert torch.equal(ctl.buffer.flat_param, before)
    assert float(ctl.optimizer.state['exp_avg'].abs().max()) == 0.0
    assert float(ctl.optimizer.state['exp_avg_sq'].abs().max()) == 0.0
    ctl.optimizer.scaler.loss_scale = 2.0 ** 12
    out = ctl.train_step([sample])
    assert out is not None and ctl.optimizer.step_count == 1 and ctl.get_num_updates() == 1
    assert not torch.equal(ctl.buffer.flat_param, before)
