"""BERT on the sm_100a kernels: autograd Functions + the glue that attaches them to the module tree.

``attach(model, buffer, args)`` leaves parameters, names and checkpoints untouched and swaps the
*computation* of BertModel / BertLayer / BertForPreTraining for the native path:

  embeddings   hs_embed_gather_sum -> hs_ln_fwd(post_dropout)                      (K8 / W8)
  per layer    QKV: ONE tcgen05 GEMM against the packed [3H,H] weight view          (K1 / W3)
               attention: fused masked-softmax(+dropout) PV kernel                  (K2-K4 / W4)
               out-proj GEMM(+bias) -> hs_ln_fwd(dropout + residual + LN)           (K5 / W5, W7)
               FFN1 GEMM + bias + GELU epilogue (pre-activation kept for backward)  (K6 / W6)
               FFN2 GEMM(+bias) -> hs_ln_fwd(dropout + residual + LN)               (K7)
  heads        gather labelled rows -> transform GEMM+GELU -> LN -> decoder GEMM against the tied
               embedding matrix -> hs_xent in place; pooler GEMM+tanh on the strided CLS rows; NSP
               GEMM -> hs_xent                                                     (K9-K13 / W9, W10)
  backward     dgrad = GEMM(K-major dY, MN-major W); wgrad = GEMM(MN-major dY, MN-major X) written (or
               accumulated) in fp32 straight into the flat main-grad buffer; FFN1's dgrad applies
               GELU' in the FFN2-dgrad epilogue; bias grads via hs_colsum                (K14 / W11)

Activations are 16-bit (fp16 or bf16), accumulation fp32. Dropout masks are never stored.
"""
from __future__ import annotations

import math
import os
from typing import Optional

import torch
import torch.nn.functional as F

from hetseq_b200.ops import lib, native, rng
from hetseq_b200.utils import nvtx
from hetseq_b200.ops.native import (ACT_ADD_AUX, ACT_DGELU_MUL, ACT_GELU, ACT_GELU_DAUX, ACT_MUL_AUX, ACT_NONE,
                                    ACT_TANH)


def _flat(p):
    """(FlatParamBuffer, index) of a parameter, or None when it is not in a flat buffer."""
    return getattr(p, '_hs_flat', None)


def _wgrad_target(params):
    """fp32 destination + accumulate flag for the gradient of one parameter, or of several
    parameters that are adjacent in the flat buffer and written by one kernel."""
    buf, i0 = params[0]._hs_flat
    acc = buf._written[i0]
    n = 0
    for p in params:
        b, i = p._hs_flat
        buf._written[i] = True
        n += p.numel()
    start = buf.offsets[i0]
    return buf.flat_grad[start:start + n], acc


def _notify(params):
    for p in params:
        buf, _ = p._hs_flat
        buf.notify_ready(p)


# ------------------------------------------------------------------------------------------------
# Linear (+bias, +activation)
# ------------------------------------------------------------------------------------------------
class LinearFn(torch.autograd.Function):
    """y = act(x W^T + b). ``wparams``/``bparams`` are the nn.Parameters behind ``weight``/``bias``
    (several when the weight is a packed view, e.g. q/k/v). x: [M, K] with row stride ``ldx``."""

    @staticmethod
    def forward(ctx, x, weight, bias, wparams, bparams, act, ldx, ldy_pad):
        M = x.shape[0]
        N, K = weight.shape
        ldx = ldx or K
        ldy = ldy_pad or N
        y = torch.empty(M, ldy, dtype=x.dtype, device=x.device)
        # GELU: the epilogue also emits gelu'(pre-activation) (they share the erf), so backward is one multiply
        aux = torch.empty(M, N, dtype=x.dtype, device=x.device) if act == ACT_GELU else None
        native.gemm(x, weight, y, M, N, K, lda=ldx, ldb=K, ldc=ldy, bias=bias,
                    act=ACT_GELU_DAUX if act == ACT_GELU else act, aux_out=aux, ldaux=N)
        ctx.save_for_backward(x, weight, aux if aux is not None else (y if act == ACT_TANH else None))
        ctx.meta = (wparams, bparams, act, ldx, ldy, M, N, K)
        return y if ldy == N else y[:, :N]

    @staticmethod
    def backward(ctx, dy):
        x, weight, saved = ctx.saved_tensors
        wparams, bparams, act, ldx, ldy, M, N, K = ctx.meta
        if act == ACT_GELU:
            dy = dy * saved                      # saved = gelu'(pre-activation), written by the forward epilogue
        elif act == ACT_TANH:
            yv = saved[:, :N].float()
            dy = (dy.float() * (1 - yv * yv)).to(x.dtype)
        dy = dy.contiguous()
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty(M, K, dtype=x.dtype, device=x.device)
            native.gemm(dy, weight, dx, M, K, N, lda=N, ldb=K, ldc=K, b_mn=True)            # dX = dY W
        wg, acc = _wgrad_target(wparams)
        native.gemm(dy, x, wg, N, K, M, lda=N, ldb=ldx, ldc=K, a_mn=True, b_mn=True, accumulate=acc)  # dW = dY^T X
        _notify(wparams)
        if bparams:
            bg, bacc = _wgrad_target(bparams)
            native.colsum(dy, M, N, N, bg, bacc)
            _notify(bparams)
        return dx, None, None, None, None, None, None, None


def linear(x, weight, bias, wparams, bparams, act=ACT_NONE, ldx=0, ldy_pad=0):
    return LinearFn.apply(x, weight, bias, wparams, bparams, act, ldx, ldy_pad)


class FfnFn(torch.autograd.Function):
    """f = (gelu(x W1^T + b1)) W2^T + b2 with both GEMM epilogues fused; backward applies GELU' inside
    the FFN2-dgrad GEMM epilogue, so no elementwise pass ever touches the [M, 4H] tensors."""

    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2, p1w, p1b, p2w, p2b):
        M, H = x.shape
        I = w1.shape[0]
        pre = torch.empty(M, I, dtype=x.dtype, device=x.device)
        h = torch.empty(M, I, dtype=x.dtype, device=x.device)
        native.gemm(x, w1, h, M, I, H, lda=H, ldb=H, ldc=I, bias=b1, act=ACT_GELU, aux_out=pre, ldaux=I)
        f = torch.empty(M, H, dtype=x.dtype, device=x.device)
        native.gemm(h, w2, f, M, H, I, lda=I, ldb=I, ldc=H, bias=b2)
        ctx.save_for_backward(x, w1, w2, pre, h)
        ctx.params = (p1w, p1b, p2w, p2b)
        return f

    @staticmethod
    def backward(ctx, df):
        x, w1, w2, pre, h = ctx.saved_tensors
        p1w, p1b, p2w, p2b = ctx.params
        M, H = x.shape
        I = w1.shape[0]
        df = df.contiguous()
        # dPre = (dF W2) * gelu'(pre)   -- one GEMM, dGELU in the epilogue
        dpre = torch.empty(M, I, dtype=x.dtype, device=x.device)
        native.gemm(df, w2, dpre, M, I, H, lda=H, ldb=I, ldc=I, b_mn=True, act=ACT_DGELU_MUL, aux_in=pre, ldaux=I)
        wg2, acc2 = _wgrad_target([p2w])
        native.gemm(df, h, wg2, H, I, M, lda=H, ldb=I, ldc=I, a_mn=True, b_mn=True, accumulate=acc2)
        _notify([p2w])
        bg2, bacc2 = _wgrad_target([p2b])
        native.colsum(df, M, H, H, bg2, bacc2)
        _notify([p2b])
        dx = torch.empty(M, H, dtype=x.dtype, device=x.device)
        native.gemm(dpre, w1, dx, M, H, I, lda=I, ldb=H, ldc=H, b_mn=True)
        wg1, acc1 = _wgrad_target([p1w])
        native.gemm(dpre, x, wg1, I, H, M, lda=I, ldb=H, ldc=H, a_mn=True, b_mn=True, accumulate=acc1)
        _notify([p1w])
        bg1, bacc1 = _wgrad_target([p1b])
        native.colsum(dpre, M, I, I, bg1, bacc1)
        _notify([p1b])
        return dx, None, None, None, None, None, None, None, None


# ------------------------------------------------------------------------------------------------
# dropout + residual + LayerNorm
# ------------------------------------------------------------------------------------------------
class DropAddLnFn(torch.autograd.Function):
    """y = LN(residual + dropout(x)); ``post`` = True means y = dropout(LN(x)) (embeddings)."""

    @staticmethod
    def forward(ctx, x, residual, gamma, beta, pg, pb, p, site, post):
        seed = rng.seed_tensor(x.device) if p > 0 else None
        y, z, mean, rstd = native.ln_fwd(x, None, residual, gamma, beta, eps=1e-12, p=p, seed=seed, site=site,
                                         post_dropout=post)
        ctx.save_for_backward(z, gamma, mean, rstd)
        ctx.meta = (pg, pb, p, site, post, residual is not None)
        return y

    @staticmethod
    def backward(ctx, dy):
        z, gamma, mean, rstd = ctx.saved_tensors
        pg, pb, p, site, post, has_res = ctx.meta
        seed = rng.seed_tensor(dy.device) if p > 0 else None
        gg, gacc = _wgrad_target([pg])
        bg, bacc = _wgrad_target([pb])
        assert gacc == bacc
        dz, dx = native.ln_bwd(dy.contiguous(), z, gamma, mean, rstd, gg, bg, gacc, p=p, seed=seed, site=site,
                               post_dropout=post)
        _notify([pg, pb])
        return dx, (dz if has_res else None), None, None, None, None, None, None, None


def drop_add_ln(x, residual, ln_module, p, site, post=False):
    return DropAddLnFn.apply(x, residual, ln_module.weight, ln_module.bias, ln_module.weight, ln_module.bias,
                             float(p), int(site), bool(post))


# ------------------------------------------------------------------------------------------------
# one autograd node per encoder layer
# ------------------------------------------------------------------------------------------------
class EncoderLayerFn(torch.autograd.Function):
    """A whole BertLayer (bert_modeling.py:351-437) as ONE autograd node with a hand-scheduled backward.

    forward : QKV GEMM -> fused attention -> out-proj GEMM -> dropout+residual+LN -> FFN1 GEMM(+GELU) ->
              FFN2 GEMM -> dropout+residual+LN                                  (7 launches)
    backward: LN bwd (+ FFN2 bias grad) -> FFN2 dgrad (x GELU', + FFN1 bias grad in the epilogue), FFN2 wgrad
              -> FFN1 dgrad (+ residual-path gradient in the epilogue), FFN1 wgrad -> LN bwd (+ out-proj bias
              grad) -> out-proj dgrad, wgrad -> attention bwd (+ q/k/v bias grad) -> QKV dgrad (+ residual-path
              gradient), QKV wgrad                                            (11 launches)
    Compared with one node per op this removes the four bias-gradient reductions and the two gradient
    additions autograd inserts where the residual stream forks. Gradients go straight into the flat fp32
    buffer; each parameter is reported ready as soon as its gradient is final (bucket overlap)."""

    @staticmethod
    def forward(ctx, x, add_mask, anchor, nb, layer, training):
        nvtx.push('encoder layer fwd')
        try:
            return EncoderLayerFn._forward(ctx, x, add_mask, anchor, nb, layer, training)
        finally:
            nvtx.pop()

    @staticmethod
    def backward(ctx, dx2):
        nvtx.push('encoder layer bwd')
        try:
            return EncoderLayerFn._backward(ctx, dx2)
        finally:
            nvtx.pop()

    @staticmethod
    def _forward(ctx, x, add_mask, anchor, nb, layer, training):
        L = lib.load()
        B, S = nb._bs
        H, I, heads = nb.H, nb.I, nb.heads
        M = x.shape[0]
        dt, dev = x.dtype, x.device
        qkv_w, qkv_b, ws, bs = nb.packed[id(layer)]
        so, inter, out = layer.attention.output, layer.intermediate.dense_act, layer.output
        p_attn = nb.p_attn if training else 0.0
        p_h = nb.p_hidden if training else 0.0
        seed = rng.seed_tensor(dev) if (p_attn > 0 or p_h > 0) else None
        s_attn, s_ao, s_ffn = (nb.sites[(id(layer), t)] for t in ('attn', 'ao', 'ffn'))

        qkv = torch.empty(M, 3 * H, dtype=dt, device=dev)
        native.gemm(x, qkv_w, qkv, M, 3 * H, H, lda=H, ldb=H, ldc=3 * H, bias=qkv_b)
        ctxv = torch.empty(M, H, dtype=dt, device=dev)
        lse = torch.empty(B * heads * S, dtype=torch.float32, device=dev)
        lib.check(L.hs_attention_fwd(qkv.data_ptr(), add_mask.data_ptr(), ctxv.data_ptr(), lse.data_ptr(), B, S, heads, H,
                                     float(p_attn), lib.ptr(seed), int(s_attn), lib.dtype_code(dt), lib.stream_ptr()),
                  'hs_attention_fwd')
        ao = torch.empty(M, H, dtype=dt, device=dev)
        native.gemm(ctxv, so.dense.weight, ao, M, H, H, lda=H, ldb=H, ldc=H, bias=so.dense.bias)
        x1, z1, mean1, rstd1 = native.ln_fwd(ao, None, x, so.LayerNorm.weight, so.LayerNorm.bias, eps=1e-12, p=p_h,
                                             seed=seed, site=s_ao)
        del ao
        # FFN1: the epilogue emits gelu(pre) AND gelu'(pre) (they share the erf); backward then only multiplies
        dact = torch.empty(M, I, dtype=dt, device=dev)
        h = torch.empty(M, I, dtype=dt, device=dev)
        native.gemm(x1, inter.weight, h, M, I, H, lda=H, ldb=H, ldc=I, bias=inter.bias, act=ACT_GELU_DAUX, aux_out=dact,
                    ldaux=I)
        f = torch.empty(M, H, dtype=dt, device=dev)
        native.gemm(h, out.dense.weight, f, M, H, I, lda=I, ldb=I, ldc=H, bias=out.dense.bias)
        x2, z2, mean2, rstd2 = native.ln_fwd(f, None, x1, out.LayerNorm.weight, out.LayerNorm.bias, eps=1e-12, p=p_h,
                                             seed=seed, site=s_ffn)
        del f
        ctx.save_for_backward(x, add_mask, qkv, ctxv, lse, z1, mean1, rstd1, x1, dact, h, z2, mean2, rstd2)
        ctx.meta = (nb, layer, p_attn, p_h, B, S)
        return x2

    @staticmethod
    def _backward(ctx, dx2):
        x, add_mask, qkv, ctxv, lse, z1, mean1, rstd1, x1, dact, h, z2, mean2, rstd2 = ctx.saved_tensors
        nb, layer, p_attn, p_h, B, S = ctx.meta
        L = lib.load()
        H, I, heads = nb.H, nb.I, nb.heads
        M = x.shape[0]
        dt, dev = x.dtype, x.device
        qkv_w, qkv_b, ws, bs = nb.packed[id(layer)]
        so, inter, out = layer.attention.output, layer.intermediate.dense_act, layer.output
        seed = rng.seed_tensor(dev) if (p_attn > 0 or p_h > 0) else None
        s_attn, s_ao, s_ffn = (nb.sites[(id(layer), t)] for t in ('attn', 'ao', 'ffn'))

        # ---- output LayerNorm: dz2 feeds the residual path, df the FFN; FFN2's bias gradient rides along
        gg, gacc = _wgrad_target([out.LayerNorm.weight])
        bg, bacc = _wgrad_target([out.LayerNorm.bias])
        b2g, b2acc = _wgrad_target([out.dense.bias])
        dz2, df = native.ln_bwd(dx2.contiguous(), z2, out.LayerNorm.weight, mean2, rstd2, gg, bg, gacc, p=p_h,
                                seed=seed, site=s_ffn, dbias=b2g, accumulate_bias=b2acc)
        assert gacc == bacc
        _notify([out.LayerNorm.weight, out.LayerNorm.bias, out.dense.bias])
        # ---- FFN
        b1g, b1acc = _wgrad_target([inter.bias])
        dpre = torch.empty(M, I, dtype=dt, device=dev)
        native.gemm(df, out.dense.weight, dpre, M, I, H, lda=H, ldb=I, ldc=I, b_mn=True, act=ACT_MUL_AUX, aux_in=dact,
                    ldaux=I, colsum_out=b1g, colsum_accumulate=b1acc)
        wg2, acc2 = _wgrad_target([out.dense.weight])
        native.gemm(df, h, wg2, H, I, M, lda=H, ldb=I, ldc=I, a_mn=True, b_mn=True, accumulate=acc2)
        _notify([out.dense.weight, inter.bias])
        dx1 = torch.empty(M, H, dtype=dt, device=dev)
        native.gemm(dpre, inter.weight, dx1, M, H, I, lda=I, ldb=H, ldc=H, b_mn=True, act=ACT_ADD_AUX, aux_in=dz2, ldaux=H)
        wg1, acc1 = _wgrad_target([inter.weight])
        native.gemm(dpre, x1, wg1, I, H, M, lda=I, ldb=H, ldc=H, a_mn=True, b_mn=True, accumulate=acc1)
        _notify([inter.weight])
        del dpre, dz2, df
        # ---- attention-output LayerNorm (+ out-proj bias gradient)
        gg, gacc = _wgrad_target([so.LayerNorm.weight])
        bg, bacc = _wgrad_target([so.LayerNorm.bias])
        bog, boacc = _wgrad_target([so.dense.bias])
        dz1, dao = native.ln_bwd(dx1, z1, so.LayerNorm.weight, mean1, rstd1, gg, bg, gacc, p=p_h, seed=seed,
                                 site=s_ao, dbias=bog, accumulate_bias=boacc)
        assert gacc == bacc
        _notify([so.LayerNorm.weight, so.LayerNorm.bias, so.dense.bias])
        dctx = torch.empty(M, H, dtype=dt, device=dev)
        native.gemm(dao, so.dense.weight, dctx, M, H, H, lda=H, ldb=H, ldc=H, b_mn=True)
        wgo, acco = _wgrad_target([so.dense.weight])
        native.gemm(dao, ctxv, wgo, H, H, M, lda=H, ldb=H, ldc=H, a_mn=True, b_mn=True, accumulate=acco)
        _notify([so.dense.weight])
        # ---- attention (+ q/k/v bias gradient)
        dqkv = torch.empty(M, 3 * H, dtype=dt, device=dev)
        bqg, bqacc = _wgrad_target(bs)
        lib.check(L.hs_attention_bwd(qkv.data_ptr(), add_mask.data_ptr(), ctxv.data_ptr(), dctx.data_ptr(), lse.data_ptr(),
                                     dqkv.data_ptr(), bqg.data_ptr(), int(bqacc), B, S, heads, H, float(p_attn),
                                     lib.ptr(seed), int(s_attn), lib.dtype_code(dt), lib.stream_ptr()), 'hs_attention_bwd')
        _notify(bs)
        dx = torch.empty(M, H, dtype=dt, device=dev)
        native.gemm(dqkv, qkv_w, dx, M, H, 3 * H, lda=3 * H, ldb=H, ldc=H, b_mn=True, act=ACT_ADD_AUX, aux_in=dz1, ldaux=H)
        wgq, accq = _wgrad_target(ws)
        native.gemm(dqkv, x, wgq, 3 * H, H, M, lda=3 * H, ldb=H, ldc=H, a_mn=True, b_mn=True, accumulate=accq)
        _notify(ws)
        return dx, None, None, None, None, None


# ------------------------------------------------------------------------------------------------
# embeddings
# ------------------------------------------------------------------------------------------------
class EmbeddingSumFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, W, P, T, ids, seg, pW, pP, pT, B, S, plan=None):
        z = native.embed_gather_sum(W, P, T, ids, seg, B * S, S)
        if plan is None and os.environ.get('HETSEQ_EMBED_ATOMIC', '0') != '1':
            plan = native.embedding_scatter_plan(ids)     # device sort; the task normally supplies a host-made plan
        if plan is None:
            ctx.save_for_backward(ids, seg)
        else:
            ctx.save_for_backward(ids, seg, plan[0].contiguous(), plan[1].contiguous())
        ctx.meta = (pW, pP, pT, B, S)
        return z

    @staticmethod
    def backward(ctx, de):
        ids, seg = ctx.saved_tensors[:2]
        pW, pP, pT, B, S = ctx.meta
        de = de.contiguous()
        H = de.shape[1]
        bufW = pW._hs_flat[0]
        dW = bufW.ensure_initialized(pW)          # the scatter adds to what the decoder's wgrad left (tied matrix)
        tg, tacc = _wgrad_target([pT])
        tg = tg.view(pT.shape)
        types_done = False
        if len(ctx.saved_tensors) == 4:
            # segmented reduction over positions sorted by token id: one writer per (token, column), fixed order;
            # the two token-type sums come out of the same pass over de
            types_done = pT.shape[0] == 2 and seg.dtype == ids.dtype and tg.is_contiguous()
            native.embed_backward_sorted(de, ids, ctx.saved_tensors[2], ctx.saved_tensors[3], dW,
                                         seg=seg if types_done else None, dT=tg if types_done else None, accumulate_T=tacc)
        else:
            native.embed_scatter_word(de, ids, dW)    # fp32 atomics (HETSEQ_EMBED_ATOMIC=1)
        _notify([pW])
        pg, pacc = _wgrad_target([pP])
        pg = pg.view(pP.shape)
        if pacc:
            native.sum_over_batch(de, B, S * H, pg[:S].reshape(-1), True)
        else:
            if S < pP.shape[0]:
                pg[S:].zero_()
            native.sum_over_batch(de, B, S * H, pg[:S].reshape(-1), False)
        _notify([pP])
        if not types_done:
            for t in range(pT.shape[0]):
                native.colsum_rows_where(de, seg, t, tg[t], tacc)
        _notify([pT])
        return None, None, None, None, None, None, None, None, None, None, None


# ------------------------------------------------------------------------------------------------
# attention (PyTorch composite until the fused kernel is attached; see attention_native.py)
# ------------------------------------------------------------------------------------------------
def attention_composite(qkv, add_mask, B, S, heads, p, training):
    """softmax(QK^T/sqrt(d) + mask) V on a packed [M, 3H] projection -- library ops, used only when
    the fused kernel is unavailable for a shape; numerically the reference formulation."""
    M, H3 = qkv.shape
    H = H3 // 3
    d = H // heads
    q, k, v = qkv.view(B, S, 3, heads, d).permute(2, 0, 3, 1, 4)
    scores = torch.matmul(q, k.transpose(-1, -2)) / math.sqrt(d) + add_mask.view(B, 1, 1, S).to(qkv.dtype)
    probs = F.dropout(F.softmax(scores, dim=-1), p=p, training=training)
    return torch.matmul(probs, v).permute(0, 2, 1, 3).reshape(M, H)


# ------------------------------------------------------------------------------------------------
# heads
# ------------------------------------------------------------------------------------------------
class GatherRowsFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, src, idx):
        ctx.save_for_backward(idx)
        ctx.shape = src.shape
        return native.gather_rows(src, idx, idx.shape[0])

    @staticmethod
    def backward(ctx, d):
        (idx,) = ctx.saved_tensors
        out = torch.zeros(ctx.shape, dtype=d.dtype, device=d.device)
        native.scatter_add_rows(d.contiguous(), idx, out)
        return out, None


class LinearXentFn(torch.autograd.Function):
    """loss = mean CE(x W^T + b, labels) with ignore_index = -1; logits live in one padded buffer that
    the CE kernel turns into their own gradient in place (MLM decoder: W is the tied embedding matrix)."""

    @staticmethod
    def forward(ctx, x, weight, bias, labels, wparams, bparams):
        M, K = x.shape
        V = weight.shape[0]
        ld = (V + 127) // 128 * 128 if V > 128 else 8
        logits = torch.empty(M, ld, dtype=x.dtype, device=x.device)
        native.gemm(x, weight, logits, M, V, K, lda=K, ldb=K, ldc=ld, bias=bias)
        # [loss, 1 / number of labelled rows]: the stored gradient is softmax - onehot, O(1) entries that
        # survive fp16; the 1/n (and the upstream gradient with the loss scale) go into the GEMMs' fp32 alpha
        out2 = torch.zeros(2, dtype=torch.float32, device=x.device)
        loss = out2[:1]
        native.xent(logits, ld, labels, M, V, loss, False, write_grad=True, inv_count_out=out2[1:])
        ctx.save_for_backward(x, weight, logits, out2)
        ctx.meta = (wparams, bparams, M, K, V, ld)
        return loss[0]

    @staticmethod
    def backward(ctx, g):
        x, weight, dlogits, out2 = ctx.saved_tensors
        wparams, bparams, M, K, V, ld = ctx.meta
        g = (g.to(torch.float32).reshape(1) * out2[1:]).contiguous()
        dx = torch.empty(M, K, dtype=x.dtype, device=x.device)
        native.gemm(dlogits, weight, dx, M, K, V, lda=ld, ldb=K, ldc=K, b_mn=True, alpha_dev=g)
        wg, acc = _wgrad_target(wparams)
        native.gemm(dlogits, x, wg, V, K, M, lda=ld, ldb=K, ldc=K, a_mn=True, b_mn=True, accumulate=acc, alpha_dev=g)
        _notify(wparams)
        if bparams:
            bg, bacc = _wgrad_target(bparams)
            tmp = torch.empty(V, dtype=torch.float32, device=x.device)
            native.colsum(dlogits, M, V, ld, tmp, False)
            if bacc:
                bg.add_(tmp * g)
            else:
                torch.mul(tmp, g, out=bg)
            _notify(bparams)
        return dx, None, None, None, None, None


# ------------------------------------------------------------------------------------------------
# attach
# ------------------------------------------------------------------------------------------------
class NativeBert(object):
    def __init__(self, model, buffer, args):
        from hetseq_b200.models import bert as bm
        self.args = args
        self.buffer = buffer
        self.bert = model.bert if hasattr(model, 'bert') else model
        cfg = self.bert.config
        self.cfg = cfg
        if cfg.hidden_act != 'gelu':
            raise RuntimeError('native BERT path supports hidden_act=gelu only (got {})'.format(cfg.hidden_act))
        self.H, self.heads, self.I = cfg.hidden_size, cfg.num_attention_heads, cfg.intermediate_size
        if self.H % 8 or self.I % 8 or (self.H // self.heads) % 8:
            raise RuntimeError('native BERT path needs hidden/intermediate/head sizes that are multiples of 8')
        self.p_hidden = float(cfg.hidden_dropout_prob)
        self.p_attn = float(cfg.attention_probs_dropout_prob)
        self.sites = {}
        # one autograd node per encoder layer (EncoderLayerFn); HETSEQ_FUSED_LAYER=0 keeps one node per op
        self.fused_layer = os.environ.get('HETSEQ_FUSED_LAYER', '1') != '0'
        self.attn_impl = None
        try:
            from hetseq_b200.ops import attention_native
            self.attn_impl = attention_native
        except ImportError:
            self.attn_impl = None
        # packed q/k/v views (adjacent in the flat buffer thanks to BertForPreTraining.param_order)
        self.packed = {}
        for li, layer in enumerate(self.bert.encoder.layer):
            sa = layer.attention.self
            ws, bs = [sa.query.weight, sa.key.weight, sa.value.weight], [sa.query.bias, sa.key.bias, sa.value.bias]
            self.packed[id(layer)] = (self._pack(ws, (3 * self.H, self.H)), self._pack(bs, (3 * self.H,)), ws, bs)
            layer.native = self
            for tag in ('attn', 'ao', 'ffn'):
                self.sites[(id(layer), tag)] = rng.new_site()
        self.site_emb = rng.new_site()
        self.bert.native = self
        # Parameters may still be arriving from the peers while the forward pass runs (the fused data-parallel
        # backend overlaps its all-gather with it, parallel/fused_dp.py): every block waits for its own slice of
        # the flat buffer right before its first kernel. The lists are in flat order, so each gate is monotone.
        self.gate_params = {id(layer): list(layer.parameters()) for layer in self.bert.encoder.layer}
        self.gate_params['emb'] = list(self.bert.embeddings.parameters())
        buffer.gated_forward = True
        if hasattr(model, 'cls') and hasattr(model, 'bert') and model.__class__.__name__ == 'BertForPreTraining':
            model.native = self
            # the tied decoder / word-embedding matrix gets two gradient contributions per backward
            # (decoder wgrad, then the embedding scatter): report it ready only after the second
            buffer.set_expected_notifications(self.bert.embeddings.word_embeddings.weight, 2)

    def _pack(self, params, shape):
        buf, i0 = params[0]._hs_flat
        start = buf.offsets[i0]
        n = sum(p.numel() for p in params)
        for k, p in enumerate(params):
            b, i = p._hs_flat
            expect = start + sum(q.numel() for q in params[:k])
            if b is not buf or buf.offsets[i] != expect:
                raise RuntimeError('q/k/v parameters are not adjacent in the flat buffer; build the buffer '
                                   'with model.param_order()')
        return buf.flat_param[start:start + n].view(shape)

    # -- encoder ------------------------------------------------------------------------------------
    def layer_forward(self, layer, x, add_mask):
        """x: [B, S, H] (or [M, H] with self._bs set); add_mask: [B, S] additive fp32 mask."""
        B, S = self._bs
        H = self.H
        self.buffer.gate(self.gate_params[id(layer)])
        x2 = x.reshape(B * S, H)
        training = layer.training
        qkv_w, qkv_b, ws, bs = self.packed[id(layer)]
        if self.fused_layer and self.attn_impl is not None and \
                self.attn_impl.supports(B, S, self.heads, H // self.heads, x2.dtype):
            y = EncoderLayerFn.apply(x2.contiguous(), add_mask.contiguous(), layer.output.LayerNorm.weight, self, layer,
                                     training)
            return y.view(x.shape)
        qkv = linear(x2, qkv_w, qkv_b, ws, bs)
        p_attn = self.p_attn if training else 0.0
        if self.attn_impl is not None and self.attn_impl.supports(B, S, self.heads, H // self.heads, qkv.dtype):
            ctxv = self.attn_impl.attention(qkv, add_mask, B, S, self.heads, p_attn, self.sites[(id(layer), 'attn')])
        else:
            if not getattr(self, '_warned_composite', False):
                self._warned_composite = True
                print('| WARNING: no fused attention kernel for sequence length {} / head size {} (needs a multiple '
                      'of 16 up to 512 and head size 64): using the PyTorch attention composite'.format(
                          S, H // self.heads), flush=True)
            ctxv = attention_composite(qkv, add_mask, B, S, self.heads, p_attn, training)
        so = layer.attention.output
        ao = linear(ctxv, so.dense.weight, so.dense.bias, [so.dense.weight], [so.dense.bias])
        p_h = self.p_hidden if training else 0.0
        x1 = drop_add_ln(ao, x2, so.LayerNorm, p_h, self.sites[(id(layer), 'ao')])
        inter, out = layer.intermediate.dense_act, layer.output
        f = FfnFn.apply(x1, inter.weight, inter.bias, out.dense.weight, out.dense.bias,
                        inter.weight, inter.bias, out.dense.weight, out.dense.bias)
        x2o = drop_add_ln(f, x1, out.LayerNorm, p_h, self.sites[(id(layer), 'ffn')])
        return x2o.view(x.shape)

    def model_forward(self, bert, input_ids, token_type_ids, attention_mask, output_all_encoded_layers,
                      checkpoint_activations):
        B, S = input_ids.shape
        self._bs = (B, S)
        emb = bert.embeddings
        ids = input_ids.reshape(-1).contiguous()
        seg = token_type_ids.reshape(-1).to(ids.dtype).contiguous()
        self.buffer.gate(self.gate_params['emb'])
        plan, self._emb_plan = getattr(self, '_emb_plan', None), None
        if plan is not None and (plan[0].numel() != ids.numel() or plan[0].device != ids.device):
            plan = None
        z = EmbeddingSumFn.apply(emb.word_embeddings.weight, emb.position_embeddings.weight,
                                 emb.token_type_embeddings.weight, ids, seg, emb.word_embeddings.weight,
                                 emb.position_embeddings.weight, emb.token_type_embeddings.weight, B, S, plan)
        p_h = self.p_hidden if bert.training else 0.0
        x = drop_add_ln(z, None, emb.LayerNorm, p_h, self.site_emb, post=True).view(B, S, self.H)
        add_mask = ((1.0 - attention_mask.to(torch.float32)) * -10000.0).contiguous()
        layers = bert.encoder(x, add_mask, output_all_encoded_layers=output_all_encoded_layers,
                              checkpoint_activations=checkpoint_activations)
        seq = layers[-1]
        pool = bert.pooler.dense_act
        if self.buffer.param_gate is not None:
            self.buffer.param_gate(self.buffer.total)    # pooler + whatever head follows: everything must be in
        # pooler reads the CLS rows in place: row stride S*H, no gather copy
        pooled = linear(seq.view(B * S, self.H)[::S], pool.weight, pool.bias, [pool.weight], [pool.bias], act=ACT_TANH,
                        ldx=S * self.H)
        if not output_all_encoded_layers:
            layers = layers[-1]
        return layers, pooled

    # -- pre-training heads ---------------------------------------------------------------------------
    def pretraining_loss(self, model, input_ids, token_type_ids, attention_mask, masked_lm_labels,
                         next_sentence_label, checkpoint_activations, mlm_rows=None, mlm_row_labels=None,
                         emb_plan=None):
        self._emb_plan = emb_plan      # (order, run_end) of the deterministic word-embedding scatter, host-made
        seq, pooled = model.bert(input_ids, token_type_ids, attention_mask, output_all_encoded_layers=False,
                                 checkpoint_activations=checkpoint_activations)
        B, S = input_ids.shape
        H = self.H
        labels = masked_lm_labels.reshape(-1)
        if model.mlm_all_positions:
            hidden, lab = seq.reshape(B * S, H), labels.contiguous()
        else:
            # labelled rows only. The task normally supplies a fixed-capacity index computed on the host
            # (no sync, static shapes); without it the index is derived here with a nonzero().
            if mlm_rows is not None:
                rows, lab = mlm_rows.contiguous(), mlm_row_labels.contiguous()
            else:
                rows = torch.nonzero(labels != -1, as_tuple=False).squeeze(1).to(torch.int32)
                lab = labels.index_select(0, rows.long()).contiguous()
            hidden = GatherRowsFn.apply(seq.reshape(B * S, H), rows)
        pred = model.cls.predictions
        tr = pred.transform
        t = linear(hidden, tr.dense_act.weight, tr.dense_act.bias, [tr.dense_act.weight], [tr.dense_act.bias], act=ACT_GELU)
        t = drop_add_ln(t, None, tr.LayerNorm, 0.0, 0)
        mlm_loss = LinearXentFn.apply(t, pred.decoder.weight, pred.bias, lab, [pred.decoder.weight], [pred.bias])
        nsp = model.cls.seq_relationship
        nsp_loss = LinearXentFn.apply(pooled.contiguous(), nsp.weight, nsp.bias, next_sentence_label.reshape(-1).contiguous(),
                                      [nsp.weight], [nsp.bias])
        return mlm_loss + nsp_loss


def attach(model, buffer, args):
    """Route ``model``'s hot ops through the native kernels (no-op for models without a BERT trunk)."""
    from hetseq_b200.models import bert as bm
    trunk = getattr(model, 'bert', None)
    if trunk is None and hasattr(model, 'head'):
        trunk = getattr(model.head, 'bert', None)
    if not isinstance(trunk, bm.BertModel):
        return None
    if buffer.param_dtype not in (torch.float16, torch.bfloat16):
        # fp32 training keeps the PyTorch ops (the tcgen05 kernels take 16-bit operands)
        return None
    owner = model if hasattr(model, 'bert') else model.head
    return NativeBert(owner, buffer, args)
