"""Autograd binding of the fused tcgen05 attention kernels (csrc/attention_sm100.cu)."""
from __future__ import annotations

import torch

from hetseq_b200.ops import lib, rng


def supports(B, S, heads, head_dim, dtype) -> bool:
    return dtype in (torch.float16, torch.bfloat16) and bool(lib.load().hs_attention_supported(int(S), int(head_dim)))


class AttentionFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, add_mask, B, S, heads, p, site):
        L = lib.load()
        M, H3 = qkv.shape
        H = H3 // 3
        out = torch.empty(M, H, dtype=qkv.dtype, device=qkv.device)
        lse = torch.empty(B * heads * S, dtype=torch.float32, device=qkv.device)
        seed = rng.seed_tensor(qkv.device) if p > 0 else None
        lib.check(L.hs_attention_fwd(qkv.data_ptr(), add_mask.data_ptr(), out.data_ptr(), lse.data_ptr(), B, S, heads, H,
                                     float(p), lib.ptr(seed), int(site), lib.dtype_code(qkv.dtype), lib.stream_ptr()),
                  'hs_attention_fwd')
        ctx.save_for_backward(qkv, add_mask, out, lse)
        ctx.meta = (B, S, heads, H, p, site)
        return out

    @staticmethod
    def backward(ctx, dout):
        qkv, add_mask, out, lse = ctx.saved_tensors
        B, S, heads, H, p, site = ctx.meta
        L = lib.load()
        dout = dout.contiguous()
        dqkv = torch.empty_like(qkv)
        seed = rng.seed_tensor(qkv.device) if p > 0 else None
        lib.check(L.hs_attention_bwd(qkv.data_ptr(), add_mask.data_ptr(), out.data_ptr(), dout.data_ptr(), lse.data_ptr(),
                                     dqkv.data_ptr(), None, 0, B, S, heads, H, float(p), lib.ptr(seed), int(site),
                                     lib.dtype_code(qkv.dtype), lib.stream_ptr()), 'hs_attention_bwd')
        return dqkv, None, None, None, None, None, None


def attention(qkv, add_mask, B, S, heads, p, site):
    """qkv: [B*S, 3H] contiguous 16-bit; add_mask: [B, S] fp32 additive; returns [B*S, H]."""
    return AttentionFn.apply(qkv.contiguous(), add_mask.contiguous(), int(B), int(S), int(heads), float(p), int(site))
