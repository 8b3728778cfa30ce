"""Thin functional wrappers over the C ABI (one Python function per kernel entry point).

Tensors must be CUDA, contiguous where stated; every call launches on the current torch stream and
returns immediately. Shape / dtype mistakes raise here or in the launcher (never a silent fallback).
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

from hetseq_b200.ops import lib
from hetseq_b200.ops.lib import (ACT_ADD_AUX, ACT_DGELU_MUL, ACT_GELU, ACT_GELU_DAUX, ACT_MUL_AUX,  # noqa: F401
                                 ACT_NONE, ACT_TANH)

_scratch = {}


def scratch(device, n_floats: int, tag: str = 'default') -> torch.Tensor:
    """Per-(device, tag) fp32 scratch that only grows; stream-ordered reuse is safe because every
    kernel that consumes it runs on the same stream as the one that fills it."""
    key = (device.index, tag)
    t = _scratch.get(key)
    if t is None or t.numel() < n_floats:
        t = torch.empty(max(n_floats, 1024), dtype=torch.float32, device=device)
        _scratch[key] = t
    return t


def gemm(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor, M: int, N: int, K: int, *, lda: int, ldb: int,
         ldc: int, a_mn: bool = False, b_mn: bool = False, bias: Optional[torch.Tensor] = None, act: int = ACT_NONE,
         aux_out: Optional[torch.Tensor] = None, aux_in: Optional[torch.Tensor] = None, ldaux: int = 0,
         accumulate: bool = False, alpha: float = 1.0, alpha_dev: Optional[torch.Tensor] = None,
         colsum_out: Optional[torch.Tensor] = None, colsum_accumulate: bool = False) -> torch.Tensor:
    """out[M,N] = epilogue(alpha * A(m,k) B(n,k)); see include/hetseq_b200.h for operand layouts."""
    L = lib.load()
    g = lib.GemmArgs()
    g.A, g.B, g.C = a.data_ptr(), b.data_ptr(), out.data_ptr()
    g.M, g.N, g.K = M, N, K
    g.lda, g.ldb, g.ldc = lda, ldb, ldc
    g.a_mn_major, g.b_mn_major = int(a_mn), int(b_mn)
    g.in_dtype = lib.dtype_code(a.dtype)
    g.out_dtype = lib.dtype_code(out.dtype)
    g.bias = lib.ptr(bias)
    g.act = act
    g.aux_out, g.aux_in, g.ldaux = lib.ptr(aux_out), lib.ptr(aux_in), ldaux
    g.accumulate = int(accumulate)
    g.alpha = float(alpha)
    g.alpha_dev = lib.ptr(alpha_dev)
    g.colsum_out, g.colsum_accumulate = lib.ptr(colsum_out), int(colsum_accumulate)
    lib.check(L.hs_gemm(ctypes.byref(g), lib.stream_ptr()), 'hs_gemm')
    return out


def ln_fwd(x, bias, residual, gamma, beta, *, eps=1e-12, p=0.0, seed=None, site=0, post_dropout=False,
           save_z=True, save_stats=True):
    """Returns (y, z, mean, rstd); x is [M, H] contiguous 16-bit."""
    L = lib.load()
    M, H = x.shape
    y = torch.empty_like(x)
    z = torch.empty_like(x) if save_z else None
    mean = torch.empty(M, dtype=torch.float32, device=x.device) if save_stats else None
    rstd = torch.empty(M, dtype=torch.float32, device=x.device) if save_stats else None
    lib.check(L.hs_ln_fwd(x.data_ptr(), lib.ptr(bias), lib.ptr(residual), gamma.data_ptr(), beta.data_ptr(),
                          y.data_ptr(), lib.ptr(z), lib.ptr(mean), lib.ptr(rstd), M, H, float(eps), float(p),
                          lib.ptr(seed), int(site), int(post_dropout), lib.dtype_code(x.dtype), lib.stream_ptr()),
              'hs_ln_fwd')
    return y, z, mean, rstd


def ln_bwd(dy, z, gamma, mean, rstd, dgamma, dbeta, accumulate_params, *, p=0.0, seed=None, site=0,
           post_dropout=False, need_dx=True, dbias=None, accumulate_bias=False):
    """Returns (dz, dx). dgamma/dbeta are fp32 views (main grads) written or accumulated in place;
    ``dbias`` (optional fp32 [H]) receives the column sums of dx."""
    L = lib.load()
    M, H = dy.shape
    dz = torch.empty_like(dy)
    dx = torch.empty_like(dy) if (need_dx and p > 0.0 and not post_dropout) else None
    lib.check(L.hs_ln_bwd(dy.data_ptr(), z.data_ptr(), gamma.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                          dz.data_ptr(), lib.ptr(dx), lib.ptr(dgamma), lib.ptr(dbeta), int(accumulate_params),
                          lib.ptr(dbias), int(accumulate_bias), M, H, float(p), lib.ptr(seed), int(site), int(post_dropout),
                          lib.dtype_code(dy.dtype), lib.stream_ptr()), 'hs_ln_bwd')
    return dz, (dx if dx is not None else dz)


def colsum(x, M, N, ld, out, accumulate):
    L = lib.load()
    lib.check(L.hs_colsum(x.data_ptr(), M, N, ld, out.data_ptr(), int(accumulate),
                          lib.dtype_code(x.dtype), lib.stream_ptr()), 'hs_colsum')
    return out


def xent(logits, ld, labels, rows, V, loss, accumulate_loss, write_grad=True, inv_count_out=None):
    """In-place: logits become d(loss)/d(logits); returns per-row losses. With ``inv_count_out`` (1-element fp32
    tensor) the stored gradient is softmax - onehot without the 1/n_valid, which lands in ``inv_count_out``."""
    L = lib.load()
    row_loss = torch.empty(rows, dtype=torch.float32, device=logits.device)
    lib.check(L.hs_xent(logits.data_ptr(), ld, labels.data_ptr(), int(labels.dtype == torch.int64), rows, V,
                        row_loss.data_ptr(), lib.ptr(loss), int(accumulate_loss), int(write_grad),
                        lib.ptr(inv_count_out), lib.dtype_code(logits.dtype), lib.stream_ptr()), 'hs_xent')
    return row_loss


def gather_rows(src2d, idx, rows):
    L = lib.load()
    H = src2d.shape[1]
    dst = torch.empty(rows, H, dtype=src2d.dtype, device=src2d.device)
    lib.check(L.hs_gather_rows(src2d.data_ptr(), idx.data_ptr(), int(idx.dtype == torch.int64), dst.data_ptr(), rows,
                               H * src2d.element_size(), lib.stream_ptr()), 'hs_gather_rows')
    return dst


def scatter_add_rows(src2d, idx, dst2d):
    L = lib.load()
    rows, H = src2d.shape
    lib.check(L.hs_scatter_add_rows(src2d.data_ptr(), idx.data_ptr(), int(idx.dtype == torch.int64), dst2d.data_ptr(),
                                    rows, H, lib.dtype_code(src2d.dtype), lib.stream_ptr()), 'hs_scatter_add_rows')
    return dst2d


def embed_gather_sum(W, P, T, ids, seg, M, S):
    L = lib.load()
    H = W.shape[1]
    z = torch.empty(M, H, dtype=W.dtype, device=W.device)
    lib.check(L.hs_embed_gather_sum(W.data_ptr(), P.data_ptr(), T.data_ptr(), ids.data_ptr(), seg.data_ptr(),
                                    int(ids.dtype == torch.int64), z.data_ptr(), M, S, H, W.shape[0], T.shape[0],
                                    lib.dtype_code(W.dtype), lib.stream_ptr()), 'hs_embed_gather_sum')
    return z


def embed_scatter_word(de, ids, dW):
    L = lib.load()
    M, H = de.shape
    lib.check(L.hs_embed_scatter_word(de.data_ptr(), ids.data_ptr(), int(ids.dtype == torch.int64), dW.data_ptr(), M, H,
                                      dW.shape[0], lib.dtype_code(de.dtype), lib.stream_ptr()), 'hs_embed_scatter_word')


def embed_backward_sorted(de, ids, order, run_end, dW, seg=None, dT=None, accumulate_T=False):
    """dW[id] += sum of the rows of de holding token id -- deterministic: segmented reduction over the positions
    sorted by token id, 16-row segments, one writer and one order of additions per slice of dW. With ``seg`` (token
    types 0 / 1) and ``dT`` ([2, H] fp32) the same pass also produces the token-type gradient."""
    L = lib.load()
    M, H = de.shape
    sc = scratch(de.device, int(L.hs_embed_backward_scratch_elems(M, H)), 'embed_bwd')
    lib.check(L.hs_embed_backward_sorted(de.data_ptr(), ids.data_ptr(), seg.data_ptr() if seg is not None else None,
                                         int(ids.dtype == torch.int64), order.data_ptr(), run_end.data_ptr(),
                                         dW.data_ptr(), dT.data_ptr() if dT is not None else None, int(accumulate_T),
                                         sc.data_ptr(), M, H, dW.shape[0], lib.dtype_code(de.dtype), lib.stream_ptr()),
              'hs_embed_backward_sorted')


def embed_scatter_word_sorted(de, ids, order, run_end, dW):
    """The word-embedding part of ``embed_backward_sorted`` alone."""
    embed_backward_sorted(de, ids, order, run_end, dW)


def embedding_scatter_plan(ids_flat: torch.Tensor):
    """(order, run_end) int32 for ``embed_scatter_word_sorted`` from the token ids, with static-shape tensor ops
    only (usable on the device inside a captured graph, or on the host in the batch-prefetch thread)."""
    M = ids_flat.numel()
    sorted_ids, order = torch.sort(ids_flat.reshape(-1), stable=True)
    idx = torch.arange(M, device=ids_flat.device)
    start = torch.ones(M, dtype=torch.bool, device=ids_flat.device)
    start[1:] = sorted_ids[1:] != sorted_ids[:-1]
    nxt = torch.where(start, idx, torch.full_like(idx, M))                # index of a run start, M elsewhere
    shifted = torch.cat([nxt[1:], torch.full((1,), M, dtype=idx.dtype, device=idx.device)])
    next_start = torch.flip(torch.cummin(torch.flip(shifted, [0]), 0).values, [0])   # first run start after slot i
    run_end = torch.where(start, next_start, torch.zeros_like(next_start))
    return order.to(torch.int32), run_end.to(torch.int32)


def sum_over_batch(x, B, inner, out, accumulate):
    L = lib.load()
    lib.check(L.hs_sum_over_batch(x.data_ptr(), B, inner, out.data_ptr(), int(accumulate), lib.dtype_code(x.dtype),
                                  lib.stream_ptr()), 'hs_sum_over_batch')


def colsum_rows_where(x, sel, want, out, accumulate):
    L = lib.load()
    M, N = x.shape
    sc = scratch(x.device, int(L.hs_colsum_scratch_elems(M, N)), 'colsum')
    lib.check(L.hs_colsum_rows_where(x.data_ptr(), sel.data_ptr(), int(sel.dtype == torch.int64), int(want), M, N,
                                     out.data_ptr(), int(accumulate), sc.data_ptr(), lib.dtype_code(x.dtype),
                                     lib.stream_ptr()), 'hs_colsum_rows_where')
