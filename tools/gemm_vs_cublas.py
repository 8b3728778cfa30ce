#!/usr/bin/env python
"""The BERT-large GEMM shapes through the tcgen05 kernel and through cuBLAS (torch.matmul) on the same RANDOM
fp16 operands: CUDA events, L2 flushed (256 MB write) between timed launches, best and mean of N runs, and each
against the measured cuBLAS burst peak of MEASURED_PEAKS.json. Markdown on stdout:

    python tools/gemm_vs_cublas.py > gpurun_out/gemm_vs_cublas.md          (on a B200)
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from hetseq_b200.ops import native
from hetseq_b200.ops.native import ACT_GELU_DAUX, ACT_NONE

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
M0, H, I, V = 4096, 1024, 4096, 30522        # 32 x 128 tokens, BERT-large
# name, kind, M, N, K   (kind: fwd = X W^T, dgrad = dY W, wgrad = dY^T X with fp32 output)
SHAPES = [
    ('QKV forward (+bias)', 'fwd', M0, 3 * H, H),
    ('attention-out forward (+bias)', 'fwd', M0, H, H),
    ('FFN1 forward (+bias, GELU, GELU\')', 'fwd_gelu', M0, I, H),
    ('FFN2 forward (+bias)', 'fwd', M0, H, I),
    ('QKV dgrad', 'dgrad', M0, H, 3 * H),
    ('attention-out dgrad', 'dgrad', M0, H, H),
    ('FFN1 dgrad', 'dgrad', M0, H, I),
    ('FFN2 dgrad', 'dgrad', M0, I, H),
    ('QKV wgrad (fp32 out)', 'wgrad', 3 * H, H, M0),
    ('attention-out wgrad (fp32 out)', 'wgrad', H, H, M0),
    ('FFN1 wgrad (fp32 out)', 'wgrad', I, H, M0),
    ('FFN2 wgrad (fp32 out)', 'wgrad', H, I, M0),
    ('MLM decoder forward (640 labelled rows)', 'fwd', 640, V, H),
    ('square 8192^3', 'fwd', 8192, 8192, 8192),
]


def timed(fn, flush, iters=12):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for i in range(iters + 3):
        flush.fill_(i)                  # evict L2 (126 MB) between launches
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        if i >= 3:
            ts.append(e0.elapsed_time(e1))
    return min(ts), sum(ts) / len(ts)


def main():
    torch.manual_seed(0)
    dev = torch.device('cuda', 0)
    peak = 1699.2
    try:
        peak = float(json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))['bf16_tflops'])
    except Exception:
        pass
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    print('# tcgen05 GEMM vs cuBLAS on the BERT-large shapes (random fp16 operands)\n')
    print('`tools/gemm_vs_cublas.py` on one B200: CUDA events around single launches, L2 flushed between launches, best of 12')
    print('(mean in brackets). "of peak" = best / %.0f TFLOP/s, the measured cuBLAS bf16 burst rate of this pool' % peak)
    print('(`MEASURED_PEAKS.json`). cuBLAS = `torch.matmul` / `torch.addmm` in fp16 on the same tensors; the tcgen05 column')
    print('includes the fused epilogue named in the first column, the cuBLAS column only bias where stated.\n')
    print('| GEMM | M x N x K | tcgen05 us | TFLOP/s | of peak | cuBLAS us | TFLOP/s | tcgen05 / cuBLAS |')
    print('| --- | --- | --- | --- | --- | --- | --- | --- |')
    for name, kind, M, N, K in SHAPES:
        fl = 2.0 * M * N * K
        if kind in ('fwd', 'fwd_gelu'):
            x = torch.randn(M, K, device=dev).half()
            w = (torch.randn(N, K, device=dev) * 0.05).half()
            bias = torch.randn(N, device=dev).half()
            ld = (N + 127) // 128 * 128 if N % 8 else N
            y = torch.empty(M, ld, device=dev, dtype=torch.float16)
            aux = torch.empty(M, N, device=dev, dtype=torch.float16) if kind == 'fwd_gelu' else None
            act = ACT_GELU_DAUX if kind == 'fwd_gelu' else ACT_NONE

            def ours():
                native.gemm(x, w, y, M, N, K, lda=K, ldb=K, ldc=ld, bias=bias, act=act, aux_out=aux, ldaux=N)

            def ref():
                torch.addmm(bias, x, w.t())
        elif kind == 'dgrad':
            dy = torch.randn(M, K, device=dev).half()          # GEMM K = the layer's output width
            w = (torch.randn(K, N, device=dev) * 0.05).half()   # [out, in] weight, read MN-major
            y = torch.empty(M, N, device=dev, dtype=torch.float16)

            def ours():
                native.gemm(dy, w, y, M, N, K, lda=K, ldb=N, ldc=N, b_mn=True)

            def ref():
                torch.matmul(dy, w)
        else:
            dy = torch.randn(K, M, device=dev).half()           # [tokens, out]
            x = torch.randn(K, N, device=dev).half()            # [tokens, in]
            y = torch.empty(M, N, device=dev, dtype=torch.float32)

            def ours():
                native.gemm(dy, x, y, M, N, K, lda=M, ldb=N, ldc=N, a_mn=True, b_mn=True)

            def ref():
                torch.matmul(dy.t(), x)
        tb, tm = timed(ours, flush)
        cb, cm = timed(ref, flush)
        print('| %s | %d x %d x %d | %.1f (%.1f) | %.0f | %.0f %% | %.1f (%.1f) | %.0f | %.2f x |' % (
            name, M, N, K, tb * 1e3, tm * 1e3, fl / tb / 1e9, 100 * fl / tb / 1e9 / peak, cb * 1e3, cm * 1e3,
            fl / cb / 1e9, cb / tb))
        del ours, ref
        torch.cuda.empty_cache()


if __name__ == '__main__':
    main()
