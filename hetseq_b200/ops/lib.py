"""ctypes binding of ``libhetseq_b200.so`` (C ABI: include/hetseq_b200.h).

The library is built in-tree by ``hetseq_b200/ops/build.py`` (``__graft_entry__.build()``); it has
no torch dependency, tensors cross the boundary as raw device pointers plus the current CUDA
stream. ``load()`` raises if the library is missing -- on a GPU box a missing kernel library is an
error, not a reason to fall back to library kernels.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

_lib: Optional[ctypes.CDLL] = None

HS_F16, HS_BF16, HS_F32 = 0, 1, 2
ACT_NONE, ACT_GELU, ACT_TANH, ACT_DGELU_MUL, ACT_ADD_AUX, ACT_GELU_DAUX, ACT_MUL_AUX = 0, 1, 2, 3, 4, 5, 6

_DT = {torch.float16: HS_F16, torch.bfloat16: HS_BF16, torch.float32: HS_F32}


def dtype_code(dt: torch.dtype) -> int:
    try:
        return _DT[dt]
    except KeyError:
        raise TypeError('unsupported dtype for native kernels: {}'.format(dt)) from None


class GemmArgs(ctypes.Structure):
    _fields_ = [
        ('A', ctypes.c_void_p), ('B', ctypes.c_void_p), ('C', ctypes.c_void_p),
        ('M', ctypes.c_int), ('N', ctypes.c_int), ('K', ctypes.c_int),
        ('lda', ctypes.c_int), ('ldb', ctypes.c_int), ('ldc', ctypes.c_int),
        ('a_mn_major', ctypes.c_int), ('b_mn_major', ctypes.c_int),
        ('in_dtype', ctypes.c_int), ('out_dtype', ctypes.c_int),
        ('bias', ctypes.c_void_p), ('act', ctypes.c_int),
        ('aux_out', ctypes.c_void_p), ('aux_in', ctypes.c_void_p), ('ldaux', ctypes.c_int),
        ('accumulate', ctypes.c_int), ('alpha', ctypes.c_float), ('alpha_dev', ctypes.c_void_p),
        ('colsum_out', ctypes.c_void_p), ('colsum_accumulate', ctypes.c_int),
    ]


def available() -> bool:
    from hetseq_b200.ops.build import CUDA_LIB
    return CUDA_LIB.exists()


def load() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    from hetseq_b200.ops.build import CUDA_LIB
    if not CUDA_LIB.exists():
        raise RuntimeError(
            'native kernel library {} not found: run `python -c "import __graft_entry__ as g; g.build()"` '
            '(or `python -m hetseq_b200.ops.build`) first, or pass --kernels torch'.format(CUDA_LIB))
    lib = ctypes.CDLL(str(CUDA_LIB))
    lib.hs_last_error.restype = ctypes.c_char_p
    lib.hs_version.restype = ctypes.c_int
    lib.hs_launch_count.restype = ctypes.c_uint64
    for fn in ('hs_sumsq_scratch_elems',):
        getattr(lib, fn).restype = ctypes.c_int
    lib.hs_colsum_scratch_elems.argtypes = [ctypes.c_int, ctypes.c_int]
    lib.hs_embed_backward_scratch_elems.argtypes = [ctypes.c_int, ctypes.c_int]
    lib.hs_embed_backward_scratch_elems.restype = ctypes.c_longlong
    lib.hs_attention_supported.argtypes = [ctypes.c_int, ctypes.c_int]
    lib.hs_dp_signal_words.argtypes = [ctypes.c_int, ctypes.c_int]
    lib.hs_dp_max_gather_buckets.restype = ctypes.c_int
    _declare(lib)
    _lib = lib
    return lib


def _declare(lib) -> None:
    """argtypes/restype for every entry point (all return int: 0 = ok)."""
    vp, ci, cf, cl, cu64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_longlong, ctypes.c_uint64
    sigs = {
        'hs_gemm': [ctypes.POINTER(GemmArgs), vp],
    }
    from hetseq_b200.ops import signatures
    sigs.update(signatures.SIGNATURES)
    for name, argtypes in sigs.items():
        fn = getattr(lib, name, None)
        if fn is None:
            raise RuntimeError('native library is stale: missing symbol {} (rebuild it)'.format(name))
        fn.argtypes = argtypes
        fn.restype = ctypes.c_int


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError('{} failed: {}'.format(what, load().hs_last_error().decode()))


def stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def launch_count() -> int:
    return int(load().hs_launch_count()) if _lib is not None or available() else 0


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()
