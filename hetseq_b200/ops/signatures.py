"""ctypes signatures of the C ABI (include/hetseq_b200.h), one entry per extern "C" function
besides hs_gemm. Kept in one table so a stale library is detected at load time."""
import ctypes

vp, ci, cf, cl, cu = ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_longlong, ctypes.c_uint

SIGNATURES = {
    'hs_sumsq': [vp, ci, cl, ci, cl, vp, vp, ci, vp],
    'hs_adam_flat': [vp, vp, vp, vp, vp, ci, cl, ci, cl, cf, cf, cf, cf, cf, cf, cf, vp, vp, vp, vp],
    'hs_ln_fwd': [vp, vp, vp, vp, vp, vp, vp, vp, vp, ci, ci, cf, cf, vp, cu, ci, ci, vp],
    'hs_ln_bwd': [vp, vp, vp, vp, vp, vp, vp, vp, vp, ci, ci, ci, cf, vp, cu, ci, ci, vp],
    'hs_colsum': [vp, ci, ci, ci, vp, ci, ci, vp],
    'hs_xent': [vp, ci, vp, ci, ci, ci, vp, vp, ci, ci, ci, vp],
    'hs_gather_rows': [vp, vp, ci, vp, ci, ci, vp],
    'hs_scatter_add_rows': [vp, vp, ci, vp, ci, ci, ci, vp],
    'hs_embed_gather_sum': [vp, vp, vp, vp, vp, ci, vp, ci, ci, ci, ci, ci, ci, vp],
    'hs_embed_scatter_word': [vp, vp, ci, vp, ci, ci, ci, ci, vp],
    'hs_sum_over_batch': [vp, ci, cl, vp, ci, ci, vp],
    'hs_colsum_rows_where': [vp, vp, ci, cl, ci, ci, vp, ci, vp, ci, vp],
    'hs_attention_fwd': [vp, vp, vp, vp, ci, ci, ci, ci, cf, vp, cu, ci, vp],
    'hs_attention_bwd': [vp, vp, vp, vp, vp, vp, ci, ci, ci, ci, cf, vp, cu, ci, vp],
    'hs_adadelta_flat': [vp, vp, vp, vp, vp, ci, cl, ci, cl, cf, cf, cf, cf, cf, vp, vp, vp, vp],
}
