"""ctypes signatures of the C ABI (include/hetseq_b200.h), one entry per extern "C" function
besides hs_gemm. Kept in one table so a stale library is detected at load time."""
import ctypes

vp, ci, cf, cl, cu = ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_longlong, ctypes.c_uint
csz = ctypes.c_size_t

SIGNATURES = {
    'hs_sumsq': [vp, ci, cl, ci, cl, vp, vp, ci, vp],
    'hs_adam_flat': [vp, vp, vp, vp, vp, ci, cl, ci, cl, cf, cf, cf, cf, cf, cf, cf, vp, vp, vp, vp],
    'hs_ln_fwd': [vp, vp, vp, vp, vp, vp, vp, vp, vp, ci, ci, cf, cf, vp, cu, ci, ci, vp],
    'hs_ln_bwd': [vp, vp, vp, vp, vp, vp, vp, vp, vp, ci, vp, ci, ci, ci, cf, vp, cu, ci, ci, vp],
    'hs_colsum': [vp, ci, ci, ci, vp, ci, ci, vp],
    'hs_xent': [vp, ci, vp, ci, ci, ci, vp, vp, ci, ci, vp, ci, vp],
    'hs_gather_rows': [vp, vp, ci, vp, ci, ci, vp],
    'hs_scatter_add_rows': [vp, vp, ci, vp, ci, ci, ci, vp],
    'hs_embed_gather_sum': [vp, vp, vp, vp, vp, ci, vp, ci, ci, ci, ci, ci, ci, vp],
    'hs_embed_scatter_word': [vp, vp, ci, vp, ci, ci, ci, ci, vp],
    'hs_embed_backward_sorted': [vp, vp, vp, ci, vp, vp, vp, vp, ci, vp, ci, ci, ci, ci, vp],
    'hs_sum_over_batch': [vp, ci, cl, vp, ci, ci, vp],
    'hs_colsum_rows_where': [vp, vp, ci, cl, ci, ci, vp, ci, vp, ci, vp],
    'hs_attention_fwd': [vp, vp, vp, vp, ci, ci, ci, ci, cf, vp, cu, ci, vp],
    'hs_attention_bwd': [vp, vp, vp, vp, vp, vp, vp, ci, ci, ci, ci, ci, cf, vp, cu, ci, vp],
    'hs_ipc_alloc': [csz, ctypes.POINTER(vp), ctypes.c_char_p],
    'hs_ipc_open': [ctypes.c_char_p, ctypes.POINTER(vp)],
    'hs_ipc_close': [vp],
    'hs_ipc_free': [vp],
    'hs_dp_set_timeout': [ctypes.c_double],
    'hs_dp_reduce_scatter': [vp, vp, vp, ci, ci, cl, cl, ci, ci, cu, vp, vp, ci, vp],
    'hs_dp_norm_exchange': [vp, ci, ci, cu, vp, vp, cf, vp, vp, cf, vp, vp, vp, vp],
    'hs_dp_update_allgather': [ci, vp, vp, vp, vp, vp, vp, vp, ci, ci, cl, ci, ci, cf, cf, cf, cf, cf, cf, vp, vp, cu,
                               vp, ci, vp, vp, ci, ci, vp],
    'hs_dp_update_local': [ci, vp, vp, vp, vp, vp, ci, ci, cl, ci, ci, cf, cf, cf, cf, cf, cf, vp, vp, ci, vp],
    'hs_dp_allgather': [vp, vp, vp, ci, ci, cl, ci, vp, cu, vp, ci, vp, vp, ci, ci, vp],
    'hs_dp_gather_bucket_of': [cl, ci, cl],
    'hs_dp_gather_count': [cl, ci],
    'hs_dp_wait_params': [vp, ci, ci, ci, ci, cu, vp, ci, vp],
    'hs_dp_emulate_step': [ci, ci, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, cl, cl, ci, ci, ci, ci,
                           cf, cf, cf, cf, cf, cf, cf, cf, cu, ci, vp],
    'hs_sumsq_noalloc': [vp, ci, cl, ci, cl, vp, ci, vp],
    'hs_adadelta_flat': [vp, vp, vp, vp, vp, ci, cl, ci, cl, cf, cf, cf, cf, cf, vp, vp, vp, vp],
}
