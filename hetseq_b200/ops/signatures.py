"""ctypes signatures of the C ABI (include/hetseq_b200.h), one entry per extern "C" function
besides hs_gemm. Kept in one table so a stale library is detected at load time."""
import ctypes

vp, ci, cf, cl = ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_longlong

SIGNATURES = {}
