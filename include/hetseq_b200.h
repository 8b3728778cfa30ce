/* hetseq_b200 -- C ABI of the sm_100a kernel library (libhetseq_b200.so).
 *
 * Every entry point takes raw device pointers plus the CUDA stream to launch on and returns 0 on
 * success; on failure it returns non-zero and hs_last_error() describes why. No entry point
 * synchronises the device. Python reaches these through ctypes (hetseq_b200/ops/lib.py); any other
 * host program can link the library directly.
 *
 * The reference (BismarckBamfo/hetseq) has no native GPU code at all: each
 * declaration cites the Python/ATen path of the reference whose math it reproduces.
 */
#ifndef HETSEQ_B200_H_
#define HETSEQ_B200_H_

#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__)
#define HS_API __attribute__((visibility("default")))
#else
#define HS_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

typedef void* hs_stream_t; /* cudaStream_t */

enum hs_dtype { HS_F16 = 0, HS_BF16 = 1, HS_F32 = 2 };

/* ---- library ---------------------------------------------------------------------------------- */
HS_API const char* hs_last_error(void);
HS_API int hs_version(void);
/* Number of kernel launches issued through this library since load (all threads). bench.py reports
 * the delta across its timed region as "gpu_launches". */
HS_API uint64_t hs_launch_count(void);

/* ---- tcgen05 GEMM ----------------------------------------------------------------------------- */
enum hs_gemm_act {
  HS_ACT_NONE = 0,
  HS_ACT_GELU = 1,      /* x*0.5*(1+erf(x/1.41421))  (bert_modeling.py:104-116) */
  HS_ACT_TANH = 2,      /* pooler (bert_modeling.py:514-515) */
  HS_ACT_DGELU_MUL = 3  /* out = acc * gelu'(aux_in): dgrad through the FFN activation */
};

typedef struct hs_gemm_args {
  /* C[M,N] = epilogue( sum_k A(m,k) * B(n,k) ), fp32 accumulation in TMEM.
   * a_mn_major = 0: A stored row-major [M, lda] (k contiguous); 1: stored [K, lda] (m contiguous).
   * b_mn_major = 0: B stored row-major [N, ldb] (k contiguous); 1: stored [K, ldb] (n contiguous).
   * Leading dimensions are in elements and must be multiples of 8; base pointers 16-byte aligned.
   * Forward linear (X*W^T): a=0,b=0. dgrad (dY*W): a=0,b=1. wgrad (dY^T*X): a=1,b=1. */
  const void* A;
  const void* B;
  void* C;
  int M, N, K;
  int lda, ldb, ldc;
  int a_mn_major, b_mn_major;
  int in_dtype;  /* HS_F16 | HS_BF16 (A, B, bias, aux) */
  int out_dtype; /* HS_F16 | HS_BF16 | HS_F32 (C); 16-bit output must equal in_dtype */
  const void* bias; /* [N] in in_dtype, or NULL */
  int act;          /* hs_gemm_act */
  void* aux_out;    /* HS_ACT_GELU: pre-activation (acc+bias) stored here when non-NULL, layout of C */
  const void* aux_in; /* HS_ACT_DGELU_MUL: pre-activation, in_dtype, row stride ldaux */
  int ldaux;
  int accumulate;   /* 1: C += result (read-modify-write); used by wgrad into the flat grad buffer */
  float alpha;      /* scales the accumulator before bias/activation */
} hs_gemm_args;

/* Replaces F.linear/cuBLAS at bert_modeling.py:352-354,387-391,409-413,423-427,514-515,525-528,547
 * and the autograd GEMMs behind them (SURVEY.md K1,K5-K7,K9-K11,K14). */
HS_API int hs_gemm(const hs_gemm_args* args, hs_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* HETSEQ_B200_H_ */
