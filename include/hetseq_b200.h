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
  HS_ACT_DGELU_MUL = 3, /* out = acc * gelu'(aux_in): dgrad through the FFN activation */
  HS_ACT_ADD_AUX = 4,   /* out = acc (+bias) + aux_in: dgrad that also collects the residual-path gradient */
  HS_ACT_GELU_DAUX = 5, /* like HS_ACT_GELU, but aux_out receives gelu'(pre-activation) instead of the
                         * pre-activation: backward then only multiplies (HS_ACT_MUL_AUX) */
  HS_ACT_MUL_AUX = 6    /* out = acc * aux_in */
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
  void* aux_out;    /* HS_ACT_GELU: pre-activation (acc+bias), HS_ACT_GELU_DAUX: gelu'(acc+bias); stored when
                     * non-NULL, layout of C */
  const void* aux_in; /* HS_ACT_DGELU_MUL: pre-activation; HS_ACT_ADD_AUX: addend; HS_ACT_MUL_AUX: factor.
                       * in_dtype, row stride ldaux */
  int ldaux;
  int accumulate;   /* 1: C += result (read-modify-write); used by wgrad into the flat grad buffer */
  float alpha;      /* scales the accumulator before bias/activation */
  const float* alpha_dev; /* optional device scalar multiplied into alpha (e.g. the upstream loss gradient) */
  float* colsum_out;      /* optional fp32 [N]: column sums of C (after the epilogue) are added here -- when C is
                           * the output gradient of a Linear this is its bias gradient, for free */
  int colsum_accumulate;  /* 0: colsum_out is cleared first (on the stream); 1: += */
} hs_gemm_args;

/* Replaces F.linear/cuBLAS at bert_modeling.py:352-354,387-391,409-413,423-427,514-515,525-528,547
 * and the autograd GEMMs behind them (SURVEY.md K1,K5-K7,K9-K11,K14). */
HS_API int hs_gemm(const hs_gemm_args* args, hs_stream_t stream);

/* ---- flat-buffer gradient norm + fused optimizer updates (csrc/optim_sm100.cu) ------------------- */
/* "Owned" elements: chunk c covers [c*stride, c*stride+chunk) of the flat gradient/parameter buffers;
 * optimizer state and fp32 master weights are compact (c*chunk + i). chunk, stride: multiples of 8. */

/* Number of floats of scratch hs_sumsq needs (per call; contents are overwritten). */
HS_API int hs_sumsq_scratch_elems(void);
/* out[0] (= or +=) sum of squares of the owned gradient elements. Two deterministic stages: block
 * partials into `scratch`, then one block reduces them. Replaces the norm loop of optim.py:65-70. */
HS_API int hs_sumsq(const void* grad, int dtype, long long nchunks, int chunk, long long stride,
                    float* scratch, float* out, int accumulate, hs_stream_t stream);
/* Same result as hs_sumsq for callers that cannot allocate device memory, so cannot supply `scratch`:
 * processes whose GPU memory is entirely pre-assigned to their own pool (a late cudaMalloc can fail, and
 * cudaMalloc/cudaFree synchronise the device in the middle of a step loop), and bare C/C++ hosts of this
 * ABI that have no caching allocator to draw from. The block partials go to an array built into the
 * library: 4096 floats (16 KiB) per GPU, held from library load until process exit. The call makes no
 * cudaMalloc/cudaFree and uses no allocating library routine. Limits and rules:
 *   - it needs min(nchunks, 8 * SM count) partials; more than hs_sumsq_noalloc_max_partials() (4096)
 *     is refused with an error -- use hs_sumsq then;
 *   - the first call on each device looks up the array's device address (cudaGetSymbolAddress, with
 *     that device current); make that first call before starting any stream capture;
 *   - all calls on a device share the one array. A call overwrites every partial it later reads, so
 *     stale contents are harmless, but two calls that run concurrently on the GPU (different streams)
 *     race on it: serialise them.
 * No counterpart in the reference (it has no native code and never computes a flat gradient norm). */
HS_API int hs_sumsq_noalloc(const void* grad, int dtype, long long nchunks, int chunk, long long stride,
                            float* out, int accumulate, hs_stream_t stream);
HS_API int hs_sumsq_noalloc_max_partials(void);
/* fairseq Adam on flat memory, math of optim.py:162-231 (decoupled decay before the update, bias
 * correction via step_factor = sqrt(1-b2^t)/(1-b1^t), eps outside the sqrt). The gradient is read
 * once and multiplied by mult_host * (*mult_dev) on the fly -- this is where multiply_grads
 * (optim.py:59-63), the clip coefficient and 1/loss_scale are applied. hyper_dev = [lr, step_factor]
 * overrides the host values when non-NULL; *skip_dev != 0 makes the call a no-op (overflow step). */
HS_API int hs_adam_flat(const float* grad, float* master, float* exp_avg, float* exp_avg_sq, void* param_out,
                        int out_dtype, long long nchunks, int chunk, long long stride, float lr, float beta1,
                        float beta2, float eps, float weight_decay, float step_factor, float mult_host,
                        const float* mult_dev, const float* hyper_dev, const int* skip_dev, hs_stream_t stream);
/* Adadelta with coupled L2 decay, math of optim.py:263-304; same conventions as hs_adam_flat. */
HS_API int hs_adadelta_flat(const float* grad, float* master, float* square_avg, float* acc_delta,
                            void* param_out, int out_dtype, long long nchunks, int chunk, long long stride,
                            float lr, float rho, float eps, float weight_decay, float mult_host,
                            const float* mult_dev, const float* hyper_dev, const int* skip_dev,
                            hs_stream_t stream);

/* ---- fused (bias +) dropout + residual + LayerNorm (csrc/layernorm_sm100.cu) --------------------- */
/* z = residual + dropout(x + bias); y = gamma*(z-mean)/sqrt(var+eps)+beta, TF-style (eps inside the
 * sqrt, biased variance: bert_modeling.py:277-289). post_dropout=1 instead computes y = dropout(LN(x))
 * (the embeddings, bert_modeling.py:313-319). bias/residual/z/mean/rstd may be NULL. Dropout masks are
 * Philox(seed[0..1], site, element) and are recomputed by hs_ln_bwd. H % 8 == 0, H <= 2048.
 * Replaces the dropout + add + ~9 LayerNorm launches of bert_modeling.py:387-391,423-427 (K5,K7,W5,W7). */
HS_API int hs_ln_fwd(const void* x, const void* bias, const void* residual, const void* gamma, const void* beta,
                     void* y, void* z, float* mean, float* rstd, int M, int H, float eps, float dropout_p,
                     const long long* seed, unsigned site, int post_dropout, int dtype, hs_stream_t stream);
/* dz = dLN/dz (residual path), dx = dz * dropout mask (GEMM path, may be NULL), dgamma/dbeta fp32
 * (= or += per accumulate_params); dbias (optional, fp32 [H], = or += per accumulate_bias) receives the
 * column sums of dx, i.e. the bias gradient of the Linear that produced x -- no separate reduction pass.
 * Persistent kernel, one CTA per SM: rows stream through per-warp shared-memory rings filled by bulk
 * async copies; per-column partials live in registers, meet in shared memory, then one atomicAdd per
 * block and column goes straight into the destination (the flat main-grad buffer). */
HS_API int hs_ln_bwd(const void* dy, const void* z, const void* gamma, const float* mean, const float* rstd,
                     void* dz, void* dx, float* dgamma, float* dbeta, int accumulate_params,
                     float* dbias, int accumulate_bias,
                     int M, int H, float dropout_p, const long long* seed, unsigned site, int post_dropout,
                     int dtype, hs_stream_t stream);
/* out[n] (= or +=) sum_m x[m,n] for a 16-bit [M,N] matrix: the bias gradients of every Linear. */
HS_API int hs_colsum(const void* x, int M, int N, int ld, float* out, int accumulate, int dtype,
                     hs_stream_t stream);
/* Floats of scratch needed by hs_colsum_rows_where (row-slab partial sums). */
HS_API int hs_colsum_scratch_elems(int M, int N);

/* ---- fused cross-entropy + MLM row gather/scatter (csrc/xent_sm100.cu) ------------------------------- */
/* CrossEntropyLoss(ignore_index=-1), mean reduction (bert_modeling.py:881-886, K13). Per row: online
 * log-sum-exp, row_loss[r]; with write_grad the logits are overwritten in place by
 * (softmax - onehot) / n_valid (ignored rows and pad columns [V, ld) become 0). loss (optional)
 * (= or +=) mean over valid rows. labels int32 or int64. With inv_count_out non-NULL the stored gradient is
 * softmax - onehot WITHOUT the 1/n_valid (O(1) values survive fp16; ~1e-7 ones do not) and *inv_count_out
 * receives 1/n_valid for the consumer to fold into its GEMM alpha. */
HS_API int hs_xent(void* logits, int ld, const void* labels, int label_is_i64, int rows, int V, float* row_loss,
                   float* loss, int accumulate_loss, int write_grad, float* inv_count_out, int dtype,
                   hs_stream_t stream);
/* dst[r,:] = src[idx[r],:] (idx < 0 -> zeros): only the labelled positions enter the MLM head (Q21/W9). */
HS_API int hs_gather_rows(const void* src, const void* idx, int idx_is_i64, void* dst, int rows, int row_bytes,
                          hs_stream_t stream);
/* dst[idx[r],:] += src[r,:] for unique valid idx (gradient of hs_gather_rows). */
HS_API int hs_scatter_add_rows(const void* src, const void* idx, int idx_is_i64, void* dst, int rows, int H,
                               int dtype, hs_stream_t stream);

/* ---- embeddings (csrc/embedding_sm100.cu; bert_modeling.py:292-320, K8/W8) --------------------------- */
HS_API int hs_embed_gather_sum(const void* W, const void* P, const void* Tt, const void* ids, const void* seg,
                               int ids_i64, void* z, int M, int S, int H, int vocab, int ntypes, int dtype,
                               hs_stream_t stream);
HS_API int hs_embed_scatter_word(const void* de, const void* ids, int ids_i64, float* dW, int M, int H, int vocab,
                                 int dtype, hs_stream_t stream);
/* Deterministic form of the same scatter-add (the reference's embedding_dense_backward is an atomic scatter,
 * bert_modeling.py:313), and the token-type gradient from the same pass over de. order[M] lists the positions
 * sorted by token id (stable), run_end[i] is, for the first sorted slot of each run of equal ids, one past the
 * run's last slot and 0 for every other slot. The sorted slots are cut into 16-row segments (one CTA each, rows
 * added in slot order); runs that cross segments are completed by a second kernel in segment order, so every
 * slice of dW has one writer and one order of additions: bit-identical from run to run.
 * seg / dT: token types (0 or 1) of the positions and the [2, H] fp32 gradient of the token-type table
 * (overwritten, or added to when accumulate_T); both NULL to leave the token types out.
 * scratch: hs_embed_backward_scratch_elems(M, H) floats. */
HS_API long long hs_embed_backward_scratch_elems(int M, int H);
HS_API int hs_embed_backward_sorted(const void* de, const void* ids, const void* seg, int ids_i64, const int* order,
                                    const int* run_end, float* dW, float* dT, int accumulate_T, float* scratch,
                                    int M, int H, int vocab, int dtype, hs_stream_t stream);
HS_API int hs_sum_over_batch(const void* x, int B, long long inner, float* out, int accumulate, int dtype,
                             hs_stream_t stream);
HS_API int hs_colsum_rows_where(const void* x, const void* sel, int sel_i64, long long want, int M, int N, float* out,
                                int accumulate, float* scratch, int dtype, hs_stream_t stream);

/* ---- fused attention on tcgen05 (csrc/attention_sm100.cu; bert_modeling.py:351-377, K2-K4/W4) -------- */
/* 1 when the fused kernels handle this (sequence length, head dim): head dim 64 and S a multiple of 16 up to
 * 512 -- S = 128 by the single-tile kernels, every other length by the multi-tile kernels
 * (attention_multi_sm100.cu: 1..4 tiles of 128 rows, the last one partial through out-of-bounds zero fill;
 * HETSEQ_ATTN_MULTI=0 disables them). Callers must check first. */
HS_API int hs_attention_supported(int S, int head_dim);
/* ctx[M,H] = merge_heads(dropout(softmax(Q K^T / sqrt(d) + mask)) V) from the packed projection
 * qkv [M, 3H] (q | k | v column blocks, heads of 64 inside each). mask: [B,S] additive fp32.
 * lse: [B*heads*S] log-sum-exp of the scaled+masked scores (kept for backward). */
HS_API int hs_attention_fwd(const void* qkv, const float* mask, void* ctx, float* lse, int B, int S, int heads,
                            int H, float dropout_p, const long long* seed, unsigned site, int dtype,
                            hs_stream_t stream);
/* dqkv [M, 3H] from dctx [M,H]; recomputes the probabilities and the dropout mask. dbias (optional, fp32
 * [3H], = or += per accumulate_bias) receives the column sums of dqkv: the q/k/v bias gradient. */
HS_API int hs_attention_bwd(const void* qkv, const float* mask, const void* ctx, const void* dctx, const float* lse,
                            void* dqkv, float* dbias, int accumulate_bias, int B, int S, int heads, int H,
                            float dropout_p, const long long* seed, unsigned site, int dtype, hs_stream_t stream);

/* ---- fused data-parallel gradient path over NVLink / NVSwitch (csrc/comm_sm100.cu) ---------------------- */
/* Replaces DDP's bucketed NCCL all-reduce (controller.py:78-86, N4) + multiply_grads + clip_grad_norm +
 * Adam.step / Adadelta.step on every rank (optim.py:59-70,162-231,263-304, K15-K18) -- SURVEY.md W1, W12.
 * W <= 16 ranks of one node. The flat gradient / parameter buffers and one signal pad per rank live in
 * symmetric memory: *_ptrs are HOST arrays of W device pointers indexed by rank (entry [rank] is the local
 * buffer, the others are peer mappings); *_mc is the NVLS multicast address of the same buffer, or NULL --
 * then the kernels use peer loads / stores instead of multimem.ld_reduce / multimem.st. Ownership is
 * cyclic by chunk: chunk c of the flat buffers belongs to rank c % W. Flags in the signal pads carry `epoch`
 * (the update number, strictly increasing); when `epoch_dev` is non-NULL the kernels read the epoch from that
 * device word instead (so a captured CUDA graph stays valid across updates). Pads are hs_dp_signal_words()
 * 32-bit words, zero-initialised. */
HS_API int hs_ipc_alloc(size_t bytes, void** ptr, unsigned char* handle64);   /* cudaMalloc + CUDA IPC handle */
HS_API int hs_ipc_open(const unsigned char* handle64, void** ptr);
HS_API int hs_ipc_close(void* ptr);
HS_API int hs_ipc_free(void* ptr);
HS_API int hs_dp_signal_words(int W, int max_buckets);
HS_API int hs_dp_max_gather_buckets(void);
/* Upper bound of every in-kernel wait on a peer, for the current device (default 1800 s, the process-group
 * timeout that bounds the reference's waits inside NCCL). Beyond it the kernel prints which peer it was
 * waiting for and traps. Synchronous: call it at set-up. */
HS_API int hs_dp_set_timeout(double seconds);
/* One bucket = chunks [chunk_begin, chunk_end). Announces the bucket to all peers, waits for theirs, then
 * reduces the W copies of every owned chunk into the local fp32 gradient and adds the squared sum into
 * *sumsq_accum. Launch on a side stream as soon as the bucket's gradients are final. */
HS_API int hs_dp_reduce_scatter(void* const* grad_ptrs, const void* grads_mc, void* const* sig_ptrs, int rank, int W,
                                long long chunk_begin, long long chunk_end, int chunk, int slot, unsigned epoch,
                                const unsigned* epoch_dev, float* sumsq_accum, int nblocks, hs_stream_t stream);
/* All ranks exchange their norm partial and their sample size through the pads and derive identical
 * out5 = [grad_norm, multiplier, overflow flag, global sum of squares, global sample size]; *skip_out = flag.
 * multiplier = mult_host * (*mult_dev) * clip coefficient, and additionally / (sum of sample sizes) when
 * `sample_size` is non-NULL (the reference's multiply_grads(world / sample_size) on DDP's mean,
 * controller.py:338-339). `pending` (optional) is set to 1: an update is waiting for hs_dp_update_allgather. */
HS_API int hs_dp_norm_exchange(void* const* sig_ptrs, int rank, int W, unsigned epoch, const unsigned* epoch_dev,
                               float* sumsq_accum, float mult_host, const float* mult_dev, const float* sample_size,
                               float max_norm, float* out5, int* skip_out, int* pending, hs_stream_t stream);
/* opt = 0: Adam (optim.py:162-231; a_coef/b_coef = beta1/beta2, state1/2 = exp_avg/exp_avg_sq),
 * opt = 1: Adadelta (optim.py:263-304; a_coef = rho, state1/2 = square_avg/acc_delta) on the owned chunks, in
 * flat order; the updated 16-bit parameters are stored into all W parameter buffers (the all-gather).
 * The buffer is cut into `ngather` slices ("gather buckets"); each is published to every peer with its own flag
 * (value epoch + epoch_offset) as soon as this rank's part of it is stored. scalars = out5 of
 * hs_dp_norm_exchange. With `pending` non-NULL the update is applied only if *pending != 0 (and *pending is
 * cleared); the flags are published either way. counters: ngather + 1 zeroed words. */
HS_API int hs_dp_update_allgather(int opt, const float* grad, float* master, float* state1, float* state2,
                                  void* const* param_ptrs, void* params_mc, void* const* sig_ptrs, int rank, int W,
                                  long long nblocks, int chunk, int out_dtype, float lr, float a_coef, float b_coef,
                                  float eps, float weight_decay, float step_factor, const float* scalars,
                                  const float* hyper_dev, unsigned epoch, const unsigned* epoch_dev, int epoch_offset,
                                  unsigned int* counters, int* pending, int ngather, int grid_blocks,
                                  hs_stream_t stream);
/* The same update split in two, which is what training uses: hs_dp_update_local runs Adam / Adadelta on the owned
 * chunks at the end of the step and stores the new 16-bit values into THIS rank's parameter buffer only;
 * hs_dp_allgather later copies the owned chunks into every peer's buffer (multimem.st or W-1 peer stores), gather
 * bucket by gather bucket with the same flags, counters and `pending` protocol as hs_dp_update_allgather. The
 * gather is a pure copy that fits next to the next forward pass's GEMMs, where it is overlapped. */
HS_API int hs_dp_update_local(int opt, const float* grad, float* master, float* state1, float* state2,
                              void* param_local, int rank, int W, long long nblocks, int chunk, int out_dtype,
                              float lr, float a_coef, float b_coef, float eps, float weight_decay, float step_factor,
                              const float* scalars, const float* hyper_dev, int grid_blocks, hs_stream_t stream);
HS_API int hs_dp_allgather(void* const* param_ptrs, void* params_mc, void* const* sig_ptrs, int rank, int W,
                           long long nblocks, int chunk, const float* scalars, unsigned epoch,
                           const unsigned* epoch_dev, int epoch_offset, unsigned int* counters, int* pending,
                           int ngather, int grid_blocks, hs_stream_t stream);
/* Gather bucket that owned block `block` (of nblocks per rank) falls into, and the number of non-empty gather
 * buckets (<= ngather: buckets hold ceil(nblocks / ngather) blocks each, so the last requested ones may be
 * empty; flags exist, and may be waited for, only for buckets [0, hs_dp_gather_count)). */
HS_API int hs_dp_gather_bucket_of(long long nblocks, int ngather, long long block);
HS_API int hs_dp_gather_count(long long nblocks, int ngather);
/* Returns (on the stream) once gather buckets [j_lo, j_hi] of update `epoch + epoch_offset` have arrived from
 * every rank. The forward pass queues one in front of the first kernel that reads a slice. */
HS_API int hs_dp_wait_params(void* const* sig_ptrs, int rank, int W, int j_lo, int j_hi, unsigned epoch,
                             const unsigned* epoch_dev, int epoch_offset, hs_stream_t stream);
/* Protocol test on ONE GPU: W virtual ranks (blocks_per_rank CTAs each) run one whole update -- the same device
 * code as the four kernels above, peer-load / peer-store variant -- inside one cooperative launch. All pointer
 * arguments are host arrays of W device pointers (one "rank" each); barrier_words: W words. split = 1 runs
 * hs_dp_update_local + hs_dp_allgather instead of the fused hs_dp_update_allgather. */
HS_API int hs_dp_emulate_step(int W, int blocks_per_rank, void* const* grad_ptrs, void* const* param_ptrs,
                              void* const* sig_ptrs, float* const* sumsq, const float* const* sample_size,
                              float* const* scalars, int* const* skip, float* const* master, float* const* state1,
                              float* const* state2, unsigned int* const* counters, int* const* pending,
                              unsigned int* barrier_words, long long total_chunks, long long chunks_per_bucket,
                              int chunk, int ngather, int opt, int out_dtype, float lr, float a_coef, float b_coef,
                              float eps, float weight_decay, float step_factor, float mult_host, float max_norm,
                              unsigned epoch, int split, hs_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* HETSEQ_B200_H_ */
