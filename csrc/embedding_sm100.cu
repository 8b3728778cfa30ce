// hetseq_b200 -- BERT embeddings: gather-sum forward and gradient scatter.
//
// Reference: BertEmbeddings (bert_modeling.py:292-320): word + position(arange) + token-type lookups,
// then LayerNorm and dropout -- 3 embedding kernels + 2 adds forward, and 3 embedding_dense_backward
// launches (atomic scatter-add into [V,H]) backward.
//   hs_embed_gather_sum    z[m,:] = W[ids[m],:] + P[m % S,:] + T[seg[m],:]        (LayerNorm + dropout
//                          follow through hs_ln_fwd(post_dropout=1), the gradient through hs_ln_bwd)
//   hs_embed_scatter_word_sorted  dW[id,:] += sum of de[m,:] over the positions m holding token id, as a segmented
//                          reduction over the positions SORTED by token id: one thread per (run of equal ids,
//                          8 columns) adds the rows of its run in ascending position order and is the only writer
//                          of its slice of dW -- deterministic, no atomics. The order and the run
//                          boundaries come from the host (computed in the batch-prefetch thread) or from a device sort.
//   hs_embed_scatter_word  the same sum with fp32 vector atomics (order of additions varies from run to run)
//   hs_sum_over_batch      dP[s,:] (+)= sum_b de[b*S+s,:]            deterministic, no atomics
//   hs_colsum_rows_where   dT[t,:] (+)= sum_{m: seg[m]==t} de[m,:]   deterministic two-stage column sum
// Token ids arrive as int32 (the trainer narrows them before the host->device copy) or int64.
#include "common.cuh"
#include "host.h"
#include "../include/hetseq_b200.h"

namespace hs {

template <typename T>
__device__ __forceinline__ void unpack8e(const uint4& u, float (&f)[8]) {
  float2 a = Cvt<T>::unpack2(u.x), b = Cvt<T>::unpack2(u.y), c = Cvt<T>::unpack2(u.z), d = Cvt<T>::unpack2(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}

template <typename T, typename IdT>
__global__ void __launch_bounds__(256)
embed_gather_sum_kernel(const T* __restrict__ W, const T* __restrict__ P, const T* __restrict__ Tt,
                        const IdT* __restrict__ ids, const IdT* __restrict__ seg, T* __restrict__ z, int M, int S,
                        int H, int vocab, int ntypes) {
  const int vpr = H >> 3;
  const long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int m = static_cast<int>(t / vpr), c = static_cast<int>(t % vpr);
  if (m >= M) return;
  long long id = static_cast<long long>(ids[m]);
  long long sg = static_cast<long long>(seg[m]);
  id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
  sg = sg < 0 ? 0 : (sg >= ntypes ? ntypes - 1 : sg);
  float a[8], b[8], d[8];
  unpack8e<T>(reinterpret_cast<const uint4*>(W + static_cast<size_t>(id) * H)[c], a);
  unpack8e<T>(reinterpret_cast<const uint4*>(P + static_cast<size_t>(m % S) * H)[c], b);
  unpack8e<T>(reinterpret_cast<const uint4*>(Tt + static_cast<size_t>(sg) * H)[c], d);
  uint4 o;
  o.x = Cvt<T>::pack2(a[0] + b[0] + d[0], a[1] + b[1] + d[1]);
  o.y = Cvt<T>::pack2(a[2] + b[2] + d[2], a[3] + b[3] + d[3]);
  o.z = Cvt<T>::pack2(a[4] + b[4] + d[4], a[5] + b[5] + d[5]);
  o.w = Cvt<T>::pack2(a[6] + b[6] + d[6], a[7] + b[7] + d[7]);
  reinterpret_cast<uint4*>(z + static_cast<size_t>(m) * H)[c] = o;
}

template <typename T, typename IdT>
__global__ void __launch_bounds__(256)
embed_scatter_word_kernel(const T* __restrict__ de, const IdT* __restrict__ ids, float* __restrict__ dW, int M, int H,
                          int vocab) {
  const int vpr = H >> 3;
  const long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int m = static_cast<int>(t / vpr), c = static_cast<int>(t % vpr);
  if (m >= M) return;
  const long long id = static_cast<long long>(ids[m]);
  if (id < 0 || id >= vocab) return;
  float f[8];
  unpack8e<T>(reinterpret_cast<const uint4*>(de + static_cast<size_t>(m) * H)[c], f);
  float4* dst = reinterpret_cast<float4*>(dW + static_cast<size_t>(id) * H + c * 8);
  atomicAdd(dst, make_float4(f[0], f[1], f[2], f[3]));
  atomicAdd(dst + 1, make_float4(f[4], f[5], f[6], f[7]));
}

// order[i]: position (row of de) that is i-th when positions are sorted by token id (stable: ties in ascending
// position). run_end[i]: for the FIRST sorted slot of a run of equal ids, one past its last slot; 0 elsewhere.
template <typename T, typename IdT>
__global__ void __launch_bounds__(256)
embed_scatter_sorted_kernel(const T* __restrict__ de, const IdT* __restrict__ ids, const int* __restrict__ order,
                            const int* __restrict__ run_end, float* __restrict__ dW, int M, int H, int vocab) {
  const int vpr = H >> 3;
  const long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int i = static_cast<int>(t / vpr), c = static_cast<int>(t % vpr);
  if (i >= M) return;
  const int e = min(run_end[i], M);
  if (e <= i) return;                     // not the first slot of a run
  const int p0 = order[i];
  if (p0 < 0 || p0 >= M) return;
  const long long id = static_cast<long long>(ids[p0]);
  if (id < 0 || id >= vocab) return;
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  constexpr int U = 8;                    // rows in flight per thread; a warp reads 512 contiguous bytes of each row
  for (int j = i; j < e; j += U) {
    int p[U];
#pragma unroll
    for (int u = 0; u < U; ++u) p[u] = j + u < e ? order[j + u] : -1;
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u)
      v[u] = (p[u] >= 0 && p[u] < M) ? reinterpret_cast<const uint4*>(de + static_cast<size_t>(p[u]) * H)[c] : make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int u = 0; u < U; ++u) {          // ascending slot order: the sum is the same on every run
      float f[8];
      unpack8e<T>(v[u], f);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += f[k];
    }
  }
  float4* dst = reinterpret_cast<float4*>(dW + static_cast<size_t>(id) * H + c * 8);
  float4 a = dst[0], b = dst[1];           // the decoder's weight gradient is already in there (tied matrix)
  a.x += acc[0]; a.y += acc[1]; a.z += acc[2]; a.w += acc[3];
  b.x += acc[4]; b.y += acc[5]; b.z += acc[6]; b.w += acc[7];
  dst[0] = a;
  dst[1] = b;
}

template <typename T>
__global__ void __launch_bounds__(256)
sum_over_batch_kernel(const T* __restrict__ x, int B, long long inner, float* __restrict__ out, int accumulate) {
  const long long j8 = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (j8 * 8 >= inner) return;
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int b = 0; b < B; ++b) {
    float f[8];
    unpack8e<T>(*reinterpret_cast<const uint4*>(x + static_cast<size_t>(b) * inner + j8 * 8), f);
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] += f[k];
  }
  float* o = out + j8 * 8;
#pragma unroll
  for (int k = 0; k < 8; ++k) o[k] = accumulate ? o[k] + acc[k] : acc[k];
}

template <typename T, typename IdT>
__global__ void __launch_bounds__(256)
colsum_where_partial_kernel(const T* __restrict__ x, const IdT* __restrict__ sel, long long want, int M, int N,
                            int rows_per_block, float* __restrict__ part) {
  const int c8 = blockIdx.x * blockDim.x + threadIdx.x;
  if (c8 * 8 >= N) return;
  const int r0 = blockIdx.y * rows_per_block, r1 = min(M, r0 + rows_per_block);
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int r = r0; r < r1; ++r) {
    if (static_cast<long long>(sel[r]) != want) continue;
    float f[8];
    unpack8e<T>(*reinterpret_cast<const uint4*>(x + static_cast<size_t>(r) * N + c8 * 8), f);
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] += f[k];
  }
  float* dst = part + static_cast<size_t>(blockIdx.y) * N + c8 * 8;
#pragma unroll
  for (int k = 0; k < 8; ++k) dst[k] = acc[k];
}

__global__ void __launch_bounds__(256)
colsum_parts2_kernel(const float* __restrict__ part, int nparts, int N, float* __restrict__ out, int accumulate) {
  const int h = blockIdx.x * blockDim.x + threadIdx.x;
  if (h >= N) return;
  float a = 0.f;
  for (int i = 0; i < nparts; ++i) a += part[static_cast<size_t>(i) * N + h];
  out[h] = accumulate ? out[h] + a : a;
}

}  // namespace hs

using namespace hs;

extern "C" int hs_embed_gather_sum(const void* W, const void* P, const void* Tt, const void* ids, const void* seg,
                                   int ids_i64, void* z, int M, int S, int H, int vocab, int ntypes, int dtype,
                                   hs_stream_t stream) {
  if (M <= 0 || S <= 0 || H % 8 != 0) { set_last_error("hs_embed_gather_sum: bad shape M=%d S=%d H=%d", M, S, H); return 1; }
  const long long total = static_cast<long long>(M) * (H / 8);
  const int grid = static_cast<int>((total + 255) / 256);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
#define HS_EG(T)                                                                                                          \
  do {                                                                                                                    \
    if (ids_i64) embed_gather_sum_kernel<T, long long><<<grid, 256, 0, s>>>((const T*)W, (const T*)P, (const T*)Tt, (const long long*)ids, (const long long*)seg, (T*)z, M, S, H, vocab, ntypes); \
    else embed_gather_sum_kernel<T, int><<<grid, 256, 0, s>>>((const T*)W, (const T*)P, (const T*)Tt, (const int*)ids, (const int*)seg, (T*)z, M, S, H, vocab, ntypes); \
  } while (0)
  if (dtype == HS_F16) HS_EG(__half);
  else if (dtype == HS_BF16) HS_EG(__nv_bfloat16);
  else { set_last_error("hs_embed_gather_sum: dtype must be f16 or bf16"); return 1; }
#undef HS_EG
  return post_launch("embed_gather_sum_kernel");
}

extern "C" int hs_embed_scatter_word(const void* de, const void* ids, int ids_i64, float* dW, int M, int H, int vocab,
                                     int dtype, hs_stream_t stream) {
  if (M <= 0 || H % 8 != 0) { set_last_error("hs_embed_scatter_word: bad shape"); return 1; }
  const long long total = static_cast<long long>(M) * (H / 8);
  const int grid = static_cast<int>((total + 255) / 256);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
#define HS_ES(T)                                                                                                  \
  do {                                                                                                            \
    if (ids_i64) embed_scatter_word_kernel<T, long long><<<grid, 256, 0, s>>>((const T*)de, (const long long*)ids, dW, M, H, vocab); \
    else embed_scatter_word_kernel<T, int><<<grid, 256, 0, s>>>((const T*)de, (const int*)ids, dW, M, H, vocab);    \
  } while (0)
  if (dtype == HS_F16) HS_ES(__half);
  else if (dtype == HS_BF16) HS_ES(__nv_bfloat16);
  else { set_last_error("hs_embed_scatter_word: dtype must be f16 or bf16"); return 1; }
#undef HS_ES
  return post_launch("embed_scatter_word_kernel");
}

extern "C" int hs_embed_scatter_word_sorted(const void* de, const void* ids, int ids_i64, const int* order,
                                            const int* run_end, float* dW, int M, int H, int vocab, int dtype,
                                            hs_stream_t stream) {
  if (M <= 0 || H % 8 != 0 || order == nullptr || run_end == nullptr) { set_last_error("hs_embed_scatter_word_sorted: bad arguments"); return 1; }
  const long long total = static_cast<long long>(M) * (H / 8);
  const int grid = static_cast<int>((total + 255) / 256);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
#define HS_ESS(T)                                                                                                 \
  do {                                                                                                            \
    if (ids_i64) embed_scatter_sorted_kernel<T, long long><<<grid, 256, 0, s>>>((const T*)de, (const long long*)ids, order, run_end, dW, M, H, vocab); \
    else embed_scatter_sorted_kernel<T, int><<<grid, 256, 0, s>>>((const T*)de, (const int*)ids, order, run_end, dW, M, H, vocab); \
  } while (0)
  if (dtype == HS_F16) HS_ESS(__half);
  else if (dtype == HS_BF16) HS_ESS(__nv_bfloat16);
  else { set_last_error("hs_embed_scatter_word_sorted: dtype must be f16 or bf16"); return 1; }
#undef HS_ESS
  return post_launch("embed_scatter_sorted_kernel");
}

extern "C" int hs_sum_over_batch(const void* x, int B, long long inner, float* out, int accumulate, int dtype,
                                 hs_stream_t stream) {
  if (B <= 0 || inner <= 0 || inner % 8 != 0) { set_last_error("hs_sum_over_batch: inner must be a multiple of 8"); return 1; }
  const int grid = static_cast<int>((inner / 8 + 255) / 256);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (dtype == HS_F16) sum_over_batch_kernel<__half><<<grid, 256, 0, s>>>((const __half*)x, B, inner, out, accumulate);
  else if (dtype == HS_BF16) sum_over_batch_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>((const __nv_bfloat16*)x, B, inner, out, accumulate);
  else { set_last_error("hs_sum_over_batch: dtype must be f16 or bf16"); return 1; }
  return post_launch("sum_over_batch_kernel");
}

extern "C" int hs_colsum_rows_where(const void* x, const void* sel, int sel_i64, long long want, int M, int N, float* out,
                                    int accumulate, float* scratch, int dtype, hs_stream_t stream) {
  if (M <= 0 || N <= 0 || N % 8 != 0 || scratch == nullptr) { set_last_error("hs_colsum_rows_where: bad arguments"); return 1; }
  const int slabs = colsum_slabs(M);
  const int rows_per_block = (M + slabs - 1) / slabs;
  dim3 grid((N / 8 + 255) / 256, slabs);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
#define HS_CW(T)                                                                                                   \
  do {                                                                                                             \
    if (sel_i64) colsum_where_partial_kernel<T, long long><<<grid, 256, 0, s>>>((const T*)x, (const long long*)sel, want, M, N, rows_per_block, scratch); \
    else colsum_where_partial_kernel<T, int><<<grid, 256, 0, s>>>((const T*)x, (const int*)sel, want, M, N, rows_per_block, scratch); \
  } while (0)
  if (dtype == HS_F16) HS_CW(__half);
  else if (dtype == HS_BF16) HS_CW(__nv_bfloat16);
  else { set_last_error("hs_colsum_rows_where: dtype must be f16 or bf16"); return 1; }
#undef HS_CW
  if (post_launch("colsum_where_partial_kernel")) return 1;
  colsum_parts2_kernel<<<(N + 255) / 256, 256, 0, s>>>(scratch, slabs, N, out, accumulate);
  return post_launch("colsum_parts2_kernel");
}
