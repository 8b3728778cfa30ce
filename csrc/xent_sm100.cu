// hetseq_b200 -- fused softmax cross-entropy (forward + gradient in one pass) and row gather/scatter.
//
// Reference: CrossEntropyLoss(ignore_index=-1), mean reduction, applied to the MLM logits
// [B*S, 30522] and the NSP logits [B, 2] (bert_modeling.py:881-886). ATen runs
// log_softmax + nll_loss forward, then materialises a second [rows, V] tensor in backward. Here one
// block per row makes two passes over the (L2-resident) row: online max / sum-exp, then it overwrites
// the logits IN PLACE with d(loss)/d(logits), zeroing ignored rows and the pad columns [V, ld). By default
// that is (softmax - onehot) / n_valid; with inv_count_out the 1/n_valid is NOT applied to the stored values
// but written to a device scalar for the consumer to fold into its accumulator scale: the stored entries are
// then O(1) instead of O(1/n_valid), which matters in fp16 (with ~600 labelled rows and V = 30522 almost the
// whole tail of softmax/n_valid is subnormal or flushes to zero). The scalar upstream gradient is applied
// later through the decoder GEMMs' alpha_dev, so the [rows, V] tensor is touched exactly twice overall. The MLM head only ever sees
// the labelled rows (hs_gather_rows / hs_scatter_add_rows implement SURVEY.md Q21 / W9).
#include "common.cuh"
#include "host.h"
#include "../include/hetseq_b200.h"

namespace hs {

constexpr int XE_THREADS = 256;

__device__ __forceinline__ float block_reduce_sum(float v, float* red) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  v = (lane < (blockDim.x >> 5)) ? red[lane] : 0.0f;
  return warp_sum(v);  // every thread of warp 0..n gets the total (all warps read the same red[])
}
__device__ __forceinline__ float block_reduce_max(float v, float* red) {
  v = warp_max(v);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  v = (lane < (blockDim.x >> 5)) ? red[lane] : -INFINITY;
  return warp_max(v);
}

template <typename T, typename LabelT>
__global__ void __launch_bounds__(XE_THREADS)
xent_fwd_bwd_kernel(T* __restrict__ logits, int ld, const LabelT* __restrict__ labels, int rows, int V,
                    float* __restrict__ row_loss, int write_grad, int normalize) {
  __shared__ float red[32];
  const int row = blockIdx.x;
  // number of valid rows (same value in every block; labels is tiny)
  float cnt = 0.f;
  for (int i = threadIdx.x; i < rows; i += blockDim.x) cnt += (labels[i] >= 0) ? 1.f : 0.f;
  cnt = block_reduce_sum(cnt, red);
  const float inv_n = normalize ? (cnt > 0.f ? 1.f / cnt : 0.f) : 1.f;

  T* x = logits + static_cast<size_t>(row) * ld;
  const long long label = static_cast<long long>(labels[row]);
  const bool valid = label >= 0 && label < V;
  const int nvec = ld >> 3;  // ld % 8 == 0 checked by the launcher; columns >= V are padding
  if (!valid) {
    if (threadIdx.x == 0) row_loss[row] = 0.f;
    if (write_grad) {
      const uint4 zero = make_uint4(0, 0, 0, 0);
      for (int i = threadIdx.x; i < nvec; i += blockDim.x) reinterpret_cast<uint4*>(x)[i] = zero;
    }
    return;
  }
  // pass 1: online max and sum of exp
  float m = -INFINITY, s = 0.f;
  for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
    const uint4 u = reinterpret_cast<const uint4*>(x)[i];
    float2 f[4] = {Cvt<T>::unpack2(u.x), Cvt<T>::unpack2(u.y), Cvt<T>::unpack2(u.z), Cvt<T>::unpack2(u.w)};
    float v[8] = {f[0].x, f[0].y, f[1].x, f[1].y, f[2].x, f[2].y, f[3].x, f[3].y};
    float lm = m;
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (i * 8 + k < V) lm = fmaxf(lm, v[k]);
    if (lm > m) { s *= __expf(m - lm); m = lm; }
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (i * 8 + k < V) s += __expf(v[k] - m);
  }
  const float gmax = block_reduce_max(m, red);
  s = (m == -INFINITY) ? 0.f : s * __expf(m - gmax);
  const float gsum = block_reduce_sum(s, red);
  const float lse = gmax + __logf(gsum);
  if (threadIdx.x == 0) row_loss[row] = lse - Cvt<T>::to_f(x[label]);
  if (!write_grad) return;
  __syncthreads();  // x[label] has been read before anyone overwrites it
  // pass 2: gradient in place
  const float inv_sum = 1.f / gsum;
  for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
    const uint4 u = reinterpret_cast<const uint4*>(x)[i];
    float2 f[4] = {Cvt<T>::unpack2(u.x), Cvt<T>::unpack2(u.y), Cvt<T>::unpack2(u.z), Cvt<T>::unpack2(u.w)};
    float v[8] = {f[0].x, f[0].y, f[1].x, f[1].y, f[2].x, f[2].y, f[3].x, f[3].y};
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int c = i * 8 + k;
      float g = 0.f;
      if (c < V) {
        g = __expf(v[k] - gmax) * inv_sum;
        if (c == label) g -= 1.f;
        g *= inv_n;
      }
      v[k] = g;
    }
    uint4 o;
    o.x = Cvt<T>::pack2(v[0], v[1]); o.y = Cvt<T>::pack2(v[2], v[3]);
    o.z = Cvt<T>::pack2(v[4], v[5]); o.w = Cvt<T>::pack2(v[6], v[7]);
    reinterpret_cast<uint4*>(x)[i] = o;
  }
}

// loss (= or +=) sum(row_loss) / count(labels >= 0); single block, deterministic
template <typename LabelT>
__global__ void __launch_bounds__(XE_THREADS)
xent_finalize_kernel(const float* __restrict__ row_loss, const LabelT* __restrict__ labels, int rows,
                     float* __restrict__ loss, int accumulate, float* __restrict__ inv_count_out) {
  __shared__ float red[32];
  float s = 0.f, c = 0.f;
  for (int i = threadIdx.x; i < rows; i += blockDim.x) {
    s += row_loss[i];
    c += (labels[i] >= 0) ? 1.f : 0.f;
  }
  s = block_reduce_sum(s, red);
  c = block_reduce_sum(c, red);
  if (threadIdx.x == 0) {
    const float v = c > 0.f ? s / c : 0.f;
    if (loss != nullptr) loss[0] = accumulate ? loss[0] + v : v;
    if (inv_count_out != nullptr) inv_count_out[0] = c > 0.f ? 1.f / c : 0.f;
  }
}

// dst[r, :] = src[idx[r], :] (idx < 0 -> zeros)
template <typename IdxT>
__global__ void __launch_bounds__(256)
gather_rows_kernel(const uint4* __restrict__ src, const IdxT* __restrict__ idx, uint4* __restrict__ dst, int rows,
                   int vec_per_row) {
  const long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int r = static_cast<int>(t / vec_per_row), c = static_cast<int>(t % vec_per_row);
  if (r >= rows) return;
  const long long s = static_cast<long long>(idx[r]);
  dst[static_cast<size_t>(r) * vec_per_row + c] =
      s >= 0 ? src[static_cast<size_t>(s) * vec_per_row + c] : make_uint4(0, 0, 0, 0);
}

// dst[idx[r], :] += src[r, :] for 16-bit rows; indices must be unique among valid rows (they are: one
// row per labelled position), so plain read-modify-write is race-free.
template <typename T, typename IdxT>
__global__ void __launch_bounds__(256)
scatter_add_rows_kernel(const uint4* __restrict__ src, const IdxT* __restrict__ idx, uint4* __restrict__ dst,
                        int rows, int vec_per_row) {
  const long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int r = static_cast<int>(t / vec_per_row), c = static_cast<int>(t % vec_per_row);
  if (r >= rows) return;
  const long long d = static_cast<long long>(idx[r]);
  if (d < 0) return;
  const uint4 a = src[static_cast<size_t>(r) * vec_per_row + c];
  uint4* p = dst + static_cast<size_t>(d) * vec_per_row + c;
  const uint4 b = *p;
  const float2 a0 = Cvt<T>::unpack2(a.x), a1 = Cvt<T>::unpack2(a.y), a2 = Cvt<T>::unpack2(a.z), a3 = Cvt<T>::unpack2(a.w);
  const float2 b0 = Cvt<T>::unpack2(b.x), b1 = Cvt<T>::unpack2(b.y), b2 = Cvt<T>::unpack2(b.z), b3 = Cvt<T>::unpack2(b.w);
  uint4 o;
  o.x = Cvt<T>::pack2(a0.x + b0.x, a0.y + b0.y); o.y = Cvt<T>::pack2(a1.x + b1.x, a1.y + b1.y);
  o.z = Cvt<T>::pack2(a2.x + b2.x, a2.y + b2.y); o.w = Cvt<T>::pack2(a3.x + b3.x, a3.y + b3.y);
  *p = o;
}

}  // namespace hs

using namespace hs;

extern "C" int hs_xent(void* logits, int ld, const void* labels, int label_is_i64, int rows, int V, float* row_loss,
                       float* loss, int accumulate_loss, int write_grad, float* inv_count_out, int dtype,
                       hs_stream_t stream) {
  if (rows <= 0 || V <= 0 || ld < V || ld % 8 != 0 || (reinterpret_cast<uintptr_t>(logits) % 16) != 0) {
    set_last_error("hs_xent: bad arguments rows=%d V=%d ld=%d (ld must be a multiple of 8, >= V; base 16B aligned)", rows, V, ld);
    return 1;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
#define HS_XENT(T, L)                                                                                           \
  do {                                                                                                          \
    xent_fwd_bwd_kernel<T, L><<<rows, XE_THREADS, 0, s>>>((T*)logits, ld, (const L*)labels, rows, V, row_loss, write_grad, inv_count_out == nullptr ? 1 : 0); \
    if (post_launch("xent_fwd_bwd_kernel")) return 1;                                                           \
    if (loss || inv_count_out) {                                                                                \
      xent_finalize_kernel<L><<<1, XE_THREADS, 0, s>>>(row_loss, (const L*)labels, rows, loss, accumulate_loss, inv_count_out);  \
      if (post_launch("xent_finalize_kernel")) return 1;                                                        \
    }                                                                                                           \
  } while (0)
  if (dtype == HS_F16) { if (label_is_i64) HS_XENT(__half, long long); else HS_XENT(__half, int); }
  else if (dtype == HS_BF16) { if (label_is_i64) HS_XENT(__nv_bfloat16, long long); else HS_XENT(__nv_bfloat16, int); }
  else { set_last_error("hs_xent: dtype must be f16 or bf16"); return 1; }
#undef HS_XENT
  return 0;
}

extern "C" int hs_gather_rows(const void* src, const void* idx, int idx_is_i64, void* dst, int rows, int row_bytes,
                              hs_stream_t stream) {
  if (rows <= 0 || row_bytes <= 0 || row_bytes % 16 != 0) { set_last_error("hs_gather_rows: row_bytes must be a multiple of 16"); return 1; }
  const int vpr = row_bytes / 16;
  const long long total = static_cast<long long>(rows) * vpr;
  const int grid = static_cast<int>((total + 255) / 256);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (idx_is_i64) gather_rows_kernel<long long><<<grid, 256, 0, s>>>((const uint4*)src, (const long long*)idx, (uint4*)dst, rows, vpr);
  else gather_rows_kernel<int><<<grid, 256, 0, s>>>((const uint4*)src, (const int*)idx, (uint4*)dst, rows, vpr);
  return post_launch("gather_rows_kernel");
}

extern "C" int hs_scatter_add_rows(const void* src, const void* idx, int idx_is_i64, void* dst, int rows, int H,
                                   int dtype, hs_stream_t stream) {
  if (rows <= 0 || H <= 0 || H % 8 != 0) { set_last_error("hs_scatter_add_rows: H must be a multiple of 8"); return 1; }
  const int vpr = H / 8;
  const long long total = static_cast<long long>(rows) * vpr;
  const int grid = static_cast<int>((total + 255) / 256);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
#define HS_SC(T)                                                                                                   \
  do {                                                                                                             \
    if (idx_is_i64) scatter_add_rows_kernel<T, long long><<<grid, 256, 0, s>>>((const uint4*)src, (const long long*)idx, (uint4*)dst, rows, vpr); \
    else scatter_add_rows_kernel<T, int><<<grid, 256, 0, s>>>((const uint4*)src, (const int*)idx, (uint4*)dst, rows, vpr); \
  } while (0)
  if (dtype == HS_F16) HS_SC(__half);
  else if (dtype == HS_BF16) HS_SC(__nv_bfloat16);
  else { set_last_error("hs_scatter_add_rows: dtype must be f16 or bf16"); return 1; }
#undef HS_SC
  return post_launch("scatter_add_rows_kernel");
}
