// hetseq_b200 -- fused (bias +) dropout + residual + LayerNorm, forward and backward.
//
//   z = residual + dropout(x + bias)          y = gamma * (z - mean) / sqrt(var + eps) + beta
//
// TF-style LayerNorm of the reference: biased variance, eps = 1e-12 *inside* the sqrt
// (bert_modeling.py:277-289). Replaces, per call, the GEMM-output dropout kernel, the residual add and
// the ~9 elementwise/reduction launches of BertLayerNorm in BertSelfOutput / BertOutput
// (bert_modeling.py:387-391, 423-427) and the plain LayerNorm of the MLM
// transform and the embeddings (bert_modeling.py:313-319, 525-528).
//
// One warp owns one row: the row lives in registers (H <= 2048), statistics are two shuffles-only
// passes over registers (mean, then centred sum of squares), every global access is a 128-bit vector.
// Dropout masks come from Philox keyed by (step seed, site, element) -- backward recomputes them, so
// no mask tensor exists. The kernels are HBM-bound: forward moves 2 reads + 2 writes of [M,H] 16-bit.
#include <cstdlib>
#include "common.cuh"
#include "host.h"
#include "../include/hetseq_b200.h"

namespace hs {

constexpr int LN_MAXV = 8;             // vectors (8 elements each) per lane -> H <= 2048
constexpr int LN_ROWS_PER_BLOCK = 4;   // one warp per row
constexpr int LN_THREADS = LN_ROWS_PER_BLOCK * 32;

template <typename T>
__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  float2 a = Cvt<T>::unpack2(u.x), b = Cvt<T>::unpack2(u.y), c = Cvt<T>::unpack2(u.z), d = Cvt<T>::unpack2(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
template <typename T>
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 u;
  u.x = Cvt<T>::pack2(f[0], f[1]); u.y = Cvt<T>::pack2(f[2], f[3]);
  u.z = Cvt<T>::pack2(f[4], f[5]); u.w = Cvt<T>::pack2(f[6], f[7]);
  return u;
}

struct LnFwdParams {
  const void* x; const void* bias; const void* residual; const void* gamma; const void* beta;
  void* y; void* z; float* mean; float* rstd;
  int M, H; float eps; float p; const long long* seed; uint32_t site;
  int post_dropout;  // 1: y = dropout(LN(z)) (embeddings) instead of z = res + dropout(x)
};

template <typename T, int NV>
__global__ void __launch_bounds__(LN_THREADS) ln_fwd_kernel(LnFwdParams P) {
  pdl_sync();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * LN_ROWS_PER_BLOCK + warp;
  if (row >= P.M) return;
  const int nvec = P.H >> 3;
  const T* x = reinterpret_cast<const T*>(P.x) + static_cast<size_t>(row) * P.H;
  const T* res = P.residual ? reinterpret_cast<const T*>(P.residual) + static_cast<size_t>(row) * P.H : nullptr;
  const bool drop = P.p > 0.f;
  const float scale = drop ? 1.f / (1.f - P.p) : 1.f;
  const uint64_t seed = drop ? philox_seed(P.seed) : 0ull;
  const uint32_t thr = dropout_thr16(P.p);
  float v[NV][8];
  float sum = 0.f;
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int vi = j * 32 + lane;
    if (vi < nvec) {
      float f[8];
      unpack8<T>(reinterpret_cast<const uint4*>(x)[vi], f);
      if (P.bias) {
        float b[8];
        unpack8<T>(reinterpret_cast<const uint4*>(P.bias)[vi], b);
#pragma unroll
        for (int k = 0; k < 8; ++k) f[k] += b[k];
      }
      if (drop && !P.post_dropout) {
        float mk[8];
        philox_keep8(seed, P.site, (static_cast<uint64_t>(row) * P.H + vi * 8) >> 3, thr, scale, mk);
#pragma unroll
        for (int k = 0; k < 8; ++k) f[k] *= mk[k];
      }
      if (res) {
        float r[8];
        unpack8<T>(reinterpret_cast<const uint4*>(res)[vi], r);
#pragma unroll
        for (int k = 0; k < 8; ++k) f[k] += r[k];
      }
      if (P.z) reinterpret_cast<uint4*>(reinterpret_cast<T*>(P.z) + static_cast<size_t>(row) * P.H)[vi] = pack8<T>(f);
#pragma unroll
      for (int k = 0; k < 8; ++k) { v[j][k] = f[k]; sum += f[k]; }
    }
  }
  const float mean = warp_sum(sum) / P.H;
  float sq = 0.f;
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    if (j * 32 + lane < nvec) {
#pragma unroll
      for (int k = 0; k < 8; ++k) { const float d = v[j][k] - mean; sq += d * d; }
    }
  }
  const float rstd = rsqrtf(warp_sum(sq) / P.H + P.eps);
  if (lane == 0) {
    if (P.mean) P.mean[row] = mean;
    if (P.rstd) P.rstd[row] = rstd;
  }
  T* y = reinterpret_cast<T*>(P.y) + static_cast<size_t>(row) * P.H;
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int vi = j * 32 + lane;
    if (vi < nvec) {
      float g[8], b[8], o[8];
      unpack8<T>(reinterpret_cast<const uint4*>(P.gamma)[vi], g);
      unpack8<T>(reinterpret_cast<const uint4*>(P.beta)[vi], b);
#pragma unroll
      for (int k = 0; k < 8; ++k) o[k] = g[k] * ((v[j][k] - mean) * rstd) + b[k];
      if (drop && P.post_dropout) {
        float mk[8];
        philox_keep8(seed, P.site, (static_cast<uint64_t>(row) * P.H + vi * 8) >> 3, thr, scale, mk);
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] *= mk[k];
      }
      reinterpret_cast<uint4*>(y)[vi] = pack8<T>(o);
    }
  }
}

struct LnBwdParams {
  const void* dy; const void* z; const void* gamma; const float* mean; const float* rstd;
  void* dz;       // gradient w.r.t. z (the residual path), 16-bit
  void* dx;       // gradient w.r.t. x = dz * dropout mask (NULL when p == 0 and dz suffices)
  float* dgamma; float* dbeta;             // fp32 destinations, accumulated with one atomicAdd per block and column
  float* dbias;   // optional fp32 [H]: column sums of dx -- the bias gradient of the Linear that produced x
  int M, H; float p; const long long* seed; uint32_t site;
  int post_dropout;  // 1: dy is first multiplied by the mask (embeddings: y = dropout(LN(z)))
};

constexpr int LNB_WARPS = 8;
constexpr int LNB_THREADS = LNB_WARPS * 32;

// Persistent: one CTA per SM, one warp per row at a time. Each warp owns a LNB_STAGES-deep ring of
// (dy row, z row) pairs in shared memory that lane 0 fills with untiled bulk copies (cp.async.bulk +
// mbarrier complete_tx), so up to 4 rows x 8 warps x 4 KB = 128 KB of loads are in flight per SM without
// holding a single register; the per-column partials (dgamma, dbeta and the bias gradient) stay in
// registers across all rows of the warp and meet in shared memory (the retired ring) at the end.
template <typename T, int NV, int LNB_STAGES>
__global__ void __launch_bounds__(LNB_THREADS, 1) ln_bwd_kernel(LnBwdParams P, uint32_t bar_offset) {
  extern __shared__ __align__(128) uint8_t lnb_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nvec = P.H >> 3;
  const uint32_t row_bytes = static_cast<uint32_t>(P.H) * sizeof(T);
  uint8_t* ring = lnb_smem + static_cast<size_t>(warp) * LNB_STAGES * 2 * row_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(lnb_smem + bar_offset) + warp * LNB_STAGES;
  const bool drop = P.p > 0.f;
  const float scale = drop ? 1.f / (1.f - P.p) : 1.f;
  const uint64_t seed = drop ? philox_seed(P.seed) : 0ull;
  const uint32_t thr = dropout_thr16(P.p);
  const int gw = blockIdx.x * LNB_WARPS + warp, GW = gridDim.x * LNB_WARPS;

  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < LNB_STAGES; ++s) mbar_init(&bars[s], 1);
    fence_mbar_init();
  }
  pdl_sync();   // dy / z of the previous kernel are visible from here on
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < LNB_STAGES; ++s) {
      const long long row = gw + static_cast<long long>(s) * GW;
      if (row < P.M) {
        mbar_expect_tx(&bars[s], 2 * row_bytes);
        bulk_load_1d(ring + (2 * s) * row_bytes, reinterpret_cast<const T*>(P.dy) + row * P.H, row_bytes, &bars[s]);
        bulk_load_1d(ring + (2 * s + 1) * row_bytes, reinterpret_cast<const T*>(P.z) + row * P.H, row_bytes, &bars[s]);
      }
    }
  }
  __syncwarp();

  float dg[NV][8], db[NV][8], dc[NV][8];
  uint4 gpk[NV];
#pragma unroll
  for (int j = 0; j < NV; ++j) {
#pragma unroll
    for (int k = 0; k < 8; ++k) { dg[j][k] = 0.f; db[j][k] = 0.f; dc[j][k] = 0.f; }
    const int vi = j * 32 + lane;
    gpk[j] = vi < nvec ? reinterpret_cast<const uint4*>(P.gamma)[vi] : make_uint4(0, 0, 0, 0);
  }

  int stage = 0;
  uint32_t phase = 0;
  for (long long row = gw; row < P.M; row += GW) {
    const float mean = P.mean[row], rstd = P.rstd[row];
    mbar_wait(&bars[stage], phase);
    const uint4* sdy = reinterpret_cast<const uint4*>(ring + (2 * stage) * row_bytes);
    const uint4* sz = reinterpret_cast<const uint4*>(ring + (2 * stage + 1) * row_bytes);
    float xh[NV][8], gy[NV][8];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int vi = j * 32 + lane;
      if (vi < nvec) {
        float d[8], zz[8], g[8];
        unpack8<T>(sdy[vi], d);
        unpack8<T>(sz[vi], zz);
        unpack8<T>(gpk[j], g);
        if (drop && P.post_dropout) {
          float mk[8];
          philox_keep8(seed, P.site, (static_cast<uint64_t>(row) * P.H + vi * 8) >> 3, thr, scale, mk);
#pragma unroll
          for (int k = 0; k < 8; ++k) d[k] *= mk[k];
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float h = (zz[k] - mean) * rstd;
          xh[j][k] = h;
          dg[j][k] += d[k] * h;
          db[j][k] += d[k];
          const float t = d[k] * g[k];
          gy[j][k] = t;
          s1 += t;
          s2 += t * h;
        }
      }
    }
    // the ring slot is consumed: refill it with the row LNB_STAGES rounds ahead
    __syncwarp();
    if (lane == 0) {
      const long long nxt = row + static_cast<long long>(LNB_STAGES) * GW;
      if (nxt < P.M) {
        mbar_expect_tx(&bars[stage], 2 * row_bytes);
        bulk_load_1d(ring + (2 * stage) * row_bytes, reinterpret_cast<const T*>(P.dy) + nxt * P.H, row_bytes, &bars[stage]);
        bulk_load_1d(ring + (2 * stage + 1) * row_bytes, reinterpret_cast<const T*>(P.z) + nxt * P.H, row_bytes, &bars[stage]);
      }
    }
    if (++stage == LNB_STAGES) { stage = 0; phase ^= 1; }
    s1 = warp_sum(s1) / P.H;
    s2 = warp_sum(s2) / P.H;
    T* dz = reinterpret_cast<T*>(P.dz) + static_cast<size_t>(row) * P.H;
    T* dx = P.dx ? reinterpret_cast<T*>(P.dx) + static_cast<size_t>(row) * P.H : nullptr;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int vi = j * 32 + lane;
      if (vi < nvec) {
        float o[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = rstd * (gy[j][k] - s1 - xh[j][k] * s2);
        reinterpret_cast<uint4*>(dz)[vi] = pack8<T>(o);
        if (dx) {
          if (drop && !P.post_dropout) {
            float mk[8];
            philox_keep8(seed, P.site, (static_cast<uint64_t>(row) * P.H + vi * 8) >> 3, thr, scale, mk);
#pragma unroll
            for (int k = 0; k < 8; ++k) o[k] *= mk[k];
          }
          reinterpret_cast<uint4*>(dx)[vi] = pack8<T>(o);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) dc[j][k] += o[k];
      }
    }
  }

  // combine the warps' per-column partials through shared memory (the ring is retired: every bulk copy this
  // warp issued has been waited for), then one atomicAdd per block and column
  __syncthreads();
  float* sm = reinterpret_cast<float*>(lnb_smem);
  const int narr = P.dbias ? 3 : 2;
  float* mine = sm + static_cast<size_t>(warp) * narr * P.H;
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int vi = j * 32 + lane;
    if (vi < nvec) {
      float4* a = reinterpret_cast<float4*>(mine + vi * 8);
      a[0] = make_float4(dg[j][0], dg[j][1], dg[j][2], dg[j][3]);
      a[1] = make_float4(dg[j][4], dg[j][5], dg[j][6], dg[j][7]);
      float4* b = reinterpret_cast<float4*>(mine + P.H + vi * 8);
      b[0] = make_float4(db[j][0], db[j][1], db[j][2], db[j][3]);
      b[1] = make_float4(db[j][4], db[j][5], db[j][6], db[j][7]);
      if (P.dbias) {
        float4* c = reinterpret_cast<float4*>(mine + 2 * P.H + vi * 8);
        c[0] = make_float4(dc[j][0], dc[j][1], dc[j][2], dc[j][3]);
        c[1] = make_float4(dc[j][4], dc[j][5], dc[j][6], dc[j][7]);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < narr * P.H; i += blockDim.x) {
    float a = 0.f;
#pragma unroll
    for (int w = 0; w < LNB_WARPS; ++w) a += sm[static_cast<size_t>(w) * narr * P.H + i];
    const int arr = i / P.H, col = i - arr * P.H;
    float* dst = arr == 0 ? P.dgamma : (arr == 1 ? P.dbeta : P.dbias);
    if (dst) atomicAdd(dst + col, a);
  }
}

// Column sums of a 16-bit [M, N] matrix (bias gradients): block = 32 lanes x 8 rows, each lane owns 8
// consecutive columns (256 columns per block), rows strided by 8 inside a slab of rows; the block's 256
// column sums are reduced in shared memory and added to the fp32 destination with one atomicAdd each.
template <typename T>
__global__ void __launch_bounds__(256)
colsum_partial_kernel(const T* __restrict__ x, int M, int N, int ld, int rows_per_block, float* __restrict__ part) {
  __shared__ float red[8][32][9];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c0 = (blockIdx.x * 32 + tx) * 8;
  const int r0 = blockIdx.y * rows_per_block, r1 = min(M, r0 + rows_per_block);
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (c0 < N) {
    for (int r = r0 + ty; r < r1; r += 8) {
      float f[8];
      unpack8<T>(*reinterpret_cast<const uint4*>(x + static_cast<size_t>(r) * ld + c0), f);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += f[k];
    }
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) red[ty][tx][k] = acc[k];
  __syncthreads();
  // 256 threads -> 256 columns of this block
  const int col = threadIdx.x;          // 0..255
  const int lx = col >> 3, lk = col & 7;
  float t = 0.f;
#pragma unroll
  for (int y = 0; y < 8; ++y) t += red[y][lx][lk];
  const int c = blockIdx.x * 256 + col;
  if (c < N) atomicAdd(part + c, t);
}

static int ln_check(const char* who, int M, int H) {
  if (M <= 0 || H <= 0 || H % 8 != 0 || H > LN_MAXV * 256) {
    set_last_error("%s: unsupported shape M=%d H=%d (H must be a multiple of 8, <= %d)", who, M, H, LN_MAXV * 256);
    return 1;
  }
  return 0;
}

}  // namespace hs

using namespace hs;

extern "C" int hs_ln_fwd(const void* x, const void* bias, const void* residual, const void* gamma, const void* beta,
                         void* y, void* z, float* mean, float* rstd, int M, int H, float eps, float dropout_p,
                         const long long* seed, unsigned site, int post_dropout, int dtype, hs_stream_t stream) {
  if (ln_check("hs_ln_fwd", M, H)) return 1;
  if (dropout_p > 0.f && seed == nullptr) { set_last_error("hs_ln_fwd: dropout needs a seed pointer"); return 1; }
  LnFwdParams P{x, bias, residual, gamma, beta, y, z, mean, rstd, M, H, eps, dropout_p, seed, site, post_dropout};
  const int grid = (M + LN_ROWS_PER_BLOCK - 1) / LN_ROWS_PER_BLOCK;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int nv = (H + 255) / 256;
#define HS_LN_FWD(T)                                                                 \
  do {                                                                               \
    if (nv <= 1) (void)launch_k(ln_fwd_kernel<T, 1>, dim3(grid), dim3(LN_THREADS), 0, s, P);                 \
    else if (nv <= 2) (void)launch_k(ln_fwd_kernel<T, 2>, dim3(grid), dim3(LN_THREADS), 0, s, P);            \
    else if (nv <= 4) (void)launch_k(ln_fwd_kernel<T, 4>, dim3(grid), dim3(LN_THREADS), 0, s, P);            \
    else (void)launch_k(ln_fwd_kernel<T, 8>, dim3(grid), dim3(LN_THREADS), 0, s, P);                         \
  } while (0)
  if (dtype == HS_F16) HS_LN_FWD(__half);
  else if (dtype == HS_BF16) HS_LN_FWD(__nv_bfloat16);
  else { set_last_error("hs_ln_fwd: dtype must be f16 or bf16"); return 1; }
#undef HS_LN_FWD
  return post_launch("ln_fwd_kernel");
}

extern "C" int hs_ln_bwd(const void* dy, const void* z, const void* gamma, const float* mean, const float* rstd,
                         void* dz, void* dx, float* dgamma, float* dbeta, int accumulate_params,
                         float* dbias, int accumulate_bias,
                         int M, int H, float dropout_p, const long long* seed, unsigned site, int post_dropout,
                         int dtype, hs_stream_t stream) {
  if (ln_check("hs_ln_bwd", M, H)) return 1;
  if (dropout_p > 0.f && seed == nullptr) { set_last_error("hs_ln_bwd: dropout needs a seed pointer"); return 1; }
  if ((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(z)) % 16) {
    set_last_error("hs_ln_bwd: dy and z must be 16-byte aligned");
    return 1;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (!accumulate_params) {   // overwrite semantics: clear the destinations, then accumulate into them
    if (dgamma) HS_CUDA_OK(cudaMemsetAsync(dgamma, 0, sizeof(float) * H, s));
    if (dbeta) HS_CUDA_OK(cudaMemsetAsync(dbeta, 0, sizeof(float) * H, s));
  }
  if (dbias && !accumulate_bias) HS_CUDA_OK(cudaMemsetAsync(dbias, 0, sizeof(float) * H, s));
  LnBwdParams P{dy, z, gamma, mean, rstd, dz, dx, dgamma, dbeta, dbias, M, H, dropout_p, seed, site, post_dropout};
  const int nv = (H + 255) / 256;
  const int stages = nv <= 4 ? 4 : 2;     // rows in flight per warp; H = 2048 leaves room for two
  const size_t esz = 2;
  const size_t ring = static_cast<size_t>(LNB_WARPS) * stages * 2 * H * esz;
  const size_t red = static_cast<size_t>(LNB_WARPS) * 3 * H * sizeof(float);
  const size_t bar_offset = ring > red ? ring : red;
  const size_t smem = bar_offset + LNB_WARPS * stages * sizeof(uint64_t);
  if (smem > 227 * 1024) { set_last_error("hs_ln_bwd: H=%d needs %zu B of shared memory", H, smem); return 1; }
  const int rows_blocks = (M + LNB_WARPS - 1) / LNB_WARPS;
  const int grid = rows_blocks < num_sms() ? rows_blocks : num_sms();
#define HS_LN_BWD_ONE(T, NVV, ST)                                                                        \
  do {                                                                                                 \
    static bool attr_done[64] = {false};                                                               \
    int dev = 0;                                                                                       \
    HS_CUDA_OK(cudaGetDevice(&dev));                                                                   \
    if (dev < 64 && !attr_done[dev]) {                                                                 \
      HS_CUDA_OK(cudaFuncSetAttribute(ln_bwd_kernel<T, NVV, ST>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024)); \
      attr_done[dev] = true;                                                                           \
    }                                                                                                  \
    (void)launch_k(ln_bwd_kernel<T, NVV, ST>, dim3(grid), dim3(LNB_THREADS), smem, s, P, static_cast<uint32_t>(bar_offset));                                    \
  } while (0)
#define HS_LN_BWD(T)                                  \
  do {                                                \
    if (nv <= 1) HS_LN_BWD_ONE(T, 1, 4);              \
    else if (nv <= 2) HS_LN_BWD_ONE(T, 2, 4);         \
    else if (nv <= 4) HS_LN_BWD_ONE(T, 4, 4);         \
    else HS_LN_BWD_ONE(T, 8, 2);                      \
  } while (0)
  if (dtype == HS_F16) HS_LN_BWD(__half);
  else if (dtype == HS_BF16) HS_LN_BWD(__nv_bfloat16);
  else { set_last_error("hs_ln_bwd: dtype must be f16 or bf16"); return 1; }
#undef HS_LN_BWD
#undef HS_LN_BWD_ONE
  return post_launch("ln_bwd_kernel");
}

namespace hs {
int colsum_slabs(int M) {
  const int s = (M + 127) / 128;
  return s < 1 ? 1 : (s > 128 ? 128 : s);
}
}  // namespace hs

extern "C" int hs_colsum_scratch_elems(int M, int N) { return colsum_slabs(M) * N; }

extern "C" int hs_colsum(const void* x, int M, int N, int ld, float* out, int accumulate, int dtype,
                         hs_stream_t stream) {
  if (M <= 0 || N <= 0 || ld % 8 != 0) {
    set_last_error("hs_colsum: bad arguments M=%d N=%d ld=%d", M, N, ld);
    return 1;
  }
  const int slabs = colsum_slabs(M);
  const int rows_per_block = (M + slabs - 1) / slabs;
  dim3 grid((N + 255) / 256, slabs);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (!accumulate) HS_CUDA_OK(cudaMemsetAsync(out, 0, sizeof(float) * N, s));
  if (dtype == HS_F16) colsum_partial_kernel<__half><<<grid, 256, 0, s>>>((const __half*)x, M, N, ld, rows_per_block, out);
  else if (dtype == HS_BF16) colsum_partial_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>((const __nv_bfloat16*)x, M, N, ld, rows_per_block, out);
  else { set_last_error("hs_colsum: dtype must be f16 or bf16"); return 1; }
  return post_launch("colsum_partial_kernel");
}
